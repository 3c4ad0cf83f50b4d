// common.cuh -- shared device helpers for the sm_100a GStex hot-path kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/gstex_b200.h"

// after every <<<>>>: count the launch (statistics) and surface launch errors (the reference checks none)
#define GSTEX_LAUNCH_CHECK()                         \
    do {                                             \
        gstex::note_launch();                        \
        cudaError_t _e = cudaGetLastError();         \
        if (_e != cudaSuccess) return (int)_e;       \
    } while (0)

#define GSTEX_CUDA_TRY(expr)                         \
    do {                                             \
        cudaError_t _e = (expr);                     \
        if (_e != cudaSuccess) return (int)_e;       \
    } while (0)

namespace gstex {

void note_launch();  // binning.cu
// SMs of the current device (cudaDevAttrMultiProcessorCount, cached per device): persistent grids are sized from it.
int sm_count();
// Test hook (gstex_debug_force_nocull): the compositing kernels evaluate every list entry for every unit.
int force_nocull_flag();

constexpr int kThreads1D = 256;  // per-splat / per-intersection kernels
constexpr float kNear = 0.01f;   // ref: t_near in get_aabb_2d.cu:25, texture.cu:213,294
constexpr float kFar = 1000.0f;  // ref: t_far

static inline int div_up(long long a, int b) { return (int)((a + b - 1) / b); }

// ------------------------------------------------------------------------------------------------
// Per-view packed splat record (128 B = one L2 line, 16-byte aligned so a batch can be staged by
// TMA bulk copies).  Everything the per-(pixel,splat) loop of the forward needs that does not
// depend on the pixel, computed once per (view, splat) instead of once per pair as the reference
// does (texture.cu:161-185 recomputes quat->R and the mean's projection for every pixel).
// ------------------------------------------------------------------------------------------------
struct __align__(16) SplatRec {
    float4 q[8];
    // q[0] = mean.xyz, tnum      tnum = dot(ax3, mean - origin): numerator of the ray/plane hit
    // q[1] = ax3.xyz,  opacity                                   (texture_helpers.cuh:302-313)
    // q[2] = ax1.xyz,  q1        q1 = (1/(scale.x*glob))^2
    // q[3] = ax2.xyz,  q2        q2 = (1/(scale.y*glob))^2
    // q[4] = xy.x, xy.y, h, w    projected mean (blur floor, texture.cu:184-188); chart dims as int bits
    // q[5] = rgb.xyz,  start     chart start texel as int bits
    // q[6] = umap.xyz, u0
    // q[7] = vmap.xyz, v0
    // The first five quads decide alpha / skip; the last three are only read for contributing pairs.
};
static_assert(sizeof(SplatRec) == 128, "SplatRec must be one 128-byte line");

// Extra per-(view, splat) data only the backward needs (32 B).
struct __align__(16) SplatRecBwd {
    float4 quat;  // (w,x,y,z) for quat_to_rotmat_vjp (helpers.cuh:187-228)
    float4 aux;   // 1/scale.x, 1/scale.y, rw = 1/(p_view.z + 1e-6) (project_pix_vjp, helpers.cuh:155-164), unused
};
static_assert(sizeof(SplatRecBwd) == 32, "SplatRecBwd must be 32 bytes");

struct Camera {
    float vm[12];   // world-to-camera, row-major [R|t]
    float c2w[12];  // camera-to-world, row-major [R|t]
    float fx, fy, cx, cy;
};

__device__ __forceinline__ float dot3(const float a0, const float a1, const float a2, const float b0,
                                      const float b1, const float b2) {
    return a0 * b0 + a1 * b1 + a2 * b2;
}

// ref: texture_helpers.cuh:336-352 get_ray
__device__ __forceinline__ void get_ray(const Camera& cam, float px, float py, float3& origin, float3& ray) {
    origin = make_float3(cam.c2w[3], cam.c2w[7], cam.c2w[11]);
    const float u = (px - cam.cx) / cam.fx;
    const float v = (py - cam.cy) / cam.fy;
    float rx = cam.c2w[0] * u + cam.c2w[1] * v + cam.c2w[2] + cam.c2w[3];
    float ry = cam.c2w[4] * u + cam.c2w[5] * v + cam.c2w[6] + cam.c2w[7];
    float rz = cam.c2w[8] * u + cam.c2w[9] * v + cam.c2w[10] + cam.c2w[11];
    rx -= origin.x;
    ry -= origin.y;
    rz -= origin.z;
    const float nrm = sqrtf(rx * rx + ry * ry + rz * rz);
    ray = make_float3(rx / nrm, ry / nrm, rz / nrm);
}

// ref: texture_helpers.cuh:302-313 (denominator clamp of compute_tt)
__device__ __forceinline__ float clamp_tt_denom(float d) {
    // 0 <= d < eps -> eps; -eps < d <= 0 -> -eps (d == -0.0 takes the first branch, as in the reference); NaN stays
    const float eps = 1e-6f;
    return (fabsf(d) < eps) ? ((d < 0.f) ? -eps : eps) : d;
}

// ref: texture_helpers.cuh:37-51 clamp_coord (replicate padding is hard-wired on, texture.cu:65)
__device__ __forceinline__ float clamp01(float v) {
    if (v <= 0.f) v = 0.f;
    if (v >= 1.f) v = 1.f;
    return v;
}

// Channel-independent part of the bilinear jagged-chart query (texture_helpers.cuh:162-215): the
// reference recomputes this once per channel.
struct TexQuery {
    float w00, w01, w10, w11;  // weights for (i,j), (i,j+1), (i+1,j), (i+1,j+1)
    int b00, b01, b10, b11;    // texel indices (start + i*w + j), to be multiplied by C
    float fx, fy;              // fractional parts (weight.x, weight.y)
};

__device__ __forceinline__ TexQuery tex_query(int h, int w, int start, float u, float v, bool bilinear) {
    TexQuery q;
    // explicit roundings: the texel coordinate is rounded before its integer part is subtracted (no fma), as compiled
    // in the reference's forward
    const float tcx = __fmul_rn((float)h, u), tcy = __fmul_rn((float)w, v);
    int ix = (int)tcx, iy = (int)tcy;
    const int jx = min(ix + 1, h - 1), jy = min(iy + 1, w - 1);
    q.fx = __fsub_rn(tcx, (float)ix);
    q.fy = __fsub_rn(tcy, (float)iy);
    ix = min(ix, h - 1);
    iy = min(iy, w - 1);
    const float ox = __fsub_rn(1.f, q.fx), oy = __fsub_rn(1.f, q.fy);
    const float wii = __fmul_rn(ox, oy), wij = __fmul_rn(ox, q.fy);
    const float wji = __fmul_rn(q.fx, oy), wjj = __fmul_rn(q.fx, q.fy);
    q.b00 = start + ix * w + iy;
    q.b01 = start + ix * w + jy;
    q.b10 = start + jx * w + iy;
    q.b11 = start + jx * w + jy;
    if (bilinear) {
        q.w00 = wii; q.w01 = wij; q.w10 = wji; q.w11 = wjj;
    } else {  // settings bit 2: nearest texel = one-hot on the largest weight (first wins ties)
        q.w00 = q.w01 = q.w10 = q.w11 = 0.f;
        if (wii >= wij && wii >= wji && wii >= wjj) q.w00 = 1.f;
        else if (wij >= wii && wij >= wji && wij >= wjj) q.w01 = 1.f;
        else if (wji >= wii && wji >= wij && wji >= wjj) q.w10 = 1.f;
        else q.w11 = 1.f;
    }
    return q;
}

}  // namespace gstex

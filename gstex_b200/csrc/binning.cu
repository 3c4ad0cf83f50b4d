// binning.cu -- stage (a) per-splat projection / bounds / tile count and the small integer kernels of
// stage (b) (prefix sum, key emit, tile ranges).  sm_100a.  The radix sort lives in sort.cu.
//
// Bit-exactness: everything here feeds the tile|depth sort keys, so the fp32 arithmetic is written
// with explicit __fmaf_rn/__fmul_rn/__fadd_rn (never re-associated or re-contracted by the
// compiler) in exactly the fused form the reference kernel gets from nvcc 12.9 for sm_100 (read back
// from its SASS; the same form is restated for the CPU in oracle/gstex_oracle.c).
#include "common.cuh"

#include <atomic>

namespace gstex {

static std::atomic<long long> g_launches{0};
void note_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

static std::atomic<int> g_sm_count[64];
int sm_count() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    int n = g_sm_count[dev].load(std::memory_order_relaxed);
    if (n <= 0) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;  // B200
        g_sm_count[dev].store(n, std::memory_order_relaxed);
    }
    return n;
}

static std::atomic<int> g_force_nocull{0};
int force_nocull_flag() { return g_force_nocull.load(std::memory_order_relaxed); }

// ---------------------------------------------------------------------------------------------
// projection helpers
// ---------------------------------------------------------------------------------------------
// torch: view = points @ viewmat.T[:3,:3] + viewmat.T[3:,:3]  (get_aabb_2d.py:23).  cuBLAS K=3 dot
// product = fma chain in k order, translation added by a separate elementwise kernel.
__device__ __forceinline__ float row_matmul(const float* m, float x, float y, float z) {
    return __fadd_rn(__fmaf_rn(z, m[2], __fmaf_rn(y, m[1], __fmul_rn(x, m[0]))), m[3]);
}
// reference kernel transform_4x3 (helpers.cuh:124-131) as compiled: fma(z,m2,fma(x,m0,y*m1)) + m3
__device__ __forceinline__ float row_pt(const float* m, float x, float y, float z) {
    return __fadd_rn(__fmaf_rn(z, m[2], __fmaf_rn(x, m[0], __fmul_rn(y, m[1]))), m[3]);
}
// ... except row 0 of corner 00: fma(z,m2,fma(y,m1,x*m0)) + m3
__device__ __forceinline__ float row_c00x(const float* m, float x, float y, float z) {
    return __fadd_rn(__fmaf_rn(z, m[2], __fmaf_rn(y, m[1], __fmul_rn(x, m[0]))), m[3]);
}

struct BoundsParams {
    const float* means;
    const float* scales;
    const float* quats;
    float glob_scale;
    const float* vm;  // 12 floats on the device, row-major [R|t]
    float fx, fy, cx, cy;
    int n;
    int img_h, img_w, bw;
    float* xys;
    float* depths;
    float* centers;
    float* extents;
    int32_t* num_tiles_hit;
};

// ref: get_aabb_2d.py:70-91
__device__ __forceinline__ int tiles_hit(float cx, float cy, float ex, float ey, int tiles_x, int tiles_y, float fbw) {
    // torch: floor((c - e) / bw).to(int32), floor((c + e) / bw + 1).to(int32), clamp, product
    int tlx = __float2int_rz(floorf(__fdiv_rn(__fsub_rn(cx, ex), fbw)));
    int tly = __float2int_rz(floorf(__fdiv_rn(__fsub_rn(cy, ey), fbw)));
    int brx = __float2int_rz(floorf(__fadd_rn(__fdiv_rn(__fadd_rn(cx, ex), fbw), 1.0f)));
    int bry = __float2int_rz(floorf(__fadd_rn(__fdiv_rn(__fadd_rn(cy, ey), fbw), 1.0f)));
    tlx = min(max(tlx, 0), tiles_x);
    tly = min(max(tly, 0), tiles_y);
    brx = min(max(brx, 0), tiles_x);
    bry = min(max(bry, 0), tiles_y);
    return (brx - tlx) * (bry - tly);
}

__device__ __forceinline__ void project_clipped(float vx, float vy, float vz, float fx, float fy, float cx, float cy,
                                                float& ox, float& oy) {
    // clip_and_project (helpers.cuh:250-257): z = max(z, t_near); project_pix (:145-152)
    const float zc = fmaxf(vz, kNear);
    const float rw = __frcp_rn(__fadd_rn(zc, 1e-6f));
    ox = __fmaf_rn(__fmul_rn(vx, rw), fx, cx);
    oy = __fmaf_rn(__fmul_rn(vy, rw), fy, cy);
}

// One thread per splat.  Fuses project_points (a1), get_aabb_2d_kernel (a2) and get_num_tiles_hit_2d (a3).
__global__ void __launch_bounds__(kThreads1D) project_bounds_kernel(const BoundsParams p) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n) return;
    const float px = p.means[3 * i], py = p.means[3 * i + 1], pz = p.means[3 * i + 2];
    float m[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) m[k] = __ldg(p.vm + k);

    if (p.xys != nullptr || p.depths != nullptr) {
        const float vz = row_matmul(m + 8, px, py, pz);
        if (p.depths) p.depths[i] = vz;
        if (p.xys) {  // _torch_impl.py:140-147: separate torch ops, nothing contracted
            const float vx = row_matmul(m + 0, px, py, pz), vy = row_matmul(m + 4, px, py, pz);
            const float rw = __fdiv_rn(1.0f, __fadd_rn(vz, 1e-6f));
            reinterpret_cast<float2*>(p.xys)[i] =
                make_float2(__fadd_rn(__fmul_rn(__fmul_rn(vx, rw), p.fx), p.cx),
                            __fadd_rn(__fmul_rn(__fmul_rn(vy, rw), p.fy), p.cy));
        }
    }
    if (p.centers == nullptr && p.extents == nullptr && p.num_tiles_hit == nullptr) return;

    // ---- get_aabb_2d_kernel, get_aabb_2d.cu:11-89
    const float vx = row_pt(m + 0, px, py, pz), vy = row_pt(m + 4, px, py, pz), vz = row_pt(m + 8, px, py, pz);
    const bool not_clipped = !(vz <= kNear);
    const float rw0 = __frcp_rn(__fadd_rn(vz, 1e-6f));
    const float xy_x = __fmaf_rn(__fmul_rn(vx, rw0), p.fx, p.cx);
    const float xy_y = __fmaf_rn(__fmul_rn(vy, rw0), p.fy, p.cy);

    const float4 q = reinterpret_cast<const float4*>(p.quats)[i];
    const float w = q.x, x = q.y, y = q.z, z = q.w;
    const float zz = __fmul_rn(z, z);
    float t;
    t = __fmaf_rn(y, y, zz);
    const float a1x = __fsub_rn(1.0f, __fadd_rn(t, t));
    t = __fmaf_rn(x, y, __fmul_rn(w, z));
    const float a1y = __fadd_rn(t, t);
    t = __fmaf_rn(x, z, -__fmul_rn(w, y));
    const float a1z = __fadd_rn(t, t);
    t = __fmaf_rn(x, y, -__fmul_rn(w, z));
    const float a2x = __fadd_rn(t, t);
    t = __fmaf_rn(x, x, zz);
    const float a2y = __fsub_rn(1.0f, __fadd_rn(t, t));
    t = __fmaf_rn(w, x, __fmul_rn(y, z));
    const float a2z = __fadd_rn(t, t);

    const float ell = __fmul_rn(p.glob_scale, 3.0f);
    const float sx3 = __fmul_rn(ell, p.scales[3 * i]), sy3 = __fmul_rn(ell, p.scales[3 * i + 1]);
    const float Ax = __fmaf_rn(a1x, sx3, px), Ay = __fmaf_rn(a1y, sx3, py), Az = __fmaf_rn(a1z, sx3, pz);
    const float Bx = __fmaf_rn(a1x, -sx3, px), By = __fmaf_rn(a1y, -sx3, py), Bz = __fmaf_rn(a1z, -sx3, pz);

    float maxx, maxy, minx, miny;
    {
        float cxs[4], cys[4], czs[4];
        cxs[0] = __fmaf_rn(a2x, sy3, Ax);  cys[0] = __fmaf_rn(a2y, sy3, Ay);  czs[0] = __fmaf_rn(a2z, sy3, Az);
        cxs[1] = __fmaf_rn(a2x, -sy3, Ax); cys[1] = __fmaf_rn(a2y, -sy3, Ay); czs[1] = __fmaf_rn(a2z, -sy3, Az);
        cxs[2] = __fmaf_rn(a2x, sy3, Bx);  cys[2] = __fmaf_rn(a2y, sy3, By);  czs[2] = __fmaf_rn(a2z, sy3, Bz);
        cxs[3] = __fmaf_rn(a2x, -sy3, Bx); cys[3] = __fmaf_rn(a2y, -sy3, By); czs[3] = __fmaf_rn(a2z, -sy3, Bz);
        float pxs[4], pys[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const float cvx = (c == 0) ? row_c00x(m + 0, cxs[c], cys[c], czs[c]) : row_pt(m + 0, cxs[c], cys[c], czs[c]);
            const float cvy = row_pt(m + 4, cxs[c], cys[c], czs[c]);
            const float cvz = row_pt(m + 8, cxs[c], cys[c], czs[c]);
            project_clipped(cvx, cvy, cvz, p.fx, p.fy, p.cx, p.cy, pxs[c], pys[c]);
        }
        maxx = fmaxf(fmaxf(pxs[0], pxs[1]), fmaxf(pxs[2], pxs[3]));
        maxy = fmaxf(fmaxf(pys[0], pys[1]), fmaxf(pys[2], pys[3]));
        minx = fminf(fminf(pxs[0], pxs[1]), fminf(pxs[2], pxs[3]));
        miny = fminf(fminf(pys[0], pys[1]), fminf(pys[2], pys[3]));
    }
    float2 center, extent;
    if (!not_clipped) {  // get_aabb_2d.cu:81-84
        center = make_float2(xy_x, xy_y);
        extent = make_float2(0.f, 0.f);
    } else {
        center = make_float2(__fmul_rn(__fadd_rn(maxx, minx), 0.5f), __fmul_rn(__fadd_rn(maxy, miny), 0.5f));
        extent = make_float2(__fmul_rn(__fsub_rn(maxx, minx), 0.5f), __fmul_rn(__fsub_rn(maxy, miny), 0.5f));
    }
    if (p.centers) reinterpret_cast<float2*>(p.centers)[i] = center;
    if (p.extents) reinterpret_cast<float2*>(p.extents)[i] = extent;
    if (p.num_tiles_hit) {
        const int tiles_x = (p.img_w + p.bw - 1) / p.bw, tiles_y = (p.img_h + p.bw - 1) / p.bw;
        p.num_tiles_hit[i] = tiles_hit(center.x, center.y, extent.x, extent.y, tiles_x, tiles_y, (float)p.bw);
    }
}

__global__ void __launch_bounds__(kThreads1D)
num_tiles_hit_kernel(const float2* __restrict__ centers, const float2* __restrict__ extents, int n, int tiles_x,
                     int tiles_y, float fbw, int32_t* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float2 c = centers[i], e = extents[i];
    out[i] = tiles_hit(c.x, c.y, e.x, e.y, tiles_x, tiles_y, fbw);
}

}  // namespace gstex

using namespace gstex;


static int check_bw(int bw) { return (bw >= 2 && bw <= 16) ? GSTEX_OK : GSTEX_E_BADARG; }

extern "C" int gstex_debug_force_nocull(int on) {
    return gstex::g_force_nocull.exchange(on ? 1 : 0, std::memory_order_relaxed);
}

extern "C" int gstex_project_bounds(const float* means, const float* scales, float glob_scale, const float* quats,
                                    const float* viewmat, float fx, float fy, float cx, float cy, int n,
                                    int img_height, int img_width, int block_width, float* xys, float* depths,
                                    float* centers, float* extents, int32_t* num_tiles_hit, void* stream) {
    if (n < 0 || means == nullptr || viewmat == nullptr) return GSTEX_E_BADARG;
    const bool need_aabb = centers || extents || num_tiles_hit;
    if (need_aabb && (scales == nullptr || quats == nullptr)) return GSTEX_E_BADARG;
    if (num_tiles_hit && (check_bw(block_width) != GSTEX_OK || img_height <= 0 || img_width <= 0)) return GSTEX_E_BADARG;
    if (n == 0) return GSTEX_OK;
    BoundsParams p;
    p.means = means; p.scales = scales; p.quats = quats; p.glob_scale = glob_scale; p.vm = viewmat;
    p.fx = fx; p.fy = fy; p.cx = cx; p.cy = cy; p.n = n;
    p.img_h = img_height; p.img_w = img_width; p.bw = block_width > 0 ? block_width : 16;
    p.xys = xys; p.depths = depths; p.centers = centers; p.extents = extents; p.num_tiles_hit = num_tiles_hit;
    project_bounds_kernel<<<div_up(n, kThreads1D), kThreads1D, 0, (cudaStream_t)stream>>>(p);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

extern "C" int gstex_project_points(const float* means, const float* viewmat, float fx, float fy, float cx, float cy,
                                    int n, float* xys, float* depths, void* stream) {
    return gstex_project_bounds(means, nullptr, 1.f, nullptr, viewmat, fx, fy, cx, cy, n, 0, 0, 0, xys, depths, nullptr,
                                nullptr, nullptr, stream);
}

extern "C" int gstex_get_aabb_2d(const float* means, const float* scales, float glob_scale, const float* quats,
                                 const float* viewmat, float fx, float fy, float cx, float cy, int n, float* centers,
                                 float* extents, void* stream) {
    if (centers == nullptr || extents == nullptr) return GSTEX_E_BADARG;
    return gstex_project_bounds(means, scales, glob_scale, quats, viewmat, fx, fy, cx, cy, n, 0, 0, 0, nullptr, nullptr,
                                centers, extents, nullptr, stream);
}

extern "C" int gstex_get_num_tiles_hit_2d(const float* centers, const float* extents, int n, int img_height,
                                          int img_width, int block_width, int32_t* num_tiles_hit, void* stream) {
    if (n < 0 || centers == nullptr || extents == nullptr || num_tiles_hit == nullptr) return GSTEX_E_BADARG;
    if (check_bw(block_width) != GSTEX_OK || img_height <= 0 || img_width <= 0) return GSTEX_E_BADARG;
    if (n == 0) return GSTEX_OK;
    const int tiles_x = (img_width + block_width - 1) / block_width, tiles_y = (img_height + block_width - 1) / block_width;
    num_tiles_hit_kernel<<<div_up(n, kThreads1D), kThreads1D, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const float2*>(centers), reinterpret_cast<const float2*>(extents), n, tiles_x, tiles_y,
        (float)block_width, num_tiles_hit);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

extern "C" int gstex_abi_version(void) { return GSTEX_B200_ABI_VERSION; }

extern "C" long long gstex_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

extern "C" const char* gstex_error_string(int code) {
    if (code == GSTEX_OK) return "ok";
    if (code == GSTEX_E_BADARG) return "gstex_b200: bad argument (null pointer, negative size or block_width outside [2,16])";
    if (code == GSTEX_E_UNSUPPORTED) return "gstex_b200: unsupported settings bits (viewer visualisation 16-29) or channel count";
    if (code == GSTEX_E_WORKSPACE) return "gstex_b200: workspace too small";
    if (code == GSTEX_E_TOO_LARGE)
        return "gstex_b200: size exceeds the fixed capacity of a no-allocation entry point (gstex_cumsum_noalloc_max_n)";
    if (code > 0) return cudaGetErrorString((cudaError_t)code);
    return "gstex_b200: unknown error";
}

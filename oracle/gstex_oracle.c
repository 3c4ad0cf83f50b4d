/*
 * gstex_oracle.c -- CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
 *
 * A plain-C restatement of the reference's hot path (nvnhat95/GStex, submodule GStex_cuda):
 *   project -> screen AABB -> tile count -> cumsum -> key emit -> stable sort -> tile bins
 *   -> textured forward compositing -> backward.
 * It follows the semantics of the reference's CUDA kernels (NOT of its approximate pure-PyTorch
 * twin): alpha clamp 0.99, blur floor (settings bit 9), NDC distortion depth (bit 10), the
 * "next_T tested before skip" quirk, final_index / depth_index as GLOBAL sorted indices.
 * Every function cites the reference file:line it restates; paths are relative to
 * /root/reference/submodules/GStex_cuda/gstex_cuda/ ("C/" = cuda/csrc/).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product (gstex_b200/) never does and has no CPU fallback.
 *
 * Parity pinning: see oracle/README.md and tests/test_oracle_golden.py -- the restatement is
 * checked (a) against golden vectors produced here by importing the reference's own Python
 * (`_torch_impl.texture_forward` + autograd, `get_num_tiles_hit_2d`, `get_aabb_2d_torch`,
 * `torch.sort`) and (b) on the GPU box against the reference's own CUDA extension built
 * unmodified into oracle/_ref/ (oracle/build_ref.py).
 *
 * Build: gcc -O2 -fopenmp -ffp-contract=off -shared -fPIC gstex_oracle.c -o _build/libgstex_oracle.so -lm
 * (-ffp-contract=off: every fused multiply-add in this file is an explicit fmaf(), so the AABB /
 *  depth arithmetic is bit-reproducible and identical to the explicit __fmaf_rn chain of the
 *  CUDA kernels.)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct { float x, y; } f2;
typedef struct { float x, y, z; } f3;
typedef struct { float x, y, z, w; } f4;

static inline float dot3(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline float minf_(float a, float b) { return a < b ? a : b; }
static inline float maxf_(float a, float b) { return a > b ? a : b; }
static inline int mini_(int a, int b) { return a < b ? a : b; }
static inline int maxi_(int a, int b) { return a > b ? a : b; }

/* CUDA __expf(x) == ex2.approx(x * log2(e)); restated with exp2f (differs from the SFU by <= 2 ulp). */
static inline float cuda_expf_fast(float x) { return exp2f(x * 1.4426950408889634f); }

/* float -> int conversion with CUDA semantics (cvt.rzi.s32.f32: saturating, NaN -> 0). */
static inline int f2i_rz_sat(float v) {
    if (v != v) return 0;
    if (v >= 2147483648.0f) return 2147483647;
    if (v <= -2147483648.0f) return (-2147483647 - 1);
    return (int)v;
}

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void orc_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* ------------------------------------------------------------------------------------------
 * (a1) project_points -- get_aabb_2d.py:22-32 + _torch_impl.py:140-147 (project_pix), clip=False.
 *   view = points @ viewmat.T[:3,:3] + viewmat.T[3:,:3];  depths = view[:,2]
 *   rw = 1/(z + 1e-6); xys = (x*rw*fx + cx, y*rw*fy + cy)   (separate torch ops: no contraction)
 * The K=3 matmul is restated as the fused chain fma(z,m2,fma(y,m1,x*m0)) followed by a separate
 * add of the translation (the add is its own torch op).  tests/test_binning_gpu.py checks on the
 * B200 that torch's own matmul produces these bits.
 * ------------------------------------------------------------------------------------------ */
static inline float row_dot_fma(const float* m, float x, float y, float z) {
    return fmaf(z, m[2], fmaf(y, m[1], x * m[0]));
}

void orc_project_points(const float* means, const float* viewmat, float fx, float fy, float cx,
                        float cy, int n, float* xys, float* depths) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < n; ++i) {
        const float x = means[3 * i], y = means[3 * i + 1], z = means[3 * i + 2];
        const float vx = row_dot_fma(viewmat + 0, x, y, z) + viewmat[3];
        const float vy = row_dot_fma(viewmat + 4, x, y, z) + viewmat[7];
        const float vz = row_dot_fma(viewmat + 8, x, y, z) + viewmat[11];
        const float rw = 1.0f / (vz + 1e-6f);
        if (xys) {
            xys[2 * i + 0] = (vx * rw) * fx + cx;
            xys[2 * i + 1] = (vy * rw) * fy + cy;
        }
        if (depths) depths[i] = vz;
    }
}

/* ------------------------------------------------------------------------------------------
 * (a2) get_aabb_2d -- C/get_aabb_2d.cu:11-89 with helpers C/helpers.cuh:124-131 (transform_4x3),
 * :145-152 (project_pix), :166-185 (quat_to_rotmat), :240-257 (clip_near_plane / clip_and_project).
 * The fused-multiply-add placement below is the one nvcc 12.9 -O3 (default -fmad=true) gives the
 * reference kernel for sm_100, read back from the SASS of the reference extension built by
 * oracle/build_ref.py (cuobjdump -sass oracle/_ref/gstex_ref_C.so); every fma is explicit so gcc
 * (-ffp-contract=off) and our CUDA kernel (__fmaf_rn / __fmul_rn / __fadd_rn) give the same bits.
 * tests/test_binning_gpu.py::test_aabb_matches_reference_extension pins it on the B200.
 * ------------------------------------------------------------------------------------------ */
static inline float row_pt(const float* m, float x, float y, float z) {
    /* mean point and corners 01/10/11, all rows: fma(z,m2, fma(x,m0, y*m1)) + m3 */
    return fmaf(z, m[2], fmaf(x, m[0], y * m[1])) + m[3];
}
static inline float row_c00x(const float* m, float x, float y, float z) {
    /* corner 00, row 0 only: fma(z,m2, fma(y,m1, x*m0)) + m3 */
    return fmaf(z, m[2], fmaf(y, m[1], x * m[0])) + m[3];
}
static inline void proj_clipped(float vx, float vy, float vz, float fx, float fy, float cx, float cy,
                                float* ox, float* oy) {
    const float zc = fmaxf(vz, 0.01f);
    const float rw = 1.0f / (zc + 1e-6f);
    *ox = fmaf(vx * rw, fx, cx);
    *oy = fmaf(vy * rw, fy, cy);
}

void orc_get_aabb_2d(const float* means, const float* scales, float glob_scale, const float* quats,
                     const float* viewmat, float fx, float fy, float cx, float cy, int n,
                     float* centers, float* extents) {
    const float* m = viewmat;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < n; ++i) {
        const float px = means[3 * i], py = means[3 * i + 1], pz = means[3 * i + 2];
        const float vx = row_pt(m + 0, px, py, pz), vy = row_pt(m + 4, px, py, pz);
        const float vz = row_pt(m + 8, px, py, pz);
        const int not_clipped = !(vz <= 0.01f); /* FSETP.GTU: NaN counts as not clipped */
        const float rw0 = 1.0f / (vz + 1e-6f);
        const float xy_x = fmaf(vx * rw0, fx, cx), xy_y = fmaf(vy * rw0, fy, cy);

        const float w = quats[4 * i], x = quats[4 * i + 1], y = quats[4 * i + 2], z = quats[4 * i + 3];
        const float zz = z * z;
        float t;
        t = fmaf(y, y, zz); const float a1x = 1.0f - (t + t);
        t = fmaf(x, y, w * z); const float a1y = t + t;
        t = fmaf(x, z, -(w * y)); const float a1z = t + t;
        t = fmaf(x, y, -(w * z)); const float a2x = t + t;
        t = fmaf(x, x, zz); const float a2y = 1.0f - (t + t);
        t = fmaf(w, x, y * z); const float a2z = t + t;
        const float ell = glob_scale * 3.0f;
        const float sx3 = ell * scales[3 * i], sy3 = ell * scales[3 * i + 1];
        const float Ax = fmaf(a1x, sx3, px), Ay = fmaf(a1y, sx3, py), Az = fmaf(a1z, sx3, pz);
        const float Bx = fmaf(a1x, -sx3, px), By = fmaf(a1y, -sx3, py), Bz = fmaf(a1z, -sx3, pz);
        float cxs[4], cys[4], czs[4];
        cxs[0] = fmaf(a2x, sy3, Ax);  cys[0] = fmaf(a2y, sy3, Ay);  czs[0] = fmaf(a2z, sy3, Az);
        cxs[1] = fmaf(a2x, -sy3, Ax); cys[1] = fmaf(a2y, -sy3, Ay); czs[1] = fmaf(a2z, -sy3, Az);
        cxs[2] = fmaf(a2x, sy3, Bx);  cys[2] = fmaf(a2y, sy3, By);  czs[2] = fmaf(a2z, sy3, Bz);
        cxs[3] = fmaf(a2x, -sy3, Bx); cys[3] = fmaf(a2y, -sy3, By); czs[3] = fmaf(a2z, -sy3, Bz);
        float pxs[4], pys[4];
        for (int c = 0; c < 4; ++c) {
            const float cvx = (c == 0) ? row_c00x(m + 0, cxs[c], cys[c], czs[c])
                                       : row_pt(m + 0, cxs[c], cys[c], czs[c]);
            const float cvy = row_pt(m + 4, cxs[c], cys[c], czs[c]);
            const float cvz = row_pt(m + 8, cxs[c], cys[c], czs[c]);
            proj_clipped(cvx, cvy, cvz, fx, fy, cx, cy, &pxs[c], &pys[c]);
        }
        const float maxx = fmaxf(fmaxf(pxs[0], pxs[1]), fmaxf(pxs[2], pxs[3]));
        const float maxy = fmaxf(fmaxf(pys[0], pys[1]), fmaxf(pys[2], pys[3]));
        const float minx = fminf(fminf(pxs[0], pxs[1]), fminf(pxs[2], pxs[3]));
        const float miny = fminf(fminf(pys[0], pys[1]), fminf(pys[2], pys[3]));
        if (!not_clipped) {
            centers[2 * i] = xy_x; centers[2 * i + 1] = xy_y;
            extents[2 * i] = 0.f; extents[2 * i + 1] = 0.f;
        } else {
            centers[2 * i] = (maxx + minx) * 0.5f; centers[2 * i + 1] = (maxy + miny) * 0.5f;
            extents[2 * i] = (maxx - minx) * 0.5f; extents[2 * i + 1] = (maxy - miny) * 0.5f;
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * (a3) get_num_tiles_hit_2d -- get_aabb_2d.py:70-91.
 *   tl = floor((c - e)/bw).int(); br = floor((c + e)/bw + 1).int(); clamp to [0, tiles]; n = dx*dy
 * ------------------------------------------------------------------------------------------ */
void orc_get_num_tiles_hit_2d(const float* centers, const float* extents, int n, int img_h,
                              int img_w, int bw, int32_t* num_tiles_hit) {
    const int tiles_x = (img_w + bw - 1) / bw, tiles_y = (img_h + bw - 1) / bw;
    const float fbw = (float)bw;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < n; ++i) {
        const float cxv = centers[2 * i], cyv = centers[2 * i + 1];
        const float ex = extents[2 * i], ey = extents[2 * i + 1];
        int tlx = f2i_rz_sat(floorf((cxv - ex) / fbw));
        int tly = f2i_rz_sat(floorf((cyv - ey) / fbw));
        int brx = f2i_rz_sat(floorf((cxv + ex) / fbw + 1.0f));
        int bry = f2i_rz_sat(floorf((cyv + ey) / fbw + 1.0f));
        tlx = mini_(maxi_(tlx, 0), tiles_x);
        tly = mini_(maxi_(tly, 0), tiles_y);
        brx = mini_(maxi_(brx, 0), tiles_x);
        bry = mini_(maxi_(bry, 0), tiles_y);
        num_tiles_hit[i] = (brx - tlx) * (bry - tly);
    }
}

/* (a4) compute_cumulative_intersects -- utils.py:40-59: inclusive int32 cumsum; returns cum[-1]. */
int32_t orc_cumsum_i32(const int32_t* in, int n, int32_t* out) {
    int32_t acc = 0;
    for (int i = 0; i < n; ++i) {
        acc += in[i];
        out[i] = acc;
    }
    return acc;
}

/* ------------------------------------------------------------------------------------------
 * (a5) map_gaussian_to_intersects -- C/forward.cu:13-71, bbox math C/helpers.cuh:37-51,75-92,
 *      zero-initialised outputs C/bindings.cu:99-102.
 *   Splats with extent.x <= 1e-4 && extent.y <= 1e-4 write nothing (their reserved slots keep
 *   key 0 / id 0: the "zero-key quirk").  The extent test compares against the DOUBLE literal
 *   1e-4 (the reference writes `extent.x <= 1e-4`), so the float is promoted.
 *   tile_min/max = clamp((int)(c/bw -/+ e/bw [+1]), 0, tiles) with truncation toward zero.
 *   Divergence (documented): if this bbox yields MORE tiles than get_num_tiles_hit_2d reserved
 *   (the two formulas round differently) the reference overruns into the next splat's slots
 *   (a data race / out-of-bounds write); here and in the CUDA product the emit stops at the
 *   reserved end.
 * ------------------------------------------------------------------------------------------ */
void orc_map_gaussian_to_intersects(int n, int num_intersects, const float* centers,
                                    const float* extents, const float* depths,
                                    const int32_t* cum_tiles_hit, int tiles_x, int tiles_y, int bw,
                                    int64_t* isect_ids, int32_t* gaussian_ids) {
    memset(isect_ids, 0, sizeof(int64_t) * (size_t)num_intersects);
    memset(gaussian_ids, 0, sizeof(int32_t) * (size_t)num_intersects);
    const float fbw = (float)bw;
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < n; ++idx) {
        const float ex = extents[2 * idx], ey = extents[2 * idx + 1];
        if ((double)ex <= 1e-4 && (double)ey <= 1e-4) continue;
        const float tcx = centers[2 * idx] / fbw, tcy = centers[2 * idx + 1] / fbw;
        const float tex = ex / fbw, tey = ey / fbw;
        const int min_x = mini_(maxi_(0, f2i_rz_sat(tcx - tex)), tiles_x);
        const int max_x = mini_(maxi_(0, f2i_rz_sat(tcx + tex + 1.0f)), tiles_x);
        const int min_y = mini_(maxi_(0, f2i_rz_sat(tcy - tey)), tiles_y);
        const int max_y = mini_(maxi_(0, f2i_rz_sat(tcy + tey + 1.0f)), tiles_y);
        int32_t cur = (idx == 0) ? 0 : cum_tiles_hit[idx - 1];
        const int32_t end = mini_(cum_tiles_hit[idx], num_intersects);
        int32_t dbits;
        memcpy(&dbits, &depths[idx], 4);
        const int64_t depth_id = (int64_t)dbits; /* sign-extended, C/forward.cu:47 */
        for (int i = min_y; i < max_y; ++i) {
            for (int j = min_x; j < max_x; ++j) {
                if (cur >= end) break;
                const int64_t tile_id = (int64_t)(i * tiles_x + j);
                isect_ids[cur] = (tile_id << 32) | depth_id;
                gaussian_ids[cur] = idx;
                ++cur;
            }
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * (a6) torch.sort(isect_ids) + torch.gather(gaussian_ids) -- utils.py:159-160.
 *   Ascending signed 64-bit sort; ties keep emission order (stable), which is what the CUDA radix
 *   path of torch.sort does.  LSD radix, 8 x 8-bit digits, constant digits skipped.
 * ------------------------------------------------------------------------------------------ */
void orc_sort_isects(int n, const int64_t* keys_in, const int32_t* ids_in, int64_t* keys_out,
                     int32_t* ids_out) {
    if (n <= 0) return;
    uint64_t* ka = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)n);
    uint64_t* kb = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)n);
    int32_t* va = (int32_t*)malloc(sizeof(int32_t) * (size_t)n);
    int32_t* vb = (int32_t*)malloc(sizeof(int32_t) * (size_t)n);
    for (int i = 0; i < n; ++i) {
        ka[i] = (uint64_t)keys_in[i] ^ 0x8000000000000000ull; /* signed order */
        va[i] = ids_in[i];
    }
    for (int pass = 0; pass < 8; ++pass) {
        size_t hist[256];
        memset(hist, 0, sizeof(hist));
        const int sh = pass * 8;
        for (int i = 0; i < n; ++i) hist[(ka[i] >> sh) & 255]++;
        int constant = 0;
        for (int d = 0; d < 256; ++d)
            if (hist[d] == (size_t)n) constant = 1;
        if (constant) continue;
        size_t sum = 0;
        for (int d = 0; d < 256; ++d) {
            size_t c = hist[d];
            hist[d] = sum;
            sum += c;
        }
        for (int i = 0; i < n; ++i) {
            size_t p = hist[(ka[i] >> sh) & 255]++;
            kb[p] = ka[i];
            vb[p] = va[i];
        }
        uint64_t* tk = ka; ka = kb; kb = tk;
        int32_t* tv = va; va = vb; vb = tv;
    }
    for (int i = 0; i < n; ++i) {
        keys_out[i] = (int64_t)(ka[i] ^ 0x8000000000000000ull);
        ids_out[i] = va[i];
    }
    free(ka); free(kb); free(va); free(vb);
}

/* (a7) get_tile_bin_edges -- C/forward.cu:76-98; tile_bins zero-initialised (C/bindings.cu:129). */
void orc_get_tile_bin_edges(int num_intersects, const int64_t* keys_sorted, int num_tiles,
                            int32_t* tile_bins) {
    memset(tile_bins, 0, sizeof(int32_t) * 2 * (size_t)num_tiles);
    for (int idx = 0; idx < num_intersects; ++idx) {
        const int32_t cur = (int32_t)(keys_sorted[idx] >> 32);
        if (idx == 0) tile_bins[2 * cur + 0] = 0;
        if (idx == num_intersects - 1) tile_bins[2 * cur + 1] = num_intersects;
        if (idx == 0) continue;
        const int32_t prev = (int32_t)(keys_sorted[idx - 1] >> 32);
        if (prev != cur) {
            tile_bins[2 * prev + 1] = idx;
            tile_bins[2 * cur + 0] = idx;
        }
    }
}

// loss.cu -- the step on the far side of the rasterizer: compositing of the rendered
// layers into the predicted image and the photometric loss of the reference trainer, forward and backward.
//
//   rgb  = clamp(out_img + out_texture[:,:,0:3] + (1 - out_alpha) * background, 0, 1)      NS/models/gstex.py:1204-1205
//   Ll1  = mean |gt - rgb|,   S = SSIM(gt, rgb)                                             NS/models/gstex.py:1301-1302
//   main = (1 - lambda) * Ll1 + lambda * (1 - S)                                            NS/models/gstex.py:1319
//
// SSIM is `pytorch_msssim.SSIM(data_range=1.0, size_average=True, channel=3)` (NS/models/gstex.py:33,351): an 11-tap
// Gaussian window (sigma 1.5) applied separably WITHOUT padding, K = (0.01, 0.03), mean of the per-channel means
// over the (H-10) x (W-10) valid positions.  pytorch_msssim is a third-party dependency of the reference
// (pyproject.toml:60, no version pinned) that is not under /root/reference; its published algorithm (v1.0.0,
// ssim.py `_fspecial_gauss_1d`, `gaussian_filter`, `_ssim`) is restated here.
//
// The reference runs this as ~45 torch kernels forward (two depthwise convolutions for each of five moments, the
// elementwise SSIM algebra, the compose/clamp/abs/mean chain) and about twice that in autograd's backward, each a pass
// over the image.  Here: one forward kernel (compose + clamp + |gt - rgb| + both separable passes in shared memory +
// the SSIM algebra + the three partial-derivative maps the backward needs), a one-block reduction of the per-block
// partial sums in a fixed order (deterministic), and one backward kernel that pulls the derivative maps back through
// the window (again separably in shared memory) and writes v_out_img / v_out_texture / v_out_alpha in the same pass.
// Algorithmic bytes: forward reads 4*(10 + C) B and writes 48 B per pixel, backward reads 4*(19 + C) B and writes
// 4*(4 + C) B; measured limiter: instruction issue (330 multiply-adds per pixel and channel in the two passes alone),
// see the kernel comments.  sm_100a.
#include "common.cuh"

namespace gstex {

constexpr int kWin = 11;               // pytorch_msssim default win_size
constexpr int kTileH = 16;             // owned pixels per block: 16 rows x 32 columns
#ifndef GSTEX_LOSS_TILE_W
#define GSTEX_LOSS_TILE_W 32
#endif
#ifndef GSTEX_LOSS_OUT2
#define GSTEX_LOSS_OUT2 8
#endif
constexpr int kTileW = GSTEX_LOSS_TILE_W;
constexpr int kHaloH = kTileH + kWin - 1;  // 26
constexpr int kHaloW = kTileW + kWin - 1;  // 42
constexpr int kHaloE = kHaloW * 3;     // 126 floats per halo row (three interleaved channels)
constexpr int kOwnE = kTileW * 3;      // 96 floats per owned row
constexpr int kPitch = kHaloE + 1;     // shared-memory row pitch, in elements
constexpr int kLossThreads = 256;
constexpr int kOut1 = 8;               // outputs per thread of the pass along H (two halves of the 16 rows)
constexpr int kOut2 = GSTEX_LOSS_OUT2; // outputs per thread of the pass along W
static_assert(kTileW % kOut2 == 0 && 3 * (kTileW / kOut2) * kTileH <= kLossThreads && 2 * kHaloE <= kLossThreads &&
              (kOwnE % 4) == 0, "loss tile geometry");

struct LossWindow {
    float g[kWin];
    float2 g2[kWin];  // (g, g): operand of the packed FFMA2 of sm_100
};

struct LossDims {
    int H, W, C;       // image size, channels of out_texture (>= 3; only the first three are composited)
    int Hv, Wv;        // valid SSIM positions: H - 10, W - 10
};

__device__ __forceinline__ float clamp_unit(float v) { return fminf(fmaxf(v, 0.f), 1.f); }

// Ground truth as the trainer holds it: fp32 in [0, 1], or the uint8 image of the data manager
// (NS/data/datamanagers/full_images_datamanager.py:316-333) converted the way `get_gt_img` does it
// (NS/models/gstex.py:1238-1247: `image.float() / 255.0`, which eager torch evaluates as a multiplication by the
// fp32 reciprocal of 255) -- same bits, without the two elementwise passes and the fp32 copy per view.
__device__ __forceinline__ float load_gt(const float* gt, size_t i) { return __ldg(gt + i); }
__device__ __forceinline__ float load_gt(const uint8_t* gt, size_t i) {
    return __fmul_rn((float)__ldg(gt + i), 1.f / 255.f);
}

__device__ __forceinline__ float block_sum_256(float v, float* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) red[w] = v;
    __syncthreads();
    float s = 0.f;
    if (threadIdx.x == 0)
        for (int i = 0; i < 8; ++i) s += red[i];
    return s;  // valid on thread 0
}

// Rows of an owned tile (kOwnE floats each) from shared memory to `dst` (row pitch `W * 3` floats), whole 16-byte
// vectors when the geometry allows it.
__device__ __forceinline__ void store_tile_rows(float* __restrict__ dst, const float* tile, int i0, int j0, const LossDims& d) {
    const bool vec = (d.W & 3) == 0 && j0 + kTileW <= d.W;
    if (vec) {
        for (int t = threadIdx.x; t < kTileH * (kOwnE / 4); t += kLossThreads) {
            const int r = t / (kOwnE / 4), q = t - r * (kOwnE / 4);
            if (i0 + r < d.H) {
                const float* s = tile + r * kOwnE + 4 * q;
                *reinterpret_cast<float4*>(dst + ((size_t)(i0 + r) * d.W + j0) * 3 + 4 * q) = make_float4(s[0], s[1], s[2], s[3]);
            }
        }
    } else {
        for (int t = threadIdx.x; t < kTileH * kOwnE; t += kLossThreads) {
            const int r = t / kOwnE, e = t - r * kOwnE;
            if (i0 + r < d.H && j0 + e / 3 < d.W) dst[((size_t)(i0 + r) * d.W + j0) * 3 + e] = tile[r * kOwnE + e];
        }
    }
}

struct LossFwdShared {
    float2 sxy[kHaloH][kPitch];      // (gt, predicted), both masked: [row][column * 3 + channel]
    float2 s01[kTileH][kPitch];      // after the pass along H: (E[x], E[y]) ...
    float2 s23[kTileH][kPitch];      // ... (E[x^2], E[y^2]) ...
    float s4[kTileH][kPitch];        // ... E[xy]
    float red[8];
};
static_assert(sizeof(float2) * kHaloH * kPitch >= sizeof(float) * 3 * kTileH * kOwnE, "the map staging tile aliases sxy");

// One block = 16x32 owned pixels (the SSIM positions whose window STARTS there) + the 26x42 pixels their windows cover.
// maps: [3][H][W][3] = d ssim / d mu_pred, d ssim / d E[pred^2], d ssim / d E[gt*pred] per position and channel,
// zero where the position is not a valid window start (written only when maps != nullptr).
// The kernel is instruction-issue bound (ncu, round 1: issue slots 65 % busy, DRAM 6 %), so everything here is about
// instructions per pixel: a wide tile (2.1 halo elements loaded per owned one, 1.3 for the pass along H), both passes
// register-tiled (a thread of the first walks 18 rows of one (column, channel) for 8 output rows of all five moments,
// a thread of the second walks 18 columns for 8 outputs), moments kept as float2 pairs in shared memory and registers
// so that four of the five multiply-adds per tap are two packed FFMA2 (each half an IEEE fma: the same bits as five
// scalar ones), no integer division in the load loop, and the three derivative maps staged through shared memory so
// that they leave as whole rows.  Pixels stay channel-interleaved in shared memory, as in HBM.
template <typename GT>
__global__ void __launch_bounds__(kLossThreads) loss_forward_kernel(LossDims d, LossWindow win, const float* __restrict__ out_img,
                                                           const float* __restrict__ out_texture,
                                                           const float* __restrict__ out_alpha,
                                                           const float* __restrict__ background,
                                                           const GT* __restrict__ gt, const float* __restrict__ mask,
                                                           float* __restrict__ rgb, float* __restrict__ maps,
                                                           float2* __restrict__ partials) {
    extern __shared__ __align__(16) unsigned char loss_smem[];
    LossFwdShared& sh = *reinterpret_cast<LossFwdShared*>(loss_smem);
    const int i0 = blockIdx.y * kTileH, j0 = blockIdx.x * kTileW;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float bg[3] = {__ldg(background), __ldg(background + 1), __ldg(background + 2)};
    float l1 = 0.f;
    for (int r = warp; r < kHaloH; r += kLossThreads / 32) {
        const int gi = i0 + r;
#pragma unroll
        for (int it = 0; it < (kHaloE + 31) / 32; ++it) {
            const int e = lane + 32 * it;
            if (e < kHaloE) {
                const int c = e / 3, k = e - c * 3;
                const int gj = j0 + c;
                float x = 0.f, y = 0.f;
                if (gi < d.H && gj < d.W) {
                    const size_t pix = (size_t)gi * d.W + gj;
                    // gstex.py:1204-1205 as torch evaluates it (no contraction: every product and sum rounded on its own)
                    const float om = __fsub_rn(1.f, __ldg(out_alpha + pix));
                    const float pre = __fadd_rn(__fadd_rn(__ldg(out_img + pix * 3 + k), __ldg(out_texture + pix * d.C + k)),
                                                __fmul_rn(om, k == 0 ? bg[0] : k == 1 ? bg[1] : bg[2]));
                    const float v = clamp_unit(pre);
                    const float m = mask ? __ldg(mask + pix) : 1.f;
                    x = __fmul_rn(load_gt(gt, pix * 3 + k), m);  // gstex.py:1297-1298 (gt_img * mask, pred_img * mask)
                    y = __fmul_rn(v, m);
                    if (r < kTileH && c < kTileW) {
                        rgb[pix * 3 + k] = v;
                        l1 += fabsf(__fsub_rn(x, y));
                    }
                }
                sh.sxy[r][e] = make_float2(x, y);
            }
        }
    }
    __syncthreads();
    // pass 1, along H (pytorch_msssim filters dim 2 first)
    if (threadIdx.x < 2 * kHaloE) {
        const int half = threadIdx.x / kHaloE, e = threadIdx.x - half * kHaloE, r0 = half * kOut1;
        float2 a01[kOut1], a23[kOut1];
        float a4[kOut1];
#pragma unroll
        for (int o = 0; o < kOut1; ++o) { a01[o] = make_float2(0.f, 0.f); a23[o] = make_float2(0.f, 0.f); a4[o] = 0.f; }
#pragma unroll
        for (int rin = 0; rin < kOut1 + kWin - 1; ++rin) {
            const float2 xy = sh.sxy[r0 + rin][e];
            const float2 sq = __fmul2_rn(xy, xy);         // (x*x, y*y)
            const float pxy = __fmul_rn(xy.x, xy.y);
#pragma unroll
            for (int o = 0; o < kOut1; ++o) {
                const int q = rin - o;
                if (q >= 0 && q < kWin) {
                    a01[o] = __ffma2_rn(win.g2[q], xy, a01[o]);
                    a23[o] = __ffma2_rn(win.g2[q], sq, a23[o]);
                    a4[o] = __fmaf_rn(win.g[q], pxy, a4[o]);
                }
            }
        }
#pragma unroll
        for (int o = 0; o < kOut1; ++o) { sh.s01[r0 + o][e] = a01[o]; sh.s23[r0 + o][e] = a23[o]; sh.s4[r0 + o][e] = a4[o]; }
    }
    __syncthreads();
    // pass 2, along W, then the SSIM algebra (_ssim of pytorch_msssim/ssim.py); 8 positions of one channel per thread.
    // sxy is dead from here on: its memory stages the three derivative maps of the owned tile.
    float* stage = reinterpret_cast<float*>(&sh.sxy[0][0]);  // [3][kTileH][kOwnE]
    float ssim = 0.f;
    if (threadIdx.x < 3 * (kTileW / kOut2) * kTileH) {
        const int k = threadIdx.x % 3, cg = (threadIdx.x / 3) % (kTileW / kOut2), r = threadIdx.x / (3 * (kTileW / kOut2));
        float2 a01[kOut2], a23[kOut2];
        float a4[kOut2];
#pragma unroll
        for (int o = 0; o < kOut2; ++o) { a01[o] = make_float2(0.f, 0.f); a23[o] = make_float2(0.f, 0.f); a4[o] = 0.f; }
#pragma unroll
        for (int cin = 0; cin < kOut2 + kWin - 1; ++cin) {
            const int e = (kOut2 * cg + cin) * 3 + k;
            const float2 v01 = sh.s01[r][e], v23 = sh.s23[r][e];
            const float v4 = sh.s4[r][e];
#pragma unroll
            for (int o = 0; o < kOut2; ++o) {
                const int q = cin - o;
                if (q >= 0 && q < kWin) {
                    a01[o] = __ffma2_rn(win.g2[q], v01, a01[o]);
                    a23[o] = __ffma2_rn(win.g2[q], v23, a23[o]);
                    a4[o] = __fmaf_rn(win.g[q], v4, a4[o]);
                }
            }
        }
        const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;  // (K1 * data_range)^2, (K2 * data_range)^2
        const int gi = i0 + r;
#pragma unroll
        for (int o = 0; o < kOut2; ++o) {
            const int gj = j0 + kOut2 * cg + o;
            const bool valid = gi < d.Hv && gj < d.Wv;
            // every operation rounded on its own, left to right (what the C oracle's -ffp-contract=off build does): the
            // compiler is free to contract plain C differently in the fp32 and the uint8 instantiation of this kernel
            const float mu1 = a01[o].x, mu2 = a01[o].y;
            const float mu1_sq = __fmul_rn(mu1, mu1), mu2_sq = __fmul_rn(mu2, mu2), mu12 = __fmul_rn(mu1, mu2);
            const float s1 = __fsub_rn(a23[o].x, mu1_sq), s2 = __fsub_rn(a23[o].y, mu2_sq), s12 = __fsub_rn(a4[o], mu12);
            const float A1 = __fadd_rn(__fmul_rn(2.f, mu12), C1), A2 = __fadd_rn(__fmul_rn(2.f, s12), C2);
            const float B1 = __fadd_rn(__fadd_rn(mu1_sq, mu2_sq), C1), B2 = __fadd_rn(__fadd_rn(s1, s2), C2);
            const float rB1 = __fdiv_rn(1.f, B1), rB2 = __fdiv_rn(1.f, B2);
            const float cs = __fmul_rn(A2, rB2), lum = __fmul_rn(A1, rB1);
            if (valid) ssim = __fadd_rn(ssim, __fmul_rn(lum, cs));
            // partial derivatives of lum*cs with E[y], E[y^2], E[xy] as the independent moments of the prediction y
            float dmu = 0.f, dyy = 0.f, dxy = 0.f;
            if (valid) {
                dyy = __fmul_rn(__fmul_rn(-lum, cs), rB2);
                dxy = __fmul_rn(__fmul_rn(2.f, lum), rB2);
                const float t = __fsub_rn(__fmul_rn(mu1, cs), __fmul_rn(__fmul_rn(mu2, lum), cs));
                dmu = __fsub_rn(__fsub_rn(__fmul_rn(__fmul_rn(2.f, rB1), t), __fmul_rn(__fmul_rn(2.f, mu2), dyy)),
                                __fmul_rn(mu1, dxy));
            }
            const int se = r * kOwnE + (kOut2 * cg + o) * 3 + k;
            stage[se] = dmu; stage[kTileH * kOwnE + se] = dyy; stage[2 * kTileH * kOwnE + se] = dxy;
        }
    }
    const float l1_sum = block_sum_256(l1, sh.red);      // (its barriers also publish the staged maps)
    const float ssim_sum = block_sum_256(ssim, sh.red);
    if (threadIdx.x == 0) partials[blockIdx.y * gridDim.x + blockIdx.x] = make_float2(l1_sum, ssim_sum);
    if (maps != nullptr) {
        const size_t plane = (size_t)d.H * d.W * 3;
#pragma unroll
        for (int m = 0; m < 3; ++m) store_tile_rows(maps + m * plane, stage + m * kTileH * kOwnE, i0, j0, d);
    }
}

// out[0] = Ll1, out[1] = SSIM, out[2] = main loss; partial sums added in block order, in double (deterministic)
__global__ void __launch_bounds__(256) loss_finalize_kernel(const float2* __restrict__ partials, int n, LossDims d,
                                                            float ssim_lambda, float* __restrict__ out) {
    __shared__ double sa[256], sb[256];
    double a = 0.0, b = 0.0;
    for (int i = threadIdx.x; i < n; i += 256) { a += partials[i].x; b += partials[i].y; }
    sa[threadIdx.x] = a; sb[threadIdx.x] = b;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) { sa[threadIdx.x] += sa[threadIdx.x + s]; sb[threadIdx.x] += sb[threadIdx.x + s]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const float l1 = (float)(sa[0] / (3.0 * d.H * d.W));
        const float ssim = (float)(sb[0] / (3.0 * d.Hv * d.Wv));
        out[0] = l1; out[1] = ssim;
        out[2] = (1.f - ssim_lambda) * l1 + ssim_lambda * (1.f - ssim);
    }
}

struct LossBwdShared {
    float2 mab[kHaloH][kPitch];   // (d/d mu, d/d E[y^2]) maps, [row][column * 3 + channel], origin (i0 - 10, j0 - 10)
    float mf[kHaloH][kPitch];     // d/d E[xy] map
    float2 uab[kTileH][kPitch];   // after the pass along H
    float uf[kTileH][kPitch];
    float sg[kTileH][kOwnE + 1];  // gradient * background per (pixel, channel)
};
static_assert(sizeof(float2) * kHaloH * kPitch >= sizeof(float) * kTileH * kOwnE, "the gradient staging tile aliases mab");

// One block = 16x32 owned pixels; the window positions that cover them start up to 10 rows/columns earlier.
// Same tiling as the forward; the transpose of pass 2 then of pass 1 (the window is symmetric, so the adjoint is the
// same filter read from 10 positions earlier).
template <typename GT>
__global__ void __launch_bounds__(kLossThreads) loss_backward_kernel(LossDims d, LossWindow win, float ssim_lambda,
                                                            const float* __restrict__ v_main,
                                                            const float* __restrict__ out_img,
                                                            const float* __restrict__ out_texture,
                                                            const float* __restrict__ out_alpha,
                                                            const float* __restrict__ background,
                                                            const GT* __restrict__ gt, const float* __restrict__ mask,
                                                            const float* __restrict__ maps, float* __restrict__ v_out_img,
                                                            float* __restrict__ v_out_texture,
                                                            float* __restrict__ v_out_alpha) {
    extern __shared__ __align__(16) unsigned char loss_smem[];
    LossBwdShared& sh = *reinterpret_cast<LossBwdShared*>(loss_smem);
    const int i0 = blockIdx.y * kTileH, j0 = blockIdx.x * kTileW;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t plane = (size_t)d.H * d.W * 3;
    for (int r = warp; r < kHaloH; r += kLossThreads / 32) {
        const int gi = i0 - (kWin - 1) + r;
#pragma unroll
        for (int it = 0; it < (kHaloE + 31) / 32; ++it) {
            const int e = lane + 32 * it;
            if (e < kHaloE) {
                const int c = e / 3;
                const int gj = j0 - (kWin - 1) + c;
                float a = 0.f, b = 0.f, f = 0.f;
                if (gi >= 0 && gj >= 0 && gi < d.H && gj < d.W) {
                    const size_t idx = ((size_t)gi * d.W + gj) * 3 + (e - c * 3);
                    a = __ldg(maps + idx); b = __ldg(maps + plane + idx); f = __ldg(maps + 2 * plane + idx);
                }
                sh.mab[r][e] = make_float2(a, b);
                sh.mf[r][e] = f;
            }
        }
    }
    __syncthreads();
    // position q = p - s contributes to pixel p with weight g[s]; in the local frame that is row (p - i0) + (10 - s)
    if (threadIdx.x < 2 * kHaloE) {
        const int half = threadIdx.x / kHaloE, e = threadIdx.x - half * kHaloE, r0 = half * kOut1;
        float2 aab[kOut1];
        float af[kOut1];
#pragma unroll
        for (int o = 0; o < kOut1; ++o) { aab[o] = make_float2(0.f, 0.f); af[o] = 0.f; }
#pragma unroll
        for (int rin = 0; rin < kOut1 + kWin - 1; ++rin) {
            const float2 ab = sh.mab[r0 + rin][e];
            const float f = sh.mf[r0 + rin][e];
#pragma unroll
            for (int o = 0; o < kOut1; ++o) {
                const int q = rin - o;
                if (q >= 0 && q < kWin) {
                    aab[o] = __ffma2_rn(win.g2[kWin - 1 - q], ab, aab[o]);
                    af[o] = __fmaf_rn(win.g[kWin - 1 - q], f, af[o]);
                }
            }
        }
#pragma unroll
        for (int o = 0; o < kOut1; ++o) { sh.uab[r0 + o][e] = aab[o]; sh.uf[r0 + o][e] = af[o]; }
    }
    __syncthreads();
    // mab is dead from here on: its memory stages the gradient of the owned tile
    float* stage = reinterpret_cast<float*>(&sh.mab[0][0]);  // [kTileH][kOwnE]
    if (threadIdx.x < 3 * (kTileW / kOut2) * kTileH) {
        const int k = threadIdx.x % 3, cg = (threadIdx.x / 3) % (kTileW / kOut2), r = threadIdx.x / (3 * (kTileW / kOut2));
        float2 aab[kOut2];
        float af[kOut2];
#pragma unroll
        for (int o = 0; o < kOut2; ++o) { aab[o] = make_float2(0.f, 0.f); af[o] = 0.f; }
#pragma unroll
        for (int cin = 0; cin < kOut2 + kWin - 1; ++cin) {
            const int e = (kOut2 * cg + cin) * 3 + k;
            const float2 ab = sh.uab[r][e];
            const float f = sh.uf[r][e];
#pragma unroll
            for (int o = 0; o < kOut2; ++o) {
                const int q = cin - o;
                if (q >= 0 && q < kWin) {
                    aab[o] = __ffma2_rn(win.g2[kWin - 1 - q], ab, aab[o]);
                    af[o] = __fmaf_rn(win.g[kWin - 1 - q], f, af[o]);
                }
            }
        }
        const float up = v_main ? __ldg(v_main) : 1.f;
        const float w_l1 = __fdiv_rn(__fmul_rn(up, __fsub_rn(1.f, ssim_lambda)), __fmul_rn(__fmul_rn(3.f, (float)d.H), (float)d.W));
        const float w_ss = __fdiv_rn(__fmul_rn(-up, ssim_lambda), __fmul_rn(__fmul_rn(3.f, (float)d.Hv), (float)d.Wv));
        const float bgk = __ldg(background + k);
        const int gi = i0 + r;
#pragma unroll
        for (int o = 0; o < kOut2; ++o) {
            const int gj = j0 + kOut2 * cg + o;
            float g = 0.f, gb = 0.f;
            if (gi < d.H && gj < d.W) {
                const size_t pix = (size_t)gi * d.W + gj;
                const float om = __fsub_rn(1.f, __ldg(out_alpha + pix));
                const float pre = __fadd_rn(__fadd_rn(__ldg(out_img + pix * 3 + k), __ldg(out_texture + pix * d.C + k)),
                                            __fmul_rn(om, bgk));
                const float m = mask ? __ldg(mask + pix) : 1.f;
                const float x = __fmul_rn(load_gt(gt, pix * 3 + k), m), y = __fmul_rn(clamp_unit(pre), m);
                const float diff = __fsub_rn(x, y);  // Ll1 = mean |gt - pred|: d/dpred = -sign(gt - pred)
                const float sgn = diff > 0.f ? -1.f : (diff < 0.f ? 1.f : 0.f);
                // explicit roundings, left to right: see the forward's SSIM algebra
                const float inner = __fadd_rn(__fadd_rn(aab[o].x, __fmul_rn(__fmul_rn(2.f, y), aab[o].y)), __fmul_rn(x, af[o]));
                g = __fadd_rn(__fmul_rn(w_l1, sgn), __fmul_rn(w_ss, inner));
                g = __fmul_rn(g, m);                       // pred_img * mask
                if (!(pre >= 0.f && pre <= 1.f)) g = 0.f;  // torch.clamp passes the gradient on [min, max] only
                gb = __fmul_rn(g, bgk);  // d/d out_alpha of (1 - out_alpha) * background, summed over the channels below
            }
            const int se = (kOut2 * cg + o) * 3 + k;
            stage[r * kOwnE + se] = g;
            sh.sg[r][se] = gb;
        }
    }
    __syncthreads();
    store_tile_rows(v_out_img, stage, i0, j0, d);
    if (d.C == 3) {
        store_tile_rows(v_out_texture, stage, i0, j0, d);
    } else {
        for (int t = threadIdx.x; t < kTileH * kTileW; t += kLossThreads) {
            const int pr = t / kTileW, pc = t - pr * kTileW;
            if (i0 + pr < d.H && j0 + pc < d.W) {
                float* o = v_out_texture + ((size_t)(i0 + pr) * d.W + j0 + pc) * d.C;
                for (int k = 0; k < 3; ++k) o[k] = stage[pr * kOwnE + pc * 3 + k];
                for (int k = 3; k < d.C; ++k) o[k] = 0.f;  // channels the composite does not read
            }
        }
    }
    for (int t = threadIdx.x; t < kTileH * kTileW; t += kLossThreads) {
        const int pr = t / kTileW, pc = t - pr * kTileW;
        if (i0 + pr < d.H && j0 + pc < d.W)
            v_out_alpha[(size_t)(i0 + pr) * d.W + j0 + pc] = ((0.f - sh.sg[pr][pc * 3]) - sh.sg[pr][pc * 3 + 1]) - sh.sg[pr][pc * 3 + 2];
    }
}

static LossWindow make_window(float sigma) {
    // _fspecial_gauss_1d(11, sigma): exp(-(x - 5)^2 / (2 sigma^2)) in fp32, normalised by the fp32 sum
    LossWindow w;
    float s = 0.f;
    for (int i = 0; i < kWin; ++i) {
        const float x = (float)(i - kWin / 2);
        w.g[i] = expf(-(x * x) / (2.f * sigma * sigma));
        s += w.g[i];
    }
    for (int i = 0; i < kWin; ++i) {
        w.g[i] /= s;
        w.g2[i] = make_float2(w.g[i], w.g[i]);
    }
    return w;
}

}  // namespace gstex

using namespace gstex;

extern "C" size_t gstex_photometric_loss_scratch_bytes(int img_height, int img_width) {
    if (img_height < 1 || img_width < 1) return 0;
    return (size_t)div_up(img_height, kTileH) * div_up(img_width, kTileW) * sizeof(float2);
}

template <typename GT>
static int loss_forward_impl(int img_height, int img_width, int channels, const float* out_img, const float* out_texture,
                             const float* out_alpha, const float* background, const GT* gt_img, const float* mask,
                             float ssim_lambda, int win_size, float win_sigma, float* rgb, float* losses, float* ssim_maps,
                             void* scratch, size_t scratch_bytes, void* stream) {
    if (img_height < kWin || img_width < kWin || channels < 3 || win_size != kWin || !(win_sigma > 0.f)) return GSTEX_E_BADARG;
    if (!out_img || !out_texture || !out_alpha || !background || !gt_img || !rgb || !losses || !scratch) return GSTEX_E_BADARG;
    if (scratch_bytes < gstex_photometric_loss_scratch_bytes(img_height, img_width)) return GSTEX_E_WORKSPACE;
    const LossDims d{img_height, img_width, channels, img_height - (kWin - 1), img_width - (kWin - 1)};
    const dim3 grid(div_up(img_width, kTileW), div_up(img_height, kTileH));
    cudaStream_t st = (cudaStream_t)stream;
    // > 48 KB of shared memory: opt in (per device, cheap: done at every launch)
    GSTEX_CUDA_TRY(cudaFuncSetAttribute(loss_forward_kernel<GT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)sizeof(LossFwdShared)));
    loss_forward_kernel<GT><<<grid, kLossThreads, sizeof(LossFwdShared), st>>>(d, make_window(win_sigma), out_img, out_texture,
                                                                              out_alpha, background, gt_img, mask, rgb, ssim_maps,
                                                                              reinterpret_cast<float2*>(scratch));
    GSTEX_LAUNCH_CHECK();
    loss_finalize_kernel<<<1, 256, 0, st>>>(reinterpret_cast<const float2*>(scratch), (int)(grid.x * grid.y), d, ssim_lambda,
                                            losses);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

template <typename GT>
static int loss_backward_impl(int img_height, int img_width, int channels, const float* out_img, const float* out_texture,
                              const float* out_alpha, const float* background, const GT* gt_img, const float* mask,
                              float ssim_lambda, int win_size, float win_sigma, const float* ssim_maps,
                              const float* v_main_loss, float* v_out_img, float* v_out_texture, float* v_out_alpha,
                              void* stream) {
    if (img_height < kWin || img_width < kWin || channels < 3 || win_size != kWin || !(win_sigma > 0.f)) return GSTEX_E_BADARG;
    if (!out_img || !out_texture || !out_alpha || !background || !gt_img || !ssim_maps || !v_out_img || !v_out_texture ||
        !v_out_alpha)
        return GSTEX_E_BADARG;
    const LossDims d{img_height, img_width, channels, img_height - (kWin - 1), img_width - (kWin - 1)};
    const dim3 grid(div_up(img_width, kTileW), div_up(img_height, kTileH));
    GSTEX_CUDA_TRY(cudaFuncSetAttribute(loss_backward_kernel<GT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)sizeof(LossBwdShared)));
    loss_backward_kernel<GT><<<grid, kLossThreads, sizeof(LossBwdShared), (cudaStream_t)stream>>>(
        d, make_window(win_sigma), ssim_lambda, v_main_loss, out_img, out_texture, out_alpha, background, gt_img, mask, ssim_maps,
        v_out_img, v_out_texture, v_out_alpha);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

#define GSTEX_LOSS_FWD_ARGS                                                                                              \
    img_height, img_width, channels, out_img, out_texture, out_alpha, background, gt_img, mask, ssim_lambda, win_size,   \
        win_sigma, rgb, losses, ssim_maps, scratch, scratch_bytes, stream
#define GSTEX_LOSS_BWD_ARGS                                                                                              \
    img_height, img_width, channels, out_img, out_texture, out_alpha, background, gt_img, mask, ssim_lambda, win_size,   \
        win_sigma, ssim_maps, v_main_loss, v_out_img, v_out_texture, v_out_alpha, stream

extern "C" int gstex_photometric_loss_forward(int img_height, int img_width, int channels, const float* out_img,
                                              const float* out_texture, const float* out_alpha, const float* background,
                                              const float* gt_img, const float* mask, float ssim_lambda, int win_size,
                                              float win_sigma, float* rgb, float* losses, float* ssim_maps, void* scratch,
                                              size_t scratch_bytes, void* stream) {
    return loss_forward_impl<float>(GSTEX_LOSS_FWD_ARGS);
}

extern "C" int gstex_photometric_loss_backward(int img_height, int img_width, int channels, const float* out_img,
                                               const float* out_texture, const float* out_alpha, const float* background,
                                               const float* gt_img, const float* mask, float ssim_lambda, int win_size,
                                               float win_sigma, const float* ssim_maps, const float* v_main_loss,
                                               float* v_out_img, float* v_out_texture, float* v_out_alpha, void* stream) {
    return loss_backward_impl<float>(GSTEX_LOSS_BWD_ARGS);
}

// The same two calls with the ground truth as the uint8 image the data manager holds (value / 255, see load_gt).
extern "C" int gstex_photometric_loss_forward_u8(int img_height, int img_width, int channels, const float* out_img,
                                                 const float* out_texture, const float* out_alpha, const float* background,
                                                 const uint8_t* gt_img, const float* mask, float ssim_lambda, int win_size,
                                                 float win_sigma, float* rgb, float* losses, float* ssim_maps, void* scratch,
                                                 size_t scratch_bytes, void* stream) {
    return loss_forward_impl<uint8_t>(GSTEX_LOSS_FWD_ARGS);
}

extern "C" int gstex_photometric_loss_backward_u8(int img_height, int img_width, int channels, const float* out_img,
                                                  const float* out_texture, const float* out_alpha, const float* background,
                                                  const uint8_t* gt_img, const float* mask, float ssim_lambda, int win_size,
                                                  float win_sigma, const float* ssim_maps, const float* v_main_loss,
                                                  float* v_out_img, float* v_out_texture, float* v_out_alpha, void* stream) {
    return loss_backward_impl<uint8_t>(GSTEX_LOSS_BWD_ARGS);
}

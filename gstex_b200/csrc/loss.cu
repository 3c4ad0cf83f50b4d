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
// By contract an HBM stream: forward reads 4*(10 + C) B and writes 48 B per pixel, backward reads 4*(19 + C) B and writes
// 4*(4 + C) B.  sm_100a.
#include "common.cuh"

namespace gstex {

constexpr int kWin = 11;              // pytorch_msssim default win_size
constexpr int kLossTile = 16;         // owned pixels per block edge
constexpr int kHalo = kLossTile + kWin - 1;  // 26
constexpr int kHaloE = kHalo * 3;     // floats per halo row (three interleaved channels)
constexpr int kRowPitch = 80;         // shared-memory row pitch: 16 banks from row to row

struct LossWindow { float g[kWin]; };

struct LossDims {
    int H, W, C;       // image size, channels of out_texture (>= 3; only the first three are composited)
    int Hv, Wv;        // valid SSIM positions: H - 10, W - 10
};

__device__ __forceinline__ float clamp_unit(float v) { return fminf(fmaxf(v, 0.f), 1.f); }

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

// One block = 16x16 owned pixels (the SSIM positions whose window STARTS there) + the 26x26 pixels their windows cover.
// maps: [3][H][W][3] = d ssim / d mu_pred, d ssim / d E[pred^2], d ssim / d E[gt*pred] per position and channel,
// zero where the position is not a valid window start (written only when maps != nullptr).
// Shared-memory traffic is what bounds a naive version (one output per thread = 2 loads per 5 FMAs; ncu: L1/shared
// pipe 87 % busy, mio_throttle the top stall), so both passes are register-tiled: a pass-1 thread walks 18 rows of one
// (column, channel) and produces 8 output rows of all five moments (36 loads for 440 FMAs), a pass-2 thread walks 14
// columns and produces 4 outputs (70 loads for 220 FMAs).  Pixels stay channel-interleaved in shared memory, as in
// HBM, so the global loads/stores are contiguous row segments and consecutive threads hit consecutive banks.
__global__ void __launch_bounds__(256) loss_forward_kernel(LossDims d, LossWindow win, const float* __restrict__ out_img,
                                                           const float* __restrict__ out_texture,
                                                           const float* __restrict__ out_alpha,
                                                           const float* __restrict__ background,
                                                           const float* __restrict__ gt, const float* __restrict__ mask,
                                                           float* __restrict__ rgb, float* __restrict__ maps,
                                                           float2* __restrict__ partials) {
    __shared__ float sx[kHalo][kRowPitch];             // gt (masked), [row][column * 3 + channel]
    __shared__ float sy[kHalo][kRowPitch];             // predicted (masked)
    __shared__ float st[5][kLossTile][kRowPitch];      // the five moments after pass 1
    __shared__ float red[8];
    const int i0 = blockIdx.y * kLossTile, j0 = blockIdx.x * kLossTile;
    float l1 = 0.f;
    for (int t = threadIdx.x; t < kHalo * kHaloE; t += 256) {
        const int r = t / kHaloE, e = t - r * kHaloE, c = e / 3, k = e - c * 3;
        const int gi = i0 + r, gj = j0 + c;
        float x = 0.f, y = 0.f;
        if (gi < d.H && gj < d.W) {
            const size_t pix = (size_t)gi * d.W + gj;
            // gstex.py:1204-1205 as torch evaluates it (no contraction: every product and sum rounded on its own)
            const float om = __fsub_rn(1.f, __ldg(out_alpha + pix));
            const float pre = __fadd_rn(__fadd_rn(__ldg(out_img + pix * 3 + k), __ldg(out_texture + pix * d.C + k)),
                                        __fmul_rn(om, __ldg(background + k)));
            const float v = clamp_unit(pre);
            const float m = mask ? __ldg(mask + pix) : 1.f;
            x = __fmul_rn(__ldg(gt + pix * 3 + k), m);  // gstex.py:1297-1298 (gt_img * mask, pred_img * mask)
            y = __fmul_rn(v, m);
            if (r < kLossTile && c < kLossTile) {
                rgb[pix * 3 + k] = v;
                l1 += fabsf(x - y);
            }
        }
        sx[r][e] = x; sy[r][e] = y;
    }
    __syncthreads();
    // pass 1, along H (pytorch_msssim filters dim 2 first)
    if (threadIdx.x < 2 * kHaloE) {
        const int half = threadIdx.x / kHaloE, e = threadIdx.x - half * kHaloE, r0 = half * 8;
        float acc[8][5];
#pragma unroll
        for (int o = 0; o < 8; ++o)
#pragma unroll
            for (int m = 0; m < 5; ++m) acc[o][m] = 0.f;
#pragma unroll
        for (int rin = 0; rin < 8 + kWin - 1; ++rin) {
            const float x = sx[r0 + rin][e], y = sy[r0 + rin][e];
            const float xx = x * x, yy = y * y, xy = x * y;
#pragma unroll
            for (int o = 0; o < 8; ++o) {
                const int q = rin - o;
                if (q >= 0 && q < kWin) {
                    const float g = win.g[q];
                    acc[o][0] = fmaf(g, x, acc[o][0]); acc[o][1] = fmaf(g, y, acc[o][1]);
                    acc[o][2] = fmaf(g, xx, acc[o][2]); acc[o][3] = fmaf(g, yy, acc[o][3]);
                    acc[o][4] = fmaf(g, xy, acc[o][4]);
                }
            }
        }
#pragma unroll
        for (int o = 0; o < 8; ++o)
#pragma unroll
            for (int m = 0; m < 5; ++m) st[m][r0 + o][e] = acc[o][m];
    }
    __syncthreads();
    // pass 2, along W, then the SSIM algebra (_ssim of pytorch_msssim/ssim.py); 4 positions of one channel per thread
    float ssim = 0.f;
    if (threadIdx.x < 3 * 4 * kLossTile) {
        const int k = threadIdx.x % 3, cg = (threadIdx.x / 3) & 3, r = threadIdx.x / 12;
        float acc[4][5];
#pragma unroll
        for (int o = 0; o < 4; ++o)
#pragma unroll
            for (int m = 0; m < 5; ++m) acc[o][m] = 0.f;
#pragma unroll
        for (int cin = 0; cin < 4 + kWin - 1; ++cin) {
            float v[5];
#pragma unroll
            for (int m = 0; m < 5; ++m) v[m] = st[m][r][(4 * cg + cin) * 3 + k];
#pragma unroll
            for (int o = 0; o < 4; ++o) {
                const int q = cin - o;
                if (q >= 0 && q < kWin) {
                    const float g = win.g[q];
#pragma unroll
                    for (int m = 0; m < 5; ++m) acc[o][m] = fmaf(g, v[m], acc[o][m]);
                }
            }
        }
        const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;  // (K1 * data_range)^2, (K2 * data_range)^2
        const int gi = i0 + r;
        const size_t plane = (size_t)d.H * d.W * 3;
#pragma unroll
        for (int o = 0; o < 4; ++o) {
            const int gj = j0 + 4 * cg + o;
            const bool valid = gi < d.Hv && gj < d.Wv;
            const float mu1 = acc[o][0], mu2 = acc[o][1];
            const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
            const float s1 = acc[o][2] - mu1_sq, s2 = acc[o][3] - mu2_sq, s12 = acc[o][4] - mu12;
            const float A1 = 2.f * mu12 + C1, A2 = 2.f * s12 + C2, B1 = mu1_sq + mu2_sq + C1, B2 = s1 + s2 + C2;
            const float rB1 = 1.f / B1, rB2 = 1.f / B2;
            const float cs = A2 * rB2, lum = A1 * rB1;
            if (valid) ssim += lum * cs;
            if (maps != nullptr && gi < d.H && gj < d.W) {
                // partial derivatives of lum*cs with E[y], E[y^2], E[xy] as the independent moments of the prediction y
                float dmu = 0.f, dyy = 0.f, dxy = 0.f;
                if (valid) {
                    dyy = -lum * cs * rB2;
                    dxy = 2.f * lum * rB2;
                    dmu = 2.f * rB1 * (mu1 * cs - mu2 * lum * cs) - 2.f * mu2 * dyy - mu1 * dxy;
                }
                const size_t idx = ((size_t)gi * d.W + gj) * 3 + k;
                maps[idx] = dmu; maps[plane + idx] = dyy; maps[2 * plane + idx] = dxy;
            }
        }
    }
    const float l1_sum = block_sum_256(l1, red);
    const float ssim_sum = block_sum_256(ssim, red);
    if (threadIdx.x == 0) partials[blockIdx.y * gridDim.x + blockIdx.x] = make_float2(l1_sum, ssim_sum);
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

// One block = 16x16 owned pixels; the window positions that cover them start up to 10 rows/columns earlier.
// Same register tiling as the forward; the transpose of pass 2 then of pass 1 (the window is symmetric, so the
// adjoint is the same filter read from 10 positions earlier).
__global__ void __launch_bounds__(256) loss_backward_kernel(LossDims d, LossWindow win, float ssim_lambda,
                                                            const float* __restrict__ v_main,
                                                            const float* __restrict__ out_img,
                                                            const float* __restrict__ out_texture,
                                                            const float* __restrict__ out_alpha,
                                                            const float* __restrict__ background,
                                                            const float* __restrict__ gt, const float* __restrict__ mask,
                                                            const float* __restrict__ maps, float* __restrict__ v_out_img,
                                                            float* __restrict__ v_out_texture,
                                                            float* __restrict__ v_out_alpha) {
    __shared__ float sm[3][kHalo][kRowPitch];      // [map][row][column * 3 + channel], origin (i0 - 10, j0 - 10)
    __shared__ float su[3][kLossTile][kRowPitch];  // after the column pass
    __shared__ float sg[kLossTile][kLossTile * 3 + 1];  // gradient * background per (pixel, channel)
    const int i0 = blockIdx.y * kLossTile, j0 = blockIdx.x * kLossTile;
    const size_t plane = (size_t)d.H * d.W * 3;
    for (int t = threadIdx.x; t < kHalo * kHaloE; t += 256) {
        const int r = t / kHaloE, e = t - r * kHaloE, c = e / 3;
        const int gi = i0 - (kWin - 1) + r, gj = j0 - (kWin - 1) + c;
        float a = 0.f, b = 0.f, f = 0.f;
        if (gi >= 0 && gj >= 0 && gi < d.H && gj < d.W) {
            const size_t idx = ((size_t)gi * d.W + gj) * 3 + (e - c * 3);
            a = __ldg(maps + idx); b = __ldg(maps + plane + idx); f = __ldg(maps + 2 * plane + idx);
        }
        sm[0][r][e] = a; sm[1][r][e] = b; sm[2][r][e] = f;
    }
    __syncthreads();
    // position q = p - s contributes to pixel p with weight g[s]; in the local frame that is row (p - i0) + (10 - s)
    if (threadIdx.x < 2 * kHaloE) {
        const int half = threadIdx.x / kHaloE, e = threadIdx.x - half * kHaloE, r0 = half * 8;
        float acc[8][3];
#pragma unroll
        for (int o = 0; o < 8; ++o) acc[o][0] = acc[o][1] = acc[o][2] = 0.f;
#pragma unroll
        for (int rin = 0; rin < 8 + kWin - 1; ++rin) {
            const float a = sm[0][r0 + rin][e], b = sm[1][r0 + rin][e], f = sm[2][r0 + rin][e];
#pragma unroll
            for (int o = 0; o < 8; ++o) {
                const int q = rin - o;
                if (q >= 0 && q < kWin) {
                    const float g = win.g[kWin - 1 - q];
                    acc[o][0] = fmaf(g, a, acc[o][0]); acc[o][1] = fmaf(g, b, acc[o][1]); acc[o][2] = fmaf(g, f, acc[o][2]);
                }
            }
        }
#pragma unroll
        for (int o = 0; o < 8; ++o) { su[0][r0 + o][e] = acc[o][0]; su[1][r0 + o][e] = acc[o][1]; su[2][r0 + o][e] = acc[o][2]; }
    }
    __syncthreads();
    if (threadIdx.x < 3 * 4 * kLossTile) {
        const int k = threadIdx.x % 3, cg = (threadIdx.x / 3) & 3, r = threadIdx.x / 12;
        float acc[4][3];
#pragma unroll
        for (int o = 0; o < 4; ++o) acc[o][0] = acc[o][1] = acc[o][2] = 0.f;
#pragma unroll
        for (int cin = 0; cin < 4 + kWin - 1; ++cin) {
            const int e = (4 * cg + cin) * 3 + k;
            const float a = su[0][r][e], b = su[1][r][e], f = su[2][r][e];
#pragma unroll
            for (int o = 0; o < 4; ++o) {
                const int q = cin - o;
                if (q >= 0 && q < kWin) {
                    const float g = win.g[kWin - 1 - q];
                    acc[o][0] = fmaf(g, a, acc[o][0]); acc[o][1] = fmaf(g, b, acc[o][1]); acc[o][2] = fmaf(g, f, acc[o][2]);
                }
            }
        }
        const float up = v_main ? __ldg(v_main) : 1.f;
        const float w_l1 = up * (1.f - ssim_lambda) / (3.f * (float)d.H * (float)d.W);
        const float w_ss = -up * ssim_lambda / (3.f * (float)d.Hv * (float)d.Wv);
        const float bgk = __ldg(background + k);
        const int gi = i0 + r;
#pragma unroll
        for (int o = 0; o < 4; ++o) {
            const int gj = j0 + 4 * cg + o;
            float gb = 0.f;
            if (gi < d.H && gj < d.W) {
                const size_t pix = (size_t)gi * d.W + gj;
                const float om = __fsub_rn(1.f, __ldg(out_alpha + pix));
                const float pre = __fadd_rn(__fadd_rn(__ldg(out_img + pix * 3 + k), __ldg(out_texture + pix * d.C + k)),
                                            __fmul_rn(om, bgk));
                const float m = mask ? __ldg(mask + pix) : 1.f;
                const float x = __fmul_rn(__ldg(gt + pix * 3 + k), m), y = __fmul_rn(clamp_unit(pre), m);
                const float diff = x - y;  // Ll1 = mean |gt - pred|: d/dpred = -sign(gt - pred)
                const float sgn = diff > 0.f ? -1.f : (diff < 0.f ? 1.f : 0.f);
                float g = w_l1 * sgn + w_ss * (acc[o][0] + 2.f * y * acc[o][1] + x * acc[o][2]);
                g *= m;                                    // pred_img * mask
                if (!(pre >= 0.f && pre <= 1.f)) g = 0.f;  // torch.clamp passes the gradient on [min, max] only
                v_out_img[pix * 3 + k] = g;
                v_out_texture[pix * d.C + k] = g;
                gb = g * bgk;  // d/d out_alpha of (1 - out_alpha) * background, summed over the channels below
            }
            sg[r][(4 * cg + o) * 3 + k] = gb;
        }
    }
    __syncthreads();
    const int pr = threadIdx.x >> 4, pc = threadIdx.x & 15;
    const int gi = i0 + pr, gj = j0 + pc;
    if (gi < d.H && gj < d.W) {
        const size_t pix = (size_t)gi * d.W + gj;
        for (int k = 3; k < d.C; ++k) v_out_texture[pix * d.C + k] = 0.f;  // channels the composite does not read
        v_out_alpha[pix] = ((0.f - sg[pr][pc * 3]) - sg[pr][pc * 3 + 1]) - sg[pr][pc * 3 + 2];
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
    for (int i = 0; i < kWin; ++i) w.g[i] /= s;
    return w;
}

}  // namespace gstex

using namespace gstex;

extern "C" size_t gstex_photometric_loss_scratch_bytes(int img_height, int img_width) {
    if (img_height < 1 || img_width < 1) return 0;
    return (size_t)div_up(img_height, kLossTile) * div_up(img_width, kLossTile) * sizeof(float2);
}

extern "C" int gstex_photometric_loss_forward(int img_height, int img_width, int channels, const float* out_img,
                                              const float* out_texture, const float* out_alpha, const float* background,
                                              const float* gt_img, const float* mask, float ssim_lambda, int win_size,
                                              float win_sigma, float* rgb, float* losses, float* ssim_maps, void* scratch,
                                              size_t scratch_bytes, void* stream) {
    if (img_height < kWin || img_width < kWin || channels < 3 || win_size != kWin || !(win_sigma > 0.f)) return GSTEX_E_BADARG;
    if (!out_img || !out_texture || !out_alpha || !background || !gt_img || !rgb || !losses || !scratch) return GSTEX_E_BADARG;
    if (scratch_bytes < gstex_photometric_loss_scratch_bytes(img_height, img_width)) return GSTEX_E_WORKSPACE;
    const LossDims d{img_height, img_width, channels, img_height - (kWin - 1), img_width - (kWin - 1)};
    const dim3 grid(div_up(img_width, kLossTile), div_up(img_height, kLossTile));
    cudaStream_t st = (cudaStream_t)stream;
    loss_forward_kernel<<<grid, 256, 0, st>>>(d, make_window(win_sigma), out_img, out_texture, out_alpha, background, gt_img,
                                              mask, rgb, ssim_maps, reinterpret_cast<float2*>(scratch));
    GSTEX_LAUNCH_CHECK();
    loss_finalize_kernel<<<1, 256, 0, st>>>(reinterpret_cast<const float2*>(scratch), (int)(grid.x * grid.y), d, ssim_lambda,
                                            losses);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

extern "C" int gstex_photometric_loss_backward(int img_height, int img_width, int channels, const float* out_img,
                                               const float* out_texture, const float* out_alpha, const float* background,
                                               const float* gt_img, const float* mask, float ssim_lambda, int win_size,
                                               float win_sigma, const float* ssim_maps, const float* v_main_loss,
                                               float* v_out_img, float* v_out_texture, float* v_out_alpha, void* stream) {
    if (img_height < kWin || img_width < kWin || channels < 3 || win_size != kWin || !(win_sigma > 0.f)) return GSTEX_E_BADARG;
    if (!out_img || !out_texture || !out_alpha || !background || !gt_img || !ssim_maps || !v_out_img || !v_out_texture ||
        !v_out_alpha)
        return GSTEX_E_BADARG;
    const LossDims d{img_height, img_width, channels, img_height - (kWin - 1), img_width - (kWin - 1)};
    const dim3 grid(div_up(img_width, kLossTile), div_up(img_height, kLossTile));
    loss_backward_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(d, make_window(win_sigma), ssim_lambda, v_main_loss, out_img,
                                                                 out_texture, out_alpha, background, gt_img, mask, ssim_maps,
                                                                 v_out_img, v_out_texture, v_out_alpha);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

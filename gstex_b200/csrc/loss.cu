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
// HBM-bound by contract: forward reads 4*(10 + C) B and writes 48 B per pixel, backward reads 4*(19 + C) B and writes
// 4*(4 + C) B.  sm_100a.
#include "common.cuh"

namespace gstex {

constexpr int kWin = 11;              // pytorch_msssim default win_size
constexpr int kLossTile = 16;         // owned pixels per block edge
constexpr int kHalo = kLossTile + kWin - 1;  // 26
constexpr int kPitch = kHalo + 1;     // shared-memory row pitch (odd: column walks do not conflict)

struct LossWindow { float g[kWin]; };

struct LossDims {
    int H, W, C;       // image size, channels of out_texture (>= 3; only the first three are composited)
    int Hv, Wv;        // valid SSIM positions: H - 10, W - 10
};

// compose one pixel's three channels exactly as torch evaluates gstex.py:1204-1205 (no contraction: every product
// and sum rounded on its own), returns the pre-clamp values
__device__ __forceinline__ void compose_pixel(const float* __restrict__ out_img, const float* __restrict__ out_texture,
                                              const float* __restrict__ out_alpha, const float bg[3], int C, size_t pix,
                                              float pre[3]) {
    const float om = __fsub_rn(1.f, out_alpha[pix]);
#pragma unroll
    for (int k = 0; k < 3; ++k)
        pre[k] = __fadd_rn(__fadd_rn(out_img[pix * 3 + k], out_texture[pix * C + k]), __fmul_rn(om, bg[k]));
}

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
__global__ void __launch_bounds__(256) loss_forward_kernel(LossDims d, LossWindow win, const float* __restrict__ out_img,
                                                           const float* __restrict__ out_texture,
                                                           const float* __restrict__ out_alpha,
                                                           const float* __restrict__ background,
                                                           const float* __restrict__ gt, const float* __restrict__ mask,
                                                           float* __restrict__ rgb, float* __restrict__ maps,
                                                           float2* __restrict__ partials) {
    __shared__ float sx[3][kHalo][kPitch];  // gt (masked)
    __shared__ float sy[3][kHalo][kPitch];  // predicted (masked)
    __shared__ float st[5][kLossTile][kPitch];
    __shared__ float red[8];
    const int i0 = blockIdx.y * kLossTile, j0 = blockIdx.x * kLossTile;
    const float bg[3] = {background[0], background[1], background[2]};
    float l1 = 0.f;
    for (int t = threadIdx.x; t < kHalo * kHalo; t += 256) {
        const int r = t / kHalo, c = t - r * kHalo;
        const int gi = i0 + r, gj = j0 + c;
        float x[3] = {0.f, 0.f, 0.f}, y[3] = {0.f, 0.f, 0.f};
        if (gi < d.H && gj < d.W) {
            const size_t pix = (size_t)gi * d.W + gj;
            float pre[3];
            compose_pixel(out_img, out_texture, out_alpha, bg, d.C, pix, pre);
            const float m = mask ? mask[pix] : 1.f;
            const bool owned = r < kLossTile && c < kLossTile;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const float v = clamp_unit(pre[k]);
                if (owned) rgb[pix * 3 + k] = v;
                x[k] = __fmul_rn(gt[pix * 3 + k], m);  // gstex.py:1297-1298 (gt_img * mask, pred_img * mask)
                y[k] = __fmul_rn(v, m);
                if (owned) l1 += fabsf(x[k] - y[k]);
            }
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) { sx[k][r][c] = x[k]; sy[k][r][c] = y[k]; }
    }
    __syncthreads();
    const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;  // (K1 * data_range)^2, (K2 * data_range)^2
    const int pr = threadIdx.x >> 4, pc = threadIdx.x & 15;  // this thread's owned position
    const int gi = i0 + pr, gj = j0 + pc;
    const bool valid = gi < d.Hv && gj < d.Wv;
    float ssim = 0.f;
    for (int k = 0; k < 3; ++k) {
        // pass 1, along H (pytorch_msssim filters dim 2 first): rows 0..15, columns 0..25
        for (int t = threadIdx.x; t < kLossTile * kHalo; t += 256) {
            const int r = t / kHalo, c = t - r * kHalo;
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, a4 = 0.f;
#pragma unroll
            for (int q = 0; q < kWin; ++q) {
                const float x = sx[k][r + q][c], y = sy[k][r + q][c], g = win.g[q];
                a0 = fmaf(g, x, a0); a1 = fmaf(g, y, a1);
                a2 = fmaf(g, x * x, a2); a3 = fmaf(g, y * y, a3); a4 = fmaf(g, x * y, a4);
            }
            st[0][r][c] = a0; st[1][r][c] = a1; st[2][r][c] = a2; st[3][r][c] = a3; st[4][r][c] = a4;
        }
        __syncthreads();
        // pass 2, along W, then the SSIM algebra (_ssim of pytorch_msssim/ssim.py)
        float mu1 = 0.f, mu2 = 0.f, exx = 0.f, eyy = 0.f, exy = 0.f;
#pragma unroll
        for (int q = 0; q < kWin; ++q) {
            const float g = win.g[q];
            mu1 = fmaf(g, st[0][pr][pc + q], mu1); mu2 = fmaf(g, st[1][pr][pc + q], mu2);
            exx = fmaf(g, st[2][pr][pc + q], exx); eyy = fmaf(g, st[3][pr][pc + q], eyy);
            exy = fmaf(g, st[4][pr][pc + q], exy);
        }
        const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
        const float s1 = exx - mu1_sq, s2 = eyy - mu2_sq, s12 = exy - mu12;
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
            const size_t plane = (size_t)d.H * d.W * 3, o = ((size_t)gi * d.W + gj) * 3 + k;
            maps[o] = dmu; maps[plane + o] = dyy; maps[2 * plane + o] = dxy;
        }
        __syncthreads();
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
    __shared__ float sm[3][3][kHalo][kPitch];  // [map][channel][row][col], rows/cols start at (i0 - 10, j0 - 10)
    __shared__ float su[3][kLossTile][kPitch];
    const int i0 = blockIdx.y * kLossTile, j0 = blockIdx.x * kLossTile;
    const size_t plane = (size_t)d.H * d.W * 3;
    for (int t = threadIdx.x; t < kHalo * kHalo * 3; t += 256) {
        const int r = t / (kHalo * 3), rem = t - r * (kHalo * 3), c = rem / 3, k = rem - c * 3;
        const int gi = i0 - (kWin - 1) + r, gj = j0 - (kWin - 1) + c;
        float a = 0.f, b = 0.f, e = 0.f;
        if (gi >= 0 && gj >= 0 && gi < d.H && gj < d.W) {
            const size_t o = ((size_t)gi * d.W + gj) * 3 + k;
            a = __ldg(maps + o); b = __ldg(maps + plane + o); e = __ldg(maps + 2 * plane + o);
        }
        sm[0][k][r][c] = a; sm[1][k][r][c] = b; sm[2][k][r][c] = e;
    }
    __syncthreads();
    const int pr = threadIdx.x >> 4, pc = threadIdx.x & 15;
    const int gi = i0 + pr, gj = j0 + pc;
    const bool inside = gi < d.H && gj < d.W;
    const size_t pix = inside ? (size_t)gi * d.W + gj : 0;
    const float up = v_main ? __ldg(v_main) : 1.f;
    const float w_l1 = up * (1.f - ssim_lambda) / (3.f * (float)d.H * (float)d.W);
    const float w_ss = -up * ssim_lambda / (3.f * (float)d.Hv * (float)d.Wv);
    const float bg[3] = {background[0], background[1], background[2]};
    float pre[3] = {0.f, 0.f, 0.f};
    float m = 1.f;
    if (inside) {
        compose_pixel(out_img, out_texture, out_alpha, bg, d.C, pix, pre);
        if (mask) m = mask[pix];
    }
    float v_alpha = 0.f;
    for (int k = 0; k < 3; ++k) {
        // transpose of pass 2 then of pass 1: position q = p - s contributes with weight g[s]; in the local frame
        // (origin 10 rows/columns before the tile) that is rows pr + (10 - s), and the window is symmetric
        for (int t = threadIdx.x; t < kLossTile * kHalo; t += 256) {
            const int r = t / kHalo, c = t - r * kHalo;
            float a0 = 0.f, a1 = 0.f, a2 = 0.f;
#pragma unroll
            for (int q = 0; q < kWin; ++q) {
                const float g = win.g[kWin - 1 - q];
                a0 = fmaf(g, sm[0][k][r + q][c], a0); a1 = fmaf(g, sm[1][k][r + q][c], a1);
                a2 = fmaf(g, sm[2][k][r + q][c], a2);
            }
            su[0][r][c] = a0; su[1][r][c] = a1; su[2][r][c] = a2;
        }
        __syncthreads();
        float d0 = 0.f, d1 = 0.f, d2 = 0.f;
#pragma unroll
        for (int q = 0; q < kWin; ++q) {
            const float g = win.g[kWin - 1 - q];
            d0 = fmaf(g, su[0][pr][pc + q], d0); d1 = fmaf(g, su[1][pr][pc + q], d1); d2 = fmaf(g, su[2][pr][pc + q], d2);
        }
        __syncthreads();
        if (inside) {
            const float v = clamp_unit(pre[k]);
            const float x = __fmul_rn(gt[pix * 3 + k], m), y = __fmul_rn(v, m);
            const float diff = x - y;  // Ll1 = mean |gt - pred|: d/dpred = -sign(gt - pred)
            const float sgn = diff > 0.f ? -1.f : (diff < 0.f ? 1.f : 0.f);
            float g = w_l1 * sgn + w_ss * (d0 + 2.f * y * d1 + x * d2);
            g *= m;                                          // pred_img * mask
            if (!(pre[k] >= 0.f && pre[k] <= 1.f)) g = 0.f;  // torch.clamp passes the gradient on [min, max] only
            v_out_img[pix * 3 + k] = g;
            v_out_texture[pix * d.C + k] = g;
            v_alpha -= g * bg[k];  // d/d out_alpha of (1 - out_alpha) * background
        }
    }
    if (inside) {
        for (int k = 3; k < d.C; ++k) v_out_texture[pix * d.C + k] = 0.f;  // channels the composite does not read
        v_out_alpha[pix] = v_alpha;
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

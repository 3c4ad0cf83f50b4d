/* gstex_oracle_loss.c -- CPU restatement of the reference trainer's compositing + photometric loss.
 * TEST INFRASTRUCTURE, NOT PRODUCT CODE (see oracle/__init__.py).
 *
 * Follows nerfstudio/models/gstex.py:1204-1205 (rgb = clamp(out_img + out_texture[:,:,0:3] + (1 - alpha) * bg, 0, 1)),
 * :1297-1298 (both images times the mask), :1301 (Ll1 = mean |gt - pred|), :1302 (simloss = 1 - SSIM(gt, pred)) and
 * :1319 (main = (1 - lambda) Ll1 + lambda simloss).  SSIM is pytorch_msssim (third-party dependency of the reference,
 * pyproject.toml:60, unpinned, not under /root/reference); its published algorithm, v1.0.0 ssim.py, is restated:
 *   _fspecial_gauss_1d: g[i] = exp(-(i - size//2)^2 / (2 sigma^2)) / sum            (fp32)
 *   gaussian_filter:    depthwise valid (no padding) convolution along H, then along W
 *   _ssim:              mu, sigma from the five filtered moments, K = (0.01, 0.03), data_range 1,
 *                       ssim_map = (2 mu1 mu2 + C1)/(mu1^2 + mu2^2 + C1) * (2 s12 + C2)/(s1 + s2 + C2),
 *                       mean over positions per channel, then over channels (size_average).
 * The backward is the hand-derived adjoint of the above (what torch.autograd produces for the same graph); it is
 * pinned against torch.autograd by tests/golden/loss_golden.npz (tests/golden/make_golden_loss.py).  Parity with the
 * pytorch_msssim PACKAGE itself is unpinned (not installable here); the golden generator restates it in torch ops.
 * Sums are accumulated in double so the oracle is the more accurate side of every comparison.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

#define WIN 11

static float clamp_unit(float v) { return v < 0.f ? 0.f : (v > 1.f ? 1.f : v); }

static void make_window(float sigma, float* g) {
    float s = 0.f;
    for (int i = 0; i < WIN; ++i) {
        const float x = (float)(i - WIN / 2);
        g[i] = expf(-(x * x) / (2.f * sigma * sigma));
        s += g[i];
    }
    for (int i = 0; i < WIN; ++i) g[i] /= s;
}

/* valid separable filter of one (H, W) plane: along H first, then along W -> (H - 10, W - 10) */
static void filter_plane(const float* in, int H, int W, const float* g, float* tmp, float* out) {
    const int Hv = H - (WIN - 1), Wv = W - (WIN - 1);
    for (int i = 0; i < Hv; ++i)
        for (int j = 0; j < W; ++j) {
            float a = 0.f;
            for (int q = 0; q < WIN; ++q) a += g[q] * in[(size_t)(i + q) * W + j];
            tmp[(size_t)i * W + j] = a;
        }
    for (int i = 0; i < Hv; ++i)
        for (int j = 0; j < Wv; ++j) {
            float a = 0.f;
            for (int q = 0; q < WIN; ++q) a += g[q] * tmp[(size_t)i * W + j + q];
            out[(size_t)i * Wv + j] = a;
        }
}

/* adjoint of filter_plane: (H - 10, W - 10) -> (H, W), ADDED to out scaled by nothing (out must be zeroed) */
static void filter_plane_adjoint(const float* in, int H, int W, const float* g, float* tmp, float* out) {
    const int Hv = H - (WIN - 1), Wv = W - (WIN - 1);
    for (size_t i = 0; i < (size_t)Hv * W; ++i) tmp[i] = 0.f;
    for (int i = 0; i < Hv; ++i)
        for (int j = 0; j < Wv; ++j)
            for (int q = 0; q < WIN; ++q) tmp[(size_t)i * W + j + q] += g[q] * in[(size_t)i * Wv + j];
    for (int i = 0; i < Hv; ++i)
        for (int j = 0; j < W; ++j)
            for (int q = 0; q < WIN; ++q) out[(size_t)(i + q) * W + j] += g[q] * tmp[(size_t)i * W + j];
}

/* out_img (H,W,3), out_texture (H,W,C), out_alpha (H,W), background (3), gt (H,W,3), mask (H,W) or NULL.
 * rgb (H,W,3); losses[3] = {Ll1, SSIM, main}; when v_out_img != NULL also the gradients of v_main * main:
 * v_out_img (H,W,3), v_out_texture (H,W,C), v_out_alpha (H,W). */
int orc_photometric_loss(int H, int W, int C, const float* out_img, const float* out_texture, const float* out_alpha,
                         const float* background, const float* gt, const float* mask, float ssim_lambda, float win_sigma,
                         float v_main, float* rgb, float* losses, float* v_out_img, float* v_out_texture,
                         float* v_out_alpha) {
    if (H < WIN || W < WIN || C < 3) return -1;
    const int Hv = H - (WIN - 1), Wv = W - (WIN - 1);
    const size_t P = (size_t)H * W, Pv = (size_t)Hv * Wv;
    float g[WIN];
    make_window(win_sigma, g);
    float* pre = (float*)malloc(P * 3 * sizeof(float));
    float* x = (float*)malloc(P * sizeof(float));   /* one channel of gt * mask */
    float* y = (float*)malloc(P * sizeof(float));   /* one channel of pred * mask */
    float* prod = (float*)malloc(P * sizeof(float));
    float* tmp = (float*)malloc(P * sizeof(float));
    float* mom[5];
    float* dmap[3];
    float* adj[3];
    for (int m = 0; m < 5; ++m) mom[m] = (float*)malloc(Pv * sizeof(float));
    for (int m = 0; m < 3; ++m) { dmap[m] = (float*)malloc(Pv * sizeof(float)); adj[m] = (float*)malloc(P * sizeof(float)); }
    const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;

    double l1_sum = 0.0, ssim_sum = 0.0;
    for (size_t p = 0; p < P; ++p) {
        const float om = 1.f - out_alpha[p];
        for (int k = 0; k < 3; ++k) {
            const float t = out_img[p * 3 + k] + out_texture[p * C + k];
            const float u = om * background[k];
            pre[p * 3 + k] = t + u;
            rgb[p * 3 + k] = clamp_unit(pre[p * 3 + k]);
        }
    }
    const float w_l1 = v_main * (1.f - ssim_lambda) / (3.f * (float)H * (float)W);
    const float w_ss = -v_main * ssim_lambda / (3.f * (float)Hv * (float)Wv);
    if (v_out_alpha) for (size_t p = 0; p < P; ++p) v_out_alpha[p] = 0.f;
    for (int k = 0; k < 3; ++k) {
        for (size_t p = 0; p < P; ++p) {
            const float m = mask ? mask[p] : 1.f;
            x[p] = gt[p * 3 + k] * m;
            y[p] = rgb[p * 3 + k] * m;
            l1_sum += fabsf(x[p] - y[p]);
        }
        filter_plane(x, H, W, g, tmp, mom[0]);
        filter_plane(y, H, W, g, tmp, mom[1]);
        for (size_t p = 0; p < P; ++p) prod[p] = x[p] * x[p];
        filter_plane(prod, H, W, g, tmp, mom[2]);
        for (size_t p = 0; p < P; ++p) prod[p] = y[p] * y[p];
        filter_plane(prod, H, W, g, tmp, mom[3]);
        for (size_t p = 0; p < P; ++p) prod[p] = x[p] * y[p];
        filter_plane(prod, H, W, g, tmp, mom[4]);
        for (size_t q = 0; q < Pv; ++q) {
            const float mu1 = mom[0][q], mu2 = mom[1][q];
            const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
            const float s1 = mom[2][q] - mu1_sq, s2 = mom[3][q] - mu2_sq, s12 = mom[4][q] - mu12;
            const float A1 = 2.f * mu12 + C1, A2 = 2.f * s12 + C2, B1 = mu1_sq + mu2_sq + C1, B2 = s1 + s2 + C2;
            const float cs = A2 / B2, lum = A1 / B1;
            ssim_sum += lum * cs;
            /* d(lum*cs) with E[y], E[y^2], E[xy] as the independent moments */
            const float dyy = -lum * cs / B2;
            const float dxy = 2.f * lum / B2;
            dmap[0][q] = 2.f * (mu1 * cs - mu2 * lum * cs) / B1 - 2.f * mu2 * dyy - mu1 * dxy;
            dmap[1][q] = dyy;
            dmap[2][q] = dxy;
        }
        if (v_out_img) {
            for (int m = 0; m < 3; ++m) {
                for (size_t p = 0; p < P; ++p) adj[m][p] = 0.f;
                filter_plane_adjoint(dmap[m], H, W, g, tmp, adj[m]);
            }
            for (size_t p = 0; p < P; ++p) {
                const float m = mask ? mask[p] : 1.f;
                const float diff = x[p] - y[p];
                const float sgn = diff > 0.f ? -1.f : (diff < 0.f ? 1.f : 0.f);
                float gr = w_l1 * sgn + w_ss * (adj[0][p] + 2.f * y[p] * adj[1][p] + x[p] * adj[2][p]);
                gr *= m;
                const float pv = pre[p * 3 + k];
                if (!(pv >= 0.f && pv <= 1.f)) gr = 0.f; /* torch.clamp's backward mask: min <= x <= max */
                v_out_img[p * 3 + k] = gr;
                v_out_texture[p * C + k] = gr;
                v_out_alpha[p] -= gr * background[k];
            }
        }
    }
    if (v_out_texture)
        for (size_t p = 0; p < P; ++p)
            for (int k = 3; k < C; ++k) v_out_texture[p * C + k] = 0.f;
    losses[0] = (float)(l1_sum / (3.0 * (double)P));
    losses[1] = (float)(ssim_sum / (3.0 * (double)Pv));
    losses[2] = (1.f - ssim_lambda) * losses[0] + ssim_lambda * (1.f - losses[1]);
    free(pre); free(x); free(y); free(prod); free(tmp);
    for (int m = 0; m < 5; ++m) free(mom[m]);
    for (int m = 0; m < 3; ++m) { free(dmap[m]); free(adj[m]); }
    return 0;
}

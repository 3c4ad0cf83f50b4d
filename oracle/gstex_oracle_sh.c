/* gstex_oracle_sh.c -- CPU restatement of the reference's spherical-harmonics colour stage.
 * TEST INFRASTRUCTURE, NOT PRODUCT CODE (see oracle/__init__.py).
 *
 * Follows gstex_cuda/cuda/csrc/sh.cuh: constants :6-33, num_sh_bases :35-45, sh_coeffs_to_color :47-119,
 * sh_coeffs_to_color_vjp :121-215, the one-thread-per-point kernels :217-253 and the zero-initialised
 * v_coeffs of bindings.cu:62-63.  Pinned against the reference's own Python
 * (_torch_impl.compute_sh_color + autograd) by tests/golden/sh_golden.npz.
 *
 * Written table-driven: every basis function above degree 1 is  K * f1 * f2  with the reference's
 * multiplication order (K*f1 first, then *f2, then the coefficient); f2 = 1 where the reference has a single
 * factor (multiplying by 1.0f is exact), so the fp32 roundings are the reference's.
 */
#include <math.h>
#include <stdint.h>

static const float K0 = 0.28209479177387814f, K1 = 0.4886025119029199f;
static const float KHI[21] = {
    /* degree 2, sh.cuh:8-13 */ 1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f,
    0.5462742152960396f,
    /* degree 3, :14-21 */ -0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
    -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f,
    /* degree 4, :22-31 */ 2.5033429417967046f, -1.7701307697799304f, 0.9461746957575601f, -0.6690465435572892f,
    0.10578554691520431f, -0.6690465435572892f, 0.47308734787878004f, -1.7701307697799304f, 0.6258357354491761f};

static int nbases(int degree) { return degree == 0 ? 1 : degree == 1 ? 4 : degree == 2 ? 9 : degree == 3 ? 16 : 25; }

/* factors f1[k], f2[k] of bases 4..24 at the unit direction (x, y, z); sh.cuh:83-117 */
static void hi_factors(float x, float y, float z, float* f1, float* f2) {
    const float xx = x * x, xy = x * y, xz = x * z, yy = y * y, yz = y * z, zz = z * z;
    int k = 0;
#define F(a, b) do { f1[k] = (a); f2[k] = (b); ++k; } while (0)
    F(xy, 1.f); F(yz, 1.f); F(2.f * zz - xx - yy, 1.f); F(xz, 1.f); F(xx - yy, 1.f);
    F(y, 3.f * xx - yy); F(xy, z); F(y, 4.f * zz - xx - yy); F(z, 2.f * zz - 3.f * xx - 3.f * yy);
    F(x, 4.f * zz - xx - yy); F(z, xx - yy); F(x, xx - 3.f * yy);
    F(xy, xx - yy); F(yz, 3.f * xx - yy); F(xy, 7.f * zz - 1.f); F(yz, 7.f * zz - 3.f);
    F(zz * (35.f * zz - 30.f) + 3.f, 1.f); F(xz, 7.f * zz - 3.f); F(xx - yy, 7.f * zz - 1.f); F(xz, xx - 3.f * yy);
    F(xx * (xx - 3.f * yy) - yy * (3.f * xx - yy), 1.f);
#undef F
}

static void unit(const float* v, float* x, float* y, float* z) { /* sh.cuh:62-67 */
    const float norm = sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    *x = v[0] / norm; *y = v[1] / norm; *z = v[2] / norm;
}

/* colors (n,3) from viewdirs (n,3), coeffs (n, nbases(degree), 3) using bases of `degrees_to_use` */
void orc_sh_forward(int n, int degree, int degrees_to_use, const float* viewdirs, const float* coeffs, float* colors) {
    const int D = nbases(degree);
    const int group_end[5] = {1, 4, 9, 16, 25};
    const int use = degrees_to_use > 4 ? 4 : degrees_to_use;
#pragma omp parallel for schedule(static)
    for (int p = 0; p < n; ++p) {
        const float* cf = coeffs + (int64_t)p * D * 3;
        float* out = colors + (int64_t)p * 3;
        for (int c = 0; c < 3; ++c) out[c] = K0 * cf[c];
        if (use < 1) continue;
        float x, y, z, f1[21], f2[21];
        unit(viewdirs + (int64_t)p * 3, &x, &y, &z);
        hi_factors(x, y, z, f1, f2);
        for (int c = 0; c < 3; ++c) {
            out[c] += K1 * (-y * cf[3 + c] + z * cf[6 + c] - x * cf[9 + c]); /* :77-79 */
            for (int g = 2; g <= use; ++g) {                                 /* one `colors[c] +=` per degree, :83-117 */
                float s = 0.f;
                for (int k = group_end[g - 1]; k < group_end[g]; ++k) {
                    const float t = KHI[k - 4] * f1[k - 4] * f2[k - 4] * cf[3 * k + c];
                    s = (k == group_end[g - 1]) ? t : s + t;
                }
                out[c] += s;
            }
        }
    }
}

/* v_coeffs (n, nbases(degree), 3) = basis * v_colors for the used bases, zero above */
void orc_sh_backward(int n, int degree, int degrees_to_use, const float* viewdirs, const float* v_colors,
                     float* v_coeffs) {
    const int D = nbases(degree);
    const int use = degrees_to_use > 4 ? 4 : degrees_to_use;
    const int used = nbases(use);
#pragma omp parallel for schedule(static)
    for (int p = 0; p < n; ++p) {
        float* vc = v_coeffs + (int64_t)p * D * 3;
        const float* g = v_colors + (int64_t)p * 3;
        for (int k = 0; k < D * 3; ++k) vc[k] = 0.f;
        for (int c = 0; c < 3; ++c) vc[c] = K0 * g[c];
        if (use < 1) continue;
        float x, y, z, f1[21], f2[21];
        unit(viewdirs + (int64_t)p * 3, &x, &y, &z);
        hi_factors(x, y, z, f1, f2);
        for (int c = 0; c < 3; ++c) {
            vc[3 + c] = (-K1 * y) * g[c]; vc[6 + c] = (K1 * z) * g[c]; vc[9 + c] = (-K1 * x) * g[c]; /* :150-155 */
            for (int k = 4; k < used; ++k) vc[3 * k + c] = (KHI[k - 4] * f1[k - 4] * f2[k - 4]) * g[c];
        }
    }
}

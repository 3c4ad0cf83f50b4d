// sh.cu -- per-splat colour stage in front of the rasterizer: view-dependent colour
// from real spherical-harmonic coefficients and its VJP with respect to the coefficients.
// Replaces compute_sh_forward / compute_sh_backward (gstex_cuda/cuda/csrc/sh.cuh:46-254, wrappers
// bindings.cu:18-75; Python gstex_cuda/sh.py:36-96).  sm_100a.
//
// Both directions are pure streams (no reuse): forward reads 12 + 12*used bytes per point and writes 12, backward
// reads 24 and writes 12*D.  The reference gives each thread one point and lets it walk its 12*D-byte row with
// 4-byte accesses at a 12*D-byte stride between neighbouring threads.  Here a block owns 128 consecutive points
// = one contiguous slab of the (N, D, 3) array: the slab moves between global and shared memory with
// fully coalesced accesses (consecutive threads, consecutive addresses; only the `used` leading bases of every
// row are read in the forward), the per-point arithmetic runs out of a row-padded shared tile (stride 3*D + 1
// words: conflict-free), and the basis functions are evaluated once per point instead of once per channel.
#include "common.cuh"

namespace gstex {

constexpr int kShPoints = 128;      // points per block = threads per block
constexpr int kShMaxBases = 25;     // degree 4

__host__ __device__ inline int sh_num_bases(int degree) {  // sh.cuh:35-45 (anything above 3 is degree 4)
    return degree == 0 ? 1 : degree == 1 ? 4 : degree == 2 ? 9 : degree == 3 ? 16 : 25;
}

// Real SH basis values b[0 .. num_bases(degree)) at direction v (normalised here, as sh.cuh:62-67 does; degree 0
// never touches v).  Constants: sh.cuh:6-33.
__device__ __forceinline__ void sh_basis(int degree, float vx, float vy, float vz, float* b) {
    b[0] = 0.28209479177387814f;
    if (degree < 1) return;
    const float norm = sqrtf(vx * vx + vy * vy + vz * vz);
    const float x = vx / norm, y = vy / norm, z = vz / norm;
    const float c1 = 0.4886025119029199f;
    b[1] = -c1 * y; b[2] = c1 * z; b[3] = -c1 * x;
    if (degree < 2) return;
    const float xx = x * x, xy = x * y, xz = x * z, yy = y * y, yz = y * z, zz = z * z;
    b[4] = 1.0925484305920792f * xy;
    b[5] = -1.0925484305920792f * yz;
    b[6] = 0.31539156525252005f * (2.f * zz - xx - yy);
    b[7] = -1.0925484305920792f * xz;
    b[8] = 0.5462742152960396f * (xx - yy);
    if (degree < 3) return;
    b[9] = -0.5900435899266435f * y * (3.f * xx - yy);
    b[10] = 2.890611442640554f * xy * z;
    b[11] = -0.4570457994644658f * y * (4.f * zz - xx - yy);
    b[12] = 0.3731763325901154f * z * (2.f * zz - 3.f * xx - 3.f * yy);
    b[13] = -0.4570457994644658f * x * (4.f * zz - xx - yy);
    b[14] = 1.445305721320277f * z * (xx - yy);
    b[15] = -0.5900435899266435f * x * (xx - 3.f * yy);
    if (degree < 4) return;
    b[16] = 2.5033429417967046f * xy * (xx - yy);
    b[17] = -1.7701307697799304f * yz * (3.f * xx - yy);
    b[18] = 0.9461746957575601f * xy * (7.f * zz - 1.f);
    b[19] = -0.6690465435572892f * yz * (7.f * zz - 3.f);
    b[20] = 0.10578554691520431f * (zz * (35.f * zz - 30.f) + 3.f);
    b[21] = -0.6690465435572892f * xz * (7.f * zz - 3.f);
    b[22] = 0.47308734787878004f * (xx - yy) * (7.f * zz - 1.f);
    b[23] = -1.7701307697799304f * xz * (xx - 3.f * yy);
    b[24] = 0.6258357354491761f * (xx * (xx - 3.f * yy) - yy * (3.f * xx - yy));
}

// colors[p][c] = sum_k basis_k(viewdirs[p]) * coeffs[p][k][c], k < num_bases(degrees_to_use)
__global__ void __launch_bounds__(kShPoints)
sh_forward_kernel(int n, int num_bases, int degrees_to_use, const float* __restrict__ viewdirs,
                  const float* __restrict__ coeffs, float* __restrict__ colors, const bool vec4) {
    __shared__ float tile[kShPoints * (3 * kShMaxBases + 1)];
    const int used3 = 3 * sh_num_bases(degrees_to_use), row3 = 3 * num_bases, pitch = used3 + 1;
    const long long p0 = (long long)blockIdx.x * kShPoints;
    const int npts = (int)min((long long)kShPoints, (long long)n - p0);
    const float* src = coeffs + p0 * row3;
    if (used3 == row3) {  // whole slab is contiguous: 16-byte loads, (point, offset) advanced without divisions
        const int total = npts * row3;
        int j = threadIdx.x * 4;
        if (vec4) {
            int pp = j / row3, kk = j - pp * row3;
            const int dp = (4 * kShPoints) / row3, dk = (4 * kShPoints) - dp * row3;
            for (; j + 3 < total; j += 4 * kShPoints) {
                const float4 x = __ldcs(reinterpret_cast<const float4*>(src + j));
                const float xs[4] = {x.x, x.y, x.z, x.w};
                int q = pp, k = kk;
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    tile[q * pitch + k] = xs[e];
                    if (++k == row3) { k = 0; ++q; }
                }
                pp += dp; kk += dk;
                if (kk >= row3) { kk -= row3; ++pp; }
            }
        }
        // scalar: everything when unaligned, else the (at most 3) trailing floats of a partial last block
        for (int i = vec4 ? (total & ~3) + threadIdx.x : threadIdx.x; i < total; i += kShPoints) {
            const int q = i / row3, k = i - q * row3;
            tile[q * pitch + k] = __ldcs(src + i);
        }
    } else {              // leading `used3` floats of every row
        for (int j = threadIdx.x; j < npts * used3; j += kShPoints) {
            const int p = j / used3, k = j - p * used3;
            tile[p * pitch + k] = __ldcs(src + (long long)p * row3 + k);
        }
    }
    __syncthreads();
    const int p = threadIdx.x;
    if (p < npts) {
        const float* v = viewdirs + (p0 + p) * 3;
        float b[kShMaxBases];
        sh_basis(degrees_to_use, __ldg(v), __ldg(v + 1), __ldg(v + 2), b);
        const float* row = tile + p * pitch;
        float c0 = 0.f, c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int k = 0; k < kShMaxBases; ++k) {
            if (3 * k < used3) {
                c0 += b[k] * row[3 * k];
                c1 += b[k] * row[3 * k + 1];
                c2 += b[k] * row[3 * k + 2];
            }
        }
        float* o = colors + (p0 + p) * 3;
        o[0] = c0; o[1] = c1; o[2] = c2;
    }
}

// v_coeffs[p][k][c] = basis_k(viewdirs[p]) * v_colors[p][c] for k < num_bases(degrees_to_use), 0 above
// (the reference zero-fills the output, bindings.cu:62-63)
__global__ void __launch_bounds__(kShPoints)
sh_backward_kernel(int n, int num_bases, int degrees_to_use, const float* __restrict__ viewdirs,
                   const float* __restrict__ v_colors, float* __restrict__ v_coeffs, const bool vec4) {
    __shared__ float tile[kShPoints * (3 * kShMaxBases + 1)];
    const int used = sh_num_bases(degrees_to_use), row3 = 3 * num_bases, pitch = row3 + 1;
    const long long p0 = (long long)blockIdx.x * kShPoints;
    const int npts = (int)min((long long)kShPoints, (long long)n - p0);
    const int p = threadIdx.x;
    if (p < npts) {
        const float* v = viewdirs + (p0 + p) * 3;
        const float* g = v_colors + (p0 + p) * 3;
        float b[kShMaxBases];
        sh_basis(degrees_to_use, __ldg(v), __ldg(v + 1), __ldg(v + 2), b);
        const float g0 = __ldg(g), g1 = __ldg(g + 1), g2 = __ldg(g + 2);
        float* row = tile + p * pitch;
#pragma unroll
        for (int k = 0; k < kShMaxBases; ++k) {
            if (k < num_bases) {
                const float bk = (k < used) ? b[k] : 0.f;
                row[3 * k] = (k < used) ? bk * g0 : 0.f;
                row[3 * k + 1] = (k < used) ? bk * g1 : 0.f;
                row[3 * k + 2] = (k < used) ? bk * g2 : 0.f;
            }
        }
    }
    __syncthreads();
    float* dst = v_coeffs + p0 * row3;
    const int total = npts * row3;
    if (vec4) {  // 16-byte stores, (point, offset) advanced without divisions
        int j = threadIdx.x * 4;
        int pp = j / row3, kk = j - pp * row3;
        const int dp = (4 * kShPoints) / row3, dk = (4 * kShPoints) - dp * row3;
        for (; j + 3 < total; j += 4 * kShPoints) {
            float xs[4];
            int q = pp, k = kk;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                xs[e] = tile[q * pitch + k];
                if (++k == row3) { k = 0; ++q; }
            }
            __stcs(reinterpret_cast<float4*>(dst + j), make_float4(xs[0], xs[1], xs[2], xs[3]));
            pp += dp; kk += dk;
            if (kk >= row3) { kk -= row3; ++pp; }
        }
    }
    for (int i = vec4 ? (total & ~3) + threadIdx.x : threadIdx.x; i < total; i += kShPoints) {
        const int q = i / row3, k = i - q * row3;
        __stcs(dst + i, tile[q * pitch + k]);
    }
}

}  // namespace gstex

using namespace gstex;

static int sh_check(int n, int degree, int degrees_to_use) {
    if (n < 0 || degree < 0 || degree > 4 || degrees_to_use < 0) return GSTEX_E_BADARG;
    if (sh_num_bases(degrees_to_use) > sh_num_bases(degree)) return GSTEX_E_BADARG;  // sh.py:52
    return GSTEX_OK;
}

extern "C" int gstex_compute_sh_forward(int n, int degree, int degrees_to_use, const float* viewdirs, const float* coeffs,
                                        float* colors, void* stream) {
    const int rc = sh_check(n, degree, degrees_to_use);
    if (rc != GSTEX_OK) return rc;
    if (n == 0) return GSTEX_OK;
    if (!viewdirs || !coeffs || !colors) return GSTEX_E_BADARG;
    sh_forward_kernel<<<div_up(n, kShPoints), kShPoints, 0, (cudaStream_t)stream>>>(n, sh_num_bases(degree), degrees_to_use,
                                                                                     viewdirs, coeffs, colors, ((uintptr_t)coeffs & 15) == 0);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

extern "C" int gstex_compute_sh_backward(int n, int degree, int degrees_to_use, const float* viewdirs,
                                         const float* v_colors, float* v_coeffs, void* stream) {
    const int rc = sh_check(n, degree, degrees_to_use);
    if (rc != GSTEX_OK) return rc;
    if (n == 0) return GSTEX_OK;
    if (!viewdirs || !v_colors || !v_coeffs) return GSTEX_E_BADARG;
    sh_backward_kernel<<<div_up(n, kShPoints), kShPoints, 0, (cudaStream_t)stream>>>(n, sh_num_bases(degree), degrees_to_use,
                                                                                      viewdirs, v_colors, v_coeffs, ((uintptr_t)v_coeffs & 15) == 0);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

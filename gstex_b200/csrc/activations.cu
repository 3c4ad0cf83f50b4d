// activations.cu -- next row f-1: the per-splat activation stage the reference model
// runs before every view as ~30 torch elementwise / reduction launches over N
// (/root/reference/nerfstudio/models/gstex.py:1062-1066 and get_uv_mapping, :975-990), as ONE kernel forward and one
// backward.  sm_100a.
//
//   quats     = quats_raw / ||quats_raw||                                   gstex.py:1062
//   scales    = (clamp(exp(s0), 1e-9), clamp(exp(s1), 1e-9), 1e-5 * mean(those two) [detached])   :1063-1065
//   opacities = sigmoid(opacities_raw)                                      :1066
//   uv0 = 0.5;  umap = mappings[:,0] * ax1;  vmap = mappings[:,1] * ax2      :975-990, axes = columns of
//         R(F.normalize(quats)) (gstex_cuda/_torch_impl.py:116-138), all detached
//
// Bit-exact with the torch ops on the same GPU: every operation is an explicit round-to-nearest intrinsic in the order
// eager torch executes it (one kernel per op there, so nothing is ever contracted into an fma), with the two forms
// that are not obvious read off the running library (scripts/probe_torch_forms.py): `norm(dim=-1)` over four elements
// adds the ROUNDED squares as (q0^2 + q2^2) + (q1^2 + q3^2) (two vector lanes, then combined), and a division by a
// tensor is a true division.  exp is libdevice's expf on both sides, sigmoid is 1 / (1 + exp(-x)).
#include "common.cuh"

namespace gstex {

__device__ __forceinline__ float norm4_torch(float a, float b, float c, float d) {
    return __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(a, a), __fmul_rn(c, c)), __fadd_rn(__fmul_rn(b, b), __fmul_rn(d, d))));
}

__global__ void __launch_bounds__(kThreads1D)
activations_forward_kernel(int n, const float4* __restrict__ quats_raw, const float* __restrict__ scales_raw,
                           const float* __restrict__ opacities_raw, const float2* __restrict__ mappings,
                           float4* __restrict__ quats, float* __restrict__ scales, float* __restrict__ opacities,
                           float2* __restrict__ uv0, float* __restrict__ umap, float* __restrict__ vmap) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 q = quats_raw[i];
    const float nq = norm4_torch(q.x, q.y, q.z, q.w);
    const float4 qn = make_float4(__fdiv_rn(q.x, nq), __fdiv_rn(q.y, nq), __fdiv_rn(q.z, nq), __fdiv_rn(q.w, nq));
    quats[i] = qn;
    const float s0 = fmaxf(expf(scales_raw[3 * i]), 1e-9f), s1 = fmaxf(expf(scales_raw[3 * i + 1]), 1e-9f);
    scales[3 * i] = s0;
    scales[3 * i + 1] = s1;
    scales[3 * i + 2] = __fmul_rn(1e-5f, __fmul_rn(__fadd_rn(s0, s1), 0.5f));
    opacities[i] = __fdiv_rn(1.f, __fadd_rn(1.f, expf(-opacities_raw[i])));
    if (uv0 == nullptr) return;
    // F.normalize(quats): x / max(||x||, 1e-12)
    const float n2 = fmaxf(norm4_torch(qn.x, qn.y, qn.z, qn.w), 1e-12f);
    const float w = __fdiv_rn(qn.x, n2), x = __fdiv_rn(qn.y, n2), y = __fdiv_rn(qn.z, n2), z = __fdiv_rn(qn.w, n2);
    const float xx = __fmul_rn(x, x), yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
    const float xy = __fmul_rn(x, y), xz = __fmul_rn(x, z), yz = __fmul_rn(y, z);
    const float wx = __fmul_rn(w, x), wy = __fmul_rn(w, y), wz = __fmul_rn(w, z);
    // first two columns of normalized_quat_to_rotmat
    const float a1x = __fsub_rn(1.f, __fmul_rn(2.f, __fadd_rn(yy, zz)));
    const float a1y = __fmul_rn(2.f, __fadd_rn(xy, wz));
    const float a1z = __fmul_rn(2.f, __fsub_rn(xz, wy));
    const float a2x = __fmul_rn(2.f, __fsub_rn(xy, wz));
    const float a2y = __fsub_rn(1.f, __fmul_rn(2.f, __fadd_rn(xx, zz)));
    const float a2z = __fmul_rn(2.f, __fadd_rn(yz, wx));
    const float2 m = mappings[i];
    uv0[i] = make_float2(0.5f, 0.5f);
    umap[3 * i] = __fmul_rn(m.x, a1x); umap[3 * i + 1] = __fmul_rn(m.x, a1y); umap[3 * i + 2] = __fmul_rn(m.x, a1z);
    vmap[3 * i] = __fmul_rn(m.y, a2x); vmap[3 * i + 1] = __fmul_rn(m.y, a2y); vmap[3 * i + 2] = __fmul_rn(m.y, a2z);
}

// VJP of the differentiable part (quats, scales[:, :2], opacities; the third scale and the uv maps are detached in the
// reference).  Formulas of torch's own backward nodes: div / norm, exp + clamp(min), sigmoid.
__global__ void __launch_bounds__(kThreads1D)
activations_backward_kernel(int n, const float4* __restrict__ quats_raw, const float* __restrict__ scales_raw,
                            const float4* __restrict__ quats, const float* __restrict__ scales,
                            const float* __restrict__ opacities, const float4* __restrict__ v_quats,
                            const float* __restrict__ v_scales, const float* __restrict__ v_opacities,
                            float4* __restrict__ v_quats_raw, float* __restrict__ v_scales_raw,
                            float* __restrict__ v_opacities_raw) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (v_quats_raw != nullptr) {
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
        if (v_quats != nullptr) {  // y = q / |q|:  v_q = (v_y - y (y . v_y)) / |q|
            const float4 q = quats_raw[i], y = quats[i], v = v_quats[i];
            const float nq = norm4_torch(q.x, q.y, q.z, q.w);
            const float dot = y.x * v.x + y.y * v.y + y.z * v.z + y.w * v.w;
            g = make_float4((v.x - y.x * dot) / nq, (v.y - y.y * dot) / nq, (v.z - y.z * dot) / nq, (v.w - y.w * dot) / nq);
        }
        v_quats_raw[i] = g;
    }
    if (v_scales_raw != nullptr) {
#pragma unroll
        for (int k = 0; k < 2; ++k) {  // clamp(exp(s), min): the gradient passes where exp(s) >= min, times exp(s)
            const float e = expf(scales_raw[3 * i + k]);
            v_scales_raw[3 * i + k] = (v_scales != nullptr && e >= 1e-9f) ? v_scales[3 * i + k] * scales[3 * i + k] : 0.f;
        }
        v_scales_raw[3 * i + 2] = 0.f;  // scales[:, 2] = 1e-5 * mean(...).detach()
    }
    if (v_opacities_raw != nullptr) {  // sigmoid: grad * (1 - y) * y
        const float y = opacities[i];
        v_opacities_raw[i] = v_opacities != nullptr ? v_opacities[i] * (1.f - y) * y : 0.f;
    }
}

}  // namespace gstex

using namespace gstex;

extern "C" int gstex_activations_forward(int n, const float* quats_raw, const float* scales_raw,
                                         const float* opacities_raw, const float* mappings, float* quats, float* scales,
                                         float* opacities, float* uv0, float* umap, float* vmap, void* stream) {
    if (n < 0) return GSTEX_E_BADARG;
    if (n == 0) return GSTEX_OK;
    if (!quats_raw || !scales_raw || !opacities_raw || !quats || !scales || !opacities) return GSTEX_E_BADARG;
    const bool maps = uv0 != nullptr || umap != nullptr || vmap != nullptr;
    if (maps && (!uv0 || !umap || !vmap || !mappings)) return GSTEX_E_BADARG;
    if (((uintptr_t)quats_raw | (uintptr_t)quats) & 15) return GSTEX_E_BADARG;
    activations_forward_kernel<<<div_up(n, kThreads1D), kThreads1D, 0, (cudaStream_t)stream>>>(
        n, reinterpret_cast<const float4*>(quats_raw), scales_raw, opacities_raw, reinterpret_cast<const float2*>(mappings),
        reinterpret_cast<float4*>(quats), scales, opacities, maps ? reinterpret_cast<float2*>(uv0) : nullptr, umap, vmap);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

extern "C" int gstex_activations_backward(int n, const float* quats_raw, const float* scales_raw, const float* quats,
                                          const float* scales, const float* opacities, const float* v_quats,
                                          const float* v_scales, const float* v_opacities, float* v_quats_raw,
                                          float* v_scales_raw, float* v_opacities_raw, void* stream) {
    if (n < 0) return GSTEX_E_BADARG;
    if (n == 0) return GSTEX_OK;
    if (!quats_raw || !scales_raw || !quats || !scales || !opacities) return GSTEX_E_BADARG;
    if (((uintptr_t)quats_raw | (uintptr_t)quats | (uintptr_t)v_quats | (uintptr_t)v_quats_raw) & 15) return GSTEX_E_BADARG;
    activations_backward_kernel<<<div_up(n, kThreads1D), kThreads1D, 0, (cudaStream_t)stream>>>(
        n, reinterpret_cast<const float4*>(quats_raw), scales_raw, reinterpret_cast<const float4*>(quats), scales, opacities,
        reinterpret_cast<const float4*>(v_quats), v_scales, v_opacities, reinterpret_cast<float4*>(v_quats_raw),
        v_scales_raw, v_opacities_raw);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

// charts.cu -- next row f-3: re-charting.  Every 100 training steps the reference model
// re-allocates the per-splat texel grids so that their total hits `pixel_num` and resamples the texture into them:
//   build_charts            /root/reference/nerfstudio/models/gstex.py:841-888  (bisection on the texel size, chart
//                           dims {h, w, start}, uv mappings)
//   JaggedTexture.init_from_dims   /root/reference/nerfstudio/models/jagged_texture.py:116-143 (+ texture_dims_to_query
//                           :23-35): every texel of the NEW charts is a bilinear sample of the OLD chart of its splat.
// The reference runs the bisection as up to 31 sequential rounds of ~8 torch launches and one blocking `.item()` each,
// and the resampling through three materialised per-texel index tensors (ids, query_dims, uv: 28 bytes per texel).
// Here: one pass over the splats scores up to 31 candidate texel sizes at once (a whole 5-level bisection subtree, so
// the 31 rounds need at most 7 passes and 7 read-backs), one kernel writes dims and mappings, and one kernel resamples
// with the splat of each new texel found by binary search over the chart starts (nothing per-texel is materialised).
// fp32 arithmetic is written in the order of the torch ops it replaces, so scores, dims, mappings and uv are the
// reference's bit for bit.  sm_100a.
#include "common.cuh"

namespace gstex {

constexpr int kMaxCandidates = 32;

struct ChartCandidates {
    float x[kMaxCandidates];  // candidate texel sizes
    unsigned recip_mask;      // bit k: candidate k divides the way `tensor / python_float` does (multiply by 1/x)
    int k;
};

// ceil(sigma_factor * length / x) as the reference computes it (gstex.py:848,876-877): length = exp(scale);
// `sigma_factor * length` is one fp32 multiply; the division is a true division when x is a tensor (every round of the
// loop: `mid / adjustments`) and a multiplication by the fp32 reciprocal when x is a python float (the first score).
__device__ __forceinline__ float chart_cells(float sl, float x, bool recip) {
    return ceilf(recip ? __fmul_rn(sl, __fdiv_rn(1.f, x)) : __fdiv_rn(sl, x));
}

// scores[k] = sum_i ceil(sf * l0_i / x_k) * ceil(sf * l1_i / x_k), accumulated in double: every term is an integer, so
// the sum is exact (and independent of the order of the atomics) as long as it stays below 2^53.
__global__ void __launch_bounds__(256)
chart_scores_kernel(int n, const float* __restrict__ scales_raw, float sigma_factor, const ChartCandidates c,
                    double* __restrict__ scores) {
    __shared__ double warp_part[8][kMaxCandidates];
    double acc[kMaxCandidates];
#pragma unroll
    for (int k = 0; k < kMaxCandidates; ++k) acc[k] = 0.0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float sl0 = __fmul_rn(sigma_factor, expf(scales_raw[3 * i]));
        const float sl1 = __fmul_rn(sigma_factor, expf(scales_raw[3 * i + 1]));
#pragma unroll
        for (int k = 0; k < kMaxCandidates; ++k) {
            if (k < c.k) {
                const bool r = (c.recip_mask >> k) & 1u;
                acc[k] += (double)__fmul_rn(chart_cells(sl0, c.x[k], r), chart_cells(sl1, c.x[k], r));
            }
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < kMaxCandidates; ++k) {
        double v = acc[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) warp_part[warp][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < c.k) {
        double v = 0.0;
        for (int w = 0; w < 8; ++w) v += warp_part[w][threadIdx.x];
        atomicAdd(scores + threadIdx.x, v);
    }
}

// gstex.py:874-883: dims {h, w} from the final texel size (a tensor there: true division), h * w for the scan that
// gives the starts, and mappings = 1 / (2 * sigma_factor * length).
__global__ void __launch_bounds__(kThreads1D)
chart_dims_kernel(int n, const float* __restrict__ scales_raw, float sigma_factor, float two_sigma, float x,
                  int32_t* __restrict__ dims, int32_t* __restrict__ hw, float2* __restrict__ mappings) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float l0 = expf(scales_raw[3 * i]), l1 = expf(scales_raw[3 * i + 1]);
    const int h = (int)chart_cells(__fmul_rn(sigma_factor, l0), x, false);
    const int w = (int)chart_cells(__fmul_rn(sigma_factor, l1), x, false);
    dims[3 * i] = h;
    dims[3 * i + 1] = w;
    hw[i] = h * w;
    if (mappings != nullptr) mappings[i] = make_float2(__fdiv_rn(1.f, __fmul_rn(two_sigma, l0)), __fdiv_rn(1.f, __fmul_rn(two_sigma, l1)));
}

// dims[:, 2] = cumsum(h * w) - h * w  (gstex.py:878-879)
__global__ void __launch_bounds__(kThreads1D)
chart_starts_kernel(int n, const int32_t* __restrict__ cum, const int32_t* __restrict__ hw, int32_t* __restrict__ dims) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dims[3 * i + 2] = cum[i] - hw[i];
}

// JaggedTexture.init_from_dims (jagged_texture.py:116-143): new texel t belongs to the splat i whose new chart contains
// it (repeat_interleave over the new h * w, :118-120); its query point is the texel's integer coordinate over the new
// chart size, uu = float(local / w) / float(h), vv = float(local % w) / float(w) (texture_dims_to_query, :23-35); the
// value is the bilinear sample of the OLD chart of the same splat at (uu, vv) (texture_sample, :138).
__global__ void __launch_bounds__(kThreads1D)
texture_resample_kernel(long long num_new_texels, int n, int C, const int32_t* __restrict__ new_dims,
                        const int32_t* __restrict__ old_dims, const float* __restrict__ old_texture,
                        float* __restrict__ new_texture) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= num_new_texels) return;
    // last splat whose start is <= t (empty charts share their start with the chart that follows them)
    int lo = 0, hi = n - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if ((long long)__ldg(new_dims + 3 * (size_t)mid + 2) <= t) lo = mid;
        else hi = mid - 1;
    }
    const int32_t* nd = new_dims + 3 * (size_t)lo;
    const int h = __ldg(nd), w = __ldg(nd + 1), start = __ldg(nd + 2);
    const int local = (int)(t - start);
    const float uu = __fdiv_rn((float)(local / w), (float)h), vv = __fdiv_rn((float)(local % w), (float)w);
    const int32_t* od = old_dims + 3 * (size_t)lo;
    const TexQuery tq = tex_query(__ldg(od), __ldg(od + 1), __ldg(od + 2), clamp01(uu), clamp01(vv), true);
    const float* t00 = old_texture + (size_t)tq.b00 * C;
    const float* t01 = old_texture + (size_t)tq.b01 * C;
    const float* t10 = old_texture + (size_t)tq.b10 * C;
    const float* t11 = old_texture + (size_t)tq.b11 * C;
    float* o = new_texture + (size_t)t * C;
    for (int k = 0; k < C; ++k)  // the sum in texture_sample_kernel's order (csrc/texture_sample.cu)
        o[k] = tq.w00 * __ldg(t00 + k) + tq.w01 * __ldg(t01 + k) + tq.w10 * __ldg(t10 + k) + tq.w11 * __ldg(t11 + k);
}

}  // namespace gstex

using namespace gstex;

extern "C" int gstex_chart_max_candidates(void) { return kMaxCandidates; }

extern "C" int gstex_chart_scores(int n, const float* scales_raw, float sigma_factor, int num_candidates,
                                  const float* texel_sizes_host, unsigned recip_mask, double* scores, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (n < 0 || num_candidates < 1 || num_candidates > kMaxCandidates || !texel_sizes_host || !scores) return GSTEX_E_BADARG;
    if (n > 0 && !scales_raw) return GSTEX_E_BADARG;
    GSTEX_CUDA_TRY(cudaMemsetAsync(scores, 0, sizeof(double) * num_candidates, st));
    if (n == 0) return GSTEX_OK;
    ChartCandidates c;
    c.k = num_candidates;
    c.recip_mask = recip_mask;
    for (int k = 0; k < kMaxCandidates; ++k) c.x[k] = k < num_candidates ? texel_sizes_host[k] : 1.f;
    int blocks = sm_count() * 4;
    if (blocks > div_up(n, 256)) blocks = div_up(n, 256);
    chart_scores_kernel<<<blocks, 256, 0, st>>>(n, scales_raw, sigma_factor, c, scores);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

extern "C" int gstex_chart_dims(int n, const float* scales_raw, float sigma_factor, float texel_size, int32_t* texture_dims,
                                int32_t* hw, float* mappings, void* stream) {
    if (n < 0) return GSTEX_E_BADARG;
    if (n == 0) return GSTEX_OK;
    if (!scales_raw || !texture_dims || !hw || !(texel_size > 0.f)) return GSTEX_E_BADARG;
    chart_dims_kernel<<<div_up(n, kThreads1D), kThreads1D, 0, (cudaStream_t)stream>>>(
        n, scales_raw, sigma_factor, (float)(2.0 * (double)sigma_factor), texel_size, texture_dims, hw,
        reinterpret_cast<float2*>(mappings));
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

extern "C" int gstex_chart_starts(int n, const int32_t* cum_hw, const int32_t* hw, int32_t* texture_dims, void* stream) {
    if (n < 0) return GSTEX_E_BADARG;
    if (n == 0) return GSTEX_OK;
    if (!cum_hw || !hw || !texture_dims) return GSTEX_E_BADARG;
    chart_starts_kernel<<<div_up(n, kThreads1D), kThreads1D, 0, (cudaStream_t)stream>>>(n, cum_hw, hw, texture_dims);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

extern "C" int gstex_texture_resample(long long num_new_texels, int n, int channels, const int32_t* new_dims,
                                      const int32_t* old_dims, const float* old_texture, float* new_texture, void* stream) {
    if (num_new_texels < 0 || n < 0 || channels < 1) return GSTEX_E_BADARG;
    if (num_new_texels == 0) return GSTEX_OK;
    if (n == 0 || !new_dims || !old_dims || !old_texture || !new_texture) return GSTEX_E_BADARG;
    const long long blocks = (num_new_texels + kThreads1D - 1) / kThreads1D;
    if (blocks > 0x7fffffffll) return GSTEX_E_BADARG;
    texture_resample_kernel<<<(unsigned)blocks, kThreads1D, 0, (cudaStream_t)stream>>>(num_new_texels, n, channels, new_dims,
                                                                                   old_dims, old_texture, new_texture);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

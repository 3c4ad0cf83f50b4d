// texture_sample.cu -- bilinear sampling of a jagged texture at arbitrary (chart, uv) queries: the kernel behind the
// reference's re-charting step (NS/models/jagged_texture.py:116-143 resamples every texel of
// the new charts from the old ones).  Replaces texture_sample_forward (gstex_cuda/cuda/csrc/texture_sample.cu:11-36,
// wrapper :77-110; helpers texture_helpers.cuh:155-237; Python gstex_cuda/texture_sample.py:13-86).  sm_100a.
//
// A pure gather: per query 20 bytes in (dims, uv), 4 corner texels of C floats, C floats out.  The reference gives a
// thread one query and loops over the channels, recomputing indices and weights per channel.  Here too a thread owns
// a query (a (query, channel)-per-thread layout was measured 1.6x slower at C = 3: it triples the per-query loads
// and index arithmetic), but the channel-independent index/weight arithmetic -- the rasterizer's own tex_query --
// runs once, the streamed inputs/outputs bypass L1 (ld.cs / st.cs) so the cache keeps texels, and three channels
// (twelve corner loads) are in flight at a time.
// uv is clamped to [0,1]^2 first (the documented contract, texture_sample.py:29-31, and what the torch twin does,
// _torch_impl.py:152-153); the reference kernel does not clamp and reads outside the chart for such queries.
#include "common.cuh"

namespace gstex {

// one thread per query; channels in groups of three (twelve independent corner loads)
__global__ void __launch_bounds__(kThreads1D)
texture_sample_kernel(int num_queries, int C, const int32_t* __restrict__ texture_dims, const float2* __restrict__ uvs,
                      const float* __restrict__ texture, float* __restrict__ out) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= num_queries) return;
    const int32_t* d = texture_dims + 3 * (size_t)q;
    const int h = __ldcs(d), w = __ldcs(d + 1), start = __ldcs(d + 2);
    const float2 uv = __ldcs(uvs + q);
    const TexQuery tq = tex_query(h, w, start, clamp01(uv.x), clamp01(uv.y), true);
    const float* t00 = texture + (size_t)tq.b00 * C;
    const float* t01 = texture + (size_t)tq.b01 * C;
    const float* t10 = texture + (size_t)tq.b10 * C;
    const float* t11 = texture + (size_t)tq.b11 * C;
    float* o = out + (size_t)q * C;
    int k = 0;
    for (; k + 3 <= C; k += 3) {  // three channels at a time: twelve independent loads in flight
        const float a0 = __ldg(t00 + k), a1 = __ldg(t00 + k + 1), a2 = __ldg(t00 + k + 2);
        const float b0 = __ldg(t01 + k), b1 = __ldg(t01 + k + 1), b2 = __ldg(t01 + k + 2);
        const float c0 = __ldg(t10 + k), c1 = __ldg(t10 + k + 1), c2 = __ldg(t10 + k + 2);
        const float d0 = __ldg(t11 + k), d1 = __ldg(t11 + k + 1), d2 = __ldg(t11 + k + 2);
        __stcs(o + k, tq.w00 * a0 + tq.w01 * b0 + tq.w10 * c0 + tq.w11 * d0);
        __stcs(o + k + 1, tq.w00 * a1 + tq.w01 * b1 + tq.w10 * c1 + tq.w11 * d1);
        __stcs(o + k + 2, tq.w00 * a2 + tq.w01 * b2 + tq.w10 * c2 + tq.w11 * d2);
    }
    for (; k < C; ++k)
        __stcs(o + k, tq.w00 * __ldg(t00 + k) + tq.w01 * __ldg(t01 + k) + tq.w10 * __ldg(t10 + k) + tq.w11 * __ldg(t11 + k));
}

}  // namespace gstex

using namespace gstex;

extern "C" int gstex_texture_sample_forward(int num_queries, int channels, const int32_t* texture_dims, const float* uvs,
                                            const float* texture, float* out, void* stream) {
    if (num_queries < 0 || channels < 1) return GSTEX_E_BADARG;
    if (num_queries == 0) return GSTEX_OK;
    if (!texture_dims || !uvs || !texture || !out) return GSTEX_E_BADARG;
    texture_sample_kernel<<<div_up(num_queries, kThreads1D), kThreads1D, 0, (cudaStream_t)stream>>>(
        num_queries, channels, texture_dims, reinterpret_cast<const float2*>(uvs), texture, out);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

// bwd_stream.cuh -- interface between the backward entry point (raster_bwd.cu) and the stream kernel.
#pragma once
#include "pair_stream.cuh"
#include "raster_common.cuh"

namespace gstex {

struct StreamBwdParams {
    RasterGeom g;
    int C;
    const SplatRec* recs;
    const SplatRecBwd* recs_bwd;
    const float* texture;  // unused by the math (samples are recorded); kept for symmetry
    const float *viewmat, *c2w, *background;
    float fx, fy, cx, cy;
    int settings;
    const float *final_Ts, *final_s;
    const int32_t* depth_idx;
    const float *v_output, *v_output_depth, *v_output_reg, *v_output_alpha, *v_output_texture, *v_output_normal;
    float *v_colors, *v_opacity, *v_means, *v_scales, *v_quats, *v_texture;
    StreamView sv;
    // padded accumulators (grad scratch): texel gradients as (T,4), splat gradients as (N,16) floats:
    // [colors3 opacity | means3 - | scales.xy - - | quats4]
    // The scratch is all-zero between calls: the unpack kernel clears exactly what the stream kernel wrote.
    float4 *tex_acc, *splat_acc;
    int n;
    long long num_texels;
    const int32_t* order;         // the forward's tile order (longest list first) and its validity flag
    const uint32_t* order_valid;  // (workspace counters[1]); units are taken in that order when set
    float tex_scale;    // slope of the forward's texel activation: the unpack sweep multiplies texel gradients by it
    int scratch_clean;  // the caller vouches the scratch is as a previous call left it: skip the clear
    int defer_unpack;   // leave the sums in the scratch (GSTEX_DEFER_UNPACK): gstex_unpack_grads moves them later
};

static inline size_t grad_scratch_bytes(int n, long long num_texels) {
    return ((size_t)(num_texels > 0 ? num_texels : 0) + 4 * (size_t)(n > 0 ? n : 0)) * sizeof(float4);
}

// Enqueues the stream backward; the kernel returns at once when the forward flagged a pool overflow.
int launch_backward_stream(const StreamBwdParams& p, cudaStream_t st);
// The unpack sweep alone: adds every non-zero accumulator word to the gradient tensors and clears it.
int launch_unpack_grads(float4* tex_acc, long long num_texels, int C, float tex_scale, float4* splat_acc, int n,
                        float* v_texture, float* v_colors, float* v_opacity, float* v_means, float* v_scales,
                        float* v_quats, cudaStream_t st);

}  // namespace gstex

// raster_fwd.cuh -- parameters shared by the two forward compositing kernels (raster_fwd.cu: general kernel, any
// block width / channel count / viewer bit; raster_fwd_dense.cu: the lane-dense training kernel).
#pragma once
#include "pair_stream.cuh"
#include "pipeline.cuh"

namespace gstex {

struct FwdParams {
    RasterGeom g;
    int C;
    const SplatRec* recs;
    const float4* boxes;
    const int32_t* ids;
    const int2* tile_bins;
    const float* texture;
    const float4* texture4;  // optional (T,4) zero-padded copy of a <= 3-channel texture: 16-byte texel fetches (else nullptr)
    const float *viewmat, *c2w, *background;
    float fx, fy, cx, cy;
    int settings;
    float *out_img, *out_depth, *out_reg, *out_texture, *out_normal, *final_Ts, *out_reg_s;
    int32_t *final_idx, *depth_idx;
    uint32_t* unit_cursor;  // zeroed before the launch
    const int32_t* order;   // tiles, longest list first (nullptr: natural order)
    float tex_scale, tex_bias;  // every fetched sample is tex_scale * sample + tex_bias (1, 0: the reference's call)
    int units_per_tile, num_units;
    int force_nocull;  // tests (gstex_debug_force_nocull): evaluate every list entry for every unit
    int dbg;           // tests (gstex_debug_forward_variant)
    StreamView sv;  // pair stream (REC kernels only)
};

// Lane-dense forward (raster_fwd_dense.cu): 16x16 tiles, <= 3 texture channels, no viewer bits.
bool forward_dense_supported(const FwdParams& p);
int launch_forward_dense(const FwdParams& p, bool record, cudaStream_t st);

}  // namespace gstex

/*
 * gstex_b200.h -- C ABI of the B200-native GStex rasterizer hot path (libgstex_b200.so).
 *
 * One entry point per native function the reference's Python binds for this path (the pybind11
 * module `_C` of nvnhat95/GStex, submodules/GStex_cuda/gstex_cuda/cuda/csrc/ext.cpp:8-24) plus the
 * three torch library calls the reference makes in between (cumsum, sort, gather).  Conventions:
 *   - every pointer is a DEVICE pointer into memory owned by the caller (PyTorch's caching
 *     allocator in the shipped host layer); nothing is allocated or freed behind the caller's back,
 *     scratch space is passed in as `workspace` and sized by the matching *_workspace_bytes();
 *   - tensors are contiguous row-major, fp32 / int32 / int64 exactly as in the reference;
 *   - every call enqueues work on `stream` (a cudaStream_t passed as void*) and returns without
 *     synchronising; the return value is 0 on success, a negative GSTEX_E_* argument error, or a
 *     positive cudaError_t from the launch (the reference checks neither: csrc/config.h:6-17 is
 *     never used);
 *   - re-entrant, no global state.
 * "ref:" comments name the reference interface each function replaces; paths are relative to
 * /root/reference/submodules/GStex_cuda/gstex_cuda/.
 */
#ifndef GSTEX_B200_H
#define GSTEX_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSTEX_B200_ABI_VERSION 2

#define GSTEX_OK 0
#define GSTEX_E_BADARG (-1)      /* null pointer, negative size, block_width outside [2,16] ... */
#define GSTEX_E_UNSUPPORTED (-2) /* settings bits 16-29 (viewer visualisation), channels > 16 */
#define GSTEX_E_WORKSPACE (-3)   /* workspace too small */
#define GSTEX_E_TOO_LARGE (-4)   /* size beyond the fixed capacity of a no-allocation entry point */

/* settings bit-field, ref: cuda/csrc/texture.cu:54-65,381-387 */
#define GSTEX_SET_NEAREST (1 << 2)
#define GSTEX_SET_UV_GRAD (1 << 8)
#define GSTEX_SET_BLUR (1 << 9)
#define GSTEX_SET_NDC_REG (1 << 10)
/* viewer "visualization_mode" (texture.cu:58,236-241): out_normal accumulates the splat normal flipped towards the camera.
 * Forward only -- the reference's backward never reads the bit (texture.cu:381-387) and neither does ours. */
#define GSTEX_SET_VIS_NORMALS (1 << 15)
#define GSTEX_SETTINGS_SUPPORTED \
    (GSTEX_SET_NEAREST | GSTEX_SET_UV_GRAD | GSTEX_SET_BLUR | GSTEX_SET_NDC_REG | GSTEX_SET_VIS_NORMALS)
#define GSTEX_MAX_CHANNELS 16

int gstex_abi_version(void);
/* Human-readable text for a return code of this library (static storage). */
const char* gstex_error_string(int code);
/* Number of CUDA kernels this library has launched in this process so far (statistics only). */
long long gstex_launch_count(void);

/* TEST HOOK: non-zero makes the compositing kernels (forward, texture_edit) evaluate every list entry for every pixel
 * block instead of culling by support box, so that tests can assert culling never changes a result.  Process-wide;
 * returns the previous value.  Not for production use (no environment variable is ever read by this library). */
int gstex_debug_force_nocull(int on);
/* TEST / TUNING HOOK: 2 selects the lane-dense forward kernel (raster_fwd_dense.cu: pair-parallel alpha evaluation;
 * 16x16 tiles, <= 3 channels, no viewer bit) where it applies, 0 the default kernel (raster_fwd.cu); the two agree bit
 * for bit (tests/test_gpu_raster.py).  4 = dense + resume-path statistics.  Returns the previous value. */
int gstex_debug_forward_variant(int v);

/* ---- stage (a): per-splat projection, bounds, tile count ------------------------------------- */

/* ref: get_aabb_2d.py:22-32 project_points(points, viewmat, intrins, clip=False).
 * viewmat: 12 floats, row-major [R|t].  xys (N,2) and/or depths (N) may be NULL. */
int gstex_project_points(const float* means, const float* viewmat, float fx, float fy, float cx, float cy, int n,
                         float* xys, float* depths, void* stream);

/* ref: cuda/csrc/get_aabb_2d.cu:91-126 get_aabb_2d_tensor -> kernel :11-89.  centers, extents: (N,2). */
int gstex_get_aabb_2d(const float* means, const float* scales, float glob_scale, const float* quats,
                      const float* viewmat, float fx, float fy, float cx, float cy, int n, float* centers,
                      float* extents, void* stream);

/* ref: get_aabb_2d.py:70-91 get_num_tiles_hit_2d(centers, extents, H, W, block_width) -> (N,) int32. */
int gstex_get_num_tiles_hit_2d(const float* centers, const float* extents, int n, int img_height, int img_width,
                               int block_width, int32_t* num_tiles_hit, void* stream);

/* The three calls above fused into one kernel (what NS/models/gstex.py:1077-1080 asks for every step).
 * Any output pointer may be NULL. */
int gstex_project_bounds(const float* means, const float* scales, float glob_scale, const float* quats,
                         const float* viewmat, float fx, float fy, float cx, float cy, int n, int img_height,
                         int img_width, int block_width, float* xys, float* depths, float* centers, float* extents,
                         int32_t* num_tiles_hit, void* stream);

/* ---- stage (b): prefix sum, key emit, radix sort, tile ranges --------------------------------- */

/* ref: utils.py:57 torch.cumsum(num_tiles_hit, dim=0, dtype=int32): inclusive scan.  `total`, if not
 * NULL, receives out[n-1] (device int32; 0 when n == 0).  workspace: gstex_cumsum_workspace_bytes(n). */
size_t gstex_cumsum_workspace_bytes(int n);
int gstex_cumsum_i32(const int32_t* in, int n, int32_t* out, int32_t* total, void* workspace, size_t workspace_bytes,
                     void* stream);

/* No counterpart in the reference.  gstex_cumsum_i32 for callers that
 * cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): the same
 * kernels and the same results, but the scratch (per-tile totals) lives in a static 64 KB device array inside the
 * library instead of a caller-provided workspace, and the call makes no cudaMalloc, cudaFree or allocating library call.
 * Limits and caveats: n <= gstex_cumsum_noalloc_max_n() = 67 108 864, larger n returns GSTEX_E_TOO_LARGE and launches
 * nothing; every GPU that loads the library holds the 64 KB until exit; ALL calls on a device share the one array, so
 * two calls that overlap on the GPU (different streams, or graph branches without an edge between them) race on it --
 * order them.  The array keeps the last call's contents; nothing needs clearing (every slot a call reads it has
 * written first).  in / out / total are the caller's, as in gstex_cumsum_i32. */
int gstex_cumsum_noalloc_max_n(void);
int gstex_cumsum_i32_noalloc(const int32_t* in, int n, int32_t* out, int32_t* total, void* stream);

/* ref: cuda/csrc/bindings.cu:77-121 map_gaussian_to_intersects_tensor -> kernel forward.cu:13-71
 * (wrapped=false).  Zero-fills isect_ids (I int64) and gaussian_ids (I int32) first, as the reference's
 * torch::zeros does, so the reserved-but-unwritten slots of clipped splats keep key 0 / id 0. */
int gstex_map_gaussian_to_intersects(int n, int num_intersects, const float* centers, const float* extents,
                                     const float* depths, const int32_t* cum_tiles_hit, int tiles_x, int tiles_y,
                                     int block_width, int64_t* isect_ids, int32_t* gaussian_ids, void* stream);

/* ref: utils.py:159-160 torch.sort(isect_ids) + torch.gather(gaussian_ids, 0, idx): stable ascending
 * LSD radix sort of the (tile<<32 | depth bits) keys carrying the splat id as payload.  Only key bits
 * [0, key_bits) are examined (32 + bits needed for tiles_x*tiles_y - 1; pass 64 to sort on everything);
 * keys must be non-negative.  Inputs are not modified. */
size_t gstex_sort_workspace_bytes(int num_intersects);
int gstex_sort_isects(int num_intersects, const int64_t* isect_ids, const int32_t* gaussian_ids,
                      int64_t* isect_ids_sorted, int32_t* gaussian_ids_sorted, int key_bits, void* workspace,
                      size_t workspace_bytes, void* stream);

/* ref: cuda/csrc/bindings.cu:123-141 get_tile_bin_edges_tensor -> kernel forward.cu:76-98.
 * Zero-fills tile_bins (num_tiles,2) int32 first. */
int gstex_get_tile_bin_edges(int num_intersects, const int64_t* isect_ids_sorted, int num_tiles, int32_t* tile_bins,
                             void* stream);

/* Fast path of stage (b) used by texture_gaussians (no counterpart in the reference, which always sorts the
 * whole list, utils.py:106-162): tile-bucketed binning.  Results are bit-identical to
 * map_gaussian_to_intersects -> gstex_sort_isects -> gstex_get_tile_bin_edges, but only
 * gaussian_ids_sorted and tile_bins are produced.
 *   gstex_tile_bucket_count: per-tile intersection counts (block_counts: gstex_tile_bucket_scratch_bytes(
 *     num_tiles) of scratch, kept for the sort call), their exclusive scan -> tile_bins (num_tiles,2) with (0,0)
 *     for empty tiles, tile_start (num_tiles int32), stats[0] = num_intersects, stats[1] = longest tile list.
 *     GSTEX_E_UNSUPPORTED when tiles_x*tiles_y exceeds gstex_tile_bucket_max_tiles().
 *   gstex_tile_bucket_sort: claims a slot per intersection in its tile's range (keys: num_intersects x 8 B
 *     scratch) and sorts every tile's range by (depth bits, splat id) in shared memory.  max_list = stats[1]
 *     read back by the host; returns GSTEX_E_UNSUPPORTED when it exceeds gstex_tile_bucket_max_list() (the
 *     caller then takes the general path above).  Depths of emitted splats must be non-negative (they are:
 *     splats with z <= 0.01 are clipped and emit nothing, get_aabb_2d.cu:25).
 *   Both take num_tiles_hit (stage a3's per-splat counts) as it is: the path needs no prefix sum over splats. */
int gstex_tile_bucket_max_list(void);
int gstex_tile_bucket_max_tiles(void);
size_t gstex_tile_bucket_scratch_bytes(int num_tiles);
int gstex_tile_bucket_count(int n, const float* centers, const float* extents, const int32_t* num_tiles_hit,
                            int tiles_x, int tiles_y, int block_width, int32_t* block_counts, int32_t* tile_bins,
                            int32_t* tile_start, int32_t* stats, void* stream);
int gstex_tile_bucket_sort(int n, int num_intersects, int max_list, const float* centers, const float* extents,
                           const float* depths, const int32_t* num_tiles_hit, int tiles_x, int tiles_y,
                           int block_width, const int32_t* block_counts, const int32_t* tile_start,
                           const int32_t* tile_bins, void* keys, int32_t* gaussian_ids_sorted, void* stream);

/* ---- stages (c), (d): textured forward compositing and backward -------------------------------- */

/* ref: cuda/csrc/texture.cu:762-901 texture_forward_tensor -> kernel :11-329.
 * Inputs as in the reference (texture_dims (N,3) int32 {h,w,start}; uv0 (N,1,2); umap, vmap (N,1,3);
 * texture (T,C); viewmat >= 12 floats; c2w >= 12 floats; background 3 floats).  All nine outputs are
 * written for every pixel (no zero-fill needed):
 *   out_img (H,W,3) out_depth (H,W) out_reg (H,W) out_texture (H,W,C) out_normal (H,W,3)
 *   final_Ts (H,W) final_idx (H,W) int32 depth_idx (H,W) int32 out_reg_s (H,W,3).
 * workspace: gstex_texture_workspace_bytes(n) (per-view packed splat records).
 *
 * pair_stream (optional, may be NULL): a device buffer of pair_stream_bytes in which the forward records
 * one 32-byte item per contributing (pixel, splat) pair, so that gstex_texture_backward can run its
 * lane-dense stream kernel instead of re-walking the tile lists (the reference has no counterpart: it
 * always recomputes, texture.cu:455-759).  Size it with gstex_pair_stream_bytes(H, W, block_width,
 * max_pairs); if the pool is too small the forward flags an overflow inside the buffer and the backward
 * transparently recomputes -- results never depend on the pool.  Recording is used when
 * gstex_pair_stream_supported(channels, n, settings) is non-zero (<= 3 channels, no uv gradients). */
size_t gstex_texture_workspace_bytes(int n);
size_t gstex_pair_stream_bytes(int img_height, int img_width, int block_width, long long max_pairs);
int gstex_pair_stream_supported(int channels, int n, int settings);
int gstex_texture_forward(int img_height, int img_width, int block_width, int channels, int n, int num_intersects,
                          const int32_t* texture_dims, const int32_t* gaussian_ids_sorted, const int32_t* tile_bins,
                          const float* colors, const float* opacities, const float* means, const float* scales,
                          float glob_scale, const float* quats, const float* uv0, const float* umap,
                          const float* vmap, const float* texture, const float* viewmat, const float* c2w, float fx,
                          float fy, float cx, float cy, int settings, const float* background, float* out_img,
                          float* out_depth, float* out_reg, float* out_texture, float* out_normal, float* final_Ts,
                          int32_t* final_idx, int32_t* depth_idx, float* out_reg_s, void* workspace,
                          size_t workspace_bytes, void* pair_stream, size_t pair_stream_bytes, void* stream);

/* ref: cuda/csrc/texture.cu:903-1053 texture_backward_tensor -> kernel :331-760.
 * The nine gradient outputs (shapes of the inputs they belong to) must be ZERO-FILLED by the caller
 * (the reference allocates them with torch::zeros); gradients are accumulated into them.
 * `workspace` may be the one the forward of the same view filled (records_valid != 0 skips re-packing).
 * `pair_stream`: the buffer the forward of the same view recorded into, or NULL to recompute.
 * `grad_scratch` (gstex_grad_scratch_bytes(n, num_texels) bytes, 16-byte aligned; needed by the stream path,
 * may be NULL otherwise): padded accumulators the stream kernel reduces into with 16-byte vector atomics; an
 * unpack kernel then adds every non-zero accumulator word to the reference's gradient tensors and clears it
 * again, so the scratch LEAVES every call all-zero.
 * num_texels = rows of `texture`.
 * Any of the six upstream gradients v_output* may be NULL = all zero (autograd's None for an output the loss does
 * not use -- the reference's trainer uses three of the six): nothing is allocated, cleared or read for it.
 * zero_fill, low two bits -- 1: the nine gradient buffers are uninitialised on entry and the library clears them
 * first; 0: they are zero-filled by the caller; 2: they hold running sums (a multi-view step's accumulated
 * gradients) and this view's gradients are ADDED to them -- what autograd's AccumulateGrad would do in a separate
 * pass over all nine tensors.  OR-ing in GSTEX_SCRATCH_CLEAN vouches that `grad_scratch` is exactly as an earlier
 * call of this function left it (or was zero-filled by the caller) and skips the clear of the scratch.
 * OR-ing in GSTEX_DEFER_UNPACK (only with zero_fill = 2, multi-view steps): the stream path leaves this view's sums in
 * `grad_scratch` instead of sweeping all T texel accumulators after every view; the scratch keeps accumulating over
 * the views of the step (pass it back with GSTEX_SCRATCH_CLEAN) and ONE gstex_unpack_grads call at the end of the
 * step adds the sums to the gradient tensors and hands the scratch back all-zero.  Views that fall back to the
 * recompute kernel (pool overflow) add to the gradient tensors directly, as without the flag. */
#define GSTEX_SCRATCH_CLEAN 4
#define GSTEX_DEFER_UNPACK 8
size_t gstex_grad_scratch_bytes(int n, long long num_texels);
/* The unpack sweep on its own (see GSTEX_DEFER_UNPACK): for every non-zero accumulator word of `grad_scratch`,
 * v_texture += tex_scale * word (channels <= 3 of each texel) / v_colors, v_opacity, v_means, v_scales.xy, v_quats +=
 * word, and the word is cleared.  v_quats must be 16-byte aligned. */
int gstex_unpack_grads(int channels, int n, long long num_texels, float tex_scale, void* grad_scratch,
                       size_t grad_scratch_size, float* v_colors, float* v_opacity, float* v_means, float* v_scales,
                       float* v_quats, float* v_texture, void* stream);
int gstex_texture_backward(int img_height, int img_width, int block_width, int channels, int n, int num_intersects,
                           const int32_t* texture_dims, const int32_t* gaussian_ids_sorted, const int32_t* tile_bins,
                           const float* colors, const float* opacities, const float* means, const float* scales,
                           float glob_scale, const float* quats, const float* uv0, const float* umap,
                           const float* vmap, const float* texture, const float* viewmat, const float* c2w, float fx,
                           float fy, float cx, float cy, int settings, const float* background,
                           const float* final_Ts, const int32_t* final_idx, const int32_t* depth_idx,
                           const float* final_s, const float* v_output, const float* v_output_depth,
                           const float* v_output_reg, const float* v_output_alpha, const float* v_output_texture,
                           const float* v_output_normal, float* v_colors, float* v_opacity, float* v_means,
                           float* v_scales, float* v_quats, float* v_uv0, float* v_umap, float* v_vmap,
                           float* v_texture, void* workspace, size_t workspace_bytes, int records_valid,
                           void* pair_stream, size_t pair_stream_bytes, void* grad_scratch, size_t grad_scratch_size,
                           long long num_texels, int zero_fill, void* stream);
/* Next row, the texel half of the caller's colour stage: the reference trainer
 * builds `texture = SH2RGB(texture_dc)` = 0.28209479 * texture_dc + 0.5 over all T texels before every call
 * (NS/models/gstex.py:1091-1121, two passes over T*C floats forward and two more in autograd's backward).  The
 * `_act` entry points take the raw texels and apply `tex_scale * sample + tex_bias` to every fetched bilinear
 * sample instead (exact for an affine map: the four weights sum to one); the backward returns the gradient with
 * respect to the RAW texels (times tex_scale, applied where the texel gradients are unpacked).  (1, 0) is the plain
 * entry point, bit for bit. */
int gstex_texture_forward_act(int img_height, int img_width, int block_width, int channels, int n, int num_intersects,
                          const int32_t* texture_dims, const int32_t* gaussian_ids_sorted, const int32_t* tile_bins,
                          const float* colors, const float* opacities, const float* means, const float* scales,
                          float glob_scale, const float* quats, const float* uv0, const float* umap,
                          const float* vmap, const float* texture, float tex_scale, float tex_bias, const float* viewmat, const float* c2w, float fx,
                          float fy, float cx, float cy, int settings, const float* background, float* out_img,
                          float* out_depth, float* out_reg, float* out_texture, float* out_normal, float* final_Ts,
                          int32_t* final_idx, int32_t* depth_idx, float* out_reg_s, void* workspace,
                          size_t workspace_bytes, void* pair_stream, size_t pair_stream_bytes, void* stream);
/* gstex_texture_forward_act with the texel fetch reading `texture4`, a (num_texels, 4) zero-padded, 16-byte aligned
 * copy of the (num_texels, channels <= 3) `texture` written by gstex_texture_pad4: one 128-bit load per bilinear
 * corner instead of `channels` scalar loads (the "128-bit texel loads" of BASELINE.json's north star).  The
 * arithmetic and its inputs are the same, so every output is bit-identical to gstex_texture_forward_act's.  The caller
 * owns the copy and refreshes it whenever `texture` changes (once per optimiser step for a multi-view batch); the
 * backward takes the unpadded texture.  No counterpart in the reference, whose kernels query the chart once per channel,
 * four scalar loads each (texture.cu:260-268, query_jagged_chart). */
int gstex_texture_forward_padded(int img_height, int img_width, int block_width, int channels, int n, int num_intersects,
                          const int32_t* texture_dims, const int32_t* gaussian_ids_sorted, const int32_t* tile_bins,
                          const float* colors, const float* opacities, const float* means, const float* scales,
                          float glob_scale, const float* quats, const float* uv0, const float* umap,
                          const float* vmap, const float* texture, const float* texture4, float tex_scale,
                          float tex_bias, const float* viewmat, const float* c2w, float fx,
                          float fy, float cx, float cy, int settings, const float* background, float* out_img,
                          float* out_depth, float* out_reg, float* out_texture, float* out_normal, float* final_Ts,
                          int32_t* final_idx, int32_t* depth_idx, float* out_reg_s, void* workspace,
                          size_t workspace_bytes, void* pair_stream, size_t pair_stream_bytes, void* stream);
/* The packing stage of gstex_texture_forward* (the per-view 128-byte records, support boxes and tile order the
 * compositing kernels read, written into `workspace`) as a call of its own, and the forward that takes them as given.
 * A multi-view caller issues the pack of view i+1 on another stream while view i's compositing kernels run: the pack is
 * an HBM stream over all splats, the compositing kernels are issue-bound, so the two overlap almost for free.  The
 * inputs of the two calls (and the workspace) must be the same and unchanged in between; the outputs are bit-identical
 * to gstex_texture_forward_padded's (`texture4` may be NULL here: scalar texel loads).  The backward takes the same
 * workspace with records_valid = 1, as after any forward.  No counterpart in the reference, which recomputes the
 * per-splat terms in every pixel thread (texture.cu:129-200). */
int gstex_texture_pack_records(int img_height, int img_width, int block_width, int n, const int32_t* texture_dims,
                               const int32_t* tile_bins, const float* colors, const float* opacities, const float* means,
                               const float* scales, float glob_scale, const float* quats, const float* uv0,
                               const float* umap, const float* vmap, const float* viewmat, const float* c2w, float fx,
                               float fy, float cx, float cy, int settings, void* workspace, size_t workspace_bytes,
                               void* stream);
int gstex_texture_forward_packed(int img_height, int img_width, int block_width, int channels, int n, int num_intersects,
                          const int32_t* texture_dims, const int32_t* gaussian_ids_sorted, const int32_t* tile_bins,
                          const float* colors, const float* opacities, const float* means, const float* scales,
                          float glob_scale, const float* quats, const float* uv0, const float* umap,
                          const float* vmap, const float* texture, const float* texture4, float tex_scale,
                          float tex_bias, const float* viewmat, const float* c2w, float fx,
                          float fy, float cx, float cy, int settings, const float* background, float* out_img,
                          float* out_depth, float* out_reg, float* out_texture, float* out_normal, float* final_Ts,
                          int32_t* final_idx, int32_t* depth_idx, float* out_reg_s, void* workspace,
                          size_t workspace_bytes, void* pair_stream, size_t pair_stream_bytes, void* stream);
/* texture4[t] = (texture[t][0], texture[t][1] or 0, texture[t][2] or 0, 0) for channels in 1..3. */
int gstex_texture_pad4(long long num_texels, int channels, const float* texture, float* texture4, void* stream);
int gstex_texture_backward_act(int img_height, int img_width, int block_width, int channels, int n, int num_intersects,
                           const int32_t* texture_dims, const int32_t* gaussian_ids_sorted, const int32_t* tile_bins,
                           const float* colors, const float* opacities, const float* means, const float* scales,
                           float glob_scale, const float* quats, const float* uv0, const float* umap,
                           const float* vmap, const float* texture, float tex_scale, float tex_bias, const float* viewmat, const float* c2w, float fx,
                           float fy, float cx, float cy, int settings, const float* background,
                           const float* final_Ts, const int32_t* final_idx, const int32_t* depth_idx,
                           const float* final_s, const float* v_output, const float* v_output_depth,
                           const float* v_output_reg, const float* v_output_alpha, const float* v_output_texture,
                           const float* v_output_normal, float* v_colors, float* v_opacity, float* v_means,
                           float* v_scales, float* v_quats, float* v_uv0, float* v_umap, float* v_vmap,
                           float* v_texture, void* workspace, size_t workspace_bytes, int records_valid,
                           void* pair_stream, size_t pair_stream_bytes, void* grad_scratch, size_t grad_scratch_size,
                           long long num_texels, int zero_fill, void* stream);

/* ---- next row: per-splat colour from spherical harmonics -------------------- */

/* ref: cuda/csrc/bindings.cu:18-43 compute_sh_forward_tensor -> sh.cuh:217-235 (sh_coeffs_to_color :47-119).
 * viewdirs (n,3) need not be normalised; coeffs (n, num_bases(degree), 3); colors (n,3), every element written.
 * degree in 0..4 fixes the row length (1/4/9/16/25 bases), degrees_to_use <= degree selects the bases summed. */
int gstex_compute_sh_forward(int n, int degree, int degrees_to_use, const float* viewdirs, const float* coeffs,
                             float* colors, void* stream);

/* ref: cuda/csrc/bindings.cu:45-75 compute_sh_backward_tensor -> sh.cuh:237-253 (sh_coeffs_to_color_vjp :121-215).
 * v_coeffs (n, num_bases(degree), 3): every element written (zeros above the used bases, as the reference's
 * torch::zeros leaves them).  No gradient with respect to viewdirs (sh.py:44-45). */
int gstex_compute_sh_backward(int n, int degree, int degrees_to_use, const float* viewdirs, const float* v_colors,
                              float* v_coeffs, void* stream);

/* ---- next row (the other half): per-splat activations + uv maps ------------------- */

/* ref: /root/reference/nerfstudio/models/gstex.py:1062-1066 (quats = raw / |raw|; scales = clamp(exp(raw[:, :2]), 1e-9)
 * with scales[:, 2] = 1e-5 * mean of those two, detached; opacities = sigmoid(raw)) and get_uv_mapping, :975-990
 * (uv0 = 0.5, umap = mappings[:, 0] * ax1, vmap = mappings[:, 1] * ax2 with ax1, ax2 the first two columns of
 * R(F.normalize(quats)), gstex_cuda/_torch_impl.py:116-138; all three detached) -- ~30 torch launches over N per view in
 * the reference, one kernel here.  quats_raw (n,4), scales_raw (n,3), opacities_raw (n,1), mappings (n,2) ->
 * quats (n,4), scales (n,3), opacities (n,1), uv0 (n,1,2), umap (n,1,3), vmap (n,1,3); every element written.  The three
 * map outputs may all be NULL (then `mappings` is not read).  Forward results are bit-identical to the torch ops on the
 * same GPU.  quats_raw / quats must be 16-byte aligned. */
int gstex_activations_forward(int n, const float* quats_raw, const float* scales_raw, const float* opacities_raw,
                              const float* mappings, float* quats, float* scales, float* opacities, float* uv0,
                              float* umap, float* vmap, void* stream);
/* Gradients with respect to the raw parameters from the gradients of quats / scales / opacities (any of the three
 * upstream pointers may be NULL = zero; any of the three outputs may be NULL = not wanted).  v_scales_raw[:, 2] = 0 and no
 * gradient flows through the uv maps, as in the reference (detached). */
int gstex_activations_backward(int n, const float* quats_raw, const float* scales_raw, const float* quats,
                               const float* scales, const float* opacities, const float* v_quats, const float* v_scales,
                               const float* v_opacities, float* v_quats_raw, float* v_scales_raw, float* v_opacities_raw,
                               void* stream);

/* ---- next row: jagged-texture sampling used by re-charting ------------------- */

/* ref: cuda/csrc/texture_sample.cu:77-110 texture_sample_forward_tensor -> kernel :11-36.
 * texture_dims (Q,3) int32 {h, w, start} of the chart each query reads, uvs (Q,2), texture (T,C) -> out (Q,C),
 * every element written.  uv is clamped to [0,1]^2 (texture_sample.py:29-31).  Forward only: the reference's
 * autograd backward returns no gradients (texture_sample.py:77-86). */
int gstex_texture_sample_forward(int num_queries, int channels, const int32_t* texture_dims, const float* uvs,
                                 const float* texture, float* out, void* stream);

/* ref: /root/reference/nerfstudio/models/gstex.py:841-888 build_charts -- the texel-size bisection and what follows it.
 *   gstex_chart_scores: for up to gstex_chart_max_candidates() = 32 candidate texel sizes x_k (a HOST array) at once,
 *     scores[k] = sum_i ceil(sigma_factor * exp(s_i0) / x_k) * ceil(sigma_factor * exp(s_i1) / x_k)  (get_score, :847-849;
 *     device doubles, exact: every term is an integer).  Bit k of recip_mask: divide the way `tensor / python_float` does
 *     (the reference's first score, :859) instead of `tensor / tensor` (every later one, :868).  scales_raw (n,3) logs.
 *   gstex_chart_dims: texture_dims[:, 0:2] = ceil(sigma_factor * exp(s) / texel_size) (:876-877), hw[i] = h * w,
 *     mappings (n,2) = 1 / (2 * sigma_factor * exp(s)) (:882-883; may be NULL).
 *   gstex_chart_starts: texture_dims[:, 2] = cum_hw - hw (:878-879; cum_hw = gstex_cumsum_i32 of hw). */
int gstex_chart_max_candidates(void);
int gstex_chart_scores(int n, const float* scales_raw, float sigma_factor, int num_candidates,
                       const float* texel_sizes_host, unsigned recip_mask, double* scores, void* stream);
int gstex_chart_dims(int n, const float* scales_raw, float sigma_factor, float texel_size, int32_t* texture_dims,
                     int32_t* hw, float* mappings, void* stream);
int gstex_chart_starts(int n, const int32_t* cum_hw, const int32_t* hw, int32_t* texture_dims, void* stream);
/* ref: /root/reference/nerfstudio/models/jagged_texture.py:116-143 JaggedTexture.init_from_dims (+ texture_dims_to_query,
 * :23-35): new_texture (num_new_texels, C), every texel the bilinear sample of its splat's OLD chart at the texel's
 * integer coordinate over the new chart size.  new_dims / old_dims (n,3) int32 {h, w, start}; nothing per-texel is
 * materialised (the reference builds ids, query_dims and uv tensors: 28 bytes per texel). */
int gstex_texture_resample(long long num_new_texels, int n, int channels, const int32_t* new_dims, const int32_t* old_dims,
                           const float* old_texture, float* new_texture, void* stream);

/* ---- next row: texture editing (splat-painting of a canvas into the charts) ----- */

/* ref: cuda/csrc/texture_edit.cu:238-346 texture_edit_tensor -> kernel :11-236 (bound at cuda/csrc/ext.cpp, called by
 * texture_edit.py:206-236).  Walks every pixel's sorted tile list with the forward's alpha / skip / termination rules
 * and, for each contributing pair whose view depth t * view_depth lies in [depth_lower, depth_upper] of its pixel,
 * adds weight * (r*m, g*m, b*m, m, 1) to channels 0..4 of the four bilinear corner texels of the pair's uv
 * (r,g,b = updated_img (H,W,3), m = updated_alpha (H,W,1)).  `channels` = texture_info.z is the row stride of
 * updated_texture (total_texels, channels), >= 5 (the model passes 5, NS/models/gstex.py:572); the call zero-fills
 * it first (the reference's torch::zeros, :308-310), so every element is written.  `settings`: bit 0 = blur floor
 * (:48; NOT the forward's bit 9), all other bits ignored as in the reference.  The reference's `background` and
 * `texture` arguments are never read by its kernel and are not part of this signature.  `workspace`:
 * gstex_texture_workspace_bytes(n) bytes of scratch (per-view packed records). */
int gstex_texture_edit(int img_height, int img_width, int block_width, int channels, int n, int num_intersects,
                       long long total_texels, const int32_t* texture_dims, const float* updated_img,
                       const float* updated_alpha, const float* depth_lower, const float* depth_upper,
                       const int32_t* gaussian_ids_sorted, const int32_t* tile_bins, const float* opacities,
                       const float* means, const float* scales, float glob_scale, const float* quats,
                       const float* uv0, const float* umap, const float* vmap, const float* viewmat,
                       const float* c2w, float fx, float fy, float cx, float cy, int settings,
                       float* updated_texture, void* workspace, size_t workspace_bytes, void* stream);

/* ---- next row: compositing + photometric loss on the far side of the path ------- */

/* ref: nerfstudio/models/gstex.py:1204-1205 (rgb = clamp(out_img + out_texture[:,:,0:3] + (1 - out_alpha) * background,
 * 0, 1)), :1297-1302 (optional mask on both images; Ll1 = mean |gt - rgb|; SSIM(gt, rgb)) and :1319 (main_loss =
 * (1 - ssim_lambda) * Ll1 + ssim_lambda * (1 - SSIM)).  SSIM = pytorch_msssim.SSIM(data_range=1, size_average=True,
 * channel=3) (gstex.py:33,351; third-party, unpinned in pyproject.toml:60): win_size taps of a Gaussian (win_sigma),
 * no padding, K = (0.01, 0.03); only win_size == 11 is built, img_height and img_width must be >= 11.
 * out_img (H,W,3), out_texture (H,W,C>=3), out_alpha (H,W), background (3), gt_img (H,W,3) already composited with
 * the background, mask (H,W) or NULL.  Outputs: rgb (H,W,3); losses[3] = {Ll1, SSIM, main_loss}; ssim_maps
 * (3,H,W,3) partial derivatives kept for the backward, or NULL when no gradient is wanted.  scratch: per-block
 * partial sums, gstex_photometric_loss_scratch_bytes(H, W) bytes; they are added in a fixed order (deterministic). */
size_t gstex_photometric_loss_scratch_bytes(int img_height, int img_width);
int gstex_photometric_loss_forward(int img_height, int img_width, int channels, const float* out_img,
                                   const float* out_texture, const float* out_alpha, const float* background,
                                   const float* gt_img, const float* mask, float ssim_lambda, int win_size,
                                   float win_sigma, float* rgb, float* losses, float* ssim_maps, void* scratch,
                                   size_t scratch_bytes, void* stream);
/* Gradients of main_loss (times *v_main_loss, a DEVICE scalar; NULL = 1) with respect to the three rendered layers:
 * v_out_img (H,W,3), v_out_texture (H,W,C; channels >= 3 zero), v_out_alpha (H,W); every element written.  This is
 * what autograd produces from gstex.py:1204-1205,1301-1319 and hands to _TextureGaussians.backward. */
int gstex_photometric_loss_backward(int img_height, int img_width, int channels, const float* out_img,
                                    const float* out_texture, const float* out_alpha, const float* background,
                                    const float* gt_img, const float* mask, float ssim_lambda, int win_size,
                                    float win_sigma, const float* ssim_maps, const float* v_main_loss, float* v_out_img,
                                    float* v_out_texture, float* v_out_alpha, void* stream);
/* The same two calls with `gt_img` as the uint8 (H,W,3) image the reference's data manager holds
 * (NS/data/datamanagers/full_images_datamanager.py:316-333): every value is converted as `get_gt_img` does
 * (NS/models/gstex.py:1238-1247, `image.float() / 255.0` = a multiplication by the fp32 reciprocal of 255), so the
 * results are bit-identical to converting first, without the two elementwise passes and the fp32 copy per view. */
int gstex_photometric_loss_forward_u8(int img_height, int img_width, int channels, const float* out_img,
                                      const float* out_texture, const float* out_alpha, const float* background,
                                      const uint8_t* gt_img, const float* mask, float ssim_lambda, int win_size,
                                      float win_sigma, float* rgb, float* losses, float* ssim_maps, void* scratch,
                                      size_t scratch_bytes, void* stream);
int gstex_photometric_loss_backward_u8(int img_height, int img_width, int channels, const float* out_img,
                                       const float* out_texture, const float* out_alpha, const float* background,
                                       const uint8_t* gt_img, const float* mask, float ssim_lambda, int win_size,
                                       float win_sigma, const float* ssim_maps, const float* v_main_loss, float* v_out_img,
                                       float* v_out_texture, float* v_out_alpha, void* stream);

/* ---- multi-GPU: the gradient sum of a view-sharded step over NVLink peer memory ----
 * Replaces the NCCL all-reduce torch DDP issues for the reference (NS/pipelines/base_pipeline.py:282) over the flat
 * gradient bucket.  `peer_bufs[world]` (a HOST array) holds, for every rank in rank order, this process's device
 * mapping of that rank's bucket of `count` floats (count % 4 == 0, 16-byte aligned; peer_bufs[rank] is the local
 * one).  The call sums slice `rank` (1/world of the bucket) over all ranks, in rank order, and stores the result into
 * that slice of every rank's bucket: once every rank has made the call, all buckets hold the full sum, bit-identical
 * on every rank.  The CALLER orders it: a barrier over the ranks before (every bucket complete) and after (every
 * store landed) -- torch symmetric memory's in gstex_b200.multiview.  `multicast_buf`, when not NULL, is an NVSwitch
 * multicast mapping of the same buckets: the slice is then reduced and replicated by the switch
 * (multimem.ld_reduce / multimem.st) and peer_bufs is not read.  `max_ctas` <= 0: the library picks (every SM for peer loads/stores, 16 CTAs for the multicast variant).
 * world <= gstex_peer_allreduce_max_world() = 8 (one NVSwitch domain). */
int gstex_peer_allreduce_max_world(void);
int gstex_peer_allreduce_f32(int rank, int world, const void* const* peer_bufs, long long count, void* multicast_buf,
                             int max_ctas, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GSTEX_B200_H */

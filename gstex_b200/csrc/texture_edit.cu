// texture_edit.cu -- next row f-4: splat-painting of an edit canvas into the charts.
// Replaces texture_edit (gstex_cuda/cuda/csrc/texture_edit.cu:11-236, wrapper :238-346).  sm_100a.
//
// WHAT (as the reference): the forward's traversal of every pixel's sorted tile list -- same alpha, skip and
// termination rules (texture_edit.cu:130-177) -- and, for each contributing pair whose view depth t * view_depth
// lies inside [depth_lower, depth_upper] of its pixel (:180-181), weight * (r*m, g*m, b*m, m, 1) added to channels
// 0..4 of the four bilinear corner texels of the pair's uv (:194-216, edit_jagged_chart,
// texture_helpers.cuh:239-250; always bilinear, replicate padding).  The blur floor is settings bit 0 here (:48).
//
// HOW (ours): the per-view packed records and conservative support boxes of pack.cu; warp-autonomous units
// (32 pixels, an 8x4 block of a 16x16 tile) handed out longest tile list first by a persistent grid; per round of 32
// list entries a warp tests the support boxes against its block, compacts the survivors and pulls their 128-byte
// records into its own double-buffered shared-memory stage (cp.async, 8 lanes x 16 B per record) one round ahead
// of the round it evaluates in list order; ids are requested three rounds ahead and boxes two, so a warp walking a
// long list alone does not sit through the id -> box -> record chain of dependent loads every round.  Culling is exact: a culled pair has alpha < 1/255 (skipped by the reference too); it could only
// END a pixel whose transmittance is within a factor 255/254 of the 1e-4 threshold, and from the moment a warp
// holds such a pixel it evaluates every remaining entry of the list.  Products that are exactly zero (untouched
// canvas: m = 0; zero bilinear weights at chart borders) are not sent to memory: x + 0 == x, so no result changes,
// and an edit that covers a small part of the image issues few atomics.
#include <cstdlib>

#include "raster_common.cuh"

namespace gstex {

struct EditParams {
    RasterGeom g;
    int C;  // channel stride of updated_texture (texture_info.z, >= 5)
    const SplatRec* recs;
    const float4* boxes;
    const int32_t* ids;
    const int2* tile_bins;
    const float *updated_img, *updated_alpha, *depth_lower, *depth_upper;
    const float *viewmat, *c2w;
    float fx, fy, cx, cy;
    int use_blur;
    float* updated_texture;
    uint32_t* unit_cursor;  // zeroed by pack_records_kernel
    const int32_t* order;
    int units_per_tile, num_units;
    int force_nocull;
};

constexpr int kEditWarps = 4;  // double-buffered 4 KB record stages: 8.5 KB per warp, static shared memory < 48 KB
// T * (1 - alpha) <= 1e-4 with alpha < 1/255 needs T <= 1e-4 * 255/254 = 1.0039e-4; a little above that
constexpr float kEditDangerT = 1.01e-4f;

struct EditWarpShared {
    SplatRec rec[2][32];  // 2 x 4 KB: the round being evaluated and the round being fetched
    int pos[2][32];       // round-local list position of each staged record
    int gid[2][32];       // and its splat
};
struct EditShared {
    EditWarpShared w[kEditWarps];
    Camera cam;
};

__device__ __forceinline__ void edit_cp_async16(void* smem_dst, const void* gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)),
                 "l"(gmem_src)
                 : "memory");
}
__device__ __forceinline__ void edit_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void edit_cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void edit_cp_async_wait_all_but_one() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }

// Compacts the entries of `mask` (bit = lane of the round) into stage `b` and starts the copy of their records
// (8 lanes x 16 B per record, cp.async -> LDGSTS: nobody waits here).  Returns the number of staged records.
__device__ __forceinline__ int edit_stage(EditWarpShared& ws, int b, unsigned mask, int gid, int lane,
                                          const SplatRec* __restrict__ recs) {
    const int ns = __popc(mask);
    if ((mask >> lane) & 1u) {
        const int rank = __popc(mask & ((1u << lane) - 1u));
        ws.pos[b][rank] = lane;
        ws.gid[b][rank] = gid;
    }
    __syncwarp();
    for (int s = lane >> 3; s < ns; s += 4) edit_cp_async16(&ws.rec[b][s].q[lane & 7], &recs[ws.gid[b][s]].q[lane & 7]);
    edit_cp_async_commit();
    return ns;
}

__device__ __forceinline__ void edit_add(float* p, float v) {
    if (v != 0.f) atomicAdd(p, v);  // result unused: RED.E.ADD.F32
}

__global__ void __launch_bounds__(kEditWarps * 32) texture_edit_kernel(const EditParams p) {
    __shared__ __align__(128) EditShared sm;
    const RasterGeom g = p.g;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    EditWarpShared& ws = sm.w[warp];
    load_camera(sm.cam, p.viewmat, p.c2w, p.fx, p.fy, p.cx, p.cy);
    __syncthreads();
    const bool use_blur = p.use_blur != 0;
    const int C = p.C;
    const unsigned lt_mask = (1u << lane) - 1u;

    while (true) {
        int unit = 0;
        if (lane == 0) unit = (int)atomicAdd(p.unit_cursor, 1u);
        unit = __shfl_sync(0xffffffffu, unit, 0);
        if (unit >= p.num_units) break;
        const int slot = unit / p.units_per_tile, w = unit - slot * p.units_per_tile;
        const int tile_id = p.order ? __ldg(p.order + slot) : slot;
        const int tile_x = tile_id % g.tiles_x, tile_y = tile_id / g.tiles_x;
        const int2 range = p.tile_bins[tile_id];
        int lx, ly;
        unit_pixel_xy(g, w, lane, lx, ly);
        const Pixel pix = make_pixel_xy(sm.cam, g, tile_x, tile_y, lx, ly);
        float X0, X1, Y0, Y1;  // pixel centres covered by this unit
        {
            int x0, x1, y0, y1;
            unit_rect(g, w, x0, x1, y0, y1);
            X0 = (float)(tile_x * g.bw + x0) + 0.5f; X1 = (float)(tile_x * g.bw + x1) + 0.5f;
            Y0 = (float)(tile_y * g.bw + y0) + 0.5f; Y1 = (float)(tile_y * g.bw + y1) + 0.5f;
        }
        // the pixel's edit: (r*m, g*m, b*m, m, 1) and its depth window (texture_edit.cu:55-56, 72-73, 194-211)
        const float m = __ldg(p.updated_alpha + pix.pix_id);
        const float v0 = __ldg(p.updated_img + 3 * pix.pix_id) * m, v1 = __ldg(p.updated_img + 3 * pix.pix_id + 1) * m,
                    v2 = __ldg(p.updated_img + 3 * pix.pix_id + 2) * m;
        const float lo = __ldg(p.depth_lower + pix.pix_id), hi = __ldg(p.depth_upper + pix.pix_id);

        bool done = !pix.inside;
        float T = 1.f;
        bool danger = p.force_nocull != 0;  // warp-uniform: culling is off from here on

        // ---- software pipeline over rounds of 32 list entries.  A warp walking a long list alone is bound by the
        // chain of dependent loads per round (id -> support box -> record): ids are requested three rounds ahead,
        // boxes two, and the culled survivors' records one round ahead (cp.async into the other stage), so that the
        // evaluation of round r waits for none of them.
        const int nrounds = (range.y - range.x + 31) >> 5;
        auto load_gid = [&](int r) { const int i = range.x + 32 * r + lane; return (r < nrounds && i < range.y) ? __ldg(p.ids + i) : -1; };
        auto load_box = [&](int g) { return g >= 0 ? __ldg(p.boxes + g) : make_float4(1.f, 1.f, 0.f, 0.f); };
        auto hits = [&](const float4 b) { return (b.x <= X1) && (b.z >= X0) && (b.y <= Y1) && (b.w >= Y0); };
        int gid_c = load_gid(0), gid_1 = load_gid(1), gid_2 = load_gid(2);
        float4 box_1 = load_box(gid_1);
        unsigned valid_c = __ballot_sync(0xffffffffu, gid_c >= 0);
        int ns_c = edit_stage(ws, 0, danger ? valid_c : __ballot_sync(0xffffffffu, gid_c >= 0 && hits(load_box(gid_c))),
                              gid_c, lane, p.recs);
        for (int r = 0; r < nrounds; ++r) {
            const int cur = r & 1;
            if (__all_sync(0xffffffffu, done)) break;  // (outstanding copies are drained below)
            // next round: cull with the boxes requested one iteration ago, start its record copies
            const unsigned valid_n = __ballot_sync(0xffffffffu, gid_1 >= 0);
            const bool culled_n = !danger;
            int ns_n = edit_stage(ws, cur ^ 1, danger ? valid_n : __ballot_sync(0xffffffffu, gid_1 >= 0 && hits(box_1)),
                                  gid_1, lane, p.recs);
            const float4 box_2 = load_box(gid_2);
            const int gid_3 = load_gid(r + 3);
            edit_cp_async_wait_all_but_one();  // this round's records have landed
            __syncwarp();
            int ns = ns_c;
            while (ns > 0) {
                unsigned pending = 0u;
                for (int s = 0; s < ns; ++s) {
                    if (!done) {
                        const SplatRec& rc = ws.rec[cur][s];
                        const float4 r0 = rc.q[0], r1 = rc.q[1], r2 = rc.q[2], r3 = rc.q[3], r4 = rc.q[4];
                        const PairEval e = eval_pair(pix, r0, r1, r2, r3, r4, use_blur);
                        const float next_T = T * (1.f - e.alpha);
                        if (next_T <= 1e-4f) done = true;  // tested before `skip`, texture_edit.cu:171-176
                        if (!e.skip && !done) {
                            const float t_view = e.t * pix.view_depth;
                            if (t_view >= lo && t_view <= hi) {
                                const float4 r5 = rc.q[5], r6 = rc.q[6], r7 = rc.q[7];
                                const float u = chart_coord(r6, e.delta);
                                const float v = chart_coord(r7, e.delta);
                                const TexQuery tq = tex_query(__float_as_int(r4.z), __float_as_int(r4.w),
                                                              __float_as_int(r5.w), u, v, true);
                                const float wq[4] = {tq.w00, tq.w01, tq.w10, tq.w11};
                                const int bq[4] = {tq.b00, tq.b01, tq.b10, tq.b11};
#pragma unroll
                                for (int c = 0; c < 4; ++c) {
                                    if (wq[c] != 0.f) {
                                        float* dst = p.updated_texture + (size_t)bq[c] * C;
                                        edit_add(dst, wq[c] * v0);
                                        edit_add(dst + 1, wq[c] * v1);
                                        edit_add(dst + 2, wq[c] * v2);
                                        edit_add(dst + 3, wq[c] * m);
                                        atomicAdd(dst + 4, wq[c]);  // weight * 1.f
                                    }
                                }
                            }
                            T = next_T;
                        }
                    }
                    if (!danger && __any_sync(0xffffffffu, !done && T <= kEditDangerT)) {
                        // a pixel of this warp can now be ended by a pair the boxes would cull: evaluate every
                        // remaining entry of the list, starting with the rest of this round
                        danger = true;
                        const int j = ws.pos[cur][s];
                        pending = valid_c & ~((2u << j) - 1u);
                        break;
                    }
                }
                ns = 0;
                if (pending) {  // restage the rest of this round unculled (the next round's copies finish with it)
                    __syncwarp();
                    ns = edit_stage(ws, cur, pending, gid_c, lane, p.recs);
                    edit_cp_async_wait_all();
                    __syncwarp();
                }
            }
            if (danger && culled_n) {  // the next round was staged culled before the flip: stage it again, in full
                edit_cp_async_wait_all();
                __syncwarp();
                ns_n = edit_stage(ws, cur ^ 1, valid_n, gid_1, lane, p.recs);
            }
            __syncwarp();
            gid_c = gid_1; valid_c = valid_n; ns_c = ns_n;
            gid_1 = gid_2; box_1 = box_2; gid_2 = gid_3;
        }
        edit_cp_async_wait_all();  // nothing of this unit may still be landing when the next unit stages
        __syncwarp();
    }
}

}  // namespace gstex

using namespace gstex;

extern "C" int gstex_texture_edit(int img_height, int img_width, int block_width, int channels, int n,
                                  int num_intersects, long long total_texels, const int32_t* texture_dims,
                                  const float* updated_img, const float* updated_alpha, const float* depth_lower,
                                  const float* depth_upper, const int32_t* gaussian_ids_sorted,
                                  const int32_t* tile_bins, const float* opacities, const float* means,
                                  const float* scales, float glob_scale, const float* quats, const float* uv0,
                                  const float* umap, const float* vmap, const float* viewmat, const float* c2w,
                                  float fx, float fy, float cx, float cy, int settings, float* updated_texture,
                                  void* workspace, size_t workspace_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (img_height <= 0 || img_width <= 0 || block_width < 2 || block_width > 16 || n < 0 || num_intersects < 0 ||
        total_texels < 0)
        return GSTEX_E_BADARG;
    if (channels < 5 || channels > GSTEX_MAX_CHANNELS) return GSTEX_E_UNSUPPORTED;  // channels 0..4 are written
    if (!tile_bins || !viewmat || !c2w || !updated_img || !updated_alpha || !depth_lower || !depth_upper || !workspace)
        return GSTEX_E_BADARG;
    if (total_texels > 0 && !updated_texture) return GSTEX_E_BADARG;
    if (n > 0 && (!texture_dims || !opacities || !means || !scales || !quats || !uv0 || !umap || !vmap))
        return GSTEX_E_BADARG;
    if (num_intersects > 0 && !gaussian_ids_sorted) return GSTEX_E_BADARG;
    if (workspace_bytes < ws_bytes(n)) return GSTEX_E_WORKSPACE;
    // the reference's torch::zeros (texture_edit.cu:308-310)
    if (total_texels > 0)
        GSTEX_CUDA_TRY(cudaMemsetAsync(updated_texture, 0, (size_t)total_texels * channels * sizeof(float), st));
    if (n == 0 || num_intersects == 0) return GSTEX_OK;
    const int use_blur = settings & 1;  // texture_edit.cu:48 (bit 0, not the forward's bit 9)
    const int num_tiles = ((img_width + block_width - 1) / block_width) * ((img_height + block_width - 1) / block_width);
    {
        // colours are not composited here: any readable (n,3) array serves as the packer's colour input
        const int rc = launch_pack_records(n, texture_dims, /*colors=*/means, opacities, means, scales, glob_scale,
                                           quats, uv0, umap, vmap, viewmat, c2w, fx, fy, cx, cy,
                                           use_blur ? GSTEX_SET_BLUR : 0, workspace, st, tile_bins, num_tiles);
        if (rc != GSTEX_OK) return rc;
    }
    EditParams p;
    p.g = make_raster_geom(img_height, img_width, block_width);
    p.C = channels;
    p.recs = ws_recs(workspace);
    p.boxes = ws_boxes(workspace, n);
    p.ids = gaussian_ids_sorted; p.tile_bins = reinterpret_cast<const int2*>(tile_bins);
    p.updated_img = updated_img; p.updated_alpha = updated_alpha; p.depth_lower = depth_lower;
    p.depth_upper = depth_upper; p.viewmat = viewmat; p.c2w = c2w;
    p.fx = fx; p.fy = fy; p.cx = cx; p.cy = cy; p.use_blur = use_blur;
    p.updated_texture = updated_texture;
    p.unit_cursor = ws_counters(workspace);
    p.order = num_tiles <= kOrderMaxTiles ? ws_order(workspace, n) : nullptr;
    p.units_per_tile = p.g.threads / 32;
    p.num_units = p.g.tiles_x * p.g.tiles_y * p.units_per_tile;
    p.force_nocull = force_nocull_flag();
    int grid = sm_count() * 6;  // persistent: units come from the cursor, surplus CTAs leave at once
    if (grid > (p.num_units + kEditWarps - 1) / kEditWarps) grid = (p.num_units + kEditWarps - 1) / kEditWarps;
    texture_edit_kernel<<<grid, kEditWarps * 32, 0, st>>>(p);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

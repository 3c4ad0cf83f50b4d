// raster_fwd.cu -- stage (c): per-tile textured forward compositing.
// Replaces texture_forward (gstex_cuda/cuda/csrc/texture.cu:11-329, wrapper :762-901).  sm_100a.
//
// Differences in HOW (not WHAT) from the reference kernel:
//   * per-splat invariants (rotation axes, inverse-scale squares, ray/plane numerator, projected mean) come
//     from a 128-byte packed record computed once per (view, splat) (pack.cu) instead of once per pixel;
//   * warp-specialised: one producer warp gathers the tile's records by sorted id with TMA bulk copies
//     (cp.async.bulk -> UBLKCP, 128 B each) into an 8-stage shared-memory ring guarded by full/empty
//     mbarriers; the eight pixel warps consume stages independently -- no block-wide barrier per batch;
//   * per stage the producer also tests each record's conservative support box against the pixel rows of
//     every warp and hands each warp a 32-bit "worth evaluating" mask, so a warp never touches splats that
//     cannot reach alpha >= 1/255 on its rows (exact: see pack.cu and the `danger` fallback below);
//   * the bilinear index/weight computation is done once per pair, not once per channel; texture
//     accumulators live in registers (channel count is a template bound), not in a float[64] local array;
//   * a warp stops as soon as all its 32 pixels are done (ballot), the producer when all warps are.
#include <cstddef>

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
    const float *viewmat, *c2w, *background;
    float fx, fy, cx, cy;
    int settings;
    float *out_img, *out_depth, *out_reg, *out_texture, *out_normal, *final_Ts, *out_reg_s;
    int32_t *final_idx, *depth_idx;
    StreamView sv;  // pair stream (REC kernels only)
};

struct FwdShared {
    SplatRec recs[kStages][kStageRecs];  // 32 KB
    uint64_t full[kStages];
    uint64_t empty[kStages];
    unsigned bits[kStages][8];  // per stage, per pixel warp: records worth evaluating
    int ids[kStages][kStageRecs];  // splat ids of the stage (recorded into the pair stream)
    Camera cam;
    int warps_done;
    volatile int stop_at;
};

template <int CMAX, bool REC>
__global__ void __launch_bounds__(kMaxTileThreads + 32, REC ? 2 : 3) texture_forward_kernel(const FwdParams p) {
    __shared__ __align__(128) FwdShared sm;
    const RasterGeom g = p.g;
    const int n_pix_warps = g.threads >> 5;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool is_producer = warp == n_pix_warps;
    load_camera(sm.cam, p.viewmat, p.c2w, p.fx, p.fy, p.cx, p.cy);
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&sm.full[s], 1);
            mbar_init(&sm.empty[s], n_pix_warps);
        }
        sm.warps_done = 0;
        sm.stop_at = 0x7fffffff;
        mbar_fence_init();
    }
    __syncthreads();
    const int tile_id = blockIdx.y * g.tiles_x + blockIdx.x;
    const int2 range = p.tile_bins[tile_id];
    const int n_stages = (range.y - range.x + kStageRecs - 1) / kStageRecs;

    if (is_producer) {
        // ===================== producer warp: gather records, build per-warp masks =====================
        const float tx0 = (float)(blockIdx.x * g.bw) + 0.5f, tx1 = (float)(blockIdx.x * g.bw + g.bw - 1) + 0.5f;
        const float ty = (float)(blockIdx.y * g.bw) + 0.5f;
        // software pipeline: ids two stages ahead, support boxes one stage ahead of the stage being published
        const float4 no_box = make_float4(1.f, 1.f, 0.f, 0.f);
        int g_next = (range.x + lane < range.y) ? __ldg(p.ids + range.x + lane) : -1;
        int g_next2 = (range.x + kStageRecs + lane < range.y) ? __ldg(p.ids + range.x + kStageRecs + lane) : -1;
        float4 box_next = (g_next >= 0) ? __ldg(p.boxes + g_next) : no_box;
        for (int n = 0; n < n_stages; ++n) {
            const int s = n % kStages;
            const int gid = g_next;
            const float4 box = box_next;
            g_next = g_next2;
            const int nxt = range.x + (n + 2) * kStageRecs + lane;
            g_next2 = (nxt < range.y) ? __ldg(p.ids + nxt) : -1;
            box_next = (g_next >= 0) ? __ldg(p.boxes + g_next) : no_box;
            if (n >= kStages) mbar_wait(&sm.empty[s], ((n / kStages) - 1) & 1);
            if (*(volatile int*)&sm.warps_done >= n_pix_warps) {  // every pixel of the tile is finished
                if (lane == 0) sm.stop_at = n;
                break;
            }
            unsigned mine = 0, any = 0;
            for (int w = 0; w < n_pix_warps; ++w) {
                int r0, r1;
                warp_rows(w, g.bw, r0, r1);
                const bool hit = gid >= 0 && box_overlaps(box, tx0, tx1, ty + (float)r0, ty + (float)r1);
                const unsigned b = __ballot_sync(0xffffffffu, hit);
                if (lane == w) mine = b;
                any |= b;
            }
            if (lane < n_pix_warps) sm.bits[s][lane] = mine;
            if (REC) sm.ids[s][lane] = gid;
            __syncwarp();
            // every record of the stage is copied, also those no warp asked for: a warp in `danger` mode (below)
            // evaluates all of them
            (void)any;
            const unsigned present = __ballot_sync(0xffffffffu, gid >= 0);
            if (lane == 0) mbar_arrive_expect_tx(&sm.full[s], (uint32_t)__popc(present) * (uint32_t)sizeof(SplatRec));
            if (gid >= 0) bulk_copy_g2s(&sm.recs[s][lane], p.recs + gid, sizeof(SplatRec), &sm.full[s]);
        }
        return;
    }

    // =============================== pixel warps: composite front to back ===============================
    const Pixel pix = make_pixel(sm.cam, g, blockIdx.x, blockIdx.y);
    const bool use_blur = p.settings & GSTEX_SET_BLUR;
    const bool use_ndc = p.settings & GSTEX_SET_NDC_REG;
    const bool bilinear = !(p.settings & GSTEX_SET_NEAREST);
    const int C = p.C;

    bool done = !pix.inside;
    float T = 1.f;
    int cur_idx = 0, pix_depth_index = -1;
    float3 rgb = make_float3(0.f, 0.f, 0.f), nrm = make_float3(0.f, 0.f, 0.f), ts = make_float3(0.f, 0.f, 0.f);
    float pix_depth = 0.f, pix_reg = 0.f;
    float tex[CMAX];
#pragma unroll
    for (int k = 0; k < CMAX; ++k) tex[k] = 0.f;

    // pair-stream cursor of this unit (REC only): current chunk, items in it, items in total
    int cur_chunk = -1, fill = kChunkItems, cnt = 0;
    bool rec = REC && (*(volatile uint32_t*)&p.sv.hdr->overflow == 0u);

    bool warp_done = __all_sync(0xffffffffu, done);
    if (warp_done && lane == 0) atomicAdd(&sm.warps_done, 1);

    for (int n = 0; n < n_stages; ++n) {
        const int s = n % kStages;
        const uint32_t parity = (n / kStages) & 1;
        bool stopped = false;
        while (!mbar_try_wait(&sm.full[s], parity)) {
            if (sm.stop_at <= n) { stopped = true; break; }
        }
        if (stopped) break;
        if (!warp_done) {
            const int batch_start = range.x + n * kStageRecs;
            const int count = min(kStageRecs, range.y - batch_start);
            unsigned remaining = (count >= 32) ? 0xffffffffu : ((1u << count) - 1u);  // records not yet passed
            unsigned todo = sm.bits[s][warp] & remaining;
            while (true) {
                // A pixel whose transmittance sits in (1e-4, 1e-4/(1-1/255)] can be terminated even by a pair
                // with alpha < 1/255 (next_T <= 1e-4 is tested before `skip`, texture.cu:216-221): then nothing
                // may be culled for this warp.
                if (__any_sync(0xffffffffu, !done && T <= 1.0041e-4f)) todo = remaining;
                if (todo == 0) break;
                const int i = __ffs(todo) - 1;
                remaining &= ~((2u << i) - 1u);
                todo &= remaining;
                const SplatRec& r = sm.recs[s][i];
                const float4 r0 = r.q[0], r1 = r.q[1], r2 = r.q[2], r3 = r.q[3], r4 = r.q[4];
                const PairEval e = eval_pair(pix, r0, r1, r2, r3, r4, use_blur);
                const float next_T = T * (1.f - e.alpha);
                bool hit = false;
                float h_T = 0.f, h_u = 0.f, h_v = 0.f, h_val[3] = {0.f, 0.f, 0.f};
                if (!done) {
                    if (next_T <= 1e-4f) {
                        done = true;  // this splat is not composited
                    } else if (!e.skip) {
                        const float vis = e.alpha * T;
                        const float4 r5 = r.q[5], r6 = r.q[6], r7 = r.q[7];
                        rgb.x += r5.x * vis; rgb.y += r5.y * vis; rgb.z += r5.z * vis;
                        nrm.x += vis * r1.x; nrm.y += vis * r1.y; nrm.z += vis * r1.z;
                        const float t_view = e.t * pix.view_depth;
                        const float u = clamp01(r6.w + (r6.x * e.delta.x + r6.y * e.delta.y + r6.z * e.delta.z));
                        const float v = clamp01(r7.w + (r7.x * e.delta.x + r7.y * e.delta.y + r7.z * e.delta.z));
                        const TexQuery q =
                            tex_query(__float_as_int(r4.z), __float_as_int(r4.w), __float_as_int(r5.w), u, v, bilinear);
                        const float* t00 = p.texture + (size_t)q.b00 * C;
                        const float* t01 = p.texture + (size_t)q.b01 * C;
                        const float* t10 = p.texture + (size_t)q.b10 * C;
                        const float* t11 = p.texture + (size_t)q.b11 * C;
#pragma unroll
                        for (int k = 0; k < CMAX; ++k) {
                            if (k < C) {
                                const float val = q.w00 * __ldg(t00 + k) + q.w01 * __ldg(t01 + k) +
                                                  q.w10 * __ldg(t10 + k) + q.w11 * __ldg(t11 + k);
                                tex[k] += vis * val;
                                if (REC && k < 3) h_val[k] = val;
                            }
                        }
                        if (REC) { hit = true; h_T = T; h_u = u; h_v = v; }
                        if (T > 0.5f) {  // median depth, texture.cu:286-291
                            pix_depth = t_view;
                            pix_depth_index = batch_start + i;
                        }
                        float d = e.t;
                        if (use_ndc) d = (kFar * t_view - kFar * kNear) / ((kFar - kNear) * t_view);
                        pix_reg += vis * (d * d * ts.x + ts.z - 2.f * d * ts.y);  // distortion(), helpers.cuh:259-264
                        ts.x += vis; ts.y += vis * d; ts.z += vis * d * d;
                        T = next_T;
                        cur_idx = batch_start + i;
                    }
                }
                if (REC) {  // append this splat's contributing pairs to the unit's stream, lane order
                    const unsigned hb = __ballot_sync(0xffffffffu, hit);
                    if (hb != 0u && rec) {
                        const int nh = __popc(hb);
                        int new_chunk = cur_chunk;
                        if (fill + nh > kChunkItems) {
                            int c = 0;
                            if (lane == 0) c = (int)atomicAdd(&p.sv.hdr->next_chunk, 1u);
                            c = __shfl_sync(0xffffffffu, c, 0);
                            if (c >= p.sv.num_chunks) {
                                if (lane == 0) p.sv.hdr->overflow = 1u;
                                rec = false;
                            } else {
                                if (lane == 0) p.sv.chunk_prev[c] = cur_chunk;
                                new_chunk = c;
                            }
                        }
                        if (rec) {
                            if (hit) {
                                int pos = fill + __popc(hb & ((1u << lane) - 1u));
                                int ch = cur_chunk;
                                if (pos >= kChunkItems) { pos -= kChunkItems; ch = new_chunk; }
                                float4* dst = reinterpret_cast<float4*>(p.sv.items + (size_t)ch * kChunkItems + pos);
                                const uint32_t gl = ((uint32_t)sm.ids[s][i] << 5) | (uint32_t)lane;
                                dst[0] = make_float4(__uint_as_float(gl), __int_as_float(batch_start + i), h_T, h_u);
                                dst[1] = make_float4(h_v, h_val[0], h_val[1], h_val[2]);
                            }
                            fill += nh;
                            if (fill > kChunkItems) { fill -= kChunkItems; cur_chunk = new_chunk; }
                            cnt += nh;
                        }
                    }
                }
                if (__all_sync(0xffffffffu, done)) {
                    warp_done = true;
                    if (lane == 0) atomicAdd(&sm.warps_done, 1);
                    break;
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.empty[s]);
    }

    if (REC) {
        const int unit0 = tile_id * kUnitsPerTile;
        if (lane == 0) p.sv.units[unit0 + warp] = make_int2(cur_chunk, rec ? cnt : 0);
        if (warp == 0 && lane >= n_pix_warps && lane < kUnitsPerTile) p.sv.units[unit0 + lane] = make_int2(-1, 0);
    }
    if (pix.inside) {
        const int id = pix.pix_id;
        p.final_Ts[id] = T;
        p.final_idx[id] = cur_idx;
        p.depth_idx[id] = pix_depth_index;
        const float bg0 = __ldg(p.background), bg1 = __ldg(p.background + 1), bg2 = __ldg(p.background + 2);
        p.out_img[3 * id + 0] = rgb.x + T * bg0;
        p.out_img[3 * id + 1] = rgb.y + T * bg1;
        p.out_img[3 * id + 2] = rgb.z + T * bg2;
        p.out_normal[3 * id + 0] = nrm.x; p.out_normal[3 * id + 1] = nrm.y; p.out_normal[3 * id + 2] = nrm.z;
        p.out_depth[id] = pix_depth;
        p.out_reg[id] = pix_reg;
        p.out_reg_s[3 * id + 0] = ts.x; p.out_reg_s[3 * id + 1] = ts.y; p.out_reg_s[3 * id + 2] = ts.z;
#pragma unroll
        for (int k = 0; k < CMAX; ++k)
            if (k < C) p.out_texture[(size_t)C * id + k] = tex[k];
    }
}

}  // namespace gstex

using namespace gstex;

extern "C" int gstex_pair_stream_supported(int channels, int n, int settings) {
    return channels >= 1 && channels <= 3 && n > 0 && n < (1 << 27) && !(settings & GSTEX_SET_UV_GRAD);
}

extern "C" size_t gstex_pair_stream_bytes(int img_height, int img_width, int block_width, long long max_pairs) {
    if (img_height <= 0 || img_width <= 0 || block_width < 2 || block_width > 16 || max_pairs < 0) return 0;
    const RasterGeom g = make_raster_geom(img_height, img_width, block_width);
    const size_t units = (size_t)g.tiles_x * g.tiles_y * kUnitsPerTile;
    // every unit can leave one partly filled chunk behind
    const size_t chunks = (size_t)(max_pairs + kChunkItems - 1) / kChunkItems + units;
    return stream_fixed_bytes((int)units) + 256 + chunks * (sizeof(PairItem) * kChunkItems + sizeof(int32_t)) + 128;
}

extern "C" int gstex_texture_forward(int img_height, int img_width, int block_width, int channels, int n,
                                     int num_intersects, const int32_t* texture_dims,
                                     const int32_t* gaussian_ids_sorted, const int32_t* tile_bins, const float* colors,
                                     const float* opacities, const float* means, const float* scales, float glob_scale,
                                     const float* quats, const float* uv0, const float* umap, const float* vmap,
                                     const float* texture, const float* viewmat, const float* c2w, float fx, float fy,
                                     float cx, float cy, int settings, const float* background, float* out_img,
                                     float* out_depth, float* out_reg, float* out_texture, float* out_normal,
                                     float* final_Ts, int32_t* final_idx, int32_t* depth_idx, float* out_reg_s,
                                     void* workspace, size_t workspace_bytes, void* pair_stream,
                                     size_t pair_stream_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (img_height <= 0 || img_width <= 0 || block_width < 2 || block_width > 16 || n < 0 || num_intersects < 0)
        return GSTEX_E_BADARG;
    if (settings & ~GSTEX_SETTINGS_SUPPORTED) return GSTEX_E_UNSUPPORTED;
    if (channels < 1 || channels > GSTEX_MAX_CHANNELS) return GSTEX_E_UNSUPPORTED;
    if (!tile_bins || !viewmat || !c2w || !background || !out_img || !out_depth || !out_reg || !out_texture ||
        !out_normal || !final_Ts || !final_idx || !depth_idx || !out_reg_s)
        return GSTEX_E_BADARG;
    if (n > 0 && (!texture_dims || !colors || !opacities || !means || !scales || !quats || !uv0 || !umap || !vmap ||
                  !texture || !workspace))
        return GSTEX_E_BADARG;
    if (num_intersects > 0 && !gaussian_ids_sorted) return GSTEX_E_BADARG;
    if (n > 0 && workspace_bytes < ws_bytes(n)) return GSTEX_E_WORKSPACE;
    if (n > 0) {
        const int rc = launch_pack_records(n, texture_dims, colors, opacities, means, scales, glob_scale, quats, uv0, umap,
                                           vmap, viewmat, c2w, fx, fy, cx, cy, settings, workspace, st);
        if (rc != GSTEX_OK) return rc;
    }
    FwdParams p;
    p.g = make_raster_geom(img_height, img_width, block_width);
    p.C = channels;
    p.recs = n > 0 ? ws_recs(workspace) : nullptr;
    p.boxes = n > 0 ? ws_boxes(workspace, n) : nullptr;
    p.ids = gaussian_ids_sorted; p.tile_bins = reinterpret_cast<const int2*>(tile_bins);
    p.texture = texture; p.viewmat = viewmat; p.c2w = c2w; p.background = background;
    p.fx = fx; p.fy = fy; p.cx = cx; p.cy = cy; p.settings = settings;
    p.out_img = out_img; p.out_depth = out_depth; p.out_reg = out_reg; p.out_texture = out_texture;
    p.out_normal = out_normal; p.final_Ts = final_Ts; p.out_reg_s = out_reg_s; p.final_idx = final_idx;
    p.depth_idx = depth_idx;
    const dim3 grid(p.g.tiles_x, p.g.tiles_y);
    const int threads = p.g.threads + 32;  // + the producer warp
    // Record the pair stream for the stream backward when the caller gave a pool and the configuration is the
    // one that backward covers (<= 3 channels, no uv gradients); otherwise the backward recomputes.
    bool record = pair_stream != nullptr && gstex_pair_stream_supported(channels, n, settings) &&
                  make_stream_view(pair_stream, pair_stream_bytes, p.g.tiles_x * p.g.tiles_y, p.sv);
    if (pair_stream != nullptr && pair_stream_bytes >= sizeof(StreamHeader)) {
        GSTEX_CUDA_TRY(cudaMemsetAsync(pair_stream, 0, sizeof(StreamHeader), st));
        if (!record)  // tell the backward there is nothing to read
            GSTEX_CUDA_TRY(cudaMemsetAsync(reinterpret_cast<char*>(pair_stream) + offsetof(StreamHeader, overflow), 1, 1, st));
    }
    if (record) texture_forward_kernel<3, true><<<grid, threads, 0, st>>>(p);
    else if (channels <= 3) texture_forward_kernel<3, false><<<grid, threads, 0, st>>>(p);
    else if (channels <= 6) texture_forward_kernel<6, false><<<grid, threads, 0, st>>>(p);
    else texture_forward_kernel<GSTEX_MAX_CHANNELS, false><<<grid, threads, 0, st>>>(p);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

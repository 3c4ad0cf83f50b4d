// raster_fwd.cu -- stage (c): textured forward compositing.
// Replaces texture_forward (gstex_cuda/cuda/csrc/texture.cu:11-329, wrapper :762-901).  sm_100a.
//
// Differences in HOW (not WHAT) from the reference kernel:
//   * per-splat invariants (rotation axes, inverse-scale squares, ray/plane numerator, projected mean) come
//     from a 128-byte packed record computed once per (view, splat) (pack.cu) instead of once per pixel;
//   * warp-autonomous units: the work item is one warp's 32 pixels of a tile (an 8x4 block for 16x16 tiles),
//     handed out by a persistent grid through an atomic cursor.  A warp walks its tile's sorted list on its own:
//     32 list entries at a time it tests the conservative support boxes (pack.cu) against its 8x4 block,
//     compacts the survivors into a small queue, and pulls their 128-byte records into its private
//     double-buffered shared-memory stage with TMA bulk copies (cp.async.bulk -> UBLKCP) completing on an
//     mbarrier, one batch ahead of the batch being composited.  Warps never wait for each other (the reference
//     and our earlier ring pipeline synchronise the 8 warps of a tile on every batch), a warp whose pixels are
//     all finished just takes the next unit;
//   * culling is exact: a culled pair has alpha < 1/255 (skipped by the reference too, texture.cu:213); only a
//     pixel whose transmittance sits in (1e-4, 1e-4/(1-1/255)] can be terminated by such a pair (next_T <= 1e-4
//     is tested before `skip`, texture.cu:216-221), and as soon as a warp sees one it drops its queue and
//     re-walks the list from that entry with culling off;
//   * the bilinear index/weight computation is done once per pair, not once per channel; texture
//     accumulators live in registers (channel count is a template bound), not in a float[64] local array;
//   * REC kernels append one 32-byte item per contributing pair to the pair stream (pair_stream.cuh) that the
//     stream backward consumes.
#include <cstddef>
#include <cstdlib>

#include "raster_fwd.cuh"

namespace gstex {

#ifdef GSTEX_STATS
__device__ unsigned long long g_fwd_stats[16];
#define STAT_ADD(k, v) st_[k] += (unsigned long long)(v)
#else
#define STAT_ADD(k, v)
#endif

// Tuning knobs (scripts/gpu_fwd_sweep.sh builds and times the alternatives; the defaults are the measured best)
#ifndef GSTEX_FWD_WARPS
#define GSTEX_FWD_WARPS 5
#endif
#ifndef GSTEX_FWD_CTAS
#define GSTEX_FWD_CTAS 4
#endif
#ifndef GSTEX_FWD_BATCH
#define GSTEX_FWD_BATCH 16
#endif
constexpr int kFwdWarps = GSTEX_FWD_WARPS;    // warps per CTA; they share nothing but the camera (4 CTAs/SM = 20 warps)
constexpr int kFwdCtasPerSm = GSTEX_FWD_CTAS;
constexpr int kHits = 64;       // ring of contributing pairs waiting for their texture fetch (deferred, lane-dense)
constexpr int kBatch = GSTEX_FWD_BATCH;       // records per stage (<= 32: one TMA bulk copy per lane)
constexpr int kQueue = 128;                   // survivor ring: batch being composited + batch in flight + the next one
                                              // filling + one culling round <= 3 * kBatch + 31 entries
static_assert(kBatch <= 32 && 3 * kBatch + 31 <= kQueue, "survivor ring too small");

struct FwdWarpShared {
    SplatRec rec[2][kBatch];  // 8 KB
    int q_gid[kQueue];
    int q_idx[kQueue];
    float4 h_e[kHits];   // (record slot | pixel lane << 5, T before the pair, u, v); after the fetch: (val0..2, vis)
    float h_vis[kHits];  // alpha * T of the pair
    uint64_t bar[2];
};

struct FwdShared {
    FwdWarpShared w[kFwdWarps];
    Camera cam;
    float4 origin;  // camera position: re-read from here in every pair evaluation instead of living in registers
};

// VIS: settings bit 15 (viewer): normals are accumulated flipped towards the camera (texture.cu:236-241); its own
// instantiations, so the training kernels carry no trace of it.
template <int CMAX, bool REC, bool VIS = false>
__global__ void __launch_bounds__(kFwdWarps * 32, kFwdCtasPerSm) texture_forward_kernel(const FwdParams p) {
    // <= 3 channels (training): the texture fetch of a contributing pair is deferred to a lane-dense phase
    constexpr bool kDefer = (CMAX == 3);
    extern __shared__ __align__(128) unsigned char fwd_smem[];
    FwdShared& sm = *reinterpret_cast<FwdShared*>(fwd_smem);
    const RasterGeom g = p.g;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    FwdWarpShared& ws = sm.w[warp];
    load_camera(sm.cam, p.viewmat, p.c2w, p.fx, p.fy, p.cx, p.cy);
    if (threadIdx.x == 32) sm.origin = make_float4(__ldg(p.c2w + 3), __ldg(p.c2w + 7), __ldg(p.c2w + 11), 0.f);
    if (lane == 0) {
        mbar_init(&ws.bar[0], 1);
        mbar_init(&ws.bar[1], 1);
        mbar_fence_init();
    }
    __syncthreads();
    const bool use_blur = p.settings & GSTEX_SET_BLUR;
    const bool use_ndc = p.settings & GSTEX_SET_NDC_REG;
    const bool bilinear = !(p.settings & GSTEX_SET_NEAREST);
    const int C = p.C;
    const unsigned lt_mask = (1u << lane) - 1u;
    const float4 no_box = make_float4(1.f, 1.f, 0.f, 0.f);
    unsigned phase = 0;  // bit b: parity to wait for on bar[b]
    bool rec_ok = REC && (*(volatile uint32_t*)&p.sv.hdr->overflow == 0u);

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
        // pixel centres covered by this unit, for the support-box test
        float X0, X1, Y0, Y1;
        {
            int x0, x1, y0, y1;
            unit_rect(g, w, x0, x1, y0, y1);
            X0 = (float)(tile_x * g.bw + x0) + 0.5f; X1 = (float)(tile_x * g.bw + x1) + 0.5f;
            Y0 = (float)(tile_y * g.bw + y0) + 0.5f; Y1 = (float)(tile_y * g.bw + y1) + 0.5f;
        }

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

        // ---- list walk state: cursor = next list position to test; ring indices are monotonic
        int cursor = range.x, last_pos = range.x - 1;
        int q_head = 0, q_issue = 0, q_tail = 0;
        bool nocull = p.force_nocull != 0, danger = false;
        int gid_pf, gid_pf2;
        float4 box_pf;
        auto prime = [&]() {  // ids two rounds ahead, support boxes one round ahead of the round being tested
            gid_pf = (cursor + lane < range.y) ? __ldg(p.ids + cursor + lane) : -1;
            gid_pf2 = (cursor + 32 + lane < range.y) ? __ldg(p.ids + cursor + 32 + lane) : -1;
            box_pf = (gid_pf >= 0) ? __ldg(p.boxes + gid_pf) : no_box;
        };
        // Test list entries until a batch of survivors is queued (or the list ends), then start the TMA copies
        // of the next `kBatch` queued records into stage `buf`.  Returns the number of records in flight.
        auto fill_and_issue = [&](int buf) -> int {
            while (q_tail - q_issue < kBatch && cursor < range.y) {
                const int gid = gid_pf;
                const float4 box = box_pf;
                const int pos = cursor + lane;
                cursor += 32;
                gid_pf = gid_pf2;
                gid_pf2 = (cursor + 32 + lane < range.y) ? __ldg(p.ids + cursor + 32 + lane) : -1;
                box_pf = (gid_pf >= 0) ? __ldg(p.boxes + gid_pf) : no_box;
                const bool hit = gid >= 0 && (nocull || box_overlaps(box, X0, X1, Y0, Y1));
                const unsigned m = __ballot_sync(0xffffffffu, hit);
                if (hit) {
                    const int slot = (q_tail + __popc(m & lt_mask)) & (kQueue - 1);
                    ws.q_gid[slot] = gid;
                    ws.q_idx[slot] = pos;
                }
                q_tail += __popc(m);
            }
            __syncwarp();
            const int n_rec = min(q_tail - q_issue, kBatch);
            if (n_rec > 0) {
                if (lane == 0) mbar_arrive_expect_tx(&ws.bar[buf], (uint32_t)n_rec * (uint32_t)sizeof(SplatRec));
                if (lane < n_rec)
                    bulk_copy_g2s(&ws.rec[buf][lane], p.recs + ws.q_gid[(q_issue + lane) & (kQueue - 1)], sizeof(SplatRec),
                                  &ws.bar[buf]);
                q_issue += n_rec;
            }
            return n_rec;
        };

#ifdef GSTEX_STATS
        unsigned long long st_[16];
        for (int k = 0; k < 16; ++k) st_[k] = 0;
        STAT_ADD(9, 1);
#endif
        // ---- deferred texture phase (kDefer): one queued pair per lane, whatever pixel it belongs to.
        // Computes the bilinear sample, appends the pair to the stream (REC) and hands vis*sample back to the
        // lane that owns the pixel, which adds its pairs in list order (so the sums are deterministic).
        int h_head = 0, h_tail = 0;
        unsigned long long mine = 0ull;  // ring slots holding pairs of this lane's pixel
        int cur = 0, n_cur = 0;
        auto fetch_hits = [&](const int m) {
            __syncwarp();
            const int hs = (h_head + lane) & (kHits - 1);
            const bool valid = lane < m;
            float4 ent = make_float4(0.f, 0.f, 0.f, 0.f);
            float vis = 0.f, val0 = 0.f, val1 = 0.f, val2 = 0.f;
            int si = 0, pl = 0;
            if (valid) {
                ent = ws.h_e[hs];
                vis = ws.h_vis[hs];
                const int pk = __float_as_int(ent.x);
                si = pk & 31; pl = (pk >> 5) & 31;
                const SplatRec& rr = ws.rec[cur][si];
                const float4 r4 = rr.q[4];
                const int start = __float_as_int(rr.q[5].w);
                const TexQuery tq = tex_query(__float_as_int(r4.z), __float_as_int(r4.w), start, ent.z, ent.w, bilinear);
                if (p.texture4 != nullptr) {  // padded layout: one 128-bit load per corner instead of C scalar loads
                    const float4 a = __ldg(p.texture4 + tq.b00), b = __ldg(p.texture4 + tq.b01);
                    const float4 c = __ldg(p.texture4 + tq.b10), d = __ldg(p.texture4 + tq.b11);
                    val0 = bilinear_ref(tq.w00, tq.w01, tq.w10, tq.w11, a.x, b.x, c.x, d.x);
                    if (C > 1) val1 = bilinear_ref(tq.w00, tq.w01, tq.w10, tq.w11, a.y, b.y, c.y, d.y);
                    if (C > 2) val2 = bilinear_ref(tq.w00, tq.w01, tq.w10, tq.w11, a.z, b.z, c.z, d.z);
                } else {
                    const float* t00 = p.texture + (size_t)tq.b00 * C;
                    const float* t01 = p.texture + (size_t)tq.b01 * C;
                    const float* t10 = p.texture + (size_t)tq.b10 * C;
                    const float* t11 = p.texture + (size_t)tq.b11 * C;
                    val0 = bilinear_ref(tq.w00, tq.w01, tq.w10, tq.w11, __ldg(t00), __ldg(t01), __ldg(t10), __ldg(t11));
                    if (C > 1) val1 = bilinear_ref(tq.w00, tq.w01, tq.w10, tq.w11, __ldg(t00 + 1), __ldg(t01 + 1), __ldg(t10 + 1), __ldg(t11 + 1));
                    if (C > 2) val2 = bilinear_ref(tq.w00, tq.w01, tq.w10, tq.w11, __ldg(t00 + 2), __ldg(t01 + 2), __ldg(t10 + 2), __ldg(t11 + 2));
                }
                // affine texel activation (SH2RGB of the caller folded into the fetch; the weights sum to one);
                // (1, 0) leaves the sample bit for bit
                val0 = __fmaf_rn(p.tex_scale, val0, p.tex_bias);
                if (C > 1) val1 = __fmaf_rn(p.tex_scale, val1, p.tex_bias);
                if (C > 2) val2 = __fmaf_rn(p.tex_scale, val2, p.tex_bias);
            }
            if (REC && rec_ok) {  // the m pairs are consecutive items of the unit's stream, list order
                int new_chunk = cur_chunk;
                if (fill + m > kChunkItems) {
                    int c = 0;
                    if (lane == 0) c = (int)atomicAdd(&p.sv.hdr->next_chunk, 1u);
                    c = __shfl_sync(0xffffffffu, c, 0);
                    if (c >= p.sv.num_chunks) {
                        if (lane == 0) p.sv.hdr->overflow = 1u;
                        rec_ok = false;
                    } else {
                        if (lane == 0) p.sv.chunk_prev[c] = cur_chunk;
                        new_chunk = c;
                    }
                }
                if (rec_ok) {
                    if (valid) {
                        int pos = fill + lane;
                        int ch = cur_chunk;
                        if (pos >= kChunkItems) { pos -= kChunkItems; ch = new_chunk; }
                        float4* dst = reinterpret_cast<float4*>(p.sv.items + (size_t)ch * kChunkItems + pos);
                        const int q = (q_head + si) & (kQueue - 1);
                        const uint32_t gl = ((uint32_t)ws.q_gid[q] << 5) | (uint32_t)pl;
                        dst[0] = make_float4(__uint_as_float(gl), __int_as_float(ws.q_idx[q]), ent.y, ent.z);
                        dst[1] = make_float4(ent.w, val0, val1, val2);
                    }
                    fill += m;
                    if (fill > kChunkItems) { fill -= kChunkItems; cur_chunk = new_chunk; }
                    cnt += m;
                }
            }
            if (valid) ws.h_e[hs] = make_float4(val0, val1, val2, vis);
            __syncwarp();
            {   // this lane's pixel: its pairs among the m just fetched, oldest first
                const int sh = h_head & (kHits - 1);
                const unsigned long long rot = sh ? ((mine >> sh) | (mine << (64 - sh))) : mine;
                const unsigned window = (m >= 32) ? 0xffffffffu : ((1u << m) - 1u);
                unsigned low = (unsigned)rot & window;
                while (low) {
                    const int j = __ffs(low) - 1;
                    low &= low - 1u;
                    const float4 e4 = ws.h_e[(h_head + j) & (kHits - 1)];
                    tex[0] = __fmaf_rn(e4.w, e4.x, tex[0]);
                    if (CMAX > 1) tex[1] = __fmaf_rn(e4.w, e4.y, tex[1]);
                    if (CMAX > 2) tex[2] = __fmaf_rn(e4.w, e4.z, tex[2]);
                }
                const unsigned long long wmask = (unsigned long long)window;
                mine &= ~(sh ? ((wmask << sh) | (wmask >> (64 - sh))) : wmask);
            }
            h_head += m;
            __syncwarp();
        };

        bool unit_done = __all_sync(0xffffffffu, done);
        if (!unit_done) {
            prime();
            n_cur = fill_and_issue(cur);
        }
        while (n_cur > 0) {
            const int n_next = fill_and_issue(cur ^ 1);  // next batch in flight while this one is composited
            mbar_wait(&ws.bar[cur], (phase >> cur) & 1u);
            phase ^= 1u << cur;
            bool flush = false;
            int i = 0;
            for (; i < n_cur; ++i) {
                // A pixel whose transmittance sits in (1e-4, 1e-4/(1-1/255)] can be terminated even by a pair with
                // alpha < 1/255: from here on nothing may be culled for this warp until that pixel is finished.
                // `danger` is warp-uniform and refreshed whenever a transmittance changed (a pixel that finishes
                // without one may leave it set a little longer, which only delays the return to culling).
                if (danger && !nocull) { flush = true; break; }
                if (!danger) nocull = p.force_nocull != 0;
                const SplatRec& r = ws.rec[cur][i];
                const int q = (q_head + i) & (kQueue - 1);
                const float4 r0 = r.q[0], r1 = r.q[1], r2 = r.q[2], r3 = r.q[3], r4 = r.q[4];
                // The ray origin is the same for every pixel; keeping its three floats alive across the whole loop cost
                // a local-memory spill whose reload stalled every iteration (7 % of the kernel's stall samples).  A
                // volatile shared-memory read at the top of the iteration keeps them short-lived instead.
                Pixel pe = pix;
                {
                    float ox, oy, oz, ow;
                    asm volatile("ld.volatile.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                                 : "=f"(ox), "=f"(oy), "=f"(oz), "=f"(ow)
                                 : "r"((unsigned)__cvta_generic_to_shared(&sm.origin)));
                    pe.origin = make_float3(ox, oy, oz);
                }
                const PairEval e = eval_pair(pe, r0, r1, r2, r3, r4, use_blur);
                const float next_T = __fmul_rn(T, __fsub_rn(1.f, e.alpha));
#ifdef GSTEX_STATS
                {
                    const unsigned live = __ballot_sync(0xffffffffu, !done);
                    const unsigned pot = __ballot_sync(0xffffffffu, !e.skip);
                    const float4 bx = __ldg(p.boxes + ws.q_gid[q]);
                    const bool inbox = pix.px >= bx.x && pix.px <= bx.z && pix.py >= bx.y && pix.py <= bx.w;
                    const unsigned inb = __ballot_sync(0xffffffffu, inbox);
                    STAT_ADD(0, 1); STAT_ADD(1, __popc(live)); STAT_ADD(2, __popc(pot)); STAT_ADD(3, __popc(pot & live));
                    STAT_ADD(4, (pot & live) == 0u); STAT_ADD(6, pot == 0u); STAT_ADD(7, __popc(inb & live));
                    STAT_ADD(8, __popc(inb)); STAT_ADD(5, (inb & live) == 0u);
                    STAT_ADD(10, nocull);
                }
#endif
                bool hit = false;
                float h_T = 0.f, h_u = 0.f, h_v = 0.f, h_vis = 0.f;
                if (!done) {
                    if (next_T <= 1e-4f) {
                        done = true;  // this splat is not composited
                    } else if (!e.skip) {
                        const int idx = ws.q_idx[q];
                        const float vis = __fmul_rn(e.alpha, T);
                        const float4 r5 = r.q[5], r6 = r.q[6], r7 = r.q[7];
                        rgb.x = __fmaf_rn(vis, r5.x, rgb.x); rgb.y = __fmaf_rn(vis, r5.y, rgb.y); rgb.z = __fmaf_rn(vis, r5.z, rgb.z);
                        if constexpr (VIS) {  // the reference tests the UNCLAMPED denominator dot(ax3, ray) > 0 (texture.cu:236-241)
                            const bool flip = dot3_ref(r1.x, r1.y, r1.z, pe.ray.x, pe.ray.y, pe.ray.z) > 0.f;
                            nrm.x = __fmaf_rn(vis, flip ? -r1.x : r1.x, nrm.x); nrm.y = __fmaf_rn(vis, flip ? -r1.y : r1.y, nrm.y);
                            nrm.z = __fmaf_rn(vis, flip ? -r1.z : r1.z, nrm.z);
                        } else {
                            nrm.x = __fmaf_rn(vis, r1.x, nrm.x); nrm.y = __fmaf_rn(vis, r1.y, nrm.y); nrm.z = __fmaf_rn(vis, r1.z, nrm.z);
                        }
                        const float t_view = __fmul_rn(e.t, pix.view_depth);
                        const float u = chart_coord(r6, e.delta);
                        const float v = chart_coord(r7, e.delta);
                        hit = true;
                        if constexpr (kDefer) {
                            h_T = T; h_u = u; h_v = v; h_vis = vis;
                        } else {
                            const TexQuery tq =
                                tex_query(__float_as_int(r4.z), __float_as_int(r4.w), __float_as_int(r5.w), u, v, bilinear);
                            const float* t00 = p.texture + (size_t)tq.b00 * C;
                            const float* t01 = p.texture + (size_t)tq.b01 * C;
                            const float* t10 = p.texture + (size_t)tq.b10 * C;
                            const float* t11 = p.texture + (size_t)tq.b11 * C;
#pragma unroll
                            for (int k = 0; k < CMAX; ++k) {
                                if (k < C) {
                                    const float val = __fmaf_rn(p.tex_scale, bilinear_ref(tq.w00, tq.w01, tq.w10, tq.w11, __ldg(t00 + k),
                                                                                          __ldg(t01 + k), __ldg(t10 + k), __ldg(t11 + k)), p.tex_bias);
                                    tex[k] = __fmaf_rn(vis, val, tex[k]);
                                }
                            }
                        }
                        if (T > 0.5f) {  // median depth, texture.cu:286-291
                            pix_depth = t_view;
                            pix_depth_index = idx;
                        }
                        const float d = use_ndc ? ndc_depth_ref(t_view) : e.t;
                        distortion_ref(d, vis, ts, pix_reg);  // distortion(), helpers.cuh:259-264
                        T = next_T;
                        cur_idx = idx;
                    }
                }
                const unsigned hb = __ballot_sync(0xffffffffu, hit);
                if (hb != 0u) {
                    danger = __any_sync(0xffffffffu, !done && T <= 1.0041e-4f);
                    if constexpr (kDefer) {  // queue this splat's contributing pairs for the lane-dense texture phase
                        if (hit) {
                            const int pos = (h_tail + __popc(hb & lt_mask)) & (kHits - 1);
                            ws.h_e[pos] = make_float4(__int_as_float(i | (lane << 5)), h_T, h_u, h_v);
                            ws.h_vis[pos] = h_vis;
                            mine |= 1ull << pos;
                        }
                        h_tail += __popc(hb);
                        if (h_tail - h_head >= 32) fetch_hits(32);
                    }
                }
                if (__all_sync(0xffffffffu, done)) { unit_done = true; break; }
            }
            __syncwarp();
            if constexpr (kDefer) {  // the queued pairs point into this batch's record stage: finish them now
                while (h_tail != h_head) fetch_hits(min(32, h_tail - h_head));
            }
            if (unit_done || flush) {
                if (n_next > 0) {  // drain the copies in flight before their stage and barrier are reused
                    mbar_wait(&ws.bar[cur ^ 1], (phase >> (cur ^ 1)) & 1u);
                    phase ^= 1u << (cur ^ 1);
                }
                if (unit_done) break;
                // re-walk from the entry after the last evaluated one, with culling off
                if (i > 0) last_pos = ws.q_idx[(q_head + i - 1) & (kQueue - 1)];
                nocull = true;
                cursor = last_pos + 1;
                q_head = q_issue = q_tail = 0;
                prime();
                n_cur = fill_and_issue(cur);
                continue;
            }
            last_pos = ws.q_idx[(q_head + n_cur - 1) & (kQueue - 1)];
            q_head += n_cur;
            n_cur = n_next;
            cur ^= 1;
            {   // cull the rest of the list against the rectangle of the pixels that are still alive
                const int big = 1 << 30;
                const int ipx = (int)pix.px, ipy = (int)pix.py;  // pixel centre = index + 0.5
                const int ax0 = __reduce_min_sync(0xffffffffu, done ? big : ipx), ax1 = __reduce_max_sync(0xffffffffu, done ? -big : ipx);
                const int ay0 = __reduce_min_sync(0xffffffffu, done ? big : ipy), ay1 = __reduce_max_sync(0xffffffffu, done ? -big : ipy);
                X0 = (float)ax0 + 0.5f; X1 = (float)ax1 + 0.5f; Y0 = (float)ay0 + 0.5f; Y1 = (float)ay1 + 0.5f;
            }
        }

#ifdef GSTEX_STATS
        if (lane == 0)
            for (int k = 0; k < 16; ++k) if (st_[k]) atomicAdd(&g_fwd_stats[k], st_[k]);
#endif
        if (REC) {
            const int unit0 = tile_id * kUnitsPerTile;
            if (lane == 0) p.sv.units[unit0 + w] = make_int2(cur_chunk, rec_ok ? cnt : 0);
            if (w == 0 && lane >= p.units_per_tile && lane < kUnitsPerTile) p.sv.units[unit0 + lane] = make_int2(-1, 0);
        }
        if (pix.inside) {
            const float bg0 = __ldg(p.background), bg1 = __ldg(p.background + 1), bg2 = __ldg(p.background + 2);
            const int id = pix.pix_id;
            p.final_Ts[id] = T;
            p.final_idx[id] = cur_idx;
            p.depth_idx[id] = pix_depth_index;
            p.out_img[3 * id + 0] = __fmaf_rn(bg0, T, rgb.x);
            p.out_img[3 * id + 1] = __fmaf_rn(bg1, T, rgb.y);
            p.out_img[3 * id + 2] = __fmaf_rn(bg2, T, rgb.z);
            p.out_normal[3 * id + 0] = nrm.x; p.out_normal[3 * id + 1] = nrm.y; p.out_normal[3 * id + 2] = nrm.z;
            p.out_depth[id] = pix_depth;
            p.out_reg[id] = pix_reg;
            p.out_reg_s[3 * id + 0] = ts.x; p.out_reg_s[3 * id + 1] = ts.y; p.out_reg_s[3 * id + 2] = ts.z;
#pragma unroll
            for (int k = 0; k < CMAX; ++k)
                if (k < C) p.out_texture[(size_t)C * id + k] = tex[k];
        }
        __syncwarp();
    }
}

}  // namespace gstex

using namespace gstex;

#ifdef GSTEX_STATS
extern "C" int gstex_debug_fwd_stats(unsigned long long* out, int reset) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, g_fwd_stats, sizeof(unsigned long long) * 16);
    if (reset) { unsigned long long z[16] = {0}; cudaMemcpyToSymbol(g_fwd_stats, z, sizeof(z)); }
    return 0;
}
#endif

static int g_forward_variant = 0;
extern "C" int gstex_debug_forward_variant(int v) {
    const int old = g_forward_variant;
    g_forward_variant = v;
    return old;
}

extern "C" int gstex_pair_stream_supported(int channels, int n, int settings) {
    return channels >= 1 && channels <= 3 && n > 0 && n < (1 << 27) &&
           !(settings & (GSTEX_SET_UV_GRAD | GSTEX_SET_VIS_NORMALS));
}

extern "C" size_t gstex_pair_stream_bytes(int img_height, int img_width, int block_width, long long max_pairs) {
    if (img_height <= 0 || img_width <= 0 || block_width < 2 || block_width > 16 || max_pairs < 0) return 0;
    const RasterGeom g = make_raster_geom(img_height, img_width, block_width);
    const size_t units = (size_t)g.tiles_x * g.tiles_y * kUnitsPerTile;
    // every unit can leave one partly filled chunk behind
    const size_t chunks = (size_t)(max_pairs + kChunkItems - 1) / kChunkItems + units;
    return stream_fixed_bytes((int)units) + 256 + chunks * (sizeof(PairItem) * kChunkItems + sizeof(int32_t)) + 128;
}

static int forward_impl(int img_height, int img_width, int block_width, int channels, int n,
                                     int num_intersects, const int32_t* texture_dims,
                                     const int32_t* gaussian_ids_sorted, const int32_t* tile_bins, const float* colors,
                                     const float* opacities, const float* means, const float* scales, float glob_scale,
                                     const float* quats, const float* uv0, const float* umap, const float* vmap,
                                     const float* texture, const float* texture4, int records_valid, float tex_scale, float tex_bias, const float* viewmat, const float* c2w, float fx, float fy,
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
    if (!workspace) return GSTEX_E_BADARG;
    if (n > 0 && (!texture_dims || !colors || !opacities || !means || !scales || !quats || !uv0 || !umap || !vmap ||
                  !texture))
        return GSTEX_E_BADARG;
    if (num_intersects > 0 && !gaussian_ids_sorted) return GSTEX_E_BADARG;
    if (workspace_bytes < ws_bytes(n)) return GSTEX_E_WORKSPACE;
    const int num_tiles = ((img_width + block_width - 1) / block_width) * ((img_height + block_width - 1) / block_width);
    if (n > 0 && !records_valid) {
        const int rc = launch_pack_records(n, texture_dims, colors, opacities, means, scales, glob_scale, quats, uv0, umap,
                                           vmap, viewmat, c2w, fx, fy, cx, cy, settings, workspace, st, tile_bins,
                                           num_tiles);
        if (rc != GSTEX_OK) return rc;
    } else {  // the pack kernel zeroes the unit cursor; without splats nobody did, and an earlier forward over the same
              // packed records has left it at the end
        GSTEX_CUDA_TRY(cudaMemsetAsync(ws_counters(workspace), 0, sizeof(uint32_t), st));
    }
    FwdParams p;
    p.g = make_raster_geom(img_height, img_width, block_width);
    p.C = channels;
    p.recs = n > 0 ? ws_recs(workspace) : nullptr;
    p.boxes = n > 0 ? ws_boxes(workspace, n) : nullptr;
    p.ids = gaussian_ids_sorted; p.tile_bins = reinterpret_cast<const int2*>(tile_bins);
    if (texture4 != nullptr && (channels > 3 || ((uintptr_t)texture4 & 15))) return GSTEX_E_BADARG;
    p.texture = texture; p.texture4 = reinterpret_cast<const float4*>(texture4); p.viewmat = viewmat; p.c2w = c2w; p.background = background;
    p.fx = fx; p.fy = fy; p.cx = cx; p.cy = cy; p.settings = settings;
    p.out_img = out_img; p.out_depth = out_depth; p.out_reg = out_reg; p.out_texture = out_texture;
    p.out_normal = out_normal; p.final_Ts = final_Ts; p.out_reg_s = out_reg_s; p.final_idx = final_idx;
    p.depth_idx = depth_idx;
    p.unit_cursor = ws_counters(workspace);
    p.tex_scale = tex_scale; p.tex_bias = tex_bias;
    p.order = (n > 0 && tile_bins != nullptr && num_tiles <= kOrderMaxTiles) ? ws_order(workspace, n) : nullptr;
    p.units_per_tile = p.g.threads / 32;
    p.num_units = p.g.tiles_x * p.g.tiles_y * p.units_per_tile;
    p.force_nocull = force_nocull_flag();
    p.dbg = g_forward_variant;
    // persistent grid: enough CTAs to fill every SM at the kernels' occupancy, never more than there are units
    int grid = sm_count() * kFwdCtasPerSm;
    if (grid > (p.num_units + kFwdWarps - 1) / kFwdWarps) grid = (p.num_units + kFwdWarps - 1) / kFwdWarps;
    const int threads = kFwdWarps * 32;
    // Record the pair stream for the stream backward when the caller gave a pool and the configuration is the
    // one that backward covers (<= 3 channels, no uv gradients); otherwise the backward recomputes.
    bool record = pair_stream != nullptr && gstex_pair_stream_supported(channels, n, settings) &&
                  make_stream_view(pair_stream, pair_stream_bytes, p.g.tiles_x * p.g.tiles_y, p.sv);
    if (pair_stream != nullptr && pair_stream_bytes >= sizeof(StreamHeader)) {
        GSTEX_CUDA_TRY(cudaMemsetAsync(pair_stream, 0, sizeof(StreamHeader), st));
        if (!record)  // tell the backward there is nothing to read
            GSTEX_CUDA_TRY(cudaMemsetAsync(reinterpret_cast<char*>(pair_stream) + offsetof(StreamHeader, overflow), 1, 1, st));
    }
    // The lane-dense kernel (raster_fwd_dense.cu; 16x16 tiles, <= 3 channels, no viewer bit) is bit-identical but measured
    // 2-15 % slower than this file's kernel on every BASELINE configuration: opt-in only.
    if (g_forward_variant >= 2 && forward_dense_supported(p)) return launch_forward_dense(p, record, st);
    // > 48 KB of shared memory: opt in (per device, cheap: done at every launch)
#define GSTEX_FWD_LAUNCH(CM, RC, VS)                                                                                \
    do {                                                                                                            \
        GSTEX_CUDA_TRY(cudaFuncSetAttribute(texture_forward_kernel<CM, RC, VS>,                                     \
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FwdShared)));  \
        texture_forward_kernel<CM, RC, VS><<<grid, threads, sizeof(FwdShared), st>>>(p);                            \
    } while (0)
    if (settings & GSTEX_SET_VIS_NORMALS) {  // never recorded (gstex_pair_stream_supported)
        if (channels <= 3) GSTEX_FWD_LAUNCH(3, false, true);
        else if (channels <= 6) GSTEX_FWD_LAUNCH(6, false, true);
        else GSTEX_FWD_LAUNCH(GSTEX_MAX_CHANNELS, false, true);
    } else if (record) GSTEX_FWD_LAUNCH(3, true, false);
    else if (channels <= 3) GSTEX_FWD_LAUNCH(3, false, false);
    else if (channels <= 6) GSTEX_FWD_LAUNCH(6, false, false);
    else GSTEX_FWD_LAUNCH(GSTEX_MAX_CHANNELS, false, false);
#undef GSTEX_FWD_LAUNCH
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

#define GSTEX_FWD_SIG_HEAD                                                                                            \
    int img_height, int img_width, int block_width, int channels, int n, int num_intersects,                         \
        const int32_t *texture_dims, const int32_t *gaussian_ids_sorted, const int32_t *tile_bins, const float *colors, \
        const float *opacities, const float *means, const float *scales, float glob_scale, const float *quats,       \
        const float *uv0, const float *umap, const float *vmap, const float *texture
#define GSTEX_FWD_SIG_TAIL                                                                                            \
    const float *viewmat, const float *c2w, float fx, float fy, float cx, float cy, int settings,                    \
        const float *background, float *out_img, float *out_depth, float *out_reg, float *out_texture,               \
        float *out_normal, float *final_Ts, int32_t *final_idx, int32_t *depth_idx, float *out_reg_s, void *workspace, \
        size_t workspace_bytes, void *pair_stream, size_t pair_stream_bytes, void *stream
#define GSTEX_FWD_ARGS_HEAD                                                                                           \
    img_height, img_width, block_width, channels, n, num_intersects, texture_dims, gaussian_ids_sorted, tile_bins,   \
        colors, opacities, means, scales, glob_scale, quats, uv0, umap, vmap, texture
#define GSTEX_FWD_ARGS_TAIL                                                                                           \
    viewmat, c2w, fx, fy, cx, cy, settings, background, out_img, out_depth, out_reg, out_texture, out_normal,        \
        final_Ts, final_idx, depth_idx, out_reg_s, workspace, workspace_bytes, pair_stream, pair_stream_bytes, stream

extern "C" int gstex_texture_forward(GSTEX_FWD_SIG_HEAD, GSTEX_FWD_SIG_TAIL) {
    return forward_impl(GSTEX_FWD_ARGS_HEAD, nullptr, 0, 1.f, 0.f, GSTEX_FWD_ARGS_TAIL);
}

extern "C" int gstex_texture_forward_act(GSTEX_FWD_SIG_HEAD, float tex_scale, float tex_bias, GSTEX_FWD_SIG_TAIL) {
    return forward_impl(GSTEX_FWD_ARGS_HEAD, nullptr, 0, tex_scale, tex_bias, GSTEX_FWD_ARGS_TAIL);
}

// The same forward with the texel fetch reading `texture4`, a (num_texels, 4) zero-padded copy of the (num_texels,
// channels <= 3) texture made by gstex_texture_pad4: four 128-bit loads per contributing pair instead of 4 * channels
// scalar ones.  Same arithmetic on the same values, so every output is bit-identical to gstex_texture_forward_act's.
extern "C" int gstex_texture_forward_padded(GSTEX_FWD_SIG_HEAD, const float* texture4, float tex_scale, float tex_bias,
                                            GSTEX_FWD_SIG_TAIL) {
    if (!texture4) return GSTEX_E_BADARG;
    return forward_impl(GSTEX_FWD_ARGS_HEAD, texture4, 0, tex_scale, tex_bias, GSTEX_FWD_ARGS_TAIL);
}

// The per-view packed records of gstex_texture_forward* as a call of their own, so that a multi-view caller can issue
// them early, on another stream, beside the compositing kernels of the view before (the pack is an HBM stream over
// all splats, ~65 us per view for 1 M splats; the compositing kernels are issue-bound).
extern "C" int gstex_texture_pack_records(int img_height, int img_width, int block_width, int n, const int32_t* texture_dims,
                                          const int32_t* tile_bins, const float* colors, const float* opacities,
                                          const float* means, const float* scales, float glob_scale, const float* quats,
                                          const float* uv0, const float* umap, const float* vmap, const float* viewmat,
                                          const float* c2w, float fx, float fy, float cx, float cy, int settings,
                                          void* workspace, size_t workspace_bytes, void* stream) {
    if (img_height <= 0 || img_width <= 0 || block_width < 2 || block_width > 16 || n < 0) return GSTEX_E_BADARG;
    if (settings & ~GSTEX_SETTINGS_SUPPORTED) return GSTEX_E_UNSUPPORTED;
    if (!workspace || !tile_bins || !viewmat || !c2w) return GSTEX_E_BADARG;
    if (workspace_bytes < ws_bytes(n)) return GSTEX_E_WORKSPACE;
    if (n == 0) return GSTEX_OK;
    if (!texture_dims || !colors || !opacities || !means || !scales || !quats || !uv0 || !umap || !vmap) return GSTEX_E_BADARG;
    const int num_tiles = ((img_width + block_width - 1) / block_width) * ((img_height + block_width - 1) / block_width);
    return launch_pack_records(n, texture_dims, colors, opacities, means, scales, glob_scale, quats, uv0, umap, vmap, viewmat,
                               c2w, fx, fy, cx, cy, settings, workspace, (cudaStream_t)stream, tile_bins, num_tiles);
}

// gstex_texture_forward_padded (`texture4` may be NULL here: scalar texel loads) over records that
// gstex_texture_pack_records has already written into `workspace` for exactly these inputs.
extern "C" int gstex_texture_forward_packed(GSTEX_FWD_SIG_HEAD, const float* texture4, float tex_scale, float tex_bias,
                                            GSTEX_FWD_SIG_TAIL) {
    return forward_impl(GSTEX_FWD_ARGS_HEAD, texture4, 1, tex_scale, tex_bias, GSTEX_FWD_ARGS_TAIL);
}

namespace gstex {
__global__ void __launch_bounds__(kThreads1D)
texture_pad4_kernel(long long num_texels, int C, const float* __restrict__ texture, float4* __restrict__ texture4) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= num_texels) return;
    const float* s = texture + (size_t)t * C;
    texture4[t] = make_float4(s[0], C > 1 ? s[1] : 0.f, C > 2 ? s[2] : 0.f, 0.f);
}
}  // namespace gstex

extern "C" int gstex_texture_pad4(long long num_texels, int channels, const float* texture, float* texture4, void* stream) {
    if (num_texels < 0 || channels < 1 || channels > 3) return GSTEX_E_BADARG;
    if (num_texels == 0) return GSTEX_OK;
    if (!texture || !texture4 || ((uintptr_t)texture4 & 15)) return GSTEX_E_BADARG;
    const long long blocks = (num_texels + kThreads1D - 1) / kThreads1D;
    if (blocks > 0x7fffffffll) return GSTEX_E_TOO_LARGE;
    gstex::texture_pad4_kernel<<<(unsigned)blocks, kThreads1D, 0, (cudaStream_t)stream>>>(
        num_texels, channels, texture, reinterpret_cast<float4*>(texture4));
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

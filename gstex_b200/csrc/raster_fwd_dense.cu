// raster_fwd_dense.cu -- stage (c), training configuration: LANE-DENSE textured forward compositing.
// Replaces texture_forward (gstex_cuda/cuda/csrc/texture.cu:11-329) for 16x16 tiles, <= 3 texture channels and no
// viewer bits; raster_fwd.cu keeps the general kernel (other block widths, more channels, settings bit 15).  sm_100a.
//
// The general kernel evaluates every surviving splat of a warp's 8x4 pixel block on all 32 lanes, although on
// average only ~10 of the 32 pixels lie inside the splat's support box and ~6 contribute: it was issue-bound on
// instructions spent on lanes that could not contribute.  alpha does not depend on the transmittance -- only the
// blend does -- so here the evaluation is PAIR-parallel, like the texture fetch already was:
//   phase A  the (survivor, pixel-inside-its-support-box) pairs of a staged batch are enumerated densely over the
//            lanes (prefix sum of the box-rectangle areas; a lane finds its survivor with one ballot, one
//            REDUX.OR and one shuffle) and evaluated with eval_pair; pairs with alpha >= 1/255 are compacted into
//            a per-warp ring, in (list position, pixel) order, with everything the blend needs (alpha, depth
//            terms, chart coordinates);
//   group    32 ring entries at a time: (i) lane-dense bilinear fetch, (ii) every pixel lane learns which entries
//            are its own (five ballots), (iii) walks only those, front to back -- transmittance, termination,
//            colour / normal / texture / median depth / distortion sums, in list order, so the sums are the
//            reference's bit for bit -- and (iv) the composited entries are appended lane-densely to the pair
//            stream the backward reads.
// Exactness of the culling: a dropped pair has alpha < 1/255 and is skipped by the reference too (texture.cu:213);
// it could only END a pixel whose transmittance lies in (1e-4, 1e-4 / (1 - 1/255)] (next_T <= 1e-4 is tested before
// `skip`, texture.cu:216-221).  A pixel entering that window is SUSPENDED at that list position; once the warp's
// other pixels are finished the suspended ones are resumed one list entry at a time with nothing dropped.
#include "raster_fwd.cuh"

namespace gstex {

constexpr int kDWarps = 5;       // warps per CTA (they share only the camera)
constexpr int kDCtasPerSm = 4;
constexpr int kDBatch = 16;      // survivors per record stage (one 128-byte TMA bulk copy each)
constexpr int kDQueue = 64;      // survivor ring: batch being evaluated + batch in flight + one culling round
constexpr int kDRing = 64;       // pair ring: < 32 pending + <= 32 new per phase-A step

struct DenseWarpShared {
    SplatRec rec[2 * kDBatch];   // 4 KB: two stages of kDBatch records (the resume path uses all 32 slots at once)
    int q_gid[kDQueue];          // splat id
    int q_idx[kDQueue];          // position in the global sorted list
    int q_rect[kDQueue];         // support box clipped to the unit: first pixel lane | width << 5 | pixel count << 9
    float pix[6][32];            // per pixel lane: ray xyz, pixel centre xy, view depth of the ray
    float4 e_a[kDRing];          // (alpha, negative = skipped pair that may still end the pixel; t; delta.x; delta.y)
    float4 e_b[kDRing];          // (delta.z; record slot | pixel lane << 5; T in front of the pair or -1 = not composited; t_view)
    float4 e_c[kDRing];          // (sample 0..2, depth term d of the distortion sum)
    uint64_t bar[2];
};

struct DenseShared {
    DenseWarpShared w[kDWarps];
    Camera cam;
};

__device__ unsigned long long g_dense_stats[8];  // dbg == 4: resume-path statistics (tests / tuning only)

// floor(i / w) for 0 <= i < 32, 1 <= w <= 8 as (i * m) >> 8
__device__ __forceinline__ int div_magic(int w) {
    return w == 1 ? 257 : (int)((0x21252B34415681ull >> (8 * (w - 2))) & 0xffull);
}

template <bool REC>
__global__ void __launch_bounds__(kDWarps * 32, kDCtasPerSm) texture_forward_dense_kernel(const FwdParams p) {
    extern __shared__ __align__(128) unsigned char dense_smem[];
    DenseShared& sm = *reinterpret_cast<DenseShared*>(dense_smem);
    const RasterGeom g = p.g;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    DenseWarpShared& ws = sm.w[warp];
    load_camera(sm.cam, p.viewmat, p.c2w, p.fx, p.fy, p.cx, p.cy);
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
    const unsigned full = 0xffffffffu;
    const unsigned lt_mask = (1u << lane) - 1u, le_mask = lt_mask | (1u << lane);
    const float3 origin = make_float3(sm.cam.c2w[3], sm.cam.c2w[7], sm.cam.c2w[11]);
    const float4 no_box = make_float4(1.f, 1.f, 0.f, 0.f);
    unsigned phase = 0;  // bit b: parity to wait for on bar[b]
    bool rec_ok = REC && (*(volatile uint32_t*)&p.sv.hdr->overflow == 0u);

    while (true) {
        int unit = 0;
        if (lane == 0) unit = (int)atomicAdd(p.unit_cursor, 1u);
        unit = __shfl_sync(full, unit, 0);
        if (unit >= p.num_units) break;
        const int tslot = unit >> 3, w = unit & 7;  // 16x16 tiles: 8 units of 8x4 pixels
        const int tile_id = p.order ? __ldg(p.order + tslot) : tslot;
        const int tile_x = tile_id % g.tiles_x, tile_y = tile_id / g.tiles_x;
        const int2 range = p.tile_bins[tile_id];
        // lane == pixel lane: x = lane & 7, y = lane >> 3 within the unit
        const Pixel pix = make_pixel_xy(sm.cam, g, tile_x, tile_y, ((w & 1) << 3) | (lane & 7), ((w >> 1) << 2) | (lane >> 3));
        const float UX = (float)(tile_x * 16 + ((w & 1) << 3)) + 0.5f, UY = (float)(tile_y * 16 + ((w >> 1) << 2)) + 0.5f;
        __syncwarp();
        ws.pix[0][lane] = pix.ray.x; ws.pix[1][lane] = pix.ray.y; ws.pix[2][lane] = pix.ray.z;
        ws.pix[3][lane] = pix.px; ws.pix[4][lane] = pix.py; ws.pix[5][lane] = pix.view_depth;
        __syncwarp();

        // ---- per-pixel state (this lane's own pixel)
        bool done = !pix.inside, susp = false;
        int susp_pos = 0;
        float T = 1.f;
        int cur_idx = 0, pix_depth_index = -1;
        float3 rgb = make_float3(0.f, 0.f, 0.f), nrm = make_float3(0.f, 0.f, 0.f), ts = make_float3(0.f, 0.f, 0.f);
        float pix_depth = 0.f, pix_reg = 0.f, tex0 = 0.f, tex1 = 0.f, tex2 = 0.f;
        unsigned done_mask = __ballot_sync(full, done);  // pixels that take no more pairs (finished or suspended)
        // live rectangle (unit-local pixel indices, inclusive) the support boxes are clipped to
        int lx0 = 0, lx1 = 7, ly0 = 0, ly1 = 3;
        auto shrink_live_rect = [&]() {
            const unsigned live = ~done_mask;
            const unsigned cols = (live | (live >> 8) | (live >> 16) | (live >> 24)) & 0xffu;
            if (cols == 0u) return;
            lx0 = __ffs(cols) - 1; lx1 = 31 - __clz(cols);
            ly0 = (__ffs(live) - 1) >> 3; ly1 = (31 - __clz(live)) >> 3;
        };
        shrink_live_rect();
        // pair-stream cursor of this unit (REC only): current chunk, items in it, items in total
        int cur_chunk = -1, fill = kChunkItems, cnt = 0;

        // ---- list walk: cursor = next list position to test; ring indices are monotonic
        int cursor = range.x;
        int q_head = 0, q_issue = 0, q_tail = 0;
        int r_head = 0, r_tail = 0;
        int cur = 0;
        int gid_pf, gid_pf2;
        float4 box_pf;
        auto prime = [&]() {  // ids two rounds ahead, support boxes one round ahead of the round being tested
            gid_pf = (cursor + lane < range.y) ? __ldg(p.ids + cursor + lane) : -1;
            gid_pf2 = (cursor + 32 + lane < range.y) ? __ldg(p.ids + cursor + 32 + lane) : -1;
            box_pf = (gid_pf >= 0) ? __ldg(p.boxes + gid_pf) : no_box;
        };
        // Test list entries until a batch of survivors is queued (or the list ends), then start the TMA copies of the
        // next kDBatch queued records into stage `buf`.  A survivor's support box is clipped to the rectangle of the
        // pixels still alive: that rectangle (first pixel lane, width, pixel count) is all phase A needs to enumerate
        // its pairs.  Returns the number of records in flight.
        auto fill_and_issue = [&](int buf) -> int {
            const bool nocull = p.force_nocull != 0;
            while (q_tail - q_issue < kDBatch && cursor < range.y) {
                const int gid = gid_pf;
                const float4 box = box_pf;
                const int pos = cursor + lane;
                cursor += 32;
                gid_pf = gid_pf2;
                gid_pf2 = (cursor + 32 + lane < range.y) ? __ldg(p.ids + cursor + 32 + lane) : -1;
                box_pf = (gid_pf >= 0) ? __ldg(p.boxes + gid_pf) : no_box;
                int xa = lx0, xb = lx1, ya = ly0, yb = ly1;
                if (!nocull) {  // pixel (x, y) of the unit has its centre at (UX + x, UY + y)
                    xa = max(xa, (int)ceilf(fmaxf(box.x - UX, -1.f)));
                    xb = min(xb, (int)floorf(fminf(box.z - UX, 8.f)));
                    ya = max(ya, (int)ceilf(fmaxf(box.y - UY, -1.f)));
                    yb = min(yb, (int)floorf(fminf(box.w - UY, 4.f)));
                }
                const bool hit = gid >= 0 && xa <= xb && ya <= yb;
                const unsigned m = __ballot_sync(full, hit);
                if (hit) {
                    const int slot = (q_tail + __popc(m & lt_mask)) & (kDQueue - 1);
                    const int wd = xb - xa + 1;
                    ws.q_gid[slot] = gid;
                    ws.q_idx[slot] = pos;
                    ws.q_rect[slot] = (ya * 8 + xa) | (wd << 5) | ((wd * (yb - ya + 1)) << 9);
                }
                q_tail += __popc(m);
            }
            __syncwarp();
            const int n_rec = min(q_tail - q_issue, kDBatch);
            if (n_rec > 0) {
                if (lane == 0) mbar_arrive_expect_tx(&ws.bar[buf], (uint32_t)n_rec * (uint32_t)sizeof(SplatRec));
                if (lane < n_rec)
                    bulk_copy_g2s(&ws.rec[buf * kDBatch + lane], p.recs + ws.q_gid[(q_issue + lane) & (kDQueue - 1)], sizeof(SplatRec),
                                  &ws.bar[buf]);
                q_issue += n_rec;
            }
            return n_rec;
        };

        // ---- the walk: ONE loop whose body is one of three steps (every step exists once in the code, so the
        // per-pixel state is loop-carried through a single back edge):
        //   GROUP    blend the oldest <= 32 ring entries            (>= 32 pending, or draining a finished batch)
        //   PRODUCE  evaluate 32 pairs and push the ones that matter (dense enumeration; or, resuming suspended
        //            pixels, one list entry with every lane its own pixel and nothing dropped)
        //   ADVANCE  next batch / switch to the resume path / finish
        int n_cur = 0, n_next = 0;
        int pa_base = 0, pa_total = 0, pa_word = 0, pa_off = 0;  // phase A progress over the batch's pairs
        bool setup = true, drain = false;
        bool slow = false;                    // resume path (see RESUME)
        int slow_base = 0, slow_start = 0;    // base of the round of 32 list entries being resumed
        unsigned slow_todo = 0u;              // suspended pixels still to evaluate against this round
        if (done_mask != full) {
            prime();
            n_cur = fill_and_issue(cur);
        }
        bool running = n_cur > 0;
        while (running) {
            if (setup) {  // begin the batch staged in rec[cur]: next batch in flight, pair offsets of this one
                setup = false;
                drain = false;
                n_next = fill_and_issue(cur ^ 1);
                mbar_wait(&ws.bar[cur], (phase >> cur) & 1u);
                phase ^= 1u << cur;
                int rect = 0, cnt_s = 0;
                if (lane < n_cur) {
                    rect = ws.q_rect[(q_head + lane) & (kDQueue - 1)];
                    cnt_s = rect >> 9;
                }
                int end_s = cnt_s;
#pragma unroll
                for (int o = 1; o < kDBatch; o <<= 1) {
                    const int up = __shfl_up_sync(full, end_s, o);
                    if (lane >= o) end_s += up;
                }
                pa_total = __shfl_sync(full, end_s, kDBatch - 1);
                pa_base = 0;
                pa_off = (lane < n_cur) ? end_s - cnt_s : 0x7fffffff;
                const int wd = (rect >> 5) & 15;
                // offset | first pixel lane << 10 | width << 15 | division magic << 19
                pa_word = (end_s - cnt_s) | ((rect & 31) << 10) | (wd << 15) | (div_magic(max(wd, 1)) << 19);
            }
            const int pending = r_tail - r_head;
            if (pending >= 32 || (drain && pending > 0)) {
                // ================= GROUP: m <= 32 ring entries, oldest first =================
                const int m = min(pending, 32);
                __syncwarp();
                const int hs = (r_head + lane) & (kDRing - 1);
                const bool valid = lane < m;
                float4 ea = make_float4(0.f, 0.f, 0.f, 0.f), eb = ea;
                if (valid) { ea = ws.e_a[hs]; eb = ws.e_b[hs]; }
                const int code = __float_as_int(eb.y);
                const int es = code & 31, epl = (code >> 5) & 31;
                // (i) lane-dense: chart coordinates, depth terms and the bilinear fetch of the pairs that can still
                // be composited
                float u = 0.f, v = 0.f, val0 = 0.f, val1 = 0.f, val2 = 0.f, dterm = 0.f, t_view = 0.f;
                if (valid && ea.x > 0.f && (slow || !((done_mask >> epl) & 1u))) {
                    const SplatRec& rr = ws.rec[cur * kDBatch + es];
                    const float3 delta = make_float3(ea.z, ea.w, eb.x);
                    u = chart_coord(rr.q[6], delta);
                    v = chart_coord(rr.q[7], delta);
                    t_view = __fmul_rn(ea.y, ws.pix[5][epl]);
                    dterm = use_ndc ? ndc_depth_ref(t_view) : ea.y;
                    const float4 r4 = rr.q[4];
                    const int start = __float_as_int(rr.q[5].w);
                    const TexQuery tq = tex_query(__float_as_int(r4.z), __float_as_int(r4.w), start, u, v, bilinear);
                    if (p.texture4 != nullptr) {  // padded layout: one 128-bit load per corner
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
                    // affine texel activation (SH2RGB of the caller folded into the fetch); (1, 0) leaves the sample as it is
                    val0 = __fmaf_rn(p.tex_scale, val0, p.tex_bias);
                    if (C > 1) val1 = __fmaf_rn(p.tex_scale, val1, p.tex_bias);
                    if (C > 2) val2 = __fmaf_rn(p.tex_scale, val2, p.tex_bias);
                }
                if (valid) {
                    ws.e_c[hs] = make_float4(val0, val1, val2, dterm);
                    ws.e_b[hs].w = t_view;
                }
                // (ii) which entries belong to this lane's pixel
                unsigned own = __ballot_sync(full, valid);
#pragma unroll
                for (int bit = 0; bit < 5; ++bit) {
                    const unsigned bal = __ballot_sync(full, valid && ((epl >> bit) & 1));
                    own &= ((lane >> bit) & 1) ? bal : ~bal;
                }
                __syncwarp();
                // (iii) owner walk, list order; straight-line body (no early exits: one loop-carried copy of the state)
                bool stop = done || (susp && !slow);
                while (own) {
                    const int j = __ffs(own) - 1;
                    own &= own - 1u;
                    const int slot = (r_head + j) & (kDRing - 1);
                    const float4 a4 = ws.e_a[slot], b4 = ws.e_b[slot];
                    const float a = a4.x;
                    const float next_T = __fmul_rn(T, __fsub_rn(1.f, fabsf(a)));
                    const bool term = !stop && next_T <= 1e-4f;  // this splat is not composited (texture.cu:216-221)
                    const bool take = !stop && !term && a > 0.f;
                    done = done || term;
                    if (take) {
                        const int s = __float_as_int(b4.y) & 31;
                        const SplatRec& r = ws.rec[cur * kDBatch + s];
                        const float4 r1 = r.q[1], r5 = r.q[5], c4 = ws.e_c[slot];
                        const int idx = ws.q_idx[(q_head + s) & (kDQueue - 1)];
                        const float vis = __fmul_rn(a, T);
                        rgb.x = __fmaf_rn(vis, r5.x, rgb.x); rgb.y = __fmaf_rn(vis, r5.y, rgb.y); rgb.z = __fmaf_rn(vis, r5.z, rgb.z);
                        nrm.x = __fmaf_rn(vis, r1.x, nrm.x); nrm.y = __fmaf_rn(vis, r1.y, nrm.y); nrm.z = __fmaf_rn(vis, r1.z, nrm.z);
                        tex0 = __fmaf_rn(vis, c4.x, tex0); tex1 = __fmaf_rn(vis, c4.y, tex1); tex2 = __fmaf_rn(vis, c4.z, tex2);
                        if (T > 0.5f) {  // median depth, texture.cu:286-291
                            pix_depth = b4.w;
                            pix_depth_index = idx;
                        }
                        distortion_ref(c4.w, vis, ts, pix_reg);  // distortion(), helpers.cuh:259-264
                        ws.e_b[slot].z = T;  // composited: a pair-stream item with this transmittance in front of it
                        T = next_T;
                        cur_idx = idx;
                        // From here a pair with alpha < 1/255 -- which PRODUCE drops -- could end this pixel: suspend it
                        // at this list position; it is resumed with nothing dropped once the others are finished.
                        if (!slow && next_T <= 1.0041e-4f) { susp = true; susp_pos = idx; }
                    }
                    stop = stop || term || (susp && !slow);
                }
                __syncwarp();
                // (iv) the composited entries, in ring order, become consecutive items of the unit's pair stream
                if (REC && rec_ok) {
                    const float Tb = valid ? ws.e_b[hs].z : -1.f;
                    const bool alive = Tb >= 0.f;
                    const unsigned am = __ballot_sync(full, alive);
                    const int na = __popc(am);
                    if (na > 0) {
                        int new_chunk = cur_chunk;
                        if (fill + na > kChunkItems) {
                            int c = 0;
                            if (lane == 0) c = (int)atomicAdd(&p.sv.hdr->next_chunk, 1u);
                            c = __shfl_sync(full, c, 0);
                            if (c >= p.sv.num_chunks) {
                                if (lane == 0) p.sv.hdr->overflow = 1u;
                                rec_ok = false;
                            } else {
                                if (lane == 0) p.sv.chunk_prev[c] = cur_chunk;
                                new_chunk = c;
                            }
                        }
                        if (rec_ok) {
                            if (alive) {
                                int pos = fill + __popc(am & lt_mask);
                                int ch = cur_chunk;
                                if (pos >= kChunkItems) { pos -= kChunkItems; ch = new_chunk; }
                                float4* dst = reinterpret_cast<float4*>(p.sv.items + (size_t)ch * kChunkItems + pos);
                                const int q = (q_head + es) & (kDQueue - 1);
                                const uint32_t gl = ((uint32_t)ws.q_gid[q] << 5) | (uint32_t)epl;
                                dst[0] = make_float4(__uint_as_float(gl), __int_as_float(ws.q_idx[q]), Tb, u);
                                dst[1] = make_float4(v, val0, val1, val2);
                            }
                            fill += na;
                            if (fill > kChunkItems) { fill -= kChunkItems; cur_chunk = new_chunk; }
                            cnt += na;
                        }
                    }
                }
                done_mask = __ballot_sync(full, done || (susp != slow));
                r_head += m;
                if (done_mask == full) {  // nobody takes pairs any more: drop what is queued, end the batch
                    r_head = r_tail;
                    pa_base = pa_total;
                    drain = true;
                }
                __syncwarp();
            } else if (slow) {
                // ================= RESUME: 32 list entries (one per lane) against ONE suspended pixel =================
                // A suspended pixel (transmittance within 255/254 of the termination threshold, where even a pair with
                // alpha < 1/255 could end it) re-walks the list from the entry after the one that suspended it, and
                // nothing may be culled or dropped for it.  alpha of different entries is independent, so the walk is
                // entry-parallel: every lane evaluates its own list entry for the pixel, straight from the packed
                // records in global memory; entries that cannot change anything (1 - alpha rounds to 1: next_T == T) are
                // left out, the rest go through the ring and GROUP like any other pair.  ~1 pixel in 200.
                drain = false;
                if (slow_todo == 0u) {
                    slow_base += 32;
                    slow_todo = __ballot_sync(full, susp && !done);
                    if (slow_base >= range.y || slow_todo == 0u) {
                        running = false;
                        continue;
                    }
                }
                const int px = __ffs(slow_todo) - 1;
                slow_todo &= slow_todo - 1u;
                const int pos = slow_base + lane;
                const bool in_list = pos < range.y;
                const int gid = in_list ? __ldg(p.ids + pos) : __ldg(p.ids + range.y - 1);
                const float4* src = p.recs[gid].q;
                const float4 r0 = __ldg(src), r1 = __ldg(src + 1), r2 = __ldg(src + 2), r3 = __ldg(src + 3), r4 = __ldg(src + 4);
                const int behind = __shfl_sync(full, susp_pos, px);
                Pixel pe;
                pe.origin = origin;
                pe.ray = make_float3(ws.pix[0][px], ws.pix[1][px], ws.pix[2][px]);
                pe.px = ws.pix[3][px]; pe.py = ws.pix[4][px];
                const PairEval e = eval_pair(pe, r0, r1, r2, r3, r4, use_blur);
                const bool small = e.alpha < 1.f / 255.f;
                const bool keep = in_list && pos > behind && (!small || __fsub_rn(1.f, e.alpha) < 1.f);
                const unsigned km = __ballot_sync(full, keep);
                if (km != 0u) {
                    if (keep) {  // GROUP reads the record, the splat id and the list position of slot `lane`
                        SplatRec& dst = ws.rec[lane];
#pragma unroll
                        for (int k = 0; k < 8; ++k) dst.q[k] = __ldg(src + k);  // (reloaded: L1 hits; keeps the step's registers few)
                        ws.q_gid[lane] = gid;
                        ws.q_idx[lane] = pos;
                        const bool skip = small || (e.t < kNear) || (e.t > kFar);
                        const int slot = (r_tail + __popc(km & lt_mask)) & (kDRing - 1);
                        ws.e_a[slot] = make_float4(skip ? -e.alpha : e.alpha, e.t, e.delta.x, e.delta.y);
                        ws.e_b[slot] = make_float4(e.delta.z, __int_as_float(lane | (px << 5)), -1.f, 0.f);
                    }
                    r_tail += __popc(km);
                    drain = true;  // blend them before the slots are reused
                }
            } else if (!drain) {
                // ================= PRODUCE: one pair per lane =================
                // dense enumeration of the batch's (survivor, pixel inside its clipped support box) pairs
                const int j = pa_base + lane;
                const bool starts_here = pa_off >= pa_base && pa_off < pa_base + 32;
                const unsigned startmask = __reduce_or_sync(full, starts_here ? (1u << (pa_off - pa_base)) : 0u);
                const int nbefore = __popc(__ballot_sync(full, pa_off < pa_base));
                const int s = min(max(nbefore + __popc(startmask & le_mask) - 1, 0), n_cur - 1);
                const int wv = __shfl_sync(full, pa_word, s);
                const int i = j - (wv & 1023);
                const int yy = (i * ((wv >> 19) & 511)) >> 8;
                int pl = ((wv >> 10) & 31) + yy * 8 + (i - yy * ((wv >> 15) & 15));
                const bool active = j < pa_total && (unsigned)pl < 32u && !((done_mask >> (pl & 31)) & 1u);
                pl = active ? pl : 0;
                pa_base += 32;
                drain = pa_base >= pa_total;
                const SplatRec& r = ws.rec[cur * kDBatch + s];
                const float4 r0 = r.q[0], r1 = r.q[1], r2 = r.q[2], r3 = r.q[3], r4 = r.q[4];
                Pixel pe;
                pe.origin = origin;
                pe.ray = make_float3(ws.pix[0][pl], ws.pix[1][pl], ws.pix[2][pl]);
                pe.px = ws.pix[3][pl]; pe.py = ws.pix[4][pl];
                const PairEval e = eval_pair(pe, r0, r1, r2, r3, r4, use_blur);
                // a dropped pair has alpha < 1/255 (skipped by the reference too); see RESUME for the one case it matters
                const bool keep = active && e.alpha >= 1.f / 255.f;
                const unsigned km = __ballot_sync(full, keep);
                if (keep) {
                    const bool skip = (e.t < kNear) || (e.t > kFar);
                    const int slot = (r_tail + __popc(km & lt_mask)) & (kDRing - 1);
                    ws.e_a[slot] = make_float4(skip ? -e.alpha : e.alpha, e.t, e.delta.x, e.delta.y);
                    ws.e_b[slot] = make_float4(e.delta.z, __int_as_float(s | (pl << 5)), -1.f, 0.f);
                }
                r_tail += __popc(km);
            } else {
                // ================= ADVANCE: the batch is evaluated and blended =================
                const bool all_done = done_mask == full;
                if (all_done && n_next > 0) {  // drain the copies in flight before their stage and barrier are reused
                    mbar_wait(&ws.bar[cur ^ 1], (phase >> (cur ^ 1)) & 1u);
                    phase ^= 1u << (cur ^ 1);
                }
                q_head += n_cur;
                n_cur = all_done ? 0 : n_next;
                cur ^= 1;
                if (n_cur == 0 && p.dbg != 3 && __any_sync(full, susp && !done)) {
                    // the main walk is over: resume the suspended pixels (RESUME above)
                    slow = true;
                    cur = 0;
                    q_head = 0;
                    slow_start = __reduce_min_sync(full, (susp && !done) ? susp_pos + 1 : 0x7fffffff);
                    slow_base = slow_start - 32;
                    slow_todo = 0u;
                    drain = false;
                    if (p.dbg == 4 && lane == 0) {
                        atomicAdd(&g_dense_stats[0], 1ull);
                        atomicAdd(&g_dense_stats[2], (unsigned long long)(range.y - slow_start));
                        atomicMax(&g_dense_stats[3], (unsigned long long)(range.y - slow_start));
                    }
                    continue;
                }
                if (n_cur > 0) {
                    shrink_live_rect();
                    setup = true;
                } else {
                    running = false;
                }
            }
        }

        if (p.dbg == 4 && slow && lane == 0) {
            atomicAdd(&g_dense_stats[4], (unsigned long long)(min(slow_base + 32, range.y) - slow_start));  // entries walked
            atomicMax(&g_dense_stats[5], (unsigned long long)(min(slow_base + 32, range.y) - slow_start));
        }
        if (REC) {
            const int unit0 = tile_id * kUnitsPerTile;
            if (lane == 0) p.sv.units[unit0 + w] = make_int2(cur_chunk, rec_ok ? cnt : 0);
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
            p.out_texture[(size_t)C * id] = tex0;
            if (C > 1) p.out_texture[(size_t)C * id + 1] = tex1;
            if (C > 2) p.out_texture[(size_t)C * id + 2] = tex2;
        }
        __syncwarp();
    }
}

extern "C" int gstex_debug_dense_stats(unsigned long long* out, int reset) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, g_dense_stats, sizeof(unsigned long long) * 8);
    if (reset) {
        unsigned long long z[8] = {0};
        cudaMemcpyToSymbol(g_dense_stats, z, sizeof(z));
    }
    return 0;
}

bool forward_dense_supported(const FwdParams& p) {
    return p.g.bw == 16 && p.C >= 1 && p.C <= 3 && !(p.settings & GSTEX_SET_VIS_NORMALS);
}

int launch_forward_dense(const FwdParams& p, bool record, cudaStream_t st) {
    int grid = sm_count() * kDCtasPerSm;
    if (grid > (p.num_units + kDWarps - 1) / kDWarps) grid = (p.num_units + kDWarps - 1) / kDWarps;
    if (grid < 1) grid = 1;
    const int threads = kDWarps * 32;
    if (record) {
        GSTEX_CUDA_TRY(cudaFuncSetAttribute(texture_forward_dense_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)sizeof(DenseShared)));
        texture_forward_dense_kernel<true><<<grid, threads, sizeof(DenseShared), st>>>(p);
    } else {
        GSTEX_CUDA_TRY(cudaFuncSetAttribute(texture_forward_dense_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)sizeof(DenseShared)));
        texture_forward_dense_kernel<false><<<grid, threads, sizeof(DenseShared), st>>>(p);
    }
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

}  // namespace gstex

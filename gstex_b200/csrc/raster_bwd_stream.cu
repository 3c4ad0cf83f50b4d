// raster_bwd_stream.cu -- stage (d), fast path: backward from the pair stream the forward recorded.
// Replaces texture_backward (gstex_cuda/cuda/csrc/texture.cu:331-760, wrapper :903-1053) for the training
// configuration (<= 3 texture channels, no uv gradients); raster_bwd.cu remains the general kernel and the
// fallback when the stream pool overflowed.  sm_100a.
//
// The reference (and raster_bwd.cu) re-walk every tile list back to front with one thread per pixel: every
// (pixel, splat) pair of the list is re-evaluated and the ~450-instruction gradient math runs with ~7 of 32
// lanes active, because a splat only touches a few pixels of a warp.  Here a warp owns one unit (32 pixels)
// and reads its recorded pairs back 32 at a time, ONE PAIR PER LANE, whatever pixel the pair belongs to:
//   * per-pixel constants come from a small shared-memory table indexed by the pair's pixel lane;
//   * the pair is re-evaluated with the very code the forward used (eval_pair), so alpha is bit-identical;
//   * the only sequential quantity, the running d(loss)/d(T) of a pixel (texture.cu:650-670), lives in shared
//     memory and is advanced splat by splat (a batch holds ~4 splats; pairs of one splat touch distinct pixels);
//   * per-splat gradients are summed with a segmented shuffle reduction over the lanes holding the same splat
//     and leave the warp as one atomicAdd per (splat, value); texel gradients are the 4*C atomics of the
//     reference, now issued by full warps.
#include "bwd_stream.cuh"

namespace gstex {


constexpr int kStreamWarps = 8;

// 16-byte vector reduction (sm_90+): one L2 atomic transaction for four floats
__device__ __forceinline__ void red_add_v4(float4* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__device__ __forceinline__ void prefetch_l1(const void* ptr) { asm volatile("prefetch.global.L1 [%0];" ::"l"(ptr)); }
// per-pixel table, structure of arrays: field f of pixel lane l at tab[f][l]
enum PixField {
    F_RAYX, F_RAYY, F_RAYZ, F_PX, F_PY, F_VDEPTH, F_VOUTX, F_VOUTY, F_VOUTZ, F_VNX, F_VNY, F_VNZ, F_VT0, F_VT1, F_VT2,
    F_VDEP, F_VREG, F_TS0, F_TS1, F_TS2, F_DFINAL, F_VTRUN, F_COUNT
};

constexpr int kRecQuads = 8;  // staged per splat: SplatRec quads 0..5 + SplatRecBwd (quat, aux)

struct StreamShared {
    Camera cam;
    float tab[kStreamWarps][F_COUNT][32];
    // records of the splats of the current / next batch, one slot per segment (run of lanes holding the same
    // splat), filled one batch ahead with cp.async by the segment's head lane
    float4 rs[kStreamWarps][2][32][kRecQuads];
};

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all_but_one() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }

__global__ void __launch_bounds__(kStreamWarps * 32, 2) texture_backward_stream_kernel(const StreamBwdParams p) {
    extern __shared__ __align__(16) unsigned char stream_smem[];
    StreamShared& sm = *reinterpret_cast<StreamShared*>(stream_smem);
    if (*(volatile uint32_t*)&p.sv.hdr->overflow != 0u) return;  // the recompute kernel handles this view
    const RasterGeom g = p.g;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    load_camera(sm.cam, p.viewmat, p.c2w, p.fx, p.fy, p.cx, p.cy);
    __syncthreads();
    const bool use_blur = p.settings & GSTEX_SET_BLUR;
    const bool use_ndc = p.settings & GSTEX_SET_NDC_REG;
    const bool bilinear = !(p.settings & GSTEX_SET_NEAREST);
    const int C = p.C;
    const float3 origin = make_float3(sm.cam.c2w[3], sm.cam.c2w[7], sm.cam.c2w[11]);
    float (*tab)[32] = sm.tab[warp];
    float4 (*rs)[32][kRecQuads] = sm.rs[warp];
    const float bg0 = __ldg(p.background), bg1 = __ldg(p.background + 1), bg2 = __ldg(p.background + 2);

    const int32_t* order = (p.order_valid != nullptr && __ldg(p.order_valid) != 0u) ? p.order : nullptr;

    while (true) {
        int unit = 0;
        if (lane == 0) unit = (int)atomicAdd(&p.sv.hdr->next_unit, 1u);
        unit = __shfl_sync(0xffffffffu, unit, 0);
        if (unit >= p.sv.num_units) break;
        if (order) unit = __ldg(order + unit / kUnitsPerTile) * kUnitsPerTile + unit % kUnitsPerTile;
        const int2 uc = p.sv.units[unit];
        const int cnt = uc.y;
        if (cnt <= 0) continue;
        const int tile_id = unit / kUnitsPerTile, w = unit % kUnitsPerTile;
        const int tile_x = tile_id % g.tiles_x, tile_y = tile_id / g.tiles_x;
        __syncwarp();
        {  // this lane's own pixel -> table column `lane`
            int lx, ly;
            unit_pixel_xy(g, w, lane, lx, ly);  // the forward's lane -> pixel mapping
            const Pixel pix = make_pixel_xy(sm.cam, g, tile_x, tile_y, lx, ly);
            const int id = pix.pix_id;
            tab[F_RAYX][lane] = pix.ray.x; tab[F_RAYY][lane] = pix.ray.y; tab[F_RAYZ][lane] = pix.ray.z;
            tab[F_PX][lane] = pix.px; tab[F_PY][lane] = pix.py; tab[F_VDEPTH][lane] = pix.view_depth;
            // a NULL upstream gradient is a zero gradient (autograd's None): nothing is read for it
            const bool has_o = p.v_output != nullptr, has_n = p.v_output_normal != nullptr, has_t = p.v_output_texture != nullptr;
            const float vo0 = has_o ? p.v_output[3 * id] : 0.f, vo1 = has_o ? p.v_output[3 * id + 1] : 0.f,
                        vo2 = has_o ? p.v_output[3 * id + 2] : 0.f;
            tab[F_VOUTX][lane] = vo0; tab[F_VOUTY][lane] = vo1; tab[F_VOUTZ][lane] = vo2;
            tab[F_VNX][lane] = has_n ? p.v_output_normal[3 * id] : 0.f;
            tab[F_VNY][lane] = has_n ? p.v_output_normal[3 * id + 1] : 0.f;
            tab[F_VNZ][lane] = has_n ? p.v_output_normal[3 * id + 2] : 0.f;
            tab[F_VT0][lane] = has_t ? p.v_output_texture[(size_t)C * id] : 0.f;
            tab[F_VT1][lane] = (has_t && C > 1) ? p.v_output_texture[(size_t)C * id + 1] : 0.f;
            tab[F_VT2][lane] = (has_t && C > 2) ? p.v_output_texture[(size_t)C * id + 2] : 0.f;
            tab[F_VDEP][lane] = p.v_output_depth ? p.v_output_depth[id] : 0.f;
            tab[F_VREG][lane] = p.v_output_reg ? p.v_output_reg[id] : 0.f;
            tab[F_TS0][lane] = p.final_s[3 * id]; tab[F_TS1][lane] = p.final_s[3 * id + 1];
            tab[F_TS2][lane] = p.final_s[3 * id + 2];
            tab[F_DFINAL][lane] = __int_as_float(p.depth_idx[id]);
            // seed of the d(loss)/d(T) recurrence, texture.cu:470-475
            tab[F_VTRUN][lane] = bg0 * vo0 + bg1 * vo1 + bg2 * vo2 - (p.v_output_alpha ? p.v_output_alpha[id] : 0.f);
        }
        __syncwarp();

        // Item loads run two batches ahead of the batch being processed and the records of the next batch are
        // pulled into L1 one batch ahead, so the HBM/L2 latency of this gather chain overlaps the math.
        int load_b = (cnt - 1) >> 5, load_chunk = uc.x;
        float4 n0[2], n1[2];
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            n0[k] = make_float4(0.f, __int_as_float(-1 - lane), 0.f, 0.f);
            n1[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        auto fetch = [&](float4& a, float4& c) {
            a = make_float4(0.f, __int_as_float(-1 - lane), 0.f, 0.f);
            if (load_b < 0) return;
            const int i = (load_b << 5) + lane;
            if (i < cnt) {
                const float4* src = reinterpret_cast<const float4*>(p.sv.items + (size_t)load_chunk * kChunkItems +
                                                                    (i & (kChunkItems - 1)));
                a = __ldcs(src);
                c = __ldcs(src + 1);
            }
            if ((load_b & 3) == 0 && load_b > 0) load_chunk = __ldg(p.sv.chunk_prev + load_chunk);
            --load_b;
        };
        // Segments = runs of lanes holding the same splat (same sorted-list position).  The segment structure of a
        // batch is worked out one batch ahead, when its items are already in registers, and each segment's head
        // lane starts the copy of the splat's record into the slot of that segment (cp.async, 8 x 16 B), so the
        // gather latency overlaps the previous batch's math instead of stalling the first use of the record.
        auto stage = [&](const float4& it0, int buf, int& seg_o, unsigned& heads_o) {
            const int idx_n = __float_as_int(it0.y);  // invalid lanes: distinct negative values (own segments)
            const int idx_p = __shfl_up_sync(0xffffffffu, idx_n, 1);
            const bool head_n = (lane == 0) || (idx_n != idx_p);
            heads_o = __ballot_sync(0xffffffffu, head_n);
            seg_o = __popc(heads_o & (0xffffffffu >> (31 - lane))) - 1;
            if (head_n && idx_n >= 0) {
                const uint32_t g_n = __float_as_uint(it0.x) >> 5;
                const SplatRec* src = p.recs + g_n;
                const SplatRecBwd* srb = p.recs_bwd + g_n;
                float4* dst = rs[buf][seg_o];
#pragma unroll
                for (int k = 0; k < 6; ++k) cp_async16(dst + k, &src->q[k]);
                cp_async16(dst + 6, &srb->quat);
                cp_async16(dst + 7, &srb->aux);
            }
            cp_async_commit();
        };
        fetch(n0[0], n1[0]);
        fetch(n0[1], n1[1]);
        int buf = 0, seg = 0, seg_nxt = 0;
        unsigned heads = 0u, heads_nxt = 0u;
        __syncwarp();  // the previous unit's reads of the record slots are done
        stage(n0[0], buf, seg, heads);
        for (int b = (cnt - 1) >> 5; b >= 0; --b) {
            const int i = (b << 5) + lane;
            const bool valid = i < cnt;
            const float4 w0 = n0[0], w1 = n1[0];
            n0[0] = n0[1]; n1[0] = n1[1];
            if (b > 0) stage(n0[0], buf ^ 1, seg_nxt, heads_nxt);  // records of batch b-1
            else cp_async_commit();                                 // keeps the group count uniform
            fetch(n0[1], n1[1]);
            const uint32_t gl = __float_as_uint(w0.x);
            const int pl = (int)(gl & 31u), gid = (int)(gl >> 5);
            const int idx = __float_as_int(w0.y);
            const float T = w0.z, u = w0.w, vv = w1.x;
            const bool head = (heads >> lane) & 1u;
            const unsigned later = (lane == 31) ? 0u : (heads >> (lane + 1));
            const int seg_end = later ? lane + __ffs(later) - 1 : 31;
            cp_async_wait_all_but_one();  // this batch's records have landed (the next batch's may be in flight)
            __syncwarp();
            const float4* rq = rs[buf][valid ? seg : 0];

            float v[13];
#pragma unroll
            for (int k = 0; k < 13; ++k) v[k] = 0.f;
            // values the sequential step needs
            float alpha = 0.f, v_vis = 0.f, v_weight = 0.f, v_alpha = 0.f;
            // record + pixel (gathered; lanes of one segment read the same record)
            const float4 r0 = rq[0], r1 = rq[1], r2 = rq[2], r3 = rq[3], r4 = rq[4], r5 = rq[5];
            Pixel pix;
            pix.origin = origin;
            pix.ray = make_float3(tab[F_RAYX][pl], tab[F_RAYY][pl], tab[F_RAYZ][pl]);
            pix.px = tab[F_PX][pl]; pix.py = tab[F_PY][pl]; pix.view_depth = tab[F_VDEPTH][pl];
            const PairEval e = eval_pair(pix, r0, r1, r2, r3, r4, use_blur);
            const float3 v_out = make_float3(tab[F_VOUTX][pl], tab[F_VOUTY][pl], tab[F_VOUTZ][pl]);
            const float3 v_out_normal = make_float3(tab[F_VNX][pl], tab[F_VNY][pl], tab[F_VNZ][pl]);
            const float v_tex0 = tab[F_VT0][pl], v_tex1 = tab[F_VT1][pl], v_tex2 = tab[F_VT2][pl];
            const float v_out_reg = tab[F_VREG][pl];
            const float3 ts_final = make_float3(tab[F_TS0][pl], tab[F_TS1][pl], tab[F_TS2][pl]);
            const float t_view = e.t * pix.view_depth;
            float v_t = 0.f, v_t_view = 0.f, vis = 0.f;
            if (valid) {
                alpha = e.alpha;
                vis = alpha * T;
                v[0] = vis * v_out.x; v[1] = vis * v_out.y; v[2] = vis * v_out.z;
                v_vis = r5.x * v_out.x + r5.y * v_out.y + r5.z * v_out.z;
                v_vis += r1.x * v_out_normal.x + r1.y * v_out_normal.y + r1.z * v_out_normal.z;
                {   // texel gradients, query_jagged_chart_vjp (texture_helpers.cuh:252-260); sample values recorded
                    const TexQuery q = tex_query(__float_as_int(r4.z), __float_as_int(r4.w), __float_as_int(r5.w), u, vv, bilinear);
                    const float vt[3] = {v_tex0, v_tex1, v_tex2};
                    const float val[3] = {w1.y, w1.z, w1.w};
#pragma unroll
                    for (int k = 0; k < 3; ++k) v_vis += val[k] * vt[k];
                    // one 16-byte vector reduction per corner into the padded (T,4) accumulator instead of C scalar
                    // atomics; corners that coincide (replicate padding, 1x1 charts) are merged, zero weights skipped
                    float c00 = q.w00, c01 = q.w01, c10 = q.w10, c11 = q.w11;
                    if (q.b01 == q.b00) { c00 += c01; c01 = 0.f; }
                    if (q.b11 == q.b10) { c10 += c11; c11 = 0.f; }
                    if (q.b10 == q.b00) { c00 += c10; c10 = 0.f; }
                    if (q.b11 == q.b01) { c01 += c11; c11 = 0.f; }
                    const float g0 = vis * vt[0], g1 = vis * vt[1], g2 = vis * vt[2];
#ifndef GSTEX_EXP_NO_TEXEL_RED  // experiment only (scripts/bwd_time.py): upper bound of what aggregating these could save
                    if (c00 != 0.f) red_add_v4(p.tex_acc + q.b00, c00 * g0, c00 * g1, c00 * g2, 0.f);
                    if (c01 != 0.f) red_add_v4(p.tex_acc + q.b01, c01 * g0, c01 * g1, c01 * g2, 0.f);
                    if (c10 != 0.f) red_add_v4(p.tex_acc + q.b10, c10 * g0, c10 * g1, c10 * g2, 0.f);
                    if (c11 != 0.f) red_add_v4(p.tex_acc + q.b11, c11 * g0, c11 * g1, c11 * g2, 0.f);
#endif
                }
                {   // distortion_vjp, helpers.cuh:266-269, with the FINAL sums
                    const float d = use_ndc ? div_normal(kFar * t_view - kFar * kNear, (kFar - kNear) * t_view) : e.t;
                    const float v_d = 2.f * (vis * d * ts_final.x - vis * ts_final.y) * v_out_reg;
                    v_weight = (d * d * ts_final.x - 2.f * d * ts_final.y + ts_final.z) * v_out_reg;
                    if (use_ndc) v_t_view += div_normal(kFar * kNear, (kFar - kNear) * t_view * t_view) * v_d;
                    else v_t += v_d;
                }
            }
            // the running d(loss)/d(T) of each pixel, back to front = segment by segment from the last one
            // (texture.cu:650-670); pairs of one segment belong to distinct pixels
            // Only pairs of the SAME pixel must be ordered: rank them within their pixel (later lane = later in the
            // list = processed first) and take one step per rank instead of one per segment.
            {
                // lanes holding a pair of the same pixel: five ballots (MATCH.ANY stalled this point for ~5 % of the
                // kernel's samples)
                unsigned same = __ballot_sync(0xffffffffu, valid);
#pragma unroll
                for (int bit = 0; bit < 5; ++bit) {
                    const bool on = (pl >> bit) & 1;
                    const unsigned bal = __ballot_sync(0xffffffffu, on);
                    same &= on ? bal : ~bal;
                }
                if (!valid) same = 1u << lane;
                const int rank = __popc(lane == 31 ? 0u : (same >> (lane + 1)));
                const int steps = __reduce_max_sync(0xffffffffu, rank) + 1;
                for (int s = 0; s < steps; ++s) {
                    if (rank == s && valid) {
                        const float run = tab[F_VTRUN][pl];
                        v_alpha = T * v_vis - T * run;
                        float v_T_cur = alpha * v_vis + (1.f - alpha) * run;
                        v_alpha += v_weight * T;
                        v_T_cur += v_weight * alpha;
                        tab[F_VTRUN][pl] = v_T_cur;
                    }
                    __syncwarp();
                }
            }
            if (valid) {
                const float4 rbq = rq[6], rba = rq[7];
                const float3 v_normal_local = make_float3(vis * v_out_normal.x, vis * v_out_normal.y, vis * v_out_normal.z);
                // texture.cu:672-707
                const float v_sigma = e.blur ? 0.f : -r1.w * e.ex * v_alpha;
                const float v_sigma_blur = e.blur ? -r1.w * e.ex * v_alpha : 0.f;
                const float v_l1 = r2.w * e.l1 * v_sigma, v_l2 = r3.w * e.l2 * v_sigma;
                v[7] = -(r2.w * e.l1 * e.l1) * v_sigma * rba.x;
                v[8] = -(r3.w * e.l2 * e.l2) * v_sigma * rba.y;
                v_t += v_l1 * (pix.ray.x * r2.x + pix.ray.y * r2.y + pix.ray.z * r2.z) +
                       v_l2 * (pix.ray.x * r3.x + pix.ray.y * r3.y + pix.ray.z * r3.z);
                const int depth_final = __float_as_int(tab[F_DFINAL][pl]);
                if (idx == depth_final && depth_final != -1) v_t_view += tab[F_VDEP][pl];
                v_t += pix.view_depth * v_t_view;
                float3 v_mean_extra = make_float3(0.f, 0.f, 0.f);
                if (e.blur) {  // project_pix_vjp + transform_4x3_rot_only_transposed, texture.cu:685-692
                    const float vxp = sm.cam.fx * (2.0f * v_sigma_blur * (r4.x - pix.px));
                    const float vyp = sm.cam.fy * (2.0f * v_sigma_blur * (r4.y - pix.py));
                    const float rw = rba.z;
                    const float nx = (r4.x - sm.cam.cx) / sm.cam.fx, ny = (r4.y - sm.cam.cy) / sm.cam.fy;
                    const float wx = vxp * rw, wy = vyp * rw, wz = -(vxp * nx + vyp * ny) * rw;
                    v_mean_extra = make_float3(sm.cam.vm[0] * wx + sm.cam.vm[4] * wy + sm.cam.vm[8] * wz,
                                               sm.cam.vm[1] * wx + sm.cam.vm[5] * wy + sm.cam.vm[9] * wz,
                                               sm.cam.vm[2] * wx + sm.cam.vm[6] * wy + sm.cam.vm[10] * wz);
                }
                // compute_tt_vjp, texture_helpers.cuh:315-334
                const float den = clamp_tt_denom(r1.x * pix.ray.x + r1.y * pix.ray.y + r1.z * pix.ray.z);
                const float vt_den = div_normal(v_t, den);
                const float3 v_normal_tt = make_float3((r0.x - pix.origin.x - pix.ray.x * e.t) * vt_den,
                                                       (r0.y - pix.origin.y - pix.ray.y * e.t) * vt_den,
                                                       (r0.z - pix.origin.z - pix.ray.z * e.t) * vt_den);
                v[4] = -(r2.x * v_l1 + r3.x * v_l2) + r1.x * vt_den + v_mean_extra.x;
                v[5] = -(r2.y * v_l1 + r3.y * v_l2) + r1.y * vt_den + v_mean_extra.y;
                v[6] = -(r2.z * v_l1 + r3.z * v_l2) + r1.z * vt_den + v_mean_extra.z;
                {  // quat_to_rotmat_vjp, helpers.cuh:187-228
                    const float qw = rbq.x, qx = rbq.y, qy = rbq.z, qz = rbq.w;
                    const float c00 = e.delta.x * v_l1, c01 = e.delta.y * v_l1, c02 = e.delta.z * v_l1;
                    const float c10 = e.delta.x * v_l2, c11 = e.delta.y * v_l2, c12 = e.delta.z * v_l2;
                    const float c20 = v_normal_tt.x + v_normal_local.x, c21 = v_normal_tt.y + v_normal_local.y,
                                c22 = v_normal_tt.z + v_normal_local.z;
                    v[9] = 2.f * (qx * (c12 - c21) + qy * (c20 - c02) + qz * (c01 - c10));
                    v[10] = 2.f * (-2.f * qx * (c11 + c22) + qy * (c01 + c10) + qz * (c02 + c20) + qw * (c12 - c21));
                    v[11] = 2.f * (qx * (c01 + c10) - 2.f * qy * (c00 + c22) + qz * (c12 + c21) + qw * (c20 - c02));
                    v[12] = 2.f * (qx * (c02 + c20) + qy * (c12 + c21) - 2.f * qz * (c00 + c11) + qw * (c01 - c10));
                }
                v[3] = e.ex * v_alpha;
            }
            // segmented reduction: after the loop the head lane of each segment holds its totals
            // (a step of offset `off` is only needed when some segment is longer than `off` lanes, i.e. when the
            // head mask has a run of `off` non-head lanes: warp-uniform, so whole steps are skipped)
            unsigned run = ~heads;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                if (run == 0u) break;
                const bool take = lane + off <= seg_end;
#pragma unroll
                for (int k = 0; k < 13; ++k) {
                    const float t = __shfl_down_sync(0xffffffffu, v[k], off);
                    if (take) v[k] += t;
                }
                run &= run >> off;
            }
            if (head && valid) {  // four 16-byte vector reductions per (unit, splat)
                float4* acc = p.splat_acc + (size_t)gid * 4;
                red_add_v4(acc, v[0], v[1], v[2], v[3]);
                red_add_v4(acc + 1, v[4], v[5], v[6], 0.f);
                red_add_v4(acc + 2, v[7], v[8], 0.f, 0.f);
                red_add_v4(acc + 3, v[9], v[10], v[11], v[12]);
            }
            __syncwarp();  // every lane has read this batch's record slots before they are refilled
            buf ^= 1; seg = seg_nxt; heads = heads_nxt;
        }
    }
}

// Moves the padded accumulators into the reference's gradient tensors and hands the scratch back all-zero.  The
// accumulators are swept once, coalesced; only words that are non-zero -- a view reaches about a third of the splats
// and of their texels -- are cleared again and ADDED to the gradient tensors, which hold zeros (zero_fill 0/1) or the
// caller's running sum over the views of a step (zero_fill 2).  So no full-size clear precedes the stream kernel and
// the gradient tensors are only touched where this view has something to add.
__global__ void __launch_bounds__(256) unpack_grads_kernel(float4* __restrict__ tex_acc, long long num_texels, int C, float tex_scale,
                                                           float4* __restrict__ splat_acc, int n,
                                                           float* __restrict__ v_texture, float* __restrict__ v_colors,
                                                           float* __restrict__ v_opacity, float* __restrict__ v_means,
                                                           float* __restrict__ v_scales, float* __restrict__ v_quats) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    auto any = [](const float4& a) { return a.x != 0.f || a.y != 0.f || a.z != 0.f || a.w != 0.f; };
    auto texel = [&](long long i, const float4 a) {  // by value: `a` must not alias the word cleared below
        if (!any(a)) return;
        tex_acc[i] = z;
        float* o = v_texture + i * C;
        o[0] += tex_scale * a.x;
        if (C > 1) o[1] += tex_scale * a.y;
        if (C > 2) o[2] += tex_scale * a.z;
    };
    long long i = t0;
    for (; i + 3 * stride < num_texels; i += 4 * stride) {  // four independent loads in flight per thread
        const float4 a0 = tex_acc[i], a1 = tex_acc[i + stride], a2 = tex_acc[i + 2 * stride], a3 = tex_acc[i + 3 * stride];
        texel(i, a0); texel(i + stride, a1); texel(i + 2 * stride, a2); texel(i + 3 * stride, a3);
    }
    for (; i < num_texels; i += stride) texel(i, tex_acc[i]);
    for (long long i = t0; i < n; i += stride) {
        float4* acc = splat_acc + 4 * i;
        const float4 a = acc[0], b = acc[1], c = acc[2], d = acc[3];
        if (!(any(a) || any(b) || any(c) || any(d))) continue;
        acc[0] = z; acc[1] = z; acc[2] = z; acc[3] = z;
        v_colors[3 * i] += a.x; v_colors[3 * i + 1] += a.y; v_colors[3 * i + 2] += a.z;
        v_opacity[i] += a.w;
        v_means[3 * i] += b.x; v_means[3 * i + 1] += b.y; v_means[3 * i + 2] += b.z;
        v_scales[3 * i] += c.x; v_scales[3 * i + 1] += c.y;
        float4* q = reinterpret_cast<float4*>(v_quats) + i;
        const float4 o = *q;
        *q = make_float4(o.x + d.x, o.y + d.y, o.z + d.z, o.w + d.w);
    }
}

int launch_backward_stream(const StreamBwdParams& p, cudaStream_t st) {
    // the work counter is the only state the backward owns in the header
    GSTEX_CUDA_TRY(cudaMemsetAsync(&p.sv.hdr->next_unit, 0, sizeof(uint32_t), st));
    if (!p.scratch_clean) GSTEX_CUDA_TRY(cudaMemsetAsync(p.tex_acc, 0, grad_scratch_bytes(p.n, p.num_texels), st));
    int blocks = sm_count() * 2;
    const int need = (p.sv.num_units + kStreamWarps - 1) / kStreamWarps;
    if (blocks > need) blocks = need;
    if (blocks < 1) blocks = 1;
    GSTEX_CUDA_TRY(cudaFuncSetAttribute(texture_backward_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)sizeof(StreamShared)));
    texture_backward_stream_kernel<<<blocks, kStreamWarps * 32, sizeof(StreamShared), st>>>(p);
    GSTEX_LAUNCH_CHECK();
    if (p.defer_unpack) return GSTEX_OK;  // the sums stay in the scratch until gstex_unpack_grads
    return launch_unpack_grads(p.tex_acc, p.num_texels, p.C, p.tex_scale, p.splat_acc, p.n, p.v_texture, p.v_colors, p.v_opacity,
                               p.v_means, p.v_scales, p.v_quats, st);
}

int launch_unpack_grads(float4* tex_acc, long long num_texels, int C, float tex_scale, float4* splat_acc, int n,
                        float* v_texture, float* v_colors, float* v_opacity, float* v_means, float* v_scales,
                        float* v_quats, cudaStream_t st) {
    unpack_grads_kernel<<<sm_count() * 8, 256, 0, st>>>(tex_acc, num_texels, C, tex_scale, splat_acc, n, v_texture, v_colors,
                                                         v_opacity, v_means, v_scales, v_quats);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

}  // namespace gstex

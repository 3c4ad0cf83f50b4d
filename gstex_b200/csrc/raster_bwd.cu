// raster_bwd.cu -- stage (d): back-to-front backward re-traversal.
// Replaces texture_backward (gstex_cuda/cuda/csrc/texture.cu:331-760, wrapper :903-1053).  sm_100a.
//
// Differences in HOW from the reference kernel:
//   * same warp-specialised TMA/mbarrier ring and per-warp support-box culling as the forward
//     (raster_fwd.cu, pipeline.cuh), walking the tile's list back to front from the block's last contributor;
//     records behind every pixel's last contributor are never loaded, and a warp only evaluates splats
//     that are not beyond its own last contributor and can reach its pixel rows;
//   * the per-splat gradients of a warp are reduced with a transposing butterfly (16 shuffles for the 13
//     values of the training configuration instead of 22 five-step reductions = 110 shuffles), after which
//     lane 2s holds the total of value s, so ONE red instruction issues all of the warp's atomics for that
//     splat instead of 22 serial ones from lane 0;
//   * the bilinear index/weight computation is done once per pair, not once per channel.
#include "bwd_stream.cuh"
#include "pipeline.cuh"

namespace gstex {

// Transposing butterfly: on entry every lane holds K partial values; on exit v[0] of lane L holds the
// warp total of value (L * K / 32).  K = 32: 31 shuffles; K = 16: 16 shuffles.
template <int K>
__device__ __forceinline__ void butterfly_reduce(float (&v)[K], const int lane) {
    static_assert(K == 16 || K == 32, "K must be 16 or 32");
    int off = 16;
#pragma unroll
    for (int half = K / 2; half >= 1; half >>= 1, off >>= 1) {
        const bool hi = lane & off;
#pragma unroll
        for (int i = 0; i < half; ++i) {
            const float keep = hi ? v[i + half] : v[i];
            const float send = hi ? v[i] : v[i + half];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
    }
    if (K == 16) v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
}

struct BwdParams {
    RasterGeom g;
    int C;
    const SplatRec* recs;
    const SplatRecBwd* recs_bwd;
    const float4* boxes;
    const int32_t* ids;
    const int2* tile_bins;
    const float* texture;
    float tex_scale, tex_bias;  // texel activation of the forward (tex_scale * texel + tex_bias)
    const float *viewmat, *c2w, *background;
    float fx, fy, cx, cy;
    int settings;
    const float *final_Ts, *final_s;
    const int32_t *final_idx, *depth_idx;
    const float *v_output, *v_output_depth, *v_output_reg, *v_output_alpha, *v_output_texture, *v_output_normal;
    float *v_colors, *v_opacity, *v_means, *v_scales, *v_quats, *v_uv0, *v_umap, *v_vmap, *v_texture;
    // non-null when a pair stream exists: this kernel only runs if the forward flagged a pool overflow there
    const uint32_t* stream_overflow;
};

struct BwdShared {
    SplatRec recs[kStages][kStageRecs];     // 32 KB
    SplatRecBwd bwd[kStages][kStageRecs];   // 8 KB
    int ids[kStages][kStageRecs];           // 1 KB
    uint64_t full[kStages];
    uint64_t empty[kStages];
    unsigned bits[kStages][8];
    Camera cam;
    int warp_final[8];
    int block_final;
};

// slot layout of the per-splat gradient vector
//  0-2 v_colors | 3 v_opacity | 4-6 v_means | 7-8 v_scales.xy | 9-12 v_quats | (UV) 13-14 v_uv0 | 15-17 v_umap | 18-20 v_vmap
template <bool UV>
__device__ __forceinline__ void slot_target(const BwdParams& p, int slot, float*& base, int& stride) {
    base = nullptr; stride = 0;
    if (slot < 3) { base = p.v_colors + slot; stride = 3; }
    else if (slot == 3) { base = p.v_opacity; stride = 1; }
    else if (slot < 7) { base = p.v_means + (slot - 4); stride = 3; }
    else if (slot < 9) { base = p.v_scales + (slot - 7); stride = 3; }
    else if (slot < 13) { base = p.v_quats + (slot - 9); stride = 4; }
    else if (UV) {
        if (slot < 15) { base = p.v_uv0 + (slot - 13); stride = 2; }
        else if (slot < 18) { base = p.v_umap + (slot - 15); stride = 3; }
        else if (slot < 21) { base = p.v_vmap + (slot - 18); stride = 3; }
    }
}

template <int CMAX, bool UV>
__global__ void __launch_bounds__(kMaxTileThreads + 32, 2) texture_backward_kernel(const BwdParams p) {
    constexpr int K = UV ? 32 : 16;
    __shared__ __align__(128) BwdShared sm;
    if (p.stream_overflow != nullptr && *(volatile const uint32_t*)p.stream_overflow == 0u) return;
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
        sm.block_final = -1;
        mbar_fence_init();
    }
    __syncthreads();
    const int tile_id = blockIdx.y * g.tiles_x + blockIdx.x;
    const int2 range = p.tile_bins[tile_id];

    // per-pixel state (pixel warps only; the producer computes a dummy pixel it never uses)
    Pixel pix;
    int bin_final = -1;
    if (!is_producer) {
        pix = make_pixel(sm.cam, g, blockIdx.x, blockIdx.y);
        bin_final = pix.inside ? p.final_idx[pix.pix_id] : -1;
        int wf = bin_final;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) wf = max(wf, __shfl_xor_sync(0xffffffffu, wf, o));
        if (lane == 0) {
            sm.warp_final[warp] = wf;
            atomicMax(&sm.block_final, wf);
        }
    }
    __syncthreads();
    // walk idx = block_final, block_final-1, ..., range.x; stage n covers [batch_end-31, batch_end]
    const int block_final = min(sm.block_final, range.y - 1);
    const int n_stages = (block_final >= range.x) ? (block_final - range.x) / kStageRecs + 1 : 0;

    if (is_producer) {
        const float tx0 = (float)(blockIdx.x * g.bw) + 0.5f, tx1 = (float)(blockIdx.x * g.bw + g.bw - 1) + 0.5f;
        const float ty = (float)(blockIdx.y * g.bw) + 0.5f;
        int g_next = (block_final - lane >= range.x && n_stages > 0) ? __ldg(p.ids + block_final - lane) : -1;
        for (int n = 0; n < n_stages; ++n) {
            const int s = n % kStages;
            const int idx = block_final - n * kStageRecs - lane;
            const int gid = g_next;
            const int nxt = idx - kStageRecs;
            g_next = (nxt >= range.x) ? __ldg(p.ids + nxt) : -1;
            const float4 box = (gid >= 0) ? __ldg(p.boxes + gid) : make_float4(1.f, 1.f, 0.f, 0.f);
            if (n >= kStages) mbar_wait(&sm.empty[s], ((n / kStages) - 1) & 1);
            unsigned mine = 0, any = 0;
            for (int w = 0; w < n_pix_warps; ++w) {
                int r0, r1;
                warp_rows(w, g.bw, r0, r1);
                const bool hit = gid >= 0 && idx <= sm.warp_final[w] &&
                                 box_overlaps(box, tx0, tx1, ty + (float)r0, ty + (float)r1);
                const unsigned b = __ballot_sync(0xffffffffu, hit);
                if (lane == w) mine = b;
                any |= b;
            }
            if (lane < n_pix_warps) sm.bits[s][lane] = mine;
            if (gid >= 0) sm.ids[s][lane] = gid;
            __syncwarp();
            if (lane == 0)
                mbar_arrive_expect_tx(&sm.full[s], (uint32_t)__popc(any) * (uint32_t)(sizeof(SplatRec) + sizeof(SplatRecBwd)));
            if ((any >> lane) & 1u) {
                bulk_copy_g2s(&sm.recs[s][lane], p.recs + gid, sizeof(SplatRec), &sm.full[s]);
                bulk_copy_g2s(&sm.bwd[s][lane], p.recs_bwd + gid, sizeof(SplatRecBwd), &sm.full[s]);
            }
        }
        return;
    }
    // =============================== pixel warps: back to front ===============================
    const bool use_blur = p.settings & GSTEX_SET_BLUR;
    const bool use_ndc = p.settings & GSTEX_SET_NDC_REG;
    const bool bilinear = !(p.settings & GSTEX_SET_NEAREST);
    const int C = p.C;
    const int id = pix.pix_id;
    float T = p.final_Ts[id];
    const float3 ts_final = make_float3(p.final_s[3 * id], p.final_s[3 * id + 1], p.final_s[3 * id + 2]);
    const int depth_final = p.depth_idx[id];
    // a NULL upstream gradient is a zero gradient (autograd's None): nothing is read for it
    const float3 v_out = p.v_output ? make_float3(p.v_output[3 * id], p.v_output[3 * id + 1], p.v_output[3 * id + 2])
                                    : make_float3(0.f, 0.f, 0.f);
    const float v_out_depth = p.v_output_depth ? p.v_output_depth[id] : 0.f, v_out_reg = p.v_output_reg ? p.v_output_reg[id] : 0.f;
    const float3 v_out_normal =
        p.v_output_normal ? make_float3(p.v_output_normal[3 * id], p.v_output_normal[3 * id + 1], p.v_output_normal[3 * id + 2])
                          : make_float3(0.f, 0.f, 0.f);
    float v_tex[CMAX];
#pragma unroll
    for (int k = 0; k < CMAX; ++k) v_tex[k] = (p.v_output_texture && k < C) ? p.v_output_texture[(size_t)C * id + k] : 0.f;
    float v_T_running = __ldg(p.background) * v_out.x + __ldg(p.background + 1) * v_out.y +
                        __ldg(p.background + 2) * v_out.z - (p.v_output_alpha ? p.v_output_alpha[id] : 0.f);

    // where this lane's reduced slot goes
    float* slot_base;
    int slot_stride;
    slot_target<UV>(p, (K == 16) ? (lane >> 1) : lane, slot_base, slot_stride);
    if (K == 16 && (lane & 1)) slot_base = nullptr;

    for (int n = 0; n < n_stages; ++n) {
        const int s = n % kStages;
        mbar_wait(&sm.full[s], (n / kStages) & 1);
        const int batch_end = block_final - n * kStageRecs;
        unsigned todo = sm.bits[s][warp];
        while (todo) {
            const int j = __ffs(todo) - 1;
            todo &= todo - 1;
            const int idx = batch_end - j;
            const SplatRec& r = sm.recs[s][j];
            const float4 r0 = r.q[0], r1 = r.q[1], r2 = r.q[2], r3 = r.q[3], r4 = r.q[4];
            const PairEval e = eval_pair(pix, r0, r1, r2, r3, r4, use_blur);
            const bool valid = (idx <= bin_final) && !e.skip;
            if (!__any_sync(0xffffffffu, valid)) continue;
            float v[K];
#pragma unroll
            for (int k = 0; k < K; ++k) v[k] = 0.f;
            if (valid) {
                const float4 r5 = r.q[5], r6 = r.q[6], r7 = r.q[7];
                const SplatRecBwd rb = sm.bwd[s][j];
                const float alpha = e.alpha;
                const float ra = 1.f / (1.f - alpha);
                T *= ra;
                const float vis = alpha * T;
                v[0] = vis * v_out.x; v[1] = vis * v_out.y; v[2] = vis * v_out.z;
                const float3 v_normal_local = make_float3(vis * v_out_normal.x, vis * v_out_normal.y, vis * v_out_normal.z);
                float v_vis = r5.x * v_out.x + r5.y * v_out.y + r5.z * v_out.z;
                v_vis += r1.x * v_out_normal.x + r1.y * v_out_normal.y + r1.z * v_out_normal.z;
                float3 v_diff = make_float3(0.f, 0.f, 0.f);
                {
                    const int th = __float_as_int(r4.z), tw = __float_as_int(r4.w);
                    const float u = chart_coord(r6, e.delta);
                    const float vv = chart_coord(r7, e.delta);
                    const TexQuery q = tex_query(th, tw, __float_as_int(r5.w), u, vv, bilinear);
                    const size_t o00 = (size_t)q.b00 * C, o01 = (size_t)q.b01 * C, o10 = (size_t)q.b10 * C, o11 = (size_t)q.b11 * C;
                    float v_u = 0.f, v_v = 0.f;
#pragma unroll
                    for (int k = 0; k < CMAX; ++k) {
                        if (k < C) {
                            const float t00 = fmaf(p.tex_scale, __ldg(p.texture + o00 + k), p.tex_bias);
                            const float t01 = fmaf(p.tex_scale, __ldg(p.texture + o01 + k), p.tex_bias);
                            const float t10 = fmaf(p.tex_scale, __ldg(p.texture + o10 + k), p.tex_bias);
                            const float t11 = fmaf(p.tex_scale, __ldg(p.texture + o11 + k), p.tex_bias);
                            const float val = q.w00 * t00 + q.w01 * t01 + q.w10 * t10 + q.w11 * t11;
                            const float v_val = vis * v_tex[k];
                            // query_jagged_chart_vjp, texture_helpers.cuh:252-260 (times the activation's slope)
                            const float v_raw = v_val * p.tex_scale;
                            atomicAdd(p.v_texture + o00 + k, q.w00 * v_raw);
                            atomicAdd(p.v_texture + o01 + k, q.w01 * v_raw);
                            atomicAdd(p.v_texture + o10 + k, q.w10 * v_raw);
                            atomicAdd(p.v_texture + o11 + k, q.w11 * v_raw);
                            if (UV && bilinear) {  // texture_helpers.cuh:262-299
                                v_u += (float)th * (v_val * (-(1.f - q.fy) * t00 - q.fy * t01 + (1.f - q.fy) * t10 + q.fy * t11));
                                v_v += (float)tw * (v_val * (-(1.f - q.fx) * t00 + (1.f - q.fx) * t01 - q.fx * t10 + q.fx * t11));
                            }
                            v_vis += val * v_tex[k];
                        }
                    }
                    if (UV) {  // get_uv_vjp, texture_helpers.cuh:62-80 (the clamp is not differentiated)
                        v[13] = v_u; v[14] = v_v;
                        v[15] = e.delta.x * v_u; v[16] = e.delta.y * v_u; v[17] = e.delta.z * v_u;
                        v[18] = e.delta.x * v_v; v[19] = e.delta.y * v_v; v[20] = e.delta.z * v_v;
                        v_diff = make_float3(r6.x * v_u + r7.x * v_v, r6.y * v_u + r7.y * v_v, r6.z * v_u + r7.z * v_v);
                    }
                }
                // texture.cu:650-670
                float v_alpha = T * v_vis - T * v_T_running;
                float v_T_cur = alpha * v_vis + (1.f - alpha) * v_T_running;
                const float t_view = e.t * pix.view_depth;
                float v_t = 0.f, v_t_view = 0.f;
                {
                    const float d = use_ndc ? div_normal(kFar * t_view - kFar * kNear, (kFar - kNear) * t_view) : e.t;
                    // distortion_vjp, helpers.cuh:266-269, with the FINAL sums
                    const float v_d = 2.f * (vis * d * ts_final.x - vis * ts_final.y) * v_out_reg;
                    const float v_weight = (d * d * ts_final.x - 2.f * d * ts_final.y + ts_final.z) * v_out_reg;
                    if (use_ndc) v_t_view += div_normal(kFar * kNear, (kFar - kNear) * t_view * t_view) * v_d;
                    else v_t += v_d;
                    v_alpha += v_weight * T;
                    v_T_cur += v_weight * alpha;
                }
                v_T_running = v_T_cur;
                // texture.cu:672-707
                const float v_sigma = e.blur ? 0.f : -r1.w * e.ex * v_alpha;
                const float v_sigma_blur = e.blur ? -r1.w * e.ex * v_alpha : 0.f;
                const float v_l1 = r2.w * e.l1 * v_sigma, v_l2 = r3.w * e.l2 * v_sigma;
                v[7] = -(r2.w * e.l1 * e.l1) * v_sigma * rb.aux.x;
                v[8] = -(r3.w * e.l2 * e.l2) * v_sigma * rb.aux.y;
                v_t += v_l1 * (pix.ray.x * r2.x + pix.ray.y * r2.y + pix.ray.z * r2.z) +
                       v_l2 * (pix.ray.x * r3.x + pix.ray.y * r3.y + pix.ray.z * r3.z);
                v_t += pix.ray.x * v_diff.x + pix.ray.y * v_diff.y + pix.ray.z * v_diff.z;
                if (idx == depth_final && depth_final != -1) v_t_view += v_out_depth;
                v_t += pix.view_depth * v_t_view;
                float3 v_mean_extra = make_float3(0.f, 0.f, 0.f);
                if (e.blur) {  // project_pix_vjp + transform_4x3_rot_only_transposed, texture.cu:685-692
                    const float vxp = sm.cam.fx * (2.0f * v_sigma_blur * (r4.x - pix.px));
                    const float vyp = sm.cam.fy * (2.0f * v_sigma_blur * (r4.y - pix.py));
                    const float rw = rb.aux.z;
                    const float nx = (r4.x - sm.cam.cx) / sm.cam.fx, ny = (r4.y - sm.cam.cy) / sm.cam.fy;  // p_view.xy * rw
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
                v[4] = -(r2.x * v_l1 + r3.x * v_l2) + r1.x * vt_den + v_mean_extra.x - v_diff.x;
                v[5] = -(r2.y * v_l1 + r3.y * v_l2) + r1.y * vt_den + v_mean_extra.y - v_diff.y;
                v[6] = -(r2.z * v_l1 + r3.z * v_l2) + r1.z * vt_den + v_mean_extra.z - v_diff.z;
                {  // quat_to_rotmat_vjp, helpers.cuh:187-228; R columns: c0 = delta*v_l1, c1 = delta*v_l2, c2
                    const float qw = rb.quat.x, qx = rb.quat.y, qy = rb.quat.z, qz = rb.quat.w;
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
            butterfly_reduce<K>(v, lane);
            if (slot_base != nullptr && v[0] != 0.f) atomicAdd(slot_base + (size_t)sm.ids[s][j] * slot_stride, v[0]);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.empty[s]);
    }
}

}  // namespace gstex

using namespace gstex;

template <int CMAX>
static void launch_bwd(const BwdParams& p, bool uv, cudaStream_t st) {
    const dim3 grid(p.g.tiles_x, p.g.tiles_y);
    const int threads = p.g.threads + 32;  // + the producer warp
    if (uv) texture_backward_kernel<CMAX, true><<<grid, threads, 0, st>>>(p);
    else texture_backward_kernel<CMAX, false><<<grid, threads, 0, st>>>(p);
}

extern "C" int gstex_pair_stream_supported(int channels, int n, int settings);

extern "C" size_t gstex_grad_scratch_bytes(int n, long long num_texels) { return grad_scratch_bytes(n, num_texels); }

extern "C" int gstex_texture_backward_act(
    int img_height, int img_width, int block_width, int channels, int n, int num_intersects,
    const int32_t* texture_dims, const int32_t* gaussian_ids_sorted, const int32_t* tile_bins, const float* colors,
    const float* opacities, const float* means, const float* scales, float glob_scale, const float* quats,
    const float* uv0, const float* umap, const float* vmap, const float* texture, float tex_scale, float tex_bias, const float* viewmat,
    const float* c2w, float fx, float fy, float cx, float cy, int settings, const float* background,
    const float* final_Ts, const int32_t* final_idx, const int32_t* depth_idx, const float* final_s,
    const float* v_output, const float* v_output_depth, const float* v_output_reg, const float* v_output_alpha,
    const float* v_output_texture, const float* v_output_normal, float* v_colors, float* v_opacity, float* v_means,
    float* v_scales, float* v_quats, float* v_uv0, float* v_umap, float* v_vmap, float* v_texture, void* workspace,
    size_t workspace_bytes, int records_valid, void* pair_stream, size_t pair_stream_bytes, void* grad_scratch,
    size_t grad_scratch_size, long long num_texels, int zero_fill, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (img_height <= 0 || img_width <= 0 || block_width < 2 || block_width > 16 || n < 0 || num_intersects < 0)
        return GSTEX_E_BADARG;
    if (settings & ~GSTEX_SETTINGS_SUPPORTED) return GSTEX_E_UNSUPPORTED;
    if (channels < 1 || channels > GSTEX_MAX_CHANNELS) return GSTEX_E_UNSUPPORTED;
    if (zero_fill < 0 || zero_fill > 15 || (zero_fill & 3) == 3) return GSTEX_E_BADARG;
    const int scratch_clean = (zero_fill & GSTEX_SCRATCH_CLEAN) != 0;
    const int defer_unpack = (zero_fill & GSTEX_DEFER_UNPACK) != 0;
    zero_fill &= 3;
    if (defer_unpack && zero_fill != 2) return GSTEX_E_BADARG;  // deferring only makes sense for running sums
    if (zero_fill && (num_texels < 0 || (n > 0 && (!v_colors || !v_opacity || !v_means || !v_scales || !v_quats ||
                                                    !v_uv0 || !v_umap || !v_vmap || !v_texture))))
        return GSTEX_E_BADARG;
    // zero_fill: the caller passed uninitialised gradient buffers (the reference allocates them with torch::zeros)
    auto zero = [&](bool also_stream_outputs) -> int {
        if (zero_fill != 1 || n == 0) return GSTEX_OK;  // 0: zeroed by the caller, 2: running sums to add to
        const size_t f = sizeof(float);
        GSTEX_CUDA_TRY(cudaMemsetAsync(v_uv0, 0, (size_t)n * 2 * f, st));
        GSTEX_CUDA_TRY(cudaMemsetAsync(v_umap, 0, (size_t)n * 3 * f, st));
        GSTEX_CUDA_TRY(cudaMemsetAsync(v_vmap, 0, (size_t)n * 3 * f, st));
        if (!also_stream_outputs) return GSTEX_OK;
        GSTEX_CUDA_TRY(cudaMemsetAsync(v_colors, 0, (size_t)n * 3 * f, st));
        GSTEX_CUDA_TRY(cudaMemsetAsync(v_opacity, 0, (size_t)n * f, st));
        GSTEX_CUDA_TRY(cudaMemsetAsync(v_means, 0, (size_t)n * 3 * f, st));
        GSTEX_CUDA_TRY(cudaMemsetAsync(v_scales, 0, (size_t)n * 3 * f, st));
        GSTEX_CUDA_TRY(cudaMemsetAsync(v_quats, 0, (size_t)n * 4 * f, st));
        GSTEX_CUDA_TRY(cudaMemsetAsync(v_texture, 0, (size_t)num_texels * channels * f, st));
        return GSTEX_OK;
    };
    if (n == 0 || num_intersects == 0) return zero(true);  // gradients are zero (texture.py:331-340)
    if (!texture_dims || !gaussian_ids_sorted || !tile_bins || !colors || !opacities || !means || !scales || !quats ||
        !uv0 || !umap || !vmap || !texture || !viewmat || !c2w || !background || !final_Ts || !final_idx ||
        !depth_idx || !final_s || !v_colors || !v_opacity || !v_means || !v_scales || !v_quats ||
        !v_uv0 || !v_umap || !v_vmap || !v_texture || !workspace)
        return GSTEX_E_BADARG;
    if (workspace_bytes < ws_bytes(n)) return GSTEX_E_WORKSPACE;
    if (!records_valid) {
        const int rc = launch_pack_records(n, texture_dims, colors, opacities, means, scales, glob_scale, quats, uv0, umap,
                                           vmap, viewmat, c2w, fx, fy, cx, cy, settings, workspace, st);
        if (rc != GSTEX_OK) return rc;
    }
    BwdParams p;
    p.g = make_raster_geom(img_height, img_width, block_width);
    p.C = channels; p.recs = ws_recs(workspace); p.recs_bwd = ws_recs_bwd(workspace, n); p.boxes = ws_boxes(workspace, n);
    p.ids = gaussian_ids_sorted;
    p.tile_bins = reinterpret_cast<const int2*>(tile_bins); p.texture = texture; p.tex_scale = tex_scale; p.tex_bias = tex_bias; p.viewmat = viewmat; p.c2w = c2w;
    p.background = background; p.fx = fx; p.fy = fy; p.cx = cx; p.cy = cy; p.settings = settings;
    p.final_Ts = final_Ts; p.final_s = final_s; p.final_idx = final_idx; p.depth_idx = depth_idx;
    p.v_output = v_output; p.v_output_depth = v_output_depth; p.v_output_reg = v_output_reg;
    p.v_output_alpha = v_output_alpha; p.v_output_texture = v_output_texture; p.v_output_normal = v_output_normal;
    p.v_colors = v_colors; p.v_opacity = v_opacity; p.v_means = v_means; p.v_scales = v_scales; p.v_quats = v_quats;
    p.v_uv0 = v_uv0; p.v_umap = v_umap; p.v_vmap = v_vmap; p.v_texture = v_texture;
    p.stream_overflow = nullptr;
    // Fast path: the forward of this view recorded its contributing pairs.  Both kernels are enqueued; exactly one
    // of them does the work, decided on the device by the overflow flag in the stream header (no host sync).
    StreamBwdParams sp;
    if (pair_stream != nullptr && grad_scratch != nullptr && num_texels > 0 && ((uintptr_t)grad_scratch & 15) == 0 &&
        ((uintptr_t)v_quats & 15) == 0 && grad_scratch_size >= grad_scratch_bytes(n, num_texels) && records_valid &&
        gstex_pair_stream_supported(channels, n, settings) &&
        make_stream_view(pair_stream, pair_stream_bytes, p.g.tiles_x * p.g.tiles_y, sp.sv)) {
        sp.g = p.g; sp.C = channels; sp.recs = p.recs; sp.recs_bwd = p.recs_bwd; sp.texture = texture;
        sp.viewmat = viewmat; sp.c2w = c2w; sp.background = background; sp.fx = fx; sp.fy = fy; sp.cx = cx; sp.cy = cy;
        sp.settings = settings; sp.final_Ts = final_Ts; sp.final_s = final_s; sp.depth_idx = depth_idx;
        sp.v_output = v_output; sp.v_output_depth = v_output_depth; sp.v_output_reg = v_output_reg;
        sp.v_output_alpha = v_output_alpha; sp.v_output_texture = v_output_texture; sp.v_output_normal = v_output_normal;
        sp.v_colors = v_colors; sp.v_opacity = v_opacity; sp.v_means = v_means; sp.v_scales = v_scales;
        sp.v_quats = v_quats; sp.v_texture = v_texture;
        sp.n = n; sp.num_texels = num_texels; sp.scratch_clean = scratch_clean; sp.tex_scale = tex_scale;
        sp.defer_unpack = defer_unpack;
        sp.order = ws_order(workspace, n); sp.order_valid = ws_counters(workspace) + 1;
        sp.tex_acc = reinterpret_cast<float4*>(grad_scratch); sp.splat_acc = sp.tex_acc + num_texels;
        // the unpack kernel of the stream path adds to the gradients wherever this view contributed
        int rc = zero(true);
        if (rc != GSTEX_OK) return rc;
        rc = launch_backward_stream(sp, st);
        if (rc != GSTEX_OK) return rc;
        p.stream_overflow = &sp.sv.hdr->overflow;
    } else {
        const int rc = zero(true);
        if (rc != GSTEX_OK) return rc;
    }
    const bool uv = (settings & GSTEX_SET_UV_GRAD) != 0;
    if (channels <= 3) launch_bwd<3>(p, uv, st);
    else if (channels <= 6) launch_bwd<6>(p, uv, st);
    else launch_bwd<GSTEX_MAX_CHANNELS>(p, uv, st);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

extern "C" int gstex_unpack_grads(int channels, int n, long long num_texels, float tex_scale, void* grad_scratch,
                                  size_t grad_scratch_size, float* v_colors, float* v_opacity, float* v_means,
                                  float* v_scales, float* v_quats, float* v_texture, void* stream) {
    if (channels < 1 || channels > 3 || n < 0 || num_texels < 0) return GSTEX_E_BADARG;
    if (n == 0 && num_texels == 0) return GSTEX_OK;
    if (!grad_scratch || ((uintptr_t)grad_scratch & 15) || ((uintptr_t)v_quats & 15) ||
        grad_scratch_size < grad_scratch_bytes(n, num_texels))
        return GSTEX_E_BADARG;
    if (!v_colors || !v_opacity || !v_means || !v_scales || !v_quats || !v_texture) return GSTEX_E_BADARG;
    float4* tex_acc = reinterpret_cast<float4*>(grad_scratch);
    return launch_unpack_grads(tex_acc, num_texels, channels, tex_scale, tex_acc + num_texels, n, v_texture, v_colors,
                               v_opacity, v_means, v_scales, v_quats, (cudaStream_t)stream);
}

extern "C" int gstex_texture_backward(
    int img_height, int img_width, int block_width, int channels, int n, int num_intersects,
    const int32_t* texture_dims, const int32_t* gaussian_ids_sorted, const int32_t* tile_bins, const float* colors,
    const float* opacities, const float* means, const float* scales, float glob_scale, const float* quats,
    const float* uv0, const float* umap, const float* vmap, const float* texture, const float* viewmat,
    const float* c2w, float fx, float fy, float cx, float cy, int settings, const float* background,
    const float* final_Ts, const int32_t* final_idx, const int32_t* depth_idx, const float* final_s,
    const float* v_output, const float* v_output_depth, const float* v_output_reg, const float* v_output_alpha,
    const float* v_output_texture, const float* v_output_normal, float* v_colors, float* v_opacity, float* v_means,
    float* v_scales, float* v_quats, float* v_uv0, float* v_umap, float* v_vmap, float* v_texture, void* workspace,
    size_t workspace_bytes, int records_valid, void* pair_stream, size_t pair_stream_bytes, void* grad_scratch,
    size_t grad_scratch_size, long long num_texels, int zero_fill, void* stream) {
    return gstex_texture_backward_act(img_height, img_width, block_width, channels, n, num_intersects, texture_dims, gaussian_ids_sorted, tile_bins, colors, opacities, means, scales, glob_scale, quats, uv0, umap, vmap, texture, 1.f, 0.f, viewmat, c2w, fx, fy, cx, cy, settings, background, final_Ts, final_idx, depth_idx, final_s, v_output, v_output_depth, v_output_reg, v_output_alpha, v_output_texture, v_output_normal, v_colors, v_opacity, v_means, v_scales, v_quats, v_uv0, v_umap, v_vmap, v_texture, workspace, workspace_bytes, records_valid, pair_stream, pair_stream_bytes, grad_scratch, grad_scratch_size, num_texels, zero_fill, stream);
}

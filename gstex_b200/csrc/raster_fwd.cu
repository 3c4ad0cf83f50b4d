// raster_fwd.cu -- stage (c): per-tile textured forward compositing, plus the per-view record packer.
// Replaces texture_forward (gstex_cuda/cuda/csrc/texture.cu:11-329, wrapper :762-901).  sm_100a.
//
// Differences in HOW (not WHAT) from the reference kernel:
//   * per-splat invariants (rotation axes, inverse-scale squares, ray/plane numerator, projected mean)
//     come from a 128-byte packed record computed once per (view, splat) instead of once per pixel;
//   * records are staged cooperatively (8 lanes x 16 B per record = whole lines) instead of ten scattered
//     4-16 B gathers per splat;
//   * the bilinear index/weight computation is done once per pair, not once per channel; texture
//     accumulators live in registers (channel count is a template bound), not in a float[64] local array;
//   * a warp stops as soon as all its 32 pixels are done (ballot), the block when all warps are.
#include "raster_common.cuh"

namespace gstex {

// ------------------------------------------------------------------------------------------------
// record packer: one thread per splat
// ------------------------------------------------------------------------------------------------
struct PackParams {
    int n;
    const int32_t* texture_dims;
    const float *colors, *opacities, *means, *scales, *quats, *uv0, *umap, *vmap, *viewmat, *c2w;
    float glob_scale, fx, fy, cx, cy;
    SplatRec* recs;
    SplatRecBwd* recs_bwd;
};

__global__ void __launch_bounds__(kThreads1D) pack_records_kernel(const PackParams p) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n) return;
    const float mx = p.means[3 * i], my = p.means[3 * i + 1], mz = p.means[3 * i + 2];
    const float4 qt = reinterpret_cast<const float4*>(p.quats)[i];
    const float w = qt.x, x = qt.y, y = qt.z, z = qt.w;
    // columns of R(quat), helpers.cuh:166-185
    const float a1x = 1.f - 2.f * (y * y + z * z), a1y = 2.f * (x * y + w * z), a1z = 2.f * (x * z - w * y);
    const float a2x = 2.f * (x * y - w * z), a2y = 1.f - 2.f * (x * x + z * z), a2z = 2.f * (y * z + w * x);
    const float a3x = 2.f * (x * z + w * y), a3y = 2.f * (y * z - w * x), a3z = 1.f - 2.f * (x * x + y * y);
    const float ox = __ldg(p.c2w + 3), oy = __ldg(p.c2w + 7), oz = __ldg(p.c2w + 11);
    const float dx = mx - ox, dy = my - oy, dz = mz - oz;
    const float tnum = a3x * dx + a3y * dy + a3z * dz;  // compute_tt numerator, texture_helpers.cuh:303,312
    const float is1 = 1.f / (p.scales[3 * i] * p.glob_scale), is2 = 1.f / (p.scales[3 * i + 1] * p.glob_scale);
    // projected mean (texture.cu:184-185): transform_4x3 + project_pix
    const float* vm = p.viewmat;
    const float vx = __ldg(vm + 0) * mx + __ldg(vm + 1) * my + __ldg(vm + 2) * mz + __ldg(vm + 3);
    const float vy = __ldg(vm + 4) * mx + __ldg(vm + 5) * my + __ldg(vm + 6) * mz + __ldg(vm + 7);
    const float vz = __ldg(vm + 8) * mx + __ldg(vm + 9) * my + __ldg(vm + 10) * mz + __ldg(vm + 11);
    const float rw = 1.f / (vz + 1e-6f);
    const float xy_x = vx * rw * p.fx + p.cx, xy_y = vy * rw * p.fy + p.cy;
    const int h = p.texture_dims[3 * i], wd = p.texture_dims[3 * i + 1], st = p.texture_dims[3 * i + 2];
    SplatRec r;
    r.q[0] = make_float4(mx, my, mz, tnum);
    r.q[1] = make_float4(a3x, a3y, a3z, p.opacities[i]);
    r.q[2] = make_float4(a1x, a1y, a1z, is1 * is1);
    r.q[3] = make_float4(a2x, a2y, a2z, is2 * is2);
    r.q[4] = make_float4(xy_x, xy_y, __int_as_float(h), __int_as_float(wd));
    r.q[5] = make_float4(p.colors[3 * i], p.colors[3 * i + 1], p.colors[3 * i + 2], __int_as_float(st));
    r.q[6] = make_float4(p.umap[3 * i], p.umap[3 * i + 1], p.umap[3 * i + 2], p.uv0[2 * i]);
    r.q[7] = make_float4(p.vmap[3 * i], p.vmap[3 * i + 1], p.vmap[3 * i + 2], p.uv0[2 * i + 1]);
#pragma unroll
    for (int k = 0; k < 8; ++k) p.recs[i].q[k] = r.q[k];
    p.recs_bwd[i].quat = qt;
    p.recs_bwd[i].aux = make_float4(1.f / p.scales[3 * i], 1.f / p.scales[3 * i + 1], rw, 0.f);
}

int launch_pack_records(int n, const int32_t* texture_dims, const float* colors, const float* opacities,
                        const float* means, const float* scales, float glob_scale, const float* quats,
                        const float* uv0, const float* umap, const float* vmap, const float* viewmat,
                        const float* c2w, float fx, float fy, float cx, float cy, SplatRec* recs, SplatRecBwd* recs_bwd,
                        cudaStream_t st) {
    PackParams p;
    p.n = n; p.texture_dims = texture_dims; p.colors = colors; p.opacities = opacities; p.means = means;
    p.scales = scales; p.quats = quats; p.uv0 = uv0; p.umap = umap; p.vmap = vmap; p.viewmat = viewmat; p.c2w = c2w;
    p.glob_scale = glob_scale; p.fx = fx; p.fy = fy; p.cx = cx; p.cy = cy; p.recs = recs; p.recs_bwd = recs_bwd;
    pack_records_kernel<<<div_up(n, kThreads1D), kThreads1D, 0, st>>>(p);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

// ------------------------------------------------------------------------------------------------
// forward compositing
// ------------------------------------------------------------------------------------------------
struct FwdParams {
    RasterGeom g;
    int C;
    const SplatRec* recs;
    const int32_t* ids;
    const int2* tile_bins;
    const float* texture;
    const float *viewmat, *c2w, *background;
    float fx, fy, cx, cy;
    int settings;
    float *out_img, *out_depth, *out_reg, *out_texture, *out_normal, *final_Ts, *out_reg_s;
    int32_t *final_idx, *depth_idx;
};

template <int CMAX>
__global__ void __launch_bounds__(kMaxTileThreads) texture_forward_kernel(const FwdParams p) {
    __shared__ SplatRec s_recs[kBatch];
    __shared__ Camera s_cam;
    const RasterGeom g = p.g;
    const int nthreads = blockDim.x;
    load_camera(s_cam, p.viewmat, p.c2w, p.fx, p.fy, p.cx, p.cy);
    __syncthreads();
    const int tile_id = blockIdx.y * g.tiles_x + blockIdx.x;
    const Pixel pix = make_pixel(s_cam, g, blockIdx.x, blockIdx.y);
    const bool use_blur = p.settings & GSTEX_SET_BLUR;
    const bool use_ndc = p.settings & GSTEX_SET_NDC_REG;
    const bool bilinear = !(p.settings & GSTEX_SET_NEAREST);
    const int C = p.C;

    const int2 range = p.tile_bins[tile_id];
    const int num_batches = (range.y - range.x + kBatch - 1) / kBatch;

    bool done = !pix.inside;
    float T = 1.f;
    int cur_idx = 0, pix_depth_index = -1;
    float3 rgb = make_float3(0.f, 0.f, 0.f), nrm = make_float3(0.f, 0.f, 0.f), ts = make_float3(0.f, 0.f, 0.f);
    float pix_depth = 0.f, pix_reg = 0.f;
    float tex[CMAX];
#pragma unroll
    for (int k = 0; k < CMAX; ++k) tex[k] = 0.f;

    for (int b = 0; b < num_batches; ++b) {
        // block-wide early exit (texture.cu:132); also protects s_recs against the previous batch's readers
        if (__syncthreads_count(done) >= nthreads) break;
        const int batch_start = range.x + kBatch * b;
        const int count = min(kBatch, range.y - batch_start);
        stage_records(s_recs, p.recs, p.ids, batch_start, count, false, nthreads);
        __syncthreads();
        for (int i = 0; i < count; ++i) {
            if (__all_sync(0xffffffffu, done)) break;  // this warp's 32 pixels are finished
            const SplatRec& r = s_recs[i];
            const float4 r0 = r.q[0], r1 = r.q[1], r2 = r.q[2], r3 = r.q[3], r4 = r.q[4];
            const PairEval e = eval_pair(pix, r0, r1, r2, r3, r4, use_blur);
            const float next_T = T * (1.f - e.alpha);
            if (!done) {
                if (next_T <= 1e-4f) {
                    done = true;  // tested before `skip`, texture.cu:216-221; this splat is not composited
                } else if (!e.skip) {
                    const float vis = e.alpha * T;
                    const float4 r5 = r.q[5], r6 = r.q[6], r7 = r.q[7];
                    rgb.x += r5.x * vis; rgb.y += r5.y * vis; rgb.z += r5.z * vis;
                    nrm.x += vis * r1.x; nrm.y += vis * r1.y; nrm.z += vis * r1.z;
                    const float t_view = e.t * pix.view_depth;
                    const float u = clamp01(r6.w + (r6.x * e.delta.x + r6.y * e.delta.y + r6.z * e.delta.z));
                    const float v = clamp01(r7.w + (r7.x * e.delta.x + r7.y * e.delta.y + r7.z * e.delta.z));
                    const TexQuery q = tex_query(__float_as_int(r4.z), __float_as_int(r4.w), __float_as_int(r5.w), u, v, bilinear);
                    const float* t00 = p.texture + (size_t)q.b00 * C;
                    const float* t01 = p.texture + (size_t)q.b01 * C;
                    const float* t10 = p.texture + (size_t)q.b10 * C;
                    const float* t11 = p.texture + (size_t)q.b11 * C;
#pragma unroll
                    for (int k = 0; k < CMAX; ++k) {
                        if (k < C) {
                            const float val = q.w00 * __ldg(t00 + k) + q.w01 * __ldg(t01 + k) + q.w10 * __ldg(t10 + k) +
                                              q.w11 * __ldg(t11 + k);
                            tex[k] += vis * val;
                        }
                    }
                    if (T > 0.5f) {  // median depth, texture.cu:286-291
                        pix_depth = t_view;
                        pix_depth_index = batch_start + i;
                    }
                    float d = e.t;
                    if (use_ndc) d = (kFar * t_view - kFar * kNear) / ((kFar - kNear) * t_view);
                    // distortion(), helpers.cuh:259-264
                    pix_reg += vis * (d * d * ts.x + ts.z - 2.f * d * ts.y);
                    ts.x += vis; ts.y += vis * d; ts.z += vis * d * d;
                    T = next_T;
                    cur_idx = batch_start + i;
                }
            }
        }
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

extern "C" size_t gstex_texture_workspace_bytes(int n) { return (size_t)(n > 0 ? n : 1) * (sizeof(SplatRec) + sizeof(SplatRecBwd)) + 256; }

extern "C" int gstex_texture_forward(int img_height, int img_width, int block_width, int channels, int n,
                                     int num_intersects, const int32_t* texture_dims,
                                     const int32_t* gaussian_ids_sorted, const int32_t* tile_bins, const float* colors,
                                     const float* opacities, const float* means, const float* scales, float glob_scale,
                                     const float* quats, const float* uv0, const float* umap, const float* vmap,
                                     const float* texture, const float* viewmat, const float* c2w, float fx, float fy,
                                     float cx, float cy, int settings, const float* background, float* out_img,
                                     float* out_depth, float* out_reg, float* out_texture, float* out_normal,
                                     float* final_Ts, int32_t* final_idx, int32_t* depth_idx, float* out_reg_s,
                                     void* workspace, size_t workspace_bytes, void* stream) {
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
    if (workspace_bytes < gstex_texture_workspace_bytes(n)) return GSTEX_E_WORKSPACE;
    SplatRec* recs = ws_recs(workspace);
    if (n > 0) {
        const int rc = launch_pack_records(n, texture_dims, colors, opacities, means, scales, glob_scale, quats, uv0, umap,
                                           vmap, viewmat, c2w, fx, fy, cx, cy, recs, ws_recs_bwd(workspace, n), st);
        if (rc != GSTEX_OK) return rc;
    }
    FwdParams p;
    p.g = make_raster_geom(img_height, img_width, block_width);
    p.C = channels; p.recs = recs; p.ids = gaussian_ids_sorted; p.tile_bins = reinterpret_cast<const int2*>(tile_bins);
    p.texture = texture; p.viewmat = viewmat; p.c2w = c2w; p.background = background;
    p.fx = fx; p.fy = fy; p.cx = cx; p.cy = cy; p.settings = settings;
    p.out_img = out_img; p.out_depth = out_depth; p.out_reg = out_reg; p.out_texture = out_texture;
    p.out_normal = out_normal; p.final_Ts = final_Ts; p.out_reg_s = out_reg_s; p.final_idx = final_idx;
    p.depth_idx = depth_idx;
    const dim3 grid(p.g.tiles_x, p.g.tiles_y);
    if (channels <= 3) texture_forward_kernel<3><<<grid, p.g.threads, 0, st>>>(p);
    else if (channels <= 6) texture_forward_kernel<6><<<grid, p.g.threads, 0, st>>>(p);
    else texture_forward_kernel<GSTEX_MAX_CHANNELS><<<grid, p.g.threads, 0, st>>>(p);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

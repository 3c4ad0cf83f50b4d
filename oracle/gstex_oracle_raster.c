/*
 * gstex_oracle_raster.c -- CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE), part 2:
 * textured forward compositing and its backward, restating the reference CUDA kernels
 *   texture_forward   C/texture.cu:11-329
 *   texture_backward  C/texture.cu:331-760
 * and the device helpers they call (C/helpers.cuh, C/texture_helpers.cuh); "C/" is
 * /root/reference/submodules/GStex_cuda/gstex_cuda/cuda/csrc/.  See gstex_oracle.c for the rules
 * about who may load this library.  fp32 arithmetic like the kernels; per-splat / per-texel
 * gradient sums are accumulated in double (the reference sums them with fp32 atomics in a
 * non-deterministic order, so no summation order is "the" reference one).
 * Viewer-only visualisation bits of `settings` (15, 16, 17-21, 24-29) are not restated.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { float x, y; } f2;
typedef struct { float x, y, z; } f3;
typedef struct { float x, y, z, w; } f4;

static inline float dot3(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline float minf_(float a, float b) { return a < b ? a : b; }
static inline int mini_(int a, int b) { return a < b ? a : b; }
static inline int maxi_(int a, int b) { return a > b ? a : b; }
static inline float cuda_expf_fast(float x) { return exp2f(x * 1.4426950408889634f); }

/* C/helpers.cuh:124-131 */
static inline f3 transform_4x3(const float* m, f3 p) {
    f3 o = {m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3], m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7],
            m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]};
    return o;
}
/* C/helpers.cuh:114-121 */
static inline f3 transform_4x3_rot_only_transposed(const float* m, f3 p) {
    f3 o = {m[0] * p.x + m[4] * p.y + m[8] * p.z, m[1] * p.x + m[5] * p.y + m[9] * p.z,
            m[2] * p.x + m[6] * p.y + m[10] * p.z};
    return o;
}
/* C/helpers.cuh:145-152 */
static inline f2 project_pix(float fx, float fy, f3 pv, float cx, float cy) {
    const float rw = 1.f / (pv.z + 1e-6f);
    f2 o = {(pv.x * rw) * fx + cx, (pv.y * rw) * fy + cy};
    return o;
}
/* C/helpers.cuh:155-164 */
static inline f3 project_pix_vjp(float fx, float fy, f3 pv, f2 v_xy) {
    const float rw = 1.f / (pv.z + 1e-6f);
    const float vpx = fx * v_xy.x, vpy = fy * v_xy.y;
    f3 o = {vpx * rw, vpy * rw, -(vpx * pv.x + vpy * pv.y) * rw * rw};
    return o;
}
/* C/helpers.cuh:166-185: columns of R(quat), quat = (w,x,y,z) assumed normalised. */
static inline void quat_axes(f4 q, f3* a1, f3* a2, f3* a3) {
    const float w = q.x, x = q.y, y = q.z, z = q.w;
    a1->x = 1.f - 2.f * (y * y + z * z); a1->y = 2.f * (x * y + w * z); a1->z = 2.f * (x * z - w * y);
    a2->x = 2.f * (x * y - w * z); a2->y = 1.f - 2.f * (x * x + z * z); a2->z = 2.f * (y * z + w * x);
    a3->x = 2.f * (x * z + w * y); a3->y = 2.f * (y * z - w * x); a3->z = 1.f - 2.f * (x * x + y * y);
}
/* C/helpers.cuh:187-228; vR[c] is the gradient w.r.t. COLUMN c of R (glm column-major). */
static inline f4 quat_to_rotmat_vjp(f4 q, const f3 vR[3]) {
    const float w = q.x, x = q.y, y = q.z, z = q.w;
#define VR(c, r) (((const float*)&vR[c])[r])
    f4 v;
    v.x = 2.f * (x * (VR(1, 2) - VR(2, 1)) + y * (VR(2, 0) - VR(0, 2)) + z * (VR(0, 1) - VR(1, 0)));
    v.y = 2.f * (-2.f * x * (VR(1, 1) + VR(2, 2)) + y * (VR(0, 1) + VR(1, 0)) +
                 z * (VR(0, 2) + VR(2, 0)) + w * (VR(1, 2) - VR(2, 1)));
    v.z = 2.f * (x * (VR(0, 1) + VR(1, 0)) - 2.f * y * (VR(0, 0) + VR(2, 2)) +
                 z * (VR(1, 2) + VR(2, 1)) + w * (VR(2, 0) - VR(0, 2)));
    v.w = 2.f * (x * (VR(0, 2) + VR(2, 0)) + y * (VR(1, 2) + VR(2, 1)) -
                 2.f * z * (VR(0, 0) + VR(1, 1)) + w * (VR(0, 1) - VR(1, 0)));
#undef VR
    return v;
}
/* C/texture_helpers.cuh:336-352 */
static inline void get_ray(const float* c2w, float fx, float fy, float cx, float cy, float px,
                           float py, f3* origin, f3* ray) {
    origin->x = c2w[3]; origin->y = c2w[7]; origin->z = c2w[11];
    const float u = (px - cx) / fx, v = (py - cy) / fy;
    f3 uv1 = {u, v, 1.f};
    f3 r = transform_4x3(c2w, uv1);
    r.x -= origin->x; r.y -= origin->y; r.z -= origin->z;
    const float nrm = sqrtf(r.x * r.x + r.y * r.y + r.z * r.z);
    ray->x = r.x / nrm; ray->y = r.y / nrm; ray->z = r.z / nrm;
}
/* C/texture_helpers.cuh:302-313 (denominator clamp shared with the vjp, :315-334) */
static inline float tt_denom(f3 ray, f3 normal) {
    float d = dot3(normal, ray);
    const float eps = 1e-6f;
    if (0.f <= d && d < eps) d = eps;
    else if (-eps < d && d <= 0.f) d = -eps;
    return d;
}
/* C/texture_helpers.cuh:37-51 (replicate pad is hard-wired on, C/texture.cu:65) */
static inline float clamp01(float v) {
    if (v <= 0.f) v = 0.f;
    if (v >= 1.f) v = 1.f;
    return v;
}

typedef struct {
    float w[4];   /* bilinear (or one-hot) weights ii, ij, ji, jj */
    int base[4];  /* texel index (si + i*w + j) for the four corners, channel 0 -> *C + k */
    float fx_, fy_; /* fractional weights (weight.x, weight.y) */
} texq;

/* C/texture_helpers.cuh:162-215 (index/weights are channel independent apart from "+ channel") */
static inline void texel_query(const int32_t* dim, f2 uv, int bilinear, texq* q) {
    const int h = dim[0], w = dim[1], si = dim[2];
    const float tcx = (float)h * uv.x, tcy = (float)w * uv.y;
    int ix = (int)tcx, iy = (int)tcy;
    const int jx = mini_(ix + 1, h - 1), jy = mini_(iy + 1, w - 1);
    const float wx = tcx - (float)ix, wy = tcy - (float)iy;
    ix = mini_(ix, h - 1);
    iy = mini_(iy, w - 1);
    const float wii = (1.f - wx) * (1.f - wy), wij = (1.f - wx) * wy;
    const float wji = wx * (1.f - wy), wjj = wx * wy;
    q->fx_ = wx; q->fy_ = wy;
    q->base[0] = si + ix * w + iy; q->base[1] = si + ix * w + jy;
    q->base[2] = si + jx * w + iy; q->base[3] = si + jx * w + jy;
    if (bilinear) {
        q->w[0] = wii; q->w[1] = wij; q->w[2] = wji; q->w[3] = wjj;
    } else {
        q->w[0] = q->w[1] = q->w[2] = q->w[3] = 0.f;
        if (wii >= wij && wii >= wji && wii >= wjj) q->w[0] = 1.f;
        else if (wij >= wii && wij >= wji && wij >= wjj) q->w[1] = 1.f;
        else if (wji >= wii && wji >= wij && wji >= wjj) q->w[2] = 1.f;
        else q->w[3] = 1.f;
    }
}

/* Everything the forward and the backward recompute for one (pixel, splat) pair:
 * C/texture.cu:161-215 and :517-558. */
typedef struct {
    f3 mean, scale, ax1, ax2, ax3, delta, p_view;
    f4 quat;
    f2 xy;
    float opac, t, l1, l2, sigma, sigma_blur, nb, bl, e_s, e_b, alpha;
    int skip;
} pair_eval;

static inline void eval_pair(int g, const float* means, const float* scales, float glob_scale,
                             const float* quats, const float* opacities, const float* viewmat,
                             float fx, float fy, float cx, float cy, f3 origin, f3 ray, float px,
                             float py, int use_blur, pair_eval* e) {
    e->mean = *(const f3*)(means + 3 * g);
    e->scale = *(const f3*)(scales + 3 * g);
    e->quat = *(const f4*)(quats + 4 * g);
    e->opac = opacities[g];
    quat_axes(e->quat, &e->ax1, &e->ax2, &e->ax3);
    f3 diff = {e->mean.x - origin.x, e->mean.y - origin.y, e->mean.z - origin.z};
    e->t = dot3(e->ax3, diff) / tt_denom(ray, e->ax3);
    f3 pos = {origin.x + e->t * ray.x, origin.y + e->t * ray.y, origin.z + e->t * ray.z};
    e->delta.x = pos.x - e->mean.x; e->delta.y = pos.y - e->mean.y; e->delta.z = pos.z - e->mean.z;
    e->l1 = dot3(e->delta, e->ax1);
    e->l2 = dot3(e->delta, e->ax2);
    const float is1 = 1.f / (e->scale.x * glob_scale), is2 = 1.f / (e->scale.y * glob_scale);
    e->sigma = 0.5f * (is1 * is1 * e->l1 * e->l1 + is2 * is2 * e->l2 * e->l2);
    e->p_view = transform_4x3(viewmat, e->mean);
    e->xy = project_pix(fx, fy, e->p_view, cx, cy);
    const float dx = e->xy.x - px, dy = e->xy.y - py;
    e->sigma_blur = 0.5f * 2.0f * (dx * dx + dy * dy);
    e->nb = 1.f; e->bl = 0.f;
    if (use_blur && e->sigma_blur < e->sigma) { e->bl = 1.f; e->nb = 0.f; }
    e->e_s = cuda_expf_fast(-e->sigma);
    e->e_b = cuda_expf_fast(-e->sigma_blur);
    e->alpha = minf_(0.99f, e->opac * (e->nb * e->e_s + e->bl * e->e_b));
    e->skip = (e->t < 0.01f || e->t > 1000.0f || e->alpha < 1.f / 255.f);
}

/* ------------------------------------------------------------------------------------------
 * (a8) texture_forward -- C/texture.cu:11-329.  One pixel at a time, splats of the pixel's tile
 * in sorted order.  Outputs must be zero-initialised by the caller (C/texture.cu:837-863).
 * ------------------------------------------------------------------------------------------ */
void orc_texture_forward(int tiles_x, int tiles_y, int bw, int img_w, int img_h, int C,
                         const int32_t* texture_dims, const int32_t* ids_sorted,
                         const int32_t* tile_bins, const float* colors, const float* opacities,
                         const float* means, const float* scales, float glob_scale,
                         const float* quats, const float* uv0s, const float* umaps,
                         const float* vmaps, const float* texture, const float* viewmat,
                         const float* c2w, float fx, float fy, float cx, float cy, int settings,
                         const float* background, float* out_img, float* out_depth, float* out_reg,
                         float* out_texture, float* out_normal, float* final_Ts, int32_t* final_idx,
                         int32_t* depth_idx, float* out_reg_s) {
    const int use_blur = (settings & (1 << 9)) != 0;
    const int use_ndc = (settings & (1 << 10)) != 0;
    const int bilinear = !(settings & (1 << 2));
    const int vis_mode = (settings & (1 << 15)) != 0; /* C/texture.cu:58 */
    const int ntiles = tiles_x * tiles_y;
#pragma omp parallel for schedule(dynamic, 1)
    for (int tile = 0; tile < ntiles; ++tile) {
        const int ty = tile / tiles_x, tx = tile % tiles_x;
        const int r0 = tile_bins[2 * tile], r1 = tile_bins[2 * tile + 1];
        float texacc[64];
        for (int li = 0; li < bw; ++li)
            for (int lj = 0; lj < bw; ++lj) {
                const int bi = ty * bw + li, bj = tx * bw + lj;
                if (bi >= img_h || bj >= img_w) continue;
                const int pix = bi * img_w + bj;
                const float px = (float)bj + 0.5f, py = (float)bi + 0.5f;
                f3 origin, ray;
                get_ray(c2w, fx, fy, cx, cy, px, py, &origin, &ray);
                const float view_depth = viewmat[8] * ray.x + viewmat[9] * ray.y + viewmat[10] * ray.z;
                float T = 1.f;
                int cur_idx = 0, pix_depth_index = -1;
                f3 rgb = {0, 0, 0}, nrm = {0, 0, 0}, ts = {0, 0, 0};
                float pix_depth = 0.f, pix_reg = 0.f;
                for (int k = 0; k < C; ++k) texacc[k] = 0.f;
                for (int idx = r0; idx < r1; ++idx) {
                    const int g = ids_sorted[idx];
                    pair_eval e;
                    eval_pair(g, means, scales, glob_scale, quats, opacities, viewmat, fx, fy, cx, cy,
                              origin, ray, px, py, use_blur, &e);
                    const float next_T = T * (1.f - e.alpha);
                    if (next_T <= 1e-4f) break; /* done (tested before skip, C/texture.cu:216-221) */
                    if (e.skip) continue;
                    const float vis = e.alpha * T;
                    rgb.x += colors[3 * g] * vis; rgb.y += colors[3 * g + 1] * vis; rgb.z += colors[3 * g + 2] * vis;
                    {
                        f3 h = e.ax3; /* C/texture.cu:236-244: flipped towards the camera in visualization mode */
                        if (vis_mode && ray.x * h.x + ray.y * h.y + ray.z * h.z > 0) { h.x = -h.x; h.y = -h.y; h.z = -h.z; }
                        nrm.x += vis * h.x; nrm.y += vis * h.y; nrm.z += vis * h.z;
                    }
                    const float t_view = e.t * view_depth;
                    f3 um = *(const f3*)(umaps + 3 * g), vm = *(const f3*)(vmaps + 3 * g);
                    f2 uv = {uv0s[2 * g] + dot3(um, e.delta), uv0s[2 * g + 1] + dot3(vm, e.delta)};
                    uv.x = clamp01(uv.x); uv.y = clamp01(uv.y);
                    texq q;
                    texel_query(texture_dims + 3 * g, uv, bilinear, &q);
                    for (int k = 0; k < C; ++k) {
                        const float val = q.w[0] * texture[q.base[0] * C + k] + q.w[1] * texture[q.base[1] * C + k] +
                                          q.w[2] * texture[q.base[2] * C + k] + q.w[3] * texture[q.base[3] * C + k];
                        texacc[k] += vis * val;
                    }
                    if (T > 0.5f) { pix_depth = t_view; pix_depth_index = idx; } /* median, :286-291 */
                    {
                        float d = e.t;
                        if (use_ndc) d = (1000.0f * t_view - 1000.0f * 0.01f) / ((1000.0f - 0.01f) * t_view);
                        /* distortion(), C/helpers.cuh:259-264 */
                        pix_reg += vis * (d * d * ts.x + ts.z - 2.f * d * ts.y);
                        ts.x += vis; ts.y += vis * d; ts.z += vis * d * d;
                    }
                    T = next_T;
                    cur_idx = idx;
                }
                final_Ts[pix] = T;
                final_idx[pix] = cur_idx;
                depth_idx[pix] = pix_depth_index;
                out_img[3 * pix] = rgb.x + T * background[0];
                out_img[3 * pix + 1] = rgb.y + T * background[1];
                out_img[3 * pix + 2] = rgb.z + T * background[2];
                out_normal[3 * pix] = nrm.x; out_normal[3 * pix + 1] = nrm.y; out_normal[3 * pix + 2] = nrm.z;
                out_depth[pix] = pix_depth;
                out_reg[pix] = pix_reg;
                out_reg_s[3 * pix] = ts.x; out_reg_s[3 * pix + 1] = ts.y; out_reg_s[3 * pix + 2] = ts.z;
                for (int k = 0; k < C; ++k) out_texture[C * pix + k] = texacc[k];
            }
    }
}

/* ------------------------------------------------------------------------------------------
 * (a9) texture_backward -- C/texture.cu:331-760 (+ vjp helpers C/texture_helpers.cuh:62-80,
 * 252-300,315-334 and C/helpers.cuh:155-164,187-228,266-269).  Back-to-front per pixel from
 * final_index; T recovered by T *= 1/(1-alpha).  The backward kernel evaluates sigma with its
 * own expression (:535-538: sigma_factor*(l1*l1/(sx*sx) + l2*l2/(sy*sy))), restated as such.
 * Gradient buffers are double and must be zero-initialised by the caller.
 * ------------------------------------------------------------------------------------------ */
static inline void dacc(double* p, float v) {
#pragma omp atomic
    *p += (double)v;
}

void orc_texture_backward(int tiles_x, int tiles_y, int bw, int img_w, int img_h, int C,
                          const int32_t* texture_dims, const int32_t* ids_sorted,
                          const int32_t* tile_bins, const float* colors, const float* opacities,
                          const float* means, const float* scales, float glob_scale,
                          const float* quats, const float* uv0s, const float* umaps,
                          const float* vmaps, const float* texture, const float* viewmat,
                          const float* c2w, float fx, float fy, float cx, float cy, int settings,
                          const float* background, const float* final_Ts, const int32_t* final_idx,
                          const int32_t* depth_idx, const float* final_s, const float* v_output,
                          const float* v_output_depth, const float* v_output_reg,
                          const float* v_output_alpha, const float* v_output_texture,
                          const float* v_output_normal, double* v_colors, double* v_opacity,
                          double* v_means, double* v_scales, double* v_quats, double* v_uv0,
                          double* v_umap, double* v_vmap, double* v_texture) {
    const int propagate_uv = (settings & (1 << 8)) != 0;
    const int use_blur = (settings & (1 << 9)) != 0;
    const int use_ndc = (settings & (1 << 10)) != 0;
    const int bilinear = !(settings & (1 << 2));
    const float t_near = 0.01f, t_far = 1000.0f;
    const int ntiles = tiles_x * tiles_y;
#pragma omp parallel for schedule(dynamic, 1)
    for (int tile = 0; tile < ntiles; ++tile) {
        const int ty = tile / tiles_x, tx = tile % tiles_x;
        const int r0 = tile_bins[2 * tile], r1 = tile_bins[2 * tile + 1];
        for (int li = 0; li < bw; ++li)
            for (int lj = 0; lj < bw; ++lj) {
                const int bi = ty * bw + li, bj = tx * bw + lj;
                if (bi >= img_h || bj >= img_w) continue;
                const int pix = bi * img_w + bj;
                const float px = (float)bj + 0.5f, py = (float)bi + 0.5f;
                f3 origin, ray;
                get_ray(c2w, fx, fy, cx, cy, px, py, &origin, &ray);
                const float view_depth = viewmat[8] * ray.x + viewmat[9] * ray.y + viewmat[10] * ray.z;
                float T = final_Ts[pix];
                const f3 ts_final = *(const f3*)(final_s + 3 * pix);
                const int depth_final = depth_idx[pix];
                const int bin_final = final_idx[pix];
                const f3 v_out = *(const f3*)(v_output + 3 * pix);
                const float v_out_depth = v_output_depth[pix], v_out_reg = v_output_reg[pix];
                const float v_out_alpha = v_output_alpha[pix];
                const f3 v_out_normal = *(const f3*)(v_output_normal + 3 * pix);
                const float* v_out_tex = v_output_texture + (size_t)C * pix;
                f3 bg = {background[0], background[1], background[2]};
                float v_T_running = dot3(bg, v_out) - v_out_alpha;
                for (int idx = mini_(r1 - 1, bin_final); idx >= r0; --idx) {
                    const int g = ids_sorted[idx];
                    pair_eval e;
                    eval_pair(g, means, scales, glob_scale, quats, opacities, viewmat, fx, fy, cx, cy,
                              origin, ray, px, py, 0, &e);
                    /* backward's own sigma / alpha, C/texture.cu:535-558 */
                    const float sigma_factor = 0.5f / (glob_scale * glob_scale);
                    const float sigma_1 = e.l1 * e.l1 / (e.scale.x * e.scale.x);
                    const float sigma_2 = e.l2 * e.l2 / (e.scale.y * e.scale.y);
                    const float sigma = sigma_factor * (sigma_1 + sigma_2);
                    float nb = 1.f, bl = 0.f;
                    if (use_blur && e.sigma_blur < sigma) { bl = 1.f; nb = 0.f; }
                    const float e_s = cuda_expf_fast(-sigma), e_b = cuda_expf_fast(-e.sigma_blur);
                    const float alpha = minf_(0.99f, e.opac * (nb * e_s + bl * e_b));
                    if (e.t < 0.01f || e.t > 1000.0f || alpha < 1.f / 255.f) continue;

                    const float ra = 1.f / (1.f - alpha);
                    T *= ra;
                    const float vis = alpha * T;
                    const f3 v_rgb_local = {vis * v_out.x, vis * v_out.y, vis * v_out.z};
                    const f3 v_normal_local = {vis * v_out_normal.x, vis * v_out_normal.y, vis * v_out_normal.z};
                    const f3 rgb = *(const f3*)(colors + 3 * g);
                    float v_vis = dot3(rgb, v_out);
                    v_vis += dot3(e.ax3, v_out_normal);
                    f3 v_diff = {0, 0, 0};
                    float v_alpha = 0.f;
                    f2 v_uv0_l = {0, 0};
                    f3 v_umap_l = {0, 0, 0}, v_vmap_l = {0, 0, 0};
                    {
                        const int32_t* dim = texture_dims + 3 * g;
                        const f3 um = *(const f3*)(umaps + 3 * g), vm = *(const f3*)(vmaps + 3 * g);
                        f2 uv = {uv0s[2 * g] + dot3(um, e.delta), uv0s[2 * g + 1] + dot3(vm, e.delta)};
                        uv.x = clamp01(uv.x); uv.y = clamp01(uv.y);
                        texq q;
                        texel_query(dim, uv, bilinear, &q);
                        f2 v_uv = {0, 0};
                        for (int k = 0; k < C; ++k) {
                            const float t00 = texture[q.base[0] * C + k], t01 = texture[q.base[1] * C + k];
                            const float t10 = texture[q.base[2] * C + k], t11 = texture[q.base[3] * C + k];
                            const float val = q.w[0] * t00 + q.w[1] * t01 + q.w[2] * t10 + q.w[3] * t11;
                            const float v_val = vis * v_out_tex[k];
                            dacc(v_texture + (size_t)q.base[0] * C + k, q.w[0] * v_val);
                            dacc(v_texture + (size_t)q.base[1] * C + k, q.w[1] * v_val);
                            dacc(v_texture + (size_t)q.base[2] * C + k, q.w[2] * v_val);
                            dacc(v_texture + (size_t)q.base[3] * C + k, q.w[3] * v_val);
                            if (bilinear && propagate_uv) {
                                const float vwx = v_val * (-(1.f - q.fy_) * t00 - q.fy_ * t01 + (1.f - q.fy_) * t10 + q.fy_ * t11);
                                const float vwy = v_val * (-(1.f - q.fx_) * t00 + (1.f - q.fx_) * t01 - q.fx_ * t10 + q.fx_ * t11);
                                v_uv.x += (float)dim[0] * vwx;
                                v_uv.y += (float)dim[1] * vwy;
                            }
                            v_vis += val * v_out_tex[k];
                        }
                        /* get_uv_vjp, C/texture_helpers.cuh:62-80 (v_uv is zero unless bit 8) */
                        v_uv0_l.x += v_uv.x; v_uv0_l.y += v_uv.y;
                        v_umap_l.x += e.delta.x * v_uv.x; v_umap_l.y += e.delta.y * v_uv.x; v_umap_l.z += e.delta.z * v_uv.x;
                        v_vmap_l.x += e.delta.x * v_uv.y; v_vmap_l.y += e.delta.y * v_uv.y; v_vmap_l.z += e.delta.z * v_uv.y;
                        if (propagate_uv) {
                            v_diff.x += um.x * v_uv.x + vm.x * v_uv.y;
                            v_diff.y += um.y * v_uv.x + vm.y * v_uv.y;
                            v_diff.z += um.z * v_uv.x + vm.z * v_uv.y;
                        }
                    }
                    float v_t_local = 0.f, v_l1 = 0.f, v_l2 = 0.f;
                    v_alpha += T * v_vis - T * v_T_running;
                    float v_T_cur = alpha * v_vis + (1 - alpha) * v_T_running;
                    float v_t_ndc_local = 0.f, v_t_view_local = 0.f;
                    const float t_view = e.t * view_depth;
                    const float t_ndc = (t_far * t_view - t_far * t_near) / ((t_far - t_near) * t_view);
                    {
                        float v_weight = 0.f;
                        const float d = use_ndc ? t_ndc : e.t;
                        /* distortion_vjp, C/helpers.cuh:266-269 */
                        const float v_val_d = 2.f * (vis * d * ts_final.x - vis * ts_final.y) * v_out_reg;
                        v_weight += (d * d * ts_final.x - 2.f * d * ts_final.y + ts_final.z) * v_out_reg;
                        if (use_ndc) v_t_ndc_local += v_val_d; else v_t_local += v_val_d;
                        v_alpha += v_weight * T;
                        v_T_cur += v_weight * alpha;
                    }
                    v_T_running = v_T_cur;
                    const float v_sigma = -nb * e.opac * e_s * v_alpha;
                    const f3 v_scale_local = {-2 * sigma_factor * sigma_1 * v_sigma / e.scale.x,
                                              -2 * sigma_factor * sigma_2 * v_sigma / e.scale.y, 0.f};
                    v_l1 += 2.f * sigma_factor * e.l1 * v_sigma / (e.scale.x * e.scale.x);
                    v_l2 += 2.f * sigma_factor * e.l2 * v_sigma / (e.scale.y * e.scale.y);
                    v_t_local += (v_l1 * dot3(ray, e.ax1) + v_l2 * dot3(ray, e.ax2));
                    v_t_local += dot3(ray, v_diff);
                    if (idx == depth_final && depth_final != -1) v_t_view_local += v_out_depth;
                    v_t_view_local += (t_far * t_near) / ((t_far - t_near) * t_view * t_view) * v_t_ndc_local;
                    v_t_local += view_depth * v_t_view_local;
                    const float v_sigma_blur = -bl * e.opac * e_b * v_alpha;
                    f2 v_xy = {2.0f * v_sigma_blur * (e.xy.x - px), 2.0f * v_sigma_blur * (e.xy.y - py)};
                    const f3 v_mean_extra =
                        transform_4x3_rot_only_transposed(viewmat, project_pix_vjp(fx, fy, e.p_view, v_xy));
                    /* compute_tt_vjp, C/texture_helpers.cuh:315-334 */
                    const float den = tt_denom(ray, e.ax3);
                    const f3 diff = {e.mean.x - origin.x, e.mean.y - origin.y, e.mean.z - origin.z};
                    const f3 v_mean_tt = {e.ax3.x * v_t_local / den, e.ax3.y * v_t_local / den, e.ax3.z * v_t_local / den};
                    const f3 v_normal_tt = {(diff.x - ray.x * e.t) * v_t_local / den,
                                            (diff.y - ray.y * e.t) * v_t_local / den,
                                            (diff.z - ray.z * e.t) * v_t_local / den};
                    const f3 v_mean_local = {
                        -(e.ax1.x * v_l1 + e.ax2.x * v_l2) + v_mean_tt.x + v_mean_extra.x - v_diff.x,
                        -(e.ax1.y * v_l1 + e.ax2.y * v_l2) + v_mean_tt.y + v_mean_extra.y - v_diff.y,
                        -(e.ax1.z * v_l1 + e.ax2.z * v_l2) + v_mean_tt.z + v_mean_extra.z - v_diff.z};
                    const f3 vR[3] = {{e.delta.x * v_l1, e.delta.y * v_l1, e.delta.z * v_l1},
                                      {e.delta.x * v_l2, e.delta.y * v_l2, e.delta.z * v_l2},
                                      {v_normal_tt.x + v_normal_local.x, v_normal_tt.y + v_normal_local.y,
                                       v_normal_tt.z + v_normal_local.z}};
                    const f4 v_quat_local = quat_to_rotmat_vjp(e.quat, vR);
                    const float v_opacity_local = (nb * e_s + bl * e_b) * v_alpha;

                    dacc(v_colors + 3 * (size_t)g, v_rgb_local.x); dacc(v_colors + 3 * (size_t)g + 1, v_rgb_local.y);
                    dacc(v_colors + 3 * (size_t)g + 2, v_rgb_local.z);
                    dacc(v_scales + 3 * (size_t)g, v_scale_local.x); dacc(v_scales + 3 * (size_t)g + 1, v_scale_local.y);
                    dacc(v_means + 3 * (size_t)g, v_mean_local.x); dacc(v_means + 3 * (size_t)g + 1, v_mean_local.y);
                    dacc(v_means + 3 * (size_t)g + 2, v_mean_local.z);
                    dacc(v_quats + 4 * (size_t)g, v_quat_local.x); dacc(v_quats + 4 * (size_t)g + 1, v_quat_local.y);
                    dacc(v_quats + 4 * (size_t)g + 2, v_quat_local.z); dacc(v_quats + 4 * (size_t)g + 3, v_quat_local.w);
                    dacc(v_opacity + g, v_opacity_local);
                    if (propagate_uv) {
                        dacc(v_uv0 + 2 * (size_t)g, v_uv0_l.x); dacc(v_uv0 + 2 * (size_t)g + 1, v_uv0_l.y);
                        dacc(v_umap + 3 * (size_t)g, v_umap_l.x); dacc(v_umap + 3 * (size_t)g + 1, v_umap_l.y);
                        dacc(v_umap + 3 * (size_t)g + 2, v_umap_l.z);
                        dacc(v_vmap + 3 * (size_t)g, v_vmap_l.x); dacc(v_vmap + 3 * (size_t)g + 1, v_vmap_l.y);
                        dacc(v_vmap + 3 * (size_t)g + 2, v_vmap_l.z);
                    }
                }
            }
    }
}

/* ------------------------------------------------------------------------------------------
 * (f-4) texture_edit -- C/texture_edit.cu:11-236: the forward's traversal (same alpha, skip and
 * termination rules, :130-177) that, instead of compositing, splats the edit canvas into the charts:
 * every contributing pair whose view depth t*view_depth lies in [depth_lower, depth_upper] of its
 * pixel (:180-181) adds, for the four bilinear corners of its uv (always bilinear, replicate
 * padding: edit_jagged_chart, C/texture_helpers.cuh:239-250), weight * (r*m, g*m, b*m, m, 1) to
 * channels 0..4 of the corner texel (:194-216; r,g,b = updated_img, m = updated_alpha).  The blur
 * floor is settings bit 0 here (:48), not bit 9; the NDC bit (1) is read and unused.  The kernel
 * adds with float atomics in an unspecified order; this restatement accumulates in double
 * (`updated_texture`, (T, C) doubles, zero-initialised by the caller; C = texture_info.z >= 5).
 * PARITY UNPINNED on the CPU: the reference has no Python twin and no test of texture_edit
 * (texture_edit.py:117 "only a CUDA implementation was written"); it is pinned on the GPU box
 * against the reference's own extension (tests/test_gpu_edit.py).
 * ------------------------------------------------------------------------------------------ */
void orc_texture_edit(int tiles_x, int tiles_y, int bw, int img_w, int img_h, int C,
                      const int32_t* texture_dims, const float* updated_img, const float* updated_alpha,
                      const float* depth_lower, const float* depth_upper, const int32_t* ids_sorted,
                      const int32_t* tile_bins, const float* opacities, const float* means,
                      const float* scales, float glob_scale, const float* quats, const float* uv0s,
                      const float* umaps, const float* vmaps, const float* viewmat, const float* c2w,
                      float fx, float fy, float cx, float cy, int settings, double* updated_texture,
                      long long* pairs_in_range) {
    const int use_blur = (settings & (1 << 0)) != 0;
    const int ntiles = tiles_x * tiles_y;
    long long npairs = 0;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : npairs)
    for (int tile = 0; tile < ntiles; ++tile) {
        const int ty = tile / tiles_x, tx = tile % tiles_x;
        const int r0 = tile_bins[2 * tile], r1 = tile_bins[2 * tile + 1];
        for (int li = 0; li < bw; ++li)
            for (int lj = 0; lj < bw; ++lj) {
                const int bi = ty * bw + li, bj = tx * bw + lj;
                if (bi >= img_h || bj >= img_w) continue;
                const int pix = bi * img_w + bj;
                const float px = (float)bj + 0.5f, py = (float)bi + 0.5f;
                f3 origin, ray;
                get_ray(c2w, fx, fy, cx, cy, px, py, &origin, &ray);
                const float view_depth = viewmat[8] * ray.x + viewmat[9] * ray.y + viewmat[10] * ray.z;
                const float m = updated_alpha[pix];
                const float vals[5] = {updated_img[3 * pix] * m, updated_img[3 * pix + 1] * m,
                                       updated_img[3 * pix + 2] * m, m, 1.f};
                const float lo = depth_lower[pix], hi = depth_upper[pix];
                float T = 1.f;
                for (int idx = r0; idx < r1; ++idx) {
                    const int g = ids_sorted[idx];
                    pair_eval e;
                    eval_pair(g, means, scales, glob_scale, quats, opacities, viewmat, fx, fy, cx, cy,
                              origin, ray, px, py, use_blur, &e);
                    const float next_T = T * (1.f - e.alpha);
                    if (next_T <= 1e-4f) break;
                    if (e.skip) continue;
                    const float t_view = e.t * view_depth;
                    if (t_view >= lo && t_view <= hi) {
                        f3 um = *(const f3*)(umaps + 3 * g), vm = *(const f3*)(vmaps + 3 * g);
                        f2 uv = {uv0s[2 * g] + dot3(um, e.delta), uv0s[2 * g + 1] + dot3(vm, e.delta)};
                        uv.x = clamp01(uv.x); uv.y = clamp01(uv.y);
                        texq q;
                        texel_query(texture_dims + 3 * g, uv, 1, &q);
                        for (int k = 0; k < 5; ++k)
                            for (int c = 0; c < 4; ++c) dacc(updated_texture + (size_t)q.base[c] * C + k, q.w[c] * vals[k]);
                        ++npairs;
                    }
                    T = next_T;
                }
            }
    }
    if (pairs_in_range) *pairs_in_range = npairs;
}

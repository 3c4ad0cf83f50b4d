// pack.cu -- per-(view, splat) record packer shared by the forward and backward compositing kernels.
// One thread per splat writes
//   * SplatRec     (128 B): everything the per-(pixel,splat) loop needs that does not depend on the pixel;
//   * SplatRecBwd  ( 32 B): quaternion, 1/scale, 1/(z_view+1e-6) for the backward-only VJPs;
//   * a support box (16 B): a CONSERVATIVE pixel-space bounding box of the pixel centres at which the splat can
//     reach alpha >= 1/255.  The compositing kernels use it to skip, per warp, splats that cannot touch the
//     warp's pixel rows.  Culling never changes a result: a culled pair has alpha < 1/255 (it is `skip`ped by
//     the reference too, texture.cu:213) and the kernels fall back to evaluating everything whenever a pixel's
//     transmittance is so close to the 1e-4 termination threshold that even such a pair could end it.
#include "raster_common.cuh"

namespace gstex {

struct PackParams {
    int n;
    const int32_t* texture_dims;
    const float *colors, *opacities, *means, *scales, *quats, *uv0, *umap, *vmap, *viewmat, *c2w;
    float glob_scale, fx, fy, cx, cy;
    int use_blur;
    SplatRec* recs;
    SplatRecBwd* recs_bwd;
    float4* boxes;
    uint32_t* counters;
    const int2* tile_bins;  // when set, the block after the last splat block orders the tiles by list length
    int num_tiles;
    int32_t* order;
};

// Support box.  alpha = opac*exp(-sigma) >= 1/255  <=>  sigma <= smax = ln(255*opac), sigma = (u^2+v^2)/2 in
// units of the splat's standard deviations, so the support is the disc u^2+v^2 <= r^2 = 2*smax of the splat
// plane.  A point of the plane is mean + u*sx*ax1 + v*sy*ax2; in camera coordinates (X,Y,Z) = a*u + b*v + c.
// The screen bounding box of the projected disc has the closed form used for 2D Gaussian splatting (bounding
// box of a conic under a homography): with f = (r^2, r^2, -1) and rows T0,T1,T2 of [a b c],
//   d = sum f*T2*T2,  centre_x = sum f*T0*T2 / d,  half_x^2 = centre_x^2 - sum f*T0*T0 / d   (same for y),
// valid when the disc lies entirely in front of the camera (d < 0, c.z > 0).  Everything in double, r^2
// inflated by 2 % + 0.02 and the box by 0.05 px so that fp32 noise of the exact per-pixel evaluation (relative
// ~1e-4..1e-3 after the cancellation in pos - mean) can never put a contributing pixel outside it.  With the
// blur floor (settings bit 9) alpha uses min(sigma, |xy - pix|^2): the box is unioned with the disc of radius
// sqrt(smax) around the projected mean.  Anything doubtful (disc crossing the camera plane, splat behind the
// camera, non-finite numbers) gets the infinite box = never culled.
__device__ __forceinline__ float4 support_box(const double a[3], const double b[3], const double c[3], double opac,
                                              double fx, double fy, double cx, double cy, bool use_blur, double mean_px,
                                              double mean_py) {
    const float inf = __int_as_float(0x7f800000);
    const float4 everything = make_float4(-inf, -inf, inf, inf);
    const float4 nothing = make_float4(inf, inf, -inf, -inf);
    if (!(opac > 0.0)) return (opac == opac) ? nothing : everything;  // alpha <= 0 < 1/255 always; NaN: no culling
    const double smax = log(255.0 * opac) * 1.02 + 0.02;
    if (smax <= 0.0) return nothing;  // opac*exp(-sigma) < 1/255 for every sigma >= 0
    const double r2 = 2.0 * smax;
    const double d = r2 * (a[2] * a[2] + b[2] * b[2]) - c[2] * c[2];
    if (!(c[2] > 0.0) || !(d < -1e-12 * c[2] * c[2])) return everything;
    const double ccx = (r2 * (a[0] * a[2] + b[0] * b[2]) - c[0] * c[2]) / d;
    const double ccy = (r2 * (a[1] * a[2] + b[1] * b[2]) - c[1] * c[2]) / d;
    double hx2 = ccx * ccx - (r2 * (a[0] * a[0] + b[0] * b[0]) - c[0] * c[0]) / d;
    double hy2 = ccy * ccy - (r2 * (a[1] * a[1] + b[1] * b[1]) - c[1] * c[1]) / d;
    if (!(hx2 == hx2) || !(hy2 == hy2)) return everything;
    const double hx = sqrt(fmax(hx2, 0.0)), hy = sqrt(fmax(hy2, 0.0));
    double x0 = fx * (ccx - hx) + cx, x1 = fx * (ccx + hx) + cx;
    double y0 = fy * (ccy - hy) + cy, y1 = fy * (ccy + hy) + cy;
    if (x0 > x1) { const double t = x0; x0 = x1; x1 = t; }
    if (y0 > y1) { const double t = y0; y0 = y1; y1 = t; }
    if (use_blur) {
        const double rb = sqrt(smax);
        x0 = fmin(x0, mean_px - rb); x1 = fmax(x1, mean_px + rb);
        y0 = fmin(y0, mean_py - rb); y1 = fmax(y1, mean_py + rb);
    }
    const double m = 0.05;
    const float4 box = make_float4(__double2float_rd(x0 - m), __double2float_rd(y0 - m), __double2float_ru(x1 + m),
                                   __double2float_ru(y1 + m));
    if (!(box.x == box.x) || !(box.y == box.y) || !(box.z == box.z) || !(box.w == box.w)) return everything;
    return box;
}

// Longest-processing-time-first order for the forward's and the stream backward's unit cursors: a persistent grid
// that meets its heaviest tiles last idles through a tail as long as one of them (a unit of a dense tile runs for
// hundreds of microseconds).  64 buckets by list length, longest first; the order inside a bucket is arbitrary and
// only ever affects scheduling.
__device__ void order_tiles_by_length(const PackParams& p) {
    __shared__ int s_max, s_hist[64];
    const int t = threadIdx.x;
    if (t == 0) s_max = 0;
    if (t < 64) s_hist[t] = 0;
    __syncthreads();
    int mx = 0;
    for (int k = t; k < p.num_tiles; k += blockDim.x) {
        const int2 r = p.tile_bins[k];
        mx = max(mx, r.y - r.x);
    }
    atomicMax(&s_max, mx);
    __syncthreads();
    const long long denom = (long long)s_max + 1;
    auto bucket = [&](int k) {
        const int2 r = p.tile_bins[k];
        const int len = max(r.y - r.x, 0);
        return 63 - (int)(((long long)len * 64) / denom);
    };
    for (int k = t; k < p.num_tiles; k += blockDim.x) atomicAdd(&s_hist[bucket(k)], 1);
    __syncthreads();
    if (t == 0) {
        int run = 0;
        for (int b = 0; b < 64; ++b) { const int c = s_hist[b]; s_hist[b] = run; run += c; }
    }
    __syncthreads();
    for (int k = t; k < p.num_tiles; k += blockDim.x) p.order[atomicAdd(&s_hist[bucket(k)], 1)] = k;
}

__global__ void __launch_bounds__(kThreads1D) pack_records_kernel(const PackParams p) {
    if (p.tile_bins != nullptr && blockIdx.x == gridDim.x - 1) {
        order_tiles_by_length(p);
        return;
    }
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n) return;
    if (i == 0) {
        p.counters[0] = 0u;  // unit cursor of the forward kernel that follows on the stream
        p.counters[1] = p.tile_bins != nullptr ? 1u : 0u;  // the tile order beside the records is valid
    }
    const float mx = p.means[3 * i], my = p.means[3 * i + 1], mz = p.means[3 * i + 2];
    const float4 qt = reinterpret_cast<const float4*>(p.quats)[i];
    const float w = qt.x, x = qt.y, y = qt.z, z = qt.w;
    // Every fp32 expression below is written with explicit rounding intrinsics, in the fused form (read from the SASS
    // of the build that matched) under which the compositing kernels reproduce the reference extension's results
    // (full-size texture_edit: 5e-7 of scale; the parity tests' 1e-5) -- ptxas fuses plain mul/add pairs as it sees
    // fit, explicit .rn operations it leaves alone: the records then carry the same bits whatever code surrounds them.  (Left to the compiler,
    // an unrelated edit of this kernel changed the contraction of the rotation columns and moved out_texture by 1e-4 of
    // its scale: u = uv0 + umap.delta amplifies one ulp of an axis by |mean| / |splat|.)
    // columns of R(quat), helpers.cuh:166-185
    // a1x: y*y and z*z are shared with a3z and a2y in the reference's forward, so ptxas keeps both products rounded:
    // 1 - 2 * (y*y + z*z) with a plain add (texture_forward SASS); every other entry fuses one product
    const float a1x = __fsub_rn(1.f, __fmul_rn(2.f, __fadd_rn(__fmul_rn(y, y), __fmul_rn(z, z))));
    const float a1y = __fmul_rn(2.f, __fmaf_rn(x, y, __fmul_rn(w, z)));
    const float a1z = __fmul_rn(2.f, __fmaf_rn(x, z, -__fmul_rn(w, y)));
    const float a2x = __fmul_rn(2.f, __fmaf_rn(x, y, -__fmul_rn(w, z)));
    const float a2y = __fsub_rn(1.f, __fmul_rn(2.f, __fmaf_rn(x, x, __fmul_rn(z, z))));
    const float a2z = __fmul_rn(2.f, __fmaf_rn(w, x, __fmul_rn(y, z)));
    const float a3x = __fmul_rn(2.f, __fmaf_rn(x, z, __fmul_rn(w, y)));
    const float a3y = __fmul_rn(2.f, __fmaf_rn(-w, x, __fmul_rn(y, z)));
    const float a3z = __fsub_rn(1.f, __fmul_rn(2.f, __fmaf_rn(x, x, __fmul_rn(y, y))));
    const float ox = __ldg(p.c2w + 3), oy = __ldg(p.c2w + 7), oz = __ldg(p.c2w + 11);
    const float dx = __fsub_rn(mx, ox), dy = __fsub_rn(my, oy), dz = __fsub_rn(mz, oz);
    // compute_tt numerator, texture_helpers.cuh:303,312
    const float tnum = __fmaf_rn(a3z, dz, __fmaf_rn(a3x, dx, __fmul_rn(a3y, dy)));
    const float sx = p.scales[3 * i], sy = p.scales[3 * i + 1];
    const float is1 = __frcp_rn(__fmul_rn(sx, p.glob_scale)), is2 = __frcp_rn(__fmul_rn(sy, p.glob_scale));
    // projected mean (texture.cu:184-185): transform_4x3 + project_pix
    float vm[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) vm[k] = __ldg(p.viewmat + k);
    const float vx = __fadd_rn(vm[3], __fmaf_rn(mz, vm[2], __fmaf_rn(mx, vm[0], __fmul_rn(my, vm[1]))));
    const float vy = __fadd_rn(vm[7], __fmaf_rn(mz, vm[6], __fmaf_rn(mx, vm[4], __fmul_rn(my, vm[5]))));
    const float vz = __fadd_rn(vm[11], __fmaf_rn(mz, vm[10], __fmaf_rn(mx, vm[8], __fmul_rn(my, vm[9]))));
    const float rw = __frcp_rn(__fadd_rn(vz, 1e-6f));
    const float xy_x = __fmaf_rn(p.fx, __fmul_rn(vx, rw), p.cx), xy_y = __fmaf_rn(p.fy, __fmul_rn(vy, rw), p.cy);
    const int h = p.texture_dims[3 * i], wd = p.texture_dims[3 * i + 1], st = p.texture_dims[3 * i + 2];
    const float opac = p.opacities[i];
    SplatRec r;
    r.q[0] = make_float4(mx, my, mz, tnum);
    r.q[1] = make_float4(a3x, a3y, a3z, opac);
    r.q[2] = make_float4(a1x, a1y, a1z, __fmul_rn(is1, is1));
    r.q[3] = make_float4(a2x, a2y, a2z, __fmul_rn(is2, is2));
    r.q[4] = make_float4(xy_x, xy_y, __int_as_float(h), __int_as_float(wd));
    r.q[5] = make_float4(p.colors[3 * i], p.colors[3 * i + 1], p.colors[3 * i + 2], __int_as_float(st));
    r.q[6] = make_float4(p.umap[3 * i], p.umap[3 * i + 1], p.umap[3 * i + 2], p.uv0[2 * i]);
    r.q[7] = make_float4(p.vmap[3 * i], p.vmap[3 * i + 1], p.vmap[3 * i + 2], p.uv0[2 * i + 1]);
#pragma unroll
    for (int k = 0; k < 8; ++k) p.recs[i].q[k] = r.q[k];
    p.recs_bwd[i].quat = qt;
    p.recs_bwd[i].aux = make_float4(__frcp_rn(sx), __frcp_rn(sy), rw, 0.f);

    // support box, in double.  The kernels (like the reference) build R from the quaternion AS GIVEN: for a non-unit
    // quaternion its columns are neither unit nor orthogonal.  The pair evaluation reads l1 = delta.a1, l2 = delta.a2
    // on the plane delta.a3 = 0, so the points with l1 = u*sx, l2 = v*sy are delta = u*sx*c1 + v*sy*c2 with c1, c2 the
    // first two columns of inverse([a1; a2; a3]) = (a2 x a3, a3 x a1) / det -- equal to a1, a2 for a unit quaternion.
    const double gsx = (double)sx * (double)p.glob_scale, gsy = (double)sy * (double)p.glob_scale;
    const double b1[3] = {a1x, a1y, a1z}, b2[3] = {a2x, a2y, a2z}, b3[3] = {a3x, a3y, a3z};
    const double c1r[3] = {b2[1] * b3[2] - b2[2] * b3[1], b2[2] * b3[0] - b2[0] * b3[2], b2[0] * b3[1] - b2[1] * b3[0]};
    const double c2r[3] = {b3[1] * b1[2] - b3[2] * b1[1], b3[2] * b1[0] - b3[0] * b1[2], b3[0] * b1[1] - b3[1] * b1[0]};
    const double det = b1[0] * c1r[0] + b1[1] * c1r[1] + b1[2] * c1r[2];
    const bool frame_ok = fabs(det) > 1e-12 && det == det;
    const double idet = frame_ok ? 1.0 / det : 0.0;
    const double A1[3] = {c1r[0] * idet * gsx, c1r[1] * idet * gsx, c1r[2] * idet * gsx};
    const double A2[3] = {c2r[0] * idet * gsy, c2r[1] * idet * gsy, c2r[2] * idet * gsy};
    double a[3], b[3], c[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        a[k] = (double)vm[4 * k] * A1[0] + (double)vm[4 * k + 1] * A1[1] + (double)vm[4 * k + 2] * A1[2];
        b[k] = (double)vm[4 * k] * A2[0] + (double)vm[4 * k + 1] * A2[1] + (double)vm[4 * k + 2] * A2[2];
        c[k] = (double)vm[4 * k] * mx + (double)vm[4 * k + 1] * my + (double)vm[4 * k + 2] * mz + (double)vm[4 * k + 3];
    }
    const float inf = __int_as_float(0x7f800000);
    p.boxes[i] = frame_ok ? support_box(a, b, c, (double)opac, p.fx, p.fy, p.cx, p.cy, p.use_blur != 0, xy_x, xy_y)
                          : make_float4(-inf, -inf, inf, inf);  // degenerate frame: never culled
}

int launch_pack_records(int n, const int32_t* texture_dims, const float* colors, const float* opacities,
                        const float* means, const float* scales, float glob_scale, const float* quats,
                        const float* uv0, const float* umap, const float* vmap, const float* viewmat,
                        const float* c2w, float fx, float fy, float cx, float cy, int settings, void* workspace,
                        cudaStream_t st, const int32_t* tile_bins, int num_tiles) {
    if (n <= 0) return GSTEX_OK;
    PackParams p;
    const bool order = tile_bins != nullptr && num_tiles > 0 && num_tiles <= kOrderMaxTiles;
    p.tile_bins = order ? reinterpret_cast<const int2*>(tile_bins) : nullptr;
    p.num_tiles = num_tiles; p.order = ws_order(workspace, n);
    p.n = n; p.texture_dims = texture_dims; p.colors = colors; p.opacities = opacities; p.means = means;
    p.scales = scales; p.quats = quats; p.uv0 = uv0; p.umap = umap; p.vmap = vmap; p.viewmat = viewmat; p.c2w = c2w;
    p.glob_scale = glob_scale; p.fx = fx; p.fy = fy; p.cx = cx; p.cy = cy;
    p.use_blur = (settings & GSTEX_SET_BLUR) ? 1 : 0;
    p.recs = ws_recs(workspace); p.recs_bwd = ws_recs_bwd(workspace, n); p.boxes = ws_boxes(workspace, n);
    p.counters = ws_counters(workspace);
    pack_records_kernel<<<div_up(n, kThreads1D) + (order ? 1 : 0), kThreads1D, 0, st>>>(p);
    GSTEX_LAUNCH_CHECK();
    return GSTEX_OK;
}

}  // namespace gstex

extern "C" size_t gstex_texture_workspace_bytes(int n) { return gstex::ws_bytes(n); }

/* gstex_oracle_sample.c -- CPU restatement of the reference's jagged-texture sampler.
 * TEST INFRASTRUCTURE, NOT PRODUCT CODE (see oracle/__init__.py).
 *
 * Follows gstex_cuda/cuda/csrc/texture_sample.cu:11-36 (one query: for every channel, query_jagged_chart with
 * bilinear = replicate_pad = true) and texture_helpers.cuh:155-160 (fix_int_coord), :162-215 (indices / weights),
 * :217-237 (the four-corner sum, in the order wii*tii + wij*tij + wji*tji + wjj*tjj).  uv is clamped to [0,1]
 * first, as the documented contract (texture_sample.py:29-31) and the torch twin (_torch_impl.py:152-153) say.
 * Pinned against the reference's own Python (_torch_impl.sample_texture) by tests/golden/sample_golden.npz.
 */
#include <stdint.h>

static float clamp_unit(float v) { return v <= 0.f ? 0.f : (v >= 1.f ? 1.f : v); }
static int imin(int a, int b) { return a < b ? a : b; }

void orc_texture_sample(int num_queries, int C, const int32_t* texture_dims, const float* uvs, const float* texture,
                        float* out) {
#pragma omp parallel for schedule(static)
    for (int q = 0; q < num_queries; ++q) {
        const int h = texture_dims[3 * q], w = texture_dims[3 * q + 1], si = texture_dims[3 * q + 2];
        const float u = clamp_unit(uvs[2 * q]), v = clamp_unit(uvs[2 * q + 1]);
        const float tx = (float)h * u, ty = (float)w * v;
        int ix = (int)tx, iy = (int)ty;
        const int jx = imin(ix + 1, h - 1), jy = imin(iy + 1, w - 1);
        const float wx = tx - (float)ix, wy = ty - (float)iy;
        ix = imin(ix, h - 1);
        iy = imin(iy, w - 1);
        const float wii = (1.f - wx) * (1.f - wy), wij = (1.f - wx) * wy, wji = wx * (1.f - wy), wjj = wx * wy;
        const int64_t bii = (int64_t)(si + ix * w + iy) * C, bij = (int64_t)(si + ix * w + jy) * C;
        const int64_t bji = (int64_t)(si + jx * w + iy) * C, bjj = (int64_t)(si + jx * w + jy) * C;
        for (int k = 0; k < C; ++k)
            out[(int64_t)q * C + k] = wii * texture[bii + k] + wij * texture[bij + k] + wji * texture[bji + k] +
                                      wjj * texture[bjj + k];
    }
}

"""CPU oracle of re-charting -- TEST INFRASTRUCTURE, never imported by the product.

numpy restatement (fp32, one rounding per operation like eager torch) of
  * GStexModel.build_charts, /root/reference/nerfstudio/models/gstex.py:841-888, and
  * JaggedTexture.init_from_dims + texture_dims_to_query, /root/reference/nerfstudio/models/jagged_texture.py:23-35,116-143
    (bilinear sampling = _torch_impl.sample_texture, gstex_cuda/_torch_impl.py:149-194).
Pinned to tests/golden/charts_golden.npz, which tests/golden/make_golden_charts.py makes by executing the reference's
build_charts source text and its real JaggedTexture class.
"""
from __future__ import annotations

import math

import numpy as np

F = np.float32


def _score(l0, l1, sf, x, recip):
    """gstex.py:847-849.  `tensor / python_float` multiplies by the fp32 reciprocal; `tensor / tensor` divides."""
    x = F(x)
    if recip:
        r = F(1) / x
        a, b = np.ceil((sf * l0) * r), np.ceil((sf * l1) * r)
    else:
        a, b = np.ceil((sf * l0) / x), np.ceil((sf * l1) / x)
    return float((a * b).astype(np.float64).sum())


def build_charts(scales_raw, pixel_num, sigma_factor=3.0, area_sum=None):
    """Returns (texture_dims (N,3) int32, mappings (N,2) f32, pixel_scale float).  `area_sum`: the reference's
    `torch.sum(sf*sf*l0*l1*adj**2).item()` (gstex.py:857) when the caller has it -- a non-integer fp32 sum whose last bits
    depend on the reduction order; numpy's own sum is used otherwise."""
    s = np.asarray(scales_raw, F)
    sf = F(sigma_factor)
    l0, l1 = np.exp(s[:, 0], dtype=F), np.exp(s[:, 1], dtype=F)
    total = float(pixel_num)
    if area_sum is None:
        area_sum = float(((F(sigma_factor * sigma_factor) * l0) * l1).sum(dtype=F))
    lo, hi = 10.0, math.sqrt(area_sum / total)                           # :856-857
    mid = 0.5 * (lo + hi)
    score = _score(l0, l1, sf, mid, True)                                # :859 (python float divisor)
    it, tol = 0, 1e-3
    while score < (1 - tol) * total or score > (1 + tol) * total:        # :863-872
        if score < (1 - tol) * total:
            lo = mid
        else:
            hi = mid
        mid = 0.5 * (lo + hi)
        score = _score(l0, l1, sf, mid, False)                           # tensor divisor: mid / adjustments
        it += 1
        if it > 30:
            break
    x = F(mid)
    dims = np.zeros((s.shape[0], 3), np.int32)
    dims[:, 0] = np.ceil((sf * l0) / x)                                  # :876-877
    dims[:, 1] = np.ceil((sf * l1) / x)
    hw = dims[:, 0] * dims[:, 1]
    dims[:, 2] = np.cumsum(hw, dtype=np.int32) - hw                      # :878-879
    two_sf = F(2.0 * float(sigma_factor))
    mappings = np.stack([F(1) / (two_sf * l0), F(1) / (two_sf * l1)], -1).astype(F)   # :882-883
    return dims, mappings, mid


def resample_texture(old_dims, old_texture, new_dims):
    """jagged_texture.py:126-141 with sample_texture's bilinear rule (replicate padding, weights from the fractional
    part of h*u, w*v; _torch_impl.py:149-194)."""
    od, nd = np.asarray(old_dims, np.int64), np.asarray(new_dims, np.int64)
    tex = np.asarray(old_texture, F)
    hw = nd[:, 0] * nd[:, 1]
    ids = np.repeat(np.arange(nd.shape[0]), hw)                          # :118-120
    q_new, q_old = nd[ids], od[ids]
    local = np.arange(int(hw.sum())) - q_new[:, 2]
    uu = (local // q_new[:, 1]).astype(F) / q_new[:, 0].astype(F)        # texture_dims_to_query, :31-33
    vv = (local % q_new[:, 1]).astype(F) / q_new[:, 1].astype(F)
    u, v = np.clip(uu, 0, 1).astype(F), np.clip(vv, 0, 1).astype(F)
    h, w, start = q_old[:, 0], q_old[:, 1], q_old[:, 2]
    tcx, tcy = h.astype(F) * u, w.astype(F) * v
    ix, iy = tcx.astype(np.int64), tcy.astype(np.int64)
    fx, fy = (tcx - ix.astype(F)).astype(F), (tcy - iy.astype(F)).astype(F)
    jx, jy = np.minimum(ix + 1, h - 1), np.minimum(iy + 1, w - 1)
    ix, iy = np.minimum(ix, h - 1), np.minimum(iy, w - 1)
    w00, w01 = ((1 - fx) * (1 - fy))[:, None], ((1 - fx) * fy)[:, None]
    w10, w11 = (fx * (1 - fy))[:, None], (fx * fy)[:, None]
    g = lambda a, b: tex[start + a * w + b]                              # noqa: E731
    return (w00 * g(ix, iy) + w01 * g(ix, jy) + w10 * g(jx, iy) + w11 * g(jx, jy)).astype(F)

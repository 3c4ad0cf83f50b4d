"""Re-charting: `build_charts` and the resampling of `JaggedTexture.init_from_dims`.

Mirrors what the reference model does every 100 steps (`retexture_after`):
  * `GStexModel.build_charts` (/root/reference/nerfstudio/models/gstex.py:841-888): bisection on the texel size until
    `sum_i ceil(sf * l0_i / x) * ceil(sf * l1_i / x)` is within 0.1 % of `pixel_num`, then `texture_dims = {h, w, start}`
    and `mappings = 1 / (2 * sf * l)`;
  * `JaggedTexture.init_from_dims` (/root/reference/nerfstudio/models/jagged_texture.py:116-143): every texel of the new
    charts is a bilinear sample of the splat's old chart.
The reference spends up to 31 rounds x (~8 torch launches + a blocking `.item()`) on the bisection and materialises three
per-texel index tensors for the resampling; here the candidates of a whole 5-level bisection subtree are scored by one
kernel pass (<= 7 passes for the reference's 31 rounds) and the resampling is one kernel (`csrc/charts.cu`).  The host
loop below takes the reference's decisions in the reference's order, so `pixel_scale`, `texture_dims` and `mappings`
come out bit for bit (scores are summed exactly; torch's fp32 sum is exact too while it stays below 2^24).
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch
from torch import Tensor

from . import _lib
from .utils import cumsum_i32

TOL = 1e-3          # gstex.py:862
MAX_ROUNDS = 30     # gstex.py:871-872
_DEPTH = 5          # bisection levels scored per pass (2^5 - 1 = 31 candidates <= gstex_chart_max_candidates())


def _scores(scales: Tensor, sigma_factor: float, xs, recip_mask: int) -> list:
    """One pass over the splats: the scores of all candidate texel sizes `xs` (python floats)."""
    import ctypes

    dev = scales.device
    k = len(xs)
    host = (ctypes.c_float * k)(*[float(x) for x in xs])
    out = torch.empty(k, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.load().gstex_chart_scores(scales.shape[0], _lib.ptr(scales), float(sigma_factor), k, host, int(recip_mask),
                                            _lib.ptr(out), _lib.stream_ptr(dev))
    _lib.check(rc, "gstex_chart_scores")
    return out.tolist()  # the one read-back of the pass (the reference reads one score back per round)


def chart_texel_size(scales: Tensor, pixel_num: float, sigma_factor: float = 3.0) -> Tuple[float, int]:
    """The bisection of gstex.py:855-873.  Returns `(pixel_scale, passes)`: the python float the reference leaves in
    `mid`, and how many kernel passes it took."""
    if not pixel_num > 0:
        raise ValueError("pixel_num must be positive (the reference divides by it, gstex.py:857)")
    sf, total = float(sigma_factor), float(pixel_num)
    with torch.no_grad():
        # gstex.py:857, the bracket's upper end, with the reference's own torch expression (adjustments == 1): a
        # non-integer fp32 sum, which only the same reduction reproduces bit for bit -- one reduction per re-chart
        length0, length1 = torch.exp(scales[:, 0]), torch.exp(scales[:, 1])
        hi = math.sqrt(torch.sum(sf * sf * length0 * length1).item() / total)
    lo = 10.0
    rounds, passes, first = 0, 0, True
    while True:
        # heap-ordered subtree of the bisection rooted at (lo, hi): child 2i+1 follows `lo = mid`, child 2i+2 `hi = mid`
        nodes = [(lo, hi)]
        for i in range((1 << (_DEPTH - 1)) - 1):
            l, h = nodes[i]
            m = 0.5 * (l + h)
            nodes += [None] * (2 * i + 3 - len(nodes))
            nodes[2 * i + 1], nodes[2 * i + 2] = (m, h), (l, m)
        mids = [0.5 * (l + h) for l, h in nodes]
        scores = _scores(scales, sf, mids, 1 if first else 0)  # only the very first score divides by a python float
        passes += 1
        first = False
        i = 0
        while True:
            score = scores[i]
            if not (score < (1 - TOL) * total or score > (1 + TOL) * total):
                return mids[i], passes                      # the reference's loop condition fails: done
            child = 2 * i + 1 if score < (1 - TOL) * total else 2 * i + 2
            l, h = nodes[i]
            lo, hi = (mids[i], h) if child == 2 * i + 1 else (l, mids[i])
            rounds += 1
            if rounds > MAX_ROUNDS:                         # gstex.py:871-872: leaves with the new mid, unscored
                return 0.5 * (lo + hi), passes
            if child >= len(nodes):
                break                                        # next pass, rooted at (lo, hi)
            i = child


def build_charts(scales: Tensor, pixel_num: float, sigma_factor: float = 3.0, update_pixel_scale: bool = True,
                 pixel_scale: float = 1.0):
    """ref: `GStexModel.build_charts` (gstex.py:841-888) up to, not including, `texture_dc.init_from_dims`.

    `scales` (N,3): the model's raw (log) scales.  Returns `(texture_dims (N,3) int32 {h, w, start}, mappings (N,2),
    pixel_scale: float, total_texels: int)`.  `update_pixel_scale=False` keeps the reference's `pixel_scales = 1`
    (gstex.py:851) or the given `pixel_scale`."""
    scales = scales.detach().contiguous()
    dev = _lib.require_cuda(scales)
    _lib.require_dtype(scales, torch.float32, "scales")
    if scales.dim() != 2 or scales.shape[1] != 3:
        raise ValueError("scales must be (N, 3)")
    n = scales.shape[0]
    if update_pixel_scale:
        pixel_scale, _ = chart_texel_size(scales, pixel_num, sigma_factor)
    dims = torch.empty((n, 3), dtype=torch.int32, device=dev)
    hw = torch.empty((n,), dtype=torch.int32, device=dev)
    mappings = torch.empty((n, 2), dtype=torch.float32, device=dev)
    lib, p = _lib.load(), _lib.ptr
    with torch.cuda.device(dev):
        # `pixel_scales = mid / adjustments` is an fp32 tensor in the reference (gstex.py:874): the texel size is fp32(mid)
        rc = lib.gstex_chart_dims(n, p(scales), float(sigma_factor), float(pixel_scale), p(dims), p(hw), p(mappings),
                                  _lib.stream_ptr(dev))
    _lib.check(rc, "gstex_chart_dims")
    if n == 0:
        return dims, mappings, float(pixel_scale), 0
    cum, total = cumsum_i32(hw)
    with torch.cuda.device(dev):
        rc = lib.gstex_chart_starts(n, p(cum), p(hw), p(dims), _lib.stream_ptr(dev))
    _lib.check(rc, "gstex_chart_starts")
    return dims, mappings, float(pixel_scale), int(total.item())


def resample_texture(old_dims: Tensor, old_texture: Tensor, new_dims: Tensor, total_size: Optional[int] = None) -> Tensor:
    """ref: the non-initial branch of `JaggedTexture.init_from_dims` (jagged_texture.py:126-141): the `(total_size, C)`
    texture of the new charts, every texel sampled bilinearly from its splat's old chart.  `old_dims` / `new_dims`
    (N,3) int32 {h, w, start}; `total_size` = sum of the new h * w (read back from `new_dims` when not given)."""
    old_dims, new_dims = old_dims.contiguous(), new_dims.contiguous()
    old_texture = old_texture.detach().contiguous()
    dev = _lib.require_cuda(old_dims, new_dims, old_texture)
    _lib.require_dtype(old_dims, torch.int32, "old_dims")
    _lib.require_dtype(new_dims, torch.int32, "new_dims")
    _lib.require_dtype(old_texture, torch.float32, "old_texture")
    assert new_dims.shape[0] == old_dims.shape[0], "the charts must belong to the same splats (jagged_texture.py:118)"
    n, C = new_dims.shape[0], old_texture.shape[-1]
    if total_size is None:
        total_size = int((new_dims[-1, 2] + new_dims[-1, 0] * new_dims[-1, 1]).item()) if n > 0 else 0
    out = torch.empty((int(total_size), C), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.load().gstex_texture_resample(int(total_size), n, C, _lib.ptr(new_dims), _lib.ptr(old_dims),
                                                _lib.ptr(old_texture), _lib.ptr(out), _lib.stream_ptr(dev))
    _lib.check(rc, "gstex_texture_resample")
    return out

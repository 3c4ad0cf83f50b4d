"""Per-splat colour stage in front of the rasterizer: `spherical_harmonics`, the drop-in for the reference's
`gstex_cuda/sh.py` (/root/reference/submodules/GStex_cuda/gstex_cuda/sh.py:10-96): same helper names, same
signature, differentiable with respect to the coefficients only.  SURVEY.md section 8f, rank 1.
"""
from __future__ import annotations

import torch
from torch import Tensor
from torch.autograd import Function

from . import _lib


def num_sh_bases(degree: int) -> int:
    """ref: sh.py:10-19."""
    if degree == 0:
        return 1
    if degree == 1:
        return 4
    if degree == 2:
        return 9
    if degree == 3:
        return 16
    return 25


def deg_from_sh(num_bases: int) -> int:
    """ref: sh.py:22-33."""
    for deg, nb in enumerate((1, 4, 9, 16, 25)):
        if num_bases == nb:
            return deg
    assert False, "Invalid number of SH bases"


def spherical_harmonics(degrees_to_use: int, viewdirs: Tensor, coeffs: Tensor) -> Tensor:
    """Colours (N,3) of N splats from `viewdirs` (N,3; normalised by the kernel) and SH `coeffs` (N,D,3), using the
    bases of degree <= `degrees_to_use`.  Differentiable with respect to `coeffs` only (ref: sh.py:36-58)."""
    assert coeffs.shape[-2] >= num_sh_bases(degrees_to_use)
    return _SphericalHarmonics.apply(degrees_to_use, viewdirs.contiguous(), coeffs.contiguous())


def _check(viewdirs: Tensor, other: Tensor, n: int, what: str):
    dev = _lib.require_cuda(viewdirs, other)
    _lib.require_dtype(viewdirs, torch.float32, "viewdirs")
    _lib.require_dtype(other, torch.float32, what)
    if viewdirs.dim() != 2 or viewdirs.shape[0] != n or viewdirs.shape[1] != 3:
        raise RuntimeError("viewdirs must have dimensions (N, 3)")  # bindings.cu:53-56
    return dev


class _SphericalHarmonics(Function):
    """ref: sh.py:61-96."""

    @staticmethod
    def forward(ctx, degrees_to_use: int, viewdirs: Tensor, coeffs: Tensor):
        if coeffs.dim() != 3 or coeffs.shape[2] != 3:
            raise RuntimeError("coeffs must have dimensions (N, D, 3)")  # bindings.cu:27-30
        num_points = coeffs.shape[0]
        degree = deg_from_sh(coeffs.shape[-2])
        dev = _check(viewdirs, coeffs, num_points, "coeffs")
        ctx.degrees_to_use = degrees_to_use
        ctx.degree = degree
        ctx.save_for_backward(viewdirs)
        colors = torch.empty((num_points, 3), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            rc = _lib.load().gstex_compute_sh_forward(num_points, degree, int(degrees_to_use), _lib.ptr(viewdirs),
                                                      _lib.ptr(coeffs), _lib.ptr(colors), _lib.stream_ptr(dev))
        _lib.check(rc, "gstex_compute_sh_forward")
        return colors

    @staticmethod
    def backward(ctx, v_colors: Tensor):
        viewdirs = ctx.saved_tensors[0]
        v_colors = v_colors.contiguous()
        num_points = v_colors.shape[0]
        if v_colors.dim() != 2 or v_colors.shape[1] != 3:
            raise RuntimeError("v_colors must have dimensions (N, 3)")  # bindings.cu:57-60
        dev = _check(viewdirs, v_colors, num_points, "v_colors")
        v_coeffs = torch.empty((num_points, num_sh_bases(ctx.degree), 3), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            rc = _lib.load().gstex_compute_sh_backward(num_points, ctx.degree, int(ctx.degrees_to_use),
                                                       _lib.ptr(viewdirs), _lib.ptr(v_colors), _lib.ptr(v_coeffs),
                                                       _lib.stream_ptr(dev))
        _lib.check(rc, "gstex_compute_sh_backward")
        return None, None, v_coeffs

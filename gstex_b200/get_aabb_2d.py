"""Stage (a): projection, screen-space bounds and tile counts.

Same names, argument order and results as the reference's `gstex_cuda/get_aabb_2d.py`
(/root/reference/submodules/GStex_cuda/gstex_cuda/get_aabb_2d.py:22-91); each call is one hand-written
sm_100a kernel behind the C ABI instead of a cuBLAS gemm plus ~25 small torch kernels.
"""
from __future__ import annotations

from typing import Tuple

import torch
from torch import Tensor

from . import _lib


def _viewmat12(viewmat: Tensor) -> Tensor:
    """The reference passes `viewmat[:3,:]` or a full 4x4; the kernels read the first 12 floats."""
    vm = viewmat.contiguous()
    if vm.numel() < 12:
        raise RuntimeError("viewmat must hold at least 3x4 floats")
    _lib.require_dtype(vm, torch.float32, "viewmat")
    return vm


def _f32c(t: Tensor, name: str) -> Tensor:
    t = t.contiguous()
    _lib.require_dtype(t, torch.float32, name)
    return t


def project_points(points: Tensor, viewmat: Tensor, intrins, clip: bool = False) -> Tuple[Tensor, Tensor]:
    """ref: get_aabb_2d.py:22-32.  Returns (xys (N,2), depths (N,)).  Not differentiable here: the
    model only feeds `depths` to the (non-differentiable) binning, NS/models/gstex.py:1077."""
    if clip:
        raise NotImplementedError("clip=True is only used by the reference's unused get_aabb_2d_torch")
    points, vm = _f32c(points, "points"), _viewmat12(viewmat)
    dev = _lib.require_cuda(points, vm)
    fx, fy, cx, cy = (float(v) for v in intrins)
    n = points.shape[0]
    xys = torch.empty((n, 2), dtype=torch.float32, device=dev)
    depths = torch.empty((n,), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.load().gstex_project_points(_lib.ptr(points), _lib.ptr(vm), fx, fy, cx, cy, n, _lib.ptr(xys),
                                              _lib.ptr(depths), _lib.stream_ptr(dev))
    _lib.check(rc, "gstex_project_points")
    return xys, depths


def get_aabb_2d(means: Tensor, scales: Tensor, glob_scale, quats: Tensor, viewmat: Tensor,
                intrins) -> Tuple[Tensor, Tensor]:
    """ref: get_aabb_2d.py:56-68 -> `_C.get_aabb_2d` (cuda/csrc/get_aabb_2d.cu:91-126).
    Returns (centers (N,2), extents (N,2)); clipped splats (view z <= 0.01) get extent 0."""
    means, scales, quats = _f32c(means.detach(), "means"), _f32c(scales.detach(), "scales"), _f32c(quats.detach(), "quats")
    vm = _viewmat12(viewmat)
    dev = _lib.require_cuda(means, scales, quats, vm)
    fx, fy, cx, cy = (float(v) for v in intrins)
    n = means.shape[0]
    centers = torch.empty((n, 2), dtype=torch.float32, device=dev)
    extents = torch.empty((n, 2), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.load().gstex_get_aabb_2d(_lib.ptr(means), _lib.ptr(scales), float(glob_scale), _lib.ptr(quats),
                                           _lib.ptr(vm), fx, fy, cx, cy, n, _lib.ptr(centers), _lib.ptr(extents),
                                           _lib.stream_ptr(dev))
    _lib.check(rc, "gstex_get_aabb_2d")
    return centers, extents


def get_num_tiles_hit_2d(centers: Tensor, extents: Tensor, img_height: int, img_width: int,
                         block_width: int) -> Tensor:
    """ref: get_aabb_2d.py:70-91.  Returns (N,) int32."""
    centers, extents = _f32c(centers, "centers"), _f32c(extents, "extents")
    dev = _lib.require_cuda(centers, extents)
    n = centers.shape[0]
    out = torch.empty((n,), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.load().gstex_get_num_tiles_hit_2d(_lib.ptr(centers), _lib.ptr(extents), n, int(img_height),
                                                    int(img_width), int(block_width), _lib.ptr(out),
                                                    _lib.stream_ptr(dev))
    _lib.check(rc, "gstex_get_num_tiles_hit_2d")
    return out


def project_bounds(means: Tensor, scales: Tensor, glob_scale, quats: Tensor, viewmat: Tensor, intrins,
                   img_height: int, img_width: int, block_width: int):
    """The three calls above in ONE kernel launch (what `GStexModel.get_outputs` needs every step,
    NS/models/gstex.py:1075-1080).  Returns (xys, depths, centers, extents, num_tiles_hit)."""
    means, scales, quats = _f32c(means.detach(), "means"), _f32c(scales.detach(), "scales"), _f32c(quats.detach(), "quats")
    vm = _viewmat12(viewmat)
    dev = _lib.require_cuda(means, scales, quats, vm)
    fx, fy, cx, cy = (float(v) for v in intrins)
    n = means.shape[0]
    xys = torch.empty((n, 2), dtype=torch.float32, device=dev)
    depths = torch.empty((n,), dtype=torch.float32, device=dev)
    centers = torch.empty((n, 2), dtype=torch.float32, device=dev)
    extents = torch.empty((n, 2), dtype=torch.float32, device=dev)
    nth = torch.empty((n,), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.load().gstex_project_bounds(_lib.ptr(means), _lib.ptr(scales), float(glob_scale), _lib.ptr(quats),
                                              _lib.ptr(vm), fx, fy, cx, cy, n, int(img_height), int(img_width),
                                              int(block_width), _lib.ptr(xys), _lib.ptr(depths), _lib.ptr(centers),
                                              _lib.ptr(extents), _lib.ptr(nth), _lib.stream_ptr(dev))
    _lib.check(rc, "gstex_project_bounds")
    return xys, depths, centers, extents, nth

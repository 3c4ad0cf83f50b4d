"""CPU oracle for the GStex hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes/numpy front end of the plain-C restatement in ``gstex_oracle.c`` / ``gstex_oracle_raster.c``
(each C function cites the reference file:line it follows).  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may
import this package; ``gstex_b200`` never does.

Function names mirror the reference's Python entry points
(``/root/reference/submodules/GStex_cuda/gstex_cuda/{get_aabb_2d,utils,texture}.py``) so that the
parity tests read like the reference's own.  All arrays are numpy, C-contiguous, fp32 / int32 /
int64.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_BUILD = os.path.join(_HERE, "_build")
_SO = os.path.join(_BUILD, "libgstex_oracle.so")
_SRCS = [os.path.join(_HERE, "gstex_oracle.c"), os.path.join(_HERE, "gstex_oracle_raster.c"),
         os.path.join(_HERE, "gstex_oracle_sh.c"), os.path.join(_HERE, "gstex_oracle_sample.c"),
         os.path.join(_HERE, "gstex_oracle_loss.c")]

_lib = None


def build(force: bool = False) -> str:
    """Compile the C restatement with gcc (seconds).  Returns the path of the shared object."""
    os.makedirs(_BUILD, exist_ok=True)
    stale = force or not os.path.exists(_SO)
    if not stale:
        so_m = os.path.getmtime(_SO)
        stale = any(os.path.exists(s) and os.path.getmtime(s) > so_m for s in _SRCS)
    if stale:
        if not all(os.path.exists(s) for s in _SRCS):
            raise FileNotFoundError("oracle sources missing")
        cmd = ["gcc", "-O2", "-fopenmp", "-ffp-contract=off", "-shared", "-fPIC", *_SRCS, "-o", _SO, "-lm"]
        subprocess.run(cmd, check=True)
    return _SO


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        _lib.orc_num_threads.restype = ctypes.c_int
        _lib.orc_cumsum_i32.restype = ctypes.c_int32
    return _lib


def num_threads() -> int:
    return int(lib().orc_num_threads())


def set_num_threads(n: int) -> None:
    lib().orc_set_num_threads(ctypes.c_int(int(n)))


def _f32(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def _i32(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a, dtype=np.int32))


def _p(a: np.ndarray):
    return a.ctypes.data_as(ctypes.c_void_p)


def _cf(x) -> ctypes.c_float:
    return ctypes.c_float(float(x))


def tile_bounds_of(img_height: int, img_width: int, block_width: int) -> Tuple[int, int, int]:
    return ((img_width + block_width - 1) // block_width, (img_height + block_width - 1) // block_width, 1)


# ---------------------------------------------------------------------------------------------
# binning (reference: get_aabb_2d.py, utils.py)
# ---------------------------------------------------------------------------------------------
def project_points(points, viewmat, intrins):
    """get_aabb_2d.py:22-32 (clip=False).  viewmat: (>=3,4) row-major."""
    points = _f32(points)
    vm = _f32(np.asarray(viewmat, dtype=np.float32).reshape(-1)[:12])
    n = points.shape[0]
    xys = np.empty((n, 2), np.float32)
    depths = np.empty((n,), np.float32)
    fx, fy, cx, cy = intrins
    lib().orc_project_points(_p(points), _p(vm), _cf(fx), _cf(fy), _cf(cx), _cf(cy), ctypes.c_int(n), _p(xys), _p(depths))
    return xys, depths


def get_aabb_2d(means, scales, glob_scale, quats, viewmat, intrins):
    """get_aabb_2d.py:56-68 -> C/get_aabb_2d.cu:11-126."""
    means, scales, quats = _f32(means), _f32(scales), _f32(quats)
    vm = _f32(np.asarray(viewmat, dtype=np.float32).reshape(-1)[:12])
    n = means.shape[0]
    centers = np.zeros((n, 2), np.float32)
    extents = np.zeros((n, 2), np.float32)
    fx, fy, cx, cy = intrins
    lib().orc_get_aabb_2d(_p(means), _p(scales), _cf(glob_scale), _p(quats), _p(vm), _cf(fx), _cf(fy), _cf(cx),
                          _cf(cy), ctypes.c_int(n), _p(centers), _p(extents))
    return centers, extents


def get_num_tiles_hit_2d(centers, extents, img_height, img_width, block_width):
    """get_aabb_2d.py:70-91."""
    centers, extents = _f32(centers), _f32(extents)
    n = centers.shape[0]
    out = np.empty((n,), np.int32)
    lib().orc_get_num_tiles_hit_2d(_p(centers), _p(extents), ctypes.c_int(n), ctypes.c_int(img_height),
                                   ctypes.c_int(img_width), ctypes.c_int(block_width), _p(out))
    return out


def compute_cumulative_intersects(num_tiles_hit):
    """utils.py:40-59."""
    nth = _i32(num_tiles_hit)
    cum = np.empty_like(nth)
    total = lib().orc_cumsum_i32(_p(nth), ctypes.c_int(nth.shape[0]), _p(cum))
    return int(total), cum


def map_gaussian_to_intersects(num_points, num_intersects, centers, extents, depths, cum_tiles_hit, tile_bounds,
                               block_size):
    """utils.py:61-104 -> C/bindings.cu:77-121 -> C/forward.cu:13-71."""
    centers, extents, depths, cum = _f32(centers), _f32(extents), _f32(depths), _i32(cum_tiles_hit)
    isect = np.zeros((num_intersects,), np.int64)
    gids = np.zeros((num_intersects,), np.int32)
    lib().orc_map_gaussian_to_intersects(ctypes.c_int(num_points), ctypes.c_int(num_intersects), _p(centers),
                                         _p(extents), _p(depths), _p(cum), ctypes.c_int(tile_bounds[0]),
                                         ctypes.c_int(tile_bounds[1]), ctypes.c_int(block_size), _p(isect), _p(gids))
    return isect, gids


def sort_isects(isect_ids, gaussian_ids):
    """torch.sort + torch.gather of utils.py:159-160 (stable ascending)."""
    k = np.ascontiguousarray(isect_ids, dtype=np.int64)
    v = _i32(gaussian_ids)
    ko, vo = np.empty_like(k), np.empty_like(v)
    lib().orc_sort_isects(ctypes.c_int(k.shape[0]), _p(k), _p(v), _p(ko), _p(vo))
    return ko, vo


def get_tile_bin_edges(num_intersects, isect_ids_sorted, tile_bounds):
    """utils.py:11-37 -> C/forward.cu:76-98."""
    k = np.ascontiguousarray(isect_ids_sorted, dtype=np.int64)
    nt = tile_bounds[0] * tile_bounds[1]
    bins = np.zeros((nt, 2), np.int32)
    lib().orc_get_tile_bin_edges(ctypes.c_int(num_intersects), _p(k), ctypes.c_int(nt), _p(bins))
    return bins


def bin_and_sort_gaussians(num_points, num_intersects, centers, extents, depths, cum_tiles_hit, tile_bounds,
                           block_size):
    """utils.py:106-162."""
    isect, gids = map_gaussian_to_intersects(num_points, num_intersects, centers, extents, depths, cum_tiles_hit,
                                             tile_bounds, block_size)
    isect_s, gids_s = sort_isects(isect, gids)
    bins = get_tile_bin_edges(num_intersects, isect_s, tile_bounds)
    return isect, gids, isect_s, gids_s, bins


# ---- spherical harmonics (gstex_cuda/sh.py:36-96 -> cuda/csrc/sh.cuh) ------------------------------------------

def num_sh_bases(degree: int) -> int:
    return {0: 1, 1: 4, 2: 9, 3: 16}.get(int(degree), 25)


def deg_from_sh(num_bases: int) -> int:
    return {1: 0, 4: 1, 9: 2, 16: 3, 25: 4}[int(num_bases)]


def spherical_harmonics(degrees_to_use: int, viewdirs: np.ndarray, coeffs: np.ndarray) -> np.ndarray:
    """colors (N,3); ref: sh.py:36-58 -> compute_sh_forward (sh.cuh:217-235)."""
    viewdirs = np.ascontiguousarray(viewdirs, np.float32)
    coeffs = np.ascontiguousarray(coeffs, np.float32)
    n = coeffs.shape[0]
    assert coeffs.shape[-2] >= num_sh_bases(degrees_to_use)
    out = np.empty((n, 3), np.float32)
    f = ctypes.POINTER(ctypes.c_float)
    lib().orc_sh_forward(ctypes.c_int(n), ctypes.c_int(deg_from_sh(coeffs.shape[-2])), ctypes.c_int(int(degrees_to_use)),
                         viewdirs.ctypes.data_as(f), coeffs.ctypes.data_as(f), out.ctypes.data_as(f))
    return out


def spherical_harmonics_vjp(degrees_to_use: int, num_bases: int, viewdirs: np.ndarray, v_colors: np.ndarray) -> np.ndarray:
    """v_coeffs (N, num_bases, 3); ref: sh.py:84-96 -> compute_sh_backward (sh.cuh:237-253)."""
    viewdirs = np.ascontiguousarray(viewdirs, np.float32)
    v_colors = np.ascontiguousarray(v_colors, np.float32)
    n = v_colors.shape[0]
    out = np.empty((n, int(num_bases), 3), np.float32)
    f = ctypes.POINTER(ctypes.c_float)
    lib().orc_sh_backward(ctypes.c_int(n), ctypes.c_int(deg_from_sh(num_bases)), ctypes.c_int(int(degrees_to_use)),
                          viewdirs.ctypes.data_as(f), v_colors.ctypes.data_as(f), out.ctypes.data_as(f))
    return out


# ---- jagged-texture sampling (gstex_cuda/texture_sample.py:13-86 -> cuda/csrc/texture_sample.cu) ---------------

def texture_sample(texture_info, texture_dims: np.ndarray, texture: np.ndarray, uvs: np.ndarray) -> np.ndarray:
    """out (Q, C); ref: texture_sample.py:13-48 -> texture_sample_forward (texture_sample.cu:11-36)."""
    texture_dims = np.ascontiguousarray(texture_dims, np.int32)
    texture = np.ascontiguousarray(texture, np.float32)
    uvs = np.ascontiguousarray(uvs, np.float32)
    C = int(texture_info[2])
    q = uvs.shape[0]
    out = np.empty((q, C), np.float32)
    f = ctypes.POINTER(ctypes.c_float)
    lib().orc_texture_sample(ctypes.c_int(q), ctypes.c_int(C), texture_dims.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)),
                             uvs.ctypes.data_as(f), texture.ctypes.data_as(f), out.ctypes.data_as(f))
    return out


# ---- compositing + photometric loss (nerfstudio/models/gstex.py:1204-1205, 1297-1302, 1319) -------------------

def photometric_loss(out_img: np.ndarray, out_texture: np.ndarray, out_alpha: np.ndarray, background: np.ndarray,
                     gt_img: np.ndarray, ssim_lambda: float = 0.2, mask=None, v_main: float = 1.0, grads: bool = True):
    """Returns dict(rgb, Ll1, ssim, main_loss[, v_out_img, v_out_texture, v_out_alpha]); SSIM as pytorch_msssim's
    `SSIM(data_range=1.0, size_average=True, channel=3)` (gstex.py:351), window 11 / sigma 1.5."""
    out_img = np.ascontiguousarray(out_img, np.float32)
    out_texture = np.ascontiguousarray(out_texture, np.float32)
    out_alpha = np.ascontiguousarray(out_alpha, np.float32)
    background = np.ascontiguousarray(background, np.float32)
    gt_img = np.ascontiguousarray(gt_img, np.float32)
    H, W, C = out_texture.shape
    f = ctypes.POINTER(ctypes.c_float)
    null = ctypes.cast(None, f)
    if mask is not None:
        mask = np.ascontiguousarray(np.asarray(mask, np.float32).reshape(H, W))
    rgb = np.empty((H, W, 3), np.float32)
    losses = np.empty(3, np.float32)
    v_img = np.empty((H, W, 3), np.float32) if grads else None
    v_tex = np.empty((H, W, C), np.float32) if grads else None
    v_alpha = np.empty((H, W), np.float32) if grads else None
    fn = lib().orc_photometric_loss
    fn.restype = ctypes.c_int
    rc = fn(ctypes.c_int(H), ctypes.c_int(W), ctypes.c_int(C), out_img.ctypes.data_as(f), out_texture.ctypes.data_as(f),
            out_alpha.ctypes.data_as(f), background.ctypes.data_as(f), gt_img.ctypes.data_as(f),
            mask.ctypes.data_as(f) if mask is not None else null, ctypes.c_float(ssim_lambda), ctypes.c_float(1.5),
            ctypes.c_float(v_main), rgb.ctypes.data_as(f), losses.ctypes.data_as(f),
            v_img.ctypes.data_as(f) if grads else null, v_tex.ctypes.data_as(f) if grads else null,
            v_alpha.ctypes.data_as(f) if grads else null)
    if rc != 0:
        raise ValueError("photometric_loss: image smaller than the 11x11 window, or fewer than 3 texture channels")
    out = dict(rgb=rgb, Ll1=float(losses[0]), ssim=float(losses[1]), main_loss=float(losses[2]))
    if grads:
        out.update(v_out_img=v_img, v_out_texture=v_tex, v_out_alpha=v_alpha)
    return out

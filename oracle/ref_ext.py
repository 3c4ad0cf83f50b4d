"""Loader for the reference's own CUDA extension built by oracle/build_ref.py (TEST INFRASTRUCTURE).

`load()` returns the pybind11 module `_C` of the reference (functions of cuda/csrc/ext.cpp:8-24:
get_aabb_2d, map_gaussian_to_intersects, get_tile_bin_edges, texture_forward, texture_backward, ...)
or None when `oracle/_ref/gstex_ref_C.so` is absent.  `RefPipeline` restates in a few lines what the
reference's Python wrappers do around those calls (gstex_cuda/utils.py:106-162, texture.py:186-289),
because the wrappers themselves cannot travel to the GPU box.
"""
from __future__ import annotations

import importlib.util
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "_ref", "gstex_ref_C.so")
_mod = None


def available() -> bool:
    return os.path.exists(SO_PATH)


def load():
    global _mod
    if _mod is not None:
        return _mod
    if not available():
        return None
    import torch  # noqa: F401  (libtorch must be loaded first)

    spec = importlib.util.spec_from_file_location("gstex_ref_C", SO_PATH)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    _mod = mod
    return mod


def ref_project_points(points, viewmat, intrins):
    """get_aabb_2d.py:22-32 + _torch_impl.py:140-147 with torch ops on the tensors' device."""
    fx, fy, cx, cy = intrins
    view_points = points @ viewmat.T[:3, :3] + viewmat.T[3:, :3]
    rw = 1.0 / (view_points[..., 2] + 1e-6)
    import torch

    pix = torch.stack([view_points[..., 0] * rw * fx + cx, view_points[..., 1] * rw * fy + cy], dim=-1)
    return pix, view_points[:, -1]


def ref_get_num_tiles_hit_2d(centers, extents, img_height, img_width, block_width):
    """get_aabb_2d.py:70-91 with torch ops."""
    import torch

    tb = ((img_width + block_width - 1) // block_width, (img_height + block_width - 1) // block_width)
    top_left = torch.floor((centers - extents) / block_width).to(torch.int32)
    bottom_right = torch.floor((centers + extents) / block_width + 1).to(torch.int32)
    tile_min = torch.stack([torch.clamp(top_left[..., 0], 0, tb[0]), torch.clamp(top_left[..., 1], 0, tb[1])], -1)
    tile_max = torch.stack([torch.clamp(bottom_right[..., 0], 0, tb[0]), torch.clamp(bottom_right[..., 1], 0, tb[1])], -1)
    return (tile_max - tile_min)[:, 0] * (tile_max - tile_min)[:, 1]


def ref_bin_and_sort(C, num_points, num_intersects, centers, extents, depths, cum_tiles_hit, tile_bounds, bw):
    """utils.py:106-162 around the reference's native calls."""
    import torch

    isect_ids, gaussian_ids = C.map_gaussian_to_intersects(num_points, num_intersects, centers, extents, depths,
                                                           cum_tiles_hit, tile_bounds, bw, False)
    isect_ids_sorted, sorted_indices = torch.sort(isect_ids)
    gaussian_ids_sorted = torch.gather(gaussian_ids, 0, sorted_indices)
    tile_bins = C.get_tile_bin_edges(num_intersects, isect_ids_sorted.contiguous(), tile_bounds)
    return isect_ids, gaussian_ids, isect_ids_sorted, gaussian_ids_sorted, tile_bins


def ref_forward_backward(C, s, centers, extents, depths, num_tiles_hit, v=None, settings=None):
    """texture.py:186-289 (+ :291-378 when `v` holds the upstream gradients) through the reference's
    native functions.  `s` is a scene dict of CUDA tensors (gstex_b200.synthetic.make_scene)."""
    import torch

    H, W, bw = s["H"], s["W"], s.get("block_width", 16)
    settings = s["settings"] if settings is None else settings
    tb = ((W + bw - 1) // bw, (H + bw - 1) // bw, 1)
    cum = torch.cumsum(num_tiles_hit, dim=0, dtype=torch.int32)
    num_intersects = int(cum[-1].item())
    isect, gids, isect_s, gids_s, bins = ref_bin_and_sort(C, s["N"], num_intersects, centers, extents, depths, cum, tb, bw)
    vm, c2w = s["viewmat"].contiguous(), s["c2w"].contiguous()
    outs = C.texture_forward(tb, (bw, bw, 1), (W, H, 1), s["texture_info"], s["texture_dims"], gids_s, bins,
                             s["colors"], s["opacities"], s["means"], s["scales"], float(s["glob_scale"]), s["quats"],
                             s["uv0"], s["umap"], s["vmap"], s["texture"], vm, c2w, s["fx"], s["fy"], s["cx"],
                             s["cy"], int(settings), s["background"])
    names = ["out_img", "out_depth", "out_reg", "out_texture", "out_normal", "final_Ts", "final_idx", "depth_idx",
             "out_reg_s"]
    res = dict(zip(names, outs))
    res.update(num_intersects=num_intersects, cum_tiles_hit=cum, isect_ids=isect, gaussian_ids=gids,
               isect_ids_sorted=isect_s, gaussian_ids_sorted=gids_s, tile_bins=bins,
               out_alpha=1 - res["final_Ts"])
    if v is not None:
        g = C.texture_backward(H, W, bw, s["texture_info"], s["texture_dims"], gids_s, bins, s["colors"],
                               s["opacities"], s["means"], s["scales"], float(s["glob_scale"]), s["quats"], s["uv0"],
                               s["umap"], s["vmap"], s["texture"], vm, c2w, s["fx"], s["fy"], s["cx"], s["cy"],
                               int(settings), s["background"], res["final_Ts"], res["final_idx"], res["depth_idx"],
                               res["out_reg_s"], v["v_out_img"], v["v_out_depth"], v["v_out_reg"], v["v_out_alpha"],
                               v["v_out_texture"], v["v_out_normal"])
        res.update(dict(zip(["v_colors", "v_opacity", "v_means", "v_scales", "v_quats", "v_uv0", "v_umap", "v_vmap",
                             "v_texture"], g)))
    return res

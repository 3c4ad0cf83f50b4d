"""Oracle front end for the forward / backward rasterizer (TEST INFRASTRUCTURE, NOT PRODUCT CODE).

``texture_forward`` / ``texture_backward`` restate the reference's native functions
(``C/texture.cu:762-901`` and ``:903-1053``); ``texture_gaussians_fwd_bwd`` strings the whole
path together the way ``gstex_cuda/texture.py:155-408`` does.
"""
from __future__ import annotations

import ctypes
from typing import Dict, Optional

import numpy as np

from . import (_cf, _f32, _i32, _p, bin_and_sort_gaussians, compute_cumulative_intersects, lib, tile_bounds_of)

SUPPORTED_SETTINGS_MASK = (1 << 2) | (1 << 8) | (1 << 9) | (1 << 10) | (1 << 15)


def _check_settings(settings: int) -> None:
    if settings & ~SUPPORTED_SETTINGS_MASK:
        raise NotImplementedError(f"oracle restates settings bits 2, 8, 9, 10, 15 only (got {settings:#x})")


def texture_forward(tile_bounds, block_width, img_size, channels, texture_dims, gaussian_ids_sorted, tile_bins,
                    colors, opacities, means, scales, glob_scale, quats, uv0, umap, vmap, texture, viewmat, c2w, fx,
                    fy, cx, cy, settings, background) -> Dict[str, np.ndarray]:
    _check_settings(settings)
    W, H = img_size[0], img_size[1]
    C = int(channels)
    assert C <= 64
    o = dict(
        out_img=np.zeros((H, W, 3), np.float32), out_depth=np.zeros((H, W), np.float32),
        out_reg=np.zeros((H, W), np.float32), out_texture=np.zeros((H, W, C), np.float32),
        out_normal=np.zeros((H, W, 3), np.float32), final_Ts=np.zeros((H, W), np.float32),
        final_idx=np.zeros((H, W), np.int32), depth_idx=np.zeros((H, W), np.int32),
        out_reg_s=np.zeros((H, W, 3), np.float32),
    )
    a = [_i32(texture_dims), _i32(gaussian_ids_sorted), _i32(tile_bins), _f32(colors), _f32(opacities), _f32(means),
         _f32(scales)]
    b = [_f32(quats), _f32(uv0), _f32(umap), _f32(vmap), _f32(texture),
         _f32(np.asarray(viewmat, np.float32).reshape(-1)[:12]), _f32(np.asarray(c2w, np.float32).reshape(-1)[:12])]
    bg = _f32(background)
    lib().orc_texture_forward(
        ctypes.c_int(tile_bounds[0]), ctypes.c_int(tile_bounds[1]), ctypes.c_int(block_width), ctypes.c_int(W),
        ctypes.c_int(H), ctypes.c_int(C), *[_p(x) for x in a], _cf(glob_scale), *[_p(x) for x in b], _cf(fx), _cf(fy),
        _cf(cx), _cf(cy), ctypes.c_int(settings), _p(bg), _p(o["out_img"]), _p(o["out_depth"]), _p(o["out_reg"]),
        _p(o["out_texture"]), _p(o["out_normal"]), _p(o["final_Ts"]), _p(o["final_idx"]), _p(o["depth_idx"]),
        _p(o["out_reg_s"]))
    return o


def texture_backward(img_height, img_width, block_width, channels, texture_dims, gaussian_ids_sorted, tile_bins,
                     colors, opacities, means, scales, glob_scale, quats, uv0, umap, vmap, texture, viewmat, c2w, fx,
                     fy, cx, cy, settings, background, final_Ts, final_idx, depth_idx, final_s, v_output,
                     v_output_depth, v_output_reg, v_output_alpha, v_output_texture,
                     v_output_normal) -> Dict[str, np.ndarray]:
    _check_settings(settings)
    H, W, C = int(img_height), int(img_width), int(channels)
    tb = tile_bounds_of(H, W, block_width)
    n = np.asarray(means).shape[0]
    T = np.asarray(texture).shape[0]
    g = dict(v_colors=np.zeros((n, 3)), v_opacity=np.zeros((n, 1)), v_means=np.zeros((n, 3)),
             v_scales=np.zeros((n, 3)), v_quats=np.zeros((n, 4)), v_uv0=np.zeros((n, 1, 2)),
             v_umap=np.zeros((n, 1, 3)), v_vmap=np.zeros((n, 1, 3)), v_texture=np.zeros((T, C)))
    a = [_i32(texture_dims), _i32(gaussian_ids_sorted), _i32(tile_bins), _f32(colors), _f32(opacities), _f32(means),
         _f32(scales)]
    b = [_f32(quats), _f32(uv0), _f32(umap), _f32(vmap), _f32(texture),
         _f32(np.asarray(viewmat, np.float32).reshape(-1)[:12]), _f32(np.asarray(c2w, np.float32).reshape(-1)[:12])]
    c = [_f32(background), _f32(final_Ts), _i32(final_idx), _i32(depth_idx), _f32(final_s), _f32(v_output),
         _f32(v_output_depth), _f32(v_output_reg), _f32(v_output_alpha), _f32(v_output_texture),
         _f32(v_output_normal)]
    lib().orc_texture_backward(
        ctypes.c_int(tb[0]), ctypes.c_int(tb[1]), ctypes.c_int(block_width), ctypes.c_int(W), ctypes.c_int(H),
        ctypes.c_int(C), *[_p(x) for x in a], _cf(glob_scale), *[_p(x) for x in b], _cf(fx), _cf(fy), _cf(cx),
        _cf(cy), ctypes.c_int(settings), *[_p(x) for x in c], _p(g["v_colors"]), _p(g["v_opacity"]),
        _p(g["v_means"]), _p(g["v_scales"]), _p(g["v_quats"]), _p(g["v_uv0"]), _p(g["v_umap"]), _p(g["v_vmap"]),
        _p(g["v_texture"]))
    return {k: v.astype(np.float32) for k, v in g.items()}


def texture_gaussians_fwd_bwd(texture_info, texture_dims, centers, extents, depths, num_tiles_hit, colors, opacity,
                              means, scales, glob_scale, quats, uv0, umap, vmap, texture, viewmat, c2w, fx, fy, cx,
                              cy, img_height, img_width, block_width, settings, background,
                              v_outs: Optional[Dict[str, np.ndarray]] = None) -> Dict[str, np.ndarray]:
    """Whole path of gstex_cuda/texture.py:155-408 on the CPU: cumsum, bin+sort, forward, and (when
    `v_outs` holds v_out_img/v_out_depth/v_out_reg/v_out_alpha/v_out_texture/v_out_normal) backward."""
    H, W = int(img_height), int(img_width)
    C = int(texture_info[2])
    tb = tile_bounds_of(H, W, block_width)
    n = np.asarray(centers).shape[0]
    num_intersects, cum = compute_cumulative_intersects(num_tiles_hit)
    res: Dict[str, np.ndarray] = {"num_intersects": np.int64(num_intersects), "cum_tiles_hit": cum}
    if num_intersects < 1:
        raise ValueError("no intersections (the reference path is ill-formed here, texture.py:197-205)")
    isect, gids, isect_s, gids_s, bins = bin_and_sort_gaussians(n, num_intersects, centers, extents, depths, cum, tb,
                                                                block_width)
    res.update(isect_ids=isect, gaussian_ids=gids, isect_ids_sorted=isect_s, gaussian_ids_sorted=gids_s,
               tile_bins=bins)
    fwd = texture_forward(tb, block_width, (W, H, 1), C, texture_dims, gids_s, bins, colors, opacity, means, scales,
                          glob_scale, quats, uv0, umap, vmap, texture, viewmat, c2w, fx, fy, cx, cy, settings,
                          background)
    res.update(fwd)
    res["out_alpha"] = (1.0 - fwd["final_Ts"]).astype(np.float32)
    if v_outs is not None:
        v_alpha = v_outs.get("v_out_alpha")
        if v_alpha is None:
            v_alpha = np.zeros((H, W), np.float32)
        bwd = texture_backward(H, W, block_width, C, texture_dims, gids_s, bins, colors, opacity, means, scales,
                               glob_scale, quats, uv0, umap, vmap, texture, viewmat, c2w, fx, fy, cx, cy, settings,
                               background, fwd["final_Ts"], fwd["final_idx"], fwd["depth_idx"], fwd["out_reg_s"],
                               v_outs["v_out_img"], v_outs["v_out_depth"], v_outs["v_out_reg"], v_alpha,
                               v_outs["v_out_texture"], v_outs["v_out_normal"])
        res.update(bwd)
    return res


def texture_edit(texture_info, texture_dims, updated_img, updated_alpha, depth_lower, depth_upper, centers, extents,
                 depths, num_tiles_hit, opacity, means, scales, glob_scale, quats, uv0, umap, vmap, viewmat, c2w, fx,
                 fy, cx, cy, img_height, img_width, block_width, settings, background=None, return_pairs=False):
    """gstex_cuda/texture_edit.py:14-239 on the CPU (binning as texture_gaussians, then `orc_texture_edit` =
    C/texture_edit.cu:11-236).  Returns updated_texture (sum h*w, C) float32 [and the number of contributing pairs
    inside their pixel's depth range]."""
    H, W, C = int(img_height), int(img_width), int(texture_info[2])
    assert C >= 5, "texture_edit writes channels 0..4 (C/texture_edit.cu:194)"
    tb = tile_bounds_of(H, W, block_width)
    n = np.asarray(centers).shape[0]
    td = _i32(texture_dims)
    total = int((td[:, 0].astype(np.int64) * td[:, 1].astype(np.int64)).sum())
    num_intersects, cum = compute_cumulative_intersects(num_tiles_hit)
    assert num_intersects >= 1  # texture_edit.py:183-184
    _, _, _, gids_s, bins = bin_and_sort_gaussians(n, num_intersects, centers, extents, depths, cum, tb, block_width)
    out = np.zeros((total, C), np.float64)
    a = [td, _f32(updated_img), _f32(updated_alpha), _f32(depth_lower), _f32(depth_upper), _i32(gids_s), _i32(bins),
         _f32(opacity), _f32(means), _f32(scales)]
    b = [_f32(quats), _f32(uv0), _f32(umap), _f32(vmap),
         _f32(np.asarray(viewmat, np.float32).reshape(-1)[:12]), _f32(np.asarray(c2w, np.float32).reshape(-1)[:12])]
    npairs = ctypes.c_longlong(0)
    lib().orc_texture_edit(
        ctypes.c_int(tb[0]), ctypes.c_int(tb[1]), ctypes.c_int(block_width), ctypes.c_int(W), ctypes.c_int(H),
        ctypes.c_int(C), *[_p(x) for x in a], _cf(glob_scale), *[_p(x) for x in b], _cf(fx), _cf(fy), _cf(cx), _cf(cy),
        ctypes.c_int(int(settings)), _p(out), ctypes.byref(npairs))
    res = out.astype(np.float32)
    return (res, int(npairs.value)) if return_pairs else res

"""`texture_edit`, the drop-in for the reference's `gstex_cuda/texture_edit.py`
(/root/reference/submodules/GStex_cuda/gstex_cuda/texture_edit.py:14-239): paints an edited RGBA canvas seen from one
camera back into the per-splat charts.  Called by `GStexModel.edit_texture` (nerfstudio/models/gstex.py:571-605) right
after a `texture_gaussians` render of the same view whose median depth gives the per-pixel depth window.
SURVEY.md section 8f, rank 4.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
from torch import Tensor

from . import _lib
from .utils import bin_gaussians_to_tiles


def texture_edit(
    texture_info: Tuple[int, int, int],
    texture_dims: Tensor,
    updated_img: Tensor,
    updated_alpha: Tensor,
    depth_lower: Tensor,
    depth_upper: Tensor,
    centers: Tensor,
    extents: Tensor,
    depths: Tensor,
    num_tiles_hit: Tensor,
    opacity: Tensor,
    means: Tensor,
    scales: Tensor,
    glob_scale,
    quats: Tensor,
    uv0: Tensor,
    umap: Tensor,
    vmap: Tensor,
    viewmat: Tensor,
    c2w: Tensor,
    fx: float,
    fy: float,
    cx: float,
    cy: float,
    img_height: int,
    img_width: int,
    block_width: int,
    settings: int,
    background: Optional[Tensor] = None,
    use_torch_impl: bool = False,
) -> Tensor:
    """Same positional signature as the reference (texture_edit.py:14-44).  Returns `updated_texture`
    (sum h*w, texture_info[2]) with channels (r*a, g*a, b*a, a, 1), each weighted by the bilinear intersection weights
    and summed over the contributing (pixel, splat) pairs inside the pixel's depth window.  `settings`: only bit 0
    (blur floor) is read, as in the reference kernel (cuda/csrc/texture_edit.cu:48-49); `background` is accepted and
    unused (the reference passes it to a kernel that never reads it).  Not differentiable, as in the reference
    (`_TextureEdit.apply` is a plain static method, texture_edit.py:150-152)."""
    assert block_width > 1 and block_width <= 16, "block_width must be between 2 and 16"
    if use_torch_impl:
        raise NotImplementedError("gstex_b200 has no PyTorch/CPU path: the CUDA kernel is the only one")
    texture_dims = texture_dims.contiguous()
    args = [t.detach().contiguous() for t in (updated_img, updated_alpha, depth_lower, depth_upper, centers, extents,
                                              depths, num_tiles_hit, opacity, means, scales, quats, uv0, umap, vmap,
                                              viewmat, c2w)]
    (updated_img, updated_alpha, depth_lower, depth_upper, centers, extents, depths, num_tiles_hit, opacity, means,
     scales, quats, uv0, umap, vmap, viewmat, c2w) = args
    dev = _lib.require_cuda(texture_dims, *args)
    _lib.require_dtype(texture_dims, torch.int32, "texture_dims")
    for t, nm in zip(args, ("updated_img", "updated_alpha", "depth_lower", "depth_upper", "centers", "extents", "depths",
                            None, "opacity", "means", "scales", "quats", "uv0", "umap", "vmap", "viewmat", "c2w")):
        if nm is not None:
            _lib.require_dtype(t, torch.float32, nm)
    H, W, C = int(img_height), int(img_width), int(texture_info[2])
    if C < 5:
        raise ValueError("texture_info[2] must be >= 5: channels 0..4 are written (cuda/csrc/texture_edit.cu:194)")
    if updated_img.numel() != H * W * 3 or updated_alpha.numel() != H * W or depth_lower.numel() != H * W or \
            depth_upper.numel() != H * W:
        raise RuntimeError("updated_img must be (H, W, 3); updated_alpha, depth_lower, depth_upper (H, W[, 1])")
    if viewmat.numel() < 12 or c2w.numel() < 12:
        raise RuntimeError("viewmat and c2w must hold at least 3x4 floats")
    num_points = centers.size(0)
    tile_bounds = ((W + block_width - 1) // block_width, (H + block_width - 1) // block_width, 1)
    # texture_edit.py:179: the size of the jagged texture is read back from the chart table (a host sync, as there)
    texture_total_size = int(torch.sum(texture_dims[:, 0].long() * texture_dims[:, 1].long()).item())
    num_intersects, gaussian_ids_sorted, tile_bins = bin_gaussians_to_tiles(
        num_points, centers, extents, depths, num_tiles_hit, tile_bounds, block_width)
    assert num_intersects >= 1  # texture_edit.py:183-184: `assert False` without intersections
    lib = _lib.load()
    updated_texture = torch.empty((texture_total_size, C), dtype=torch.float32, device=dev)  # zero-filled by the call
    ws = torch.empty((lib.gstex_texture_workspace_bytes(num_points),), dtype=torch.uint8, device=dev)
    p = _lib.ptr
    with torch.cuda.device(dev):
        rc = lib.gstex_texture_edit(
            H, W, int(block_width), C, num_points, int(num_intersects), texture_total_size, p(texture_dims),
            p(updated_img), p(updated_alpha), p(depth_lower), p(depth_upper), p(gaussian_ids_sorted), p(tile_bins),
            p(opacity), p(means), p(scales), float(glob_scale), p(quats), p(uv0), p(umap), p(vmap), p(viewmat), p(c2w),
            float(fx), float(fy), float(cx), float(cy), int(settings), p(updated_texture), p(ws), ws.numel(),
            _lib.stream_ptr(dev))
    _lib.check(rc, "gstex_texture_edit")
    return updated_texture

"""Stage (b): prefix sum, key emission, radix sort, tile ranges.

Same names, argument order and results as the reference's `gstex_cuda/utils.py`
(/root/reference/submodules/GStex_cuda/gstex_cuda/utils.py:11-162).  `torch.cumsum`, `torch.sort` and
`torch.gather` of the reference are replaced by hand-written kernels behind the C ABI.
"""
from __future__ import annotations

from typing import Tuple

import torch
from torch import Tensor

from . import _lib


def _workspace(nbytes: int, dev: torch.device) -> Tensor:
    return torch.empty((max(int(nbytes), 16),), dtype=torch.uint8, device=dev)


def key_bits_for(tile_bounds) -> int:
    """Significant bits of the (tile<<32 | depth) keys: 32 depth bits + bits of the largest tile id."""
    num_tiles = int(tile_bounds[0]) * int(tile_bounds[1])
    return 32 + max(1, (max(num_tiles, 1) - 1).bit_length())


def get_tile_bin_edges(num_intersects: int, isect_ids_sorted: Tensor, tile_bounds) -> Tensor:
    """ref: utils.py:11-37 -> `_C.get_tile_bin_edges` (cuda/csrc/bindings.cu:123-141).  (tiles,2) int32;
    untouched tiles stay (0,0)."""
    isect_ids_sorted = isect_ids_sorted.contiguous()
    dev = _lib.require_cuda(isect_ids_sorted)
    _lib.require_dtype(isect_ids_sorted, torch.int64, "isect_ids_sorted")
    num_tiles = int(tile_bounds[0]) * int(tile_bounds[1])
    tile_bins = torch.empty((num_tiles, 2), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.load().gstex_get_tile_bin_edges(int(num_intersects), _lib.ptr(isect_ids_sorted), num_tiles,
                                                  _lib.ptr(tile_bins), _lib.stream_ptr(dev))
    _lib.check(rc, "gstex_get_tile_bin_edges")
    return tile_bins


def cumsum_i32(num_tiles_hit: Tensor) -> Tuple[Tensor, Tensor]:
    """Inclusive int32 prefix sum; returns (cum, total) with `total` a 1-element device tensor (no sync)."""
    nth = num_tiles_hit.contiguous()
    dev = _lib.require_cuda(nth)
    _lib.require_dtype(nth, torch.int32, "num_tiles_hit")
    n = nth.numel()
    cum = torch.empty((n,), dtype=torch.int32, device=dev)
    total = torch.empty((1,), dtype=torch.int32, device=dev)
    lib = _lib.load()
    ws = _workspace(lib.gstex_cumsum_workspace_bytes(n), dev)
    with torch.cuda.device(dev):
        rc = lib.gstex_cumsum_i32(_lib.ptr(nth), n, _lib.ptr(cum), _lib.ptr(total), _lib.ptr(ws), ws.numel(),
                                  _lib.stream_ptr(dev))
    _lib.check(rc, "gstex_cumsum_i32")
    return cum, total


def compute_cumulative_intersects(num_tiles_hit: Tensor) -> Tuple[int, Tensor]:
    """ref: utils.py:40-59.  Returns (num_intersects: int, cum_tiles_hit (N,) int32).  The python int is a
    blocking device->host read, exactly as in the reference (`.item()`, utils.py:58)."""
    if num_tiles_hit.numel() == 0:
        return 0, torch.empty((0,), dtype=torch.int32, device=num_tiles_hit.device)
    cum, total = cumsum_i32(num_tiles_hit.reshape(-1))
    return int(total.item()), cum


def map_gaussian_to_intersects(num_points: int, num_intersects: int, centers: Tensor, extents: Tensor,
                               depths: Tensor, cum_tiles_hit: Tensor, tile_bounds, block_size: int,
                               wrapped: bool = False) -> Tuple[Tensor, Tensor]:
    """ref: utils.py:61-104 -> `_C.map_gaussian_to_intersects` (cuda/csrc/bindings.cu:77-121).
    Returns (isect_ids (I,) int64, gaussian_ids (I,) int32)."""
    if wrapped:
        raise NotImplementedError("wrapped=True is never used by the reference's callers (utils.py:70,115)")
    centers, extents, depths = centers.contiguous(), extents.contiguous(), depths.contiguous()
    cum = cum_tiles_hit.contiguous()
    dev = _lib.require_cuda(centers, extents, depths, cum)
    for t, nm in ((centers, "centers"), (extents, "extents"), (depths, "depths")):
        _lib.require_dtype(t, torch.float32, nm)
    _lib.require_dtype(cum, torch.int32, "cum_tiles_hit")
    isect_ids = torch.empty((num_intersects,), dtype=torch.int64, device=dev)
    gaussian_ids = torch.empty((num_intersects,), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.load().gstex_map_gaussian_to_intersects(
            int(num_points), int(num_intersects), _lib.ptr(centers), _lib.ptr(extents), _lib.ptr(depths),
            _lib.ptr(cum), int(tile_bounds[0]), int(tile_bounds[1]), int(block_size), _lib.ptr(isect_ids),
            _lib.ptr(gaussian_ids), _lib.stream_ptr(dev))
    _lib.check(rc, "gstex_map_gaussian_to_intersects")
    return isect_ids, gaussian_ids


def sort_isects(isect_ids: Tensor, gaussian_ids: Tensor, key_bits: int = 64) -> Tuple[Tensor, Tensor]:
    """Stable ascending radix sort of the keys with the ids as payload (= torch.sort + torch.gather of
    utils.py:159-160).  `key_bits`: how many low key bits are significant."""
    isect_ids, gaussian_ids = isect_ids.contiguous(), gaussian_ids.contiguous()
    dev = _lib.require_cuda(isect_ids, gaussian_ids)
    _lib.require_dtype(isect_ids, torch.int64, "isect_ids")
    _lib.require_dtype(gaussian_ids, torch.int32, "gaussian_ids")
    n = isect_ids.numel()
    keys_out = torch.empty_like(isect_ids)
    ids_out = torch.empty_like(gaussian_ids)
    lib = _lib.load()
    ws = _workspace(lib.gstex_sort_workspace_bytes(n), dev)
    with torch.cuda.device(dev):
        rc = lib.gstex_sort_isects(n, _lib.ptr(isect_ids), _lib.ptr(gaussian_ids), _lib.ptr(keys_out),
                                   _lib.ptr(ids_out), int(key_bits), _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev))
    _lib.check(rc, "gstex_sort_isects")
    return keys_out, ids_out


def bin_and_sort_gaussians(num_points: int, num_intersects: int, centers: Tensor, extents: Tensor, depths: Tensor,
                           cum_tiles_hit: Tensor, tile_bounds, block_size: int, wrapped: bool = False):
    """ref: utils.py:106-162.  Returns (isect_ids, gaussian_ids, isect_ids_sorted, gaussian_ids_sorted,
    tile_bins)."""
    isect_ids, gaussian_ids = map_gaussian_to_intersects(num_points, num_intersects, centers, extents, depths,
                                                         cum_tiles_hit, tile_bounds, block_size, wrapped=wrapped)
    isect_ids_sorted, gaussian_ids_sorted = sort_isects(isect_ids, gaussian_ids, key_bits_for(tile_bounds))
    tile_bins = get_tile_bin_edges(num_intersects, isect_ids_sorted, tile_bounds)
    return isect_ids, gaussian_ids, isect_ids_sorted, gaussian_ids_sorted, tile_bins


def bin_gaussians_to_tiles(num_points: int, centers: Tensor, extents: Tensor, depths: Tensor, num_tiles_hit: Tensor,
                           tile_bounds, block_size: int):
    """What `texture_gaussians` needs from `compute_cumulative_intersects` + `bin_and_sort_gaussians`
    (ref: utils.py:40-59,106-162): `(num_intersects, gaussian_ids_sorted (I,) int32, tile_bins (tiles,2) int32)`,
    bit-identical to them, through the tile-bucket path (csrc/tile_bucket.cu) -- per-tile counts, one scan, an
    atomic-cursor emit and a per-tile shared-memory sort by (depth bits, splat id) -- instead of a radix sort of
    the whole list.  One blocking read of two ints (`num_intersects`, longest tile list) replaces the
    reference's `.item()` (utils.py:58).  A tile list longer than the shared-memory capacity sends the call
    down the general path."""
    centers, extents, depths = centers.contiguous(), extents.contiguous(), depths.contiguous()
    nth = num_tiles_hit.contiguous().reshape(-1)
    dev = _lib.require_cuda(centers, extents, depths, nth)
    for t, nm in ((centers, "centers"), (extents, "extents"), (depths, "depths")):
        _lib.require_dtype(t, torch.float32, nm)
    num_tiles = int(tile_bounds[0]) * int(tile_bounds[1])
    i32 = dict(dtype=torch.int32, device=dev)
    if nth.numel() == 0:
        return 0, torch.zeros((0,), **i32), torch.zeros((num_tiles, 2), **i32)
    cum, total = cumsum_i32(nth)
    lib = _lib.load()

    def general(num_intersects):
        if num_intersects < 1:
            return num_intersects, torch.zeros((0,), **i32), torch.zeros((num_tiles, 2), **i32)
        (_, _, _, ids_sorted, bins) = bin_and_sort_gaussians(num_points, num_intersects, centers, extents, depths, cum,
                                                             tile_bounds, block_size)
        return num_intersects, ids_sorted, bins

    if num_tiles > lib.gstex_tile_bucket_max_tiles():
        return general(int(total.item()))
    block_counts = _workspace(lib.gstex_tile_bucket_scratch_bytes(num_tiles), dev)
    small = torch.empty((num_tiles + 2,), **i32)  # tile_start | stats
    tile_start, stats = small[:num_tiles], small[num_tiles:]
    tile_bins = torch.empty((num_tiles, 2), **i32)
    p, st = _lib.ptr, _lib.stream_ptr(dev)
    with torch.cuda.device(dev):
        rc = lib.gstex_tile_bucket_count(int(num_points), p(centers), p(extents), p(cum), int(tile_bounds[0]),
                                         int(tile_bounds[1]), int(block_size), p(block_counts), p(tile_bins),
                                         p(tile_start), p(stats), st)
    _lib.check(rc, "gstex_tile_bucket_count")
    num_intersects, max_list = stats.tolist()  # the one host sync of the binning stage
    if num_intersects < 1 or max_list > lib.gstex_tile_bucket_max_list():
        return general(num_intersects)
    keys = torch.empty((num_intersects,), dtype=torch.int64, device=dev)
    ids_sorted = torch.empty((num_intersects,), **i32)
    with torch.cuda.device(dev):
        rc = lib.gstex_tile_bucket_sort(int(num_points), int(num_intersects), int(max_list), p(centers), p(extents),
                                        p(depths), p(cum), int(tile_bounds[0]), int(tile_bounds[1]), int(block_size),
                                        p(block_counts), p(tile_start), p(tile_bins), p(keys), p(ids_sorted), st)
    _lib.check(rc, "gstex_tile_bucket_sort")
    return num_intersects, ids_sorted, tile_bins

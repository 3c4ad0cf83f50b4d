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


class TileBinning:
    """State of a binning request between `begin_binning` (asynchronous: per-tile counts, a pinned
    read-back of `num_intersects` and the longest tile list) and `finish_binning` (waits for that read-back only,
    then emits and sorts).  Splitting the two lets a multi-view step launch view i+1's `begin_binning` between
    view i's forward and backward: the host learns view i+1's sizes while the backward still runs, so the one
    host synchronisation of the stage (the reference's `.item()`, utils.py:58) no longer drains the GPU."""

    __slots__ = ("num_points", "centers", "extents", "depths", "nth", "cum", "total", "tile_bounds", "block_size", "num_tiles",
                 "block_counts", "tile_start", "tile_bins", "stats_host", "event", "fast", "result", "records")


_ISECT_QUANTUM = 1 << 20  # allocation granularity of the intersection lists (elements)
_PINNED_FREE = []  # pinned 2-int read-back buffers not owned by any outstanding request


def _pinned_stats() -> Tensor:
    """A pinned read-back buffer owned by ONE binning request until `finish_binning` has read it (then it returns to
    the free list): any number of requests may be outstanding without one's async copy landing in another's buffer."""
    try:
        return _PINNED_FREE.pop()
    except IndexError:
        return torch.empty(2, dtype=torch.int32).pin_memory()


def begin_binning(num_points: int, centers: Tensor, extents: Tensor, depths: Tensor, num_tiles_hit: Tensor,
                  tile_bounds, block_size: int) -> TileBinning:
    """Asynchronous first half of `bin_gaussians_to_tiles` (no host synchronisation)."""
    centers, extents, depths = centers.contiguous(), extents.contiguous(), depths.contiguous()
    nth = num_tiles_hit.contiguous().reshape(-1)
    dev = _lib.require_cuda(centers, extents, depths, nth)
    for t, nm in ((centers, "centers"), (extents, "extents"), (depths, "depths")):
        _lib.require_dtype(t, torch.float32, nm)
    if nth.dtype != torch.int32:  # the reference takes any integer dtype (torch.cumsum, utils.py:57)
        if nth.dtype.is_floating_point or nth.dtype == torch.bool:
            raise RuntimeError(f"gstex_b200: num_tiles_hit must be an integer tensor, got {nth.dtype}")
        nth = nth.to(torch.int32)
    if nth.numel() != int(num_points):
        raise RuntimeError(f"gstex_b200: num_tiles_hit has {nth.numel()} entries for {int(num_points)} points")
    b = TileBinning()
    b.num_points, b.centers, b.extents, b.depths = int(num_points), centers, extents, depths
    b.tile_bounds, b.block_size = tile_bounds, int(block_size)
    b.num_tiles = int(tile_bounds[0]) * int(tile_bounds[1])
    b.result = None
    b.records = None  # texture.PackedRecords made ahead of the forward (multiview.BinningStream.finish(pack=...))
    b.event = b.stats_host = b.block_counts = b.tile_start = b.tile_bins = None
    i32 = dict(dtype=torch.int32, device=dev)
    if nth.numel() == 0:
        b.result = (0, torch.zeros((0,), **i32), torch.zeros((b.num_tiles, 2), **i32))
        return b
    b.nth, b.cum, b.total = nth, None, None
    lib = _lib.load()
    b.fast = b.num_tiles <= lib.gstex_tile_bucket_max_tiles()
    if not b.fast:
        b.cum, b.total = cumsum_i32(nth)
        return b
    b.block_counts = _workspace(lib.gstex_tile_bucket_scratch_bytes(b.num_tiles), dev)
    small = torch.empty((b.num_tiles + 2,), **i32)  # tile_start | stats
    b.tile_start, stats = small[:b.num_tiles], small[b.num_tiles:]
    b.tile_bins = torch.empty((b.num_tiles, 2), **i32)
    p, st = _lib.ptr, _lib.stream_ptr(dev)
    with torch.cuda.device(dev):
        rc = lib.gstex_tile_bucket_count(b.num_points, p(centers), p(extents), p(nth), int(tile_bounds[0]),
                                         int(tile_bounds[1]), b.block_size, p(b.block_counts), p(b.tile_bins),
                                         p(b.tile_start), p(stats), st)
    _lib.check(rc, "gstex_tile_bucket_count")
    b.stats_host = _pinned_stats()
    b.stats_host.copy_(stats, non_blocking=True)
    b.event = torch.cuda.Event()
    b.event.record(torch.cuda.current_stream(dev))
    return b


def finish_binning(b: TileBinning):
    """Second half: `(num_intersects, gaussian_ids_sorted, tile_bins)`.  Blocks the host until the sizes recorded by
    `begin_binning` have arrived (not until the stream is idle); idempotent."""
    if b.result is not None:
        return b.result
    dev = b.centers.device
    i32 = dict(dtype=torch.int32, device=dev)
    lib = _lib.load()

    def general(num_intersects):
        if num_intersects < 1:
            return num_intersects, torch.zeros((0,), **i32), torch.zeros((b.num_tiles, 2), **i32)
        if b.cum is None:  # the tile-bucket path needs no prefix sum; the general one does
            b.cum, b.total = cumsum_i32(b.nth)
        (_, _, _, ids_sorted, bins) = bin_and_sort_gaussians(b.num_points, num_intersects, b.centers, b.extents, b.depths,
                                                             b.cum, b.tile_bounds, b.block_size)
        return num_intersects, ids_sorted, bins

    if not b.fast:
        b.result = general(int(b.total.item()))
        return b.result
    b.event.synchronize()  # the one host wait of the binning stage
    num_intersects, max_list = b.stats_host.tolist()
    if len(_PINNED_FREE) < 64:
        _PINNED_FREE.append(b.stats_host)
    b.stats_host = None
    if num_intersects < 1 or max_list > lib.gstex_tile_bucket_max_list():
        b.result = general(num_intersects)
        return b.result
    # Sized in steps of 1 M intersections: the views of a scene differ by a few per cent in `num_intersects`, and a
    # caching allocator that is asked for a slightly different size every view sooner or later finds none of its free
    # blocks large enough and goes to the driver in the middle of a step (a cudaMalloc of 12 MB was seen to take from
    # 1 to 87 ms there: scripts/gpu_clk.sh).  The list handed back is the first `num_intersects` elements.
    cap = (int(num_intersects) + _ISECT_QUANTUM - 1) // _ISECT_QUANTUM * _ISECT_QUANTUM
    keys = torch.empty((cap,), dtype=torch.int64, device=dev)
    ids_sorted = torch.empty((cap,), **i32)[:num_intersects]
    p = _lib.ptr
    with torch.cuda.device(dev):
        rc = lib.gstex_tile_bucket_sort(b.num_points, int(num_intersects), int(max_list), p(b.centers), p(b.extents),
                                        p(b.depths), p(b.nth), int(b.tile_bounds[0]), int(b.tile_bounds[1]), b.block_size,
                                        p(b.block_counts), p(b.tile_start), p(b.tile_bins), p(keys), p(ids_sorted),
                                        _lib.stream_ptr(dev))
    _lib.check(rc, "gstex_tile_bucket_sort")
    b.result = (num_intersects, ids_sorted, b.tile_bins)
    b.block_counts = b.tile_start = None
    return b.result


def bin_gaussians_to_tiles(num_points: int, centers: Tensor, extents: Tensor, depths: Tensor, num_tiles_hit: Tensor,
                           tile_bounds, block_size: int):
    """What `texture_gaussians` needs from `compute_cumulative_intersects` + `bin_and_sort_gaussians`
    (ref: utils.py:40-59,106-162): `(num_intersects, gaussian_ids_sorted (I,) int32, tile_bins (tiles,2) int32)`,
    bit-identical to them, through the tile-bucket path (csrc/tile_bucket.cu) -- per-tile counts, one scan, a
    shared-memory-cursor emit and a per-tile shared-memory sort by (depth bits, splat id) -- instead of a radix sort
    of the whole list.  One blocking read of two ints (`num_intersects`, longest tile list) replaces the
    reference's `.item()` (utils.py:58).  A tile list longer than the shared-memory capacity (or more tiles than
    the counting kernels hold) sends the call down the general path.  = `finish_binning(begin_binning(...))`."""
    return finish_binning(begin_binning(num_points, centers, extents, depths, num_tiles_hit, tile_bounds, block_size))

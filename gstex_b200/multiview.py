"""View-parallel sharding of a multi-view training step.

The hot path has no exchange step inside a view: views are independent given replicated splat / texture
parameters and their gradients simply add.  So a step over V views on n ranks is: every rank renders its
own views, accumulates the nine gradient tensors locally, and one all-reduce(sum) per tensor combines them
(NCCL over NVLink on the GPUs; the same code runs on gloo for the CPU tests).  The reference has no
multi-view or multi-GPU mode at all (one camera per call, `assert camera.shape[0] == 1`,
nerfstudio/models/gstex.py:1006).
"""
from __future__ import annotations

from typing import Iterable, List, Optional

import torch

from . import _lib


def shard_views(num_views: int, rank: int, world: int) -> List[int]:
    """Round-robin: rank r renders views r, r+world, ...  (equal counts when world divides num_views)."""
    if not (0 <= rank < world):
        raise ValueError("rank outside [0, world)")
    return list(range(rank, num_views, world))


def allreduce_gradients(grads: Iterable[Optional[torch.Tensor]], world: int, group=None) -> None:
    """Sum the gradient tensors over the ranks in place (no-op for a single rank).  `None` entries (a rank
    that rendered no view) are rejected: every rank must contribute a tensor of the same shape."""
    if world <= 1:
        return
    import torch.distributed as dist

    for g in grads:
        if g is None:
            raise ValueError("allreduce_gradients: a gradient is None on this rank")
        dist.all_reduce(g, op=dist.ReduceOp.SUM, group=group)


class HostViewFeeder:
    """Feeds per-view HOST inputs (camera matrices, upstream-gradient images, ... in pinned memory) to the
    device one view ahead of the view being rendered, and takes per-view results back to pinned host
    buffers behind the compute: host->device copies run on their own stream, device->host copies on a
    second one, both ordered against the compute stream with events only (no host synchronisation until
    `finish()`).  The reference has nothing comparable: nerfstudio moves one camera + one image per step with
    blocking `.to(device)` calls (nerfstudio/data/datamanagers/full_images_datamanager.py).

        feeder = HostViewFeeder(device)
        feeder.prefetch(host_inputs[0])
        for i in range(V):
            dev_inputs = feeder.take()                  # compute stream now waits for view i's copies
            if i + 1 < V: feeder.prefetch(host_inputs[i + 1])
            outs = render(dev_inputs)
            feeder.send_back(outs[0].detach(), host_out[i])
        feeder.finish()                                 # host may read host_out[*]
    """

    def __init__(self, device, slots: int = 2):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("HostViewFeeder needs a CUDA device (there is no CPU path)")
        self.h2d = torch.cuda.Stream(self.device)
        self.d2h = torch.cuda.Stream(self.device)
        self._pending = None
        # Device staging buffers, `slots` sets used round-robin and allocated once per (position, shape, dtype):
        # nothing is allocated in steady state (a cudaMalloc in the middle of a step synchronises the device).
        # The tensors `take()` returns for a view stay valid until `slots` further views have been prefetched.
        self._slots = [dict() for _ in range(max(2, int(slots)))]
        self._turn = 0

    @staticmethod
    def _map(obj, fn, path=()):
        if isinstance(obj, torch.Tensor):
            return fn(obj, path)
        if isinstance(obj, dict):
            return {k: HostViewFeeder._map(v, fn, path + (k,)) for k, v in obj.items()}
        if isinstance(obj, (list, tuple)):
            return type(obj)(HostViewFeeder._map(v, fn, path + (i,)) for i, v in enumerate(obj))
        return obj

    def prefetch(self, host_inputs):
        """Start the host->device copies of one view's inputs (any nest of dict/list/tuple of pinned tensors)."""
        if self._pending is not None:
            raise RuntimeError("HostViewFeeder.prefetch: the previous view has not been taken")
        slot = self._slots[self._turn % len(self._slots)]
        self._turn += 1
        # the slot's previous user (a view rendered `slots` views ago) was enqueued on the compute stream before now
        free = torch.cuda.Event()
        free.record(torch.cuda.current_stream(self.device))
        self.h2d.wait_event(free)

        def stage(t, path):
            buf = slot.get(path)
            if buf is None or buf.shape != t.shape or buf.dtype != t.dtype:
                buf = slot[path] = torch.empty(t.shape, dtype=t.dtype, device=self.device)
            buf.copy_(t, non_blocking=True)
            return buf

        with torch.cuda.stream(self.h2d):
            dev = self._map(host_inputs, stage)
            ev = torch.cuda.Event()
            ev.record(self.h2d)
        self._pending = (dev, ev)

    def take(self, also=()):
        """Device copies of the prefetched view; the current (compute) stream waits for them, and so does every
        stream in `also` (e.g. a `BinningStream`'s, which reads the camera matrix)."""
        if self._pending is None:
            raise RuntimeError("HostViewFeeder.take: nothing was prefetched")
        dev, ev = self._pending
        self._pending = None
        torch.cuda.current_stream(self.device).wait_event(ev)
        for st in also:
            st.wait_event(ev)
        return dev

    def send_back(self, dev_tensor, host_tensor):
        """Copy a result into a pinned host tensor behind the compute stream (asynchronous)."""
        cur = torch.cuda.current_stream(self.device)
        ev = torch.cuda.Event()
        ev.record(cur)
        self.d2h.wait_event(ev)
        dev_tensor.record_stream(self.d2h)
        with torch.cuda.stream(self.d2h):
            host_tensor.copy_(dev_tensor, non_blocking=True)

    def finish(self):
        """Block the host until every copy issued so far has landed."""
        self.h2d.synchronize()
        self.d2h.synchronize()
        torch.cuda.current_stream(self.device).synchronize()


class BinningStream:
    """Projection + tile binning of the NEXT view on a CUDA stream of their own, beside the compositing kernels of
    the current view.  The forward and the stream backward are persistent grids with a tail in which most SMs idle,
    and the binning kernels (project, count, emit, per-tile sort: ~0.2 ms per view at the headline size) are short
    and latency-bound; run back to back on one stream neither fills the GPU.  On two streams the binning of view
    i+1 runs in the shadow of view i's forward/backward (measured at the headline size: 1.48 -> 1.36 ms/view).
    The reference has one camera per call and one (legacy default) stream.

        bs = BinningStream(device)
        pre = bs.begin(means, scales, 1, quats, viewmats[0], intr, H, W, 16); bs.finish(pre)
        for i in range(V):
            outs = texture_gaussians(info, dims, *pre.args, colors, ..., binning=pre.binning)
            nxt = bs.begin(..., viewmats[i + 1], ..., wait_for_current=False) if i + 1 < V else None
            loss(outs).backward()
            if nxt is not None: bs.finish(nxt)   # host waits for view i+1's sizes only; its sort goes to the stream
            pre = nxt
    """

    class Prepared:
        __slots__ = ("centers", "extents", "depths", "num_tiles_hit", "binning", "ready", "view")

        @property
        def args(self):
            """The four positional arguments of `texture_gaussians` this stage produces."""
            return self.centers, self.extents, self.depths, self.num_tiles_hit

    def __init__(self, device, own_stream: bool = True):
        """`own_stream=False` keeps the same call sequence on the CURRENT stream (no overlap, fewer host-side stream
        and event operations per view): the A/B switch of `bench.py --binning-stream current`."""
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("BinningStream needs a CUDA device (there is no CPU path)")
        self.stream = torch.cuda.Stream(self.device) if own_stream else torch.cuda.current_stream(self.device)

    def begin(self, means, scales, glob_scale, quats, viewmat, intrins, img_height: int, img_width: int,
              block_width: int, *, wait_for_current: bool = True) -> "BinningStream.Prepared":
        """`project_bounds` + `begin_binning` of one view on the binning stream (no host synchronisation).
        `wait_for_current`: the binning stream first waits for everything enqueued so far on the current stream --
        needed whenever an input was produced there (parameters after an optimizer step).  Pass False for the later
        views of a step whose parameters have not changed since its first view, so that the binning is free to
        overlap the compositing kernels already enqueued."""
        from .get_aabb_2d import project_bounds
        from .utils import begin_binning

        if wait_for_current:
            self.stream.wait_stream(torch.cuda.current_stream(self.device))
        pre = BinningStream.Prepared()
        with torch.cuda.stream(self.stream):
            _, pre.depths, pre.centers, pre.extents, pre.num_tiles_hit = project_bounds(
                means, scales, glob_scale, quats, viewmat, intrins, img_height, img_width, block_width)
            tile_bounds = ((img_width + block_width - 1) // block_width, (img_height + block_width - 1) // block_width, 1)
            pre.binning = begin_binning(means.shape[0], pre.centers, pre.extents, pre.depths, pre.num_tiles_hit,
                                        tile_bounds, block_width)
        pre.ready = False
        pre.view = (means, scales, glob_scale, quats, viewmat, intrins, img_height, img_width, block_width)
        return pre

    def finish(self, pre: "BinningStream.Prepared", pack: Optional[dict] = None):
        """Second half (`finish_binning`: the host waits for the view's sizes, emit + per-tile sort go to the binning
        stream), then the CURRENT stream is made to wait for the result.  Returns `(num_intersects,
        gaussian_ids_sorted, tile_bins)`; `texture_gaussians(..., binning=pre.binning)` picks the same result up.

        `pack`: the remaining inputs of `texture_gaussians` the forward's per-view records are made from --
        `dict(texture_dims=, colors=, opacity=, uv0=, umap=, vmap=, c2w=, settings=)` -- to have them packed here, on the
        binning stream (`texture.pack_view_records`: an HBM pass over all splats, 65 us for 1 M, that otherwise
        sits on the current stream in front of the compositing kernel).  `texture_gaussians` uses the packed records
        when it is then called with exactly the tensors given here and to `begin`, unchanged; otherwise it packs again."""
        from .utils import finish_binning

        if pre.ready:
            return pre.binning.result
        cur = torch.cuda.current_stream(self.device)
        with torch.cuda.stream(self.stream):
            result = finish_binning(pre.binning)
            if pack is not None and result[0] > 0:
                from .texture import pack_view_records

                means, scales, glob_scale, quats, viewmat, intrins, H, W, bw = pre.view
                pre.binning.records = pack_view_records(
                    pack["texture_dims"], result[2], pack["colors"], pack["opacity"], means, scales, glob_scale, quats,
                    pack["uv0"], pack["umap"], pack["vmap"], viewmat, pack["c2w"], *intrins, H, W, bw, pack["settings"])
            ev = torch.cuda.Event()
            ev.record(self.stream)
        # allocated on the binning stream, consumed on the current one: keep the allocator from recycling them early
        extra = [pre.binning.records.workspace] if pre.binning.records is not None else []
        for t in [pre.centers, pre.extents, pre.depths, pre.num_tiles_hit, result[1], result[2]] + extra:
            if t.is_cuda and t.numel() > 0:
                t.record_stream(cur)
        cur.wait_event(ev)
        pre.ready = True
        return result


class GradientBucket:
    """All parameter gradients of a multi-view step in ONE flat buffer: `p.grad` of every parameter is a view into
    it, autograd accumulates the per-view gradients in place, and the step ends with a single all-reduce over
    the whole buffer (208 MB at the headline size) instead of one collective per tensor.  Slices start on
    128-byte boundaries.  Works on any device / backend (NCCL on the GPUs, gloo in the CPU tests).

    `peer_group` (a process group whose ranks share one NVLink / NVSwitch domain, e.g. `dist.group.WORLD` on one
    node): the buffer is allocated as torch symmetric memory, every rank maps every other rank's bucket, and
    `allreduce` becomes ONE kernel of ours over that peer memory (`gstex_peer_allreduce_f32`, csrc/peer_reduce.cu:
    each rank sums its 1/W slice over all ranks and stores it to all ranks) between two cross-rank barriers, instead of
    an NCCL all-reduce.  `self.peer` says which one is in use; when the mapping cannot be set up (no peer access, an
    old torch) the constructor raises unless `peer_fallback=True`, in which case it warns and uses NCCL."""

    def __init__(self, params, peer_group=None, peer_fallback: bool = False, multicast: Optional[bool] = None):
        self.params = list(params)
        if not self.params:
            raise ValueError("GradientBucket: no parameters")
        dev, dt = self.params[0].device, self.params[0].dtype
        offs, total = [], 0
        for p in self.params:
            if p.device != dev or p.dtype != dt:
                raise ValueError("GradientBucket: parameters must share device and dtype")
            offs.append(total)
            total += (p.numel() + 31) // 32 * 32
        total = max((total + 3) // 4 * 4, 4)
        self.peer = None  # _PeerMapping when the all-reduce runs over NVLink peer memory
        self.flat = None
        if peer_group is not None:
            try:
                self.peer = _PeerMapping(total, dt, dev, peer_group, multicast)
                self.flat = self.peer.local
            except Exception as exc:
                if not peer_fallback:
                    raise
                import warnings

                warnings.warn(f"GradientBucket: no peer-memory mapping ({exc!r}); the all-reduce goes through NCCL")
                self.peer = None
        if self.flat is None:
            self.flat = torch.zeros(total, dtype=dt, device=dev)
        self.views = [self.flat[o:o + p.numel()].view(p.shape) for o, p in zip(offs, self.params)]

    def zero_(self) -> None:
        """Zero every gradient and (re)attach the views as `p.grad`."""
        from .texture import GradientTargets

        GradientTargets.flush()  # sums a previous step left deferred must not leak into this one
        self.flat.zero_()
        for p, v in zip(self.params, self.views):
            if p.grad is not v:
                p.grad = v

    def register(self, defer: bool = True) -> None:
        """Let `texture_gaussians`' backward add each view's gradients straight into the bucket
        (`texture.GradientTargets`) instead of handing them to autograd's accumulation pass.  `defer`: the per-view
        sums stay in the backward's padded accumulators and reach the bucket once per step (`allreduce` / `flush`)
        instead of after every view."""
        from .texture import GradientTargets

        GradientTargets.set(self.params, self.views, defer=defer)

    def flush(self) -> None:
        """Make the bucket hold every gradient accumulated so far (no-op unless sums are deferred)."""
        from .texture import GradientTargets

        GradientTargets.flush()

    def allreduce(self, world: int, group=None) -> None:
        """Sum over the ranks (no-op for one rank).  A gradient produced outside the bucket (a `p.grad` that was
        replaced by a fresh tensor) is ADDED to the bucket's slice -- which may already hold the sums the fused
        accumulation (`GradientTargets`) wrote for other views -- before the slice becomes `p.grad` again."""
        self.flush()
        for p, v in zip(self.params, self.views):
            if p.grad is None:
                raise ValueError("GradientBucket.allreduce: a gradient is None on this rank")
            if p.grad is not v:
                v.add_(p.grad)
                p.grad = v
        if world <= 1:
            return
        if self.peer is not None:
            if group is not None and group is not self.peer.group:
                raise ValueError("GradientBucket.allreduce: the bucket is mapped over another process group")
            self.peer.allreduce()
            return
        import torch.distributed as dist

        dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)


class _PeerMapping:
    """The plumbing of the peer-memory all-reduce: a symmetric-memory allocation of `count` floats on every rank of
    `group`, the device addresses under which this process sees each rank's copy, and the cross-rank barriers.
    torch.distributed._symmetric_memory does the handle exchange (CUDA VMM handles over the group's store); the data
    path is `gstex_peer_allreduce_f32`."""

    # NVLink bytes per GPU and direction for a bucket of B bytes: 2 B (W-1)/W with peer loads/stores (the slice is read
    # from and written to W-1 peers, and the peers do the same to this GPU), B (1 + 1/W) with the switch-reduced variant
    # (all of this GPU's data is pulled once, every store comes back replicated).  Equal at W = 3, the multicast variant
    # wins from W = 4 on and loses at W = 2 (measured: profiles/peer_allreduce_r2.txt).
    MULTICAST_FROM_WORLD = 4

    def __init__(self, count: int, dtype, device, group, multicast: Optional[bool] = None):
        import ctypes

        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem

        if dtype != torch.float32:
            raise ValueError("peer-memory all-reduce: float32 buckets only")
        if device.type != "cuda":
            raise ValueError("peer-memory all-reduce needs CUDA devices")
        self.group = group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        if self.world > _lib.load().gstex_peer_allreduce_max_world():
            raise ValueError(f"peer-memory all-reduce: at most {_lib.load().gstex_peer_allreduce_max_world()} ranks")
        self.count = int(count)
        with torch.cuda.device(device):
            self.local = symm_mem.empty(self.count, dtype=dtype, device=device)
            self.handle = symm_mem.rendezvous(self.local, group)
        self.local.zero_()
        ptrs = [int(p) for p in self.handle.buffer_ptrs]
        if len(ptrs) != self.world or ptrs[self.rank] != self.local.data_ptr():
            raise RuntimeError("peer-memory all-reduce: unexpected symmetric-memory mapping")
        self.ptrs = (ctypes.c_void_p * self.world)(*ptrs)
        self.multicast_ptr = None
        mc = int(getattr(self.handle, "multicast_ptr", 0) or 0)
        if multicast is None:  # automatic: the NVSwitch-reduced variant where it exists and pays
            multicast = mc != 0 and self.world >= self.MULTICAST_FROM_WORLD
        if multicast:
            if mc == 0:
                raise RuntimeError("peer-memory all-reduce: no NVSwitch multicast mapping on this system")
            self.multicast_ptr = ctypes.c_void_p(mc)
        self.device = device
        self.max_ctas = 0
        self.handle.barrier(channel=0)  # every rank's zeros are in place before anyone's first reduce

    def allreduce(self) -> None:
        dev = self.device
        self.handle.barrier(channel=0)  # every rank's bucket is complete (its stream reached this point)
        with torch.cuda.device(dev):
            rc = _lib.load().gstex_peer_allreduce_f32(self.rank, self.world, self.ptrs, self.count, self.multicast_ptr,
                                                      int(self.max_ctas), _lib.stream_ptr(dev))
        _lib.check(rc, "gstex_peer_allreduce_f32")
        self.handle.barrier(channel=1)  # every rank's stores have landed before anyone reads or zeroes its bucket

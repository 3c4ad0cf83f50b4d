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

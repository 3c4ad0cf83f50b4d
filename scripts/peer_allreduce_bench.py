#!/usr/bin/env python
"""Exchange step alone (GPU box, under torchrun): the flat gradient bucket of the headline config (52 M floats, 208 MB)
summed over the ranks by gstex_peer_allreduce_f32 (P2P and NVSwitch-multimem variants) and by NCCL's all-reduce.
CUDA events on every rank around {barrier, kernel, barrier} (what a step pays) and around the kernel alone; max over
ranks, median of 30.   torchrun --nproc-per-node N scripts/peer_allreduce_bench.py [count]"""
import os, statistics, sys
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gstex_b200 import _lib
from gstex_b200.multiview import GradientBucket

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
count = int(sys.argv[1]) if len(sys.argv) > 1 else 52_000_000
lib = _lib.load()

def maxr(x):
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())

def run(name, fn, kernel=None):
    whole, kern = [], []
    for it in range(35):
        flat.normal_()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        _lib.Timing.start()
        e[0].record(); fn(); e[1].record()
        rec = _lib.Timing.stop()
        if it >= 5:
            whole.append(maxr(e[0].elapsed_time(e[1])))
            if kernel: kern.append(maxr(rec[kernel][0]))
    w = statistics.median(whole)
    nv = count * 4 * (world - 1) / world
    if rank == 0:
        k = statistics.median(kern) if kern else float("nan")
        print(f"{name:10s} W={world} {count * 4 / 1e6:7.1f} MB  step-visible {w:.3f} ms  kernel {k:.3f} ms  "
              f"NVLink in (= out) per GPU {nv / 1e6:.0f} MB -> {nv / (w * 1e-3) / 1e9:.0f} GB/s visible, {nv / (k * 1e-3) / 1e9 if kern else float('nan'):.0f} GB/s kernel", flush=True)

p = [torch.zeros(count, device=dev, requires_grad=True)]
for name, mc in (("p2p", False), ("multimem", True)):
    try:
        b = GradientBucket(p, peer_group=dist.group.WORLD, multicast=mc)
    except Exception as exc:
        if rank == 0: print(name, "unavailable:", exc)
        continue
    flat = b.flat
    sweep = os.environ.get("PEER_SWEEP")
    if mc and sweep:
        for ctas in (8, 16, 32, 64, 148, 296):
            b.peer.max_ctas = ctas
            run(f"mm/{ctas}", b.peer.allreduce, "gstex_peer_allreduce_f32")
    for ctas in ((0,) if mc else (0, 74, 37)):
        b.peer.max_ctas = ctas
        run(f"{name}/{ctas or 'default'}", b.peer.allreduce, "gstex_peer_allreduce_f32")
    del b
flat = torch.zeros(count, device=dev)
run("nccl", lambda: dist.all_reduce(flat))
dist.barrier(); dist.destroy_process_group()

#!/usr/bin/env python
"""Host-side cost of one multi-view step: enqueue time vs GPU time, and a cProfile of the enqueue path (GPU box).
Measured r2 (cfg5, 8 views): host enqueue 8.3 ms of a 10.1 ms step -- the host keeps up, but only just."""
import cProfile, os, pstats, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from gstex_b200.multiview import shard_views
from gstex_b200.synthetic import make_scene, make_upstream_grads

cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg5"
V = 8
s = make_scene(cfg, views=V); v = make_upstream_grads(s)
wl = bench.Workload(s, v, torch.device("cuda", 0), shard_views(V, 0, 1))
for _ in range(3): wl.step(1)
torch.cuda.synchronize()
enq, tot = [], []
for _ in range(10):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    wl.step(1)
    t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    enq.append((t1 - t0) * 1e3); tot.append((t2 - t0) * 1e3)
print(f"{cfg}: host enqueue {sorted(enq)[5]:.2f} ms/step, enqueue+drain {sorted(tot)[5]:.2f} ms/step ({V} views)")
pr = cProfile.Profile(); pr.enable()
for _ in range(5): wl.step(1)
pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)

"""Forward-kernel SIMT statistics (library built with -DGSTEX_STATS).  python scripts/fwd_stats.py [cfg]"""
import ctypes, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from gstex_b200 import _lib
from gstex_b200.multiview import shard_views
class A: pass
a = A(); a.config = sys.argv[1] if len(sys.argv) > 1 else "cfg5"; a.n = None; a.size = None
s, v = bench.build_scene(a, 2)
wl = bench.Workload(s, v, torch.device("cuda", 0), [0, 1])
lib = _lib.load()
cd = lib._cdll
out = (ctypes.c_ulonglong * 16)()
wl.step(1)
cd.gstex_debug_fwd_stats(out, 1)
wl.step(1)
cd.gstex_debug_fwd_stats(out, 1)
names = ["eval_iters", "sum_live", "sum_pot", "sum_pot_live", "iters_no_potlive", "iters_no_inbox_live", "iters_no_pot",
         "sum_inbox_live", "sum_inbox", "units", "iters_nocull"]
d = {n: out[i] / 2 for i, n in enumerate(names)}
for k, x in d.items(): print(f"{k:22s} {x:14.0f}  per-iter {x / d['eval_iters']:.3f}")

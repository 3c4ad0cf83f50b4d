#!/usr/bin/env python
"""Forward-only timing of one view (kernel iteration aid, GPU box):  python scripts/fwd_time.py cfg5 [variant ...]
Times gstex_texture_forward (pack + compositing kernel, pair stream recorded) with CUDA events, median of 20."""
import os, statistics, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gstex_b200 as G
from gstex_b200 import _lib, texture as T
from gstex_b200.synthetic import make_scene

name = sys.argv[1] if len(sys.argv) > 1 else "cfg5"
variants = [int(x) for x in sys.argv[2:]] or [0, 2]   # 0 = default kernel, 2 = lane-dense kernel
s = make_scene(name, views=8 if name == "cfg4" else 1)
sd = {k: (x.cuda() if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
intr = (s["fx"], s["fy"], s["cx"], s["cy"]); H, W, bw = s["H"], s["W"], 16
_, depths, centers, extents, nth = G.project_bounds(sd["means"], sd["scales"], 1, sd["quats"], sd["viewmat"], intr, H, W, bw)
tb = ((W + bw - 1) // bw, (H + bw - 1) // bw, 1)
I, ids, bins = G.bin_gaussians_to_tiles(s["N"], centers, extents, depths, nth, tb, bw)
lib = _lib.load()
padded = G.pad_texture(sd["texture"]) if s["texture_info"][2] <= 3 else None
def fwd(rec, pad=False):
    return T._raster_forward(s["texture_info"], sd["texture_dims"], ids, bins, sd["colors"], sd["opacities"], sd["means"], sd["scales"], 1,
                             sd["quats"], sd["uv0"], sd["umap"], sd["vmap"], sd["texture"], sd["viewmat"], sd["c2w"], *intr, H, W, bw,
                             s["settings"], sd["background"], I, record_pairs=rec, texture_padded=padded if pad else None)
ts = []
for it in range(25):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); G.pad_texture(sd["texture"], out=padded); e1.record(); torch.cuda.synchronize()
    if it >= 5: ts.append(e0.elapsed_time(e1))
print(f"{name}: pad_texture of {sd['texture'].shape[0]} texels median {statistics.median(ts):.3f} ms", flush=True)
for pad in (False, True, False, True):
    ts = []
    for it in range(45):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); out = fwd(True, pad); e1.record(); torch.cuda.synchronize()
        if it >= 5: ts.append(e0.elapsed_time(e1))
    print(f"{name} texel loads {'128-bit' if pad else 'scalar '}: forward entry median {statistics.median(ts):.4f} ms  min {min(ts):.4f}", flush=True)
for v in variants:
    lib.gstex_debug_forward_variant(v)
    for rec in (True, False):
        ts = []
        for it in range(25):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); out = fwd(rec); e1.record(); torch.cuda.synchronize()
            if it >= 5: ts.append(e0.elapsed_time(e1))
        print(f"{name} variant {v} record {int(rec)}: forward entry median {statistics.median(ts):.3f} ms  min {min(ts):.3f}", flush=True)
    if v == 4:
        import ctypes
        buf = (ctypes.c_ulonglong * 8)()
        lib._cdll.gstex_debug_dense_stats(buf, 1)
        nf = 50.0
        print("resume stats per forward: units %.0f, entries left %.0f (max %d), entries walked %.0f (max %d)" %
              (buf[0] / nf, buf[2] / nf, buf[3], buf[4] / nf, buf[5]))
lib.gstex_debug_forward_variant(0)

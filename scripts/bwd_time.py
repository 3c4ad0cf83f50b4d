#!/usr/bin/env python
"""Backward-only timing of one view (kernel iteration aid, GPU box):  python scripts/bwd_time.py cfg5
Runs the forward once, then the backward entry point 25 times on the same pair stream; CUDA events, median of 20."""
import os, statistics, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gstex_b200 as G
from gstex_b200 import _lib
from gstex_b200.synthetic import make_scene, make_upstream_grads

name = sys.argv[1] if len(sys.argv) > 1 else "cfg5"
s = make_scene(name, views=8 if name == "cfg4" else 1)
v = {k: x.cuda() for k, x in make_upstream_grads(s).items()}
sd = {k: (x.cuda() if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
intr = (s["fx"], s["fy"], s["cx"], s["cy"]); H, W, bw = s["H"], s["W"], 16
keys = ["colors", "opacities", "means", "scales", "quats", "uv0", "umap", "vmap", "texture"]
p = {k: sd[k].clone().requires_grad_(True) for k in keys}
_, depths, centers, extents, nth = G.project_bounds(sd["means"], sd["scales"], 1, sd["quats"], sd["viewmat"], intr, H, W, bw)
for rep in range(2):   # the second forward has a pool sized from the first
    outs = G.texture_gaussians(s["texture_info"], sd["texture_dims"], centers, extents, depths, nth, p["colors"], p["opacities"],
                               p["means"], p["scales"], 1, p["quats"], p["uv0"], p["umap"], p["vmap"], p["texture"], sd["viewmat"],
                               sd["c2w"], *intr, H, W, bw, s["settings"], sd["background"])
    torch.cuda.synchronize()
vs = [v["v_out_img"], v["v_out_depth"], v["v_out_reg"], v["v_out_alpha"], v["v_out_texture"], v["v_out_normal"]]
_lib.load()
_lib.Timing.start()
for it in range(25):
    torch.autograd.backward(list(outs), vs, retain_graph=True)
t = _lib.Timing.stop()
for k, x in t.items():
    print(f"{name} {k}: median {statistics.median(x[5:]):.3f} ms  min {min(x[5:]):.3f}")

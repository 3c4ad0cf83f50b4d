#!/usr/bin/env python
"""How close to bit-identical is the forward to the reference's kernels?  (GPU box, test tooling.)
    python scripts/parity_probe.py cfg3
Prints the share of bit-identical elements per output and details of the pixels whose index outputs differ."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gstex_b200.synthetic import make_scene  # noqa: E402
from oracle import parity, ref_ext  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
s = make_scene(name, views=8 if name == "cfg4" else 1)
dev = torch.device("cuda")
sd = {k: (x.to(dev) if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
C = ref_ext.load()
ours = parity.run_ours_view(sd, None)
ref = parity.run_reference_view(C, sd, None)
ours["final_Ts"] = 1 - ours["out_alpha"]
for k in ["out_img", "out_depth", "out_reg", "out_alpha", "out_texture", "out_normal"]:
    a, b = ours[k].contiguous().view(torch.int32), ref[k].contiguous().view(torch.int32)
    d = (a - b).abs()
    print(f"{k:12s} bit-identical {float((a == b).float().mean()):.6f}  <=1ulp {float((d <= 1).float().mean()):.6f} "
          f"<=4ulp {float((d <= 4).float().mean()):.6f}")
for k in ["final_idx", "depth_idx"]:
    bad = (ours[k] != ref[k]).nonzero()
    print(k, "mismatches:", bad.shape[0])
    for y, x in bad.tolist()[:8]:
        t = (y // 16) * ((s["W"] + 15) // 16) + x // 16
        r = ref["tile_bins"][t].tolist()
        print(f"  pixel ({y},{x}) tile {t} range {r} ours {int(ours[k][y, x])} ref {int(ref[k][y, x])} "
              f"alpha ours {float(ours['out_alpha'][y, x]):.9g} ref {float(ref['out_alpha'][y, x]):.9g} "
              f"T ref {float(ref['final_Ts'][y, x]):.6g} depth ours {float(ours['out_depth'][y, x]):.9g} ref {float(ref['out_depth'][y, x]):.9g}")

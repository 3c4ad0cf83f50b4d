"""Small forward through the dense kernel (debugging aid for compute-sanitizer)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gstex_b200.synthetic import make_scene
from oracle import parity
name, n, size = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
s = make_scene(name, n=n, width=size, height=size)
sd = {k: (x.cuda() if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
from gstex_b200 import _lib
_lib.load().gstex_debug_forward_variant(int(sys.argv[4]) if len(sys.argv) > 4 else 0)
r = parity.run_ours_view(sd, None)
torch.cuda.synchronize()
print("ok", float(r["out_img"].sum()))

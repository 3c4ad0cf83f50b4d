import torch, sys
sys.path.insert(0, '/root/repo')
import gstex_b200 as G
from gstex_b200.synthetic import make_scene
s = make_scene("cfg5", views=1)
sd = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in s.items()}
intr = (s["fx"], s["fy"], s["cx"], s["cy"])
_, depths, c, e, nth = G.project_bounds(sd["means"], sd["scales"], 1, sd["quats"], sd["viewmats"][0].cuda(), intr, 800, 800, 16)
I, ids, bins = G.bin_gaussians_to_tiles(s["N"], c, e, depths, nth, (50, 50, 1), 16)
cnt = (bins[:, 1] - bins[:, 0]).cpu()
print("I", I, "max", int(cnt.max()), "nonempty", int((cnt > 0).sum()))
for lo, hi in ((1, 512), (513, 1024), (1025, 2048), (2049, 3072), (3073, 4096), (4097, 6144), (6145, 8192), (8193, 20000)):
    m = (cnt >= lo) & (cnt <= hi)
    print(lo, hi, int(m.sum()), int(cnt[m].sum()))

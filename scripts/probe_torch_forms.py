"""Which fp32 forms do torch's CUDA ops use?  (GPU box, test tooling: decides the forms csrc/activations.cu pins.)
Result on torch 2.11 / B200: norm(dim=-1) over 4 floats = sqrt((r(q0^2) + r(q2^2)) + (r(q1^2) + r(q3^2))), r = round to
fp32; tensor / tensor is a true division; F.normalize = x / norm.clamp_min(1e-12); sigmoid = 1 / (1 + exp(-x));
mean over 2 = (a + b) * 0.5 -- each 0 mismatches over 1e6 samples."""
import itertools

import torch

torch.manual_seed(0)
q = torch.randn(1_000_000, 4, device="cuda")
n = q.norm(dim=-1)
sq = q.double() * q.double()
r = lambda t: t.float().double()  # noqa: E731
for p in itertools.permutations(range(4)):
    a, b, c, e = (sq[:, i] for i in p)
    forms = {"seq-rounded": r(r(r(r(a) + r(b)) + r(c)) + r(e)), "fma-chain": r(e + r(c + r(b + r(a)))),
             "pair-rounded": r(r(r(a) + r(b)) + r(r(c) + r(e)))}
    for nm, s in forms.items():
        mm = int((torch.sqrt(s.float()) != n).sum())
        if mm == 0:
            print(f"norm == {nm} {p}")
qn = q / n[:, None]
print("q / n is a true division:", int((qn != (q.double() / n.double()[:, None]).float()).sum()) == 0,
      "  == q * (1 / n):", int((qn != q * (1.0 / n[:, None])).sum()) == 0)
x = torch.randn(1_000_000, device="cuda") * 3
print("sigmoid == 1 / (1 + exp(-x)):", int((torch.sigmoid(x) != 1.0 / (1.0 + torch.exp(-x))).sum()) == 0)
sc = torch.rand(1_000_000, 2, device="cuda")
print("mean over 2 == (a + b) * 0.5:", int((sc.mean(-1) != (sc[:, 0] + sc[:, 1]) * 0.5).sum()) == 0)
fn = torch.nn.functional.normalize(qn, dim=-1)
print("F.normalize == x / norm.clamp_min(1e-12):", int((fn != qn / qn.norm(dim=-1, keepdim=True).clamp_min(1e-12)).sum()) == 0)

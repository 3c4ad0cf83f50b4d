"""Golden vectors for the per-splat activation stage, from the REFERENCE's own source text (this container only).

    python tests/golden/make_golden_act.py

`nerfstudio/models/gstex.py` cannot be imported here (nerfstudio's dependencies are absent), so the script reads the
file where it lies, cuts out the statements of `get_outputs` that activate the raw parameters (gstex.py:1062-1066) and
the whole `get_uv_mapping` method (:975-990), and executes that text unchanged on CPU tensors with a stub `self`
(`quat_to_rotmat` is imported from the reference's `gstex_cuda/_torch_impl.py`).  Recorded: the six activated tensors
and autograd's gradients of the raw parameters for seeded upstream gradients.
"""
import os
import sys
import textwrap
import types

import numpy as np
import torch

sys.path.insert(0, "/root/reference/submodules/GStex_cuda")
from gstex_cuda._torch_impl import quat_to_rotmat  # noqa: E402

SRC = "/root/reference/nerfstudio/models/gstex.py"
HERE = os.path.dirname(os.path.abspath(__file__))


def reference_text():
    lines = open(SRC).read().split("\n")
    act = lines[1061:1066]                                   # gstex.py:1062-1066 (1-based)
    assert act[0].strip().startswith("quats = self.quats / self.quats.norm") and act[-1].strip().startswith("opacities = torch.sigmoid"), act
    start = next(i for i, l in enumerate(lines) if l.strip().startswith("def get_uv_mapping("))
    end = next(i for i in range(start + 1, len(lines)) if lines[i].strip().startswith("def "))
    return textwrap.dedent("\n".join(act)), textwrap.dedent("\n".join(lines[start:end]))


def run_reference(quats_raw, scales_raw, opac_raw, means, mappings):
    act_src, uv_src = reference_text()
    ns = {"torch": torch, "quat_to_rotmat": quat_to_rotmat}
    exec(uv_src, ns)
    self = types.SimpleNamespace(quats=quats_raw, scales=scales_raw, opacities=opac_raw, mappings=mappings)
    env = dict(ns, self=self, means=means)
    exec(act_src, env)
    uv0, umap, vmap = ns["get_uv_mapping"](self, means, env["quats"], mappings)
    return env["quats"], env["scales"], env["opacities"], uv0, umap, vmap


def main():
    g = torch.Generator().manual_seed(2024)
    n = 257
    quats_raw = (torch.randn(n, 4, generator=g) * torch.exp(torch.randn(n, 1, generator=g))).requires_grad_(True)
    scales_raw = torch.randn(n, 3, generator=g) * 1.5 - 4.0
    scales_raw[:3, 0] = -30.0                                # exp underflows the 1e-9 clamp: the gradient must be cut
    scales_raw.requires_grad_(True)
    opac_raw = (torch.randn(n, 1, generator=g) * 3).requires_grad_(True)
    means = torch.randn(n, 3, generator=g)
    mappings = torch.rand(n, 2, generator=g) * 40 + 0.5
    quats, scales, opac, uv0, umap, vmap = run_reference(quats_raw, scales_raw, opac_raw, means, mappings)
    v = [torch.randn(t.shape, generator=g) for t in (quats, scales, opac)]
    torch.autograd.backward([quats, scales, opac], v)
    out = dict(quats_raw=quats_raw, scales_raw=scales_raw, opacities_raw=opac_raw, mappings=mappings, quats=quats,
               scales=scales, opacities=opac, uv0=uv0, umap=umap, vmap=vmap, v_quats=v[0], v_scales=v[1],
               v_opacities=v[2], g_quats_raw=quats_raw.grad, g_scales_raw=scales_raw.grad, g_opacities_raw=opac_raw.grad)
    np.savez_compressed(os.path.join(HERE, "act_golden.npz"), **{k: t.detach().numpy() for k, t in out.items()})
    print("wrote act_golden.npz", {k: tuple(t.shape) for k, t in out.items()})


if __name__ == "__main__":
    main()

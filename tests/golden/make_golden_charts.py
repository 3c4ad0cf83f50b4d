"""Golden vectors for re-charting, from the REFERENCE's own code (this container only).

    python tests/golden/make_golden_charts.py

* `build_charts`: `nerfstudio/models/gstex.py` cannot be imported here (nerfstudio's dependencies are absent), so the
  method's source text (gstex.py:841-888) is cut out of the file and executed unchanged on CPU tensors with a stub `self`.
* `JaggedTexture.init_from_dims`: `nerfstudio/models/jagged_texture.py` only needs torch and the submodule, so the real
  class is loaded from its file and called (its CPU branch samples with `_torch_impl.sample_texture`).
"""
import importlib.util
import os
import sys
import textwrap
import types

import numpy as np
import torch

sys.path.insert(0, "/root/reference/submodules/GStex_cuda")
HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/nerfstudio/models/gstex.py"


def reference_build_charts():
    lines = open(SRC).read().split("\n")
    start = next(i for i, l in enumerate(lines) if l.strip().startswith("def build_charts("))
    indent = lines[start][: len(lines[start]) - len(lines[start].lstrip())]
    end = next(i for i in range(start + 1, len(lines)) if lines[i].startswith(indent + "def "))   # next METHOD (a nested def is inside)
    assert any("self.mappings[:,1]" in l for l in lines[start:end])
    ns = {"torch": torch, "np": np}
    exec(textwrap.dedent("\n".join(lines[start:end])), ns)
    return ns["build_charts"]


def run_build_charts(scales, pixel_num, sigma_factor=3.0):
    n = scales.shape[0]
    seen = {}
    self = types.SimpleNamespace(
        config=types.SimpleNamespace(pixel_num=pixel_num, sigma_factor=sigma_factor), scales=scales, num_points=n,
        pixel_scale=torch.zeros(1), texture_dims=torch.zeros(n, 3, dtype=torch.int32), mappings=torch.zeros(n, 2),
        texture_dc=types.SimpleNamespace(init_from_dims=lambda d: seen.setdefault("dims", d.clone())),
        update_edit_texture=lambda: None, edit_texture=None)
    reference_build_charts()(self)
    l0, l1 = torch.exp(scales[:, 0]), torch.exp(scales[:, 1])
    area_sum = torch.sum(sigma_factor * sigma_factor * l0 * l1 * (torch.ones_like(l0) ** 2)).item()
    return dict(scales=scales, texture_dims=self.texture_dims, mappings=self.mappings, pixel_scale=self.pixel_scale.double(),
                area_sum=torch.tensor(area_sum, dtype=torch.float64), pixel_num=torch.tensor(float(pixel_num), dtype=torch.float64))


def run_init_from_dims(g, n, channels):
    spec = importlib.util.spec_from_file_location("jagged_texture", "/root/reference/nerfstudio/models/jagged_texture.py")
    jt_mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(jt_mod)

    def dims(lo, hi):
        d = torch.stack([torch.randint(lo, hi, (n,), generator=g, dtype=torch.int32),
                         torch.randint(lo, hi, (n,), generator=g, dtype=torch.int32), torch.zeros(n, dtype=torch.int32)], -1)
        hw = d[:, 0] * d[:, 1]
        d[:, 2] = torch.cumsum(hw, 0) - hw
        return d

    old, new = dims(1, 7), dims(1, 9)
    jt = jt_mod.JaggedTexture(old.clone(), out_dim=channels)
    jt.texture.data = torch.rand(jt.texture.shape, generator=g)
    old_tex = jt.get_texture().detach().clone()
    jt.init_from_dims(new.clone())
    return dict(old_dims=old, new_dims=new, old_texture=old_tex, new_texture=jt.get_texture().detach().clone())


def main():
    g = torch.Generator().manual_seed(4711)
    out = {}
    for name, n, pixel_num, spread in (("small", 500, 5000.0, 0.5), ("wide", 20000, 300000.0, 1.2), ("tiny", 40, 40.0, 0.3)):
        scales = torch.randn(n, 3, generator=g) * spread + np.log(0.006)
        for k, v in run_build_charts(scales, pixel_num).items():
            out[f"charts_{name}_{k}"] = v.detach().numpy()
    for name, n, c in (("rgb", 300, 3), ("six", 64, 6)):
        for k, v in run_init_from_dims(g, n, c).items():
            out[f"resample_{name}_{k}"] = v.numpy()
    np.savez_compressed(os.path.join(HERE, "charts_golden.npz"), **out)
    print({k: (v.shape if v.ndim else float(v)) for k, v in out.items() if "pixel_scale" in k or "new_texture" in k})


if __name__ == "__main__":
    main()

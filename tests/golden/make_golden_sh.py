"""Golden vectors for the spherical-harmonics colour stage, from the REFERENCE's own Python (this container only).

    python tests/golden/make_golden_sh.py

Imports `/root/reference/submodules/GStex_cuda/gstex_cuda/_torch_impl.py` unchanged and records, for seeded
inputs of every degree 0..4, `compute_sh_color(viewdirs, coeffs)` (_torch_impl.py:12-22, bases :62-135) and
the `torch.autograd` gradient of `(colors * w).sum()` with respect to the coefficients -- the same pair of
quantities the reference's tests/test_sh.py compares its CUDA kernels with.  The torch implementation wants unit
directions and always uses all bases of the coefficient tensor (degrees_to_use == degree).
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference/submodules/GStex_cuda")
import gstex_cuda._torch_impl as T  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    out = {}
    g = torch.Generator().manual_seed(2024)
    for degree, nb in enumerate((1, 4, 9, 16, 25)):
        n = 257
        v = torch.randn(n, 3, generator=g)
        v = v / v.norm(dim=-1, keepdim=True)
        c = (torch.rand(n, nb, 3, generator=g) - 0.5).requires_grad_(True)
        w = torch.randn(n, 3, generator=g)
        col = T.compute_sh_color(v, c)
        (col * w).sum().backward()
        out[f"viewdirs_{degree}"] = v.numpy()
        out[f"coeffs_{degree}"] = c.detach().numpy()
        out[f"w_{degree}"] = w.numpy()
        out[f"colors_{degree}"] = col.detach().numpy()
        out[f"v_coeffs_{degree}"] = c.grad.numpy()
    np.savez_compressed(os.path.join(HERE, "sh_golden.npz"), **out)
    print("wrote sh_golden.npz", {k: v.shape for k, v in out.items() if k.startswith("colors")})


if __name__ == "__main__":
    main()

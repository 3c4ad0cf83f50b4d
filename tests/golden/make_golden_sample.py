"""Golden vectors for `texture_sample`, from the REFERENCE's own Python (this container only).

    python tests/golden/make_golden_sample.py

Imports `/root/reference/submodules/GStex_cuda/gstex_cuda/_torch_impl.py` unchanged and records
`sample_texture(texture_dims, texture, uvs)` (_torch_impl.py:149-194) for seeded jagged textures built the way the
reference's tests/test_sample.py builds them (random chart sizes, 10 channels), plus a 3-channel case with 1x1
charts and uv values on, at and beyond the [0,1] borders (the torch implementation clamps them).
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference/submodules/GStex_cuda")
import gstex_cuda._torch_impl as T  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def case(g, charts, channels, queries, hmax, wmax, border):
    dims = torch.stack([torch.randint(hmax, (charts,), generator=g, dtype=torch.int32) + (1 if border else 2),
                        torch.randint(wmax, (charts,), generator=g, dtype=torch.int32) + (1 if border else 2),
                        torch.zeros(charts, dtype=torch.int32)], -1)
    hw = dims[:, 0] * dims[:, 1]
    dims[:, 2] = torch.cumsum(hw, 0) - hw
    tex = torch.rand(int(hw.sum()), channels, generator=g)
    uvs = torch.rand(queries, 2, generator=g)
    if border:
        uvs = uvs * 1.4 - 0.2
        uvs[:8] = torch.tensor([[0., 0.], [1., 1.], [0., 1.], [1., 0.], [0.5, 1.], [1., 0.5], [-0.1, 0.3], [0.3, 1.2]])
    ids = torch.randint(charts, (queries,), generator=g)
    qd = dims[ids]
    return dict(texture_dims=qd.numpy(), texture=tex.numpy(), uvs=uvs.numpy(), out=T.sample_texture(qd, tex, uvs).numpy())


def main():
    g = torch.Generator().manual_seed(77)
    out = {}
    for name, c in (("ref_test", case(g, 5, 10, 100, 6, 7, False)), ("border", case(g, 40, 3, 600, 5, 5, True)),
                    ("wide", case(g, 300, 6, 5000, 30, 30, False))):
        for k, v in c.items():
            out[f"{name}_{k}"] = v
    np.savez_compressed(os.path.join(HERE, "sample_golden.npz"), **out)
    print("wrote sample_golden.npz", {k: v.shape for k, v in out.items() if k.endswith("_out")})


if __name__ == "__main__":
    main()

"""GPU parity of `texture_sample` (gstex_b200/texture_sample.py, csrc/texture_sample.cu): against the CPU oracle, the
golden vectors made by the reference's Python, the reference's CUDA extension when oracle/_ref holds it, and -- at
the headline texture size -- size-independent properties.  Mirrors the reference's tests/test_sample.py."""
import os

import numpy as np
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "sample_golden.npz")


def _jagged(g, charts, channels, hmax=6, wmax=7):
    dims = torch.stack([torch.randint(hmax, (charts,), generator=g, dtype=torch.int32) + 1,
                        torch.randint(wmax, (charts,), generator=g, dtype=torch.int32) + 1,
                        torch.zeros(charts, dtype=torch.int32)], -1)
    hw = dims[:, 0] * dims[:, 1]
    dims[:, 2] = torch.cumsum(hw, 0) - hw
    return dims, torch.rand(int(hw.sum()), channels, generator=g)


@pytest.mark.parametrize("name", ["ref_test", "border", "wide"])
def test_sample_matches_golden(name):
    import gstex_b200 as G

    d = np.load(GOLD)
    tex = torch.from_numpy(d[f"{name}_texture"]).cuda()
    out = G.texture_sample((0, 1, tex.shape[1]), torch.from_numpy(d[f"{name}_texture_dims"]).cuda(), tex,
                           torch.from_numpy(d[f"{name}_uvs"]).cuda())
    torch.testing.assert_close(out.cpu(), torch.from_numpy(d[f"{name}_out"]))


@pytest.mark.parametrize("channels,queries", [(1, 1), (3, 1000), (6, 255), (10, 100), (16, 4097), (3, 0)])
def test_sample_matches_oracle_and_reference_test(channels, queries):
    """tests/test_sample.py:8-33 with our kernel, the oracle beside it."""
    import gstex_b200 as G

    g = torch.Generator().manual_seed(channels * 1000 + queries)
    dims, tex = _jagged(g, 37, channels)
    ids = torch.randint(37, (queries,), generator=g)
    uvs = torch.rand(queries, 2, generator=g)
    qd = dims[ids]
    out = G.texture_sample((37, 1, channels), qd.cuda(), tex.cuda(), uvs.cuda())
    assert out.shape == (queries, channels)
    if queries:
        ref = oracle.texture_sample((37, 1, channels), qd.numpy(), tex.numpy(), uvs.numpy())
        torch.testing.assert_close(out.cpu(), torch.from_numpy(ref))


def test_sample_matches_reference_extension():
    from oracle import ref_ext

    C = ref_ext.load()
    if C is None:
        pytest.skip("oracle/_ref/gstex_ref_C.so not built (python oracle/build_ref.py)")
    import gstex_b200 as G

    g = torch.Generator().manual_seed(3)
    for channels in (3, 10):
        dims, tex = _jagged(g, 500, channels, 12, 9)
        ids = torch.randint(500, (200_000,), generator=g)
        uvs = torch.rand(200_000, 2, generator=g)
        qd, texd, uvd = dims[ids].cuda().contiguous(), tex.cuda(), uvs.cuda()
        mine = G.texture_sample((500, 1, channels), qd, texd, uvd)
        ref = C.texture_sample_forward((500, 1, channels), qd, uvd, texd)
        torch.testing.assert_close(mine, ref)


def test_sample_full_size_properties_and_errors():
    """1e7 texels, one query per texel centre (the re-charting access pattern): a sample taken on the texel grid is
    the texel itself, so resampling a texture onto its own charts is the identity; a constant texture stays constant;
    no autograd gradients (as the reference); loud errors."""
    import gstex_b200 as G

    n = 400_000
    g = torch.Generator(device="cuda").manual_seed(2)
    h = torch.randint(1, 9, (n,), device="cuda", generator=g, dtype=torch.int32)
    w = torch.randint(1, 9, (n,), device="cuda", generator=g, dtype=torch.int32)
    hw = (h * w).long()
    start = (torch.cumsum(hw, 0) - hw).int()
    T = int(hw.sum())
    tex = torch.rand(T, 3, device="cuda", generator=g)
    chart = torch.repeat_interleave(torch.arange(n, device="cuda"), hw)
    local = torch.arange(T, device="cuda") - start[chart].long()
    i, j = local // w[chart].long(), local % w[chart].long()
    qd = torch.stack([h[chart], w[chart], start[chart]], -1).contiguous()
    uv = torch.stack([i.float() / h[chart].float(), j.float() / w[chart].float()], -1)
    out = G.texture_sample((n, 1, 3), qd, tex, uv)
    torch.testing.assert_close(out, tex, rtol=1e-5, atol=1e-6)
    const = G.texture_sample((n, 1, 3), qd, torch.full_like(tex, 0.25), torch.rand(T, 2, device="cuda", generator=g))
    torch.testing.assert_close(const, torch.full_like(tex, 0.25), rtol=1e-6, atol=1e-7)
    t2 = tex.clone().requires_grad_(True)
    o = G.texture_sample((n, 1, 3), qd, t2, uv)
    assert not o.requires_grad or torch.autograd.grad(o.sum(), t2, allow_unused=True)[0] is None
    with pytest.raises(RuntimeError):
        G.texture_sample((n, 1, 3), qd.cpu(), tex.cpu(), uv.cpu())
    with pytest.raises(RuntimeError):
        G.texture_sample((n, 1, 4), qd, tex, uv)
    with pytest.raises(NotImplementedError):
        G.texture_sample((n, 1, 3), qd, tex, uv, use_torch_impl=True)

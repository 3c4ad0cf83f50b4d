"""GPU parity of re-charting (`gstex_b200.build_charts`, `resample_texture`; csrc/charts.cu) through the C ABI: against the
reference's op sequence on the same GPU (bit for bit), the goldens made by the reference's own code, and the oracle."""
import os

import numpy as np
import pytest
import torch

from common import GOLDEN_DIR

pytestmark = pytest.mark.gpu


def reference_build_charts(scales, pixel_num, sigma_factor=3.0):
    """gstex.py:841-883 restated statement for statement with torch ops on the tensors' device (the module itself needs
    nerfstudio, absent on the GPU box).  Returns (texture_dims int32, mappings, mid, rounds)."""
    total_pixel_num = pixel_num
    length0 = torch.exp(scales[:, 0])
    length1 = torch.exp(scales[:, 1])

    def get_score(x):
        return torch.sum(torch.ceil(sigma_factor * length0 / x) * torch.ceil(sigma_factor * length1 / x)).item()

    adjustments = torch.ones_like(length0)
    adjustments = torch.sqrt(adjustments**2 / torch.mean(adjustments**2))
    lo = 10.0
    hi = np.sqrt(torch.sum(sigma_factor * sigma_factor * length0 * length1 * (adjustments ** 2)).item() / total_pixel_num)
    mid = 0.5 * (lo + hi)
    score = get_score(mid)
    iter_num, tol = 0, 1e-3
    while score < (1 - tol) * total_pixel_num or score > (1 + tol) * total_pixel_num:
        if score < (1 - tol) * total_pixel_num:
            lo = mid
        else:
            hi = mid
        mid = 0.5 * (lo + hi)
        score = get_score(mid / adjustments)
        iter_num += 1
        if iter_num > 30:
            break
    pixel_scales = mid / adjustments
    dims = torch.zeros(scales.shape[0], 3, dtype=torch.int32, device=scales.device)
    dims[:, 0] = torch.ceil(sigma_factor * length0 / pixel_scales)
    dims[:, 1] = torch.ceil(sigma_factor * length1 / pixel_scales)
    hws = dims[:, 0] * dims[:, 1]
    dims[:, 2] = torch.cumsum(hws, dim=0) - hws
    mappings = torch.stack([1 / (2.0 * sigma_factor * length0), 1 / (2.0 * sigma_factor * length1)], -1)
    return dims, mappings, mid, iter_num


def reference_init_from_dims(old_dims, old_texture, new_dims):
    """jagged_texture.py:118-138 (+ texture_dims_to_query :23-35) with torch ops; sampling by our texture_sample, which
    tests/test_gpu_sample.py holds to the reference's kernel."""
    import gstex_b200 as G

    idxs = torch.arange(new_dims.shape[0], dtype=torch.int64, device=new_dims.device)
    hws = (new_dims[:, 0] * new_dims[:, 1]).long()
    ids = torch.repeat_interleave(idxs, hws, dim=0)
    query_dims = old_dims[ids, :]
    qn = new_dims[ids, :]
    local = torch.arange(int(hws.sum()), dtype=torch.int64, device=new_dims.device) - qn[:, 2]
    uu = (local // qn[:, 1]).float() / qn[:, 0].float()
    vv = (local % qn[:, 1]).float() / qn[:, 1].float()
    return G.texture_sample((1, 1, old_texture.shape[-1]), query_dims.contiguous(), old_texture, torch.stack([uu, vv], -1))


@pytest.mark.parametrize("n,pixel_num,spread", [(500, 5000.0, 0.5), (20000, 300000.0, 1.2), (40, 40.0, 0.3),
                                                 (300_000, 1e7, 0.5), (1_000_000, 1e7, 0.5), (1_000_000, 2e6, 0.7)])
def test_build_charts_bit_identical_to_reference_ops(n, pixel_num, spread):
    import gstex_b200 as G

    g = torch.Generator().manual_seed(n)
    scales = (torch.randn(n, 3, generator=g) * spread + np.log(0.006)).cuda()
    dims, mappings, pixel_scale, total = G.build_charts(scales, pixel_num)
    r_dims, r_map, r_mid, rounds = reference_build_charts(scales, pixel_num)
    assert pixel_scale == r_mid, f"texel size {pixel_scale!r} vs the reference's {r_mid!r} after {rounds} rounds"
    assert dims.dtype == torch.int32 and torch.equal(dims, r_dims), f"{int((dims != r_dims).sum())} chart entries differ"
    assert torch.equal(mappings, r_map)
    assert total == int(r_dims[-1, 2] + r_dims[-1, 0] * r_dims[-1, 1])
    assert abs(total - pixel_num) <= 1e-3 * pixel_num + 1 or rounds > 30
    _, passes = G.chart_texel_size(scales, pixel_num)
    assert passes <= (rounds + 1 + 4) // 5 + 1, (passes, rounds)   # a 5-level subtree per pass instead of one round


@pytest.mark.parametrize("name", ["small", "wide", "tiny"])
def test_build_charts_matches_golden_and_oracle(name):
    import gstex_b200 as G
    import oracle.charts as C

    gold = dict(np.load(os.path.join(GOLDEN_DIR, "charts_golden.npz")))
    g = {k[len(f"charts_{name}_"):]: v for k, v in gold.items() if k.startswith(f"charts_{name}_")}
    dims, mappings, pixel_scale, total = G.build_charts(torch.from_numpy(g["scales"]).cuda(), float(g["pixel_num"]))
    o_dims, o_map, o_ps = C.build_charts(g["scales"], float(g["pixel_num"]))
    for ref_dims, ref_map, ref_ps, what in ((g["texture_dims"], g["mappings"], float(g["pixel_scale"][0]), "golden"),
                                            (o_dims, o_map, o_ps, "oracle")):
        # CPU exp / CPU reduction order differ from the GPU's in the last ulp: same bisection path, 1-ulp lengths
        assert abs(pixel_scale - ref_ps) <= 1e-6 * ref_ps, what
        assert float((dims.cpu().numpy() != ref_dims).mean()) <= 2e-3, what
        np.testing.assert_allclose(mappings.cpu().numpy(), ref_map, rtol=5e-7, err_msg=what)
    assert abs(total - float(g["pixel_num"])) <= 1e-3 * float(g["pixel_num"]) + 1


@pytest.mark.parametrize("name", ["rgb", "six"])
def test_resample_matches_reference_jagged_texture_golden(name):
    import gstex_b200 as G
    import oracle.charts as C

    gold = dict(np.load(os.path.join(GOLDEN_DIR, "charts_golden.npz")))
    g = {k[len(f"resample_{name}_"):]: v for k, v in gold.items() if k.startswith(f"resample_{name}_")}
    t = lambda k: torch.from_numpy(g[k]).cuda()  # noqa: E731
    out = G.resample_texture(t("old_dims"), t("old_texture"), t("new_dims"))
    torch.testing.assert_close(out.cpu(), torch.from_numpy(g["new_texture"]))     # the reference test's own tolerance
    torch.testing.assert_close(out.cpu(), torch.from_numpy(C.resample_texture(g["old_dims"], g["old_texture"], g["new_dims"])))
    same = G.resample_texture(t("old_dims"), t("old_texture"), t("new_dims"), total_size=out.shape[0])
    assert torch.equal(same, out)


def test_rechart_at_full_size_matches_the_unfused_sequence():
    """1M splats, 1e7 texels -> 1.2e7 texels: the fused resampling equals the reference's materialised sequence
    (repeat_interleave + uv tensors + texture_sample), and re-charting onto the SAME charts returns the texture itself up
    to rounding (uv = (i/h, j/w) lands on texel (i, j); h * fp32(i / h) can miss i by an ulp, then the neighbour enters
    with a weight of ~1e-7 -- in the reference as here)."""
    import gstex_b200 as G

    g = torch.Generator().manual_seed(3)
    scales = (torch.randn(1_000_000, 3, generator=g) * 0.5 + np.log(0.006)).cuda()
    old_dims, _, _, T0 = G.build_charts(scales, 1e7)
    tex = torch.rand(T0, 3, generator=g).cuda()
    new_dims, _, _, T1 = G.build_charts(scales + 0.1, 1.2e7)
    out = G.resample_texture(old_dims, tex, new_dims, T1)
    ref = reference_init_from_dims(old_dims, tex, new_dims)
    assert out.shape == ref.shape == (T1, 3)
    assert float((out - ref).abs().max()) <= 1e-6
    same = G.resample_texture(old_dims, tex, old_dims, T0)
    assert float((same - tex).abs().max()) <= 2e-6 and float((same != tex).float().mean()) < 0.2
    with pytest.raises(RuntimeError):
        G.build_charts(scales.cpu(), 1e7)
    with pytest.raises(ValueError):
        G.build_charts(scales, 0)

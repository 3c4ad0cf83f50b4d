"""GPU parity of the fused per-splat activation stage (`gstex_b200.activations`, csrc/activations.cu) through the C ABI:
against the golden vectors made from the reference's own source text, against the numpy oracle, and -- bit for bit --
against the very torch ops the reference model runs (nerfstudio/models/gstex.py:1062-1066, 975-990) on the same GPU."""
import os

import numpy as np
import pytest
import torch

from common import GOLDEN_DIR

pytestmark = pytest.mark.gpu


def reference_ops(quats_raw, scales_raw, opac_raw, mappings):
    """gstex.py:1062-1066 and get_uv_mapping (:975-990) with `quat_to_rotmat` of gstex_cuda/_torch_impl.py:116-138,
    restated op for op (the module itself needs nerfstudio, absent on the GPU box)."""
    quats = quats_raw / quats_raw.norm(dim=-1, keepdim=True)
    scales = torch.zeros_like(scales_raw)
    scales[:, :-1] = torch.clamp(torch.exp(scales_raw[:, :-1]), min=1e-9)
    scales[:, -1] = 1e-5 * torch.mean(scales[:, :-1], dim=-1).detach()
    opacities = torch.sigmoid(opac_raw)
    w, x, y, z = torch.unbind(torch.nn.functional.normalize(quats.detach(), dim=-1), dim=-1)
    R = torch.stack([1 - 2 * (y**2 + z**2), 2 * (x * y - w * z), 2 * (x * z + w * y), 2 * (x * y + w * z),
                     1 - 2 * (x**2 + z**2), 2 * (y * z - w * x), 2 * (x * z - w * y), 2 * (y * z + w * x),
                     1 - 2 * (x**2 + y**2)], dim=-1).reshape(-1, 3, 3)
    uv0 = (0.5 * torch.ones_like(quats_raw[:, :2])).unsqueeze(1)
    umap = (mappings[:, 0, None].detach() * R[:, :, 0].detach()).unsqueeze(1)
    vmap = (mappings[:, 1, None].detach() * R[:, :, 1].detach()).unsqueeze(1)
    return quats, scales, opacities, uv0, umap, vmap


def _inputs(n, seed):
    g = torch.Generator().manual_seed(seed)
    q = (torch.randn(n, 4, generator=g) * torch.exp(torch.randn(n, 1, generator=g))).cuda()
    s = (torch.randn(n, 3, generator=g) * 1.5 - 4.0)
    s[: min(n, 5), 0] = -30.0
    o = (torch.randn(n, 1, generator=g) * 3).cuda()
    m = (torch.rand(n, 2, generator=g) * 40 + 0.5).cuda()
    return q, s.cuda(), o, m


@pytest.mark.parametrize("n", [1, 257, 100_003, 1_000_000])
def test_activations_bit_identical_to_the_reference_torch_ops(n):
    import gstex_b200 as G

    q, s, o, m = _inputs(n, 11 + n)
    leaves = [t.clone().requires_grad_(True) for t in (q, s, o)]
    ours = G.activations(*leaves, m)
    ref_leaves = [t.clone().requires_grad_(True) for t in (q, s, o)]
    ref = reference_ops(*ref_leaves, m)
    for nm, a, b in zip(["quats", "scales", "opacities", "uv0", "umap", "vmap"], ours, ref):
        assert a.shape == b.shape, nm
        assert torch.equal(a, b), f"{nm}: {int((a != b).sum())} of {a.numel()} elements differ from the torch ops"
    assert not ours[3].requires_grad and not ours[4].requires_grad and not ours[5].requires_grad
    g = torch.Generator().manual_seed(5)
    v = [torch.randn(t.shape, generator=g).cuda() for t in ours[:3]]
    torch.autograd.backward(list(ours[:3]), v)
    torch.autograd.backward(list(ref[:3]), v)
    for nm, a, b in zip(["quats", "scales", "opacities"], leaves, ref_leaves):
        scale = float(b.grad.abs().max())
        err = float((a.grad - b.grad).abs().max())
        assert err <= 2e-6 * scale + 1e-30, f"grad {nm}: {err:.3e} at scale {scale:.3e}"
    assert float(leaves[1].grad[:, 2].abs().max()) == 0 and float(leaves[1].grad[0, 0]) == 0   # detached / clamped


def test_activations_match_golden_and_oracle():
    import gstex_b200 as G
    import oracle.activations as A

    d = dict(np.load(os.path.join(GOLDEN_DIR, "act_golden.npz")))
    t = lambda k: torch.from_numpy(d[k]).cuda()  # noqa: E731
    leaves = [t(k).requires_grad_(True) for k in ("quats_raw", "scales_raw", "opacities_raw")]
    outs = G.activations(*leaves, t("mappings"))
    orc = A.forward(d["quats_raw"], d["scales_raw"], d["opacities_raw"], d["mappings"])
    for nm, a, b in zip(["quats", "scales", "opacities", "uv0", "umap", "vmap"], outs, orc):
        for ref, what in ((d[nm], "golden"), (b, "oracle")):
            scale = float(np.abs(ref).max())
            err = float(np.abs(a.detach().cpu().numpy().astype(np.float64) - ref).max())
            assert err <= 1e-6 * scale, f"{nm} vs {what}: {err:.3e} at scale {scale:.3e}"
    torch.autograd.backward(list(outs[:3]), [t("v_quats"), t("v_scales"), t("v_opacities")])
    for nm, leaf in zip(["g_quats_raw", "g_scales_raw", "g_opacities_raw"], leaves):
        scale = float(np.abs(d[nm]).max())
        err = float(np.abs(leaf.grad.cpu().numpy().astype(np.float64) - d[nm]).max())
        assert err <= 3e-6 * scale, f"{nm}: {err:.3e} at scale {scale:.3e}"


def test_activations_feed_the_rasterizer_and_errors_are_loud():
    """The activated tensors go straight into project_bounds / texture_gaussians; gradients reach the raw parameters."""
    import gstex_b200 as G
    from gstex_b200.synthetic import make_scene

    s = make_scene("cfg2", n=4000, width=128, height=128, texels=40000)
    sd = {k: (x.cuda() if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
    raw_q = (sd["quats"] * 1.7).requires_grad_(True)
    raw_s = torch.log(sd["scales"]).detach().requires_grad_(True)
    raw_o = torch.logit(sd["opacities"]).detach().requires_grad_(True)
    mappings = 1.0 / (2.0 * 3.0 * sd["scales"][:, :2])
    quats, scales, opac, uv0, umap, vmap = G.activations(raw_q, raw_s, raw_o, mappings)
    intr = (s["fx"], s["fy"], s["cx"], s["cy"])
    _, depths, centers, extents, nth = G.project_bounds(sd["means"], scales, 1, quats, sd["viewmat"], intr, 128, 128, 16)
    outs = G.texture_gaussians(s["texture_info"], sd["texture_dims"], centers, extents, depths, nth, sd["colors"], opac,
                               sd["means"], scales, 1, quats, uv0, umap, vmap, sd["texture"], sd["viewmat"], sd["c2w"],
                               *intr, 128, 128, 16, s["settings"], sd["background"])
    (outs[0].sum() + outs[4].sum()).backward()
    for nm, p in (("quats", raw_q), ("scales", raw_s), ("opacities", raw_o)):
        assert p.grad is not None and torch.isfinite(p.grad).all() and float(p.grad.abs().max()) > 0, nm
    torch.testing.assert_close(umap, sd["umap"], rtol=2e-5, atol=1e-4)   # the scene's own maps, built the same way
    with pytest.raises(RuntimeError):
        G.activations(raw_q.detach().cpu(), raw_s.detach().cpu(), raw_o.detach().cpu())   # no CPU path
    with pytest.raises(ValueError):
        G.activations(raw_q.detach()[:, :3], raw_s.detach(), raw_o.detach())
    q0 = G.activations(raw_q.detach()[:0], raw_s.detach()[:0], raw_o.detach()[:0])[0]
    assert q0.shape == (0, 4)

"""GPU parity of the spherical-harmonics colour stage (gstex_b200/sh.py, csrc/sh.cu): against the CPU oracle, the
golden vectors made by the reference's Python, the reference's CUDA extension when oracle/_ref holds it, and --
at the headline size -- size-independent properties.  Mirrors the reference's tests/test_sh.py (same tolerance:
torch.testing.assert_close defaults for fp32, rtol 1.3e-6 / atol 1e-5)."""
import os

import numpy as np
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "sh_golden.npz")
TOL = dict(rtol=1.3e-6, atol=1e-5)


@pytest.mark.parametrize("degree", [0, 1, 2, 3, 4])
def test_sh_matches_golden(degree):
    import gstex_b200 as G

    d = np.load(GOLD)
    v = torch.from_numpy(d[f"viewdirs_{degree}"]).cuda()
    c = torch.from_numpy(d[f"coeffs_{degree}"]).cuda().requires_grad_(True)
    w = torch.from_numpy(d[f"w_{degree}"]).cuda()
    col = G.spherical_harmonics(degree, v, c)
    (col * w).sum().backward()
    torch.testing.assert_close(col.detach().cpu(), torch.from_numpy(d[f"colors_{degree}"]), **TOL)
    torch.testing.assert_close(c.grad.cpu(), torch.from_numpy(d[f"v_coeffs_{degree}"]), **TOL)


@pytest.mark.parametrize("n", [1, 127, 128, 129, 5000])
@pytest.mark.parametrize("degree,use", [(0, 0), (1, 1), (2, 1), (3, 3), (4, 2), (4, 4), (4, 0)])
def test_sh_matches_oracle(n, degree, use):
    import gstex_b200 as G

    g = torch.Generator().manual_seed(100 * n + 10 * degree + use)
    v = torch.randn(n, 3, generator=g) * 3.0  # not normalised
    c = torch.randn(n, G.num_sh_bases(degree), 3, generator=g)
    w = torch.randn(n, 3, generator=g)
    cd = c.cuda().requires_grad_(True)
    col = G.spherical_harmonics(use, v.cuda(), cd)
    col.backward(w.cuda())
    torch.testing.assert_close(col.detach().cpu(), torch.from_numpy(oracle.spherical_harmonics(use, v.numpy(), c.numpy())), **TOL)
    ref_vc = oracle.spherical_harmonics_vjp(use, c.shape[1], v.numpy(), w.numpy())
    torch.testing.assert_close(cd.grad.cpu(), torch.from_numpy(ref_vc), **TOL)
    assert bool((cd.grad[:, G.num_sh_bases(use):] == 0).all())


@pytest.mark.parametrize("degree,use", [(0, 0), (2, 2), (3, 1), (4, 4), (4, 3)])
def test_sh_matches_reference_extension(degree, use):
    from oracle import ref_ext

    C = ref_ext.load()
    if C is None:
        pytest.skip("oracle/_ref/gstex_ref_C.so not built (python oracle/build_ref.py)")
    import gstex_b200 as G

    n = 40001
    g = torch.Generator().manual_seed(7 + degree)
    v = (torch.randn(n, 3, generator=g) * 2).cuda()
    c = torch.randn(n, G.num_sh_bases(degree), 3, generator=g).cuda()
    w = torch.randn(n, 3, generator=g).cuda()
    mine = G.spherical_harmonics(use, v, c)
    ref = C.compute_sh_forward(n, degree, use, v, c)
    torch.testing.assert_close(mine, ref, **TOL)
    cd = c.clone().requires_grad_(True)
    G.spherical_harmonics(use, v, cd).backward(w)
    torch.testing.assert_close(cd.grad, C.compute_sh_backward(n, degree, use, v, w), **TOL)


def test_sh_reference_test_loop():
    """The reference's own test (tests/test_sh.py:11-48), shortened: Adam drives degree-4 coefficients of one point
    to a constant colour through our forward/backward."""
    import gstex_b200 as G

    torch.manual_seed(0)
    dev = torch.device("cuda:0")
    gt = torch.ones(1, 3, device=dev) * 0.5
    v = torch.randn(1, 3, device=dev)
    v /= torch.linalg.norm(v, dim=-1, keepdim=True)
    c = torch.rand(1, G.num_sh_bases(4), 3, device=dev, requires_grad=True)
    opt = torch.optim.Adam([c], lr=1e-2)
    for _ in range(1000):
        opt.zero_grad()
        col = G.spherical_harmonics(4, v, c)
        torch.square(col - gt).mean().backward()
        opt.step()
    torch.testing.assert_close(G.spherical_harmonics(4, v, c).detach(), gt, rtol=1e-3, atol=1e-3)


def test_sh_full_size_properties():
    """1M points, degree 3 (the model's default): linearity in the coefficients, degree-0 closed form, gradient =
    basis (x) upstream, zeros above the used bases, errors for bad shapes / CPU tensors."""
    import gstex_b200 as G

    n = 1_000_000
    g = torch.Generator(device="cuda").manual_seed(1)
    v = torch.randn(n, 3, device="cuda", generator=g)
    c = torch.randn(n, 16, 3, device="cuda", generator=g)
    a = G.spherical_harmonics(3, v, c)
    b = G.spherical_harmonics(3, v, 2 * c)
    torch.testing.assert_close(a + b, G.spherical_harmonics(3, v, 3 * c), rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(G.spherical_harmonics(0, v, c), 0.28209479177387814 * c[:, 0], rtol=1e-6, atol=1e-7)
    cd = c.clone().requires_grad_(True)
    G.spherical_harmonics(2, v, cd).backward(torch.ones(n, 3, device="cuda"))
    assert bool((cd.grad[:, 9:] == 0).all())
    basis = cd.grad[:, :, 0]  # upstream ones: the gradient is the basis itself
    torch.testing.assert_close((basis[:, :9, None] * c[:, :9]).sum(1), G.spherical_harmonics(2, v, c), rtol=1e-5, atol=1e-5)
    with pytest.raises(RuntimeError):
        G.spherical_harmonics(1, v.cpu(), c.cpu())
    with pytest.raises(RuntimeError):
        G.spherical_harmonics(1, v, c[:, :, :2].contiguous())
    with pytest.raises(AssertionError):
        G.spherical_harmonics(4, v, c)

"""GPU parity of `photometric_loss` (gstex_b200/loss.py, csrc/loss.cu): against the golden
vectors made from the reference's own lines + torch.autograd, against the CPU oracle, against the same torch-op chain
run on the GPU at the headline image size, and wired to `texture_gaussians` the way the trainer wires it."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

import oracle
from test_oracle_loss import CASES, check_against_golden, load_case

pytestmark = pytest.mark.gpu


def _run(c, lam, v_main=1.0, grads=True):
    import gstex_b200 as G

    t = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in c.items()
         if k in ("out_img", "out_texture", "out_alpha", "background", "gt_img", "mask")}
    leaves = [t[k].clone().requires_grad_(grads) for k in ("out_img", "out_texture", "out_alpha")]
    rgb, main, l1, ssim = G.photometric_loss(leaves[0], leaves[1], leaves[2], t["background"], t["gt_img"], lam,
                                             t.get("mask"))
    got = dict(rgb=rgb.cpu().numpy(), Ll1=float(l1), ssim=float(ssim), main_loss=float(main))
    if grads:
        (main * v_main).backward()
        got.update(v_out_img=leaves[0].grad.cpu().numpy(), v_out_texture=leaves[1].grad.cpu().numpy(),
                   v_out_alpha=leaves[2].grad.cpu().numpy())
    return got


@pytest.mark.parametrize("name", CASES)
def test_loss_matches_golden(name):
    c = load_case(name)
    check_against_golden(_run(c, float(c["params"][0]), float(c["params"][1])), c)


def _scene(H, W, C, seed, bg, use_mask):
    g = torch.Generator().manual_seed(seed)
    base = torch.rand(H // 8 + 2, W // 8 + 2, 3, generator=g)
    gt = F.interpolate(base.permute(2, 0, 1)[None], size=(H, W), mode="bilinear", align_corners=False)[0].permute(1, 2, 0).contiguous()
    alpha = torch.rand(H, W, generator=g).clamp(0.05, 1.0)
    alpha[torch.rand(H, W, generator=g) < 0.2] = 0.0
    c = dict(out_img=(gt * 0.8 + 0.2 * torch.randn(H, W, 3, generator=g)) * alpha[..., None],
             out_texture=0.2 * torch.randn(H, W, C, generator=g) * alpha[..., None], out_alpha=alpha,
             background=torch.tensor(bg, dtype=torch.float32), gt_img=gt)
    if use_mask:
        c["mask"] = (torch.rand(H, W, generator=g) > 0.3).float()
    return {k: v.numpy() for k, v in c.items()}


@pytest.mark.parametrize("H,W,C,bg,use_mask", [(11, 11, 3, [0, 0, 0], False), (16, 16, 3, [1, 1, 1], False),
                                               (27, 26, 3, [0, 0, 0], True), (97, 130, 6, [0.1, 0.4, 0.8], False),
                                               (256, 256, 3, [0, 0, 0], False)])
def test_loss_matches_oracle(H, W, C, bg, use_mask):
    """ragged image sizes around the 16-pixel tile and the 11-tap window, 3 and 6 channels, with and without a mask"""
    c = _scene(H, W, C, 100 + H, bg, use_mask)
    want = oracle.photometric_loss(c["out_img"], c["out_texture"], c["out_alpha"], c["background"], c["gt_img"], 0.2,
                                   c.get("mask"), v_main=1.0)
    want["losses"] = np.array([want["Ll1"], want["ssim"], want["main_loss"]], np.float32)
    check_against_golden(_run(c, 0.2), want)


def _torch_chain(out_img, out_texture, out_alpha, background, gt, lam):
    """gstex.py:1204-1205, 1301-1302, 1319 with pytorch_msssim v1.0.0's ops, on whatever device the inputs are"""
    rgb = torch.clamp(out_img + out_texture[:, :, 0:3] + (1 - out_alpha[:, :, None]) * background[None, None, :], 0.0, 1.0)
    coords = torch.arange(11, dtype=torch.float) - 5
    g = torch.exp(-(coords ** 2) / (2 * 1.5 ** 2))
    win = (g / g.sum()).to(rgb.device)[None, None, None, :].repeat(3, 1, 1, 1)

    def filt(x):
        return F.conv2d(F.conv2d(x, win.transpose(2, 3), groups=3), win, groups=3)

    X, Y = gt.permute(2, 0, 1)[None], rgb.permute(2, 0, 1)[None]
    mu1, mu2 = filt(X), filt(Y)
    s1, s2, s12 = filt(X * X) - mu1 * mu1, filt(Y * Y) - mu2 * mu2, filt(X * Y) - mu1 * mu2
    ssim = (((2 * mu1 * mu2 + 1e-4) / (mu1 * mu1 + mu2 * mu2 + 1e-4)) * ((2 * s12 + 9e-4) / (s1 + s2 + 9e-4))).flatten(2).mean(-1).mean()
    l1 = torch.abs(gt - rgb).mean()
    return rgb, l1, ssim, (1 - lam) * l1 + lam * (1 - ssim)


def test_loss_headline_size_against_torch_ops():
    """800x800 (the headline view): the kernels against the reference's torch-op chain + autograd on the same GPU,
    and bit-reproducible from run to run (fixed-order reduction, no atomics)."""
    import gstex_b200 as G

    c = _scene(800, 800, 3, 7, [0, 0, 0], False)
    t = {k: torch.from_numpy(v).cuda() for k, v in c.items()}
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        leaves = [t[k].clone().requires_grad_(True) for k in ("out_img", "out_texture", "out_alpha")]
        rgb_t, l1_t, ssim_t, main_t = _torch_chain(*leaves, t["background"], t["gt_img"], 0.2)
        main_t.backward()
    finally:
        torch.backends.cudnn.allow_tf32 = old
    runs = []
    for _ in range(2):
        mine = [t[k].clone().requires_grad_(True) for k in ("out_img", "out_texture", "out_alpha")]
        rgb, main, l1, ssim = G.photometric_loss(*mine, t["background"], t["gt_img"], 0.2)
        main.backward()
        runs.append((rgb, main.detach(), l1, ssim, *[m.grad for m in mine]))
    for a, b in zip(*runs):
        assert torch.equal(a, b)
    rgb, main, l1, ssim, *grads = runs[0]
    assert torch.equal(rgb, rgb_t.detach())
    for a, b in ((main, main_t), (l1, l1_t), (ssim, ssim_t)):
        assert abs(float(a) - float(b)) <= 1e-5 * abs(float(b))
    for a, b, nm in zip(grads, [x.grad for x in leaves], ["v_out_img", "v_out_texture", "v_out_alpha"]):
        scale = float(b.abs().max())
        bad = ((a - b).abs() > 1e-4 * scale).float().mean()
        assert float(bad) <= 1e-3, f"{nm}: share {float(bad):.2e} beyond 1e-4 of scale {scale:.3e}"


def test_loss_feeds_the_rasterizer_backward():
    """the trainer's wiring: texture_gaussians -> photometric_loss -> backward reaches the splat and texel parameters,
    and equals the torch-op chain's gradients through the same rasterizer"""
    import gstex_b200 as G
    from gstex_b200.synthetic import make_scene

    s = make_scene("cfg2", n=20000, width=256, height=256)
    sd = {k: (x.cuda() if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
    intr = (s["fx"], s["fy"], s["cx"], s["cy"])
    H, W, bw = s["H"], s["W"], 16
    keys = ["colors", "opacities", "means", "scales", "quats", "texture"]
    vm, c2w = s["viewmats"][0].cuda().contiguous(), s["c2ws"][0].cuda().contiguous()
    gt = torch.rand(H, W, 3, generator=torch.Generator().manual_seed(5)).cuda()
    bg = torch.zeros(3, device="cuda")

    def run(fused):
        pd = {k: sd[k].clone().requires_grad_(True) for k in keys}
        _, depths, centers, extents, nth = G.project_bounds(pd["means"].detach(), pd["scales"].detach(), 1,
                                                            pd["quats"].detach(), vm, intr, H, W, bw)
        outs = G.texture_gaussians(s["texture_info"], sd["texture_dims"], centers, extents, depths, nth, pd["colors"],
                                   pd["opacities"], pd["means"], pd["scales"], 1, pd["quats"], sd["uv0"], sd["umap"],
                                   sd["vmap"], pd["texture"], vm, c2w, *intr, H, W, bw, s["settings"], bg)
        out_img, _, _, out_alpha, out_texture, _ = outs
        if fused:
            main = G.photometric_loss(out_img, out_texture, out_alpha, bg, gt, 0.2)[1]
        else:
            main = _torch_chain(out_img, out_texture, out_alpha, bg, gt, 0.2)[3]
        main.backward()
        return float(main), [pd[k].grad for k in keys]

    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        (m_ref, g_ref), (m_fus, g_fus) = run(False), run(True)
    finally:
        torch.backends.cudnn.allow_tf32 = old
    assert abs(m_ref - m_fus) <= 1e-5 * abs(m_ref)
    for k, a, b in zip(keys, g_ref, g_fus):
        scale = float(a.abs().max())
        assert scale > 0
        err = float((a - b).abs().max())
        assert err <= 2e-4 * scale, f"{k}: {err:.3e} at scale {scale:.3e}"


def test_loss_argument_errors():
    import gstex_b200 as G

    z = torch.zeros(10, 12, 3, device="cuda")
    with pytest.raises(RuntimeError):
        G.photometric_loss(z, z, z[..., 0], torch.zeros(3, device="cuda"), z)  # smaller than the window
    z = torch.zeros(12, 12, 3)
    with pytest.raises(RuntimeError):
        G.photometric_loss(z, z, z[..., 0], torch.zeros(3), z)  # CPU tensors: there is no CPU path


@pytest.mark.parametrize("size", [(97, 130), (256, 320)])
def test_loss_takes_the_uint8_ground_truth_as_it_is(size):
    """`gt_img` as the uint8 image the data manager holds == `get_gt_img`'s `image.float() / 255.0` first
    (NS/models/gstex.py:1244-1245), bit for bit: values, losses and all three gradients."""
    from gstex_b200.loss import photometric_loss

    H, W = size
    g = torch.Generator().manual_seed(H * 1000 + W)
    dev = torch.device("cuda")
    gt_u8 = torch.randint(0, 256, (H, W, 3), generator=g, dtype=torch.uint8).to(dev)
    gt_f = gt_u8.float() / 255.0
    base = [torch.rand(H, W, 3, generator=g).to(dev) * 0.6, torch.rand(H, W, 3, generator=g).to(dev) * 0.5,
            torch.rand(H, W, generator=g).to(dev)]
    bg = torch.tensor([0.2, 0.7, 0.4], device=dev)
    mask = (torch.rand(H, W, generator=g) > 0.2).float().to(dev)

    def run(gt, m):
        ins = [t.clone().requires_grad_(True) for t in base]
        rgb, main, l1, ssim = photometric_loss(ins[0], ins[1], ins[2], bg, gt, 0.2, mask=m)
        (main * 1.7).backward()
        return [rgb, main.detach(), l1, ssim] + [t.grad for t in ins]

    for m in (None, mask):
        for a, b in zip(run(gt_f, m), run(gt_u8, m)):
            assert torch.equal(a, b)
    with pytest.raises(RuntimeError):
        photometric_loss(base[0], base[1], base[2], bg, gt_u8.to(torch.int32), 0.2)

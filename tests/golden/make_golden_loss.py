"""Golden vectors for the compositing + photometric-loss row.

    python tests/golden/make_golden_loss.py

The reference computes this inline in `GStexModel.get_outputs` / `get_loss_dict`
(/root/reference/nerfstudio/models/gstex.py:1204-1205, 1297-1302, 1319) with `pytorch_msssim.SSIM(data_range=1.0,
size_average=True, channel=3)` (gstex.py:33,351).  Neither nerfstudio nor pytorch_msssim is importable in this
container (pytorch_msssim is an unpinned third-party dependency, pyproject.toml:60, absent from the wheelhouse), so
the lines are restated here with the same torch ops, and SSIM with pytorch_msssim v1.0.0's own ops
(`_fspecial_gauss_1d`, `gaussian_filter` = depthwise `F.conv2d` without padding along H then W, `_ssim`).
Gradients come from torch.autograd.  Parity with the pytorch_msssim package itself therefore stays unpinned.
"""
import os

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))


def fspecial_gauss_1d(size=11, sigma=1.5):
    coords = torch.arange(size, dtype=torch.float)
    coords -= size // 2
    g = torch.exp(-(coords ** 2) / (2 * sigma ** 2))
    g /= g.sum()
    return g.unsqueeze(0).unsqueeze(0)


def gaussian_filter(inp, win):
    C = inp.shape[1]
    out = inp
    for i, s in enumerate(inp.shape[2:]):
        assert s >= win.shape[-1]
        out = F.conv2d(out, weight=win.transpose(2 + i, -1), stride=1, padding=0, groups=C)
    return out


def ssim(X, Y, data_range=1.0, K=(0.01, 0.03)):
    win = fspecial_gauss_1d().repeat([X.shape[1]] + [1] * (len(X.shape) - 1))
    C1, C2 = (K[0] * data_range) ** 2, (K[1] * data_range) ** 2
    mu1, mu2 = gaussian_filter(X, win), gaussian_filter(Y, win)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    sigma1_sq = gaussian_filter(X * X, win) - mu1_sq
    sigma2_sq = gaussian_filter(Y * Y, win) - mu2_sq
    sigma12 = gaussian_filter(X * Y, win) - mu1_mu2
    cs_map = (2 * sigma12 + C2) / (sigma1_sq + sigma2_sq + C2)
    ssim_map = ((2 * mu1_mu2 + C1) / (mu1_sq + mu2_sq + C1)) * cs_map
    return torch.flatten(ssim_map, 2).mean(-1).mean()


def reference_lines(out_img, out_texture, out_alpha, background, gt_img, ssim_lambda, mask):
    rgb = out_img + out_texture[:, :, 0:3] + (1 - out_alpha[:, :, None]) * background[None, None, :]  # gstex.py:1204
    rgb = torch.clamp(rgb, min=0.0, max=1.0)                                                          # :1205
    pred_img, gt = rgb, gt_img
    if mask is not None:                                                                              # :1292-1298
        gt = gt * mask
        pred_img = pred_img * mask
    Ll1 = torch.abs(gt - pred_img).mean()                                                             # :1301
    s = ssim(gt.permute(2, 0, 1)[None, ...], pred_img.permute(2, 0, 1)[None, ...])                    # :1302
    main = (1 - ssim_lambda) * Ll1 + ssim_lambda * (1 - s)                                            # :1319
    return rgb, Ll1, s, main


def case(g, H, W, C, lam, use_mask, bg, v_main):
    """Rendered layers that look like the rasterizer's: smooth-ish image, sparse alpha, values that leave [0, 1]
    on both sides (so the clamp mask matters) and empty pixels that are exactly zero."""
    base = torch.rand(H // 4 + 2, W // 4 + 2, 3, generator=g)
    gt = F.interpolate(base.permute(2, 0, 1)[None], size=(H, W), mode="bilinear", align_corners=False)[0].permute(1, 2, 0).contiguous()
    out_alpha = torch.rand(H, W, generator=g).clamp(0.05, 1.0)
    empty = torch.rand(H, W, generator=g) < 0.15
    out_alpha[empty] = 0.0
    out_img = (gt * 0.7 + 0.25 * torch.randn(H, W, 3, generator=g)) * out_alpha[..., None]
    out_texture = 0.3 * torch.randn(H, W, C, generator=g) * out_alpha[..., None]
    background = torch.tensor(bg, dtype=torch.float32)
    mask = (torch.rand(H, W, 1, generator=g) > 0.2).float() if use_mask else None
    leaves = [t.clone().requires_grad_(True) for t in (out_img, out_texture, out_alpha)]
    rgb, Ll1, s, main = reference_lines(leaves[0], leaves[1], leaves[2], background, gt, lam, mask)
    (main * v_main).backward()
    d = dict(out_img=out_img, out_texture=out_texture, out_alpha=out_alpha, background=background, gt_img=gt,
             rgb=rgb.detach(), losses=torch.stack([Ll1, s, main]).detach(), v_out_img=leaves[0].grad,
             v_out_texture=leaves[1].grad, v_out_alpha=leaves[2].grad,
             params=torch.tensor([lam, v_main], dtype=torch.float32))
    if mask is not None:
        d["mask"] = mask
    return {k: v.numpy() for k, v in d.items()}


def main():
    g = torch.Generator().manual_seed(2024)
    out = {}
    cases = {"small": case(g, 24, 37, 3, 0.2, False, [0.0, 0.0, 0.0], 1.0),
             "masked": case(g, 48, 40, 3, 0.2, True, [1.0, 1.0, 1.0], 1.0),
             "viewer6": case(g, 33, 50, 6, 0.35, False, [0.2, 0.5, 0.9], 2.5),
             "min11": case(g, 11, 11, 3, 0.2, False, [0.0, 0.0, 0.0], 1.0)}
    for name, c in cases.items():
        for k, v in c.items():
            out[f"{name}_{k}"] = v
    np.savez_compressed(os.path.join(HERE, "loss_golden.npz"), **out)
    print("wrote loss_golden.npz", {n: c["losses"] for n, c in cases.items()})


if __name__ == "__main__":
    main()

"""Compositing + photometric loss on the far side of the rasterizer.

Mirrors what `GStexModel` does with the rasterizer's outputs
(/root/reference/nerfstudio/models/gstex.py:1204-1205, 1297-1302, 1319):

    rgb       = clamp(out_img + out_texture[:, :, 0:3] + (1 - out_alpha[:, :, None]) * background, 0, 1)
    Ll1       = |gt_img - rgb|.mean()
    simloss   = 1 - SSIM(gt_img, rgb)            # pytorch_msssim.SSIM(data_range=1.0, size_average=True, channel=3)
    main_loss = (1 - ssim_lambda) * Ll1 + ssim_lambda * simloss

as one forward kernel, a one-block reduction and one backward kernel (`csrc/loss.cu`) instead of ~45 torch kernels
forward and about twice that in autograd's backward.  The gradients that come out of the backward -- `v_out_img`,
`v_out_texture`, `v_out_alpha` -- are exactly what `_TextureGaussians.backward` takes as upstream gradients.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
from torch import Tensor
from torch.autograd import Function

from . import _lib

WIN_SIZE = 11     # pytorch_msssim.SSIM defaults, as the reference constructs it (gstex.py:351)
WIN_SIGMA = 1.5


def photometric_loss(out_img: Tensor, out_texture: Tensor, out_alpha: Tensor, background: Tensor, gt_img: Tensor,
                     ssim_lambda: float = 0.2, mask: Optional[Tensor] = None) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """Returns `(rgb, main_loss, Ll1, ssim)`.

    `out_img` (H,W,3), `out_texture` (H,W,C>=3), `out_alpha` (H,W) are `texture_gaussians`' outputs; `background` (3,);
    `gt_img` (H,W,3) is the ground truth already composited with the background (gstex.py:1286) -- fp32, or the uint8
    image as the data manager holds it, which the kernels convert as `get_gt_img` does (`image.float() / 255.0`,
    gstex.py:1244-1245; same bits, no fp32 copy); `mask` (H,W) or
    (H,W,1) multiplies both images before the two loss terms (gstex.py:1292-1298).  `main_loss` (0-dim) carries the
    gradient to the three rendered layers; `rgb` (H,W,3, clamped, unmasked), `Ll1` and `ssim` are detached values
    for logging/metrics (gstex.py:1384-1388)."""
    if mask is not None:
        mask = mask.reshape(mask.shape[0], mask.shape[1]).to(torch.float32).contiguous()
    return _PhotometricLoss.apply(out_img.contiguous(), out_texture.contiguous(), out_alpha.contiguous(),
                                  background.contiguous(), gt_img.contiguous(), float(ssim_lambda), mask)


class _PhotometricLoss(Function):
    @staticmethod
    def forward(ctx, out_img, out_texture, out_alpha, background, gt_img, ssim_lambda, mask):
        tensors = [out_img, out_texture, out_alpha, background, gt_img] + ([mask] if mask is not None else [])
        dev = _lib.require_cuda(*tensors)
        for t, nm in zip(tensors, ["out_img", "out_texture", "out_alpha", "background", "gt_img", "mask"]):
            if nm == "gt_img" and t.dtype == torch.uint8:
                continue
            _lib.require_dtype(t, torch.float32, nm)
        u8 = gt_img.dtype == torch.uint8
        if out_img.dim() != 3 or out_img.shape[2] != 3:
            raise RuntimeError("out_img must be (H, W, 3)")
        H, W = int(out_img.shape[0]), int(out_img.shape[1])
        C = int(out_texture.shape[-1])
        if tuple(out_texture.shape) != (H, W, C) or C < 3:
            raise RuntimeError("out_texture must be (H, W, C) with C >= 3")
        if tuple(out_alpha.shape) != (H, W) or tuple(gt_img.shape) != (H, W, 3) or background.numel() != 3:
            raise RuntimeError("out_alpha must be (H, W), gt_img (H, W, 3), background (3,)")
        if mask is not None and tuple(mask.shape) != (H, W):
            raise RuntimeError("mask must be (H, W) or (H, W, 1)")
        if H < WIN_SIZE or W < WIN_SIZE:
            raise RuntimeError(f"SSIM needs an image of at least {WIN_SIZE}x{WIN_SIZE} pixels")
        lib = _lib.load()
        want_grad = any(ctx.needs_input_grad[:3])
        rgb = torch.empty((H, W, 3), dtype=torch.float32, device=dev)
        losses = torch.empty((3,), dtype=torch.float32, device=dev)
        maps = torch.empty((3, H, W, 3), dtype=torch.float32, device=dev) if want_grad else None
        scratch = torch.empty((int(lib.gstex_photometric_loss_scratch_bytes(H, W)),), dtype=torch.uint8, device=dev)
        p = _lib.ptr
        with torch.cuda.device(dev):
            fwd = lib.gstex_photometric_loss_forward_u8 if u8 else lib.gstex_photometric_loss_forward
            rc = fwd(H, W, C, p(out_img), p(out_texture), p(out_alpha), p(background),
                     p(gt_img), p(mask), ssim_lambda, WIN_SIZE, WIN_SIGMA, p(rgb), p(losses),
                     p(maps), p(scratch), scratch.numel(), _lib.stream_ptr(dev))
        _lib.check(rc, "gstex_photometric_loss_forward")
        ctx.ssim_lambda = ssim_lambda
        ctx.has_mask = mask is not None
        ctx.save_for_backward(*tensors, *([maps] if want_grad else []))
        ctx.want_grad = want_grad
        main_loss, l1, ssim = losses[2], losses[0], losses[1]
        ctx.mark_non_differentiable(rgb, l1, ssim)
        return rgb, main_loss, l1, ssim

    @staticmethod
    def backward(ctx, _v_rgb, v_main, _v_l1, _v_ssim):
        if not ctx.want_grad:
            return (None,) * 7
        saved = ctx.saved_tensors
        out_img, out_texture, out_alpha, background, gt_img = saved[:5]
        mask = saved[5] if ctx.has_mask else None
        maps = saved[-1]
        dev = out_img.device
        H, W, C = out_img.shape[0], out_img.shape[1], out_texture.shape[-1]
        v_out_img = torch.empty_like(out_img)
        v_out_texture = torch.empty_like(out_texture)
        v_out_alpha = torch.empty_like(out_alpha)
        v_main = v_main.to(torch.float32).contiguous()
        p = _lib.ptr
        with torch.cuda.device(dev):
            lib = _lib.load()
            bwd = lib.gstex_photometric_loss_backward_u8 if gt_img.dtype == torch.uint8 else lib.gstex_photometric_loss_backward
            rc = bwd(
                int(H), int(W), int(C), p(out_img), p(out_texture), p(out_alpha), p(background), p(gt_img), p(mask),
                ctx.ssim_lambda, WIN_SIZE, WIN_SIGMA, p(maps), p(v_main), p(v_out_img), p(v_out_texture), p(v_out_alpha),
                _lib.stream_ptr(dev))
        _lib.check(rc, "gstex_photometric_loss_backward")
        return v_out_img, v_out_texture, v_out_alpha, None, None, None, None

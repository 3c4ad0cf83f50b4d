"""Per-splat activation stage of the caller, fused.

What `GStexModel.get_outputs` does with its raw parameters before every view
(/root/reference/nerfstudio/models/gstex.py:1062-1066, 1069 -> get_uv_mapping :975-990):

    quats     = self.quats / self.quats.norm(dim=-1, keepdim=True)
    scales    = zeros_like(self.scales); scales[:, :-1] = clamp(exp(self.scales[:, :-1]), min=1e-9)
    scales[:, -1] = 1e-5 * mean(scales[:, :-1], dim=-1).detach()
    opacities = sigmoid(self.opacities)
    uv0, umap, vmap = get_uv_mapping(means, quats, self.mappings)      # all detached

~30 elementwise / reduction launches over N forward and about as many in autograd's backward; here one kernel each way
(`csrc/activations.cu`), forward bit-identical to the torch ops on the same GPU.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
from torch import Tensor
from torch.autograd import Function

from . import _lib


def activations(quats: Tensor, scales: Tensor, opacities: Tensor,
                mappings: Optional[Tensor] = None) -> Tuple[Tensor, Tensor, Tensor, Optional[Tensor], Optional[Tensor], Optional[Tensor]]:
    """Returns `(quats, scales, opacities, uv0, umap, vmap)` -- the activated tensors `texture_gaussians`,
    `get_aabb_2d` and `project_points` take -- from the model's raw parameters `quats (N,4)`, `scales (N,3)` (logs; the
    third column is ignored, as in the reference), `opacities (N,1)` (logits) and `mappings (N,2)`.  Gradients flow to
    `quats`, `scales[:, :2]` and `opacities`; `uv0 (N,1,2)`, `umap`, `vmap (N,1,3)` are detached (None when `mappings`
    is None)."""
    if quats.dim() != 2 or quats.shape[1] != 4:
        raise ValueError("quats must be (N, 4)")
    n = quats.shape[0]
    if tuple(scales.shape) != (n, 3):
        raise ValueError("scales must be (N, 3)")
    if opacities.numel() != n:
        raise ValueError("opacities must be (N, 1)")
    if mappings is not None and tuple(mappings.shape) != (n, 2):
        raise ValueError("mappings must be (N, 2)")
    return _Activations.apply(quats.contiguous(), scales.contiguous(), opacities.contiguous(),
                              None if mappings is None else mappings.detach().contiguous())


class _Activations(Function):
    @staticmethod
    def forward(ctx, quats_raw, scales_raw, opac_raw, mappings):
        tensors = [quats_raw, scales_raw, opac_raw] + ([mappings] if mappings is not None else [])
        dev = _lib.require_cuda(*tensors)
        for t, nm in zip(tensors, ["quats", "scales", "opacities", "mappings"]):
            _lib.require_dtype(t, torch.float32, nm)
        n = quats_raw.shape[0]
        f32 = dict(dtype=torch.float32, device=dev)
        quats = torch.empty((n, 4), **f32)
        scales = torch.empty((n, 3), **f32)
        opac = torch.empty_like(opac_raw)
        uv0 = umap = vmap = None
        if mappings is not None:
            uv0, umap, vmap = torch.empty((n, 1, 2), **f32), torch.empty((n, 1, 3), **f32), torch.empty((n, 1, 3), **f32)
        p = _lib.ptr
        with torch.cuda.device(dev):
            rc = _lib.load().gstex_activations_forward(n, p(quats_raw), p(scales_raw), p(opac_raw), p(mappings), p(quats),
                                                       p(scales), p(opac), p(uv0), p(umap), p(vmap), _lib.stream_ptr(dev))
        _lib.check(rc, "gstex_activations_forward")
        ctx.save_for_backward(quats_raw, scales_raw, quats, scales, opac)
        ctx.mark_non_differentiable(*[t for t in (uv0, umap, vmap) if t is not None])
        return quats, scales, opac, uv0, umap, vmap

    @staticmethod
    def backward(ctx, v_quats, v_scales, v_opac, _v_uv0, _v_umap, _v_vmap):
        quats_raw, scales_raw, quats, scales, opac = ctx.saved_tensors
        dev = quats_raw.device
        n = quats_raw.shape[0]
        need = ctx.needs_input_grad
        c = lambda g: None if g is None else g.contiguous().float()  # noqa: E731
        v_quats, v_scales, v_opac = c(v_quats), c(v_scales), c(v_opac)
        g_q = torch.empty_like(quats_raw) if need[0] else None
        g_s = torch.empty_like(scales_raw) if need[1] else None
        g_o = torch.empty_like(opac) if need[2] else None
        p = _lib.ptr
        with torch.cuda.device(dev):
            rc = _lib.load().gstex_activations_backward(n, p(quats_raw), p(scales_raw), p(quats), p(scales), p(opac),
                                                        p(v_quats), p(v_scales), p(v_opac), p(g_q), p(g_s), p(g_o),
                                                        _lib.stream_ptr(dev))
        _lib.check(rc, "gstex_activations_backward")
        return g_q, g_s, g_o, None

"""`texture_sample`, the drop-in for the reference's `gstex_cuda/texture_sample.py`
(/root/reference/submodules/GStex_cuda/gstex_cuda/texture_sample.py:13-86): bilinear queries into a jagged texture,
used by the model's re-charting (nerfstudio/models/jagged_texture.py:116-143).  SURVEY.md section 8f, rank 3.
"""
from __future__ import annotations

import torch
from torch import Tensor
from torch.autograd import Function

from . import _lib


def texture_sample(texture_info, texture_dims: Tensor, texture: Tensor, uvs: Tensor, use_torch_impl: bool = False) -> Tensor:
    """Samples `texture` (sum H*W, C) at `uvs` (Q,2) in the charts given per query by `texture_dims` (Q,3) =
    (height, width, start index); returns (Q, C).  `texture_info = (#charts, 1, C)`, only C is read.  As in the
    reference the autograd backward yields no gradients (texture_sample.py:77-86)."""
    if use_torch_impl:
        raise NotImplementedError("gstex_b200 has no PyTorch/CPU sampler: the CUDA kernel is the only path")
    return _TextureSample.apply(texture_info, texture_dims.contiguous(), texture.contiguous(), uvs.contiguous())


class _TextureSample(Function):
    """ref: texture_sample.py:50-86."""

    @staticmethod
    def forward(ctx, texture_info, texture_dims, texture, uvs) -> Tensor:
        dev = _lib.require_cuda(texture_dims, texture, uvs)
        _lib.require_dtype(texture_dims, torch.int32, "texture_dims")
        _lib.require_dtype(texture, torch.float32, "texture")
        _lib.require_dtype(uvs, torch.float32, "uvs")
        C = int(texture_info[2])
        if texture.dim() != 2 or texture.shape[1] != C:
            raise RuntimeError("texture must have dimensions (sum H*W, C) with C = texture_info[2]")
        num_queries = uvs.shape[0]
        if uvs.dim() != 2 or uvs.shape[1] != 2 or texture_dims.dim() != 2 or tuple(texture_dims.shape) != (num_queries, 3):
            raise RuntimeError("uvs must be (Q, 2) and texture_dims (Q, 3)")
        out = torch.empty((num_queries, C), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            rc = _lib.load().gstex_texture_sample_forward(num_queries, C, _lib.ptr(texture_dims), _lib.ptr(uvs),
                                                          _lib.ptr(texture), _lib.ptr(out), _lib.stream_ptr(dev))
        _lib.check(rc, "gstex_texture_sample_forward")
        ctx.texture_info = texture_info
        ctx.save_for_backward(texture_dims, texture, uvs)
        return out

    @staticmethod
    def backward(ctx, v_out_value):
        return None, None, None, None

"""Stages (c)+(d): `texture_gaussians`, the drop-in for the reference's `gstex_cuda/texture.py`.

Same positional signature (27 arguments), the same six outputs
`(out_img, out_depth, out_reg, out_alpha, out_texture, out_normal)` and the same autograd contract
(gradients for colors, opacity, means, scales, quats, uv0, umap, vmap, texture; `None` elsewhere) as
/root/reference/submodules/GStex_cuda/gstex_cuda/texture.py:14-408.
"""
from __future__ import annotations

import os
import weakref
from typing import Optional, Tuple

import torch
from torch import Tensor
from torch.autograd import Function

from . import _lib
from .utils import TileBinning, bin_gaussians_to_tiles, finish_binning

_SCRATCH_CLEAN = 4   # GSTEX_SCRATCH_CLEAN, include/gstex_b200.h
_DEFER_UNPACK = 8    # GSTEX_DEFER_UNPACK
_GRAD_SCRATCH = {}   # (device index, stream, bytes) -> all-zero gradient scratch left by the last backward

SUPPORTED_SETTINGS = (1 << 2) | (1 << 8) | (1 << 9) | (1 << 10) | (1 << 15)


class PairStreamPolicy:
    """Sizing of the pair stream (the forward's record of contributing (pixel, splat) pairs the stream backward
    reads, include/gstex_b200.h).  No counterpart in the reference, which always recomputes.  The pool is sized
    from the pairs the previous views actually produced (read back asynchronously, never synchronising); a pool
    that turns out too small only costs speed: the backward then recomputes and the next pool is larger.  The size moves
    in steps of 4 M pairs (128 MB) and only grows, unless 32 views in a row needed less than a quarter of it: views of one
    scene differ by tens of per cent in their pair counts, and a pool that followed every view would ask the caching
    allocator for a new size every time (a `cudaMalloc`, sometimes a synchronising `cudaFree`, in the middle of a
    step)."""

    enabled = os.environ.get("GSTEX_B200_PAIR_STREAM", "1") != "0"
    pairs_per_pixel = float(os.environ.get("GSTEX_B200_PAIRS_PER_PIXEL", "40"))
    headroom = 1.5
    quantum = 1 << 22  # pairs
    max_bytes = None  # optional hard cap of the pool in bytes (tests use it to provoke the overflow fallback)
    _state = {}  # (device index, H, W, block_width) -> {"pairs": int, "pending": (pinned tensor, event) | None}

    @classmethod
    def max_pairs(cls, dev, H, W, bw) -> int:
        st = cls._state.setdefault((dev.index, H, W, bw), {"pairs": int(cls.pairs_per_pixel * H * W), "pending": None})
        pend = st["pending"]
        if pend is not None and pend[1].query():
            hdr = pend[0]
            used, overflow = int(hdr[0]) * 128, int(hdr[1])
            st["last"] = (used, overflow)  # pairs' worth of chunks taken, overflow flag of the last reported view
            if overflow and pend[2]:
                want = st["pairs"]  # the pool was cut by `max_bytes`, not by this policy: nothing to learn
            else:
                want = 2 * st["pairs"] if overflow else int(cls.headroom * used)
            want = max((want + cls.quantum - 1) // cls.quantum, 1) * cls.quantum
            # every contributing pair has alpha >= 1/255, so a pixel is finished (T <= 1e-4) after at most
            # ln(1e4) / -ln(1 - 1/255) = 2345 of them: no view can need more than that per pixel
            want = min(want, max(2400 * H * W, cls.quantum))
            if want > st["pairs"]:
                st["pairs"], st["low"] = want, 0
            elif want <= st["pairs"] // 4:
                st["low"] = st.get("low", 0) + 1
                if st["low"] >= 32:
                    st["pairs"], st["low"] = want, 0
            else:
                st["low"] = 0
            st["pending"] = None
        return st["pairs"]

    @classmethod
    def report(cls, dev, H, W, bw, stream_buf, capped: bool = False) -> None:
        """Queue an async read-back of the stream header (chunks used, overflow flag).  `capped`: the pool was smaller
        than this policy asked for (`max_bytes`)."""
        st = cls._state.get((dev.index, H, W, bw))
        if st is None or st["pending"] is not None:
            return
        host = st.get("host")
        if host is None:
            host = st["host"] = torch.empty(4, dtype=torch.int32).pin_memory()
        host.copy_(stream_buf[:16].view(torch.int32), non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        st["pending"] = (host, ev, bool(capped))


class GradientTargets:
    """Opt-in fused gradient accumulation for multi-view steps (no counterpart in the reference).

    `GradientTargets.set(params, grads)` registers, for leaf tensors passed to `texture_gaussians` as colors / opacity /
    means / scales / quats / uv0 / umap / vmap / texture, the tensor their gradient is accumulated in (normally their
    `.grad`, e.g. the views of a `multiview.GradientBucket`).  When all nine inputs of a call are registered, the
    backward ADDS the view's gradients to those tensors inside its own unpack kernel (C ABI `zero_fill = 2`) and
    returns `None` for them, so autograd's AccumulateGrad -- nine extra read-modify-write passes over 208 MB per view
    at the headline size -- has nothing left to do.  The sums are bit-identical to autograd's.  `clear()` restores
    the plain behaviour."""

    _map = {}  # (data_ptr, shape) -> (weak reference to the registered parameter, its gradient target)
    # `defer`: the stream backward leaves each view's sums in the padded scratch (C ABI GSTEX_DEFER_UNPACK) and
    # `flush()` -- called by `GradientBucket.allreduce` / `zero_` -- moves them into the targets once per step, instead
    # of one sweep over all texel accumulators per view.
    defer = False
    _pending = {}  # scratch key -> (scratch, C, n, T, tex_scale, the six stream-path targets)

    @classmethod
    def flush(cls) -> None:
        """Move the sums deferred so far into their targets (on the current stream of each scratch's device)."""
        pend, cls._pending = cls._pending, {}
        for key, (scratch, C, n, T, tex_scale, tg) in pend.items():
            dev = scratch.device
            with torch.cuda.device(dev):
                rc = _lib.load().gstex_unpack_grads(C, n, T, tex_scale, _lib.ptr(scratch), scratch.numel(),
                                                    *[_lib.ptr(t) for t in tg], _lib.stream_ptr(dev))
            _lib.check(rc, "gstex_unpack_grads")

    @classmethod
    def set(cls, params, grads, defer: bool = False) -> None:
        cls.flush()
        cls.defer = bool(defer)
        m = {}
        for p, g in zip(params, grads):
            if g.shape != p.shape or g.dtype != torch.float32 or not g.is_contiguous() or g.device != p.device:
                raise ValueError("GradientTargets: each target must be a contiguous fp32 tensor shaped like its parameter")
            m[(p.data_ptr(), tuple(p.shape))] = (weakref.ref(p), g)
        cls._map = m

    @classmethod
    def clear(cls) -> None:
        cls.flush()
        cls._map = {}
        cls.defer = False

    @classmethod
    def lookup(cls, tensors):
        """The nine targets if every tensor is (an alias of) a registered parameter that is STILL ALIVE at the address
        it was registered with, else None.  A registration whose parameter has been freed or re-allocated (the model
        replaces its Parameters on densification / re-charting, NS/models/gstex.py:749,890-895, and the caching
        allocator recycles their addresses) is dropped on sight, so an unrelated tensor that happens to land on a stale
        address can never be mistaken for it."""
        if not cls._map:
            return None
        out = []
        for t in tensors:
            key = (t.data_ptr(), tuple(t.shape))
            ent = cls._map.get(key)
            if ent is None:
                return None
            owner = ent[0]()
            if owner is None or owner.data_ptr() != key[0] or tuple(owner.shape) != key[1] or owner.device != t.device:
                del cls._map[key]
                return None
            out.append(ent[1])
        return out


def texture_gaussians(
    texture_info: Tuple[int, int, int],
    texture_dims: Tensor,
    centers: Tensor,
    extents: Tensor,
    depths: Tensor,
    num_tiles_hit: Tensor,
    colors: Tensor,
    opacity: Tensor,
    means: Tensor,
    scales: Tensor,
    glob_scale,
    quats: Tensor,
    uv0: Tensor,
    umap: Tensor,
    vmap: Tensor,
    texture: Tensor,
    viewmat: Tensor,
    c2w: Tensor,
    fx: float,
    fy: float,
    cx: float,
    cy: float,
    img_height: int,
    img_width: int,
    block_width: int,
    settings: int,
    background: Optional[Tensor] = None,
    use_torch_impl: bool = False,
    *,
    binning: Optional[TileBinning] = None,
    texture_affine: Optional[Tuple[float, float]] = None,
    texture_padded: Optional[Tensor] = None,
):
    """Rasterizes textured 2D gaussians (ref: texture.py:14-150).

    `binning` (not in the reference): a request started earlier with `gstex_b200.utils.begin_binning` on the same
    `centers / extents / depths / num_tiles_hit`; the call then only waits for that request's sizes instead of
    synchronising the stream (multi-view pipelining, see `TileBinning`).

    `texture_affine=(scale, bias)` (not in the reference): `texture` holds RAW texels and every fetched sample is
    `scale * sample + bias`; the gradient returned for `texture` is with respect to the raw texels.  With
    `(C0, 0.5)` = `SH2RGB` this replaces the caller's `texture[:, 0:3] = SH2RGB(texture_dc)` (NS/models/gstex.py:
    1091-1121) -- two passes over all texels forward and two more backward per view -- by one multiply-add per
    fetched sample.

    `texture_padded` (not in the reference): `pad_texture(texture)`, the `(T, 4)` copy the forward then fetches with
    one 128-bit load per bilinear corner; the outputs and gradients are bit-identical with or without it.  The caller
    refreshes it whenever `texture` changes -- a stale copy renders the stale texels."""
    assert block_width > 1 and block_width <= 16, "block_width must be between 2 and 16"
    if use_torch_impl:
        raise NotImplementedError("gstex_b200 has no PyTorch/CPU rasterizer: the CUDA kernels are the only path")
    if colors.dtype == torch.uint8:
        colors = colors.float() / 255
    if background is not None:
        assert background.shape[0] == colors.shape[-1], (
            f"incorrect shape of background color tensor, expected shape {colors.shape[-1]}")
    else:
        background = torch.ones(colors.shape[-1], dtype=torch.float32, device=colors.device)
    if colors.ndimension() != 2:
        raise ValueError("colors must have dimensions (N, D)")
    if colors.shape[-1] != 3:
        raise ValueError("colors must have 3 channels (the reference kernels read float3, texture.cu:18)")
    if int(settings) & ~SUPPORTED_SETTINGS:
        raise NotImplementedError(
            f"settings {int(settings):#x}: only bits 2, 8, 9, 10, 15 are implemented (16-29 are viewer visualisation)")
    return _TextureGaussians.apply(
        texture_info, texture_dims.contiguous(), centers.contiguous(), extents.contiguous(), depths.contiguous(),
        num_tiles_hit.contiguous(), colors.contiguous(), opacity.contiguous(), means.contiguous(),
        scales.contiguous(), glob_scale, quats.contiguous(), uv0.contiguous(), umap.contiguous(), vmap.contiguous(),
        texture.contiguous(), viewmat.contiguous(), c2w.contiguous(), fx, fy, cx, cy, img_height, img_width,
        block_width, settings, background.contiguous(), binning,
        (1.0, 0.0) if texture_affine is None else (float(texture_affine[0]), float(texture_affine[1])),
        texture_padded)


def pad_texture(texture: Tensor, out: Optional[Tensor] = None) -> Tensor:
    """`(T, 4)` zero-padded copy of a `(T, C <= 3)` texture for `texture_gaussians(..., texture_padded=...)`: one pass
    over the texels (28 bytes each at C = 3), once per optimiser step for all the views of a batch.  Detached."""
    texture = texture.detach().contiguous()
    dev = _lib.require_cuda(texture)
    _lib.require_dtype(texture, torch.float32, "texture")
    if texture.dim() != 2 or not 1 <= texture.shape[1] <= 3:
        raise ValueError("pad_texture takes a (T, C) texture with 1 <= C <= 3")
    T = texture.shape[0]
    if out is None:
        out = torch.empty((T, 4), dtype=torch.float32, device=dev)
    elif tuple(out.shape) != (T, 4) or out.dtype != torch.float32 or out.device != dev or not out.is_contiguous():
        raise ValueError("out must be a contiguous float32 (T, 4) tensor on the texture's device")
    with torch.cuda.device(dev):
        rc = _lib.load().gstex_texture_pad4(T, texture.shape[1], _lib.ptr(texture), _lib.ptr(out), _lib.stream_ptr(dev))
    _lib.check(rc, "gstex_texture_pad4")
    return out


class PackedRecords:
    """The per-view packed records of the forward (`gstex_texture_pack_records`), made ahead of `texture_gaussians` --
    typically by `BinningStream.finish(pre, pack=...)` on the binning stream, beside the compositing kernels of the
    view before.  Holds the workspace and the identity (address, shape, version counter) of every tensor it was
    computed from; the forward uses it only if it is handed exactly those tensors, unchanged, and packs again
    otherwise.  "Unchanged" is judged by torch's version counters: an in-place write through `.data` (which bypasses
    them) between the pack and the forward goes unnoticed -- pack after such writes, not before."""

    __slots__ = ("workspace", "signature")

    @staticmethod
    def _sig(tensors, scalars):
        return tuple((t.data_ptr(), tuple(t.shape), t._version) for t in tensors) + tuple(scalars)

    def matches(self, tensors, scalars) -> bool:
        return self.signature == PackedRecords._sig(tensors, scalars)


def pack_view_records(texture_dims, tile_bins, colors, opacity, means, scales, glob_scale, quats, uv0, umap, vmap,
                      viewmat, c2w, fx, fy, cx, cy, img_height, img_width, block_width, settings) -> PackedRecords:
    """ref: the per-splat part of `texture_forward_kernel` (cuda/csrc/texture.cu:129-200), once per (view, splat)
    instead of once per (pixel, splat).  Runs on the current stream; see `PackedRecords`."""
    tensors = [t.detach() for t in (texture_dims, tile_bins, colors, opacity, means, scales, quats, uv0, umap, vmap,
                                    viewmat, c2w)]
    dev = _lib.require_cuda(*tensors)
    _lib.require_dtype(tensors[0], torch.int32, "texture_dims")
    _lib.require_dtype(tensors[1], torch.int32, "tile_bins")
    for t in tensors[2:]:
        _lib.require_dtype(t, torch.float32, "pack_view_records input")
    if colors.dim() != 2 or colors.shape[-1] != 3:
        raise ValueError("colors must be (N, 3)")
    n = means.shape[0]
    lib, p = _lib.load(), _lib.ptr
    rec = PackedRecords()
    rec.workspace = torch.empty((lib.gstex_texture_workspace_bytes(n),), dtype=torch.uint8, device=dev)
    scalars = (float(glob_scale), float(fx), float(fy), float(cx), float(cy), int(img_height), int(img_width),
               int(block_width), int(settings))
    with torch.cuda.device(dev):
        rc = lib.gstex_texture_pack_records(int(img_height), int(img_width), int(block_width), n, p(tensors[0]),
                                            p(tensors[1]), p(tensors[2]), p(tensors[3]), p(tensors[4]), p(tensors[5]),
                                            float(glob_scale), p(tensors[6]), p(tensors[7]), p(tensors[8]),
                                            p(tensors[9]), p(tensors[10]), p(tensors[11]), float(fx), float(fy),
                                            float(cx), float(cy), int(settings), p(rec.workspace),
                                            rec.workspace.numel(), _lib.stream_ptr(dev))
    _lib.check(rc, "gstex_texture_pack_records")
    rec.signature = PackedRecords._sig(tensors, scalars)
    return rec


def _raster_forward(texture_info, texture_dims, gaussian_ids_sorted, tile_bins, colors, opacity, means, scales,
                    glob_scale, quats, uv0, umap, vmap, texture, viewmat, c2w, fx, fy, cx, cy, img_height, img_width,
                    block_width, settings, background, num_intersects, record_pairs=False, texture_affine=(1.0, 0.0),
                    texture_padded=None, packed=None):
    """ref: `_C.texture_forward` (cuda/csrc/texture.cu:762-901).  Returns the 9 tensors + the workspace
    holding the per-view packed records (re-used by the backward)."""
    dev = _lib.require_cuda(texture_dims, colors, opacity, means, scales, quats, uv0, umap, vmap, texture, viewmat,
                            c2w, background, tile_bins)
    _lib.require_dtype(texture_dims, torch.int32, "texture_dims")
    for t, nm in ((colors, "colors"), (opacity, "opacity"), (means, "means"), (scales, "scales"), (quats, "quats"),
                  (uv0, "uv0"), (umap, "umap"), (vmap, "vmap"), (texture, "texture"), (viewmat, "viewmat"),
                  (c2w, "c2w"), (background, "background")):
        _lib.require_dtype(t, torch.float32, nm)
    H, W, C = int(img_height), int(img_width), int(texture_info[2])
    n = means.shape[0]
    f32 = dict(dtype=torch.float32, device=dev)
    i32 = dict(dtype=torch.int32, device=dev)
    out_img = torch.empty((H, W, 3), **f32)
    out_depth = torch.empty((H, W), **f32)
    out_reg = torch.empty((H, W), **f32)
    out_texture = torch.empty((H, W, C), **f32)
    out_normal = torch.empty((H, W, 3), **f32)
    final_Ts = torch.empty((H, W), **f32)
    final_idx = torch.empty((H, W), **i32)
    depth_idx = torch.empty((H, W), **i32)
    out_reg_s = torch.empty((H, W, 3), **f32)
    lib = _lib.load()
    if packed is not None and not packed.matches(
            (texture_dims, tile_bins, colors, opacity, means, scales, quats, uv0, umap, vmap, viewmat, c2w),
            (float(glob_scale), float(fx), float(fy), float(cx), float(cy), H, W, int(block_width), int(settings))):
        packed = None  # made from other tensors (or the tensors have changed since): pack again
    ws = (packed.workspace if packed is not None else
          torch.empty((lib.gstex_texture_workspace_bytes(n),), dtype=torch.uint8, device=dev))
    pairs = None
    if (record_pairs and PairStreamPolicy.enabled and num_intersects > 0
            and lib.gstex_pair_stream_supported(C, n, int(settings))):
        nbytes = lib.gstex_pair_stream_bytes(H, W, int(block_width), PairStreamPolicy.max_pairs(dev, H, W, int(block_width)))
        pool_capped = PairStreamPolicy.max_bytes is not None and nbytes > int(PairStreamPolicy.max_bytes)
        if pool_capped:
            nbytes = int(PairStreamPolicy.max_bytes)
        pairs = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
    p = _lib.ptr
    if texture_padded is not None:
        if (tuple(texture_padded.shape) != (texture.shape[0], 4) or C > 3 or texture_padded.dtype != torch.float32
                or texture_padded.device != dev or not texture_padded.is_contiguous()):
            raise ValueError("texture_padded must be pad_texture(texture): contiguous float32 (T, 4) on the same device")
    if packed is not None:
        fwd, tex_args = lib.gstex_texture_forward_packed, (p(texture), p(texture_padded))
    elif texture_padded is not None:
        fwd, tex_args = lib.gstex_texture_forward_padded, (p(texture), p(texture_padded))
    else:
        fwd, tex_args = lib.gstex_texture_forward_act, (p(texture),)
    with torch.cuda.device(dev):
        rc = fwd(
            H, W, int(block_width), C, n, int(num_intersects), p(texture_dims), p(gaussian_ids_sorted), p(tile_bins),
            p(colors), p(opacity), p(means), p(scales), float(glob_scale), p(quats), p(uv0), p(umap), p(vmap),
            *tex_args, float(texture_affine[0]), float(texture_affine[1]), p(viewmat), p(c2w), float(fx), float(fy),
            float(cx), float(cy), int(settings), p(background),
            p(out_img), p(out_depth), p(out_reg), p(out_texture), p(out_normal), p(final_Ts), p(final_idx),
            p(depth_idx), p(out_reg_s), p(ws), ws.numel(), p(pairs), pairs.numel() if pairs is not None else 0,
            _lib.stream_ptr(dev))
    _lib.check(rc, "gstex_texture_forward")
    if pairs is not None:
        PairStreamPolicy.report(dev, H, W, int(block_width), pairs, capped=pool_capped)
    return (out_img, out_depth, out_reg, out_texture, out_normal, final_Ts, final_idx, depth_idx, out_reg_s), ws, pairs


class _TextureGaussians(Function):
    """ref: texture.py:152-408."""

    @staticmethod
    def forward(ctx, texture_info, texture_dims, centers, extents, depths, num_tiles_hit, colors, opacity, means,
                scales, glob_scale, quats, uv0, umap, vmap, texture, viewmat, c2w, fx, fy, cx, cy, img_height,
                img_width, block_width, settings, background=None, binning=None, texture_affine=(1.0, 0.0),
                texture_padded=None):
        ctx.set_materialize_grads(False)  # unused outputs reach backward as None, not as zero images
        num_points = centers.size(0)
        tile_bounds = ((img_width + block_width - 1) // block_width, (img_height + block_width - 1) // block_width, 1)
        dev = centers.device
        # The reference leaves most outputs undefined without intersections (texture.py:197-205, a latent
        # NameError guarded by the model at NS/models/gstex.py:1127); we return a well-formed empty render.
        if binning is not None:
            num_intersects, gaussian_ids_sorted, tile_bins = finish_binning(binning)
        else:
            num_intersects, gaussian_ids_sorted, tile_bins = bin_gaussians_to_tiles(
                num_points, centers, extents, depths, num_tiles_hit, tile_bounds, block_width)
        outputs, ws, pairs = _raster_forward(texture_info, texture_dims, gaussian_ids_sorted, tile_bins, colors,
                                             opacity, means, scales, glob_scale, quats, uv0, umap, vmap, texture,
                                             viewmat, c2w, fx, fy, cx, cy, img_height, img_width, block_width,
                                             settings, background, max(num_intersects, 0),
                                             record_pairs=any(ctx.needs_input_grad), texture_affine=texture_affine,
                                             texture_padded=texture_padded,
                                             packed=getattr(binning, "records", None))
        out_img, out_depth, out_reg, out_texture, out_normal, final_Ts, final_idx, depth_idx, out_reg_s = outputs
        ctx.img_width, ctx.img_height = img_width, img_height
        ctx.texture_info = texture_info
        ctx.fx, ctx.fy, ctx.cx, ctx.cy = fx, fy, cx, cy
        ctx.glob_scale = glob_scale
        ctx.num_intersects = num_intersects
        ctx.block_width = block_width
        ctx.settings = settings
        ctx.texture_affine = texture_affine
        ctx.workspace = ws  # packed per-view records, valid as long as the inputs are unchanged
        ctx.pair_stream = pairs  # contributing pairs recorded by the forward (None: the backward recomputes)
        # fused accumulation (GradientTargets): only when every gradient is wanted and every input is registered
        ctx.grad_targets = (GradientTargets.lookup((colors, opacity, means, scales, quats, uv0, umap, vmap, texture))
                            if all(ctx.needs_input_grad[i] for i in (6, 7, 8, 9, 11, 12, 13, 14, 15)) else None)
        ctx.save_for_backward(texture_dims, gaussian_ids_sorted, tile_bins, colors, opacity, means, scales, quats,
                              uv0, umap, vmap, texture, viewmat, c2w, background, final_Ts, final_idx, depth_idx,
                              out_reg_s)
        out_alpha = 1 - final_Ts
        return out_img, out_depth, out_reg, out_alpha, out_texture, out_normal

    @staticmethod
    def backward(ctx, v_out_img, v_out_depth, v_out_reg, v_out_alpha, v_out_texture, v_out_normal):
        (texture_dims, gaussian_ids_sorted, tile_bins, colors, opacity, means, scales, quats, uv0, umap, vmap,
         texture, viewmat, c2w, background, final_Ts, final_idx, depth_idx, out_reg_s) = ctx.saved_tensors
        H, W = ctx.img_height, ctx.img_width
        C = int(ctx.texture_info[2])
        dev = means.device

        def grad_or_none(g, shape):
            # an output the loss does not use arrives as None (set_materialize_grads(False) in forward) and goes to
            # the library as a NULL pointer = zero: no zero image is allocated, filled and read per view
            if g is None:
                return None
            if tuple(g.shape) != shape:
                raise ValueError(f"upstream gradient of shape {tuple(g.shape)}, expected {shape}")
            return g.contiguous().float()

        v_out_img = grad_or_none(v_out_img, (H, W, 3))
        v_out_depth = grad_or_none(v_out_depth, (H, W))
        v_out_reg = grad_or_none(v_out_reg, (H, W))
        v_out_alpha = grad_or_none(v_out_alpha, (H, W))  # ref: texture.py:306-307
        v_out_texture = grad_or_none(v_out_texture, (H, W, C))
        v_out_normal = grad_or_none(v_out_normal, (H, W, 3))

        targets = ctx.grad_targets
        if targets is not None:
            (v_colors, v_opacity, v_means, v_scales, v_quats, v_uv0, v_umap, v_vmap, v_texture) = targets
        # uninitialised: the library defines every element (zero_fill=1), which saves nine memsets on the stream path
        else:
            v_colors = torch.empty_like(colors)
            v_opacity = torch.empty_like(opacity)
            v_means = torch.empty_like(means)
            v_scales = torch.empty_like(scales)
            v_quats = torch.empty_like(quats)
            v_uv0 = torch.empty_like(uv0)
            v_umap = torch.empty_like(umap)
            v_vmap = torch.empty_like(vmap)
            v_texture = torch.empty_like(texture)
        lib = _lib.load()
        ws = ctx.workspace
        p = _lib.ptr
        pairs = ctx.pair_stream
        # The gradient scratch leaves every backward all-zero (the unpack kernel clears what the stream kernel
        # wrote), so one buffer per (device, stream, size) is kept and its 200 MB clear paid once.  It is taken out
        # of the cache for the duration of the call: a call that raises never puts a half-written buffer back.
        scratch, scratch_key, clean, defer = None, None, 0, 0
        if pairs is not None:
            nbytes = lib.gstex_grad_scratch_bytes(means.shape[0], texture.shape[0])
            scratch_key = (dev.index, torch.cuda.current_stream(dev).cuda_stream, nbytes)
            want_defer = targets is not None and GradientTargets.defer
            pend = GradientTargets._pending.get(scratch_key)
            if pend is not None:
                # The scratch of this (device, stream, size) holds deferred sums.  They may only keep accumulating for
                # the same targets and activation slope; any other use of the scratch moves them out first.
                mine = (v_colors, v_opacity, v_means, v_scales, v_quats, v_texture)
                if not (want_defer and pend[4] == float(ctx.texture_affine[0]) and all(a is b for a, b in zip(pend[5], mine))):
                    GradientTargets.flush()
                    pend = None
            scratch = _GRAD_SCRATCH.pop(scratch_key, None)
            if scratch is None and pend is not None:
                scratch = pend[0]
            if scratch is None:
                scratch = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
            else:
                clean = _SCRATCH_CLEAN
            if want_defer:
                defer = _DEFER_UNPACK
        with torch.cuda.device(dev):
            rc = lib.gstex_texture_backward_act(
                int(H), int(W), int(ctx.block_width), C, means.shape[0], max(int(ctx.num_intersects), 0),
                p(texture_dims), p(gaussian_ids_sorted), p(tile_bins), p(colors), p(opacity), p(means), p(scales),
                float(ctx.glob_scale), p(quats), p(uv0), p(umap), p(vmap), p(texture), float(ctx.texture_affine[0]),
                float(ctx.texture_affine[1]), p(viewmat), p(c2w),
                float(ctx.fx), float(ctx.fy), float(ctx.cx), float(ctx.cy), int(ctx.settings), p(background),
                p(final_Ts), p(final_idx), p(depth_idx), p(out_reg_s), p(v_out_img), p(v_out_depth), p(v_out_reg),
                p(v_out_alpha), p(v_out_texture), p(v_out_normal), p(v_colors), p(v_opacity), p(v_means),
                p(v_scales), p(v_quats), p(v_uv0), p(v_umap), p(v_vmap), p(v_texture), p(ws), ws.numel(), 1,
                p(pairs), pairs.numel() if pairs is not None else 0, p(scratch),
                scratch.numel() if scratch is not None else 0, int(texture.shape[0]),
                (2 if targets is not None else 1) | clean | defer, _lib.stream_ptr(dev))
        _lib.check(rc, "gstex_texture_backward")
        if defer:
            GradientTargets._pending[scratch_key] = (scratch, C, means.shape[0], int(texture.shape[0]),
                                                     float(ctx.texture_affine[0]),
                                                     (v_colors, v_opacity, v_means, v_scales, v_quats, v_texture))
        if scratch is not None:
            if len(_GRAD_SCRATCH) >= 4:  # a handful of model sizes at most; do not hoard memory
                _GRAD_SCRATCH.clear()
            _GRAD_SCRATCH[scratch_key] = scratch
        if targets is not None:  # already added to the registered tensors
            return (None,) * 30
        return (None, None, None, None, None, None, v_colors, v_opacity, v_means, v_scales, None, v_quats, v_uv0,
                v_umap, v_vmap, v_texture, None, None, None, None, None, None, None, None, None, None, None, None, None,
                None)

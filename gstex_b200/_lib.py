"""ctypes binding of the C-ABI CUDA library (include/gstex_b200.h).

There is NO fallback: if `lib/libgstex_b200.so` cannot be loaded (or built from csrc/ with nvcc), every
entry point raises.  Nothing under `oracle/` is ever imported from here.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_float, c_int, c_size_t, c_void_p
from typing import Optional

import torch

from . import build as _build

_LIB: Optional[ctypes.CDLL] = None

P, F, I, SZ = c_void_p, c_float, c_int, c_size_t

_SIGNATURES = {
    "gstex_abi_version": (I, []),
    "gstex_error_string": (ctypes.c_char_p, [I]),
    "gstex_launch_count": (ctypes.c_longlong, []),
    "gstex_debug_force_nocull": (I, [I]),
    "gstex_debug_forward_variant": (I, [I]),
    "gstex_project_points": (I, [P, P, F, F, F, F, I, P, P, P]),
    "gstex_get_aabb_2d": (I, [P, P, F, P, P, F, F, F, F, I, P, P, P]),
    "gstex_get_num_tiles_hit_2d": (I, [P, P, I, I, I, I, P, P]),
    "gstex_project_bounds": (I, [P, P, F, P, P, F, F, F, F, I, I, I, I, P, P, P, P, P, P]),
    "gstex_cumsum_workspace_bytes": (SZ, [I]),
    "gstex_cumsum_i32": (I, [P, I, P, P, P, SZ, P]),
    "gstex_cumsum_noalloc_max_n": (I, []),
    "gstex_cumsum_i32_noalloc": (I, [P, I, P, P, P]),
    "gstex_map_gaussian_to_intersects": (I, [I, I, P, P, P, P, I, I, I, P, P, P]),
    "gstex_sort_workspace_bytes": (SZ, [I]),
    "gstex_sort_isects": (I, [I, P, P, P, P, I, P, SZ, P]),
    "gstex_get_tile_bin_edges": (I, [I, P, I, P, P]),
    "gstex_tile_bucket_max_list": (I, []),
    "gstex_tile_bucket_count": (I, [I, P, P, P, I, I, I, P, P, P, P, P]),
    "gstex_tile_bucket_max_tiles": (I, []),
    "gstex_tile_bucket_scratch_bytes": (SZ, [I]),
    "gstex_tile_bucket_sort": (I, [I, I, I, P, P, P, P, I, I, I, P, P, P, P, P, P]),
    "gstex_texture_workspace_bytes": (SZ, [I]),
    "gstex_pair_stream_bytes": (SZ, [I, I, I, ctypes.c_longlong]),
    "gstex_pair_stream_supported": (I, [I, I, I]),
    "gstex_texture_forward": (I, [I, I, I, I, I, I] + [P] * 7 + [F] + [P] * 7 + [F, F, F, F, I] + [P] * 10 +
                              [P, SZ, P, SZ, P]),
    "gstex_texture_backward": (I, [I, I, I, I, I, I] + [P] * 7 + [F] + [P] * 7 + [F, F, F, F, I] + [P] * 11 + [P] * 9 +
                               [P, SZ, I, P, SZ, P, SZ, ctypes.c_longlong, I, P]),
    "gstex_texture_forward_act": (I, [I, I, I, I, I, I] + [P] * 7 + [F] + [P] * 5 + [F, F] + [P] * 2 + [F, F, F, F, I] +
                                  [P] * 10 + [P, SZ, P, SZ, P]),
    "gstex_texture_forward_padded": (I, [I, I, I, I, I, I] + [P] * 7 + [F] + [P] * 6 + [F, F] + [P] * 2 + [F, F, F, F, I] +
                                     [P] * 10 + [P, SZ, P, SZ, P]),
    "gstex_peer_allreduce_max_world": (I, []),
    "gstex_peer_allreduce_f32": (I, [I, I, P, ctypes.c_longlong, P, I, P]),
    "gstex_texture_forward_packed": (I, [I, I, I, I, I, I] + [P] * 7 + [F] + [P] * 6 + [F, F] + [P] * 2 + [F, F, F, F, I] +
                                     [P] * 10 + [P, SZ, P, SZ, P]),
    "gstex_texture_pack_records": (I, [I, I, I, I] + [P] * 6 + [F] + [P] * 6 + [F, F, F, F, I] + [P, SZ, P]),
    "gstex_texture_pad4": (I, [ctypes.c_longlong, I, P, P, P]),
    "gstex_texture_backward_act": (I, [I, I, I, I, I, I] + [P] * 7 + [F] + [P] * 5 + [F, F] + [P] * 2 + [F, F, F, F, I] +
                                   [P] * 11 + [P] * 9 + [P, SZ, I, P, SZ, P, SZ, ctypes.c_longlong, I, P]),
    "gstex_grad_scratch_bytes": (SZ, [I, ctypes.c_longlong]),
    "gstex_unpack_grads": (I, [I, I, ctypes.c_longlong, F, P, SZ, P, P, P, P, P, P, P]),
    "gstex_compute_sh_forward": (I, [I, I, I, P, P, P, P]),
    "gstex_compute_sh_backward": (I, [I, I, I, P, P, P, P]),
    "gstex_texture_sample_forward": (I, [I, I, P, P, P, P, P]),
    "gstex_chart_max_candidates": (I, []),
    "gstex_chart_scores": (I, [I, P, F, I, P, ctypes.c_uint, P, P]),
    "gstex_chart_dims": (I, [I, P, F, F, P, P, P, P]),
    "gstex_chart_starts": (I, [I, P, P, P, P]),
    "gstex_texture_resample": (I, [ctypes.c_longlong, I, I, P, P, P, P, P]),
    "gstex_activations_forward": (I, [I] + [P] * 11),
    "gstex_activations_backward": (I, [I] + [P] * 12),
    "gstex_texture_edit": (I, [I, I, I, I, I, I, ctypes.c_longlong] + [P] * 10 + [F] + [P] * 6 + [F, F, F, F, I] + [P, P, SZ, P]),
    "gstex_photometric_loss_scratch_bytes": (SZ, [I, I]),
    "gstex_photometric_loss_forward": (I, [I, I, I, P, P, P, P, P, P, F, I, F, P, P, P, P, SZ, P]),
    "gstex_photometric_loss_backward": (I, [I, I, I, P, P, P, P, P, P, F, I, F, P, P, P, P, P, P]),
    "gstex_photometric_loss_forward_u8": (I, [I, I, I, P, P, P, P, P, P, F, I, F, P, P, P, P, SZ, P]),
    "gstex_photometric_loss_backward_u8": (I, [I, I, I, P, P, P, P, P, P, F, I, F, P, P, P, P, P, P]),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)


class GstexNativeError(RuntimeError):
    pass


class Timing:
    """Optional per-entry-point device timing (CUDA events on the launching stream around each C-ABI call).
    Used by bench.py for the live roofline numbers; off by default (zero overhead on the hot path)."""
    enabled = False
    records = {}

    @classmethod
    def start(cls):
        cls.records = {}
        cls.enabled = True

    @classmethod
    def stop(cls):
        """Returns {entry point: [ms per call, ...]} (synchronises)."""
        cls.enabled = False
        torch.cuda.synchronize()
        out = {k: [a.elapsed_time(b) for a, b in v] for k, v in cls.records.items()}
        cls.records = {}
        return out


class _TimedLib:
    """Thin proxy over the ctypes library: identical calls, optionally bracketed by CUDA events."""

    def __init__(self, cdll):
        object.__setattr__(self, "_cdll", cdll)

    def __getattr__(self, name):
        fn = getattr(self._cdll, name)
        if not name.startswith("gstex_") or name.endswith("_bytes") or name in ("gstex_abi_version", "gstex_error_string",
                                                                               "gstex_launch_count", "gstex_debug_force_nocull", "gstex_debug_forward_variant",
                                                                               "gstex_pair_stream_supported",
                                                                               "gstex_tile_bucket_max_list", "gstex_tile_bucket_max_tiles",
                                                                               "gstex_cumsum_noalloc_max_n", "gstex_chart_max_candidates",
                                                                               "gstex_peer_allreduce_max_world"):
            return fn

        def call(*args):
            if not Timing.enabled:
                return fn(*args)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rc = fn(*args)
            e1.record()
            Timing.records.setdefault(name, []).append((e0, e1))
            return rc

        return call


def lib_path() -> str:
    return _build.LIB_PATH


def load() -> ctypes.CDLL:
    """Load (building first if stale and nvcc is present) the CUDA library; raise if impossible."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = _build.LIB_PATH
    try:
        if _build.sources() and _build.is_stale():
            _build.build()
    except Exception as exc:  # nvcc missing on a box that only has the prebuilt .so is fine
        if not os.path.exists(path):
            raise GstexNativeError(f"gstex_b200: CUDA library missing and cannot be built: {exc}") from exc
    if not os.path.exists(path):
        raise GstexNativeError(f"gstex_b200: CUDA library not found at {path}; run `python -m gstex_b200.build`")
    lib = ctypes.CDLL(path)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export what the header declares
        fn.restype = res
        fn.argtypes = args
    if lib.gstex_abi_version() != 2:
        raise GstexNativeError("gstex_b200: ABI version mismatch between the Python layer and the CUDA library")
    _LIB = _TimedLib(lib)
    return _LIB


def check(code: int, what: str) -> None:
    if code != 0:
        msg = load().gstex_error_string(code)
        raise RuntimeError(f"{what} failed: {msg.decode() if msg else code} (code {code})")


def ptr(t: Optional[torch.Tensor]):
    if t is None:
        return None
    return c_void_p(t.data_ptr())


def stream_ptr(device: torch.device):
    return c_void_p(torch.cuda.current_stream(device).cuda_stream)


def require_cuda(*tensors: torch.Tensor) -> torch.device:
    """CHECK_INPUT of the reference (bindings.h:10-15): CUDA + contiguous, else RuntimeError."""
    dev = None
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError("gstex_b200: tensor must be a CUDA tensor (there is no CPU path)")
        if not t.is_contiguous():
            raise RuntimeError("gstex_b200: tensor must be contiguous")
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise RuntimeError("gstex_b200: all tensors must be on the same device")
    if dev is None:
        raise RuntimeError("gstex_b200: no tensor given")
    return dev


def require_dtype(t: torch.Tensor, dtype: torch.dtype, name: str) -> None:
    if t.dtype != dtype:
        raise RuntimeError(f"gstex_b200: {name} must be {dtype}, got {t.dtype}")

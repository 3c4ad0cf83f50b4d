"""CPU oracle of the per-splat activation stage -- TEST INFRASTRUCTURE, never imported by the product.

numpy restatement (fp32 arithmetic, one rounding per operation like eager torch) of what the reference model does with its
raw parameters before every view: /root/reference/nerfstudio/models/gstex.py:1062-1066 (quats / scales / opacities) and
get_uv_mapping, :975-990 (uv0, umap, vmap from the first two columns of R(F.normalize(quats)),
gstex_cuda/_torch_impl.py:116-138), plus the VJP torch.autograd derives for it.  Pinned to tests/golden/act_golden.npz,
which tests/golden/make_golden_act.py makes by executing the reference's own source text.
"""
from __future__ import annotations

import numpy as np

F = np.float32


def _norm4(q):
    """torch's CUDA norm(dim=-1) over four floats adds the rounded squares as (q0^2 + q2^2) + (q1^2 + q3^2); on the CPU
    the order differs in the last ulp only (the golden comparison allows for it)."""
    s = q * q
    return np.sqrt((s[:, 0] + s[:, 2]) + (s[:, 1] + s[:, 3]), dtype=F)[:, None]


def forward(quats_raw, scales_raw, opacities_raw, mappings=None):
    q = np.asarray(quats_raw, F)
    s = np.asarray(scales_raw, F)
    o = np.asarray(opacities_raw, F)
    quats = (q / _norm4(q)).astype(F)                                                    # gstex.py:1062
    scales = np.zeros_like(s)
    scales[:, :2] = np.maximum(np.exp(s[:, :2], dtype=F), F(1e-9))                        # :1063-1064
    scales[:, 2] = F(1e-5) * ((scales[:, 0] + scales[:, 1]) * F(0.5))                     # :1065
    opac = (F(1) / (F(1) + np.exp(-o, dtype=F))).astype(F)                                # :1066
    if mappings is None:
        return quats, scales, opac, None, None, None
    m = np.asarray(mappings, F)
    qn = quats / np.maximum(_norm4(quats), F(1e-12))                                      # F.normalize, _torch_impl.py:138
    w, x, y, z = qn[:, 0], qn[:, 1], qn[:, 2], qn[:, 3]
    ax1 = np.stack([1 - 2 * (y * y + z * z), 2 * (x * y + w * z), 2 * (x * z - w * y)], -1).astype(F)   # column 0 of R
    ax2 = np.stack([2 * (x * y - w * z), 1 - 2 * (x * x + z * z), 2 * (y * z + w * x)], -1).astype(F)   # column 1 of R
    n = q.shape[0]
    uv0 = np.full((n, 1, 2), 0.5, F)                                                      # gstex.py:976,986
    umap = (m[:, 0:1] * ax1)[:, None, :].astype(F)                                        # :983,987
    vmap = (m[:, 1:2] * ax2)[:, None, :].astype(F)                                        # :984,988
    return quats, scales, opac, uv0, umap, vmap


def backward(quats_raw, scales_raw, quats, scales, opacities, v_quats, v_scales, v_opacities):
    q, y = np.asarray(quats_raw, np.float64), np.asarray(quats, np.float64)
    nq = np.sqrt((q * q).sum(-1, keepdims=True))
    v = np.asarray(v_quats, np.float64)
    g_q = (v - y * (y * v).sum(-1, keepdims=True)) / nq
    e = np.exp(np.asarray(scales_raw, F)[:, :2], dtype=F)
    g_s = np.zeros(np.asarray(scales_raw).shape, np.float64)
    g_s[:, :2] = np.where(e >= F(1e-9), np.asarray(v_scales, np.float64)[:, :2] * np.asarray(scales, np.float64)[:, :2], 0.0)
    o = np.asarray(opacities, np.float64)
    g_o = np.asarray(v_opacities, np.float64) * (1.0 - o) * o
    return g_q.astype(F), g_s.astype(F), g_o.astype(F)

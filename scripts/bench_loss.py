#!/usr/bin/env python
"""Measurement of `photometric_loss` to the hot path's bar: the headline 800x800 view,
forward + backward, against the HBM roofline, with the reference's torch-op chain (+ autograd) on the same GPU and
the CPU oracle beside it.

    python scripts/bench_loss.py [--height 800] [--width 800] [--channels 3]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

import bench  # noqa: E402
from bench_sh import timed  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--height", type=int, default=800)
    ap.add_argument("--width", type=int, default=800)
    ap.add_argument("--channels", type=int, default=3)
    ap.add_argument("--iters", type=int, default=50)
    a = ap.parse_args()
    import gstex_b200 as G
    from gstex_b200 import _lib
    from test_gpu_loss import _scene, _torch_chain

    H, W, C = a.height, a.width, a.channels
    c = _scene(H, W, C, 7, [0, 0, 0], False)
    t = {k: torch.from_numpy(v).cuda() for k, v in c.items()}
    leaves = [t[k].clone().requires_grad_(True) for k in ("out_img", "out_texture", "out_alpha")]
    flush = torch.empty(64 << 20, device="cuda")

    def ours():
        for x in leaves:
            x.grad = None
        G.photometric_loss(*leaves, t["background"], t["gt_img"], 0.2)[1].backward()

    def torch_ops():
        for x in leaves:
            x.grad = None
        _torch_chain(*leaves, t["background"], t["gt_img"], 0.2)[3].backward()

    ms = timed(ours, a.iters, flush)
    _lib.Timing.start()
    for _ in range(10):
        flush.zero_()
        ours()
    per = {k: sum(v) / len(v) for k, v in _lib.Timing.stop().items()}
    torch.backends.cudnn.allow_tf32 = False
    ms_torch = timed(torch_ops, a.iters, flush)
    P = H * W
    alg_f = P * (4 * (10 + C) + 12 + 36)   # layers + gt read; rgb + three derivative maps written
    alg_b = P * (4 * (10 + C) + 36 + 4 * (4 + C))  # layers + gt + maps read; three gradients written
    peak, src = bench.measured_peak_gbs()
    kf, kb = per.get("gstex_photometric_loss_forward"), per.get("gstex_photometric_loss_backward")
    out = {"metric": "compose + L1 + SSIM loss fwd+bwd Mpix/s", "value": P / (ms * 1e-3) / 1e6, "unit": "Mpix/s", "dtype": "f32",
           "config": {"workload": f"{W}x{H} view, C={C}, ssim_lambda 0.2, through autograd (gstex_b200.photometric_loss)",
                      "l2": "256 MB buffer rewritten between timed calls"},
           "ms": ms, "per_entry_ms": per,
           "roofline": {"bound": "hbm", "kernel": "loss_forward_kernel + loss_backward_kernel",
                        "achieved": (alg_f + alg_b) / ((kf + kb) * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                        "frac": (alg_f + alg_b) / ((kf + kb) * 1e-3) / 1e9 / peak, "peak_source": src,
                        "alg_bytes": alg_f + alg_b, "forward_gbs": alg_f / (kf * 1e-3) / 1e9,
                        "backward_gbs": alg_b / (kb * 1e-3) / 1e9, "traffic": None},
           "torch_ops_same_gpu": {"ms": ms_torch, "value": P / (ms_torch * 1e-3) / 1e6, "unit": "Mpix/s",
                                  "what": "gstex.py:1204-1205,1301-1319 with pytorch_msssim's ops (F.conv2d) + autograd"}}
    import oracle

    t0 = time.perf_counter()
    oracle.photometric_loss(c["out_img"], c["out_texture"], c["out_alpha"], c["background"], c["gt_img"], 0.2)
    dt = time.perf_counter() - t0
    out["cpu_baseline"] = {"value": P / dt / 1e6, "unit": "Mpix/s", "cores": 1, "kind": "port",
                           "sample": f"the same view once, fwd+bwd, {dt * 1e3:.0f} ms"}
    print(json.dumps(out))


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Measurement of the SH colour stage to the hot path's bar: one JSON line with the
stage's throughput, its HBM roofline fraction (algorithmic bytes / CUDA-event time / measured copy peak), the CPU
port beside it and the reference's own CUDA kernels on the same GPU.

    python scripts/bench_sh.py [--n 1000000] [--degree 3] [--iters 50]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402


def timed(fn, iters, flush):
    for _ in range(3):
        fn()
    ms = []
    for _ in range(iters):
        flush.add_(1.0)  # > L2: evicts the stage's inputs between timed calls
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    ms.sort()
    return ms[len(ms) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1_000_000)
    ap.add_argument("--degree", type=int, default=3)
    ap.add_argument("--iters", type=int, default=50)
    a = ap.parse_args()
    import gstex_b200 as G
    from gstex_b200 import _lib

    dev = torch.device("cuda", 0)
    D = G.num_sh_bases(a.degree)
    g = torch.Generator(device=dev).manual_seed(1)
    v = torch.randn(a.n, 3, device=dev, generator=g)
    c = torch.randn(a.n, D, 3, device=dev, generator=g)
    w = torch.randn(a.n, 3, device=dev, generator=g)
    flush = torch.empty(64 << 20, device=dev)  # 256 MB
    lib = _lib.load()
    col = torch.empty(a.n, 3, device=dev)
    vc = torch.empty(a.n, D, 3, device=dev)
    p, st = _lib.ptr, _lib.stream_ptr(dev)
    fwd = lambda: lib.gstex_compute_sh_forward(a.n, a.degree, a.degree, p(v), p(c), p(col), st)
    bwd = lambda: lib.gstex_compute_sh_backward(a.n, a.degree, a.degree, p(v), p(w), p(vc), st)
    ms_f, ms_b = timed(fwd, a.iters, flush), timed(bwd, a.iters, flush)
    bytes_f, bytes_b = a.n * (12 + 12 * D + 12), a.n * (24 + 12 * D)
    peak, src = bench.measured_peak_gbs()
    out = {"metric": "SH colour stage fwd+bwd Mpoints/s", "value": a.n / ((ms_f + ms_b) * 1e-3) / 1e6, "unit": "Mpoints/s",
           "config": {"workload": f"{a.n} points, degree {a.degree} ({D} bases), fwd + coefficient VJP",
                      "l2": "256 MB buffer rewritten between timed calls"},
           "dtype": "f32", "ms_forward": ms_f, "ms_backward": ms_b,
           "roofline": {"bound": "hbm", "kernel": "sh_forward_kernel + sh_backward_kernel",
                        "achieved": (bytes_f + bytes_b) / ((ms_f + ms_b) * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                        "frac": (bytes_f + bytes_b) / ((ms_f + ms_b) * 1e-3) / 1e9 / peak, "peak_source": src,
                        "forward": {"alg_bytes": bytes_f, "gbs": bytes_f / (ms_f * 1e-3) / 1e9},
                        "backward": {"alg_bytes": bytes_b, "gbs": bytes_b / (ms_b * 1e-3) / 1e9}, "traffic": None}}
    # CPU port (oracle), all host threads
    import oracle

    vn, cn, wn = v.cpu().numpy(), c.cpu().numpy(), w.cpu().numpy()
    t0 = time.perf_counter()
    oracle.spherical_harmonics(a.degree, vn, cn)
    oracle.spherical_harmonics_vjp(a.degree, D, vn, wn)
    dt = time.perf_counter() - t0
    out["cpu_baseline"] = {"value": a.n / dt / 1e6, "unit": "Mpoints/s", "cores": oracle.num_threads(), "kind": "port",
                           "sample": f"the same {a.n} points once, {dt * 1e3:.0f} ms"}
    try:
        from oracle import ref_ext

        C = ref_ext.load()
        if C is not None:
            rf = timed(lambda: C.compute_sh_forward(a.n, a.degree, a.degree, v, c), a.iters, flush)
            rb = timed(lambda: C.compute_sh_backward(a.n, a.degree, a.degree, v, w), a.iters, flush)
            out["reference_cuda_same_gpu"] = {"ms_forward": rf, "ms_backward": rb, "value": a.n / ((rf + rb) * 1e-3) / 1e6,
                                              "unit": "Mpoints/s", "what": "reference sh.cuh kernels incl. their output allocation"}
    except Exception as exc:
        out["reference_cuda_same_gpu"] = {"error": str(exc)[:200]}
    print(json.dumps(out))


if __name__ == "__main__":
    main()

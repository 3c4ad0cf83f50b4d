#!/usr/bin/env python
"""Measurement of `texture_sample` to the hot path's bar: re-charting workload -- one
query per texel of a 1e7-texel, 3-channel jagged texture (random chart per query block, random uv).

    python scripts/bench_sample.py [--texels 10000000] [--charts 1000000] [--channels 3]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))
import torch  # noqa: E402

import bench  # noqa: E402
from bench_sh import timed  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--texels", type=int, default=10_000_000)
    ap.add_argument("--charts", type=int, default=1_000_000)
    ap.add_argument("--channels", type=int, default=3)
    ap.add_argument("--iters", type=int, default=30)
    a = ap.parse_args()
    import gstex_b200 as G

    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(4)
    side = max(1, int(round((a.texels / a.charts) ** 0.5)))
    h = torch.randint(max(1, side - 1), side + 2, (a.charts,), device=dev, generator=g, dtype=torch.int32)
    w = torch.randint(max(1, side - 1), side + 2, (a.charts,), device=dev, generator=g, dtype=torch.int32)
    hw = (h * w).long()
    start = (torch.cumsum(hw, 0) - hw).int()
    T = int(hw.sum())
    tex = torch.rand(T, a.channels, device=dev, generator=g)
    chart = torch.repeat_interleave(torch.arange(a.charts, device=dev), hw)  # query t samples its own chart
    qd = torch.stack([h[chart], w[chart], start[chart]], -1).contiguous()
    uv = torch.rand(T, 2, device=dev, generator=g)
    info = (a.charts, 1, a.channels)
    flush = torch.empty(64 << 20, device=dev)
    ms = timed(lambda: G.texture_sample(info, qd, tex, uv), a.iters, flush)
    alg = T * (12 + 8 + 4 * a.channels) + T * 4 * a.channels  # dims + uv + out per query, every texel read once
    peak, src = bench.measured_peak_gbs()
    out = {"metric": "texture_sample Mqueries/s", "value": T / (ms * 1e-3) / 1e6, "unit": "Mqueries/s", "dtype": "f32",
           "config": {"workload": f"{T} queries (one per texel, own chart, random uv), {a.charts} charts, C={a.channels}",
                      "l2": "256 MB buffer rewritten between timed calls"},
           "ms": ms,
           "roofline": {"bound": "hbm", "kernel": "texture_sample_kernel", "achieved": alg / (ms * 1e-3) / 1e9, "peak": peak,
                        "unit": "GB/s", "frac": alg / (ms * 1e-3) / 1e9 / peak, "peak_source": src, "alg_bytes": alg,
                        "traffic": None}}
    import oracle

    nq = min(T, 2_000_000)
    qn, tn, un = qd[:nq].cpu().numpy(), tex.cpu().numpy(), uv[:nq].cpu().numpy()
    t0 = time.perf_counter()
    oracle.texture_sample(info, qn, tn, un)
    dt = time.perf_counter() - t0
    out["cpu_baseline"] = {"value": nq / dt / 1e6, "unit": "Mqueries/s", "cores": oracle.num_threads(), "kind": "port",
                           "sample": f"the first {nq} queries once, {dt * 1e3:.0f} ms"}
    try:
        from oracle import ref_ext

        C = ref_ext.load()
        if C is not None:
            rms = timed(lambda: C.texture_sample_forward(info, qd, uv, tex), a.iters, flush)
            out["reference_cuda_same_gpu"] = {"ms": rms, "value": T / (rms * 1e-3) / 1e6, "unit": "Mqueries/s",
                                              "what": "reference texture_sample_forward incl. its output allocation"}
    except Exception as exc:
        out["reference_cuda_same_gpu"] = {"error": str(exc)[:200]}
    print(json.dumps(out))


if __name__ == "__main__":
    main()

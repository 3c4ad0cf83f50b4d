#!/usr/bin/env python
"""Measurement of `texture_edit` to the hot path's bar, on the headline scene
(cfg5: 1M splats, 1e7 texels, 800x800): the model's edit sequence (NS/models/gstex.py:528-605) -- a render of the view,
a +-1e-2 window around its median depth, and the edit of a canvas painted over `--painted` of the image.

    python scripts/bench_edit.py [--config cfg5] [--painted 0.25] [--iters 20]
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
    ap.add_argument("--config", default="cfg5")
    ap.add_argument("--painted", type=float, default=0.25, help="share of the image area that is painted (alpha > 0)")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--no-cpu", action="store_true")
    a = ap.parse_args()
    import gstex_b200 as G
    from gstex_b200 import _lib
    from gstex_b200.synthetic import make_scene

    dev = torch.device("cuda", 0)
    s = make_scene(a.config)
    sd = {k: (x.to(dev) if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
    intr = (s["fx"], s["fy"], s["cx"], s["cy"])
    H, W, N = s["H"], s["W"], s["N"]
    _, depths = G.project_points(sd["means"], sd["viewmat"][:3, :], intr)
    centers, extents = G.get_aabb_2d(sd["means"], sd["scales"], 1, sd["quats"], sd["viewmat"][:3, :], intr)
    nth = G.get_num_tiles_hit_2d(centers, extents, H, W, 16)
    outs = G.texture_gaussians(s["texture_info"], sd["texture_dims"], centers, extents, depths, nth, sd["colors"],
                               sd["opacities"], sd["means"], sd["scales"], 1, sd["quats"], sd["uv0"], sd["umap"],
                               sd["vmap"], sd["texture"], sd["viewmat"], sd["c2w"], *intr, H, W, 16, s["settings"],
                               torch.zeros(3, device=dev))
    depth = outs[1]
    lo, hi = depth - 1e-2, depth + 1e-2
    g = torch.Generator(device=dev).manual_seed(11)
    img = torch.rand(H, W, 3, device=dev, generator=g)
    alpha = torch.zeros(H, W, 1, device=dev)
    side = a.painted ** 0.5
    h0, w0 = int(H * (1 - side) / 2), int(W * (1 - side) / 2)
    h1, w1 = h0 + max(1, int(H * side)), w0 + max(1, int(W * side))
    alpha[h0:h1, w0:w1] = 0.05 + 0.95 * torch.rand(h1 - h0, w1 - w0, 1, device=dev, generator=g)
    info = (N, 1, 5)
    bg = torch.zeros(3, device=dev)

    def ours():
        return G.texture_edit(info, sd["texture_dims"], img, alpha, lo, hi, centers, extents, depths, nth, sd["opacities"],
                              sd["means"], sd["scales"], 1, sd["quats"], sd["uv0"], sd["umap"], sd["vmap"],
                              sd["viewmat"][:3, :], sd["c2w"], *intr, H, W, 16, 1 << 13, background=bg)

    out = ours()
    pairs = float(out[:, 4].double().sum())  # every pair in its window carries weight one
    T = out.shape[0]
    flush = torch.empty(64 << 20, device=dev)
    ms = timed(ours, a.iters, flush)
    _lib.Timing.start()
    for _ in range(5):
        ours()
    per_entry = {k: sum(v) / len(v) for k, v in _lib.Timing.stop().items()}
    kernel_ms = per_entry.get("gstex_texture_edit", float("nan"))
    I = int(nth.sum())
    # algorithmic bytes of the edit entry point: zero-fill of the result (T*5*4 written), ids (4 I), splat inputs read by
    # the packer (N * 88), the canvas + window (P * 24), and one read-modify-write of every touched texel row
    # (<= 4 corners * 20 B * 2 per pair; counted as 40 B per touched texel)
    touched = int((out[:, 4] > 0).sum())
    alg = T * 20 + 4 * I + 88 * N + 24 * H * W + 40 * touched
    peak, src = bench.measured_peak_gbs()
    res = {"metric": "texture_edit Mpix/s (binning + edit of one 800x800 view)", "value": H * W / (ms * 1e-3) / 1e6,
           "unit": "Mpix/s", "dtype": "f32", "ms": ms, "per_entry_ms": per_entry,
           "config": {"workload": f"{a.config}: {N} splats, {T} texels, {W}x{H}, window = median depth +-1e-2, "
                                  f"{a.painted:.0%} of the image painted, C=5", "intersections": I,
                      "pairs_in_window": pairs, "touched_texels": touched,
                      "l2": "256 MB buffer rewritten between timed calls"},
           "roofline": {"bound": "hbm", "kernel": "gstex_texture_edit (memset + pack_records_kernel + texture_edit_kernel)",
                        "achieved": alg / (kernel_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                        "frac": alg / (kernel_ms * 1e-3) / 1e9 / peak, "peak_source": src, "alg_bytes": alg,
                        "avg_launch_ms": kernel_ms, "traffic": None}}
    if not a.no_cpu:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle
        from common import run_oracle_edit, scene_to_numpy

        d = scene_to_numpy(s)
        e = dict(updated_img=img.cpu().numpy(), updated_alpha=alpha.cpu().numpy(), depth_lower=lo.cpu().numpy(),
                 depth_upper=hi.cpu().numpy())
        t0 = time.perf_counter()
        ref = run_oracle_edit(d, e, 1 << 13)
        dt = time.perf_counter() - t0
        res["cpu_baseline"] = {"value": H * W / dt / 1e6, "unit": "Mpix/s", "cores": oracle.num_threads(), "kind": "port",
                               "sample": f"the same view once incl. projection + binning + sort, {dt:.2f} s"}
        o = out.cpu().numpy()
        scale = float(abs(ref).max())
        res["max_abs_err_vs_oracle_over_scale"] = float(abs(o - ref).max() / scale)
    try:
        from oracle import ref_ext

        C = ref_ext.load()
        if C is not None:
            tb = ((W + 15) // 16, (H + 15) // 16, 1)
            vm3, c2w = sd["viewmat"][:3, :].contiguous(), sd["c2w"].contiguous()

            def ref_call():  # texture_edit.py:169-236 around the reference's native calls
                total = int(torch.sum(sd["texture_dims"][:, 0] * sd["texture_dims"][:, 1]).item())
                cum = torch.cumsum(nth, dim=0, dtype=torch.int32)
                ni = int(cum[-1].item())
                _, _, _, gids_s, bins = ref_ext.ref_bin_and_sort(C, N, ni, centers, extents, depths, cum, tb, 16)
                return C.texture_edit(tb, (16, 16, 1), (W, H, 1), info, total, sd["texture_dims"], img, alpha, lo, hi,
                                      gids_s, bins, sd["opacities"], sd["means"], sd["scales"], 1.0, sd["quats"],
                                      sd["uv0"], sd["umap"], sd["vmap"], vm3, c2w, *intr, 1 << 13, bg)

            r = ref_call()
            rms = timed(ref_call, a.iters, flush)
            res["reference_cuda_same_gpu"] = {"ms": rms, "value": H * W / (rms * 1e-3) / 1e6, "unit": "Mpix/s",
                                              "what": "reference map/sort/bin-edges + texture_edit kernel, unmodified",
                                              "max_abs_diff_over_scale": float((out - r).abs().max() / r.abs().max())}
    except Exception as exc:
        res["reference_cuda_same_gpu"] = {"error": str(exc)[:300]}
    print(json.dumps(res))


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""bench.py -- headline benchmark of the GStex hot path on B200 (contract: see the task brief).

    python bench.py --gpus 1 --steps K --warmup W            # our CUDA path
    python bench.py --impl reference --steps K --warmup W    # the reference algorithm on the host cores

Workload (BASELINE.json: "fwd+bwd ms/iter & Mpix/s, 1M textured splats @800^2"): the cfg5 synthetic scene --
1M textured 2D Gaussians, 1e7 texels (C=3), 800x800 views, settings (1<<9)|(1<<10) as in training.  One
"step" = one multi-view training step of the hot path: for each of `views_per_gpu` camera views
project/bounds/tile-count -> scan -> key emit -> radix sort -> tile ranges -> textured forward -> backward
with fixed upstream gradients, parameter gradients accumulated over the views; with N>1 GPUs every rank
renders its own views (weak scaling) and the splat+texel gradients are summed with one NCCL all-reduce.
`value` = Mpix/s = views x 800 x 800 / time over all ranks, inputs resident in HBM.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from gstex_b200.multiview import GradientBucket, shard_views  # noqa: E402

METRIC = "fwd+bwd Mpix/s, 1M textured splats @800x800 (project+bin+sort+forward+backward)"
UNIT = "Mpix/s"
OWN_BINNING_STREAM = True  # --binning-stream own|current
TEXEL_LOADS_128 = True     # --texel-loads scalar|128
PACK_EARLY = False         # --pack early|forward
ALLREDUCE = "auto"         # --allreduce auto|peer|multimem|nccl (N > 1)
PARAM_KEYS = ["colors", "opacities", "means", "scales", "quats", "uv0", "umap", "vmap", "texture"]
VOUT_KEYS = ["v_out_img", "v_out_depth", "v_out_reg", "v_out_alpha", "v_out_texture", "v_out_normal"]


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="cfg5")
    ap.add_argument("--views-per-gpu", type=int, default=8)
    ap.add_argument("--n", type=int, default=None, help="override splat count (debug)")
    ap.add_argument("--size", type=int, default=None, help="override image size (debug)")
    ap.add_argument("--no-ref-cuda", action="store_true", help="skip timing the reference CUDA extension")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--binning-stream", default="own", choices=["own", "current"],
                    help="next view's projection + binning on a stream of their own (default) or on the current stream (A/B)")
    ap.add_argument("--allreduce", default="auto", choices=["auto", "peer", "multimem", "nccl"],
                    help="N > 1: the gradient sum at the end of a step as our kernel over NVLink peer memory -- plain peer "
                         "loads/stores, the NVSwitch-multicast variant, or whichever the library picks for N (default) -- "
                         "or NCCL's all-reduce (A/B)")
    ap.add_argument("--pack", default="forward", choices=["early", "forward"],
                    help="forward: the per-view records are packed inside the forward call, on the current stream (default); "
                         "early: on the binning stream, ahead of the forward (A/B: measured 3 %% slower per step, DESIGN.md 3)")
    ap.add_argument("--texel-loads", default="128", choices=["scalar", "128"],
                    help="128: the texture is padded to (T, 4) once per step (inside the timed region) and every view's "
                         "forward fetches texels with 128-bit loads (gstex_b200.pad_texture)")
    return ap.parse_args()


def build_scene(args, views):
    from gstex_b200.synthetic import make_scene, make_upstream_grads

    kw = {}
    if args.n:
        kw["n"] = args.n
    if args.size:
        kw["width"] = kw["height"] = args.size
    s = make_scene(args.config, views=views, **kw)
    v = make_upstream_grads(s)
    return s, v


def alg_bytes(N, I, T, C, P, tiles):
    """Algorithmic bytes per view: every datum counted once per stage that must touch it."""
    b = {
        "project_bounds": 64 * N,
        "scan_emit": 24 * N + 12 * I,
        "sort": 24 * I,
        "tile_bins": 8 * I + 8 * tiles,
        "forward": 4 * I + 100 * N + 4 * T * C + P * (56 + 4 * C),
        "backward": 4 * I + 100 * N + 4 * T * C + P * (60 + 4 * C) + 88 * N + 8 * T * C,
    }
    b["total"] = sum(b.values())
    return b


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.proc = None
        self.lines = []
        self.first = 0  # samples before this one were taken before the timed region

    def start(self):
        # the recipe's interval (200 ms); BENCH_CLOCKS=smi<ms>|off is a diagnosis knob (step outliers of 10-80 ms were
        # seen once per ~100 steps with and without the sampler: scripts/gpu_clk.sh)
        mode = os.environ.get("BENCH_CLOCKS", "smi200")
        if mode == "off":
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                                          str(self.idx), "-lms", mode[3:]], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def mark(self):
        """The timed region starts here.  nvidia-smi is started before the warm-up steps, so that its start-up (NVML
        attaches to every GPU of the box) is over by now; only the samples taken from now on are reported."""
        self.first = len(self.lines)

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return None
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines[self.first:] or self.lines[-1:]:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return None
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


class Workload:
    """Parameters resident on the device (they are the model state, like weights), a list of local camera
    views, fixed upstream gradients; `step()` runs the whole hot path forward+backward over the views."""

    def __init__(self, s, v, dev, local_view_ids):
        import torch

        self.torch = torch
        self.s, self.dev = s, dev
        self.params = {k: s[k].to(dev).detach().clone().requires_grad_(True) for k in PARAM_KEYS}
        self.texture_dims = s["texture_dims"].to(dev)
        self.background = s["background"].to(dev)
        self.viewmats = [s["viewmats"][i].contiguous().to(dev) for i in local_view_ids]
        self.c2ws = [s["c2ws"][i].contiguous().to(dev) for i in local_view_ids]
        self.v = {k: x.to(dev) for k, x in v.items()}
        self.intr = (s["fx"], s["fy"], s["cx"], s["cy"])
        self.H, self.W, self.bw = s["H"], s["W"], 16
        self.last_I = 0
        self.feeder = None
        self.binning = None  # BinningStream: the next view's projection + binning beside this view's compositing
        self.bucket = None  # one flat gradient buffer: fused accumulation, one all-reduce per step
        self.reduce_events = None  # list while the resident leg is timed
        self.padded = None  # --texel-loads 128: (T, 4) copy of the texture, refreshed at the start of every step

    def prep_view(self, viewmat, first=False):
        """Projection + the asynchronous half of the binning of one view (no host synchronisation), on the binning
        stream (gstex_b200.multiview.BinningStream): it runs beside the compositing kernels of the view before.
        Only the first view of a step orders itself after the current stream (the parameters may have changed)."""
        if self.binning is None:
            from gstex_b200.multiview import BinningStream

            self.binning = BinningStream(self.dev, own_stream=OWN_BINNING_STREAM)
        p = self.params
        return self.binning.begin(p["means"], p["scales"], 1, p["quats"], viewmat, self.intr, self.H, self.W, self.bw,
                                  wait_for_current=first)

    def finish_view(self, pre, c2w):
        """The host waits for the view's sizes; emit, sort and (--pack early) the forward's per-view records go to the
        binning stream."""
        p = self.params
        pack = None
        if PACK_EARLY:
            pack = dict(texture_dims=self.texture_dims, colors=p["colors"], opacity=p["opacities"], uv0=p["uv0"],
                        umap=p["umap"], vmap=p["vmap"], c2w=c2w, settings=self.s["settings"])
        self.binning.finish(pre, pack=pack)

    def forward_view(self, pre, viewmat, c2w):
        """The reference's call (NS/models/gstex.py:1075-1170) on a prepared view."""
        import gstex_b200 as G

        p, s = self.params, self.s
        centers, extents, depths, nth, binning = *pre.args, pre.binning
        return G.texture_gaussians(s["texture_info"], self.texture_dims, centers, extents, depths, nth, p["colors"],
                                   p["opacities"], p["means"], p["scales"], 1, p["quats"], p["uv0"], p["umap"],
                                   p["vmap"], p["texture"], viewmat, c2w, *self.intr, self.H, self.W, self.bw,
                                   s["settings"], self.background, binning=binning, texture_padded=self.padded)

    def backward_view(self, outs, v):
        self.torch.autograd.backward(list(outs), [v["v_out_img"], v["v_out_depth"], v["v_out_reg"], v["v_out_alpha"],
                                                  v["v_out_texture"], v["v_out_normal"]])

    def zero_grad(self, world=1):
        """Gradients of all nine parameter tensors live in one flat bucket; the backward of every view adds into it
        (GradientTargets), the step ends with one all-reduce when there are several ranks."""
        if self.bucket is None:
            group = None
            if world > 1 and ALLREDUCE != "nccl":
                import torch.distributed as dist

                group = dist.group.WORLD
            self.bucket = GradientBucket(list(self.params.values()), peer_group=group, peer_fallback=True,
                                         multicast={"auto": None, "peer": False, "multimem": True}.get(ALLREDUCE))
            if self.bucket.peer is not None and os.environ.get("BENCH_PEER_CTAS"):  # experiment knob
                self.bucket.peer.max_ctas = int(os.environ["BENCH_PEER_CTAS"])
            self.bucket.register()
        self.bucket.zero_()
        if TEXEL_LOADS_128:  # the parameters change between steps: the padded copy is part of every step's work
            import gstex_b200 as G

            self.padded = G.pad_texture(self.params["texture"], out=self.padded)

    def reduce_grads(self, world):
        if self.reduce_events is not None:  # the resident leg times the exchange step (flush + sum over the ranks)
            e0, e1 = self.torch.cuda.Event(enable_timing=True), self.torch.cuda.Event(enable_timing=True)
            e0.record()
            self.bucket.allreduce(world)
            e1.record()
            self.reduce_events.append((e0, e1))
            return
        self.bucket.allreduce(world)

    def step(self, world):
        """All local views, fwd+bwd through the public API.  View i+1's projection and binning counts are launched
        between view i's forward and backward, so the host reads view i+1's `num_intersects` while the backward runs
        (the reference's blocking `.item()` per view would drain the GPU each time)."""
        self.zero_grad(world)
        nv = len(self.viewmats)
        pre = self.prep_view(self.viewmats[0], first=True)
        self.finish_view(pre, self.c2ws[0])
        for i in range(nv):
            outs = self.forward_view(pre, self.viewmats[i], self.c2ws[i])
            pre = self.prep_view(self.viewmats[i + 1]) if i + 1 < nv else None
            self.backward_view(outs, self.v)
            if pre is not None:
                self.finish_view(pre, self.c2ws[i + 1])
        self.reduce_grads(world)

    def _e2e_loop(self, world, host, view_in, backward, results):
        """Common skeleton of the two end-to-end legs: every view's HOST inputs are prefetched one view ahead on a copy
        stream (`HostViewFeeder`), its results go back behind the compute, the host waits once at the end of the step."""
        if self.feeder is None:
            from gstex_b200.multiview import HostViewFeeder

            self.feeder = HostViewFeeder(self.dev)
        fd, nv = self.feeder, len(self.viewmats)
        self.zero_grad(world)
        if self.binning is None:
            from gstex_b200.multiview import BinningStream

            self.binning = BinningStream(self.dev, own_stream=OWN_BINNING_STREAM)
        side = (self.binning.stream,)  # reads each view's camera matrix
        fd.prefetch(view_in(0))
        d = fd.take(also=side)
        pre = self.prep_view(d["vm"], first=True)
        self.finish_view(pre, d["c2w"])
        if nv > 1:
            fd.prefetch(view_in(1))
        for i in range(nv):
            outs = self.forward_view(pre, d["vm"], d["c2w"])
            d_next = fd.take(also=side) if i + 1 < nv else None
            pre = self.prep_view(d_next["vm"]) if d_next is not None else None
            res = backward(outs, d)
            if pre is not None:
                self.finish_view(pre, d_next["c2w"])
            if i + 2 < nv:  # view i's staging slot is free once its backward is enqueued
                fd.prefetch(view_in(i + 2))
            for dev_t, host_t in results(i, outs, res):
                fd.send_back(dev_t, host_t)
            d = d_next
        self.reduce_grads(world)
        fd.finish()  # the host consumes the result every step

    def step_e2e(self, world, host):
        """The training step as the reference's trainer runs it (NS/pipelines/base_pipeline.py:291-304): per view the
        camera and the uint8 ground-truth image come from pinned HOST memory (full_images_datamanager.py:316-333), the
        image is rendered, composited and compared on the device (`gstex_b200.photometric_loss` = gstex.py:1204-1205,
        1297-1319), its backward feeds `texture_gaussians`' backward, and the three loss values go back to the host."""
        import gstex_b200 as G

        def backward(outs, d):
            # the uint8 image goes in as it is: the loss kernels convert it as get_gt_img does (NS/models/gstex.py:1238-1247)
            _, main, l1, ssim = G.photometric_loss(outs[0], outs[4], outs[3], self.background, d["gt"], 0.2)
            main.backward()
            return self.torch.stack([main.detach(), l1, ssim])

        self._e2e_loop(world, host, lambda i: {"vm": host["viewmats"][i], "c2w": host["c2ws"][i], "gt": host["gt"][i]},
                       backward, lambda i, outs, res: [(res, host["losses"][i])])

    def step_e2e_upstream(self, world, host):
        """Round-1 leg, kept for continuity: fixed upstream-gradient images (six fp32 images, 30.7 MB per view) from
        pinned host memory instead of a ground-truth image; rendered image + a scalar back to the host."""
        def backward(outs, d):
            self.backward_view(outs, d["v"])

        self._e2e_loop(world, host, lambda i: {"vm": host["viewmats"][i], "c2w": host["c2ws"][i], "v": host["v"]},
                       backward, lambda i, outs, res: [(outs[0].detach(), host["out_img"][i]),
                                                       (outs[3].detach().sum(), host["scalar"][i])])

    def make_host(self):
        torch = self.torch
        nv = len(self.viewmats)
        g = torch.Generator().manual_seed(4242)
        host = {
            "viewmats": [m.cpu().pin_memory() for m in self.viewmats],
            "c2ws": [m.cpu().pin_memory() for m in self.c2ws],
            "v": {k: x.cpu().pin_memory() for k, x in self.v.items()},
            "gt": [torch.randint(0, 256, (self.H, self.W, 3), dtype=torch.uint8, generator=g).pin_memory() for _ in range(nv)],
            "out_img": [torch.empty(self.H, self.W, 3).pin_memory() for _ in range(nv)],
            "scalar": [torch.empty(()).pin_memory() for _ in range(nv)],
            "losses": [torch.empty(3).pin_memory() for _ in range(nv)],
        }
        cam = 2 * 16 * 4
        bytes_ = {"loss": (nv * (cam + self.H * self.W * 3), nv * 12),
                  "upstream": (nv * (cam + sum(x.numel() * 4 for x in host["v"].values())), nv * (self.H * self.W * 3 * 4 + 4))}
        return host, bytes_

    def stats(self):
        """Intersections and depth complexity of view 0 (data dependent; reported with every number)."""
        import gstex_b200 as G

        p = self.params
        with self.torch.no_grad():
            _, _, c, e, nth = G.project_bounds(p["means"], p["scales"], 1, p["quats"], self.viewmats[0], self.intr, self.H,
                                               self.W, self.bw)
        return int(nth.sum().item())


def host_threads():
    """Host threads this process may run on (cgroup/affinity aware)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_reference_view(s, v, threads=None):
    """One view (view 0) fwd+bwd of the same scene through the CPU restatement of the reference algorithm
    (oracle/, OpenMP over tiles).  Returns (seconds, num_intersects, threads)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle
    from common import run_oracle_pipeline, scene_to_numpy

    # torchrun exports OMP_NUM_THREADS=1 to every rank: the CPU arm sets its own thread count, always
    oracle.set_num_threads(threads or host_threads())
    d = scene_to_numpy(s)
    vn = {k: x.numpy() for k, x in v.items()}
    t0 = time.perf_counter()
    r = run_oracle_pipeline(d, vn)
    return time.perf_counter() - t0, int(r["num_intersects"]), oracle.num_threads()


def run_reference_arm(args):
    """--impl reference: the reference algorithm on the host cores (the reference's own CPU path is Python
    that cannot travel to the GPU box and its CUDA path is the thing being replaced, so this is the C port
    of it in oracle/ -- kind "port"), all host threads, one view per step as the bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    s, v = build_scene(args, 1)
    P = s["H"] * s["W"]
    times, I, threads = [], 0, 0
    for i in range(args.warmup + args.steps):
        dt, I, threads = cpu_reference_view(s, v)
        if i >= args.warmup:
            times.append(dt)
    t = sum(times) / len(times)
    val = P / t / 1e6
    sample = f"1 view ({s['W']}x{s['H']}) of the {s['N']}-splat scene per step, fwd+bwd incl. binning+sort, I={I}"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.config}: {s['N']} splats, {s['T']} texels, {s['W']}x{s['H']}, 1 view/step on CPU"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def time_reference_cuda(wl, s):
    """The reference's own CUDA kernels (oracle/_ref, built unmodified for sm_100) on the same inputs and the
    same GPU: fwd+bwd of view 0, median of 5.  Reported beside our numbers; None if the .so is absent."""
    import torch

    from oracle import ref_ext

    C = ref_ext.load()
    if C is None:
        return None
    import gstex_b200 as G

    sd = {k: (x.to(wl.dev) if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
    sd["viewmat"], sd["c2w"] = wl.viewmats[0], wl.c2ws[0]
    vm3 = wl.viewmats[0][:3, :].contiguous()
    times = []
    for it in range(7):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _, depths = ref_ext.ref_project_points(sd["means"], vm3, wl.intr)
        c, e = C.get_aabb_2d(sd["means"], sd["scales"], 1.0, sd["quats"], vm3, *wl.intr)
        nth = ref_ext.ref_get_num_tiles_hit_2d(c, e, wl.H, wl.W, 16).to(torch.int32)
        ref_ext.ref_forward_backward(C, sd, c, e, depths.contiguous(), nth, wl.v)
        e1.record()
        torch.cuda.synchronize()
        if it >= 2:
            times.append(e0.elapsed_time(e1))
    return statistics.median(times)


def stats3(xs):
    return {"min": min(xs), "median": statistics.median(xs), "max": max(xs)}


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, STREAM-style copy)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def kernel_source_stamp():
    """sha256 over the CUDA sources of the hot path (every kernel the bench's timed region launches): ties
    profiles/traffic.json (written by scripts/ncu_traffic.py from an ncu capture) to the build it was captured from -- the
    GPU box has no .git to ask.  The sources of the next-row kernels the timed region never launches are left out."""
    import glob
    import hashlib

    other = {"loss.cu", "sh.cu", "texture_edit.cu", "texture_sample.cu", "activations.cu", "charts.cu", "peer_reduce.cu"}
    h = hashlib.sha256()
    for f in sorted(glob.glob(os.path.join(ROOT, "gstex_b200", "csrc", "*.cu*"))):
        if os.path.basename(f) in other:
            continue
        h.update(os.path.basename(f).encode())
        h.update(open(f, "rb").read())
    return h.hexdigest()[:16]


def ncu_capture(kernel_key):
    """{"dram_bytes", "inst"} per launch of the entry point's dominant kernel from the committed ncu capture, or None
    when profiles/traffic.json was captured from other kernel sources than the ones being timed."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            d = json.load(f)
        if d.get("source_stamp") != kernel_source_stamp():
            return None
        return d.get(kernel_key)
    except Exception:
        return None


def main():
    global OWN_BINNING_STREAM, TEXEL_LOADS_128, ALLREDUCE, PACK_EARLY
    args = parse_args()
    # second stream at every N: measured on one 8xB200 box (profiles/bench_r2_s1_n8_{own,current}.json) 3594 vs 3467 Mpix/s at
    # 8 ranks, 498 vs 470 at one (round 1 needed a switch at 8 ranks; gone with the cheaper per-view host path)
    OWN_BINNING_STREAM = args.binning_stream == "own"
    TEXEL_LOADS_128 = args.texel_loads == "128"
    ALLREDUCE = args.allreduce
    PACK_EARLY = args.pack == "early"
    # stdout carries ONE JSON line and nothing else: libraries that log to file descriptor 1 (NCCL's version banner and
    # NCCL_DEBUG output go to stdout) are sent to stderr for the whole run, the line is written to the saved descriptor
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w", buffering=1)
    if args.impl == "reference":
        run_reference_arm(args)
        return
    import torch

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU path in the product)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    from gstex_b200 import _lib

    lib = _lib.load()
    total_views = args.views_per_gpu * world
    s, v = build_scene(args, total_views)
    wl = Workload(s, v, dev, shard_views(total_views, rank, world))
    N, T, C, H, W = s["N"], s["T"], s["C"], s["H"], s["W"]
    P = H * W
    tiles = ((W + 15) // 16) * ((H + 15) // 16)

    def timed(step_fn, before=None, first=None):
        """W warm-up steps, then EXACTLY K steps between barrier + synchronize on both sides; one CUDA event per step
        boundary.  Returns (ms per step from the first to the last event, max over ranks; per-step ms of this rank)."""
        if first is not None:
            first()
        for _ in range(max(args.warmup, 3 if args.steps > 1 else 1)):
            step_fn()
        barrier()
        if before is not None:
            before()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
        ms0 = torch.cuda.memory_stats(dev)
        ev[0].record()
        grow = []  # (step, cudaMalloc calls, bytes reserved) whenever the allocator went to the driver (diagnosis only)
        dump = bool(os.environ.get("BENCH_DUMP_STEPS"))
        for k in range(args.steps):
            step_fn()
            ev[k + 1].record()
            if dump:
                m = torch.cuda.memory_stats(dev)
                if m["num_device_alloc"] != (grow[-1][1] if grow else ms0["num_device_alloc"]):
                    grow.append((k, m["num_device_alloc"], m["reserved_bytes.all.current"]))
        barrier()
        per = [ev[k].elapsed_time(ev[k + 1]) for k in range(args.steps)]
        if os.environ.get("BENCH_DUMP_STEPS"):  # diagnosis: the per-step times of this rank, allocator traffic to the driver
            ms1 = torch.cuda.memory_stats(dev)
            print("per-step ms:", " ".join(f"{x:.2f}" for x in per), file=sys.stderr)
            print("cudaMalloc / cudaFree calls inside the timed region:", ms1["num_device_alloc"] - ms0["num_device_alloc"],
                  ms1["num_device_free"] - ms0["num_device_free"], "reserved MB at start", ms0["reserved_bytes.all.current"] >> 20,
                  "growth (step, calls, reserved MB):", [(k, c, r >> 20) for k, c, r in grow], file=sys.stderr)
        return max_over_ranks(ev[0].elapsed_time(ev[-1])) / args.steps, per

    clocks = ClockSampler(local_rank)
    launches0 = [0]

    def begin_timed_region():
        if rank == 0:
            clocks.mark()
        launches0[0] = lib.gstex_launch_count()
        _lib.Timing.start()

    def begin_resident():
        begin_timed_region()
        wl.reduce_events = []

    ms_per_step, per_step = timed(lambda: wl.step(world), before=begin_resident,
                                  first=clocks.start if rank == 0 else None)
    per_call = _lib.Timing.stop()
    reduce_ms = [a.elapsed_time(b) for a, b in wl.reduce_events]
    wl.reduce_events = None
    launches = lib.gstex_launch_count() - launches0[0]
    clk = clocks.stop() if rank == 0 else None
    value = total_views * P / (ms_per_step * 1e-3) / 1e6

    # ---- end to end through the public API with host buffers
    host, hb = wl.make_host()
    t0 = time.perf_counter()
    e2e_ms, e2e_per = timed(lambda: wl.step_e2e(world, host))
    wall = time.perf_counter() - t0
    e2e_value = total_views * P / (e2e_ms * 1e-3) / 1e6
    e2e_up_ms, _ = timed(lambda: wl.step_e2e_upstream(world, host))
    h2d, d2h = hb["loss"]

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (live CUDA-event durations of the timed region)
    I = wl.stats()
    ab = alg_bytes(N, I, T, C, P, tiles)
    entry_ms = {k: (sum(x) / len(x), len(x), sum(x)) for k, x in per_call.items() if x}
    dom = max(entry_ms, key=lambda k: entry_ms[k][2])
    # the `_act` / `_padded` entry points run the same kernels (scale 1, bias 0; texels from the (T,4) copy)
    base = dom[:-4] if dom.endswith("_act") else dom[:-7] if dom.endswith(("_padded", "_packed")) else dom
    key = {"gstex_texture_backward": "backward", "gstex_texture_forward": "forward", "gstex_sort_isects": "sort"}.get(base, "total")
    peak, peak_src = measured_peak_gbs()
    achieved = ab[key] / (entry_ms[dom][0] * 1e-3) / 1e9
    cap = ncu_capture(base) or {}
    step_ms_kernels = sum(x[2] for x in entry_ms.values()) / args.steps
    roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "frac_of_nominal_8tbs": achieved / 8000.0,  # both denominators are reported
                "traffic": cap.get("dram_bytes"), "peak_source": peak_src, "alg_bytes_per_launch": ab[key],
                # instruction-issue roofline of the same kernel: warp-instructions (ncu smsp__inst_executed.sum of the capture
                # this build was profiled with) / (SMs x 4 schedulers x SM clock) / the kernel's own duration under ncu
                "issue_frac": cap.get("issue_frac"), "inst_per_launch": cap.get("inst"),
                "traffic_source": "profiles/traffic.json (ncu capture of this build)" if cap else
                                  "none: profiles/traffic.json was captured from different kernel sources",
                "avg_launch_ms": entry_ms[dom][0],
                "share_of_step": entry_ms[dom][2] / args.steps / ms_per_step,
                "whole_path": {"alg_bytes_per_view": ab["total"],
                               "achieved_gbs": ab["total"] * args.views_per_gpu / (ms_per_step * 1e-3) / 1e9,
                               "frac": ab["total"] * args.views_per_gpu / (ms_per_step * 1e-3) / 1e9 / peak,
                               "frac_of_nominal_8tbs": ab["total"] * args.views_per_gpu / (ms_per_step * 1e-3) / 1e9 / 8000.0},
                "per_entry_ms_per_view": {k: x[2] / args.steps / args.views_per_gpu for k, x in entry_ms.items()},
                "streams": ("2: the next view's projection + binning (gstex_b200.multiview.BinningStream) run beside this "
                            "view's forward/backward, so the per-entry CUDA-event times overlap, include the wait for SMs "
                            "and sum to more than the step") if OWN_BINNING_STREAM else "1 (--binning-stream current)"}

    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": f"{args.config}: {N} splats, {T} texels (C={C}), {W}x{H}, {args.views_per_gpu} views/GPU/step, "
                               f"fwd+bwd, settings {s['settings']}", "views_per_gpu": args.views_per_gpu,
                   "intersections_view0": I, "mean_tile_list": I / tiles, "ms_per_view": ms_per_step / args.views_per_gpu,
                   "pack": "per-view records packed on the binning stream (BinningStream.finish(pack=...))" if PACK_EARLY else "inside the forward call",
                   "texel_loads": "128-bit from a (T,4) copy padded once per step inside the timed region" if TEXEL_LOADS_128 else "scalar",
                   "l2": "working set (texture+texel grads+records ~400 MB) exceeds the 126 MB L2; no explicit flush",
                   "parallelism": (f"views sharded over {world} GPU(s); the flat gradient bucket (9 tensors, "
                                   f"{wl.bucket.flat.numel() * 4 / 1e6:.0f} MB) is summed over the ranks by "
                                   + ("one NCCL all-reduce" if wl.bucket.peer is None else
                                      "one kernel of ours over NVLink peer memory (gstex_peer_allreduce_f32"
                                      + (", NVSwitch multimem variant)" if wl.bucket.peer.multicast_ptr else ")")))
                   if world > 1 else "single GPU"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": e2e_ms, "step_ms": stats3(e2e_per),
                "what": "per view: camera + uint8 ground-truth image from pinned host memory, render, on-device compose + "
                        "L1/SSIM loss (gstex_b200.photometric_loss) and its backward into texture_gaussians' backward, the "
                        "three loss values back to the host (copy streams one view ahead/behind the compute: gstex_b200."
                        "multiview.HostViewFeeder); splat/texture parameters are model state resident in HBM"},
        "e2e_upstream_grads": {"value": total_views * P / (e2e_up_ms * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": e2e_up_ms,
                               "h2d_bytes_per_step": hb["upstream"][0], "d2h_bytes_per_step": hb["upstream"][1],
                               "what": "round-1 leg: six fp32 upstream-gradient images per view from the host, rendered image back"},
        "step_ms": stats3(per_step),
        # the exchange step of rank 0: deferred-gradient flush (unpack sweep) + the sum over the ranks, CUDA events
        "exchange_ms": stats3(reduce_ms) if reduce_ms else None,
        "gpu_launches": int(launches),
        "kernel_ms_per_step": step_ms_kernels,
        "roofline": roofline,
        "clocks": clk,
    }
    if not args.no_cpu_baseline:  # rank 0, every N (about 1 s on 16 host threads)
        dt, Icpu, threads = cpu_reference_view(s, v)
        out["cpu_baseline"] = {"value": P / dt / 1e6, "unit": UNIT, "cores": threads, "kind": "port",
                               "sample": f"1 view of the same scene, fwd+bwd incl. binning+sort, {dt:.1f} s, I={Icpu}"}
    if world == 1 and not args.no_ref_cuda:
        try:
            ms = time_reference_cuda(wl, s)
            if ms is not None:
                out["reference_cuda_same_gpu"] = {"ms_per_view": ms, "value": P / (ms * 1e-3) / 1e6, "unit": UNIT,
                                                  "what": "reference CUDA kernels + torch cumsum/sort, unmodified, sm_100 build"}
        except Exception as exc:  # the checker must never break the bench
            out["reference_cuda_same_gpu"] = {"error": str(exc)[:200]}
        try:  # parity of view 0 against the reference's kernels at the benchmark's own size (outside the timed region)
            from gstex_b200.texture import GradientTargets
            from oracle import parity

            GradientTargets.clear()
            rep = parity.compare_view(s, v, dev, view=0)
            if rep is not None:
                out["parity"] = {"img_max_rel": rep["img_max_rel"], "grad_max_rel": rep["grad_max_rel"],
                                 "idx_mismatch": rep["idx_mismatch"], "bin_mismatch": rep["bin_mismatch"],
                                 "fwd_bit_mismatch": rep["fwd_bit_mismatch"],
                                 "outliers": rep["outliers"], "exact": rep["exact"],
                                 "what": "view 0 vs the reference's own CUDA kernels on the same tensors: max |err| / tensor "
                                         "scale over the six images / nine gradients, share of elements outside 1e-5 / 1e-4, "
                                         "mismatching elements of sorted ids + tile bins + final_idx + depth_idx (idx) and of "
                                         "depths + centers + extents + tile counts (bin), elements of the six images that are not "
                                         "bit-identical (fwd_bit); idx, bin and fwd_bit must be 0"}
        except Exception as exc:
            out["parity"] = {"error": str(exc)[:200]}
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

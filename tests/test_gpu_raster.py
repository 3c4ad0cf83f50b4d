"""GPU parity, stages (c)+(d): forward compositing and backward, through the public
`texture_gaussians` (-> C ABI).  Checked against the CPU oracle, the golden vectors of the reference's
torch rasterizer, and the reference's own CUDA extension when oracle/_ref holds it.
Tolerances (north_star): images 1e-5 relative, gradients 1e-4 relative, indices exact."""
import numpy as np
import pytest
import torch

from common import (FWD_KEYS, GRAD_KEYS, GRAD_RTOL, IMG_RTOL, VOUT_KEYS, golden_files, load_golden,
                    run_oracle_pipeline, scene_to_numpy)

pytestmark = pytest.mark.gpu

# discrete decisions (alpha < 1/255, T <= 1e-4, T > 0.5, blur floor) can flip for isolated pixels between two
# fp32 implementations; such pixels are counted and bounded, everything else must meet the tolerance
MAX_FLIP_FRAC = 2e-4

SCENES = [
    dict(name="cfg1", n=400, width=72, height=56, texels=400 * 16, settings=1 << 8, C=3),
    dict(name="cfg1", n=400, width=72, height=56, texels=400 * 16, settings=0, C=3),
    dict(name="cfg2", n=6000, width=160, height=160, texels=60000, settings=(1 << 9) | (1 << 10), C=3),
    dict(name="cfg2", n=6000, width=160, height=160, texels=0, settings=(1 << 9) | (1 << 10), C=3),  # 2DGS 1x1 charts
    dict(name="cfg2", n=3000, width=96, height=128, texels=30000, settings=(1 << 8) | (1 << 9) | (1 << 10), C=6),
    dict(name="cfg2", n=3000, width=96, height=128, texels=30000, settings=(1 << 2) | (1 << 9), C=3),  # nearest
    dict(name="cfg4", n=20000, width=240, height=136, texels=200000, settings=(1 << 9) | (1 << 10), C=3),
    # settings bit 15, the viewer's visualization_mode: normals flipped towards the camera (forward only)
    dict(name="cfg2", n=6000, width=160, height=160, texels=60000, settings=(1 << 9) | (1 << 10) | (1 << 15), C=6),
    dict(name="cfg1", n=400, width=72, height=56, texels=400 * 16, settings=(1 << 15), C=3),
    dict(name="cfg1", n=300, width=40, height=40, texels=300 * 9, settings=1 << 9, C=10),
]


def _id(c):
    return f"{c['name']}-n{c['n']}-{c['width']}x{c['height']}-s{c['settings']}-C{c['C']}-T{c['texels']}"


@pytest.fixture(params=["stream", "recompute", "overflow"], autouse=True)
def backward_path(request):
    """Every test of this file runs three times: with the forward recording the pair stream (stream backward),
    with recording off (recompute backward), and with a pool so small that it overflows (device-side fallback)."""
    from gstex_b200.texture import PairStreamPolicy

    old = (PairStreamPolicy.enabled, PairStreamPolicy.max_bytes)
    PairStreamPolicy.enabled = request.param != "recompute"
    PairStreamPolicy.max_bytes = 40000 if request.param == "overflow" else None
    yield request.param
    PairStreamPolicy.enabled, PairStreamPolicy.max_bytes = old


def run_ours(s, v=None, block_width=16):
    """The reference's call sequence (example.py:145-184 / NS/models/gstex.py:1075-1170) on our package."""
    import gstex_b200 as G

    sd = {k: (x.cuda() if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
    intr = (s["fx"], s["fy"], s["cx"], s["cy"])
    H, W = s["H"], s["W"]
    leaves = {}
    for k in ["colors", "opacities", "means", "scales", "quats", "uv0", "umap", "vmap", "texture"]:
        leaves[k] = sd[k].clone().requires_grad_(v is not None)
    _, depths = G.project_points(leaves["means"].detach(), sd["viewmat"][:3, :], intr)
    centers, extents = G.get_aabb_2d(leaves["means"].detach(), leaves["scales"].detach(), 1, leaves["quats"].detach(),
                                     sd["viewmat"][:3, :], intr)
    nth = G.get_num_tiles_hit_2d(centers, extents, H, W, block_width)
    outs = G.texture_gaussians(s["texture_info"], sd["texture_dims"], centers, extents, depths, nth, leaves["colors"],
                               leaves["opacities"], leaves["means"], leaves["scales"], 1, leaves["quats"],
                               leaves["uv0"], leaves["umap"], leaves["vmap"], leaves["texture"], sd["viewmat"],
                               sd["c2w"], *intr, H, W, block_width, s["settings"], sd["background"])
    res = dict(zip(FWD_KEYS, outs))
    if v is not None:
        vd = {k: x.cuda() for k, x in v.items()}
        loss = sum((res[k] * vd["v_" + k]).sum() for k in FWD_KEYS)
        loss.backward()
        for k, gk in zip(["colors", "opacities", "means", "scales", "quats", "uv0", "umap", "vmap", "texture"], GRAD_KEYS):
            res[gk] = leaves[k].grad
    res.update(centers=centers, extents=extents, depths=depths, num_tiles_hit=nth)
    return {k: (x.detach().cpu().numpy() if isinstance(x, torch.Tensor) else x) for k, x in res.items()}, sd


# Tolerances.  Against the reference's own CUDA extension (same GPU, same compiler's FMA contraction): the
# north-star bars, 1e-5 images / 1e-4 gradients.  Against the CPU oracle (gcc, no contraction, libm exp2f):
# one decade looser, because `delta = origin + t*ray - mean` cancels ~|mean|/|delta| ~ 1e2-1e3 and turns the
# 1-ulp differences between two fp32 evaluation orders into ~1e-5 relative differences of sigma and alpha.
ORACLE_FWD_RTOL, ORACLE_GRAD_RTOL = 1e-4, 1e-3


def compare(ours, ref, settings, fwd_rtol=IMG_RTOL, grad_rtol=GRAD_RTOL, what=""):
    from common import compare_all

    return compare_all(ours, ref, settings, fwd_rtol, grad_rtol, what, MAX_FLIP_FRAC)


@pytest.mark.parametrize("cfg", SCENES, ids=_id)
def test_forward_backward_matches_oracle(cfg):
    from gstex_b200.synthetic import make_scene, make_upstream_grads

    s = make_scene(cfg["name"], n=cfg["n"], width=cfg["width"], height=cfg["height"], texels=cfg["texels"],
                   settings=cfg["settings"], channels=cfg["C"])
    v = make_upstream_grads(s)
    ours, _ = run_ours(s, v)
    d = scene_to_numpy(s)
    ref = run_oracle_pipeline(d, {k: x.numpy() for k, x in v.items()})
    compare(ours, ref, cfg["settings"], ORACLE_FWD_RTOL, ORACLE_GRAD_RTOL, what="oracle:")


@pytest.mark.parametrize("cfg", SCENES[:9], ids=_id)
def test_forward_backward_matches_reference_extension(cfg):
    from oracle import ref_ext

    C = ref_ext.load()
    if C is None:
        pytest.skip("oracle/_ref/gstex_ref_C.so not built (python oracle/build_ref.py)")
    from gstex_b200.synthetic import make_scene, make_upstream_grads

    s = make_scene(cfg["name"], n=cfg["n"], width=cfg["width"], height=cfg["height"], texels=cfg["texels"],
                   settings=cfg["settings"], channels=cfg["C"])
    v = make_upstream_grads(s)
    ours, sd = run_ours(s, v)
    vd = {k: x.cuda() for k, x in v.items()}
    t = lambda k: torch.from_numpy(ours[k]).cuda()  # noqa: E731
    ref = ref_ext.ref_forward_backward(C, sd, t("centers"), t("extents"), t("depths"), t("num_tiles_hit"), vd)
    ref = {k: (x.detach().cpu().numpy() if isinstance(x, torch.Tensor) else x) for k, x in ref.items()}
    compare(ours, ref, cfg["settings"], what="refext:")


@pytest.mark.parametrize("path", golden_files(), ids=lambda p: p.split("golden_")[-1][:-4])
def test_matches_golden_of_reference_torch_rasterizer(path):
    d = load_golden(path)
    s = {k: torch.from_numpy(d[k]) for k in ["means", "scales", "quats", "opacities", "colors", "texture_dims",
                                             "texture", "uv0", "umap", "vmap", "viewmat", "c2w", "background"]}
    s.update(fx=d["fx"], fy=d["fy"], cx=d["cx"], cy=d["cy"], H=d["H"], W=d["W"], settings=d["settings"], C=d["C"],
             texture_info=(len(d["means"]), 1, d["C"]))
    v = {k: torch.from_numpy(d[k]) for k in VOUT_KEYS}
    ours, _ = run_ours(s, v)
    ref = {k: d["ref_" + k] for k in ["out_img", "out_depth", "out_reg", "out_texture", "out_normal"] + GRAD_KEYS}
    ref["out_alpha"] = 1.0 - d["ref_final_Ts"]
    # the torch rasterizer ran on the CPU without FMA contraction: oracle-class tolerances (see above)
    compare(ours, ref, d["settings"], ORACLE_FWD_RTOL, ORACLE_GRAD_RTOL, what="golden:")


@pytest.mark.parametrize("bw", [2, 5, 8, 13])
def test_block_widths(bw):
    from gstex_b200.synthetic import make_scene, make_upstream_grads

    s = make_scene("cfg2", n=1500, width=50, height=38, texels=15000, settings=(1 << 9) | (1 << 10))
    s["block_width"] = bw
    v = make_upstream_grads(s)
    ours, _ = run_ours(s, v, block_width=bw)
    d = scene_to_numpy(s)
    d["bw"] = bw
    ref = run_oracle_pipeline(d, {k: x.numpy() for k, x in v.items()})
    compare(ours, ref, s["settings"], ORACLE_FWD_RTOL, ORACLE_GRAD_RTOL, what=f"bw{bw}:")


def test_no_intersections_is_well_formed():
    """The reference is ill-formed here (texture.py:197-205); we return background and zero gradients."""
    from gstex_b200.synthetic import make_scene, make_upstream_grads

    s = make_scene("cfg2", n=64, width=48, height=48, texels=640)
    s["means"] = s["means"] + torch.tensor([1000.0, 0.0, 0.0])  # everything off-screen
    s["background"] = torch.tensor([0.2, 0.4, 0.6])
    ours, _ = run_ours(s, make_upstream_grads(s))
    assert np.allclose(ours["out_img"], np.array([0.2, 0.4, 0.6], np.float32)[None, None, :])
    assert np.all(ours["out_alpha"] == 0) and np.all(ours["v_means"] == 0) and np.all(ours["v_texture"] == 0)


def test_errors_are_loud():
    import gstex_b200 as G
    from gstex_b200.synthetic import make_scene

    s = make_scene("cfg2", n=32, width=32, height=32, texels=320)
    intr = (s["fx"], s["fy"], s["cx"], s["cy"])
    with pytest.raises(RuntimeError):  # CPU tensors: there is no CPU path
        G.get_aabb_2d(s["means"], s["scales"], 1, s["quats"], s["viewmat"], intr)
    sd = {k: (x.cuda() if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
    c, e = G.get_aabb_2d(sd["means"], sd["scales"], 1, sd["quats"], sd["viewmat"], intr)
    _, dep = G.project_points(sd["means"], sd["viewmat"], intr)
    nth = G.get_num_tiles_hit_2d(c, e, 32, 32, 16)
    args = [s["texture_info"], sd["texture_dims"], c, e, dep, nth, sd["colors"], sd["opacities"], sd["means"],
            sd["scales"], 1, sd["quats"], sd["uv0"], sd["umap"], sd["vmap"], sd["texture"], sd["viewmat"], sd["c2w"],
            *intr, 32, 32]
    with pytest.raises(AssertionError):
        G.texture_gaussians(*args, 32, 0, sd["background"])  # block_width > 16
    with pytest.raises(NotImplementedError):
        G.texture_gaussians(*args, 16, 1 << 16, sd["background"])  # viewer visualisation bit


def test_pair_stream_is_recorded_and_overflow_is_flagged(backward_path):
    """The stream backward must really be the path taken when recording is on (and the fallback when the pool is
    too small): read the stream header the forward left behind."""
    from gstex_b200.synthetic import make_scene, make_upstream_grads
    from gstex_b200.texture import PairStreamPolicy

    s = make_scene("cfg2", n=6000, width=160, height=160, texels=60000, settings=(1 << 9) | (1 << 10))
    key = (torch.cuda.current_device(), 160, 160, 16)
    PairStreamPolicy._state.pop(key, None)
    run_ours(s, make_upstream_grads(s))
    torch.cuda.synchronize()
    if backward_path == "recompute":
        assert key not in PairStreamPolicy._state
        return
    PairStreamPolicy.max_pairs(torch.device("cuda", key[0]), 160, 160, 16)  # consumes the pending header read-back
    used, overflow = PairStreamPolicy._state[key]["last"]
    if backward_path == "stream":
        assert overflow == 0 and used > 0
    else:
        assert overflow == 1


def test_prepared_binning_gives_identical_results():
    """`texture_gaussians(..., binning=begin_binning(...))` (multi-view pipelining hook) == the plain call, bit for
    bit, including when other work is enqueued between the two halves; a request can be finished twice."""
    import gstex_b200 as G
    from gstex_b200.synthetic import make_scene

    s = make_scene("cfg2", n=20000, width=320, height=320)
    sd = {k: (x.cuda() if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
    intr = (s["fx"], s["fy"], s["cx"], s["cy"])
    H, W, bw = s["H"], s["W"], 16
    _, depths, centers, extents, nth = G.project_bounds(sd["means"], sd["scales"], 1, sd["quats"], sd["viewmat"], intr,
                                                        H, W, bw)
    args = (s["texture_info"], sd["texture_dims"], centers, extents, depths, nth, sd["colors"], sd["opacities"],
            sd["means"], sd["scales"], 1, sd["quats"], sd["uv0"], sd["umap"], sd["vmap"], sd["texture"], sd["viewmat"],
            sd["c2w"], *intr, H, W, bw, s["settings"], sd["background"])
    plain = G.texture_gaussians(*args)
    tb = ((W + bw - 1) // bw, (H + bw - 1) // bw, 1)
    h = G.begin_binning(s["N"], centers, extents, depths, nth, tb, bw)
    junk = torch.randn(1 << 22, device="cuda").sort().values  # unrelated work between the halves
    prepared = G.texture_gaussians(*args, binning=h)
    for a, b in zip(plain, prepared):
        assert torch.equal(a, b)
    r1, r2 = G.finish_binning(h), G.finish_binning(h)
    assert r1[0] == r2[0] and r1[1] is r2[1] and r1[2] is r2[2]
    ref = G.bin_gaussians_to_tiles(s["N"], centers, extents, depths, nth, tb, bw)
    assert ref[0] == r1[0] and torch.equal(ref[1], r1[1]) and torch.equal(ref[2], r1[2])
    assert junk.numel() == 1 << 22


def _scale_quats(s, lo=0.6, hi=1.5, seed=5):
    """Non-unit quaternions: the kernels (like the reference, helpers.cuh:166-185) build R from the quaternion as given,
    so its columns shrink and skew; the model normalises (gstex.py:1062) but the drop-in API does not promise it."""
    g = torch.Generator().manual_seed(seed)
    s["quats"] = (s["quats"] * (lo + (hi - lo) * torch.rand(s["quats"].shape[0], 1, generator=g))).contiguous()
    return s


@pytest.mark.parametrize("cfg", [dict(name="cfg2", n=30000, size=320), dict(name="cfg5", n=250000, size=400),
                                 dict(name="cfg1", n=3000, size=200), dict(name="cfg2", n=30000, size=320, nonunit=True)],
                         ids=lambda c: c["name"] + ("-nonunit" if c.get("nonunit") else ""))
def test_culling_never_changes_a_result(cfg, monkeypatch):
    """The support-box / live-rectangle culling of the forward is claimed to be exact: with culling
    forced off (test hook `gstex_debug_force_nocull`: every list entry evaluated for every pixel block) every output must be
    bit-identical, the index outputs included."""
    import gstex_b200 as G
    from gstex_b200.synthetic import make_scene

    s = make_scene(cfg["name"], n=cfg["n"], width=cfg["size"], height=cfg["size"])
    if cfg.get("nonunit"):
        _scale_quats(s)
    sd = {k: (x.cuda() if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
    intr = (s["fx"], s["fy"], s["cx"], s["cy"])
    H, W, bw = s["H"], s["W"], 16
    _, depths, centers, extents, nth = G.project_bounds(sd["means"], sd["scales"], 1, sd["quats"], sd["viewmat"], intr,
                                                        H, W, bw)
    args = (s["texture_info"], sd["texture_dims"], centers, extents, depths, nth, sd["colors"], sd["opacities"],
            sd["means"], sd["scales"], 1, sd["quats"], sd["uv0"], sd["umap"], sd["vmap"], sd["texture"], sd["viewmat"],
            sd["c2w"], *intr, H, W, bw, s["settings"], sd["background"])
    from gstex_b200 import _lib

    culled = G.texture_gaussians(*args)
    _lib.load().gstex_debug_force_nocull(1)
    try:
        plain = G.texture_gaussians(*args)
    finally:
        _lib.load().gstex_debug_force_nocull(0)
    names = ["out_img", "out_depth", "out_reg", "out_alpha", "out_texture", "out_normal"]
    for nm, a, b in zip(names, culled, plain):
        assert torch.equal(a, b), f"{nm}: {int((a != b).sum())} elements differ between culled and unculled forward"


@pytest.mark.parametrize("cfg", [dict(name="cfg2", n=30000, size=320), dict(name="cfg5", n=250000, size=400),
                                 dict(name="cfg1", n=3000, size=200), dict(name="cfg4", n=100000, size=384),
                                 dict(name="cfg2", n=20000, size=250, settings=1 << 2), dict(name="cfg2", n=20000, size=250, C=1)],
                         ids=lambda c: f"{c['name']}-{c['n']}-s{c.get('settings', 'd')}-C{c.get('C', 3)}")
def test_dense_forward_equals_general_forward(cfg):
    """The lane-dense training kernel (raster_fwd_dense.cu) and the general kernel (raster_fwd.cu) apply the same pairs
    to every pixel in the same order with the same arithmetic: all nine forward outputs, the index images included,
    are bit-identical, and the pair streams they record give the same gradients."""
    from gstex_b200 import _lib
    from gstex_b200.synthetic import make_scene, make_upstream_grads

    s = make_scene(cfg["name"], n=cfg["n"], width=cfg["size"], height=cfg["size"] - 24, channels=cfg.get("C", 3),
                   settings=cfg.get("settings"))
    v = make_upstream_grads(s)
    sd = {k: (x.cuda() if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
    from oracle import parity

    lib = _lib.load()
    vd = {k: x.cuda() for k, x in v.items()}
    general = parity.run_ours_view(sd, vd)
    lib.gstex_debug_forward_variant(2)
    try:
        dense = parity.run_ours_view(sd, vd)
    finally:
        lib.gstex_debug_forward_variant(0)
    for k in parity.IMG_KEYS + ["final_idx", "depth_idx"]:
        assert torch.equal(dense[k], general[k]), f"{k}: {int((dense[k] != general[k]).sum())} elements differ"
    for k in parity.GRAD_KEYS:
        a, b = dense[k], general[k]
        scale = float(b.abs().max())
        assert float((a - b).abs().max()) <= 2e-5 * scale + 1e-30, f"{k}: {float((a - b).abs().max()):.3e} at scale {scale:.3e}"
    assert float(dense["v_texture"].abs().max()) > 0


def test_non_unit_quaternions_match_reference_extension():
    """Quaternions scaled by 0.6..1.5: support boxes are derived from the actual (skewed) frame, so culling stays exact
    and every output and gradient still matches the reference's kernels."""
    from oracle import ref_ext

    C = ref_ext.load()
    if C is None:
        pytest.skip("oracle/_ref/gstex_ref_C.so not built (python oracle/build_ref.py)")
    from gstex_b200.synthetic import make_scene, make_upstream_grads

    s = _scale_quats(make_scene("cfg2", n=6000, width=160, height=160, texels=60000, settings=(1 << 9) | (1 << 10)))
    v = make_upstream_grads(s)
    ours, sd = run_ours(s, v)
    vd = {k: x.cuda() for k, x in v.items()}
    t = lambda k: torch.from_numpy(ours[k]).cuda()  # noqa: E731
    ref = ref_ext.ref_forward_backward(C, sd, t("centers"), t("extents"), t("depths"), t("num_tiles_hit"), vd)
    ref = {k: (x.detach().cpu().numpy() if isinstance(x, torch.Tensor) else x) for k, x in ref.items()}
    compare(ours, ref, s["settings"], what="refext-nonunit:")


def test_graft_smoke_entry():
    import __graft_entry__ as g

    g.smoke()


@pytest.mark.parametrize("defer", [False, True], ids=["unpack-per-view", "deferred-unpack"])
def test_fused_gradient_accumulation_matches_autograd(defer, backward_path):
    """GradientTargets / GradientBucket: two views accumulated by the backward's own unpack kernel (zero_fill = 2)
    give the sums autograd's AccumulateGrad gives over the same two views (the additions are the same; the backward
    itself reduces with floating-point atomics, so two runs agree to rounding, not bit for bit)."""
    import gstex_b200 as G
    from gstex_b200.multiview import GradientBucket
    from gstex_b200.synthetic import make_scene, make_upstream_grads
    from gstex_b200.texture import GradientTargets

    from gstex_b200 import texture as T

    s = make_scene("cfg2", n=20000, width=256, height=256, views=3)
    v = {k: x.cuda() for k, x in make_upstream_grads(s).items()}
    sd = {k: (x.cuda() if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
    intr = (s["fx"], s["fy"], s["cx"], s["cy"])
    H, W, bw = s["H"], s["W"], 16
    keys = ["colors", "opacities", "means", "scales", "quats", "uv0", "umap", "vmap", "texture"]

    def run(fused):
        params = [sd[k].clone().requires_grad_(True) for k in keys]
        pd = dict(zip(keys, params))
        bucket = None
        if fused:
            bucket = GradientBucket(params)
            bucket.zero_()
            bucket.register(defer=defer)
        for i in range(3):
            vm, c2w = s["viewmats"][i].cuda().contiguous(), s["c2ws"][i].cuda().contiguous()
            _, depths, centers, extents, nth = G.project_bounds(pd["means"].detach(), pd["scales"].detach(), 1,
                                                                pd["quats"].detach(), vm, intr, H, W, bw)
            outs = G.texture_gaussians(s["texture_info"], sd["texture_dims"], centers, extents, depths, nth, pd["colors"],
                                       pd["opacities"], pd["means"], pd["scales"], 1, pd["quats"], pd["uv0"], pd["umap"],
                                       pd["vmap"], pd["texture"], vm, c2w, *intr, H, W, bw, s["settings"], sd["background"])
            torch.autograd.backward(list(outs), [v["v_out_img"], v["v_out_depth"], v["v_out_reg"], v["v_out_alpha"],
                                                 v["v_out_texture"], v["v_out_normal"]])
        if fused:
            assert all(p.grad is g for p, g in zip(params, bucket.views))
            if defer and backward_path == "stream":  # the sums are still in the padded scratch, not in the bucket
                assert len(GradientTargets._pending) == 1 and float(bucket.views[8].abs().max()) == 0.0
            bucket.allreduce(1)
            assert not GradientTargets._pending
            for buf in T._GRAD_SCRATCH.values():  # handed back all-zero by the one unpack sweep
                assert int(torch.count_nonzero(buf)) == 0
        GradientTargets.clear()
        return [p.grad.clone() for p in params]

    try:
        plain, fused = run(False), run(True)
    finally:
        GradientTargets.clear()
    for k, a, b in zip(keys, plain, fused):
        scale = float(a.abs().max())
        err = float((a - b).abs().max())
        assert err <= 2e-5 * scale + 1e-30, f"{k}: fused accumulation differs from autograd's ({err:.3e} at scale {scale:.3e})"
    assert float(plain[8].abs().max()) > 0 and float(plain[2].abs().max()) > 0


def test_gradient_scratch_is_returned_all_zero(backward_path):
    """The backward's scratch (padded accumulators + touched flags) is cached across calls and never cleared again:
    every call must hand it back all-zero, over views that touch different splats, and repeated backward passes over
    the cached buffer must reproduce the first one (to the rounding of the floating-point atomics)."""
    import gstex_b200 as G
    from gstex_b200 import texture as T
    from gstex_b200.synthetic import make_scene, make_upstream_grads

    s = make_scene("cfg2", n=20000, width=256, height=256, views=3)
    v = {k: x.cuda() for k, x in make_upstream_grads(s).items()}
    sd = {k: (x.cuda() if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
    intr = (s["fx"], s["fy"], s["cx"], s["cy"])
    H, W, bw = s["H"], s["W"], 16
    keys = ["colors", "opacities", "means", "scales", "quats", "uv0", "umap", "vmap", "texture"]
    T._GRAD_SCRATCH.clear()

    def grads(i):
        params = [sd[k].clone().requires_grad_(True) for k in keys]
        pd = dict(zip(keys, params))
        vm, c2w = s["viewmats"][i].cuda().contiguous(), s["c2ws"][i].cuda().contiguous()
        _, depths, centers, extents, nth = G.project_bounds(pd["means"].detach(), pd["scales"].detach(), 1,
                                                            pd["quats"].detach(), vm, intr, H, W, bw)
        outs = G.texture_gaussians(s["texture_info"], sd["texture_dims"], centers, extents, depths, nth, pd["colors"],
                                   pd["opacities"], pd["means"], pd["scales"], 1, pd["quats"], pd["uv0"], pd["umap"],
                                   pd["vmap"], pd["texture"], vm, c2w, *intr, H, W, bw, s["settings"], sd["background"])
        torch.autograd.backward(list(outs), [v["v_out_img"], v["v_out_depth"], v["v_out_reg"], v["v_out_alpha"],
                                             v["v_out_texture"], v["v_out_normal"]])
        return [p.grad.clone() for p in params]

    first = grads(0)                       # fresh scratch: cleared by the library
    if backward_path == "recompute":       # no pair stream, no scratch
        assert len(T._GRAD_SCRATCH) == 0
        return
    assert len(T._GRAD_SCRATCH) == 1
    for i in (1, 2, 0):                    # cached scratch, other views in between
        again = grads(i)
        (buf,) = T._GRAD_SCRATCH.values()
        assert int(torch.count_nonzero(buf)) == 0, f"view {i}: the scratch came back with non-zero bytes"
    for k, a, b in zip(keys, first, again):
        scale = float(a.abs().max())
        err = float((a - b).abs().max())
        assert err <= 2e-5 * scale + 1e-30, f"{k}: backward over the cached scratch differs ({err:.3e} at scale {scale:.3e})"
    assert float(first[8].abs().max()) > 0


def test_texture_affine_matches_activated_texture(backward_path):
    """`texture_affine=(C0, 0.5)` on raw texels == the reference caller's `texture = SH2RGB(texture_dc)` followed by
    the plain call (NS/models/gstex.py:1091-1121), forward and -- through the chain rule -- backward."""
    import gstex_b200 as G
    from gstex_b200.synthetic import make_scene, make_upstream_grads

    C0 = 0.28209479177387814
    s = make_scene("cfg2", n=20000, width=256, height=256)
    v = {k: x.cuda() for k, x in make_upstream_grads(s).items()}
    sd = {k: (x.cuda() if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
    intr = (s["fx"], s["fy"], s["cx"], s["cy"])
    H, W, bw = s["H"], s["W"], 16
    keys = ["colors", "opacities", "means", "scales", "quats", "uv0", "umap", "vmap", "texture"]
    raw = (sd["texture"] - 0.5) / C0          # texture_dc whose SH2RGB is the scene's texture
    vm, c2w = s["viewmats"][0].cuda().contiguous(), s["c2ws"][0].cuda().contiguous()

    def run(fused):
        params = [sd[k].clone().requires_grad_(True) for k in keys]
        pd = dict(zip(keys, params))
        pd["texture"] = params[8] = raw.clone().requires_grad_(True)
        tex = pd["texture"] if fused else pd["texture"] * C0 + 0.5
        _, depths, centers, extents, nth = G.project_bounds(pd["means"].detach(), pd["scales"].detach(), 1,
                                                            pd["quats"].detach(), vm, intr, H, W, bw)
        outs = G.texture_gaussians(s["texture_info"], sd["texture_dims"], centers, extents, depths, nth, pd["colors"],
                                   pd["opacities"], pd["means"], pd["scales"], 1, pd["quats"], pd["uv0"], pd["umap"],
                                   pd["vmap"], tex, vm, c2w, *intr, H, W, bw, s["settings"], sd["background"],
                                   texture_affine=(C0, 0.5) if fused else None)
        torch.autograd.backward(list(outs), [v["v_out_img"], v["v_out_depth"], v["v_out_reg"], v["v_out_alpha"],
                                             v["v_out_texture"], v["v_out_normal"]])
        return [o.detach() for o in outs], [p.grad for p in params]

    (o_ref, g_ref), (o_fus, g_fus) = run(False), run(True)
    for nm, a, b in zip(["out_img", "out_depth", "out_reg", "out_alpha", "out_texture", "out_normal"], o_ref, o_fus):
        err = float((a - b).abs().max())
        assert err <= 2e-6 * max(float(a.abs().max()), 1e-30), f"{nm}: {err:.3e}"
    for k, a, b in zip(keys, g_ref, g_fus):
        scale = float(a.abs().max())
        err = float((a - b).abs().max())
        assert err <= 3e-5 * scale + 1e-30, f"{k}: fused activation differs ({err:.3e} at scale {scale:.3e})"
    assert float(g_ref[8].abs().max()) > 0


@pytest.mark.parametrize("channels", [3, 1])
@pytest.mark.parametrize("affine", [None, (0.28209479177387814, 0.5)], ids=["plain", "affine"])
def test_padded_texture_forward_is_bit_identical(channels, affine, backward_path):
    """`texture_padded=pad_texture(texture)` only changes HOW the four bilinear corners are loaded (one 128-bit load
    each instead of `channels` scalar ones): all six images bit for bit, and the gradients -- computed from the same
    recorded pairs -- within the atomics' reordering noise."""
    import gstex_b200 as G
    from gstex_b200.synthetic import make_scene, make_upstream_grads

    s = make_scene("cfg2", n=20000, width=256, height=256)
    v = {k: x.cuda() for k, x in make_upstream_grads(s).items()}
    sd = {k: (x.cuda() if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
    intr = (s["fx"], s["fy"], s["cx"], s["cy"])
    H, W, bw = s["H"], s["W"], 16
    keys = ["colors", "opacities", "means", "scales", "quats", "uv0", "umap", "vmap", "texture"]
    info = (s["texture_info"][0], s["texture_info"][1], channels)
    tex0 = sd["texture"][:, :channels].contiguous()
    vm, c2w = s["viewmats"][0].cuda().contiguous(), s["c2ws"][0].cuda().contiguous()
    padded = G.pad_texture(tex0)
    assert padded.shape == (tex0.shape[0], 4)
    assert torch.equal(padded[:, :channels], tex0) and not padded[:, channels:].any()

    def run(use_padded):
        pd = {k: sd[k].clone().requires_grad_(True) for k in keys}
        pd["texture"] = tex0.clone().requires_grad_(True)
        _, depths, centers, extents, nth = G.project_bounds(pd["means"].detach(), pd["scales"].detach(), 1,
                                                            pd["quats"].detach(), vm, intr, H, W, bw)
        outs = G.texture_gaussians(info, sd["texture_dims"], centers, extents, depths, nth, pd["colors"],
                                   pd["opacities"], pd["means"], pd["scales"], 1, pd["quats"], pd["uv0"], pd["umap"],
                                   pd["vmap"], pd["texture"], vm, c2w, *intr, H, W, bw, s["settings"], sd["background"],
                                   texture_affine=affine, texture_padded=padded if use_padded else None)
        torch.autograd.backward(list(outs), [v["v_out_img"], v["v_out_depth"], v["v_out_reg"], v["v_out_alpha"],
                                             v["v_out_texture"][..., :channels].contiguous(), v["v_out_normal"]])
        return [o.detach() for o in outs], [pd[k].grad for k in keys]

    (o_ref, g_ref), (o_pad, g_pad) = run(False), run(True)
    for nm, a, b in zip(["out_img", "out_depth", "out_reg", "out_alpha", "out_texture", "out_normal"], o_ref, o_pad):
        assert torch.equal(a, b), f"{nm}: {int((a != b).sum())} elements differ with the padded texture"
    for k, a, b in zip(keys, g_ref, g_pad):
        scale = float(a.abs().max())
        assert float((a - b).abs().max()) <= 3e-5 * scale + 1e-30, k
    assert float(g_ref[8].abs().max()) > 0


def test_padded_texture_errors_are_loud():
    import gstex_b200 as G
    from gstex_b200 import _lib

    with pytest.raises(ValueError):
        G.pad_texture(torch.zeros(10, 4, device="cuda"))
    with pytest.raises(ValueError):
        G.pad_texture(torch.zeros(10, 3, device="cuda"), out=torch.zeros(9, 4, device="cuda"))
    assert G.pad_texture(torch.zeros(0, 3, device="cuda")).shape == (0, 4)
    lib = _lib.load()
    t = torch.zeros(8, 4, device="cuda")
    assert lib.gstex_texture_pad4(2, 4, _lib.ptr(t), _lib.ptr(t), None) == -1            # channels > 3
    assert lib.gstex_texture_pad4(1, 3, _lib.ptr(t), t.data_ptr() + 4, None) == -1       # unaligned destination


def test_records_packed_ahead_give_identical_results(backward_path):
    """`BinningStream.finish(pre, pack=...)` moves the forward's per-view packing to the binning stream; the forward then
    takes the records as given (`gstex_texture_forward_packed`).  Same kernels on the same inputs: images bit for bit,
    gradients within the atomics' noise -- also when the same packed records are rendered twice, and when a tensor
    differs from the one that was packed (the forward must notice and pack again)."""
    import gstex_b200 as G
    from gstex_b200 import _lib
    from gstex_b200.multiview import BinningStream
    from gstex_b200.synthetic import make_scene, make_upstream_grads

    s = make_scene("cfg2", n=20000, width=256, height=256)
    v = {k: x.cuda() for k, x in make_upstream_grads(s).items()}
    sd = {k: (x.cuda() if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
    intr = (s["fx"], s["fy"], s["cx"], s["cy"])
    H, W, bw = s["H"], s["W"], 16
    keys = ["colors", "opacities", "means", "scales", "quats", "uv0", "umap", "vmap", "texture"]
    vm, c2w = s["viewmats"][0].cuda().contiguous(), s["c2ws"][0].cuda().contiguous()
    bs = BinningStream(torch.device("cuda"))
    calls = []
    lib = _lib.load()

    def run(mode):
        pd = {k: sd[k].clone().requires_grad_(True) for k in keys}
        pre = bs.begin(pd["means"], pd["scales"], 1, pd["quats"], vm, intr, H, W, bw)
        pack = dict(texture_dims=sd["texture_dims"], colors=pd["colors"], opacity=pd["opacities"], uv0=pd["uv0"],
                    umap=pd["umap"], vmap=pd["vmap"], c2w=c2w, settings=s["settings"])
        bs.finish(pre, pack=None if mode == "plain" else pack)
        assert (pre.binning.records is not None) == (mode != "plain")
        colors = pd["colors"]
        if mode == "stale":  # another tensor with other values than the one that was packed
            with torch.no_grad():
                pd["colors"] = (pd["colors"] * 0.5).requires_grad_(True)
            colors = pd["colors"]
        outs = None
        for _ in range(2 if mode == "twice" else 1):
            n0 = lib.gstex_launch_count()
            outs = G.texture_gaussians(s["texture_info"], sd["texture_dims"], *pre.args, colors, pd["opacities"],
                                       pd["means"], pd["scales"], 1, pd["quats"], pd["uv0"], pd["umap"], pd["vmap"],
                                       pd["texture"], vm, c2w, *intr, H, W, bw, s["settings"], sd["background"],
                                       binning=pre.binning)
            calls.append((mode, lib.gstex_launch_count() - n0))
        torch.autograd.backward(list(outs), [v["v_out_img"], v["v_out_depth"], v["v_out_reg"], v["v_out_alpha"],
                                             v["v_out_texture"], v["v_out_normal"]])
        return [o.detach() for o in outs], [pd[k].grad for k in keys]

    (o_ref, g_ref), (o_pk, g_pk), (o_tw, g_tw) = run("plain"), run("packed"), run("twice")
    for nm, a, b, c in zip(["out_img", "out_depth", "out_reg", "out_alpha", "out_texture", "out_normal"], o_ref, o_pk, o_tw):
        assert torch.equal(a, b), f"{nm}: {int((a != b).sum())} elements differ with records packed ahead"
        assert torch.equal(a, c), f"{nm}: {int((a != c).sum())} elements differ on the second render of the same records"
    for k, a, b, c in zip(keys, g_ref, g_pk, g_tw):
        scale = float(a.abs().max())
        assert float((a - b).abs().max()) <= 3e-5 * scale + 1e-30, k
        assert float((a - c).abs().max()) <= 3e-5 * scale + 1e-30, k
    launches = dict((m, n) for m, n in calls)
    assert launches["packed"] == launches["plain"] - 1, f"the packed forward must not pack again: {calls}"
    # stale records: the colours given to the forward are not the packed ones -> same result as a plain call on them
    sd2 = dict(sd)
    o_stale, _ = run("stale")
    sd["colors"] = sd["colors"] * 0.5
    o_half, _ = run("plain")
    sd.update(colors=sd2["colors"])
    assert torch.equal(o_stale[0], o_half[0]) and not torch.equal(o_stale[0], o_ref[0])


def test_unused_outputs_need_no_zero_gradients(backward_path):
    """A loss that reads only some of the six outputs (the reference trainer's reads rgb, texture and alpha) hands None
    for the others; they reach the library as NULL = zero.  Same gradients as with explicit zero images."""
    import gstex_b200 as G
    from gstex_b200.synthetic import make_scene, make_upstream_grads

    s = make_scene("cfg2", n=20000, width=256, height=256)
    v = {k: x.cuda() for k, x in make_upstream_grads(s).items()}
    sd = {k: (x.cuda() if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
    intr = (s["fx"], s["fy"], s["cx"], s["cy"])
    H, W, bw = s["H"], s["W"], 16
    keys = ["colors", "opacities", "means", "scales", "quats", "uv0", "umap", "vmap", "texture"]
    vm, c2w = s["viewmats"][0].cuda().contiguous(), s["c2ws"][0].cuda().contiguous()

    def run(used):
        pd = {k: sd[k].clone().requires_grad_(True) for k in keys}
        _, depths, centers, extents, nth = G.project_bounds(pd["means"].detach(), pd["scales"].detach(), 1,
                                                            pd["quats"].detach(), vm, intr, H, W, bw)
        outs = G.texture_gaussians(s["texture_info"], sd["texture_dims"], centers, extents, depths, nth, pd["colors"],
                                   pd["opacities"], pd["means"], pd["scales"], 1, pd["quats"], pd["uv0"], pd["umap"],
                                   pd["vmap"], pd["texture"], vm, c2w, *intr, H, W, bw, s["settings"], sd["background"])
        names = ["out_img", "out_depth", "out_reg", "out_alpha", "out_texture", "out_normal"]
        if used is None:  # all six, zeros for the unused ones
            grads = [v["v_" + nm] if nm in ("out_img", "out_alpha") else torch.zeros_like(o) for nm, o in zip(names, outs)]
            torch.autograd.backward(list(outs), grads)
        else:
            sum((o * v["v_" + nm]).sum() for nm, o in zip(names, outs) if nm in used).backward()
        return [pd[k].grad for k in keys]

    g_zero, g_none = run(None), run(("out_img", "out_alpha"))
    for k, a, b in zip(keys, g_zero, g_none):
        scale = float(a.abs().max())
        assert float((a - b).abs().max()) <= 3e-5 * scale + 1e-30, k
    assert float(g_none[2].abs().max()) > 0

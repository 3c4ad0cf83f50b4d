"""GPU parity, next row f-4: `texture_edit` (gstex_b200/texture_edit.py -> C ABI gstex_texture_edit) against the CPU
oracle and against the reference's own CUDA extension (oracle/_ref) on the same inputs.  The kernel adds with float
atomics in no fixed order (so does the reference's), hence tolerances: 1e-5 of the tensor's scale per element, with a
bounded share of texels allowed outside it where a pair's discrete decision (alpha < 1/255, T <= 1e-4, the depth
window's edges) flipped between two fp32 implementations."""
import os

import numpy as np
import pytest
import torch

from common import err_stats, make_edit_inputs, run_oracle_edit, scene_to_numpy

pytestmark = pytest.mark.gpu

CASES = [
    dict(name="cfg2", n=40000, width=128, height=128, texels=300000, settings=0, bw=16, window="median"),
    dict(name="cfg2", n=40000, width=128, height=128, texels=300000, settings=1, bw=16, window="median"),
    dict(name="cfg2", n=40000, width=128, height=128, texels=300000, settings=1 << 13, bw=16, window="all"),
    dict(name="cfg2", n=30000, width=100, height=75, texels=0, settings=0, bw=16, window="all"),  # 1x1 charts, ragged
    dict(name="cfg1", n=400, width=72, height=56, texels=400 * 16, settings=1, bw=7, window="all"),
    dict(name="cfg4", n=200000, width=240, height=136, texels=1000000, settings=0, bw=16, window="median"),
]


def _id(c):
    return f"{c['name']}-n{c['n']}-{c['width']}x{c['height']}-s{c['settings']}-bw{c['bw']}-{c['window']}"


def _scene(c):
    from gstex_b200.synthetic import make_scene

    s = make_scene(c["name"], n=c["n"], width=c["width"], height=c["height"], texels=c["texels"], settings=0)
    s["block_width"] = c["bw"]
    return s


def run_ours_edit(s, e, settings, channels=5, bw=16):
    """The model's call sequence (NS/models/gstex.py:528-605) on our package."""
    import gstex_b200 as G

    sd = {k: (x.cuda() if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
    ed = {k: torch.from_numpy(x).cuda() for k, x in e.items()}
    intr = (s["fx"], s["fy"], s["cx"], s["cy"])
    H, W = s["H"], s["W"]
    _, depths = G.project_points(sd["means"], sd["viewmat"][:3, :], intr)
    centers, extents = G.get_aabb_2d(sd["means"], sd["scales"], 1, sd["quats"], sd["viewmat"][:3, :], intr)
    nth = G.get_num_tiles_hit_2d(centers, extents, H, W, bw)
    out = G.texture_edit((s["N"], 1, channels), sd["texture_dims"], ed["updated_img"], ed["updated_alpha"],
                         ed["depth_lower"], ed["depth_upper"], centers, extents, depths, nth, sd["opacities"],
                         sd["means"], sd["scales"], 1, sd["quats"], sd["uv0"], sd["umap"], sd["vmap"],
                         sd["viewmat"][:3, :], sd["c2w"], *intr, H, W, bw, settings,
                         background=torch.zeros(3, device="cuda"))
    return out, (sd, ed, centers, extents, depths, nth)


def check(ours, ref, what, rtol=1e-5, max_frac=2e-3):
    ours = np.asarray(ours)
    assert ours.shape == ref.shape
    rows = []
    for k in range(5):
        frac, relmax, scale = err_stats(ours[:, k], ref[:, k], rtol)
        rows.append(f"{what} ch{k}: scale {scale:.3e} max|err|/scale {relmax:.2e} outside {frac:.2e}")
    print("\n".join(rows))
    for k in range(5):
        frac, _, _ = err_stats(ours[:, k], ref[:, k], rtol)
        assert frac <= max(max_frac, 3.0 / ref.shape[0]), rows[k]
    # totals (every pair carries weight one): flips may move single pairs in or out
    t_ours, t_ref = float(ours[:, 4].astype(np.float64).sum()), float(ref[:, 4].astype(np.float64).sum())
    assert abs(t_ours - t_ref) <= 1e-3 * t_ref + 2, (what, t_ours, t_ref)


@pytest.mark.parametrize("c", CASES, ids=_id)
def test_edit_matches_oracle(c):
    s = _scene(c)
    d = scene_to_numpy(s)
    e = make_edit_inputs(d, window=c["window"])
    ref, npairs = run_oracle_edit(d, e, c["settings"], return_pairs=True)
    assert npairs > 0
    ours, _ = run_ours_edit(s, e, c["settings"], bw=c["bw"])
    check(ours.cpu().numpy(), ref, "oracle:", rtol=1e-4)


@pytest.mark.parametrize("c", CASES, ids=_id)
def test_edit_matches_reference_extension(c):
    from oracle import ref_ext

    C = ref_ext.load()
    if C is None:
        pytest.skip("oracle/_ref/gstex_ref_C.so not built (python oracle/build_ref.py)")
    s = _scene(c)
    d = scene_to_numpy(s)
    e = make_edit_inputs(d, window=c["window"])
    bw = c["bw"]
    ours, (sd, ed, centers, extents, depths, nth) = run_ours_edit(s, e, c["settings"], bw=bw)
    # texture_edit.py:169-236 around the reference's native calls
    H, W = s["H"], s["W"]
    tb = ((W + bw - 1) // bw, (H + bw - 1) // bw, 1)
    cum = torch.cumsum(nth, dim=0, dtype=torch.int32)
    num_intersects = int(cum[-1].item())
    _, _, _, gids_s, bins = ref_ext.ref_bin_and_sort(C, s["N"], num_intersects, centers, extents, depths, cum, tb, bw)
    total = int(torch.sum(sd["texture_dims"][:, 0] * sd["texture_dims"][:, 1]).item())
    ref = C.texture_edit(tb, (bw, bw, 1), (W, H, 1), (s["N"], 1, 5), total, sd["texture_dims"], ed["updated_img"],
                         ed["updated_alpha"], ed["depth_lower"], ed["depth_upper"], gids_s, bins, sd["opacities"],
                         sd["means"], sd["scales"], 1.0, sd["quats"], sd["uv0"], sd["umap"], sd["vmap"],
                         sd["viewmat"][:3, :].contiguous(), sd["c2w"].contiguous(), s["fx"], s["fy"], s["cx"], s["cy"],
                         int(c["settings"]), torch.zeros(3, device="cuda"))
    torch.cuda.synchronize()
    check(ours.cpu().numpy(), ref.cpu().numpy(), "reference extension:", rtol=1e-5)


def test_edit_culling_changes_nothing():
    """The test hook `gstex_debug_force_nocull` evaluates every list entry for every unit; the support-box culling must give the same sums
    (up to the order of the float atomics)."""
    c = CASES[1]
    s = _scene(c)
    e = make_edit_inputs(scene_to_numpy(s), window="all")
    a, _ = run_ours_edit(s, e, c["settings"])
    from gstex_b200 import _lib

    _lib.load().gstex_debug_force_nocull(1)
    try:
        b, _ = run_ours_edit(s, e, c["settings"])
    finally:
        _lib.load().gstex_debug_force_nocull(0)
    a, b = a.cpu().numpy(), b.cpu().numpy()
    assert float(a[:, 4].sum()) > 0
    np.testing.assert_allclose(a, b, rtol=2e-6, atol=2e-6 * float(np.abs(b).max()))


def test_edit_properties_and_model_blend():
    """Channel identities (texture_edit.cu:194-211) and the blend the model builds from the result
    (NS/models/gstex.py:601-604) at a larger size than the oracle cases."""
    from gstex_b200.synthetic import make_scene

    s = make_scene("cfg2", n=60000, width=400, height=400, texels=600000, settings=0)
    d = scene_to_numpy(s)
    H, W = s["H"], s["W"]
    e = dict(updated_img=np.broadcast_to(np.array([0.25, 0.5, 0.75], np.float32), (H, W, 3)).copy(),
             updated_alpha=np.ones((H, W, 1), np.float32), depth_lower=np.full((H, W), -1e30, np.float32),
             depth_upper=np.full((H, W), 1e30, np.float32))
    out, _ = run_ours_edit(s, e, 1 << 13)
    out = out.cpu().numpy()
    assert out.shape == (int(d["texture_dims"][:, 0].astype(np.int64) @ d["texture_dims"][:, 1]), 5)
    assert (out >= 0).all() and out[:, 4].sum() > H * W * 0.5
    tol = dict(rtol=1e-5, atol=1e-5 * float(out[:, 4].max()))
    np.testing.assert_allclose(out[:, 3], out[:, 4], **tol)
    for k, col in enumerate((0.25, 0.5, 0.75)):
        np.testing.assert_allclose(out[:, k], col * out[:, 4], **tol)
    # the model's blend: fully painted texels take the canvas colour, untouched texels keep theirs
    cur = d["texture"]
    weight = out[:, 3:4] / (out[:, 4:] + 1e-6)
    rgb = out[:, :3] / (out[:, 3:4] + 1e-6)
    new = rgb * weight + cur * (1 - weight)
    hit = out[:, 4] > 1e-2
    np.testing.assert_allclose(new[hit], np.broadcast_to([0.25, 0.5, 0.75], new[hit].shape), rtol=0, atol=1e-3)
    untouched = out[:, 4] == 0
    assert untouched.any()
    np.testing.assert_array_equal(new[untouched], cur[untouched])
    # empty window / no paint
    e["depth_upper"][:] = -1e30
    z, _ = run_ours_edit(s, e, 0)
    assert not z.any().item()


def test_edit_argument_errors():
    import gstex_b200 as G

    c = CASES[0]
    s = _scene(c)
    e = make_edit_inputs(scene_to_numpy(s), window="all")
    with pytest.raises(ValueError):
        run_ours_edit(s, e, 0, channels=3)
    with pytest.raises(NotImplementedError):  # use_torch_impl=True: there is no PyTorch/CPU path
        G.texture_edit((1, 1, 5), *([None] * 12), 1, *([None] * 6), 1.0, 1.0, 0.0, 0.0, 8, 8, 16, 0, None, True)
    cpu = {k: (x.clone() if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
    with pytest.raises(RuntimeError):  # CPU tensors: there is no CPU path
        z = torch.zeros
        G.texture_edit((s["N"], 1, 5), cpu["texture_dims"], z(8, 8, 3), z(8, 8, 1), z(8, 8), z(8, 8), z(s["N"], 2),
                       z(s["N"], 2), z(s["N"]), z(s["N"], dtype=torch.int32), cpu["opacities"], cpu["means"],
                       cpu["scales"], 1, cpu["quats"], cpu["uv0"], cpu["umap"], cpu["vmap"], cpu["viewmat"][:3],
                       cpu["c2w"], 1.0, 1.0, 4.0, 4.0, 8, 8, 16, 0, z(3))

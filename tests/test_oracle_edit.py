"""CPU: the texture_edit oracle (oracle/gstex_oracle_raster.c: orc_texture_edit).  The reference has no Python twin and
no test of texture_edit (texture_edit.py:117), so there are no golden vectors: PARITY UNPINNED on the CPU.  What can be
held here are the properties the kernel's definition implies (cuda/csrc/texture_edit.cu:180-216); the oracle itself is
pinned on the GPU box against the reference's own extension (tests/test_gpu_edit.py)."""
import numpy as np

from common import make_edit_inputs, run_oracle_edit, run_oracle_pipeline, scene_to_numpy


def _scene(settings=0, bw=16):
    from gstex_b200.synthetic import make_scene

    s = make_scene("cfg2", n=12000, width=80, height=64, texels=60000, settings=settings)
    d = scene_to_numpy(s)
    d["bw"] = bw
    return d


def test_edit_oracle_weights_sum_to_pair_count_and_channels_are_consistent():
    d = _scene()
    e = make_edit_inputs(d, window="all")
    out, npairs = run_oracle_edit(d, e, 0, return_pairs=True)
    assert npairs > 1000
    total = int((d["texture_dims"][:, 0].astype(np.int64) * d["texture_dims"][:, 1]).sum())
    assert out.shape == (total, 5)
    # the four bilinear weights of a pair sum to one: channel 4 sums to the number of pairs in range
    assert abs(float(out[:, 4].astype(np.float64).sum()) - npairs) <= 1e-4 * npairs
    # fully painted canvas with alpha 1 and one colour: channel 3 == channel 4, rgb channels = colour * channel 4
    e["updated_alpha"][:] = 1.0
    e["updated_img"][:] = np.array([0.25, 0.5, 0.75], np.float32)
    out = run_oracle_edit(d, e, 0)
    np.testing.assert_allclose(out[:, 3], out[:, 4], rtol=1e-6, atol=1e-7)
    for k, c in enumerate((0.25, 0.5, 0.75)):
        np.testing.assert_allclose(out[:, k], c * out[:, 4], rtol=1e-6, atol=1e-7)
    assert (out >= 0).all()


def test_edit_oracle_depth_window_and_mask():
    d = _scene()
    none = run_oracle_edit(d, make_edit_inputs(d, window="none"), 0)
    assert not none.any()
    full, n_all = run_oracle_edit(d, make_edit_inputs(d, window="all"), 0, return_pairs=True)
    e = make_edit_inputs(d, window="median")
    med, n_med = run_oracle_edit(d, e, 0, return_pairs=True)
    assert 0 < n_med < n_all
    assert (med[:, 4] <= full[:, 4] * (1 + 1e-6) + 1e-7).all()
    # a pixel that composited anything has its median-depth pair inside the window: at least one pair per such pixel
    alpha = run_oracle_pipeline(d)["out_alpha"]
    assert n_med >= int((alpha > 0.6).sum())
    # untouched canvas (alpha 0 outside the painted rectangle): channels 0..3 receive only from painted pixels
    e0 = dict(e, updated_alpha=np.zeros_like(e["updated_alpha"]))
    z = run_oracle_edit(d, e0, 0)
    assert not z[:, :4].any() and z[:, 4].any()
    np.testing.assert_allclose(z[:, 4], med[:, 4], rtol=1e-6, atol=1e-7)


def test_edit_oracle_is_linear_in_the_canvas_and_reads_blur_from_bit_0():
    d = _scene()
    e = make_edit_inputs(d, window="all")
    a = run_oracle_edit(d, e, 0)
    e2 = dict(e, updated_img=(0.5 * e["updated_img"]).astype(np.float32))
    b = run_oracle_edit(d, e2, 0)
    np.testing.assert_allclose(b[:, :3], 0.5 * a[:, :3], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(b[:, 3:], a[:, 3:], rtol=1e-6, atol=1e-7)
    # settings: only bit 0 is read (texture_edit.cu:48); the forward's blur bit 9 and the model's 1<<13 change nothing
    for s in (1 << 9, 1 << 13, (1 << 9) | (1 << 10)):
        np.testing.assert_array_equal(run_oracle_edit(d, e, s), a)
    assert np.abs(run_oracle_edit(d, e, 1) - a).max() > 0  # the blur floor widens small splats to a pixel


def test_oracle_visualization_normals_bit_15_only_flips_normals():
    """settings bit 15 (C/texture.cu:58,236-241): every output but out_normal is unchanged; the accumulated normal is a
    sum of camera-facing normals, so its component along each pixel's ray is never positive."""
    from gstex_b200.synthetic import make_scene

    d = scene_to_numpy(make_scene("cfg1", n=400, width=72, height=56, texels=6400, settings=0))
    a, b = run_oracle_pipeline(d, settings=0), run_oracle_pipeline(d, settings=1 << 15)
    for k in ("out_img", "out_depth", "out_reg", "out_alpha", "out_texture"):
        np.testing.assert_array_equal(a[k], b[k])
    assert np.abs(a["out_normal"] - b["out_normal"]).max() > 0.1
    H, W = d["H"], d["W"]
    ys, xs = np.mgrid[0:H, 0:W]
    cam = np.stack([(xs + 0.5 - d["cx"]) / d["fx"], (ys + 0.5 - d["cy"]) / d["fy"], np.ones((H, W))], -1)
    rays = cam @ d["c2w"][:3, :3].T
    assert ((rays * b["out_normal"]).sum(-1) <= 1e-6).all()
    assert ((rays * a["out_normal"]).sum(-1) > 1e-3).any()

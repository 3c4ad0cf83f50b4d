"""CPU: the compositing + photometric-loss oracle (oracle/gstex_oracle_loss.c) against the golden vectors made from the
reference's own lines (nerfstudio/models/gstex.py:1204-1205, 1297-1302, 1319) and torch.autograd
(tests/golden/make_golden_loss.py)."""
import os

import numpy as np
import pytest

import oracle

GOLD = os.path.join(os.path.dirname(__file__), "golden", "loss_golden.npz")
CASES = ["small", "masked", "viewer6", "min11"]


def load_case(name):
    d = np.load(GOLD)
    return {k[len(name) + 1:]: d[k] for k in d.files if k.startswith(name + "_")}


def check_against_golden(got, c):
    """rgb bit for bit; the scalar losses to 1e-5 relative (north-star image bar); gradients to 1e-4 of their scale
    (north-star gradient bar), with the handful of pixels whose |gt - pred| sign or clamp decision sits on an fp32
    rounding boundary excluded (at most 1e-3 of the elements)."""
    np.testing.assert_array_equal(got["rgb"], c["rgb"])
    for i, k in enumerate(["Ll1", "ssim", "main_loss"]):
        assert abs(got[k] - c["losses"][i]) <= 1e-5 * abs(c["losses"][i]) + 1e-7, (k, got[k], c["losses"][i])
    for k in ["v_out_img", "v_out_texture", "v_out_alpha"]:
        a, b = np.asarray(got[k]), c[k]
        assert a.shape == b.shape
        scale = np.abs(b).max()
        bad = np.abs(a - b) > 1e-4 * scale
        assert bad.mean() <= 1e-3, f"{k}: {bad.sum()} of {bad.size} elements beyond 1e-4 of scale {scale:.3e}"


@pytest.mark.parametrize("name", CASES)
def test_loss_oracle_matches_reference_lines(name):
    c = load_case(name)
    lam, v_main = float(c["params"][0]), float(c["params"][1])
    got = oracle.photometric_loss(c["out_img"], c["out_texture"], c["out_alpha"], c["background"], c["gt_img"], lam,
                                  c.get("mask"), v_main=v_main)
    check_against_golden(got, c)


def test_loss_oracle_properties():
    g = np.random.default_rng(3)
    H, W = 20, 31
    gt = g.random((H, W, 3)).astype(np.float32)
    zeros3, zeros1 = np.zeros((H, W, 3), np.float32), np.zeros((H, W), np.float32)
    bg = np.zeros(3, np.float32)
    # a perfect prediction: Ll1 = 0, SSIM = 1, main = 0; the L1 term has zero gradient at 0 (sign(0) = 0)
    r = oracle.photometric_loss(gt, zeros3, np.ones((H, W), np.float32), bg, gt, 0.2)
    assert r["Ll1"] == 0.0 and abs(r["ssim"] - 1.0) < 1e-6 and abs(r["main_loss"]) < 1e-6
    assert np.abs(r["v_out_img"]).max() < 1e-6
    # everything outside [0, 1] before the clamp: no gradient reaches the layers
    r = oracle.photometric_loss(zeros3 + 2.0, zeros3, zeros1 + 1.0, bg, gt, 0.2)
    assert np.all(r["rgb"] == 1.0) and not r["v_out_img"].any() and not r["v_out_alpha"].any()
    # empty pixels over a black background are exactly 0 = the clamp's lower bound: torch passes the gradient there
    r = oracle.photometric_loss(zeros3, zeros3, zeros1, bg, gt, 0.0)
    assert np.allclose(r["v_out_img"], -1.0 / (3 * H * W))
    # the alpha gradient is -sum_k grad_k * background_k
    bg2 = np.array([0.3, 0.6, 0.9], np.float32)
    r = oracle.photometric_loss(gt * 0.5, zeros3, zeros1 + 0.5, bg2, gt, 0.2)
    np.testing.assert_allclose(r["v_out_alpha"], -(r["v_out_img"] * bg2).sum(-1), rtol=1e-5, atol=1e-12)
    with pytest.raises(ValueError):
        oracle.photometric_loss(zeros3[:10], zeros3[:10], zeros1[:10], bg, gt[:10])

"""Shared helpers for the parity tests: golden loading, the oracle pipeline and tolerance checks."""
import glob
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN_DIR = os.path.join(HERE, "golden")

IMG_RTOL = 1e-5   # north_star: rendered images within 1e-5 relative in fp32
GRAD_RTOL = 1e-4  # north_star: splat / texel gradients within 1e-4 relative

FWD_KEYS = ["out_img", "out_depth", "out_reg", "out_alpha", "out_texture", "out_normal"]
GRAD_KEYS = ["v_colors", "v_opacity", "v_means", "v_scales", "v_quats", "v_uv0", "v_umap", "v_vmap", "v_texture"]
VOUT_KEYS = ["v_out_img", "v_out_depth", "v_out_reg", "v_out_alpha", "v_out_texture", "v_out_normal"]


def golden_files():
    return sorted(glob.glob(os.path.join(GOLDEN_DIR, "golden_*.npz")))


def load_golden(path):
    d = dict(np.load(path))
    fx, fy, cx, cy, H, W, bw, settings, C = d["scalars"]
    d.update(fx=float(fx), fy=float(fy), cx=float(cx), cy=float(cy), H=int(H), W=int(W), bw=int(bw),
             settings=int(settings), C=int(C))
    return d


def rel_err(a, b):
    """max |a-b| / max|b|: error relative to the tensor's scale (what 'x relative in fp32' can mean for
    sums of signed terms; element-wise relative error is unbounded where terms cancel)."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    scale = float(np.abs(b).max())
    if scale == 0.0:
        return float(np.abs(a).max())
    return float(np.abs(a - b).max() / scale)


def assert_close_frac(a, b, rtol, name, max_outlier_frac=0.0, atol_scale=1.0):
    """|a-b| <= rtol * (|b| + atol_scale * scale) element-wise for all but `max_outlier_frac` of the
    elements (discrete decisions -- alpha < 1/255, T <= 1e-4, T > 0.5 -- may flip for a handful of
    pixels between two fp32 implementations; those are counted, bounded and reported)."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    assert a.shape == b.shape, (name, a.shape, b.shape)
    scale = float(np.abs(b).max()) if b.size else 0.0
    bad = np.abs(a - b) > rtol * (np.abs(b) + atol_scale * scale) + 1e-30
    frac = float(bad.mean()) if b.size else 0.0
    assert frac <= max_outlier_frac, (
        f"{name}: {int(bad.sum())}/{b.size} elements ({frac:.2e}) outside rtol={rtol}; "
        f"max abs err {np.abs(a - b).max():.3e} at scale {scale:.3e}")
    return frac


def run_oracle_pipeline(d, v=None, settings=None):
    """Full oracle path on a golden / scene dict of numpy arrays."""
    import oracle
    import oracle.raster as R

    intr = (d["fx"], d["fy"], d["cx"], d["cy"])
    H, W, bw = d["H"], d["W"], d.get("bw", 16)
    _, depths = oracle.project_points(d["means"], d["viewmat"], intr)
    centers, extents = oracle.get_aabb_2d(d["means"], d["scales"], 1.0, d["quats"], d["viewmat"], intr)
    nth = oracle.get_num_tiles_hit_2d(centers, extents, H, W, bw)
    n = d["means"].shape[0]
    res = R.texture_gaussians_fwd_bwd((n, 1, d["C"]), d["texture_dims"], centers, extents, depths, nth, d["colors"],
                                      d["opacities"], d["means"], d["scales"], 1.0, d["quats"], d["uv0"], d["umap"],
                                      d["vmap"], d["texture"], d["viewmat"], d["c2w"], *intr, H, W, bw,
                                      d["settings"] if settings is None else settings, d["background"], v)
    res.update(centers=centers, extents=extents, depths=depths, num_tiles_hit=nth)
    return res


def scene_to_numpy(s):
    import torch

    out = {}
    for k, x in s.items():
        out[k] = x.detach().cpu().numpy() if isinstance(x, torch.Tensor) else x
    out["bw"] = s.get("block_width", 16)
    return out


def err_stats(a, b, rtol, atol_abs=0.0):
    """(fraction outside tolerance, max |a-b| / scale, scale) with tolerance rtol*(|b| + scale) + atol_abs."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    if b.size == 0:
        return 0.0, 0.0, 0.0
    scale = float(np.abs(b).max())
    bad = np.abs(a - b) > rtol * (np.abs(b) + scale) + atol_abs + 1e-30
    return float(bad.mean()), float(np.abs(a - b).max() / scale) if scale > 0 else float(np.abs(a).max()), scale


def reg_atol(ref, settings):
    """Absolute rounding floor of out_reg = sum vis*(d^2*S0 + S2 - 2*d*S1): the three terms are O(d^2) each
    and cancel to a value that can be 1e5 times smaller, so fp32 leaves ~n*eps*d^2 of noise whatever the
    implementation (d ~ 1 with NDC depths, settings bit 10; d = t otherwise)."""
    if settings & (1 << 10):
        d2 = 1.0
    else:
        d2 = max(1.0, float(np.abs(ref["out_depth"]).max()) ** 2)
    return 2e-6 * d2


def compare_all(ours, ref, settings, fwd_rtol, grad_rtol, what, max_flip_frac=2e-4, min_flip_count=3, verbose=True):
    """Check every forward output and gradient; collect all violations, print an error table, assert once.
    A `max_flip_frac` share of elements (at least `min_flip_count`) may sit outside the tolerance: pixels
    where a discrete decision (alpha < 1/255, T <= 1e-4, T > 0.5, blur floor) flipped between the two
    fp32 implementations, or where a grazing ray makes t = num/den ill-conditioned."""
    fails, rows = [], []
    for k in FWD_KEYS + GRAD_KEYS:
        if k not in ref or k not in ours:
            continue
        is_grad = k in GRAD_KEYS
        rtol = grad_rtol if is_grad else fwd_rtol
        atol = reg_atol(ref, settings) if k == "out_reg" else 0.0
        frac, relmax, scale = err_stats(ours[k], ref[k], rtol, atol)
        size = np.asarray(ref[k]).size
        # round 2: gradients get the same allowance as images (round 1 allowed ten times more)
        allowed = max(max_flip_frac, min_flip_count / max(size, 1))
        rows.append(f"{what}{k:12s} scale {scale:9.3e} max|err|/scale {relmax:9.2e} outside {frac:8.2e} (allowed {allowed:7.1e})")
        if frac > allowed:
            fails.append(rows[-1])
    if verbose:
        print("\n".join(rows))
    assert not fails, "\n" + "\n".join(fails)
    return rows


def make_edit_inputs(d, seed=7, window="median", painted_frac=0.4):
    """Edit canvas + per-pixel depth window for a scene dict of numpy arrays (`scene_to_numpy`).
    window = "median": [out_depth - 1e-2, out_depth + 1e-2] around the oracle forward's median depth, as the model's
    edit_texture does (NS/models/gstex.py:566-568); "all": every depth; "none": an empty window.  A rectangle covering
    `painted_frac` of each image side is painted with random colours and alpha in (0, 1]; the rest has alpha 0."""
    g = np.random.default_rng(seed)
    H, W = d["H"], d["W"]
    img = g.random((H, W, 3)).astype(np.float32)
    alpha = np.zeros((H, W, 1), np.float32)
    h0, w0 = int(H * 0.2), int(W * 0.3)
    h1, w1 = min(H, h0 + max(1, int(H * painted_frac))), min(W, w0 + max(1, int(W * painted_frac)))
    alpha[h0:h1, w0:w1, 0] = (0.05 + 0.95 * g.random((h1 - h0, w1 - w0))).astype(np.float32)
    if window == "all":
        lo = np.full((H, W), -1e30, np.float32)
        hi = np.full((H, W), 1e30, np.float32)
    elif window == "none":
        lo = np.full((H, W), 1.0, np.float32)
        hi = np.full((H, W), -1.0, np.float32)
    else:
        depth = run_oracle_pipeline(d)["out_depth"]
        lo, hi = (depth - 1e-2).astype(np.float32), (depth + 1e-2).astype(np.float32)
    return dict(updated_img=img, updated_alpha=alpha, depth_lower=lo, depth_upper=hi)


def run_oracle_edit(d, e, settings, channels=5, return_pairs=False):
    """oracle.raster.texture_edit on a scene dict + edit inputs (binning by the oracle as well)."""
    import oracle
    import oracle.raster as R

    intr = (d["fx"], d["fy"], d["cx"], d["cy"])
    H, W, bw = d["H"], d["W"], d.get("bw", 16)
    _, depths = oracle.project_points(d["means"], d["viewmat"], intr)
    centers, extents = oracle.get_aabb_2d(d["means"], d["scales"], 1.0, d["quats"], d["viewmat"], intr)
    nth = oracle.get_num_tiles_hit_2d(centers, extents, H, W, bw)
    n = d["means"].shape[0]
    return R.texture_edit((n, 1, channels), d["texture_dims"], e["updated_img"], e["updated_alpha"], e["depth_lower"],
                          e["depth_upper"], centers, extents, depths, nth, d["opacities"], d["means"], d["scales"], 1.0,
                          d["quats"], d["uv0"], d["umap"], d["vmap"], d["viewmat"], d["c2w"], *intr, H, W, bw, settings,
                          return_pairs=return_pairs)

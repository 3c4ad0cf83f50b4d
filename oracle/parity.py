"""Full-size parity of the product path against the reference's own CUDA extension (TEST INFRASTRUCTURE).

`compare_view(scene, upstream, device, view)` runs one camera view of a scene through
  * gstex_b200's public entry points (project_bounds -> bin_gaussians_to_tiles -> texture_gaussians + autograd), and
  * the reference's unmodified kernels (oracle/_ref/gstex_ref_C.so built by oracle/build_ref.py) driven exactly as the
    reference's Python drives them (gstex_cuda/get_aabb_2d.py:22-91, utils.py:40-162, texture.py:186-378),
on the same tensors and the same GPU, and returns the error report the north-star bars are read from:
tile/intersection lists, sort order and the two index images (`final_idx`, `depth_idx`, texture.cu:837-863) compared
bit for bit, the six images at 1e-5 and the nine gradients at 1e-4 relative with the share of elements outside.

Used by tests/test_gpu_fullsize.py and by bench.py's `reference_cuda_same_gpu` leg (outside the timed region).
Never imported by the product.
"""
from __future__ import annotations

IMG_KEYS = ["out_img", "out_depth", "out_reg", "out_alpha", "out_texture", "out_normal"]
GRAD_KEYS = ["v_colors", "v_opacity", "v_means", "v_scales", "v_quats", "v_uv0", "v_umap", "v_vmap", "v_texture"]
PARAM_KEYS = ["colors", "opacities", "means", "scales", "quats", "uv0", "umap", "vmap", "texture"]
EXACT_KEYS = ["depths", "centers", "extents", "num_tiles_hit", "gaussian_ids_sorted", "tile_bins", "final_idx", "depth_idx"]
IMG_RTOL, GRAD_RTOL = 1e-5, 1e-4


def _err(a, b, rtol, atol_abs=0.0):
    """(max |a-b| / scale, share of elements with |a-b| > rtol*(|b| + scale) + atol_abs, scale); scale = max |b|."""
    import torch

    a, b = a.double().reshape(-1), b.double().reshape(-1)
    if b.numel() == 0:
        return 0.0, 0.0, 0.0
    scale = float(b.abs().max())
    d = (a - b).abs()
    bad = d > rtol * (b.abs() + scale) + atol_abs + 1e-30
    rel = float(d.max()) / scale if scale > 0 else float(a.abs().max())
    return rel, float(bad.double().mean()), scale


def run_ours_view(sd, v, settings=None):
    """One view through the public API; returns every tensor the comparison reads (all on the device)."""
    import torch

    import gstex_b200 as G
    from gstex_b200 import texture as T

    intr = (sd["fx"], sd["fy"], sd["cx"], sd["cy"])
    H, W, bw = sd["H"], sd["W"], sd.get("block_width", 16)
    settings = sd["settings"] if settings is None else settings
    vm, c2w = sd["viewmat"].contiguous(), sd["c2w"].contiguous()
    _, depths, centers, extents, nth = G.project_bounds(sd["means"], sd["scales"], 1, sd["quats"], vm, intr, H, W, bw)
    tb = ((W + bw - 1) // bw, (H + bw - 1) // bw, 1)
    num_intersects, ids_sorted, tile_bins = G.bin_gaussians_to_tiles(sd["N"], centers, extents, depths, nth, tb, bw)
    res = dict(depths=depths, centers=centers, extents=extents, num_tiles_hit=nth, gaussian_ids_sorted=ids_sorted,
               tile_bins=tile_bins, num_intersects=num_intersects)
    # the index images are not part of texture_gaussians' return value: take them from the same forward call it makes
    outs9, _, _ = T._raster_forward(sd["texture_info"], sd["texture_dims"], ids_sorted, tile_bins, sd["colors"],
                                    sd["opacities"], sd["means"], sd["scales"], 1, sd["quats"], sd["uv0"], sd["umap"],
                                    sd["vmap"], sd["texture"], vm, c2w, *intr, H, W, bw, settings, sd["background"],
                                    max(num_intersects, 0))
    res["final_idx"], res["depth_idx"] = outs9[6], outs9[7]
    leaves = {k: sd[k].detach().clone().requires_grad_(v is not None) for k in PARAM_KEYS}
    outs = G.texture_gaussians(sd["texture_info"], sd["texture_dims"], centers, extents, depths, nth, leaves["colors"],
                               leaves["opacities"], leaves["means"], leaves["scales"], 1, leaves["quats"], leaves["uv0"],
                               leaves["umap"], leaves["vmap"], leaves["texture"], vm, c2w, *intr, H, W, bw, settings,
                               sd["background"])
    res.update(zip(IMG_KEYS, [o.detach() for o in outs]))
    # the forward the public call ran must be the one whose indices were read above
    assert torch.equal(outs9[0], res["out_img"]), "two forwards of the same view differ"
    if v is not None:
        torch.autograd.backward(list(outs), [v["v_" + k] for k in IMG_KEYS])
        for k, gk in zip(PARAM_KEYS, GRAD_KEYS):
            res[gk] = leaves[k].grad
    return res


def run_reference_view(C, sd, v, settings=None):
    """The same view through the reference's kernels + the torch ops its Python wraps around them."""
    import torch

    from . import ref_ext

    intr = (sd["fx"], sd["fy"], sd["cx"], sd["cy"])
    H, W, bw = sd["H"], sd["W"], sd.get("block_width", 16)
    vm3 = sd["viewmat"][:3, :].contiguous()
    _, depths = ref_ext.ref_project_points(sd["means"], vm3, intr)
    depths = depths.contiguous()
    centers, extents = C.get_aabb_2d(sd["means"], sd["scales"], 1.0, sd["quats"], vm3, *intr)
    nth = ref_ext.ref_get_num_tiles_hit_2d(centers, extents, H, W, bw).to(torch.int32)
    res = ref_ext.ref_forward_backward(C, sd, centers, extents, depths, nth, v, settings=settings)
    res.update(depths=depths, centers=centers, extents=extents, num_tiles_hit=nth)
    return res


def report(ours, ref, settings, have_grads=True):
    """Error report: exact mismatch counts, per-tensor max error / outlier share, and the four headline numbers."""
    import torch

    rep = {"exact": {}, "images": {}, "grads": {}, "num_intersects": int(ref["num_intersects"])}
    for k in EXACT_KEYS:
        a, b = ours[k], ref[k]
        if a.shape != b.shape:
            rep["exact"][k] = int(max(a.numel(), b.numel()))
            continue
        if a.dtype.is_floating_point:  # bit for bit (NaN-safe, -0.0 != 0.0)
            a, b = a.contiguous().view(torch.int32), b.contiguous().view(torch.int32)
        rep["exact"][k] = int((a != b.to(a.dtype)).sum())
    d2 = 1.0 if settings & (1 << 10) else max(1.0, float(ref["out_depth"].abs().max()) ** 2)
    for k in IMG_KEYS:
        rel, frac, scale = _err(ours[k], ref[k], IMG_RTOL, 2e-6 * d2 if k == "out_reg" else 0.0)
        a, b = ours[k].contiguous().view(torch.int32), ref[k].contiguous().view(torch.int32)
        rep["images"][k] = {"max_rel": rel, "outside": frac, "scale": scale, "bit_mismatch": int((a != b).sum())}
    if have_grads:
        for k in GRAD_KEYS:
            rel, frac, scale = _err(ours[k], ref[k], GRAD_RTOL)
            rep["grads"][k] = {"max_rel": rel, "outside": frac, "scale": scale}
    rep["img_max_rel"] = max(x["max_rel"] for x in rep["images"].values())
    rep["grad_max_rel"] = max([x["max_rel"] for x in rep["grads"].values()] or [0.0])
    rep["idx_mismatch"] = sum(rep["exact"][k] for k in ("gaussian_ids_sorted", "tile_bins", "final_idx", "depth_idx"))
    rep["bin_mismatch"] = sum(rep["exact"][k] for k in ("depths", "centers", "extents", "num_tiles_hit"))
    rep["fwd_bit_mismatch"] = sum(x["bit_mismatch"] for x in rep["images"].values())
    rep["outliers"] = {"img": max(x["outside"] for x in rep["images"].values()),
                       "grad": max([x["outside"] for x in rep["grads"].values()] or [0.0])}
    return rep


def compare_view(s, v, device, view=0, settings=None):
    """`s`: scene dict of gstex_b200.synthetic.make_scene (any device); `v`: upstream gradients or None.
    Returns the report, or None when the reference extension is not built."""
    import torch

    from . import ref_ext

    C = ref_ext.load()
    if C is None:
        return None
    dev = torch.device(device)
    sd = {k: (x.to(dev) if isinstance(x, torch.Tensor) else x) for k, x in s.items()}
    if "viewmats" in sd:
        sd["viewmat"], sd["c2w"] = sd["viewmats"][view].contiguous(), sd["c2ws"][view].contiguous()
    vd = None if v is None else {k: x.to(dev) for k, x in v.items()}
    st = s["settings"] if settings is None else settings
    ours = run_ours_view(sd, vd, settings)
    ref = run_reference_view(C, sd, vd, settings)
    return report(ours, ref, st, have_grads=v is not None)

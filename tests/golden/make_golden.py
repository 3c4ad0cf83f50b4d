"""Generate golden vectors by running the REFERENCE's own Python on the CPU (this container only).

    python tests/golden/make_golden.py

Imports `/root/reference/submodules/GStex_cuda/gstex_cuda/{_torch_impl,get_aabb_2d}.py` unchanged and
records, for small seeded scenes, the inputs and what the reference computes:
  * `project_points` (depths, xys), `get_num_tiles_hit_2d`, `get_aabb_2d_torch`   (get_aabb_2d.py:22-91)
  * `torch.sort` order of the intersection keys                                    (utils.py:159-160)
  * `_torch_impl.texture_forward` outputs and `torch.autograd` gradients           (_torch_impl.py:196-382)
The reference's binning kernels are CUDA-only, so the sorted id list / tile bins fed to its torch
rasterizer come from the oracle (they are stored too, and pinned on the GPU box against the reference
extension).  Scenes are chosen inside the region where the torch rasterizer and the CUDA kernels share
semantics: opacity <= 0.98, no blur / NDC bits, nothing near the camera plane,
square image (the torch rasterizer indexes (x,y) into an (H,W) buffer), and -- for the uv-gradient
case (settings bit 8) -- a uv map shrunk so that no contributing sample is clamped (the CUDA backward
passes gradients through the clamp, autograd does not).
The vectors cannot be regenerated on the GPU box (no /root/reference there); they are committed.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference/submodules/GStex_cuda")

import gstex_cuda._torch_impl as T  # noqa: E402
from gstex_cuda.get_aabb_2d import get_aabb_2d_torch, get_num_tiles_hit_2d, project_points  # noqa: E402

import oracle  # noqa: E402
from gstex_b200.synthetic import make_scene, make_upstream_grads  # noqa: E402

CASES = [
    dict(tag="slab_s0", name="cfg1", n=160, size=40, texels=160 * 16, settings=0, uv_shrink=1.0, seed=11),
    dict(tag="slab_s256", name="cfg1", n=160, size=40, texels=160 * 16, settings=1 << 8, uv_shrink=0.8, seed=12),
    dict(tag="shell_s0", name="cfg2", n=1500, size=48, texels=1500 * 9, settings=0, uv_shrink=1.0, seed=13),
    # settings bit 2 (nearest texel) is not implemented by the reference torch rasterizer: no golden for it
]
TENSOR_KEYS = ["means", "scales", "quats", "opacities", "colors", "texture_dims", "texture", "uv0", "umap", "vmap",
               "viewmat", "c2w", "background"]


def run_case(c):
    s = make_scene(c["name"], seed=c["seed"], n=c["n"], width=c["size"], height=c["size"], texels=c["texels"],
                   settings=c["settings"])
    s["umap"] = s["umap"] * c["uv_shrink"]
    s["vmap"] = s["vmap"] * c["uv_shrink"]
    H = W = c["size"]
    bw = 16
    intr = (s["fx"], s["fy"], s["cx"], s["cy"])
    out = {k: s[k].numpy() for k in TENSOR_KEYS}
    out["scalars"] = np.array([s["fx"], s["fy"], s["cx"], s["cy"], H, W, bw, s["settings"], s["C"]], np.float64)
    # ---- reference binning pieces that run on the CPU
    xys, depths = project_points(s["means"], s["viewmat"][:3, :], intr)
    out["ref_xys"], out["ref_depths"] = xys.numpy(), depths.numpy()
    ct, et = get_aabb_2d_torch(s["means"], s["scales"], 1.0, s["quats"], s["viewmat"][:3, :], intr)
    out["ref_aabb_torch_centers"], out["ref_aabb_torch_extents"] = ct.numpy(), et.numpy()
    centers, extents = oracle.get_aabb_2d(out["means"], out["scales"], 1.0, out["quats"], out["viewmat"], intr)
    out["centers"], out["extents"] = centers, extents
    nth = get_num_tiles_hit_2d(torch.from_numpy(centers), torch.from_numpy(extents), H, W, bw)
    out["ref_num_tiles_hit"] = nth.numpy().astype(np.int32)
    cum = torch.cumsum(nth, dim=0, dtype=torch.int32)
    num_intersects = int(cum[-1])
    tb = oracle.tile_bounds_of(H, W, bw)
    isect, gids = oracle.map_gaussian_to_intersects(c["n"], num_intersects, centers, extents, out["ref_depths"],
                                                    cum.numpy(), tb, bw)
    ks, idx = torch.sort(torch.from_numpy(isect), stable=True)
    out["isect_ids"], out["gaussian_ids"] = isect, gids
    out["ref_isect_ids_sorted"] = ks.numpy()
    out["ref_gaussian_ids_sorted"] = torch.gather(torch.from_numpy(gids), 0, idx).numpy()
    bins = oracle.get_tile_bin_edges(num_intersects, out["ref_isect_ids_sorted"], tb)
    out["tile_bins"] = bins
    # ---- reference torch rasterizer + autograd
    leaf = {k: s[k].clone().requires_grad_(True) for k in
            ["colors", "opacities", "means", "scales", "quats", "uv0", "umap", "vmap", "texture"]}
    outs = T.texture_forward(tb, (bw, bw, 1), (W, H, 1), s["texture_info"], s["texture_dims"],
                             torch.from_numpy(out["ref_gaussian_ids_sorted"]).long(), torch.from_numpy(bins),
                             leaf["colors"], leaf["opacities"], leaf["means"], leaf["scales"], 1.0, leaf["quats"],
                             leaf["uv0"], leaf["umap"], leaf["vmap"], leaf["texture"], s["viewmat"], s["c2w"], *intr,
                             s["settings"], s["background"])
    for nm, o in zip(["out_img", "out_depth", "out_reg", "out_texture", "out_normal", "final_Ts"], outs):
        out["ref_" + nm] = o.detach().numpy().copy()
    v = make_upstream_grads(s, seed=c["seed"] + 100)
    loss = sum((o * v[k]).sum() for o, k in
               zip(outs[:5], ["v_out_img", "v_out_depth", "v_out_reg", "v_out_texture", "v_out_normal"]))
    loss.backward()
    for k, x in v.items():
        out[k] = x.numpy()
    for k, gk in [("colors", "v_colors"), ("opacities", "v_opacity"), ("means", "v_means"), ("scales", "v_scales"),
                  ("quats", "v_quats"), ("uv0", "v_uv0"), ("umap", "v_umap"), ("vmap", "v_vmap"),
                  ("texture", "v_texture")]:
        g = leaf[k].grad
        out["ref_" + gk] = (g if g is not None else torch.zeros_like(leaf[k])).numpy().copy()
    return out


if __name__ == "__main__":
    for c in CASES:
        data = run_case(c)
        path = os.path.join(HERE, f"golden_{c['tag']}.npz")
        np.savez_compressed(path, **data)
        print(path, os.path.getsize(path) // 1024, "KiB", "I =", len(data["isect_ids"]))

"""GPU parity, stages (a)+(b): projection, bounds, tile counts, scan, key emit, radix sort, tile ranges.
Bit-exact against the CPU oracle, the committed golden vectors, torch's own ops on the device (the
third-party arithmetic the reference calls) and -- when oracle/_ref holds it -- the reference's own
CUDA extension."""
import numpy as np
import pytest
import torch

from common import golden_files, load_golden, scene_to_numpy

pytestmark = pytest.mark.gpu

SCENES = [
    dict(name="cfg1", n=3000, width=200, height=136),          # ragged tiles, huge splats, many tiles per splat
    dict(name="cfg2", n=20000, width=320, height=320),
    dict(name="cfg4", n=50000, width=480, height=270),
]


def _dev(s):
    return {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in s.items()}


def _scene(cfg, behind_camera=False):
    from gstex_b200.synthetic import make_scene

    s = make_scene(cfg["name"], n=cfg["n"], width=cfg["width"], height=cfg["height"])
    if behind_camera:  # push a slice of the splats behind / onto the camera plane: clipped, zero-key quirk
        vm = s["viewmat"]
        cam_pos = s["c2w"][:3, 3]
        fwd = vm[2, :3]
        k = max(8, cfg["n"] // 50)
        s["means"][:k] = cam_pos + fwd * torch.linspace(-0.5, 0.02, k)[:, None] + 0.05 * torch.randn(k, 3)
        s["means"][0] = cam_pos - fwd * 0.3  # splat 0 itself clipped
    return s


def _bits(t):
    return t.detach().cpu().numpy().view(np.int32)


@pytest.mark.parametrize("cfg", SCENES, ids=lambda c: c["name"])
@pytest.mark.parametrize("behind", [False, True], ids=["plain", "clipped"])
def test_binning_matches_oracle_and_torch(cfg, behind):
    import gstex_b200 as G
    import oracle

    s = _scene(cfg, behind)
    d = scene_to_numpy(s)
    sd = _dev(s)
    intr = (s["fx"], s["fy"], s["cx"], s["cy"])
    H, W, bw = s["H"], s["W"], 16
    # --- kernel under test
    xys, depths = G.project_points(sd["means"], sd["viewmat"][:3, :], intr)
    centers, extents = G.get_aabb_2d(sd["means"], sd["scales"], 1, sd["quats"], sd["viewmat"][:3, :], intr)
    nth = G.get_num_tiles_hit_2d(centers, extents, H, W, bw)
    fused = G.project_bounds(sd["means"], sd["scales"], 1, sd["quats"], sd["viewmat"], intr, H, W, bw)
    for a, b in zip(fused, (xys, depths, centers, extents, nth)):
        assert torch.equal(a, b)
    # --- oracle
    o_xys, o_depths = oracle.project_points(d["means"], d["viewmat"], intr)
    o_c, o_e = oracle.get_aabb_2d(d["means"], d["scales"], 1.0, d["quats"], d["viewmat"], intr)
    o_nth = oracle.get_num_tiles_hit_2d(o_c, o_e, H, W, bw)
    assert np.array_equal(_bits(depths), o_depths.view(np.int32))
    assert np.array_equal(_bits(xys), o_xys.view(np.int32))
    assert np.array_equal(_bits(centers), o_c.view(np.int32))
    assert np.array_equal(_bits(extents), o_e.view(np.int32))
    assert np.array_equal(nth.cpu().numpy(), o_nth)
    # --- torch on the device = what the reference's Python computes (get_aabb_2d.py:22-32, 70-91)
    from oracle.ref_ext import ref_get_num_tiles_hit_2d, ref_project_points

    t_xys, t_depths = ref_project_points(sd["means"], sd["viewmat"][:3, :], intr)
    assert torch.equal(depths, t_depths), "depth bits differ from torch's matmul on this GPU"
    assert torch.equal(xys, t_xys)
    assert torch.equal(nth, ref_get_num_tiles_hit_2d(centers, extents, H, W, bw).to(torch.int32))
    # --- scan / emit / sort / tile ranges
    total, cum = G.compute_cumulative_intersects(nth)
    assert torch.equal(cum, torch.cumsum(nth, 0, dtype=torch.int32)) and total == int(cum[-1])
    tb = oracle.tile_bounds_of(H, W, bw)
    isect, gids, isect_s, gids_s, bins = G.bin_and_sort_gaussians(s["N"], total, centers, extents, depths, cum, tb, bw)
    o = oracle.bin_and_sort_gaussians(s["N"], total, o_c, o_e, o_depths, cum.cpu().numpy(), tb, bw)
    for a, b, nm in zip((isect, gids, isect_s, gids_s, bins), o, ("isect", "gids", "isect_sorted", "gids_sorted", "bins")):
        assert np.array_equal(a.cpu().numpy(), b), nm
    # torch.sort + gather (utils.py:159-160)
    ts, ti = torch.sort(isect)
    assert torch.equal(isect_s, ts) and torch.equal(gids_s, torch.gather(gids, 0, ti))
    if behind:
        assert int((isect == 0).sum()) > 0, "scene should exercise the zero-key quirk"


@pytest.mark.parametrize("cfg", SCENES, ids=lambda c: c["name"])
@pytest.mark.parametrize("behind", [False, True], ids=["plain", "clipped"])
def test_binning_matches_reference_extension(cfg, behind):
    from oracle import ref_ext

    C = ref_ext.load()
    if C is None:
        pytest.skip("oracle/_ref/gstex_ref_C.so not built (python oracle/build_ref.py)")
    import gstex_b200 as G

    s = _scene(cfg, behind)
    sd = _dev(s)
    intr = (s["fx"], s["fy"], s["cx"], s["cy"])
    H, W, bw = s["H"], s["W"], 16
    centers, extents = G.get_aabb_2d(sd["means"], sd["scales"], 1, sd["quats"], sd["viewmat"][:3, :], intr)
    r_c, r_e = C.get_aabb_2d(sd["means"], sd["scales"], 1.0, sd["quats"], sd["viewmat"][:3, :].contiguous(), *intr)
    assert torch.equal(centers, r_c), f"{int((centers != r_c).sum())} center floats differ from the reference kernel"
    assert torch.equal(extents, r_e)
    _, depths = G.project_points(sd["means"], sd["viewmat"][:3, :], intr)
    nth = G.get_num_tiles_hit_2d(centers, extents, H, W, bw)
    total, cum = G.compute_cumulative_intersects(nth)
    tb = ((W + bw - 1) // bw, (H + bw - 1) // bw, 1)
    mine = G.bin_and_sort_gaussians(s["N"], total, centers, extents, depths, cum, tb, bw)
    ref = ref_ext.ref_bin_and_sort(C, s["N"], total, centers, extents, depths, cum, tb, bw)
    for a, b, nm in zip(mine, ref, ("isect", "gids", "isect_sorted", "gids_sorted", "bins")):
        assert torch.equal(a, b), nm


@pytest.mark.parametrize("path", golden_files(), ids=lambda p: p.split("golden_")[-1][:-4])
def test_binning_matches_golden(path):
    import gstex_b200 as G

    d = load_golden(path)
    t = {k: torch.from_numpy(d[k]).cuda() for k in ["means", "scales", "quats", "viewmat"]}
    intr = (d["fx"], d["fy"], d["cx"], d["cy"])
    xys, depths = G.project_points(t["means"], t["viewmat"][:3, :], intr)
    assert np.array_equal(_bits(depths), d["ref_depths"].view(np.int32))
    assert np.array_equal(_bits(xys), d["ref_xys"].view(np.int32))
    centers, extents = G.get_aabb_2d(t["means"], t["scales"], 1, t["quats"], t["viewmat"], intr)
    assert np.array_equal(centers.cpu().numpy(), d["centers"]) and np.array_equal(extents.cpu().numpy(), d["extents"])
    nth = G.get_num_tiles_hit_2d(centers, extents, d["H"], d["W"], d["bw"])
    assert np.array_equal(nth.cpu().numpy(), d["ref_num_tiles_hit"])
    total, cum = G.compute_cumulative_intersects(nth)
    tb = ((d["W"] + 15) // 16, (d["H"] + 15) // 16, 1)
    isect, gids, isect_s, gids_s, bins = G.bin_and_sort_gaussians(len(d["means"]), total, centers, extents, depths,
                                                                  cum, tb, d["bw"])
    assert np.array_equal(isect_s.cpu().numpy(), d["ref_isect_ids_sorted"])
    assert np.array_equal(gids_s.cpu().numpy(), d["ref_gaussian_ids_sorted"])
    assert np.array_equal(bins.cpu().numpy(), d["tile_bins"])


@pytest.mark.parametrize("n", [0, 1, 31, 4096, 4097, 1_000_003])
def test_cumsum_sizes(n):
    import gstex_b200 as G

    g = torch.Generator().manual_seed(n)
    x = torch.randint(0, 7, (n,), generator=g, dtype=torch.int32).cuda()
    total, cum = G.compute_cumulative_intersects(x)
    assert torch.equal(cum, torch.cumsum(x, 0, dtype=torch.int32))
    assert total == (int(x.sum()) if n else 0)


@pytest.mark.parametrize("n,bits", [(1, 64), (255, 40), (4096, 44), (4097, 45), (300_001, 45), (2_500_000, 46), (70_000, 64)])
def test_radix_sort_stable(n, bits):
    """Sortedness + stability + permutation (size-independent properties) and equality with torch.sort."""
    import gstex_b200 as G

    g = torch.Generator().manual_seed(n)
    hi = torch.randint(0, 1 << min(max(bits - 32, 1), 31), (n,), generator=g, dtype=torch.int64) if bits > 32 else torch.zeros(n, dtype=torch.int64)
    lo = torch.randint(0, 1 << 20, (n,), generator=g, dtype=torch.int64) * 37  # many duplicate keys
    keys = ((hi << 32) | (lo & 0x7FFFFFFF)).cuda()
    vals = torch.arange(n, dtype=torch.int32).cuda()
    ks, vs = G.sort_isects(keys, vals, bits)
    ts, ti = torch.sort(keys, stable=True)
    assert torch.equal(ks, ts)
    assert torch.equal(vs, vals[ti]), "payload order differs from a stable sort"
    assert torch.equal(keys, keys.clone()) and bool((ks[1:] >= ks[:-1]).all())


def _general_and_fast(centers, extents, depths, nth, N, H, W, bw):
    import gstex_b200 as G
    from gstex_b200.utils import bin_gaussians_to_tiles

    tb = ((W + bw - 1) // bw, (H + bw - 1) // bw, 1)
    total, cum = G.compute_cumulative_intersects(nth)
    I, ids_f, bins_f = bin_gaussians_to_tiles(N, centers, extents, depths, nth, tb, bw)
    assert I == total
    if total < 1:
        assert ids_f.numel() == 0 and int(bins_f.abs().sum()) == 0
        return None
    _, _, isect_s, ids_g, bins_g = G.bin_and_sort_gaussians(N, total, centers, extents, depths, cum, tb, bw)
    assert torch.equal(bins_f, bins_g), "tile_bins of the tile-bucket path differ from the general path"
    assert torch.equal(ids_f, ids_g), "sorted ids of the tile-bucket path differ from the general path"
    return isect_s, ids_g, bins_g


@pytest.mark.parametrize("cfg", SCENES, ids=lambda c: c["name"])
@pytest.mark.parametrize("behind", [False, True], ids=["plain", "clipped"])
@pytest.mark.parametrize("bw", [16, 7, 2])
def test_tile_bucket_path_matches_general(cfg, behind, bw):
    """texture_gaussians' binning (csrc/tile_bucket.cu) == emit + stable radix sort + bin edges, bit for bit."""
    import gstex_b200 as G

    if bw == 2 and cfg["name"] != "cfg1":
        pytest.skip("block width 2 only on the small scene (tile count)")
    s = _scene(cfg, behind)
    sd = _dev(s)
    intr = (s["fx"], s["fy"], s["cx"], s["cy"])
    H, W = s["H"], s["W"]
    _, depths, centers, extents, nth = G.project_bounds(sd["means"], sd["scales"], 1, sd["quats"], sd["viewmat"], intr,
                                                        H, W, bw)
    r = _general_and_fast(centers, extents, depths, nth, s["N"], H, W, bw)
    assert r is not None
    if behind:
        assert int((r[0] == 0).sum()) > 0, "scene should exercise the zero-key quirk"


def test_tile_bucket_ties_long_lists_and_empty():
    """Equal depths (ties resolved by splat index), every splat on every tile (long lists), no intersections."""
    g = torch.Generator().manual_seed(5)
    H, W, bw = 64, 96, 16
    for n, ext, depth_vals in ((5000, 200.0, 3), (30000, 40.0, 2), (100, 0.0, 1), (7000, 200.0, 1 << 20),
                              (12000, 200.0, 7), (2048, 200.0, 1 << 20), (2049, 200.0, 1 << 20), (600, 200.0, 40)):
        centers = (torch.rand(n, 2, generator=g) * torch.tensor([W, H])).cuda()
        extents = torch.full((n, 2), ext).cuda()
        depths = (torch.randint(0, depth_vals, (n,), generator=g).float() * 0.25 + 1.0).cuda()
        import gstex_b200 as G

        nth = G.get_num_tiles_hit_2d(centers, extents, H, W, bw)
        if ext == 0.0:
            nth = torch.zeros_like(nth)
        r = _general_and_fast(centers, extents, depths, nth, n, H, W, bw)
        if r is not None:
            keys = r[0]
            assert bool((keys[1:] >= keys[:-1]).all())


def test_tile_bucket_falls_back_beyond_capacity():
    """A tile list longer than the shared-memory capacity takes the general path (same results)."""
    import gstex_b200 as G
    from gstex_b200 import _lib

    cap = _lib.load().gstex_tile_bucket_max_list()
    n = cap + 1500
    g = torch.Generator().manual_seed(9)
    H, W, bw = 32, 32, 16
    centers = (torch.rand(n, 2, generator=g) * 32).cuda()
    extents = torch.full((n, 2), 64.0).cuda()
    depths = (torch.rand(n, generator=g) + 1.0).cuda()
    nth = G.get_num_tiles_hit_2d(centers, extents, H, W, bw)
    assert int(nth.max()) == 4
    assert _general_and_fast(centers, extents, depths, nth, n, H, W, bw) is not None


def test_tile_bucket_full_size_properties():
    """cfg5-sized scene (1M splats, 800x800): per-tile lists sorted by (depth, id), ranges tile the list."""
    import gstex_b200 as G
    from gstex_b200.synthetic import make_scene
    from gstex_b200.utils import bin_gaussians_to_tiles

    s = make_scene("cfg5", views=1)
    sd = _dev(s)
    intr = (s["fx"], s["fy"], s["cx"], s["cy"])
    H, W, bw = s["H"], s["W"], 16
    _, depths, centers, extents, nth = G.project_bounds(sd["means"], sd["scales"], 1, sd["quats"], sd["viewmats"][0],
                                                        intr, H, W, bw)
    tb = ((W + bw - 1) // bw, (H + bw - 1) // bw, 1)
    I, ids, bins = bin_gaussians_to_tiles(s["N"], centers, extents, depths, nth, tb, bw)
    assert I == int(nth.sum()) and ids.numel() == I
    cnt = (bins[:, 1] - bins[:, 0]).long()
    assert int(cnt.sum()) == I
    ne = cnt > 0
    starts = torch.cumsum(cnt, 0) - cnt
    assert torch.equal(bins[ne, 0].long(), starts[ne])
    tile_of = torch.repeat_interleave(torch.arange(bins.shape[0], device=ids.device), cnt)
    key = (tile_of << 32) | depths.view(torch.int32)[ids.long()].long()
    assert bool((key[1:] >= key[:-1]).all()), "list not sorted by (tile, depth)"
    same = key[1:] == key[:-1]
    assert bool((ids[1:][same] > ids[:-1][same]).all()), "ties must keep ascending splat index"
    _general_and_fast(centers, extents, depths, nth, s["N"], H, W, bw)

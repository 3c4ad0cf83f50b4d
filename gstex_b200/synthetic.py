"""Seeded synthetic scenes shaped like the workloads of BASELINE.json.

No dataset or checkpoint is available offline, so every benchmark and parity test runs on these.
The generators follow the parameterisation the reference's model hands to the rasterizer
(`nerfstudio/models/gstex.py:1059-1069,975-990`: flat disks with scales[:,2] = 1e-5*mean(scales[:, :2]),
unit quaternions, uv0 = 0.5, umap/vmap = axes / (2*3*scale); chart sizes by the bisection of
`build_charts`, `gstex.py:841-880`) and the demo of the submodule (`example.py:69-119`).
Everything is generated on the CPU with a seeded `torch.Generator` (bit-reproducible across
machines) and then moved to `device`.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch

SIGMA_FACTOR = 3.0  # GStexModelConfig.sigma_factor


def _rotmat(quats: torch.Tensor) -> torch.Tensor:
    w, x, y, z = quats.unbind(-1)
    return torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
        2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
        2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], -1).reshape(-1, 3, 3)


def _quat_from_normal(normal: torch.Tensor, spin: torch.Tensor) -> torch.Tensor:
    """Unit quaternion (w,x,y,z) whose third rotation axis is `normal`, with in-plane angle `spin`."""
    zaxis = torch.tensor([0.0, 0.0, 1.0]).expand_as(normal)
    n = torch.nn.functional.normalize(normal, dim=-1)
    axis = torch.cross(zaxis, n, dim=-1)
    s = axis.norm(dim=-1, keepdim=True).clamp_min(1e-8)
    c = (zaxis * n).sum(-1, keepdim=True).clamp(-1, 1)
    half = 0.5 * torch.atan2(s, c)
    q1 = torch.cat([torch.cos(half), torch.sin(half) * axis / s], -1)  # z -> n
    q2 = torch.cat([torch.cos(0.5 * spin)[:, None], torch.zeros(len(spin), 2), torch.sin(0.5 * spin)[:, None]], -1)
    w1, x1, y1, z1 = q1.unbind(-1)
    w2, x2, y2, z2 = q2.unbind(-1)
    q = torch.stack([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                     w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2], -1)
    return torch.nn.functional.normalize(q, dim=-1)


def chart_dims_for_budget(len0: torch.Tensor, len1: torch.Tensor, total_texels: int) -> torch.Tensor:
    """(N,3) int32 {h, w, start}: the `build_charts` bisection on the texel size (gstex.py:841-880)."""
    n = len0.shape[0]
    if total_texels <= n:  # "pixel-num 0" plain-2DGS variant: every splat a 1x1 chart
        h = torch.ones(n, dtype=torch.int64)
        w = torch.ones(n, dtype=torch.int64)
    else:
        l0, l1 = len0.double(), len1.double()

        def score(x: float) -> float:
            return float((torch.ceil(SIGMA_FACTOR * l0 / x) * torch.ceil(SIGMA_FACTOR * l1 / x)).sum())

        hi = math.sqrt(float((SIGMA_FACTOR * SIGMA_FACTOR * l0 * l1).sum()) / total_texels)
        lo = 0.0
        while score(hi) > total_texels:  # ceil() only adds texels: grow the texel until under budget
            lo, hi = hi, hi * 2
        mid = hi
        for _ in range(60):
            mid = 0.5 * (lo + hi)
            s = score(mid)
            if abs(s - total_texels) <= 1e-3 * total_texels:
                break
            if s > total_texels:
                lo = mid
            else:
                hi = mid
        h = torch.ceil(SIGMA_FACTOR * l0 / mid).long()
        w = torch.ceil(SIGMA_FACTOR * l1 / mid).long()
    hw = h * w
    start = torch.cumsum(hw, 0) - hw
    assert int(start[-1] + hw[-1]) < 2 ** 31
    return torch.stack([h, w, start], -1).to(torch.int32)


def look_at_viewmat(eye: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    """World-to-camera (4,4), camera looks down +z (gsplat/OpenCV convention, as `example.py:100-109`)."""
    fwd = torch.nn.functional.normalize(target - eye, dim=0)
    up = torch.tensor([0.0, 1.0, 0.0])
    if abs(float((fwd * up).sum())) > 0.99:
        up = torch.tensor([1.0, 0.0, 0.0])
    right = torch.nn.functional.normalize(torch.cross(up, fwd, dim=0), dim=0)
    down = torch.cross(fwd, right, dim=0)
    R = torch.stack([right, down, fwd], 0)  # rows: camera axes in world coords
    vm = torch.eye(4)
    vm[:3, :3] = R
    vm[:3, 3] = -R @ eye
    return vm


_CONFIGS = {
    # name: N, texels, W, H, fx, fy, cx, cy, distance, kind, settings
    "cfg1": dict(n=10_000, texels=160_000, W=256, H=256, fov_x=math.pi / 2, dist=8.0, kind="slab16",
                 settings=1 << 8),
    "cfg2": dict(n=300_000, texels=10_000_000, W=800, H=800, fx=1111.1, fy=1111.1, cx=400.0, cy=400.0, dist=4.03,
                 kind="shell", settings=(1 << 9) | (1 << 10)),
    "cfg2_2dgs": dict(n=300_000, texels=0, W=800, H=800, fx=1111.1, fy=1111.1, cx=400.0, cy=400.0, dist=4.03,
                      kind="shell", settings=(1 << 9) | (1 << 10)),
    "cfg3": dict(n=500_000, texels=10_000_000, W=1600, H=1200, fx=2892.3, fy=2892.3, cx=823.2, cy=619.1, dist=2.2,
                 kind="shell", shell_scale=0.33, settings=(1 << 9) | (1 << 10)),
    "cfg4": dict(n=3_000_000, texels=40_000_000, W=1920, H=1080, fov_x=math.pi / 3, dist=3.0, kind="slab",
                 settings=(1 << 9) | (1 << 10)),
    "cfg5": dict(n=1_000_000, texels=10_000_000, W=800, H=800, fx=1111.1, fy=1111.1, cx=400.0, cy=400.0, dist=4.03,
                 kind="shell", settings=(1 << 9) | (1 << 10)),
}


def config_names():
    return list(_CONFIGS)


def make_scene(name: str = "cfg2", seed: Optional[int] = None, device="cpu", n: Optional[int] = None,
               texels: Optional[int] = None, width: Optional[int] = None, height: Optional[int] = None,
               channels: int = 3, settings: Optional[int] = None, views: int = 1) -> Dict[str, object]:
    """Build one synthetic scene.  Optional overrides shrink a config for the parity tests (focal
    lengths scale with the image so the framing is kept).  `views` > 1 adds cameras rotated about the
    y axis (returned as `viewmats` / `c2ws` of shape (V,4,4); `viewmat` / `c2w` are view 0)."""
    cfg = dict(_CONFIGS[name])
    base_w, base_h = cfg["W"], cfg["H"]
    W = int(width or base_w)
    H = int(height or base_h)
    N = int(n or cfg["n"])
    T_budget = cfg["texels"] if texels is None else int(texels)
    if texels is None and n is not None and cfg["texels"] > 0:
        T_budget = max(N, int(cfg["texels"] * (N / cfg["n"])))
    if seed is None:
        seed = 1000 + list(_CONFIGS).index(name)
    g = torch.Generator().manual_seed(int(seed))

    def U(*shape):
        return torch.rand(*shape, generator=g)

    def Nrm(*shape):
        return torch.randn(*shape, generator=g)

    if "fov_x" in cfg:
        fx = fy = 0.5 * W / math.tan(0.5 * cfg["fov_x"])
        cx, cy = W / 2, H / 2
    else:
        sx, sy = W / base_w, H / base_h
        fx, fy, cx, cy = cfg["fx"] * sx, cfg["fy"] * sy, cfg["cx"] * sx, cfg["cy"] * sy

    kind = cfg["kind"]
    if kind == "slab16":  # example.py:69-89
        means = U(N, 3) - 0.5
        means[:, :2] *= 16
        scales2 = torch.exp(0.5 * math.log(1.0 / N) * U(N, 2))
        u, v, w = U(N), U(N), U(N)
        quats = torch.stack([torch.sqrt(1 - u) * torch.sin(2 * math.pi * v), torch.sqrt(1 - u) * torch.cos(2 * math.pi * v),
                             torch.sqrt(u) * torch.sin(2 * math.pi * w), torch.sqrt(u) * torch.cos(2 * math.pi * w)], -1)
        quats = torch.nn.functional.normalize(quats, dim=-1)
        opac = torch.sigmoid(-2 + 5 * U(N, 1)).clamp_max(0.98)
    elif kind == "shell":
        k = float(cfg.get("shell_scale", 1.0))
        dirs = torch.nn.functional.normalize(Nrm(N, 3), dim=-1)
        radius = (0.6 + 0.4 * U(N, 1)) * k
        means = dirs * radius
        scales2 = torch.exp(math.log(0.006 * k) + 0.5 * Nrm(N, 2))
        normal = torch.nn.functional.normalize(dirs + 0.3 * Nrm(N, 3), dim=-1)
        quats = _quat_from_normal(normal, 2 * math.pi * U(N))
        opac = torch.sigmoid(2 + 1.5 * Nrm(N, 1)).clamp_max(0.98)
    elif kind == "slab":  # deep slab: depth complexity >= 30
        means = torch.stack([(U(N) - 0.5) * 6.0, (U(N) - 0.5) * 3.4, (U(N) - 0.5) * 2.0], -1)
        scales2 = torch.exp(math.log(0.004) + 0.5 * Nrm(N, 2))
        normal = torch.nn.functional.normalize(torch.tensor([0.0, 0.0, -1.0]) + 0.5 * Nrm(N, 3), dim=-1)
        quats = _quat_from_normal(normal, 2 * math.pi * U(N))
        opac = torch.sigmoid(0.5 + 1.5 * Nrm(N, 1)).clamp_max(0.98)
    else:
        raise ValueError(kind)

    scales = torch.zeros(N, 3)
    scales[:, :2] = scales2.clamp_min(1e-9)
    scales[:, 2] = 1e-5 * scales[:, :2].mean(-1)  # gstex.py:1062-1064
    colors = U(N, 3)
    texture_dims = chart_dims_for_budget(scales[:, 0], scales[:, 1], T_budget)
    T = int(texture_dims[-1, 2] + texture_dims[-1, 0] * texture_dims[-1, 1])
    texture = U(T, channels)
    R = _rotmat(quats)
    uv0 = torch.full((N, 1, 2), 0.5)
    umap = (R[:, :, 0] / (2 * SIGMA_FACTOR * scales[:, 0:1])).unsqueeze(1)  # gstex.py:882-883,983-984
    vmap = (R[:, :, 1] / (2 * SIGMA_FACTOR * scales[:, 1:2])).unsqueeze(1)

    d = float(cfg["dist"])
    viewmats, c2ws = [], []
    for vi in range(views):
        ang = 2 * math.pi * vi / max(views, 1)
        jitter = (U(3) - 0.5) * 0.05 * d if vi > 0 else torch.zeros(3)
        eye = torch.tensor([d * math.sin(ang), 0.0, -d * math.cos(ang)]) + jitter
        vm = look_at_viewmat(eye, torch.zeros(3))
        viewmats.append(vm)
        c2ws.append(torch.linalg.inv(vm))
    viewmats_t, c2ws_t = torch.stack(viewmats), torch.stack(c2ws)

    dev = torch.device(device)
    scene = dict(
        name=name, seed=int(seed), N=N, T=T, C=int(channels), H=H, W=W, fx=float(fx), fy=float(fy), cx=float(cx),
        cy=float(cy), glob_scale=1.0, block_width=16,
        settings=int(cfg["settings"] if settings is None else settings),
        texture_info=(N, 1, int(channels)),
        means=means.float().contiguous().to(dev), scales=scales.float().contiguous().to(dev),
        quats=quats.float().contiguous().to(dev), opacities=opac.float().contiguous().to(dev),
        colors=colors.float().contiguous().to(dev), texture_dims=texture_dims.contiguous().to(dev),
        texture=texture.float().contiguous().to(dev), uv0=uv0.contiguous().to(dev),
        umap=umap.float().contiguous().to(dev), vmap=vmap.float().contiguous().to(dev),
        viewmat=viewmats_t[0].contiguous().to(dev), c2w=c2ws_t[0].contiguous().to(dev),
        viewmats=viewmats_t.contiguous().to(dev), c2ws=c2ws_t.contiguous().to(dev),
        background=torch.zeros(3).to(dev),
    )
    return scene


def make_upstream_grads(scene: Dict[str, object], seed: int = 7, device=None) -> Dict[str, torch.Tensor]:
    """Fixed upstream gradients of SURVEY.md section 8d: v_out_img, v_out_texture ~ N(0,1)/P; depth/reg/normal
    ~ N(0,1)*1e-3/P; v_out_alpha = 0."""
    H, W, C = scene["H"], scene["W"], scene["C"]
    P = H * W
    g = torch.Generator().manual_seed(int(seed))
    dev = torch.device(device) if device is not None else scene["means"].device
    out = dict(
        v_out_img=torch.randn(H, W, 3, generator=g) / P,
        v_out_depth=torch.randn(H, W, generator=g) * 1e-3 / P,
        v_out_reg=torch.randn(H, W, generator=g) * 1e-3 / P,
        v_out_alpha=torch.zeros(H, W),
        v_out_texture=torch.randn(H, W, C, generator=g) / P,
        v_out_normal=torch.randn(H, W, 3, generator=g) * 1e-3 / P,
    )
    return {k: v.float().contiguous().to(dev) for k, v in out.items()}

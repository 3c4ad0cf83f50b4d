"""World-size-2 gloo test (CPU) of the multi-GPU host logic: view sharding + gradient all-reduce.
The per-view "renderer" here is the CPU oracle on a tiny scene, so the test also checks that the sum of
per-rank gradient accumulations equals the single-process sum over all views."""
import os
import socket
import sys

import numpy as np
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GRAD_KEYS = ["v_colors", "v_opacity", "v_means", "v_scales", "v_quats", "v_uv0", "v_umap", "v_vmap", "v_texture"]
NUM_VIEWS = 4


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _view_grads(s, v, view):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from common import run_oracle_pipeline, scene_to_numpy

    d = scene_to_numpy(s)
    d["viewmat"] = s["viewmats"][view].numpy()
    d["c2w"] = s["c2ws"][view].numpy()
    r = run_oracle_pipeline(d, {k: x.numpy() for k, x in v.items()})
    return [torch.from_numpy(r[k].astype(np.float64)) for k in GRAD_KEYS]


def _scene():
    from gstex_b200.synthetic import make_scene, make_upstream_grads

    s = make_scene("cfg2", n=600, width=48, height=48, texels=6000, views=NUM_VIEWS)
    return s, make_upstream_grads(s)


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist

    from gstex_b200.multiview import allreduce_gradients, shard_views

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    s, v = _scene()
    acc = None
    for view in shard_views(NUM_VIEWS, rank, world):
        g = _view_grads(s, v, view)
        acc = g if acc is None else [a + b for a, b in zip(acc, g)]
    # the same sums through the flat bucket: autograd accumulates into views of one buffer, one all-reduce
    from gstex_b200.multiview import GradientBucket

    params = [torch.zeros_like(a).requires_grad_(True) for a in acc]
    import warnings

    with warnings.catch_warnings(record=True) as caught:  # no peer memory on the CPU: says so, then sums through the group
        warnings.simplefilter("always")
        bucket = GradientBucket(params, peer_group=dist.group.WORLD, peer_fallback=True)
    assert bucket.peer is None and any("peer-memory" in str(w.message) for w in caught)
    try:
        GradientBucket(params, peer_group=dist.group.WORLD)
        raise AssertionError("a peer-memory bucket on CPU tensors must fail loudly without peer_fallback")
    except ValueError:
        pass
    bucket.zero_()
    for view in shard_views(NUM_VIEWS, rank, world):
        g = _view_grads(s, v, view)
        torch.autograd.backward([(p * gi).sum() for p, gi in zip(params, g)])  # d/dp = gi, accumulated in place
    assert all(p.grad is v for p, v in zip(params, bucket.views))
    bucket.allreduce(world)
    allreduce_gradients(acc, world)
    for p, a in zip(params, acc):
        assert torch.allclose(p.grad, a, rtol=1e-12, atol=1e-14)
    torch.save(acc, os.path.join(out_dir, f"rank{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_views():
    sys.path.insert(0, ROOT)
    from gstex_b200.multiview import shard_views

    for world in (1, 2, 4, 8):
        seen = sorted(v for r in range(world) for v in shard_views(64, r, world))
        assert seen == list(range(64))
        assert all(len(shard_views(64, r, world)) == 64 // world for r in range(world))


def test_two_rank_gradient_allreduce(tmp_path):
    sys.path.insert(0, ROOT)
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    s, v = _scene()
    total = None
    for view in range(NUM_VIEWS):
        g = _view_grads(s, v, view)
        total = g if total is None else [a + b for a, b in zip(total, g)]
    for rank in range(world):
        got = torch.load(os.path.join(str(tmp_path), f"rank{rank}.pt"))
        for a, b, k in zip(got, total, GRAD_KEYS):
            assert torch.allclose(a, b, rtol=1e-12, atol=1e-14), k
    assert float(total[2].abs().max()) > 0  # the views do produce gradients

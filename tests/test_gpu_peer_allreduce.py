"""The exchange step of the view-sharded path as our own kernel over NVLink peer memory:
`GradientBucket(peer_group=...)` + `gstex_peer_allreduce_f32` against NCCL's all-reduce of the same buckets.
Needs at least two GPUs in one NVLink domain (skipped on a one-GPU box; run with `gpurun --gpus 2`)."""
import os
import socket
import sys

import pytest
import torch
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# odd sizes on purpose: slices that do not divide by the world size, a last rank with a short or empty slice
SHAPES = [(1000003, 3), (70001, 1), (40007, 3), (40007, 4), (5,), (1,)]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, multicast, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist

    from gstex_b200.multiview import GradientBucket

    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world, device_id=dev)
    params = [torch.zeros(sh, device=dev, requires_grad=True) for sh in SHAPES]
    try:
        bucket = GradientBucket(params, peer_group=dist.group.WORLD, multicast=multicast)
    except RuntimeError as exc:
        if multicast and "multicast" in str(exc):
            open(os.path.join(out_dir, f"skip{rank}"), "w").write(str(exc))
            dist.destroy_process_group()
            return
        raise
    assert bucket.peer is not None and bucket.flat.data_ptr() == bucket.peer.local.data_ptr()
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    for it in range(3):  # several steps: the barriers must keep one step's stores apart from the next step's zeroing
        bucket.zero_()
        for p in params:
            p.grad.copy_(torch.randn(p.shape, device=dev, generator=g) * (10.0 ** (it - 1)))
        expect = bucket.flat.clone()
        dist.all_reduce(expect, op=dist.ReduceOp.SUM)
        bucket.allreduce(world)
        got = bucket.flat
        if world == 2:
            assert torch.equal(got, expect), f"step {it}: {int((got != expect).sum())} elements differ from NCCL's sum"
        else:
            scale = float(expect.abs().max())
            assert float((got - expect).abs().max()) <= 4e-7 * scale * world
        # every rank holds the same bits (one rank computes each element)
        mine = got.clone()
        dist.broadcast(mine, src=0)
        assert torch.equal(mine, got)
        for p, v in zip(params, bucket.views):
            assert p.grad is v
    torch.cuda.synchronize()
    open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("multicast", [False, True], ids=["p2p", "multimem"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_peer_allreduce_equals_nccl(world, multicast, tmp_path):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    mp.spawn(_worker, args=(world, _free_port(), multicast, str(tmp_path)), nprocs=world, join=True)
    done = os.listdir(str(tmp_path))
    if any(f.startswith("skip") for f in done):
        pytest.skip("no NVSwitch multicast mapping on this system")
    assert sorted(done) == sorted(f"ok{r}" for r in range(world))


def test_peer_allreduce_rejects_bad_arguments():
    from gstex_b200 import _lib

    lib = _lib.load()
    assert lib.gstex_peer_allreduce_max_world() == 8
    t = torch.zeros(64, device="cuda")
    import ctypes

    one = (ctypes.c_void_p * 2)(t.data_ptr(), t.data_ptr())
    assert lib.gstex_peer_allreduce_f32(0, 1, None, 64, None, 0, None) == 0          # one rank: nothing to do
    assert lib.gstex_peer_allreduce_f32(0, 9, one, 64, None, 0, None) == -1          # more ranks than one NVSwitch domain
    assert lib.gstex_peer_allreduce_f32(2, 2, one, 64, None, 0, None) == -1          # rank outside the world
    assert lib.gstex_peer_allreduce_f32(0, 2, one, 62, None, 0, None) == -1          # count % 4
    assert lib.gstex_peer_allreduce_f32(0, 2, None, 64, None, 0, None) == -1         # no mappings
    bad = (ctypes.c_void_p * 2)(t.data_ptr() + 4, t.data_ptr())
    assert lib.gstex_peer_allreduce_f32(0, 2, bad, 64, None, 0, None) == -1          # unaligned
    # two "ranks" mapped onto one device: rank 0 then rank 1 reduce their halves of the same pair of buffers
    a, b = torch.arange(64, device="cuda", dtype=torch.float32), torch.ones(64, device="cuda")
    both = (ctypes.c_void_p * 2)(a.data_ptr(), b.data_ptr())
    expect = a + b
    assert lib.gstex_peer_allreduce_f32(0, 2, both, 64, None, 0, None) == 0
    assert lib.gstex_peer_allreduce_f32(1, 2, both, 64, None, 0, None) == 0
    torch.cuda.synchronize()
    assert torch.equal(a, expect) and torch.equal(b, expect)

"""HostViewFeeder (gstex_b200/multiview.py): copies overlap the compute but the values are exact."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_feeder_round_trip():
    from gstex_b200.multiview import HostViewFeeder

    dev = torch.device("cuda", 0)
    fd = HostViewFeeder(dev)
    g = torch.Generator().manual_seed(7)
    views = [{"a": torch.randn(257, 33, generator=g).pin_memory(), "n": {"b": torch.randn(5, generator=g).pin_memory()},
              "k": 3} for _ in range(4)]
    outs = [torch.empty(257, 33).pin_memory() for _ in range(4)]
    fd.prefetch(views[0])
    with pytest.raises(RuntimeError):
        fd.prefetch(views[1])
    for i in range(4):
        d = fd.take()
        if i + 1 < 4:
            fd.prefetch(views[i + 1])
        assert d["k"] == 3 and d["a"].is_cuda and d["n"]["b"].is_cuda
        y = d["a"] * 2 + d["n"]["b"][0]
        fd.send_back(y, outs[i])
    with pytest.raises(RuntimeError):
        fd.take()
    fd.finish()
    for i in range(4):
        assert torch.equal(outs[i], views[i]["a"] * 2 + views[i]["n"]["b"][0])


def test_feeder_rejects_cpu():
    from gstex_b200.multiview import HostViewFeeder

    with pytest.raises(RuntimeError):
        HostViewFeeder("cpu")

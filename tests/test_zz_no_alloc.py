"""The no-allocation entry point `gstex_cumsum_i32_noalloc` (include/gstex_b200.h; no counterpart in the reference)
against the existing `gstex_cumsum_i32`, through the C ABI.

Named test_zz_* so that pytest collects it LAST: a failure here must not hide
any other test.  It passes alone in a fresh process and after the rest of the suite (the static scratch array is shared
by every call and keeps the previous call's contents, which is exactly what two calls in a row exercise).

Integer prefix sums: the bar is bit-exact, as for every other binning test.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _existing(lib, _lib, x):
    """gstex_cumsum_i32 with a caller-provided workspace (the existing entry point)."""
    n = x.numel()
    out = torch.empty_like(x)
    total = torch.empty(1, dtype=torch.int32, device=x.device)
    ws = torch.empty(max(int(lib.gstex_cumsum_workspace_bytes(n)), 16), dtype=torch.uint8, device=x.device)
    rc = lib.gstex_cumsum_i32(_lib.ptr(x), n, _lib.ptr(out), _lib.ptr(total), _lib.ptr(ws), ws.numel(), _lib.stream_ptr(x.device))
    _lib.check(rc, "gstex_cumsum_i32")
    return out, total


def _noalloc(lib, _lib, x, out, total):
    """gstex_cumsum_i32_noalloc: inputs and outputs are the caller's, there is no workspace argument."""
    rc = lib.gstex_cumsum_i32_noalloc(_lib.ptr(x), x.numel(), _lib.ptr(out), _lib.ptr(total), _lib.stream_ptr(x.device))
    _lib.check(rc, "gstex_cumsum_i32_noalloc")


def test_noalloc_cumsum_matches_existing_entry_point_twice_in_a_row():
    from gstex_b200 import _lib

    lib = _lib.load()
    dev = torch.device("cuda", torch.cuda.current_device())
    g = torch.Generator().manual_seed(99)
    # two calls in a row with DIFFERENT sizes and data: the second runs over the scratch the first left behind
    # (more tiles first, fewer second, so stale slots beyond the second call's tiles are really there)
    xs = [torch.randint(0, 40, (n,), generator=g, dtype=torch.int32).to(dev) for n in (1_000_003, 257_001, 1, 4096, 4097)]
    outs = [torch.empty_like(x) for x in xs]
    totals = [torch.empty(1, dtype=torch.int32, device=dev) for _ in xs]
    refs = [_existing(lib, _lib, x) for x in xs]
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats(dev).get("num_alloc_retries", 0), torch.cuda.memory_allocated(dev)
    for x, o, t in zip(xs, outs, totals):      # back to back on one stream, nothing allocated in between
        _noalloc(lib, _lib, x, o, t)
    assert (torch.cuda.memory_stats(dev).get("num_alloc_retries", 0), torch.cuda.memory_allocated(dev)) == before
    torch.cuda.synchronize()
    for x, o, t, (ro, rt) in zip(xs, outs, totals, refs):
        assert torch.equal(o, ro), f"n={x.numel()}: {int((o != ro).sum())} prefix sums differ from gstex_cumsum_i32"
        assert torch.equal(t, rt) and int(t) == int(x.sum())
        assert torch.equal(o, torch.cumsum(x, 0, dtype=torch.int32))   # = the reference's torch.cumsum, utils.py:57
    # the same call again over the used scratch gives the same answer
    _noalloc(lib, _lib, xs[0], outs[0], totals[0])
    torch.cuda.synchronize()
    assert torch.equal(outs[0], refs[0][0]) and torch.equal(totals[0], refs[0][1])


def test_noalloc_cumsum_rejects_sizes_beyond_its_capacity_and_bad_arguments():
    from gstex_b200 import _lib

    lib = _lib.load()
    dev = torch.device("cuda", torch.cuda.current_device())
    cap = lib.gstex_cumsum_noalloc_max_n()
    assert cap == 16384 * 4096
    x = torch.ones(8, dtype=torch.int32, device=dev)
    out = torch.full((8,), -7, dtype=torch.int32, device=dev)
    st = _lib.stream_ptr(dev)
    rc = lib.gstex_cumsum_i32_noalloc(_lib.ptr(x), cap + 1, _lib.ptr(out), None, st)   # rejected before anything is read
    assert rc == -4 and b"capacity" in lib.gstex_error_string(rc)
    with pytest.raises(RuntimeError, match="capacity"):
        _lib.check(rc, "gstex_cumsum_i32_noalloc")
    assert lib.gstex_cumsum_i32_noalloc(_lib.ptr(x), -1, _lib.ptr(out), None, st) == -1
    assert lib.gstex_cumsum_i32_noalloc(None, 8, _lib.ptr(out), None, st) == -1
    torch.cuda.synchronize()
    assert int((out != -7).sum()) == 0, "a rejected call must not launch anything"
    total = torch.full((1,), 5, dtype=torch.int32, device=dev)
    assert lib.gstex_cumsum_i32_noalloc(None, 0, None, _lib.ptr(total), st) == 0      # empty input: total = 0
    torch.cuda.synchronize()
    assert int(total) == 0

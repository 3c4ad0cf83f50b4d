"""Host logic of the pair-pool sizing policy (gstex_b200.texture.PairStreamPolicy): no GPU needed -- the asynchronous
header read-back is replaced by a ready-made (header, event) pair."""
import torch


class _Done:
    def query(self):
        return True


def _feed(P, key, used_pairs, overflow=0, capped=False):
    P._state[key]["pending"] = (torch.tensor([used_pairs // 128, overflow, 0, 0], dtype=torch.int32), _Done(), capped)
    return P.max_pairs(torch.device("cuda", key[0]), key[1], key[2], key[3])


def test_pool_size_is_quantised_sticky_and_bounded():
    from gstex_b200.texture import PairStreamPolicy as P

    key = (0, 800, 800, 16)
    P._state.pop(key, None)
    dev = torch.device("cuda", 0)
    q = P.quantum
    first = P.max_pairs(dev, 800, 800, 16)
    assert first == int(P.pairs_per_pixel * 800 * 800)
    # views that need 7..12 M pairs: the pool never follows them down, and moves up only in whole quanta
    sizes = {_feed(P, key, u) for u in (7_000_000, 12_000_000, 9_500_000, 11_000_000, 8_000_000) * 4}
    assert sizes == {first}
    grown = _feed(P, key, 30_000_000)
    assert grown % q == 0 and grown >= 45_000_000 and grown - q < 45_000_000
    assert {_feed(P, key, u) for u in (9_000_000, 30_500_000, 20_000_000)} == {grown}
    # a genuine overflow doubles the pool; an overflow of a pool that `max_bytes` had cut teaches nothing
    assert _feed(P, key, 1_000_000, overflow=1, capped=True) == grown
    doubled = _feed(P, key, grown, overflow=1)
    assert doubled == 2 * grown
    # ... but never beyond what a view can produce: alpha >= 1/255 per pair ends a pixel within 2345 pairs
    for _ in range(40):
        last = _feed(P, key, P._state[key]["pairs"], overflow=1)
    assert last <= 2400 * 800 * 800 and last == P._state[key]["pairs"]
    # only a long run of views that need less than a quarter of the pool shrinks it
    for i in range(31):
        assert _feed(P, key, 9_000_000) == last, i
    shrunk = _feed(P, key, 9_000_000)
    assert shrunk % q == 0 and 13_500_000 <= shrunk < 13_500_000 + q
    P._state.pop(key, None)

"""Both onesweep rankings (thrs_set_ranking) against the oracle, and the device probe that
chooses between them.  The counter-return ranking relies on a hardware order the probe measures
(csrc/probe.cuh); the match-mask ranking does not.  Every case here is duplicate-heavy with
values = input index, so any instability shows up as a value mismatch against
std::stable_sort (oracle.sort, unittest.cpp:358-377)."""
import numpy as np
import pytest

import oracle
import tinyhipradixsort_b200 as thrs
from helpers import check_against_oracle, make_sorter

pytestmark = pytest.mark.gpu

MATCH_MASK = thrs.RadixSort.RANKING_MATCH_MASK
COUNTER_RETURN = thrs.RadixSort.RANKING_COUNTER_RETURN


def test_probe_and_default_choice():
    s = make_sorter("u32", 4)
    in_use, in_order, groups = s.getRanking()
    assert groups > 1_000_000, "the probe must have checked many multi-lane groups"
    # AUTO follows the probe; either outcome is a GPU path
    assert in_use == (COUNTER_RETURN if in_order else MATCH_MASK)
    s.setRanking(MATCH_MASK)
    assert s.getRanking()[0] == MATCH_MASK
    s.setRanking(COUNTER_RETURN)
    assert s.getRanking()[0] == COUNTER_RETURN
    with pytest.raises(Exception):
        s.setRanking(7)


def _dup_heavy(n, kind, seed):
    r = oracle.splitmix64_at(np.arange(n, dtype=np.uint64) + np.uint64(seed * 7919))
    if kind == "equal":
        return np.full(n, 0xDEADBEEF, dtype=np.uint32)
    if kind == "two":
        return np.where((r >> np.uint64(17)) & np.uint64(1), 0x01010101, 0xFEFEFEFE).astype(np.uint32)
    if kind == "unique8":
        t = np.array([0xDEADBEEF, 0, 0xFFFFFFFF, 0x80000000, 0x00010000, 0x12345678, 0x00FF00FF, 0x7FFFFFFF], dtype=np.uint32)
        return t[(r & np.uint64(7)).astype(np.int64)]
    if kind == "nibbles":  # every digit has 16 values: about two lanes per counter in every warp step
        return (r & np.uint64(0x0F0F0F0F)).astype(np.uint32)
    if kind == "runs":  # runs of equal keys: neighbouring lanes collide
        return ((np.arange(n, dtype=np.uint64) // np.uint64(5)) * np.uint64(0x9E3779B1) & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    return (r & np.uint64(0xFFFFFFFF)).astype(np.uint32)


@pytest.mark.parametrize("ranking", [MATCH_MASK, COUNTER_RETURN])
@pytest.mark.parametrize("kind", ["equal", "two", "unique8", "nibbles", "runs", "uniform"])
def test_u32_pairs_stability_under_both_rankings(ranking, kind):
    s = make_sorter("u32", 4)
    s.setRanking(ranking)
    t = s.tileItems()
    for i, n in enumerate((1, 33, t - 1, t, t + 1, 3 * t + 17, 1 << 20)):
        keys = _dup_heavy(n, kind, i)
        check_against_oracle("u32", keys, np.arange(n, dtype=np.uint32), 0, 32, False, s)


@pytest.mark.parametrize("ranking", [MATCH_MASK, COUNTER_RETURN])
def test_other_types_under_both_rankings(ranking):
    rng = oracle.SplitMix64(11)
    n = 300_001
    for key, vb, desc, window in (("u32", None, False, (0, 32)), ("f32", 4, True, (0, 32)), ("u64", 8, False, (0, 64)),
                                  ("f64", 16, False, (16, 48)), ("u32", 4, False, (8, 24)), ("u64", None, True, (3, 35))):
        s = make_sorter(key, vb, desc)
        s.setRanking(ranking)
        keys = rng.randomize(key, n, mode=1)
        # few distinct digits in the sorted window: mask most bits
        if key in ("u32", "f32"):
            keys &= np.uint32(0xF0F0F0FF)
        else:
            keys &= np.uint64(0xF0F0F0F0F0F0F0FF)
        vals = None
        if vb == 4:
            vals = np.arange(n, dtype=np.uint32)
        elif vb == 8:
            vals = np.arange(n, dtype=np.uint64)
        elif vb == 16:
            vals = np.zeros((n, 4), dtype=np.uint32)
            vals[:, 2] = np.arange(n)
        check_against_oracle(key, keys, vals, window[0], window[1], desc, s)


def test_rankings_produce_identical_bytes_at_2pow24():
    """Size-independent property: two independent ranking schemes must give the same permutation."""
    import torch
    from helpers import gpu_sort
    n = 1 << 24
    keys = _dup_heavy(n, "nibbles", 3)
    vals = np.arange(n, dtype=np.uint32)
    outs = []
    for ranking in (MATCH_MASK, COUNTER_RETURN):
        s = make_sorter("u32", 4)
        s.setRanking(ranking)
        outs.append(gpu_sort(s, keys, vals, 0, 32, "u32"))
    assert np.array_equal(outs[0][0], outs[1][0])
    assert np.array_equal(outs[0][1], outs[1][1])
    k = outs[1][0]
    assert np.all(k[:-1] <= k[1:])
    same = k[:-1] == k[1:]
    assert np.all(outs[1][1][:-1][same] < outs[1][1][1:][same]), "equal keys must keep input order"
    torch.cuda.synchronize()

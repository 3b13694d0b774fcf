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


def test_production_check_ran_and_decides():
    """AUTO needs BOTH checks: the isolated probe and the production-kernel comparison (capi.cu
    run_production_rank_check: the real kernels under both rankings on a duplicate-heavy input)."""
    for key, vb in (("u32", 4), ("u64", 8), ("f32", 16)):
        s = make_sorter(key, vb)
        in_use, in_order, _ = s.getRanking()
        ok, items = s.getRankingCheck()
        if in_order:
            assert items >= 2 * (1 << 21) + 2 * (1 << 19), "keys-only and pairs, both tile shapes"
        assert in_use == (COUNTER_RETURN if (in_order and ok) else MATCH_MASK)


def test_forced_mismatch_falls_back_to_match_mask():
    """THRS_FORCE_RANK_MISMATCH=1 makes the production check report a difference: a fresh process
    must then resolve AUTO to the order-free ranking and still sort exactly."""
    import os
    import subprocess
    import sys
    code = (
        "import numpy as np, sys\n"
        "sys.path.insert(0, 'tests')\n"
        "import tinyhipradixsort_b200 as thrs\n"
        "from helpers import make_sorter, check_against_oracle\n"
        "s = make_sorter('u32', 4)\n"
        "in_use, in_order, _ = s.getRanking()\n"
        "ok, items = s.getRankingCheck()\n"
        "assert not ok, 'forced mismatch not seen'\n"
        "assert in_use == thrs.RadixSort.RANKING_MATCH_MASK, in_use\n"
        "n = 300001\n"
        "keys = (np.arange(n, dtype=np.uint64) * 2654435761 % 16).astype(np.uint32) * 0x01010101\n"
        "check_against_oracle('u32', keys, np.arange(n, dtype=np.uint32), 0, 32, False, s)\n"
        "print('FALLBACK-OK')\n"
    )
    env = dict(os.environ, THRS_FORCE_RANK_MISMATCH="1")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "FALLBACK-OK" in out.stdout, out.stdout + out.stderr


def test_verify_mode_repeats_every_sort():
    """THRS_VERIFY=1: every counter-return sort is repeated with match-mask on a copy and compared."""
    import os
    import subprocess
    import sys
    code = (
        "import numpy as np, sys\n"
        "sys.path.insert(0, 'tests')\n"
        "from helpers import make_sorter, check_against_oracle\n"
        "for key, vb in (('u32', 4), ('u32', None), ('u64', 8)):\n"
        "    s = make_sorter(key, vb)\n"
        "    for n in (5000, 40000, 300001, 1 << 21):\n"
        "        dt = np.uint32 if key == 'u32' else np.uint64\n"
        "        keys = (np.arange(n, dtype=np.uint64) * 2654435761 % 251).astype(dt) * dt(0x01000101)\n"
        "        vals = None if vb is None else np.arange(n, dtype=np.uint32 if vb == 4 else np.uint64)\n"
        "        check_against_oracle(key, keys, vals, 0, None, False, s)\n"
        "print('VERIFY-OK')\n"
    )
    env = dict(os.environ, THRS_VERIFY="1")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=900)
    assert out.returncode == 0 and "VERIFY-OK" in out.stdout, out.stdout + out.stderr


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

"""Mid-size inputs (4 096 < n <= 2^22) run the onesweep launches with a quarter-size tile
(thrs_set_mid_max, launch.cuh MidTuning).  The same inputs through both tile shapes must equal the
oracle bit for bit (the reference's own test sizes, 1 ... 99 999 at unittest.cpp:138, all fall in or
below this range)."""
import numpy as np
import pytest

import oracle
import tinyhipradixsort_b200 as thrs
from helpers import KEY_BITS, check_against_oracle, make_sorter
from test_small_gpu import _values

pytestmark = pytest.mark.gpu

# around the quarter-size tile sizes (2 048 / 4 096), a partial last tile, several look-back groups
SIZES = (4097, 8192, 8193, 12289, 99999, 300001, (1 << 20) - 5)


@pytest.mark.parametrize("mid", [True, False])
@pytest.mark.parametrize("key,vb,desc", [("u32", None, False), ("u32", 4, False), ("f32", 4, True), ("u64", 8, False),
                                         ("f64", 16, False), ("u64", None, True), ("u32", 16, True), ("u32", 8, False)])
def test_mid_inputs_both_tile_shapes(key, vb, desc, mid):
    s = make_sorter(key, vb, desc)
    s.setMidMax(thrs.RadixSort.MID_SORT_MAX if mid else 0)
    s.setClusterPath(False)  # the sizes up to 32 768 would otherwise take the one-cluster path
    rng = oracle.SplitMix64(77)
    for n in SIZES:
        keys = rng.randomize(key, n, mode=1)
        keys[::3] = keys[0]  # duplicates: stability is visible in the values
        check_against_oracle(key, keys, _values(n, vb), 0, KEY_BITS[key], desc, s)


@pytest.mark.parametrize("ranking", [thrs.RadixSort.RANKING_MATCH_MASK, thrs.RadixSort.RANKING_COUNTER_RETURN])
def test_mid_shape_skewed_digits_and_windows(ranking):
    s = make_sorter("u32", 4)
    s.setRanking(ranking)
    s.setClusterPath(False)
    rng = oracle.SplitMix64(9)
    n = 200003
    vals = np.arange(n, dtype=np.uint32)
    base = rng.randomize("u32", n, mode=1)
    table8 = np.array([0xDEADBEEF, 0, 0xFFFFFFFF, 0x80000000, 0x00010000, 0x12345678, 0x00FF00FF, 0x7FFFFFFF], dtype=np.uint32)
    for keys in (np.full(n, 0xDEADBEEF, dtype=np.uint32), base & np.uint32(0xFFFF), table8[base & np.uint32(7)],
                 np.sort(base), np.sort(base)[::-1].copy()):
        check_against_oracle("u32", keys, vals, 0, 32, False, s)
    for start, end in ((8, 24), (3, 19), (24, 40), (0, 8)):
        check_against_oracle("u32", base, vals, start, end, False, s)


def test_mid_limit_boundary_and_clamp():
    s = make_sorter("u32", 4)
    s.setMidMax(1 << 30)  # clamped to MID_SORT_MAX: the temp layout covers no more
    rng = oracle.SplitMix64(3)
    for n in (thrs.RadixSort.MID_SORT_MAX, thrs.RadixSort.MID_SORT_MAX + 1):
        keys = rng.randomize("u32", n, mode=1)
        check_against_oracle("u32", keys, np.arange(n, dtype=np.uint32), 0, 32, False, s)

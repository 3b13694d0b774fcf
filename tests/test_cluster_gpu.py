"""The one-cluster path (4 096 < n <= 8 x 4 096, cluster_sort.cuh): one launch of one thread-block
cluster, items exchanged through distributed shared memory.  Same inputs through the cluster path and
through the multi-block path must equal the oracle bit for bit (the reference's own test sizes,
1 ... 99 999 at unittest.cpp:138, mostly fall in this range)."""
import numpy as np
import pytest

import oracle
import tinyhipradixsort_b200 as thrs
from helpers import KEY_BITS, check_against_oracle, make_sorter
from test_small_gpu import _values

pytestmark = pytest.mark.gpu

# every cluster size (2, 4, 8 blocks), block boundaries, a last block with one item, the limit and beyond
SIZES = (4097, 8191, 8192, 8193, 12289, 16384, 16385, 24577, 30000, 32767, 32768, 32769, 50001)


@pytest.mark.parametrize("key,vb,desc", [("u32", None, False), ("u32", 4, False), ("f32", 4, True), ("u64", 8, False),
                                         ("f64", 16, False), ("u64", None, True), ("u32", 16, True), ("u32", 8, False),
                                         ("f64", None, True), ("u64", 4, False)])
def test_cluster_path_equals_oracle(key, vb, desc):
    s = make_sorter(key, vb, desc)
    limit = s.clusterSortMax()
    assert limit == 8 * 4096, limit
    rng = oracle.SplitMix64(123)
    for n in SIZES:
        keys = rng.randomize(key, n, mode=1)
        keys[::3] = keys[0]  # duplicates: stability is visible in the values
        check_against_oracle(key, keys, _values(n, vb), 0, KEY_BITS[key], desc, s)
        assert (s.lastLaunchCount() == 1) == (n <= limit), (n, limit, s.lastLaunchCount())


def test_cluster_path_skew_windows_and_float_specials():
    s = make_sorter("u32", 4)
    rng = oracle.SplitMix64(17)
    n = 30001
    vals = np.arange(n, dtype=np.uint32)
    base = rng.randomize("u32", n, mode=1)
    table8 = np.array([0xDEADBEEF, 0, 0xFFFFFFFF, 0x80000000, 0x00010000, 0x12345678, 0x00FF00FF, 0x7FFFFFFF], dtype=np.uint32)
    for keys in (np.full(n, 0xDEADBEEF, dtype=np.uint32), base & np.uint32(0xFFFF), table8[base & np.uint32(7)],
                 np.sort(base), np.sort(base)[::-1].copy()):
        check_against_oracle("u32", keys, vals, 0, 32, False, s)
        assert s.lastLaunchCount() == 1
    for start, end in ((8, 24), (3, 19), (24, 40), (0, 8), (32, 40), (5, 5)):
        check_against_oracle("u32", base, vals, start, end, False, s)
    f = make_sorter("f32", 4)
    sp = np.array([0x80000000, 0, 0x7FC00000, 0xFFC00000, 0x7F800000, 0xFF800000, 1, 0x80000001, 0x7F7FFFFF, 0xFF7FFFFF],
                  dtype=np.uint32)
    keys = np.tile(sp, 3001)
    check_against_oracle("f32", keys, np.arange(keys.shape[0], dtype=np.uint32), 0, 32, False, f)
    assert f.lastLaunchCount() == 1


def test_cluster_path_switch():
    s = make_sorter("u32", 4)
    rng = oracle.SplitMix64(2)
    n = 20000
    keys = rng.randomize("u32", n, mode=1)
    s.setClusterPath(False)
    assert s.clusterSortMax() == 0
    check_against_oracle("u32", keys, np.arange(n, dtype=np.uint32), 0, 32, False, s)
    assert s.lastLaunchCount() == 6  # zero + histogram + 4 passes
    s.setClusterPath(True)
    check_against_oracle("u32", keys, np.arange(n, dtype=np.uint32), 0, 32, False, s)
    assert s.lastLaunchCount() == 1

"""The one-block path (n <= 4096, small_sort.cuh) and the multi-block path on the SAME small inputs:
both must equal the oracle (the reference's own test sizes are 1 ... 99 999, unittest.cpp:138)."""
import numpy as np
import pytest

import oracle
import tinyhipradixsort_b200 as thrs
from helpers import KEY_BITS, check_against_oracle, make_sorter

pytestmark = pytest.mark.gpu

SIZES = (1, 2, 31, 32, 33, 255, 256, 257, 1000, 2049, 4095, 4096, 4097)


def _values(n, vb):
    if vb is None:
        return None
    if vb == 4:
        return np.arange(n, dtype=np.uint32)
    if vb == 8:
        return np.arange(n, dtype=np.uint64) * np.uint64(3)
    v = np.zeros((n, 4), dtype=np.uint32)
    v[:, 0] = np.arange(n)
    v[:, 3] = 0xABCD0000
    return v


@pytest.mark.parametrize("small", [True, False])
@pytest.mark.parametrize("key,vb,desc", [("u32", None, False), ("u32", 4, False), ("f32", 4, True), ("u64", 8, False),
                                         ("f64", 16, False), ("u64", None, True), ("u32", 16, True), ("f32", None, False)])
def test_small_inputs_both_paths(key, vb, desc, small):
    s = make_sorter(key, vb, desc)
    s.setSmallPath(small)
    rng = oracle.SplitMix64(31)
    for n in SIZES:
        keys = rng.randomize(key, n, mode=1)
        if n >= 32:
            keys[::3] = keys[0]  # duplicates: stability is visible in the values
        check_against_oracle(key, keys, _values(n, vb), 0, KEY_BITS[key], desc, s)
        # one launch: one block up to SMALL_SORT_MAX, one cluster up to clusterSortMax()
        expect_one = small and n <= max(thrs.RadixSort.SMALL_SORT_MAX, s.clusterSortMax())
        assert (s.lastLaunchCount() == 1) == expect_one, (n, s.lastLaunchCount())


@pytest.mark.parametrize("key", ["u32", "u64"])
def test_small_inputs_bit_windows_and_specials(key):
    s = make_sorter(key, 4)
    rng = oracle.SplitMix64(5)
    bits = KEY_BITS[key]
    for n in (77, 3000):
        keys = rng.randomize(key, n, mode=1)
        for start, end in ((0, 8), (8, 24), (3, 19), (bits - 8, bits), (bits - 8, bits + 8), (bits, bits + 8), (5, 5)):
            check_against_oracle(key, keys, np.arange(n, dtype=np.uint32), start, end, False, s)
    f = make_sorter("f32", 4)
    sp = np.array([0x80000000, 0, 0x7FC00000, 0xFFC00000, 0x7F800000, 0xFF800000, 1, 0x80000001, 0x7F7FFFFF, 0xFF7FFFFF],
                  dtype=np.uint32)
    keys = np.tile(sp, 37)
    check_against_oracle("f32", keys, np.arange(keys.shape[0], dtype=np.uint32), 0, 32, False, f)
    assert f.lastLaunchCount() == 1


def test_small_path_latency_is_one_launch_and_needs_no_scratch():
    import torch
    s = make_sorter("u32", 4)
    n = 2048
    keys = oracle.SplitMix64(2).randomize("u32", n, mode=1)
    dk = torch.from_numpy(keys.view(np.int32)).cuda()
    dv = torch.arange(n, dtype=torch.int32, device="cuda")
    tmp = torch.full((s.getTemporaryBufferBytes(n).getTemporaryBufferBytesForSortPairs(),), 0x77, dtype=torch.uint8, device="cuda")
    s.sortPairs(dk, dv, n, tmp, 0, 32)
    torch.cuda.synchronize()
    assert s.lastLaunchCount() == 1
    assert bool((tmp == 0x77).all()), "the one-block path must not touch the temp buffer"
    ek, ev = oracle.sort("u32", keys, np.arange(n, dtype=np.uint32))
    assert np.array_equal(dk.cpu().numpy().view(np.uint32), ek) and np.array_equal(dv.cpu().numpy().view(np.uint32), ev)

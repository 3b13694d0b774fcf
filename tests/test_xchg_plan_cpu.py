"""Host logic of the exchange layer (thrs_xchg_plan, csrc/capi_xchg.cu) WITHOUT a GPU: the C planning
function against the numpy reference (plan_partition), and a CPU emulation of the whole multi-GPU
algorithm built on the plan — top-digit partition per rank, digit-major / source-minor placement at the
owner, per-bucket sort on the low bits — which must equal the oracle's stable sort of the concatenated
slices (std::stable_sort, unittest.cpp:358-377).  The oracle stands in for the kernels here (tests only)."""
import ctypes

import numpy as np
import pytest

import oracle
from tinyhipradixsort_b200 import _lib
from tinyhipradixsort_b200.multigpu import plan_partition


def _plan(world, groups, counts):
    lib = _lib.load()
    c = np.ascontiguousarray(counts, dtype=np.uint64).reshape(world * 256)
    owner = np.zeros(256, dtype=np.int32)
    group = np.zeros(256, dtype=np.int32)
    ngroups = np.zeros(world, dtype=np.int32)
    recv = np.zeros(world, dtype=np.uint64)
    rc = lib.thrs_xchg_plan(world, groups, c.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64)),
                            owner.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), recv.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64)),
                            group.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), ngroups.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)))
    assert rc == 0
    return owner, recv, group, ngroups


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
def test_plan_matches_numpy_reference(world):
    rng = np.random.default_rng(world)
    for case in range(6):
        counts = rng.integers(0, 1 << 20, size=(world, 256)).astype(np.uint64)
        if case == 1:
            counts[:, 5:] = 0          # everything in a few buckets
        if case == 2:
            counts[:] = 0
            counts[0, 200] = 12345     # one bucket only
        if case == 3:
            counts[:] = 0              # empty input
        if case == 4:
            counts[:, ::2] = 0
        owner, recv, group, ngroups = _plan(world, 4, counts)
        dest, send = plan_partition(counts.astype(np.int64), world)
        assert np.array_equal(owner, dest)
        assert np.array_equal(recv.astype(np.int64), send.sum(axis=0))
        assert np.all(np.diff(owner) >= 0), "owners must be monotone in the digit"
        for r in range(world):  # groups restart at every owner, step by at most one, and hold <= 256 / world digits
            g = group[owner == r]
            assert ngroups[r] == (0 if g.size == 0 else g.max() + 1)
            if g.size:
                assert g[0] == 0 and np.all(np.diff(g) >= 0) and np.all(np.diff(g) <= 1) and g.max() < 4 + world
                assert np.bincount(g).max() <= 256 // world, "a group is sorted as world x digits <= 256 chunks"


@pytest.mark.parametrize("world,key", [(2, "u32"), (3, "f32"), (8, "u32"), (4, "u64")])
def test_cpu_emulation_of_the_exchange_equals_stable_sort(world, key):
    bits = 32 if key in ("u32", "f32") else 64
    rng = oracle.SplitMix64(7 * world)
    sizes = [5003, 17, 0, 9000, 2500, 1, 7777, 4096][:world]
    slices = [rng.randomize(key, n, mode=1) for n in sizes]
    if world == 8:
        for s in slices:
            s[: len(s) // 2] &= s.dtype.type(0x0FFFFFFF)  # skew
    base = np.cumsum([0] + sizes)
    vals = [np.arange(n, dtype=np.uint32) + np.uint32(base[i]) for i, n in enumerate(sizes)]
    # per rank: stable pass on the top digit (thrs_partition_pairs) + its histogram
    part, counts = [], np.zeros((world, 256), dtype=np.uint64)
    for r in range(world):
        pk, pv = oracle.sort(key, slices[r], vals[r], bits - 8, bits)
        d = np.array([oracle.field(key, int(x), bits - 8, bits) for x in pk], dtype=np.int64)
        counts[r] = np.bincount(d, minlength=256)
        part.append((pk, pv, d))
    owner, recv, group, ngroups = _plan(world, 4, counts)
    out_k, out_v = [], []
    for r in range(world):
        rk_parts, rv_parts = [], []
        for g in range(int(ngroups[r])):
            digits = [d for d in range(256) if owner[d] == r and group[d] == g]
            # what the copy engines assemble in rank r's window for this group: SOURCE-major, digit-minor
            # (one contiguous copy per source), then thrs_sort_buckets_ex: bucket d = its chunks in source
            # order, sorted stably on the low bits
            for d in digits:
                ck = [part[s][0][part[s][2] == d] for s in range(world)]
                cv = [part[s][1][part[s][2] == d] for s in range(world)]
                bk, bv = np.concatenate(ck), np.concatenate(cv)
                if bk.size:
                    bk, bv = oracle.sort(key, bk, bv, 0, bits - 8)
                rk_parts.append(bk)
                rv_parts.append(bv)
        rk = np.concatenate(rk_parts) if rk_parts else np.zeros(0, dtype=slices[0].dtype)
        rv = np.concatenate(rv_parts) if rv_parts else np.zeros(0, dtype=np.uint32)
        assert rk.shape[0] == int(recv[r])
        out_k.append(rk)
        out_v.append(rv)
    ek, ev = oracle.sort(key, np.concatenate(slices), np.concatenate(vals), 0, bits)
    assert np.array_equal(np.concatenate(out_k), ek)
    assert np.array_equal(np.concatenate(out_v), ev)

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
    recv = np.zeros(world, dtype=np.uint64)
    rc = lib.thrs_xchg_plan(world, groups, c.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64)),
                            owner.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)), recv.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64)),
                            group.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)))
    assert rc == 0
    return owner, recv, group


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
        owner, recv, group = _plan(world, 4, counts)
        dest, send = plan_partition(counts.astype(np.int64), world)
        assert np.array_equal(owner, dest)
        assert np.array_equal(recv.astype(np.int64), send.sum(axis=0))
        assert np.all(np.diff(owner) >= 0), "owners must be monotone in the digit"
        for r in range(world):  # groups restart at every owner and are monotone inside it
            g = group[owner == r]
            assert g.size == 0 or (g[0] == 0 and np.all(np.diff(g) >= 0) and g.max() < 4)


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
    owner, recv, group = _plan(world, 4, counts)
    out_k, out_v = [], []
    for r in range(world):
        # digit-major / source-minor: what the copy engines assemble in rank r's window
        ck, cv, seg = [], [], [0]
        for d in range(256):
            if owner[d] != r:
                continue
            for s in range(world):
                m = part[s][2] == d
                ck.append(part[s][0][m])
                cv.append(part[s][1][m])
            seg.append(seg[-1] + int(counts[:, d].sum()))
        rk = np.concatenate(ck) if ck else np.zeros(0, dtype=slices[0].dtype)
        rv = np.concatenate(cv) if cv else np.zeros(0, dtype=np.uint32)
        assert rk.shape[0] == int(recv[r])
        for a, b in zip(seg[:-1], seg[1:]):  # segmented sort on the low bits (thrs_sort_segments_ex)
            if b > a:
                rk[a:b], rv[a:b] = oracle.sort(key, rk[a:b], rv[a:b], 0, bits - 8)
        out_k.append(rk)
        out_v.append(rv)
    ek, ev = oracle.sort(key, np.concatenate(slices), np.concatenate(vals), 0, bits)
    assert np.array_equal(np.concatenate(out_k), ek)
    assert np.array_equal(np.concatenate(out_v), ev)

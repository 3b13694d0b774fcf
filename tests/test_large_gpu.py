"""GPU: BASELINE.json's full sizes, checked through size-independent properties computed on the
device (sortedness on the getKeyBits field, stability, permutation), plus an exact comparison
against the reference's own GPU kernels at 2^28."""
import numpy as np
import pytest
import torch

import oracle
from helpers import make_sorter

pytestmark = pytest.mark.gpu

M64 = (1 << 64) - 1


def _mix_torch(i):
    """splitmix64_at(i) on the device with int64 wrap-around arithmetic."""
    def c(x):  # python int -> signed 64-bit constant
        return x - (1 << 64) if x >= (1 << 63) else x

    def shr(z, s):  # logical shift right on int64
        return (z >> s) & ((1 << (64 - s)) - 1)

    z = (i + 1) * c(0x9E3779B97F4A7C15)
    z = (z ^ shr(z, 30)) * c(0xBF58476D1CE4E5B9)
    z = (z ^ shr(z, 27)) * c(0x94D049BB133111EB)
    return z ^ shr(z, 31)


def gen_u32_pairs(n, device="cuda", chunk=1 << 26):
    keys = torch.empty(n, dtype=torch.int32, device=device)
    for a in range(0, n, chunk):
        b = min(n, a + chunk)
        z = _mix_torch(torch.arange(a, b, dtype=torch.int64, device=device))
        keys[a:b] = (z & 0xFFFFFFFF).to(torch.int32)  # wraps to the same 32 bits
    vals = torch.arange(n, dtype=torch.int64, device=device).to(torch.int32)
    return keys, vals


def f32_field(k32):
    """getKeyBits(float) of fpKey.hpp:23-30 on int32 bit patterns, as int64 in [0, 2^32)."""
    u = k32.to(torch.int64) & 0xFFFFFFFF
    u = torch.where((u & 0x7FFFFFFF) == 0, torch.zeros_like(u), u)
    flip = torch.where(u >= 0x80000000, torch.full_like(u, 0xFFFFFFFF), torch.full_like(u, 0x80000000))
    return u ^ flip


def assert_stable_sorted(keys_out, vals_out, field_fn, gen_keys_fn, chunk=1 << 26):
    """(1) key_out[j] == original key at index value_out[j]  (2) (field, value) strictly increasing
    => a permutation, sorted, ties in input order == std::stable_sort."""
    n = keys_out.shape[0]
    prev_f = prev_v = None
    for a in range(0, n, chunk):
        b = min(n, a + chunk)
        k = keys_out[a:b]
        v = vals_out[a:b].to(torch.int64) & 0xFFFFFFFF
        assert torch.equal(gen_keys_fn(v), k), "key/value association broken"
        f = field_fn(k)
        ok = (f[1:] > f[:-1]) | ((f[1:] == f[:-1]) & (v[1:] > v[:-1]))
        assert bool(ok.all()), "not sorted / not stable"
        if prev_f is not None:
            assert (f[0] > prev_f) or (f[0] == prev_f and v[0] > prev_v)
        prev_f, prev_v = int(f[-1]), int(v[-1])


def test_c2_2pow28_u32_pairs():
    n = 1 << 28
    keys, vals = gen_u32_pairs(n)
    sorter = make_sorter("u32", 4)
    tmp = torch.empty(sorter.getTemporaryBufferBytes(n).getTemporaryBufferBytesForSortPairs(), dtype=torch.uint8,
                      device="cuda")
    sorter.sortPairs(keys, vals, n, tmp, 0, 32)
    torch.cuda.synchronize()
    assert_stable_sorted(keys, vals, lambda k: k.to(torch.int64) & 0xFFFFFFFF,
                         lambda v: (_mix_torch(v) & 0xFFFFFFFF).to(torch.int32))
    # exact comparison with the reference's own kernels on the same input
    if oracle.RefGpuSorter.available():
        del tmp
        rk, rv = gen_u32_pairs(n)
        ref = oracle.RefGpuSorter("u32", 4, False)
        rtmp = torch.empty(ref.temp_bytes(n, True), dtype=torch.uint8, device="cuda")
        ref.sort(rk.data_ptr(), rv.data_ptr(), n, rtmp.data_ptr(), 0, 32, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        assert torch.equal(rk, keys) and torch.equal(rv, vals), "differs from the reference GPU path"


def test_c3_2pow28_f32_pairs_with_specials():
    n = 1 << 28
    keys, vals = gen_u32_pairs(n)  # raw bits as float patterns: NaN, Inf, denormals occur naturally
    special = torch.tensor([0x00000000, -0x80000000, 0x7F800000, -0x00800000, 0x7FC00000, -0x00400000, 0x7F800001,
                            -0x007FFFFF, 0x00000001, -0x7FFFFFFF, 0x007FFFFF, -0x7F800001, 0x7F7FFFFF, -0x00800001],
                           dtype=torch.int32, device="cuda")
    idx = torch.arange(0, n, 4096, device="cuda")
    orig_special = special[torch.arange(idx.shape[0], device="cuda") % special.shape[0]]
    keys[idx] = orig_special

    def regen(v):
        k = (_mix_torch(v) & 0xFFFFFFFF).to(torch.int32)
        m = (v % 4096) == 0
        return torch.where(m, special[(v // 4096) % special.shape[0]], k)

    sorter = make_sorter("f32", 4)
    tmp = torch.empty(sorter.getTemporaryBufferBytes(n).getTemporaryBufferBytesForSortPairs(), dtype=torch.uint8,
                      device="cuda")
    sorter.sortPairs(keys, vals, n, tmp, 0, 32)
    torch.cuda.synchronize()
    assert_stable_sorted(keys, vals, f32_field, regen)


@pytest.mark.parametrize("dist", ["all_equal", "range16", "unique8", "window_8_24"])
@pytest.mark.parametrize("pairs", [False, True])
def test_c4_2pow27_adversarial(dist, pairs):
    n = 1 << 27
    base, vals = gen_u32_pairs(n)
    table8 = torch.tensor([-0x21524111, 0, -1, -0x80000000, 0x00010000, 0x12345678, 0x00FF00FF, 0x7FFFFFFF],
                          dtype=torch.int32, device="cuda")
    start, end = 0, 32
    if dist == "all_equal":
        xform = lambda k: torch.full_like(k, -0x21524111)  # 0xDEADBEEF
    elif dist == "range16":
        xform = lambda k: k & 0xFFFF
    elif dist == "unique8":
        xform = lambda k: table8[(k & 7).to(torch.int64)]
    else:
        xform = lambda k: k
        start, end = 8, 24
    keys = xform(base)
    del base
    orig = keys.clone() if not pairs else None
    sorter = make_sorter("u32", 4 if pairs else None)
    d = sorter.getTemporaryBufferBytes(n)
    tmp = torch.empty(d.getTemporaryBufferBytesForSortPairs() if pairs else d.getTemporaryBufferBytesForSortKeys(),
                      dtype=torch.uint8, device="cuda")
    field = lambda k: ((k.to(torch.int64) & 0xFFFFFFFF) >> start) & ((1 << (end - start)) - 1)
    if pairs:
        sorter.sortPairs(keys, vals, n, tmp, start, end)
        torch.cuda.synchronize()
        assert_stable_sorted(keys, vals, field, lambda v: xform((_mix_torch(v) & 0xFFFFFFFF).to(torch.int32)))
    else:
        sorter.sortKeys(keys, n, tmp, start, end)
        torch.cuda.synchronize()
        f = field(keys)
        assert bool((f[1:] >= f[:-1]).all())
        if dist == "window_8_24":
            # keys-only stability is observable here: compare with a stable torch sort on the field
            of = field(orig)
            _, perm = torch.sort(of, stable=True)
            assert torch.equal(orig[perm], keys)
        else:
            assert torch.equal(torch.sort(orig)[0] if dist == "all_equal" else
                               torch.sort(orig.to(torch.int64) & 0xFFFFFFFF)[0].to(torch.int32), keys)


def test_u32_large_2pow31_plus_100_keys():
    """UTEST(SortKeys, u32Large), unittest.cpp:688-717: 32-bit index overflow guard."""
    n = (1 << 31) + 100
    sorter = make_sorter("u32")
    keys = torch.empty(n, dtype=torch.int32, device="cuda")
    chunk = 1 << 27
    for a in range(0, n, chunk):
        b = min(n, a + chunk)
        keys[a:b] = (_mix_torch(torch.arange(a, b, dtype=torch.int64, device="cuda")) & 0xFFFFFFFF).to(torch.int32)
    # input multiset fingerprint: per-digit histograms of every byte + wrap-around sum
    def fingerprint(t):
        s = 0
        h = torch.zeros(4, 256, dtype=torch.int64, device="cuda")
        for a in range(0, n, chunk):
            u = t[a:min(n, a + chunk)].to(torch.int64) & 0xFFFFFFFF
            s = (s + int(u.sum())) & M64
            for b in range(4):
                h[b] += torch.bincount((u >> (8 * b)) & 0xFF, minlength=256)
        return s, h
    s0, h0 = fingerprint(keys)
    tmp = torch.empty(sorter.getTemporaryBufferBytes(n).getTemporaryBufferBytesForSortKeys(), dtype=torch.uint8,
                      device="cuda")
    sorter.sortKeys(keys, n, tmp, 0, 32)
    torch.cuda.synchronize()
    del tmp
    prev = None
    for a in range(0, n, chunk):
        u = keys[a:min(n, a + chunk)].to(torch.int64) & 0xFFFFFFFF
        assert bool((u[1:] >= u[:-1]).all())
        if prev is not None:
            assert int(u[0]) >= prev
        prev = int(u[-1])
    s1, h1 = fingerprint(keys)
    assert s0 == s1 and torch.equal(h0, h1)

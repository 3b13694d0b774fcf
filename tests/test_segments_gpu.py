"""Segmented sort (thrs_sort_segments_*): every segment must come out exactly as the oracle sorts it
alone (std::stable_sort on the ordered field, unittest.cpp:358-377), items outside the segments must
stay untouched.  Covers empty segments, one-item segments, segments of exactly / just under / just over
one tile, more than 256 segments (several batches), both rankings, odd and even pass counts, all key
types and value widths, and a bucket layout like the one the multi-GPU layer produces."""
import numpy as np
import pytest
import torch

import oracle
import tinyhipradixsort_b200 as thrs
from helpers import make_sorter, to_device, to_host

pytestmark = pytest.mark.gpu


def _values(n, vbytes):
    if vbytes == 4:
        return np.arange(n, dtype=np.uint32)
    if vbytes == 8:
        return np.arange(n, dtype=np.uint64) * np.uint64(0x100000001)
    v = np.zeros((n, 4), dtype=np.uint32)
    v[:, 0] = np.arange(n, dtype=np.uint32)
    v[:, 3] = ~np.arange(n, dtype=np.uint32)
    return v


def _check(sorter, key, keys, vals, begin, start, end, descending=False):
    n = keys.shape[0]
    dk = to_device(keys)
    dv = to_device(vals) if vals is not None else None
    nbytes = sorter.getSegmentsTemporaryBufferBytes(begin[-1], len(begin) - 1, vals is not None)
    tmp = torch.empty(max(nbytes, 16), dtype=torch.uint8, device="cuda")
    tmp.fill_(0xA5)
    sorter.sortSegments(dk, dv, begin, tmp, start, end)
    torch.cuda.synchronize()
    got_k = to_host(dk, keys)
    got_v = to_host(dv, vals) if vals is not None else None
    exp_k, exp_v = keys.copy(), (vals.copy() if vals is not None else None)
    for a, b in zip(begin[:-1], begin[1:]):
        if b > a:
            ek, ev = oracle.sort(key, keys[a:b], None if vals is None else vals[a:b], start, end, descending)
            exp_k[a:b] = ek
            if vals is not None:
                exp_v[a:b] = ev
    assert np.array_equal(got_k, exp_k), "keys differ from the per-segment oracle"
    if vals is not None:
        assert np.array_equal(got_v, exp_v), "values differ from the per-segment oracle (stability)"
    assert n >= begin[-1]


@pytest.mark.parametrize("ranking", [thrs.RadixSort.RANKING_MATCH_MASK, thrs.RadixSort.RANKING_COUNTER_RETURN])
def test_segment_shapes_u32_pairs(ranking):
    s = make_sorter("u32", 4)
    s.setRanking(ranking)
    t = s.tileItems()
    rng = oracle.SplitMix64(3)
    sizes = [0, 1, 0, 33, t - 1, t, t + 1, 0, 16 * t, 16 * t + 1, 17 * t - 1, 5, 40 * t + 123, 0]
    begin = [7]  # the arrays start before the first segment: those items must not move
    for c in sizes:
        begin.append(begin[-1] + c)
    n = begin[-1] + 11
    keys = rng.randomize("u32", n, mode=1)
    keys[::3] &= np.uint32(0x0F0F0F0F)  # duplicates: stability is observable
    _check(s, "u32", keys, _values(n, 4), begin, 0, 32)
    _check(s, "u32", keys, _values(n, 4), begin, 0, 24)  # odd pass count (what the multi-GPU layer runs)
    _check(s, "u32", keys, None, begin, 8, 32)


def test_many_segments_in_several_batches():
    s = make_sorter("u32", 4)
    rng = oracle.SplitMix64(5)
    counts = [int(rng.next() % 3000) for _ in range(700)]
    counts[100] = 0
    counts[256] = 50000
    begin = [0]
    for c in counts:
        begin.append(begin[-1] + c)
    n = begin[-1]
    keys = rng.randomize("u32", n, mode=1) & np.uint32(0x00FFFFFF)
    _check(s, "u32", keys, _values(n, 4), begin, 0, 24)
    _check(s, "u32", keys, _values(n, 4), begin, 0, 32)


@pytest.mark.parametrize("key,vbytes,desc,window", [("f32", 4, True, (0, 32)), ("u64", 8, False, (0, 64)), ("f64", 16, False, (16, 40)),
                                                     ("u32", 16, False, (0, 32)), ("u64", None, True, (8, 64)), ("u32", 8, False, (0, 24))])
def test_type_matrix(key, vbytes, desc, window):
    s = make_sorter(key, vbytes, desc)
    rng = oracle.SplitMix64(9)
    t = s.tileItems()
    begin = [0, 5 * t + 3, 5 * t + 3, 9 * t, 9 * t + 1, 30 * t + 77]
    n = begin[-1]
    keys = rng.randomize(key, n, mode=1)
    vals = None if vbytes is None else _values(n, vbytes)
    _check(s, key, keys, vals, begin, window[0], window[1], desc)


def test_top_digit_buckets_like_the_exchange():
    """What the multi-GPU layer does on a rank: the array is partitioned by the top digit (a stable pass
    on bits [24,32)), then ONE segmented 3-pass sort on bits [0,24) finishes the job: the result must be
    the full 32-bit stable sort."""
    s = make_sorter("u32", 4)
    n = (1 << 22) + 12345
    rng = oracle.SplitMix64(17)
    keys = rng.randomize("u32", n, mode=1)
    keys[: n // 2] &= np.uint32(0x3FFFFFFF)  # uneven buckets, some empty
    vals = _values(n, 4)
    pk, pv = oracle.sort("u32", keys, vals, 24, 32)
    hist = np.bincount((pk >> np.uint32(24)).astype(np.int64), minlength=256)
    begin = [0] + [int(x) for x in np.cumsum(hist)]
    dk, dv = to_device(pk), to_device(pv)
    tmp = torch.empty(s.getSegmentsTemporaryBufferBytes(n, 256, True), dtype=torch.uint8, device="cuda")
    tmp.fill_(0x5A)
    s.sortSegments(dk, dv, begin, tmp, 0, 24)
    torch.cuda.synchronize()
    ek, ev = oracle.sort("u32", keys, vals, 0, 32)
    assert np.array_equal(to_host(dk, pk), ek)
    assert np.array_equal(to_host(dv, pv), ev)


def test_large_segments_property_2pow27():
    """Size-independent check at scale: 2^27 pairs in 37 uneven segments; every segment sorted on its
    window, equal fields keep ascending values, every (key, value) pair still belongs together."""
    s = make_sorter("u32", 4)
    n = 1 << 27
    g = torch.Generator(device="cuda")
    g.manual_seed(11)
    k0 = torch.randint(-(1 << 31), (1 << 31) - 1, (n,), dtype=torch.int64, device="cuda", generator=g).to(torch.int32)
    cuts = sorted(set([0, n] + [int(x) for x in torch.randint(0, n, (36,), generator=torch.Generator().manual_seed(3))]))
    k, v = k0.clone(), torch.arange(n, dtype=torch.int32, device="cuda")
    tmp = torch.empty(s.getSegmentsTemporaryBufferBytes(n, len(cuts) - 1, True), dtype=torch.uint8, device="cuda")
    s.sortSegments(k, v, cuts, tmp, 0, 24)
    torch.cuda.synchronize()
    f = (k.to(torch.int64) & 0xFFFFFF)
    ok = (f[1:] > f[:-1]) | ((f[1:] == f[:-1]) & (v[1:] > v[:-1]))
    boundary = torch.zeros(n - 1, dtype=torch.bool, device="cuda")
    for c in cuts[1:-1]:
        boundary[c - 1] = True
    assert bool((ok | boundary).all())
    assert bool((k0[v.to(torch.int64)] == k).all())
    seg_of = torch.bucketize(torch.arange(n, device="cuda"), torch.tensor(cuts[1:-1], device="cuda"), right=True)
    assert bool((seg_of[v.to(torch.int64)] == seg_of).all()), "an item left its segment"


@pytest.mark.parametrize("nsrc,nb,ranking", [(2, 128, thrs.RadixSort.RANKING_COUNTER_RETURN), (8, 32, thrs.RadixSort.RANKING_COUNTER_RETURN),
                                            (3, 7, thrs.RadixSort.RANKING_MATCH_MASK), (1, 5, thrs.RadixSort.RANKING_COUNTER_RETURN)])
def test_buckets_gather_source_major_chunks(nsrc, nb, ranking):
    """thrs_sort_buckets_ex: the layout a rank of the multi-GPU sort receives — nsrc sources, each a run of nb
    chunks (one per bucket) — must come out as nb buckets, bucket b = stable sort on bits [0,24) of its
    chunks concatenated in source order.  Chunk sizes are uneven, some empty, some larger than a tile."""
    s = make_sorter("u32", 4)
    s.setRanking(ranking)
    t = s.tileItems()
    rng = oracle.SplitMix64(100 * nsrc + nb)
    sizes = [[int(rng.next() % (3 * t // 2)) if (rng.next() % 5) else 0 for _ in range(nb)] for _ in range(nsrc)]
    sizes[0][0] = 2 * t + 5
    chunk_begin, chunk_bucket = [13], []
    for src in range(nsrc):
        for b in range(nb):
            chunk_begin.append(chunk_begin[-1] + sizes[src][b])
            chunk_bucket.append(b)
    bucket_begin = [13]
    for b in range(nb):
        bucket_begin.append(bucket_begin[-1] + sum(sizes[src][b] for src in range(nsrc)))
    n = chunk_begin[-1] + 9
    keys = rng.randomize("u32", n, mode=1)
    keys[::2] &= np.uint32(0x00030303)  # heavy duplicates on the sorted bits: cross-chunk stability is observable
    vals = _values(n, 4)
    dk, dv = to_device(keys), to_device(vals)
    ak, av = torch.empty_like(dk), torch.empty_like(dv)
    ak.fill_(-1); av.fill_(-1)
    sb = s.getSegmentsScratchBytes(n, 256)
    scratch = torch.empty(sb, dtype=torch.uint8, device="cuda")
    scratch.fill_(0xA5)
    in_alt = s.sortBucketsOutOfPlace(dk, dv, ak, av, scratch, sb, chunk_begin, chunk_bucket, bucket_begin, 0, 24)
    torch.cuda.synchronize()
    assert in_alt, "three passes end in the alternates"
    got_k, got_v = to_host(ak, keys), to_host(av, vals)
    for b in range(nb):
        idx = np.concatenate([np.arange(chunk_begin[src * nb + b], chunk_begin[src * nb + b + 1]) for src in range(nsrc)]).astype(np.int64)
        if idx.size == 0:
            continue
        ek, ev = oracle.sort("u32", keys[idx], vals[idx], 0, 24)
        assert np.array_equal(got_k[bucket_begin[b]:bucket_begin[b + 1]], ek), f"bucket {b}: keys"
        assert np.array_equal(got_v[bucket_begin[b]:bucket_begin[b + 1]], ev), f"bucket {b}: values (stability across chunks)"

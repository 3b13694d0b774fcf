"""thrs_scatter_pairs (the fused partition + exchange pass of the multi-GPU layer) on ONE device:
three emulated ranks scatter their slices into three receive buffers on the same GPU through
per-digit destination pointers; the result must equal the source-ordered all-to-all layout, and
a local sort of each buffer must give the oracle's stable sort of everything."""
import numpy as np
import pytest
import torch

import oracle
from helpers import make_sorter, to_device
from tinyhipradixsort_b200.multigpu import plan_partition, scatter_tables

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("key,vbytes", [("u32", 4), ("f32", 4), ("u64", 8)])
@pytest.mark.parametrize("ranking", [1, 2])
def test_scatter_pairs_emulated_ranks(key, vbytes, ranking):
    world = 3
    bits = 32 if key in ("u32", "f32") else 64
    rng = oracle.SplitMix64(21)
    sorter = make_sorter(key, vbytes)
    sorter.setRanking(ranking)
    t = sorter.tileItems()
    sizes = [3 * t + 17, t - 1, 200_003]
    slices = [rng.randomize(key, n, mode=1) for n in sizes]
    slices[1][::3] = slices[1][0]  # duplicates
    vdt = np.uint32 if vbytes == 4 else np.uint64
    base = np.cumsum([0] + sizes)
    vals = [np.arange(n, dtype=vdt) + vdt(base[i]) for i, n in enumerate(sizes)]
    shift = bits - 8
    digit = [np.array([oracle.field(key, int(x), shift, shift + 8) for x in s], dtype=np.int64) for s in slices]
    h = np.stack([np.bincount(d, minlength=256) for d in digit]).astype(np.int64)
    dest, send = plan_partition(h, world)
    recv_n = [int(send[:, r].sum()) for r in range(world)]
    kdt = torch.int32 if bits == 32 else torch.int64
    rk = [torch.full((max(n, 1),), -1, dtype=kdt, device="cuda") for n in recv_n]
    rv = [torch.full((max(n, 1),), -1, dtype=torch.int32 if vbytes == 4 else torch.int64, device="cuda") for n in recv_n]
    for me in range(world):
        bin_first, first = scatter_tables(dest, send, me)
        kd = np.array([t_.data_ptr() for t_ in rk], dtype=np.uint64)
        vd = np.array([t_.data_ptr() for t_ in rv], dtype=np.uint64)
        dk, dv = to_device(slices[me]), to_device(vals[me])
        n = sizes[me]
        tmp = torch.empty(sorter.getTemporaryBufferBytes(n).getTemporaryBufferBytesForSortPairs(), dtype=torch.uint8, device="cuda")
        tmp.fill_(0x5A)
        sorter.scatterPairs(dk, dv, n, tmp, shift, bin_first, kd, vd, first)
        assert sorter.lastLaunchCount() == 1
    torch.cuda.synchronize()
    out_k, out_v = [], []
    for r in range(world):
        n = recv_n[r]
        got_k = rk[r][:n].cpu().numpy().view(slices[0].dtype)
        got_v = rv[r][:n].cpu().numpy().view(vdt)
        # expected layout: chunks in source order, input order inside a chunk
        ek, ev = [], []
        for me in range(world):
            sel = dest[digit[me]] == r
            ek.append(slices[me][sel])
            ev.append(vals[me][sel])
        assert np.array_equal(got_k, np.concatenate(ek)), f"receive buffer {r}: keys"
        assert np.array_equal(got_v, np.concatenate(ev)), f"receive buffer {r}: values"
        if n:
            tmp = torch.empty(sorter.getTemporaryBufferBytes(n).getTemporaryBufferBytesForSortPairs(), dtype=torch.uint8, device="cuda")
            sorter.sortPairs(rk[r], rv[r], n, tmp, 0, bits)
            torch.cuda.synchronize()
        out_k.append(rk[r][:n].cpu().numpy().view(slices[0].dtype))
        out_v.append(rv[r][:n].cpu().numpy().view(vdt))
    exp_k, exp_v = oracle.sort(key, np.concatenate(slices), np.concatenate(vals))
    assert np.array_equal(np.concatenate(out_k), exp_k)
    assert np.array_equal(np.concatenate(out_v), exp_v)


def test_scatter_pairs_rejects_unsupported():
    s = make_sorter("u32", 16)
    z = np.zeros(2, dtype=np.uint64)
    d = torch.zeros(64, dtype=torch.int32, device="cuda")
    tmp = torch.empty(s.getTemporaryBufferBytes(16).getTemporaryBufferBytesForSortPairs(), dtype=torch.uint8, device="cuda")
    with pytest.raises(Exception):
        s.scatterPairs(d, d, 4, tmp, 24, np.array([0, 128], dtype=np.uint32), z, z, np.zeros(2, dtype=np.uint32))
    s4 = make_sorter("u32", 4)
    tmp = torch.empty(s4.getTemporaryBufferBytes(16).getTemporaryBufferBytesForSortPairs(), dtype=torch.uint8, device="cuda")
    ptrs = np.array([d.data_ptr(), d.data_ptr()], dtype=np.uint64)
    with pytest.raises(Exception):  # bins must start at digit 0 and ascend
        s4.scatterPairs(d, d, 4, tmp, 24, np.array([5, 3], dtype=np.uint32), ptrs, ptrs, np.zeros(2, dtype=np.uint32))

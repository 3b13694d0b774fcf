"""Host logic of the partitioned multi-GPU sort on CPU: world_size-2 and -3 gloo groups, the
oracle standing in for the device kernels (tests only — the product default is CudaLocalOps).
Checks splitter planning, count matrices, the source-ordered all-to-all and global stability."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from tinyhipradixsort_b200.multigpu import KEY_BITS, DistributedRadixSort, plan_partition

_NP = {"u32": np.uint32, "f32": np.uint32, "u64": np.uint64, "f64": np.uint64}


class OracleLocalOps:
    """LocalOps over CPU tensors: oracle.sort / oracle.key_bits in place of the CUDA kernels."""

    def __init__(self, key):
        self.key = key

    def _np(self, t):
        return t.numpy().view(_NP[self.key])

    def top_digit_histogram(self, keys, n):
        k = self._np(keys)[:n]
        shift = KEY_BITS[self.key] - 8
        digits = np.array([oracle.field(self.key, int(x), shift, shift + 8) for x in k], dtype=np.int64)
        return torch.from_numpy(np.bincount(digits, minlength=256).astype(np.int64))

    def sort_pairs(self, keys, vals, n, start, end):
        if n == 0:
            return
        k, v = oracle.sort(self.key, self._np(keys)[:n], vals.numpy()[:n], start, end)
        self._np(keys)[:n] = k
        vals.numpy()[:n] = v


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _make_slices(key, world, case):
    rng = oracle.SplitMix64(7 + len(case))
    out = []
    for r in range(world):
        n = [3001, 1777, 2500][r % 3]
        if case == "empty_rank" and r == 1:
            n = 0
        k = rng.randomize(key, n, mode=1)
        if case == "all_equal":
            k[:] = 0xDEADBEEF
        elif case == "skew":
            k[: n // 2] &= 0xFF
        elif case == "specials" and n:
            sp = [0x80000000, 0, 0x7FC00000, 0xFFC00000, 0x7F800000, 0xFF800000, 1, 0x80000001]
            k[::5] = np.array(sp, dtype=k.dtype)[np.arange(len(k[::5])) % len(sp)]
        out.append(k)
    return out


def _worker(rank, world, port, key, case, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    slices = _make_slices(key, world, case)
    base = sum(len(s) for s in slices[:rank])
    k = slices[rank].copy()
    v = (np.arange(len(k), dtype=np.uint32) + np.uint32(base))
    tk = torch.from_numpy(k.view(np.int32 if k.dtype == np.uint32 else np.int64))
    tv = torch.from_numpy(v.view(np.int32))
    ds = DistributedRadixSort(dist, OracleLocalOps(key), key)
    rk, rv, cnt = ds.sort_pairs(tk, tv)
    q.put((rank, rk[:cnt].numpy().view(_NP[key]).copy(), rv[:cnt].numpy().view(np.uint32).copy(),
           ds.last_plan[0].copy(), ds.last_plan[1].copy()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,key,case", [(2, "u32", "uniform"), (2, "f32", "specials"), (3, "u32", "skew"),
                                            (2, "u32", "all_equal"), (2, "u32", "empty_rank"), (2, "u64", "uniform")])
def test_distributed_sort_gloo(world, key, case):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, key, case, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted([q.get(timeout=120) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    slices = _make_slices(key, world, case)
    all_k = np.concatenate(slices)
    all_v = np.arange(len(all_k), dtype=np.uint32)
    exp_k, exp_v = oracle.sort(key, all_k, all_v)
    out_k = np.concatenate([g[1] for g in got])
    out_v = np.concatenate([g[2] for g in got])
    assert np.array_equal(out_k, exp_k), "concatenated keys differ from the stable sort of the whole input"
    assert np.array_equal(out_v, exp_v), "global stability broken"
    dest, send = got[0][3], got[0][4]
    for g in got[1:]:
        assert np.array_equal(g[3], dest) and np.array_equal(g[4], send), "ranks disagree on the plan"
    assert np.all(np.diff(dest) >= 0) and dest.max() < world
    assert send.sum() == len(all_k) and [len(g[1]) for g in got] == list(send.sum(axis=0))


def test_plan_partition_balances_uniform_and_keeps_buckets_whole():
    h = np.full((8, 256), 1 << 20, dtype=np.int64)
    dest, send = plan_partition(h, 8)
    assert list(np.bincount(dest, minlength=8)) == [32] * 8
    assert np.all(send == (1 << 20) * 32)
    h = np.zeros((4, 256), dtype=np.int64)
    h[:, 17] = 5  # one bucket only: all of it goes to one rank
    dest, send = plan_partition(h, 4)
    assert send[:, dest[17]].sum() == 20 and send.sum() == 20
    dest, send = plan_partition(np.zeros((2, 256), dtype=np.int64), 2)
    assert send.sum() == 0
    big = np.full((8, 256), (1 << 29) // 256, dtype=np.int64)  # 2^32 items in total: 64-bit counts
    dest, send = plan_partition(big, 8)
    assert send.sum() == 1 << 32


def test_scatter_tables_give_source_ordered_disjoint_chunks():
    """The fused exchange places each (source, destination) chunk directly: chunks of one
    destination must tile its receive buffer in source-rank order, and the bins must be the
    digit ranges of plan_partition."""
    from tinyhipradixsort_b200.multigpu import scatter_tables

    rng = np.random.default_rng(3)
    for world in (1, 2, 3, 8):
        h = rng.integers(0, 50, size=(world, 256)).astype(np.int64)
        h[:, rng.integers(0, 256, size=40)] = 0
        if world == 3:
            h[:, 40:] = 0  # late ranks may end up empty
        dest, send = plan_partition(h, world)
        cursor = np.zeros(world, dtype=np.int64)
        for me in range(world):
            bin_first, first = scatter_tables(dest, send, me)
            assert bin_first[0] == 0 and np.all(np.diff(bin_first) >= 0) and bin_first.max() <= 256
            for d in range(256):  # bin of digit d = number of thresholds <= d (what the kernel computes)
                assert int(np.sum(bin_first[1:] <= d)) == dest[d]
            for r in range(world):
                assert first[r] == cursor[r], "chunks must follow each other in source-rank order"
                cursor[r] += send[me, r]
        assert np.array_equal(cursor, send.sum(axis=0))

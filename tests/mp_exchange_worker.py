"""Worker of tests/test_exchange_gpu.py: one process per GPU (torchrun), DistributedRadixSort over the
exchange layer (thrs_xchg_*).  Every rank sorts its slice of a seeded input; rank 0 gathers the shares
and compares their concatenation with the oracle's stable sort of the whole input."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from tinyhipradixsort_b200.multigpu import CudaExchangeOps, DistributedRadixSort  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=dev)
    ok = True
    for case, (key, groups) in enumerate((("u32", 4), ("u32", 1), ("f32", 3), ("u32", 4), ("u32", 16))):
        n = [300_001, 123_457, 1 << 20, 50_000, 1 << 22][case] + 1000 * rank
        rng = oracle.SplitMix64(100 * case + rank)
        keys = rng.randomize(key, n, mode=1)
        if case == 1:
            keys &= np.uint32(0x0F0F0F0F)          # heavy duplicates: global stability is observable
        if case == 3:
            keys[:] = 0xDEADBEEF                   # all equal: one rank receives everything (capacity regrowth, collective)
        sizes = [None] * world
        dist.all_gather_object(sizes, n)
        first = sum(sizes[:rank])
        vals = (np.arange(n, dtype=np.uint64) + np.uint64(first)).astype(np.uint32)
        ops = CudaExchangeOps(key, 4, groups=groups)
        ds = DistributedRadixSort(dist, ops, key)
        dk = torch.from_numpy(keys.view(np.int32)).to(dev)
        dv = torch.from_numpy(vals.view(np.int32)).to(dev)
        for rep in range(3):  # repeated sorts on one context: epochs, buffer reuse
            rk, rv, cnt = ds.sort_pairs(dk, dv)
            torch.cuda.synchronize()
        assert bool((dk.cpu() == torch.from_numpy(keys.view(np.int32))).all()), "the inputs are only read"
        shares = [None] * world
        dist.all_gather_object(shares, (rk.cpu().numpy().view(np.uint32).copy(), rv.cpu().numpy().view(np.uint32).copy(),
                                        keys, vals))
        if rank == 0:
            got_k = np.concatenate([s[0] for s in shares])
            got_v = np.concatenate([s[1] for s in shares])
            ek, ev = oracle.sort(key, np.concatenate([s[2] for s in shares]), np.concatenate([s[3] for s in shares]))
            good = np.array_equal(got_k, ek) and np.array_equal(got_v, ev)
            print(f"case {case} key={key} groups={groups} shares={[len(s[0]) for s in shares]} -> {'OK' if good else 'MISMATCH'}", flush=True)
            ok = ok and good
        ops.close()
        dist.barrier()
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    if rank == 0:
        print("EXCHANGE-OK" if ok else "EXCHANGE-FAILED", flush=True)
    return 0 if int(flag.item()) else 1


if __name__ == "__main__":
    sys.exit(main())

"""dev_xchg_sweep.py — dev tool (not product, not the judged bench): ONE torchrun launch that times the
multi-GPU sort under several exchange configurations back to back on the same box, so that a sweep costs
one process start instead of one per configuration.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29501 \
        scripts/dev_xchg_sweep.py [--log2n-per-gpu 28] [--steps 5] [--warmup 3] CONFIG ...
CONFIG = scatter | x:<splits>[:lanes[:sort_sms]]   e.g.  x:1   x:0.5,1:1   x:eq4:1:120
(splits = cumulative fractions of an owner's items per group, THRS_XCHG_SPLITS, or eqN = N equal groups;
lanes = THRS_XCHG_LANES, 0 / absent = one stream per destination; sort_sms = THRS_XCHG_SORT_SMS).
Every configuration is verified with the complete on-GPU check of the bench."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--log2n-per-gpu", type=int, default=28)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("configs", nargs="+")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist

    import bench
    from tinyhipradixsort_b200 import multigpu as mg

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    g = vars(bench)
    n_l = 1 << args.log2n_per_gpu
    total = n_l * world
    keys0, vals0 = bench.gen_pairs(torch, rank * n_l, n_l, dev)
    keys, vals = torch.empty_like(keys0), torch.empty_like(vals0)
    xops = None
    sops = None
    for cfg in args.configs:
        for k in ("THRS_XCHG_SPLITS", "THRS_XCHG_LANES", "THRS_XCHG_SORT_SMS"):
            os.environ.pop(k, None)
        if cfg == "scatter":
            if sops is None:
                sops = mg.CudaLocalOps("u32", 4)
                sops.reserve(int(n_l * 1.05) + 4096)
            ops = sops
        else:
            parts = cfg.split(":")
            sp = parts[1]
            if sp.startswith("eq"):
                m = int(sp[2:])
                sp = ",".join("%.4f" % ((j + 1) / m) for j in range(m - 1)) + ("," if m > 1 else "") + "1"
            os.environ["THRS_XCHG_SPLITS"] = sp
            if len(parts) > 2:
                os.environ["THRS_XCHG_LANES"] = parts[2]
            if len(parts) > 3:
                os.environ["THRS_XCHG_SORT_SMS"] = parts[3]
            if xops is None:
                xops = mg.CudaExchangeOps("u32", 4, groups=3)
            ops = xops
        ds = mg.DistributedRadixSort(dist, ops, "u32")
        times, stage_sum = [], {}
        for i in range(args.warmup + args.steps):
            keys.copy_(keys0)
            vals.copy_(vals0)
            torch.cuda.synchronize()
            dist.barrier()
            t = mg._StageTimer(torch)
            rk, rv, cnt = ds.sort_pairs(keys, vals, t)
            st = t.read()
            ms = torch.tensor([st["total"]], dtype=torch.float64, device=dev)
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            if i >= args.warmup:
                times.append(float(ms.item()))
                for k, v in st.items():
                    stage_sum[k] = stage_sum.get(k, 0.0) + v
        ok = mg.verify_distributed(torch, dist, g, rk, rv, cnt, total)
        if rank == 0:
            avg = sum(times) / len(times)
            print(json.dumps({"config": cfg, "n_gpus": world, "ms_per_step": round(avg, 3), "min_ms": round(min(times), 3),
                              "gkeys_per_s": round(total / (avg * 1e-3) / 1e9, 2), "verified": ok,
                              "stage_ms_rank0": {k[:24]: round(v / len(times), 3) for k, v in stage_sum.items()}}), flush=True)
        del ds
    dist.barrier()
    for o in (xops, sops):
        if o is not None and hasattr(o, "close"):
            o.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

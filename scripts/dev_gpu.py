"""Developer check on the GPU box: quick parity + per-stage timing of C2 (+ reference GPU path)."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle  # noqa: E402
import tinyhipradixsort_b200 as thrs  # noqa: E402
from helpers import check_against_oracle, make_sorter  # noqa: E402
from test_large_gpu import gen_u32_pairs  # noqa: E402


def quick_parity():
    rng = oracle.SplitMix64(1)
    for key, vb in (("u32", None), ("u32", 4), ("f32", 4), ("u64", 8), ("f64", 16)):
        s = make_sorter(key, vb)
        for n in (1, 1000, 8192, 8193, 100000, 1 << 20):
            keys = rng.randomize(key, n, mode=1)
            vals = None
            if vb == 4:
                vals = np.arange(n, dtype=np.uint32)
            elif vb == 8:
                vals = np.arange(n, dtype=np.uint64)
            elif vb == 16:
                vals = np.zeros((n, 4), dtype=np.uint32); vals[:, 1] = np.arange(n)
            check_against_oracle(key, keys, vals, 0, None, False, s)
        print("parity ok", key, vb, flush=True)


def perf(n=1 << 28, pairs=True, reps=5, key="u32"):
    keys0, vals0 = gen_u32_pairs(n)
    s = make_sorter(key, 4 if pairs else None)
    d = s.getTemporaryBufferBytes(n)
    tmp = torch.empty(d.getTemporaryBufferBytesForSortPairs() if pairs else d.getTemporaryBufferBytesForSortKeys(),
                      dtype=torch.uint8, device="cuda")
    s.profile(True)
    keys, vals = keys0.clone(), vals0.clone()
    best = None
    for r in range(reps + 2):
        keys.copy_(keys0); vals.copy_(vals0)
        if pairs:
            s.sortPairs(keys, vals, n, tmp, 0, 32)
        else:
            s.sortKeys(keys, n, tmp, 0, 32)
        p = s.readProfile()
        if r >= 2 and (best is None or p["total_ms"] < best["total_ms"]):
            best = p
    bytes_total = n * 4 + 4 * 2 * n * (8 if pairs else 4)
    print(f"n=2^{int(np.log2(n))} pairs={pairs} {key}: total {best['total_ms']:.3f} ms  "
          f"memset {best['memset_ms']:.3f} hist {best['histogram_ms']:.3f} passes "
          + " ".join(f"{x:.3f}" for x in best["pass_ms"])
          + f"  => {n / best['total_ms'] / 1e6:.2f} Gkeys/s, {bytes_total / best['total_ms'] / 1e6:.0f} GB/s algorithmic",
          flush=True)
    return best


def ref_perf(n=1 << 28):
    if not oracle.RefGpuSorter.available():
        return
    keys0, vals0 = gen_u32_pairs(n)
    ref = oracle.RefGpuSorter("u32", 4, False)
    tmp = torch.empty(ref.temp_bytes(n, True), dtype=torch.uint8, device="cuda")
    keys, vals = keys0.clone(), vals0.clone()
    st = torch.cuda.current_stream().cuda_stream
    ts = []
    for r in range(4):
        keys.copy_(keys0); vals.copy_(vals0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        ref.sort(keys.data_ptr(), vals.data_ptr(), n, tmp.data_ptr(), 0, 32, st)
        b.record(); b.synchronize()
        ts.append(a.elapsed_time(b))
    print(f"reference GPU kernels (kernel.cu, sm_100a) n=2^{int(np.log2(n))} pairs: best {min(ts):.3f} ms "
          f"=> {n / min(ts) / 1e6:.2f} Gkeys/s", flush=True)


def perf_types(log2n=26):
    """Per-pass time of the other key/value instantiations on uniform random keys."""
    n = 1 << log2n
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    for key, vb in (("u32", 8), ("u32", 16), ("u64", None), ("u64", 4), ("u64", 8), ("f64", 16), ("f32", 4)):
        kb = 4 if key in ("u32", "f32") else 8
        k0 = torch.randint(-2**31, 2**31 - 1, (n * kb // 4,), dtype=torch.int32, device="cuda", generator=g)
        v0 = torch.zeros(n * (vb or 4) // 4, dtype=torch.int32, device="cuda")
        s = make_sorter(key, vb)
        d = s.getTemporaryBufferBytes(n)
        tmp = torch.empty(d.getTemporaryBufferBytesForSortPairs() if vb else d.getTemporaryBufferBytesForSortKeys(),
                          dtype=torch.uint8, device="cuda")
        s.profile(True)
        k, v = k0.clone(), v0.clone()
        best = None
        for r in range(5):
            k.copy_(k0); v.copy_(v0)
            if vb:
                s.sortPairs(k, v, n, tmp, 0, kb * 8)
            else:
                s.sortKeys(k, n, tmp, 0, kb * 8)
            p = s.readProfile()
            if r >= 1 and (best is None or p["total_ms"] < best["total_ms"]):
                best = p
        per_pass = sum(best["pass_ms"]) / len(best["pass_ms"])
        gbs = 2 * n * (kb + (vb or 0)) / per_pass / 1e6
        print(f"{key}+{vb}: n=2^{log2n} total {best['total_ms']:.3f} ms, {per_pass:.3f} ms/pass = {gbs:.0f} GB/s per pass, "
              f"{n / best['total_ms'] / 1e6:.2f} Gkeys/s (tile {s.tileItems()})", flush=True)


if __name__ == "__main__":
    what = sys.argv[1:] or ["parity", "perf", "ref"]
    print(torch.cuda.get_device_name(0), flush=True)
    if "parity" in what:
        quick_parity()
    if "perf" in what:
        perf(1 << 28, True)
        perf(1 << 28, False)
        perf(1 << 20, False)
    if "types" in what:
        perf_types()
    if "ref" in what:
        ref_perf()

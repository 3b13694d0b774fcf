#!/usr/bin/env python
"""bench.py — the judged benchmark of the radix-sort hot path.

A "step" is one stable sort of one batch of synthetic input through the C ABI of
libthrs_b200.so (thrs_sort_pairs: 1 histogram launch + 1 onesweep launch per 8-bit pass).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

N = 1 : config C2 of BASELINE.json — 2^28 uniform-random u32 keys + u32 values, bits [0,32).
N > 1 : the partitioned sort of config C5 (MSD splitters + all-to-all + local onesweep),
        2^28 pairs per GPU by default (weak scaling; --log2n-per-gpu 29 at N=8 is 2^32 total).

`value` is sorted keys/s with the input resident in HBM (CUDA events around the sort call
only, input restored outside the timed region, inputs of 2 GiB are far larger than the
126 MB L2).  `e2e` is the same metric through thrs_sort_host with HOST buffers: H2D copy of
keys+values, sort, D2H copy of keys+values inside the timed region.

--impl reference times the reference's own CPU check path (std::stable_sort on
pair<key,value>, unittest.cpp:358-377; all host threads via __gnu_parallel) from oracle/ on a
bounded sample of the same workload.  oracle/ is only ever used here as the CPU baseline and
as the checker of the last step, never inside the measured GPU path.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "sorted keys/sec (uint32 key-value, 2^28)"
UNIT = "keys/s"
KEY_BYTES = 4
VALUE_BYTES = 4
PASSES = 4


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def _traffic():
    """dram bytes per onesweep launch from the committed ncu capture, or None."""
    p = os.path.join(ROOT, "profiles", "ncu_summary.json")
    try:
        with open(p) as f:
            return json.load(f).get("onesweep_dram_bytes_per_launch")
    except Exception:
        return None


class ClockSampler:
    """Samples SM clock and throttle reasons with NVML while the timed region runs."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._nv = None

    def _run(self):
        nv = self._nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self._h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.01)

    def start(self):
        if self._nv is not None:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()

    def stop(self):
        self._stop.set()
        if self._thread is not None:
            self._thread.join()
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def _mix_torch(torch, i):
    """splitmix64_at(i) (unittest.cpp:24-35, counter form) on int64 tensors."""
    def c(x):
        return x - (1 << 64) if x >= (1 << 63) else x

    def shr(z, s):
        return (z >> s) & ((1 << (64 - s)) - 1)

    z = (i + 1) * c(0x9E3779B97F4A7C15)
    z = (z ^ shr(z, 30)) * c(0xBF58476D1CE4E5B9)
    z = (z ^ shr(z, 27)) * c(0x94D049BB133111EB)
    return z ^ shr(z, 31)


def gen_pairs(torch, first, count, device):
    """keys[i] = (u32)splitmix64_at(first+i), values[i] = (u32)(first+i)  (SURVEY.md 8d)."""
    keys = torch.empty(count, dtype=torch.int32, device=device)
    vals = torch.empty(count, dtype=torch.int32, device=device)
    chunk = 1 << 26
    for a in range(0, count, chunk):
        b = min(count, a + chunk)
        idx = torch.arange(first + a, first + b, dtype=torch.int64, device=device)
        keys[a:b] = (_mix_torch(torch, idx) & 0xFFFFFFFF).to(torch.int32)
        vals[a:b] = (idx & 0xFFFFFFFF).to(torch.int32)
    return keys, vals


def check_sorted_stable(torch, keys, vals):
    """Size-independent check of the timed result: regenerate the key from the value
    (permutation + association) and require (key, value) strictly increasing."""
    n = keys.shape[0]
    chunk = 1 << 26
    prev = None
    for a in range(0, n, chunk):
        b = min(n, a + chunk)
        k = keys[a:b].to(torch.int64) & 0xFFFFFFFF
        v = vals[a:b].to(torch.int64) & 0xFFFFFFFF
        if not bool(((_mix_torch(torch, v) & 0xFFFFFFFF) == k).all()):
            return False
        ok = (k[1:] > k[:-1]) | ((k[1:] == k[:-1]) & (v[1:] > v[:-1]))
        if not bool(ok.all()):
            return False
        if prev is not None and not (int(k[0]) > prev[0] or (int(k[0]) == prev[0] and int(v[0]) > prev[1])):
            return False
        prev = (int(k[-1]), int(v[-1]))
    return True


def cpu_sample(log2n):
    import numpy as np
    import oracle

    n = 1 << log2n
    keys = (oracle.splitmix64_at(np.arange(n)) & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    vals = np.arange(n, dtype=np.uint32)
    return keys, vals


def run_reference(args):
    """--impl reference: the reference's CPU check path on the host cores, bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import __graft_entry__ as ge

    ge.build(quiet=True, only_missing=True)
    import oracle

    # torchrun exports OMP_NUM_THREADS=1: count the cores this process may run on instead
    try:
        threads = len(os.sched_getaffinity(0))
    except AttributeError:
        threads = os.cpu_count() or 1
    threads = max(threads, oracle.max_threads())
    log2n = args.cpu_log2n
    keys0, vals0 = cpu_sample(log2n)
    n = keys0.shape[0]
    times = []
    for i in range(args.warmup + args.steps):
        k, v = keys0.copy(), vals0.copy()
        t = oracle.time_stable_sort_pairs_u32(k, v, threads)
        if i >= args.warmup:
            times.append(t)
    total = sum(times)
    value = n * len(times) / total
    sample = (f"2^{log2n} u32 key + u32 value pairs (same generator as the GPU workload), "
              f"__gnu_parallel::stable_sort on pair<key,value> by key, {threads} threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": f"C2 sample: stable sort of 2^{log2n} u32 key-value pairs on host cores",
                   "n_per_step": n, "bits": [0, 32]},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def run_ours(args):
    import torch

    import __graft_entry__ as ge

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libthrs_b200.so has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if rank == 0:
        ge.build(quiet=True, only_missing=True)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)
        dist.barrier()
    import tinyhipradixsort_b200 as thrs

    if world > 1:
        from tinyhipradixsort_b200 import multigpu

        return multigpu.bench(args, dist, rank, world, dev, globals())

    n = 1 << args.log2n
    keys0, vals0 = gen_pairs(torch, 0, n, dev)
    cfg = thrs.RadixSort.Config().configureWithKeyPair("u32", 4)
    sorter = thrs.RadixSort([], cfg)
    tmp = torch.empty(sorter.getTemporaryBufferBytes(n).getTemporaryBufferBytesForSortPairs(), dtype=torch.uint8,
                      device=dev)
    keys, vals = torch.empty_like(keys0), torch.empty_like(vals0)
    stream = torch.cuda.current_stream()

    def step():
        """One sort exactly as a user runs it (no events between the launches: the launches of one
        sort are chained by programmatic dependent launch)."""
        keys.copy_(keys0)
        vals.copy_(vals0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        sorter.sortPairs(keys, vals, n, tmp, 0, 32, stream)
        b.record(stream)
        launches = sorter.lastLaunchCount()
        b.synchronize()
        return a.elapsed_time(b), launches

    def profiled_step():
        """The same sort with CUDA events between its launches, on the launching stream (per-kernel
        durations for the roofline; programmatic dependent launch is off while events sit between
        the launches, so these steps are not the ones `value` is computed from)."""
        keys.copy_(keys0)
        vals.copy_(vals0)
        sorter.sortPairs(keys, vals, n, tmp, 0, 32, stream)
        return sorter.readProfile()  # blocks until this sort finished

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    sampler = ClockSampler(local_rank)
    sampler.start()
    step_ms, pass_ms, hist_ms, launches = [], [], [], 0
    for _ in range(args.steps):
        ms, l = step()
        step_ms.append(ms)
        launches += l
    torch.cuda.synchronize()
    sorter.profile(True)
    profiled_step()
    for _ in range(args.steps):
        prof = profiled_step()
        pass_ms.extend(prof["pass_ms"])
        hist_ms.append(prof["histogram_ms"])
    sorter.profile(False)
    torch.cuda.synchronize()
    clocks = sampler.stop()
    verified = check_sorted_stable(torch, keys, vals)

    total_ms = sum(step_ms)
    value = n * args.steps / (total_ms * 1e-3)

    # ---- end to end through the C ABI with host buffers (thrs_sort_host) ----
    hk0 = keys0.cpu().pin_memory()
    hv0 = vals0.cpu().pin_memory()
    hk, hv = torch.empty_like(hk0).pin_memory(), torch.empty_like(hv0).pin_memory()
    del keys, vals, tmp
    torch.cuda.empty_cache()
    sorter.profile(False)
    e2e_times = []
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    for i in range(1 + e2e_steps):
        hk.copy_(hk0)
        hv.copy_(hv0)
        t0 = time.perf_counter()
        sorter.sortHost(hk, hv, 0, 32)  # H2D keys+values, sort, D2H keys+values, blocking
        t1 = time.perf_counter()
        if i > 0:
            e2e_times.append(t1 - t0)
    e2e_ok = bool((hk[1:].to(torch.int64) & 0xFFFFFFFF >= hk[:-1].to(torch.int64) & 0xFFFFFFFF).all())
    e2e_value = n * len(e2e_times) / sum(e2e_times)

    # ---- CPU baseline: the oracle's restatement of the reference's CPU check path ----
    import oracle

    ck, cv = cpu_sample(args.cpu_log2n)
    cpu_s = oracle.time_stable_sort_pairs_u32(ck, cv, 1)
    cpu = {"value": ck.shape[0] / cpu_s, "unit": UNIT, "cores": 1, "kind": "port",
           "sample": f"2^{args.cpu_log2n} pairs of the same generator, std::stable_sort on pair<u32,u32> by key "
                     f"(unittest.cpp:358-377), 1 thread of {os.cpu_count()} host cores, {cpu_s:.2f} s"}

    peak, peak_src = _peaks()
    bytes_per_launch = 2.0 * n * (KEY_BYTES + VALUE_BYTES)
    avg_pass_ms = sum(pass_ms) / len(pass_ms)
    achieved = bytes_per_launch / (avg_pass_ms * 1e-3) / 1e9
    job_bytes = n * KEY_BYTES + PASSES * bytes_per_launch
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32", "data": "synthetic",
        "config": {"workload": f"C2: 2^{args.log2n} uniform-random u32 keys + u32 values (splitmix64 counter stream), "
                               "stable, bits [0,32), single B200",
                   "n_per_step": n, "bits": [0, 32], "passes": PASSES,
                   "l2": "inputs (2 GiB keys+values) are larger than L2; input restored outside the timed events",
                   "timing": "CUDA events on the launching stream around each sort call, summed over steps (launches of one sort chained by programmatic dependent launch); roofline.avg_launch_ms from the same number of additional steps with events between the launches"},
        "verified": bool(verified and e2e_ok),
        "roofline": {"bound": "hbm", "kernel": "onesweep_kernel (one 8-bit pass)", "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak, "traffic": _traffic(), "peak_source": peak_src,
                     "bytes_per_launch": bytes_per_launch, "avg_launch_ms": avg_pass_ms,
                     "histogram_ms": sum(hist_ms) / len(hist_ms),
                     "whole_sort_algorithmic_gbs": job_bytes / (total_ms / args.steps * 1e-3) / 1e9,
                     "whole_sort_frac": job_bytes / (total_ms / args.steps * 1e-3) / 1e9 / peak},
        "cpu_baseline": cpu,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": n * (KEY_BYTES + VALUE_BYTES),
                "d2h_bytes_per_step": n * (KEY_BYTES + VALUE_BYTES), "ms_per_step": 1e3 * sum(e2e_times) / len(e2e_times),
                "api": "thrs_sort_host (pinned host buffers)"},
        "gpu_launches": launches,
        "clocks": clocks,
    }
    print(json.dumps(line), flush=True)
    return 0 if verified and e2e_ok else 1


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--log2n", type=int, default=28, help="items per step at N=1")
    ap.add_argument("--log2n-per-gpu", type=int, default=28, help="items per GPU at N>1")
    ap.add_argument("--cpu-log2n", type=int, default=26, help="size of the bounded CPU sample")
    ap.add_argument("--e2e-steps", type=int, default=3)
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())

#!/usr/bin/env python
"""bench.py — the judged benchmark of the radix-sort hot path.

A "step" is one stable sort of one batch of synthetic input through the C ABI of
libthrs_b200.so (thrs_sort_pairs: 1 histogram launch + 1 onesweep launch per 8-bit pass).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

N = 1 : config C2 of BASELINE.json — 2^28 uniform-random u32 keys + u32 values, bits [0,32).
N > 1 : the partitioned sort of config C5 (top-digit pass = the sort's last radix pass, copy-engine bucket
        exchange over NVLink, segmented 3-pass sort of the received buckets; no collective on the data path),
        2^28 pairs per GPU by default (weak scaling); --log2n-total 32 is C5 as stated (2^32 pairs, strong).

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
    """(dram bytes per onesweep launch, where the number comes from).  The figure is STATIC: it is read
    from the committed `ncu --set full` capture (profiles/ncu_summary.json), not measured by this run."""
    p = os.path.join(ROOT, "profiles", "ncu_summary.json")
    try:
        with open(p) as f:
            d = json.load(f)
        return d.get("onesweep_dram_bytes_per_launch"), f"static: profiles/ncu_summary.json ({d.get('captured', 'round 1')}, ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum of one onesweep launch at C2)"
    except Exception:
        return None, "no committed ncu capture"


NORTH_STAR_PEAK_GBS = 8000.0  # the ~8 TB/s BASELINE.json's north_star quotes; the measured copy peak is lower


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


def _build_oracle_only():
    """The reference arm must not load the product: build (if missing) and use oracle/ alone."""
    import subprocess

    if not os.path.exists(os.path.join(ROOT, "oracle", "liboracle.so")):
        r = subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "liboracle.so"], capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("building oracle/liboracle.so failed")


def run_reference(args):
    """--impl reference: the reference's CPU check path on the host cores, at the SAME config as our
    arm (2^28 pairs per step by default; --cpu-log2n shrinks it for quick local runs)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    _build_oracle_only()
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
    # a few minutes at most: one full-size sort takes seconds on all host cores
    steps = args.steps if log2n < 28 else min(args.steps, 10)
    warmup = args.warmup if log2n < 28 else min(args.warmup, 1)
    for i in range(warmup + steps):
        k, v = keys0.copy(), vals0.copy()
        t = oracle.time_stable_sort_pairs_u32(k, v, threads)
        if i >= warmup:
            times.append(t)
    total = sum(times)
    value = n * len(times) / total
    sample = (f"2^{log2n} u32 key + u32 value pairs (same generator and size as the GPU workload), "
              f"__gnu_parallel::stable_sort on pair<key,value> by key, {threads} threads, {len(times)} timed sorts")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": len(times), "warmup": warmup, "ms_per_step": 1e3 * total / len(times),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": (f"C2: stable sort of 2^{log2n} uniform-random u32 key + u32 value pairs on host cores, bits [0,32)"
                                if args.gpus <= 1 else
                                f"C5 shape: each step sorts a bounded sample (2^{log2n} pairs, one GPU's slice) of the {args.gpus} x 2^{args.log2n_per_gpu}"
                                "-pair workload of the GPU arm on the host cores, bits [0,32); keys/s of a CPU sort falls slowly with n, so the sample flatters the CPU"),
                   "n_per_step": n, "bits": [0, 32]},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def _gen_case(torch, name, n, dev):
    """Inputs of SURVEY.md 8d's configs other than C2 (keys as int32 bit patterns, values = index)."""
    idx = torch.arange(0, n, dtype=torch.int64, device=dev)
    r = _mix_torch(torch, idx)
    vals = (idx & 0xFFFFFFFF).to(torch.int32)
    if name == "equal":
        keys = torch.full((n,), 0xDEADBEEF - (1 << 32), dtype=torch.int32, device=dev)
    elif name == "range16":
        keys = (r & 0xFFFF).to(torch.int32)
    elif name == "unique8":
        t8 = torch.tensor([0xDEADBEEF, 0, 0xFFFFFFFF, 0x80000000, 0x00010000, 0x12345678, 0x00FF00FF, 0x7FFFFFFF],
                          dtype=torch.int64, device=dev)
        keys = t8[(r & 7)].to(torch.int32)
    elif name == "f32_specials":
        # raw 32 random bits as float patterns (NaN/Inf/denormals occur naturally) + a sprinkle of specials
        k = r & 0xFFFFFFFF
        sp = torch.tensor([0x00000000, 0x80000000, 0x7F800000, 0xFF800000, 0x7FC00000, 0xFFC00000, 0x7F800001,
                           0xFF800001, 0x00000001, 0x80000001, 0x007FFFFF, 0x807FFFFF, 0x7F7FFFFF, 0xFF7FFFFF],
                          dtype=torch.int64, device=dev)
        k[::4096] = sp[(idx[::4096] >> 12) % sp.shape[0]]
        keys = k.to(torch.int32)
    else:  # uniform
        keys = (r & 0xFFFFFFFF).to(torch.int32)
    return keys, vals


def extra_cases(torch, thrs, dev, peak, reps=5):
    """Device-timed side cases: C1, C2 keys-only, C3 and the eight C4 cases.  Each is
    checked on the device (sorted on the window; pairs: equal fields keep ascending values) and
    reported as best-of-reps ms, algorithmic GB/s and the fraction of the measured copy peak."""
    out = {}

    def field(keys, kind, lo, hi):
        k = keys.to(torch.int64) & 0xFFFFFFFF
        if kind == "f32":
            k = torch.where(((k << 1) & 0xFFFFFFFF) == 0, torch.zeros_like(k), k)
            k = torch.where(k >= 0x80000000, k ^ 0xFFFFFFFF, k | 0x80000000)
        return (k >> lo) & ((1 << (hi - lo)) - 1)

    def run(tag, kind, gen, log2n, pairs, lo, hi):
        n = 1 << log2n
        keys0, vals0 = _gen_case(torch, gen, n, dev)
        cfg = thrs.RadixSort.Config()
        cfg = cfg.configureWithKeyPair(kind, 4) if pairs else cfg.configureWithKey(kind)
        srt = thrs.RadixSort([], cfg)
        d = srt.getTemporaryBufferBytes(n)
        tmp = torch.empty(d.getTemporaryBufferBytesForSortPairs() if pairs else d.getTemporaryBufferBytesForSortKeys(),
                          dtype=torch.uint8, device=dev)
        keys, vals = torch.empty_like(keys0), torch.empty_like(vals0)
        st = torch.cuda.current_stream()
        best = None
        for _ in range(reps + 2):
            keys.copy_(keys0)
            vals.copy_(vals0)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(st)
            if pairs:
                srt.sortPairs(keys, vals, n, tmp, lo, hi, st)
            else:
                srt.sortKeys(keys, n, tmp, lo, hi, st)
            b.record(st)
            b.synchronize()
            ms = a.elapsed_time(b)
            best = ms if best is None or ms < best else best
        f = field(keys, kind, lo, hi)
        ok = bool((f[1:] >= f[:-1]).all())
        if pairs and ok:
            v = vals.to(torch.int64) & 0xFFFFFFFF
            ok = bool(((f[1:] > f[:-1]) | (v[1:] > v[:-1])).all())
            # association: the key that travelled with value v is the generated key of index v
            ok = ok and bool((keys0[v] == keys).all())
        passes = (hi - lo) // 8
        alg = n * 4 + passes * 2 * n * (8 if pairs else 4)
        out[tag] = {"n": n, "pairs": pairs, "bits": [lo, hi], "ms": best, "gbs_algorithmic": alg / (best * 1e-3) / 1e9,
                    "frac_of_measured_peak": alg / (best * 1e-3) / 1e9 / peak, "gkeys_per_s": n / (best * 1e-3) / 1e9,
                    "verified": ok}
        del keys, vals, keys0, vals0, tmp
        torch.cuda.empty_cache()
        return ok

    ok = run("C1_2pow20_u32_keys", "u32", "uniform", 20, False, 0, 32)
    ok &= run("C2_2pow28_u32_keys_only", "u32", "uniform", 28, False, 0, 32)
    ok &= run("C3_2pow28_f32_pairs_specials", "f32", "f32_specials", 28, True, 0, 32)
    for gen in ("equal", "range16", "unique8"):
        for pairs in (False, True):
            ok &= run(f"C4_2pow27_{gen}_{'pairs' if pairs else 'keys'}", "u32", gen, 27, pairs, 0, 32)
    for pairs in (False, True):
        ok &= run(f"C4_2pow27_window_8_24_{'pairs' if pairs else 'keys'}", "u32", "uniform", 27, pairs, 8, 24)
    return out, bool(ok)


def reference_gpu_leg(torch, dev, keys0, vals0, n, ours_keys, ours_vals):
    """The reference's own GPU kernels (kernel.cu compiled unmodified into oracle/_ref/kernel_u32_v4_asc.cubin,
    driven by oracle/ref_gpu.cu with the launch shapes of tinyhipradixsort.hpp:864-943) on the same
    input and GPU, OUTSIDE our timed region: ms of one sort (best of 3, CUDA events) and a bit-for-bit
    comparison with our result.  None where the cubins were not built (no /root/reference)."""
    try:
        import oracle

        if not oracle.RefGpuSorter.available():
            return None
        ref = oracle.RefGpuSorter("u32", 4, False)
    except Exception as e:  # noqa: BLE001
        return {"unavailable": str(e)[:200]}
    tmp = torch.empty(ref.temp_bytes(n, True), dtype=torch.uint8, device=dev)
    k, v = torch.empty_like(keys0), torch.empty_like(vals0)
    st = torch.cuda.current_stream()
    best = None
    for _ in range(3):
        k.copy_(keys0)
        v.copy_(vals0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        ref.sort(k.data_ptr(), v.data_ptr(), n, tmp.data_ptr(), 0, 32, st.cuda_stream)
        b.record(st)
        b.synchronize()
        ms = a.elapsed_time(b)
        best = ms if best is None or ms < best else best
    same = bool((k == ours_keys).all()) and bool((v == ours_vals).all())
    return {"ms": best, "keys_per_s": n / (best * 1e-3), "kernel": "reference kernel.cu, unmodified, sm_100a cubin (oracle/_ref)",
            "launches": 12, "launch_shapes": "tinyhipradixsort.hpp:864-943 (blockCount, prefixSumExclusiveInplace, reorderKeyPair per pass)",
            "identical_to_ours": same}


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

        g = dict(globals())
        g["_traffic"] = lambda: _traffic()[0]
        return multigpu.bench(args, dist, rank, world, dev, g)

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
    # the result of the LAST TIMED step (programmatic dependent launch chain) is what gets checked ...
    verified = check_sorted_stable(torch, keys, vals)
    ref_gpu = None if args.no_reference_gpu else reference_gpu_leg(torch, dev, keys0, vals0, n, keys, vals)
    sorter.profile(True)
    profiled_step()
    for _ in range(args.steps):
        prof = profiled_step()
        pass_ms.extend(prof["pass_ms"])
        hist_ms.append(prof["histogram_ms"])
    sorter.profile(False)
    torch.cuda.synchronize()
    clocks = sampler.stop()
    # ... and so is the result of the last profiled step (events between the launches, no PDL)
    verified = verified and check_sorted_stable(torch, keys, vals)
    if ref_gpu and "identical_to_ours" in ref_gpu:
        verified = verified and ref_gpu["identical_to_ours"]
    ranking = sorter.getRanking()
    rcheck = sorter.getRankingCheck()

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
    e2e_block_value = n * len(e2e_times) / sum(e2e_times)
    e2e_block_ms = 1e3 * sum(e2e_times) / len(e2e_times)
    e2e_value, e2e_ms, e2e_api = e2e_block_value, e2e_block_ms, "thrs_sort_host (pinned host buffers), one blocking call per step"
    e2e_stream = None

    # ---- the same, streaming: thrs_sort_host_submit / _wait with 3 sorts in flight.  Every step still uploads its
    # own unsorted 2 GiB from pinned host memory and downloads its own sorted 2 GiB inside the timed region (which
    # includes filling and draining the pipeline); the upload of step i+1 overlaps the download of step i-1.
    # Distinct pinned input buffers per step, so that no step sorts data a previous step already sorted.
    if args.e2e_stream_steps > 0:
        depth = min(3, thrs.RadixSort.HOST_SLOTS)
        bufs = []
        for _ in range(args.e2e_stream_steps):
            bk, bv = torch.empty_like(hk0).pin_memory(), torch.empty_like(hv0).pin_memory()
            bk.copy_(hk0)
            bv.copy_(hv0)
            bufs.append((bk, bv))
        for sl in range(depth):  # untimed: the slots' device workspaces are allocated on first use
            hk.copy_(hk0)
            hv.copy_(hv0)
            sorter.sortHostSubmit(sl, hk, hv, 0, 32)
            sorter.sortHostWait(sl)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i, (bk, bv) in enumerate(bufs):
            sl = i % depth
            sorter.sortHostWait(sl)  # the batch submitted `depth` steps ago is back in its host buffers
            sorter.sortHostSubmit(sl, bk, bv, 0, 32)
        for sl in range(depth):
            sorter.sortHostWait(sl)
        t1 = time.perf_counter()
        stream_ok = True
        for bk, bv in bufs:  # every step's result, complete check on the device
            stream_ok = stream_ok and check_sorted_stable(torch, bk.to(dev), bv.to(dev))
        del bufs
        e2e_stream = {"value": n * args.e2e_stream_steps / (t1 - t0), "ms_per_step": 1e3 * (t1 - t0) / args.e2e_stream_steps,
                      "steps": args.e2e_stream_steps, "in_flight": depth, "verified": bool(stream_ok)}
        e2e_ok = e2e_ok and stream_ok
        e2e_value, e2e_ms = e2e_stream["value"], e2e_stream["ms_per_step"]
        e2e_api = (f"thrs_sort_host_submit / thrs_sort_host_wait (pinned host buffers), {depth} sorts in flight: every step's H2D and D2H "
                   f"are inside the timed region (pipeline fill and drain included, {args.e2e_stream_steps} steps, distinct unsorted inputs); "
                   "the upload of step i+1 overlaps the download of step i-1")

    peak, peak_src = _peaks()
    extra, extra_ok = ({}, True) if args.no_extra else extra_cases(torch, thrs, dev, peak)

    # ---- CPU baseline: the oracle's restatement of the reference's CPU check path, at the full C2 size ----
    import oracle

    ck, cv = cpu_sample(args.cpu_log2n)
    cpu_s = oracle.time_stable_sort_pairs_u32(ck, cv, 1)
    cpu = {"value": ck.shape[0] / cpu_s, "unit": UNIT, "cores": 1, "kind": "port",
           "sample": f"2^{args.cpu_log2n} pairs of the same generator ({'the full C2 workload, timed once' if args.cpu_log2n == args.log2n else 'a sample'}), "
                     f"std::stable_sort on pair<u32,u32> by key (unittest.cpp:358-377), 1 thread of {os.cpu_count()} host cores, {cpu_s:.2f} s"}
    del ck, cv

    traffic, traffic_src = _traffic()
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
                   "extra": extra,
                   "timing": "CUDA events on the launching stream around each sort call, summed over steps (launches of one sort chained by programmatic dependent launch); roofline.avg_launch_ms from the same number of additional steps with events between the launches"},
        "verified": bool(verified and e2e_ok and extra_ok),
        "roofline": {"bound": "hbm", "kernel": "onesweep_kernel (one 8-bit pass)", "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                     "peak_source": peak_src, "frac_of_8TBs": achieved / NORTH_STAR_PEAK_GBS,
                     "bytes_per_launch": bytes_per_launch, "avg_launch_ms": avg_pass_ms,
                     "histogram_ms": sum(hist_ms) / len(hist_ms),
                     "whole_sort_algorithmic_gbs": job_bytes / (total_ms / args.steps * 1e-3) / 1e9,
                     "whole_sort_frac": job_bytes / (total_ms / args.steps * 1e-3) / 1e9 / peak,
                     "whole_sort_frac_of_8TBs": job_bytes / (total_ms / args.steps * 1e-3) / 1e9 / NORTH_STAR_PEAK_GBS},
        "reference_gpu": ref_gpu,
        "ranking": {"in_use": "counter-return" if ranking[0] == thrs.RadixSort.RANKING_COUNTER_RETURN else "match-mask",
                    "probe_in_order": ranking[1], "probe_groups": ranking[2], "production_check_ok": rcheck[0],
                    "production_check_items": rcheck[1]},
        "cpu_baseline": cpu,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": n * (KEY_BYTES + VALUE_BYTES),
                "d2h_bytes_per_step": n * (KEY_BYTES + VALUE_BYTES), "ms_per_step": e2e_ms, "api": e2e_api,
                "blocking": {"value": e2e_block_value, "ms_per_step": e2e_block_ms, "steps": len(e2e_times),
                             "api": "thrs_sort_host: H2D, sort, D2H strictly one after the other, one call per step"},
                "streaming": e2e_stream},
        "gpu_launches": launches,
        "clocks": clocks,
    }
    print(json.dumps(line), flush=True)
    return 0 if verified and e2e_ok and extra_ok else 1


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--log2n", type=int, default=28, help="items per step at N=1")
    ap.add_argument("--log2n-per-gpu", type=int, default=28, help="items per GPU at N>1")
    ap.add_argument("--cpu-log2n", type=int, default=28, help="size of the CPU baseline sort (28 = the full C2 workload)")
    ap.add_argument("--no-extra", action="store_true", help="skip the device-timed side cases (config.extra)")
    ap.add_argument("--no-reference-gpu", action="store_true", help="skip the reference's GPU kernels leg")
    ap.add_argument("--log2n-total", type=int, default=0, help="N>1: total items over all GPUs (strong scaling, C5 = 32)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--e2e-stream-steps", type=int, default=6,
                    help="steps of the streaming end-to-end measurement (thrs_sort_host_submit/_wait; 2 GiB of pinned host memory each); 0 = blocking calls only")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())

"""Partitioned sort over the GPUs of one NVSwitch box (config C5 of BASELINE.json): one process
per GPU, `torch.distributed` for the plumbing, libthrs_b200.so for every pass over the data.

The reference has no multi-GPU path (single device 0: unittest.cpp:22,51); this layer follows
SURVEY.md 8(e):

  1. every rank histograms the TOP 8-bit digit of getKeyBits(key) of its slice
     (thrs_digit_histogram: the sort's own single-read histogram kernel);
  2. all_gather of the 256-bin histograms (2 KiB per rank) -> identical table on every rank;
  3. splitters on digit boundaries: bucket b goes to rank floor(P * items_before(b) / N), so
     equal keys never straddle ranks and the destination is monotone in the digit;
  4. one stable onesweep pass on the top digit groups the slice by destination without
     reordering anything inside a destination;
  5. exchange; chunks land in SOURCE-RANK order.  On GPUs steps 4+5 are ONE kernel: the pass of
     step 4 stores every digit run straight into its destination rank's receive buffer over
     NVLink (thrs_scatter_pairs with CUDA-IPC peer pointers, `scatter_exchange`), followed by a
     one-word all_reduce as the "everyone has written" barrier.  Hosts without peer memory (the
     gloo CPU tests) run step 4 locally and then an all_to_all_single per array;
  6. local 4-pass onesweep over what was received.

Concatenating the ranks' outputs gives exactly std::stable_sort of the concatenated inputs:
ties are ordered by (source rank, position in the source slice) because step 4 and 6 are stable
and step 5 keeps source order.  Counts are 64-bit at this layer (2^32 pairs do not fit the
uint32 n of the single-GPU API); a rank's share must stay below 2^32 items.

The local device work is injected (`LocalOps`) so that the host logic — splitters, count
matrices, the exchange — can be exercised on CPU with gloo in tests/ (where a CPU checker stands in
for the kernels).  The product default is `CudaLocalOps`; nothing falls back automatically.
"""
import numpy as np

KEY_BITS = {"u32": 32, "f32": 32, "u64": 64, "f64": 64}


def plan_partition(all_hists, world):
    """all_hists: int64 array [world][256] (digit histograms in sort order).
    Returns (dest_of_bucket[256], send_matrix[src][dst]) with 64-bit counts."""
    h = np.asarray(all_hists, dtype=np.int64).reshape(world, 256)
    g = h.sum(axis=0)
    total = int(g.sum())
    before = np.concatenate(([0], np.cumsum(g)[:-1]))
    if total == 0:
        dest = np.zeros(256, dtype=np.int64)
    else:
        dest = np.minimum(world - 1, (before * world) // total).astype(np.int64)
    send = np.zeros((world, world), dtype=np.int64)
    for d in range(world):
        send[:, d] = h[:, dest == d].sum(axis=1)
    return dest, send


def scatter_tables(dest, send, me):
    """Tables of the fused exchange for source rank `me`: (bin_first_digit[world], first[world]).
    Bin r = the digits whose destination is rank r (dest is monotone in the digit); rank me's
    items for rank r land in r's receive buffer from element sum(send[:me, r]) on, in input order,
    so chunks lie in source-rank order."""
    world = send.shape[0]
    bin_first = np.array([int(np.searchsorted(dest, r, side="left")) for r in range(world)], dtype=np.int64)
    first = np.array([int(send[:me, r].sum()) for r in range(world)], dtype=np.int64)
    return bin_first, first


class CudaLocalOps:
    """Device work of one rank, through the C ABI (Python mirror of thrs::RadixSort)."""

    def __init__(self, key="u32", value_bytes=4, descending=False):
        import torch

        from . import RadixSort, SortOrder

        self.torch = torch
        self.key = key
        cfg = RadixSort.Config().configureWithKeyPair(key, value_bytes)
        cfg.sortOrder = SortOrder.Descending if descending else SortOrder.Ascending
        self.sorter = RadixSort([], cfg)
        self._tmp = None

    def _temp(self, n):
        need = self.sorter.getTemporaryBufferBytes(n).getTemporaryBufferBytesForSortPairs()
        if self._tmp is None or self._tmp.numel() < need:
            self._tmp = None
            self._tmp = self.torch.empty(need, dtype=self.torch.uint8, device="cuda")
        return self._tmp

    def reserve(self, n):
        self._temp(n)

    def top_digit_histogram(self, keys, n):
        out = self.torch.zeros(256, dtype=self.torch.int32, device=keys.device)
        if n:
            self.sorter.digitHistogram(keys, n, KEY_BITS[self.key] - 8, out, self._temp(n))
        return out.to(self.torch.int64) & 0xFFFFFFFF

    def sort_pairs(self, keys, vals, n, start, end):
        if n:
            self.sorter.sortPairs(keys, vals, n, self._temp(n), start, end)

    def launches(self):
        return self.sorter.lastLaunchCount()

    # ---- peer-memory exchange (one process per GPU, CUDA IPC) ----
    def exchange_alloc(self, cap_items, kbytes, vbytes):
        """cudaMalloc'ed receive buffer [keys cap | values cap] and its IPC handle."""
        from .radixsort import Buffer, ipc_export

        self._xcap = int(cap_items)
        self._xkoff = 0
        self._xvoff = (self._xcap * kbytes + 255) & ~255
        self._xbuf = Buffer(self._xvoff + self._xcap * vbytes)
        return ipc_export(self._xbuf.data())

    def exchange_open(self, handles, me):
        """Device addresses of every rank's receive buffer in this process (own one included)."""
        from .radixsort import ipc_close, ipc_open

        for p in getattr(self, "_xpeers_opened", []):
            ipc_close(p)
        self._xpeers_opened = []
        bases = []
        for r, h in enumerate(handles):
            if r == me:
                bases.append(self._xbuf.data())
            else:
                p = ipc_open(h)
                self._xpeers_opened.append(p)
                bases.append(p)
        self._xbases = bases
        return bases

    def exchange_views(self, count, key_dtype, val_dtype):
        """torch views of the first `count` received keys / values."""
        torch = self.torch

        class _Raw:
            def __init__(self, ptr, n, typestr):
                self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}

        ts = {torch.int32: "<i4", torch.int64: "<i8"}
        k = torch.as_tensor(_Raw(self._xbuf.data() + self._xkoff, max(count, 1), ts[key_dtype]), device="cuda")
        v = torch.as_tensor(_Raw(self._xbuf.data() + self._xvoff, max(count, 1), ts[val_dtype]), device="cuda")
        return k[:count], v[:count]

    def scatter_exchange(self, keys, vals, n, start_bit, bin_first, first):
        """Steps 4+5 fused: this rank's items for rank r -> r's receive buffer from element first[r] on."""
        kd = np.array([b + self._xkoff for b in self._xbases], dtype=np.uint64)
        vd = np.array([b + self._xvoff for b in self._xbases], dtype=np.uint64)
        if n:
            self.sorter.scatterPairs(keys, vals, n, self._temp(n), start_bit, bin_first, kd, vd, first)


class DistributedRadixSort:
    """sort_pairs(keys, values) over all ranks of `group`.  Returns (keys_out, values_out, count):
    views of this rank's receive buffers holding its sorted share."""

    def __init__(self, dist, local_ops, key="u32", group=None):
        self.dist = dist
        self.ops = local_ops
        self.key = key
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.last_plan = None
        self._recv = None
        self._xcap = 0
        self.fused = hasattr(local_ops, "scatter_exchange")

    def _recv_buffers(self, like_k, like_v, count):
        import torch

        if self._recv is None or self._recv[0].shape[0] < count:
            self._recv = None
            cap = int(count * 1.02) + 1024  # a little slack so repeated sorts do not reallocate
            self._recv = (torch.empty(cap, dtype=like_k.dtype, device=like_k.device),
                          torch.empty((cap,) + tuple(like_v.shape[1:]), dtype=like_v.dtype, device=like_v.device))
        return self._recv

    def sort_pairs(self, keys, vals, timers=None):
        import torch

        dist, ops, P = self.dist, self.ops, self.world
        n = int(keys.shape[0])
        bits = KEY_BITS[self.key]

        def mark(name):
            if timers is not None:
                timers.mark(name)

        # 1-2: top-digit histogram of the slice, shared with everyone
        hist = ops.top_digit_histogram(keys, n)
        gathered = [torch.empty_like(hist) for _ in range(P)]
        dist.all_gather(gathered, hist, group=self.group)
        all_h = torch.stack(gathered).cpu().numpy()
        mark("histogram+allgather")
        # 3: splitters on digit boundaries
        dest, send = plan_partition(all_h, P)
        self.last_plan = (dest, send)
        send_counts = [int(x) for x in send[self.rank]]
        recv_counts = [int(x) for x in send[:, self.rank]]
        total_recv = sum(recv_counts)
        if total_recv >= 1 << 32:
            raise ValueError("a rank's share must stay below 2^32 items")
        if self.fused:
            return self._fused_exchange_and_sort(keys, vals, n, bits, all_h, dest, send, total_recv, mark)
        # 4: stable pass on the top digit = stable partition by destination
        ops.sort_pairs(keys, vals, n, bits - 8, bits)
        mark("partition pass")
        # 5: all-to-all(v); chunks land in source-rank order
        rk, rv = self._recv_buffers(keys, vals, total_recv)
        rk, rv = rk[:total_recv], rv[:total_recv]
        dist.all_to_all_single(rk, keys[:n], recv_counts, send_counts, group=self.group)
        dist.all_to_all_single(rv, vals[:n], recv_counts, send_counts, group=self.group)
        mark("all-to-all")
        # 6: local sort of the received share
        ops.sort_pairs(rk, rv, total_recv, 0, bits)
        mark("local sort")
        return rk, rv, total_recv


    def _fused_exchange_and_sort(self, keys, vals, n, bits, all_h, dest, send, total_recv, mark):
        import torch

        dist, ops, P = self.dist, self.ops, self.world
        # receive buffers: same capacity on every rank (a pure function of the shared send matrix),
        # re-exported only when they have to grow
        need = int(send.sum(axis=0).max())
        if need > self._xcap:
            cap = int(need * 1.02) + 1024
            kb, vb = keys.element_size(), vals.element_size() * int(np.prod(vals.shape[1:], dtype=np.int64))
            torch.cuda.synchronize()
            handle = ops.exchange_alloc(cap, kb, vb)
            handles = [None] * P
            dist.all_gather_object(handles, handle, group=self.group)
            ops.exchange_open(handles, self.rank)
            self._xcap = cap
        if P > 8:
            raise ValueError("the fused exchange splits into at most 8 destinations (one NVSwitch box)")
        bin_first, first = scatter_tables(dest, send, self.rank)
        # 4+5: the partition pass stores straight into the destination ranks' receive buffers
        ops.scatter_exchange(keys, vals, n, bits - 8, bin_first, first)
        # "everyone has written": a one-word all_reduce ordered after the scatter kernel on every rank
        flag = torch.ones(1, dtype=torch.int32, device=keys.device)
        dist.all_reduce(flag, group=self.group)
        mark("partition+exchange (peer stores)")
        rk, rv = ops.exchange_views(total_recv, keys.dtype, vals.dtype)
        # 6: local sort of the received share
        ops.sort_pairs(rk, rv, total_recv, 0, bits)
        mark("local sort")
        return rk, rv, total_recv


class _StageTimer:
    """CUDA events between the stages of one distributed sort, on the launching stream."""

    def __init__(self, torch):
        self.torch = torch
        self.events = [("start", self._ev())]

    def _ev(self):
        e = self.torch.cuda.Event(enable_timing=True)
        e.record(self.torch.cuda.current_stream())
        return e

    def mark(self, name):
        self.events.append((name, self._ev()))

    def read(self):
        self.events[-1][1].synchronize()
        out = {}
        for (_, a), (name, b) in zip(self.events[:-1], self.events[1:]):
            out[name] = a.elapsed_time(b)
        out["total"] = self.events[0][1].elapsed_time(self.events[-1][1])
        return out


def verify_distributed(torch, dist, g, rk, rv, count, total):
    """Complete check of SURVEY.md 8(c) for C5: key_out == gen(value_out), (key, value) strictly
    increasing inside every rank and across rank boundaries, counts add up to N."""
    ok = g["check_sorted_stable"](torch, rk[:count], rv[:count]) if count else True
    world, rank = dist.get_world_size(), dist.get_rank()
    edge = torch.zeros(5, dtype=torch.int64, device=rk.device)
    if count:
        edge[0] = rk[0].to(torch.int64) & 0xFFFFFFFF
        edge[1] = rv[0].to(torch.int64) & 0xFFFFFFFF
        edge[2] = rk[count - 1].to(torch.int64) & 0xFFFFFFFF
        edge[3] = rv[count - 1].to(torch.int64) & 0xFFFFFFFF
    edge[4] = count
    edges = [torch.empty_like(edge) for _ in range(world)]
    dist.all_gather(edges, edge)
    e = torch.stack(edges).cpu().numpy()
    prev = None
    for r in range(world):
        if e[r][4] == 0:
            continue
        first, last = (int(e[r][0]), int(e[r][1])), (int(e[r][2]), int(e[r][3]))
        if prev is not None and not first > prev:
            ok = False
        prev = last
    if int(e[:, 4].sum()) != total:
        ok = False
    flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=rk.device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    return bool(int(flag.item()))


def bench(args, dist, rank, world, dev, g):
    """bench.py at N > 1: the partitioned sort, `--log2n-per-gpu` pairs per GPU (weak scaling)."""
    import json
    import time

    import torch

    n_l = 1 << args.log2n_per_gpu
    total = n_l * world
    keys0, vals0 = g["gen_pairs"](torch, rank * n_l, n_l, dev)
    ops = CudaLocalOps("u32", 4)
    ops.reserve(int(n_l * 1.05) + 4096)
    ds = DistributedRadixSort(dist, ops, "u32")
    keys, vals = torch.empty_like(keys0), torch.empty_like(vals0)
    stream = torch.cuda.current_stream()

    def step():
        keys.copy_(keys0)
        vals.copy_(vals0)
        torch.cuda.synchronize()
        dist.barrier()
        t = _StageTimer(torch)
        rk, rv, cnt = ds.sort_pairs(keys, vals, t)
        stages = t.read()
        ms = torch.tensor([stages["total"]], dtype=torch.float64, device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)  # max over ranks, timed on the device
        return float(ms.item()), stages, rk, rv, cnt

    warm = max(args.warmup, 3)
    for _ in range(warm):
        step()
    sampler = g["ClockSampler"](dev.index)
    sampler.start()
    step_ms, stage_sum = [], {}
    for _ in range(args.steps):
        ms, stages, rk, rv, cnt = step()
        step_ms.append(ms)
        for k, v in stages.items():
            stage_sum[k] = stage_sum.get(k, 0.0) + v
    torch.cuda.synchronize()
    clocks = sampler.stop()
    verified = verify_distributed(torch, dist, g, rk, rv, cnt, total)
    total_ms = sum(step_ms)
    value = total * args.steps / (total_ms * 1e-3)

    # per-pass device time of the local sort (dominant kernel), measured live with the sorter's events
    ops.sorter.profile(True)
    keys.copy_(keys0)
    vals.copy_(vals0)
    ds.sort_pairs(keys, vals)
    prof = ops.sorter.readProfile()
    ops.sorter.profile(False)
    launches_per_step = 1 + 1 + 6  # top-digit histogram, scatter pass, local sort (zero + histogram + 4)

    # end to end: pinned host slices in, sorted share out (H2D + sort + D2H inside the timed region)
    hk0, hv0 = keys0.cpu().pin_memory(), vals0.cpu().pin_memory()
    hk_out = torch.empty(int(n_l * 1.05) + 4096, dtype=hk0.dtype).pin_memory()
    hv_out = torch.empty_like(hk_out).pin_memory()
    e2e = []
    for i in range(1 + max(1, min(args.steps, args.e2e_steps))):
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        keys.copy_(hk0, non_blocking=True)
        vals.copy_(hv0, non_blocking=True)
        rk, rv, cnt = ds.sort_pairs(keys, vals)
        hk_out[:cnt].copy_(rk[:cnt], non_blocking=True)
        hv_out[:cnt].copy_(rv[:cnt], non_blocking=True)
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        if i > 0:
            e2e.append(float(dt.item()))
    e2e_value = total * len(e2e) / sum(e2e)

    if rank == 0:
        peak, peak_src = g["_peaks"]()
        per_gpu_bytes = 2.0 * cnt * 8
        avg_pass = sum(prof["pass_ms"]) / len(prof["pass_ms"])
        achieved = per_gpu_bytes / (avg_pass * 1e-3) / 1e9
        send_bytes = n_l * 8 * (world - 1) / world
        fused = "partition+exchange (peer stores)" in stage_sum
        a2a_ms = stage_sum.get("partition+exchange (peer stores)" if fused else "all-to-all", 0.0) / args.steps
        line = {
            "metric": g["METRIC"], "value": value, "unit": g["UNIT"], "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": f"C5 shape: {world} x 2^{args.log2n_per_gpu} uniform-random u32 key + u32 value pairs, "
                                   "MSD splitters + partition pass storing into peer memory over NVLink + local onesweep, stable, bits [0,32)",
                       "n_total": total, "n_per_gpu": n_l, "parallelism": f"partitioned sort over {world} ranks (NCCL)",
                       "l2": "per-GPU inputs (>= 2 GiB) are larger than L2; input restored outside the timed events",
                       "stage_ms_rank0": {k: v / args.steps for k, v in stage_sum.items()},
                       "all_to_all": {"how": "fused into the partition kernel (P2P stores, CUDA IPC)" if fused else "NCCL all_to_all_single",
                                      "bytes_sent_per_gpu": send_bytes, "ms": a2a_ms,
                                      "GBps_per_direction_per_gpu": (send_bytes / (a2a_ms * 1e-3) / 1e9) if a2a_ms else None,
                                      "nvlink_peak_GBps": 900.0}},
            "verified": verified,
            "roofline": {"bound": "hbm", "kernel": "onesweep_kernel (one 8-bit pass of the local sort, rank 0)",
                         "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": g["_traffic"](), "peak_source": peak_src, "bytes_per_launch": per_gpu_bytes,
                         "avg_launch_ms": avg_pass},
            "e2e": {"value": e2e_value, "unit": g["UNIT"], "h2d_bytes_per_step": total * 8, "d2h_bytes_per_step": total * 8,
                    "ms_per_step": 1e3 * sum(e2e) / len(e2e), "api": "DistributedRadixSort.sort_pairs (pinned host slices)"},
            "gpu_launches": launches_per_step * args.steps * world,
            "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0 if verified else 1


class SingleProcessMultiGpuSort:
    """thrs_mg_*: the same partitioned sort driven from ONE process over peer-capable devices
    (device ids may repeat, which emulates several ranks on one GPU for tests)."""

    def __init__(self, device_ids, key="u32", value_bytes=4, descending=False):
        import ctypes

        from . import _lib
        from .radixsort import KeyType, SortOrder, _check

        self._ct = ctypes
        self._lib = _lib.load()
        cfg = _lib.thrs_config()
        self._lib.thrs_config_default(ctypes.byref(cfg))
        cfg.key_type = {"u32": 0, "u64": 1, "f32": 2, "f64": 3}[key] if not isinstance(key, KeyType) else int(key)
        cfg.value_type = {4: 0, 8: 1, 16: 2}[value_bytes]
        cfg.sort_order = 1 if descending else 0
        self.ids = list(device_ids)
        arr = (ctypes.c_int * len(self.ids))(*self.ids)
        self._h = ctypes.c_void_p()
        rc = self._lib.thrs_mg_create(len(self.ids), arr, ctypes.byref(cfg), ctypes.byref(self._h))
        if rc:
            raise RuntimeError(f"thrs_mg_create rc={rc}: {self._lib.thrs_mg_last_error().decode()}")

    def sort_pairs(self, keys, values, out_keys, out_values):
        """keys/values: lists of CUDA tensors (one per device slot); out_*: lists of output tensors.
        Returns the list of item counts that landed in each output."""
        ct = self._ct
        n = len(self.ids)
        vp = ct.c_void_p * n
        u64 = ct.c_uint64 * n
        k = vp(*[t.data_ptr() for t in keys])
        v = vp(*[t.data_ptr() for t in values])
        ok = vp(*[t.data_ptr() for t in out_keys])
        ov = vp(*[t.data_ptr() for t in out_values])
        counts = u64(*[int(t.shape[0]) for t in keys])
        cap = u64(*[int(t.shape[0]) for t in out_keys])
        out = u64()
        rc = self._lib.thrs_mg_sort_pairs(self._h, k, v, counts, ok, ov, cap, out)
        self.last_needed = [int(x) for x in out]
        if rc:
            raise RuntimeError(f"thrs_mg_sort_pairs rc={rc}: {self._lib.thrs_mg_last_error().decode()}")
        return [int(x) for x in out]

    def close(self):
        if self._h:
            self._lib.thrs_mg_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

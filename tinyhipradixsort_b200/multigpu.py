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


class CudaExchangeOps:
    """Device work of one rank through the exchange layer of the C ABI (thrs_xchg_*, csrc/capi_xchg.cu):
    the top-digit pass IS the sort's last radix pass, every (digit, source) chunk travels to its owner
    with the copy engines, and the owner sorts the buckets it received with a segmented sort on the low
    bits while later groups are still in flight.  torch.distributed is used ONCE per buffer size, to
    pass the CUDA-IPC handles of the windows around; a sort itself involves no collective and no
    Python-side arithmetic (histogram exchange, splitters and copy lists are C++)."""

    def __init__(self, key="u32", value_bytes=4, descending=False, groups=3):
        import ctypes

        import torch

        from . import _lib

        self.torch, self.ct, self._lib = torch, ctypes, _lib.load()
        self.key, self.vbytes, self.groups = key, value_bytes, groups
        cfg = _lib.thrs_config()
        self._lib.thrs_config_default(ctypes.byref(cfg))
        cfg.key_type = {"u32": 0, "u64": 1, "f32": 2, "f64": 3}[key]
        cfg.value_type = {4: 0, 8: 1}[value_bytes]
        cfg.sort_order = 1 if descending else 0
        self._cfg = cfg
        self._x = None
        self._opened = []
        self.max_local = self.capacity = 0

    def _err(self, what, rc):
        return RuntimeError(f"{what} rc={rc}: {self._lib.thrs_xchg_last_error().decode()}")

    def close(self):
        from .radixsort import ipc_close

        if self._x:
            self.torch.cuda.synchronize()
            for p in self._opened:
                ipc_close(p)
            self._opened = []
            self._lib.thrs_xchg_destroy(self._x)
            self._x = None

    def reserve(self, dist, group, max_local, capacity):
        """COLLECTIVE: (re)build the context for slices of up to max_local items and shares of up to
        capacity items (identical numbers on every rank) and pass the window handles around."""
        from .radixsort import ipc_export, ipc_open

        ct = self.ct
        self.close()
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        x = ct.c_void_p()
        rc = self._lib.thrs_xchg_create(ct.byref(self._cfg), rank, world, int(max_local), int(capacity), self.groups, ct.byref(x))
        if rc:
            raise self._err("thrs_xchg_create", rc)
        self._x = x
        base, nbytes = ct.c_void_p(), ct.c_uint64()
        self._lib.thrs_xchg_window(x, ct.byref(base), ct.byref(nbytes))
        handles = [None] * world
        dist.all_gather_object(handles, ipc_export(base.value), group=group)
        bases = []
        for r, h in enumerate(handles):
            if r == rank:
                bases.append(base.value)
            else:
                p = ipc_open(h)
                self._opened.append(p)
                bases.append(p)
        arr = (ct.c_void_p * world)(*bases)
        rc = self._lib.thrs_xchg_set_peers(x, arr)
        if rc:
            raise self._err("thrs_xchg_set_peers", rc)
        self.max_local, self.capacity = int(max_local), int(capacity)
        dist.barrier(group=group)

    def begin(self, keys, vals, n, stream):
        rc = self._lib.thrs_xchg_begin(self._x, keys.data_ptr(), vals.data_ptr(), int(n), stream)
        if rc:
            raise self._err("thrs_xchg_begin", rc)

    def finish(self, stream):
        """-> (keys_ptr, values_ptr, count, needed_capacity, ok)"""
        ct = self.ct
        ok_, ov_, cnt, need = ct.c_void_p(), ct.c_void_p(), ct.c_uint64(), ct.c_uint64()
        rc = self._lib.thrs_xchg_finish(self._x, stream, ct.byref(ok_), ct.byref(ov_), ct.byref(cnt), ct.byref(need))
        if rc and need.value <= self.capacity:
            raise self._err("thrs_xchg_finish", rc)
        return ok_.value, ov_.value, int(cnt.value), int(need.value), rc == 0

    def stats(self):
        ct = self.ct
        recv, sent, launches, copies = ct.c_uint64(), ct.c_uint64(), ct.c_uint32(), ct.c_uint32()
        self._lib.thrs_xchg_stats(self._x, ct.byref(recv), ct.byref(sent), ct.byref(launches), ct.byref(copies))
        return {"received_items": int(recv.value), "sent_remote_bytes": int(sent.value), "kernel_launches": int(launches.value),
                "bucket_copies": int(copies.value)}

    def views(self, kptr, vptr, count, key_dtype, val_dtype):
        torch = self.torch

        class _Raw:
            def __init__(self, ptr, n, typestr):
                self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}

        ts = {torch.int32: "<i4", torch.int64: "<i8"}
        k = torch.as_tensor(_Raw(kptr, max(count, 1), ts[key_dtype]), device="cuda")
        v = torch.as_tensor(_Raw(vptr, max(count, 1), ts[val_dtype]), device="cuda")
        return k[:count], v[:count]


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
        self.exchange = hasattr(local_ops, "begin")  # CudaExchangeOps: the whole sort is C++ (thrs_xchg_*)

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

        if self.exchange:
            return self._exchange_sort(keys, vals, n, mark)

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


    def reserve(self, max_local, capacity=None):
        """COLLECTIVE (exchange path): size the per-rank buffers for slices of up to max_local items and
        shares of up to `capacity` items (default: 6 % above max_local) — same numbers on every rank."""
        capacity = int(max_local * 1.06) + 4096 if capacity is None else int(capacity)
        self.ops.reserve(self.dist, self.group, int(max_local), capacity)

    def _exchange_sort(self, keys, vals, n, mark):
        import torch

        ops = self.ops
        if ops._x is None or n > ops.max_local:
            # first use (or a larger slice): agree on the sizes once; later sorts involve no collective
            sizes = [None] * self.world
            self.dist.all_gather_object(sizes, n, group=self.group)
            self.reserve(max(sizes))
        stream = torch.cuda.current_stream().cuda_stream
        for attempt in range(2):
            ops.begin(keys, vals, n, stream)
            mark("top-digit histogram + partition pass (sort pass 4 of 4)")
            kptr, vptr, cnt, need, ok = ops.finish(stream)
            if ok:
                mark("bucket exchange (copy engines) overlapped with the segmented 3-pass sort")
                rk, rv = ops.views(kptr, vptr, cnt, keys.dtype, vals.dtype)
                return rk, rv, cnt
            # skewed keys: a share did not fit.  Every rank saw the same table and failed alike, before anything
            # was copied: rebuild with the exact need (collective) and run the sort again
            torch.cuda.synchronize()
            self.reserve(ops.max_local, int(need * 1.02) + 4096)
        raise RuntimeError("exchange capacity could not be sized")

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


def _prefer_gpu_numa_node(torch, dev):
    """Best effort, never raises: make this process's NEXT host allocations prefer the NUMA node its GPU hangs off
    (set_mempolicy(MPOL_PREFERRED), which falls back to other nodes when that one is full), so that the pinned
    slices of the end-to-end measurement do not all land on the node the launcher happens to run on.  Returns the
    node, or None where the node is unknown (-1 in sysfs), the platform is not x86-64 Linux, or the call is refused."""
    try:
        import ctypes
        import platform

        if platform.system() != "Linux" or platform.machine() != "x86_64":
            return None
        p = torch.cuda.get_device_properties(dev)
        bus = "%04x:%02x:%02x.0" % (int(p.pci_domain_id), int(p.pci_bus_id), int(p.pci_device_id))
        with open(f"/sys/bus/pci/devices/{bus}/numa_node") as f:
            node = int(f.read().strip())
        if node < 0 or node >= 64:
            return None
        libc = ctypes.CDLL(None, use_errno=True)
        mask = ctypes.c_ulong(1 << node)
        rc = libc.syscall(238, 1, ctypes.byref(mask), 65)  # SYS_set_mempolicy, MPOL_PREFERRED
        return node if rc == 0 else None
    except Exception:  # noqa: BLE001
        return None


def _default_numa_policy():
    try:
        import ctypes
        import platform

        if platform.system() == "Linux" and platform.machine() == "x86_64":
            ctypes.CDLL(None, use_errno=True).syscall(238, 0, None, 0)  # MPOL_DEFAULT
    except Exception:  # noqa: BLE001
        pass


def bench(args, dist, rank, world, dev, g):
    """bench.py at N > 1: the partitioned sort.  Weak scaling by default (`--log2n-per-gpu` pairs per
    GPU); `--log2n-total T` fixes the total instead (strong scaling; T = 32 is config C5)."""
    import json
    import os
    import time

    import torch

    from . import RadixSort

    strong = bool(getattr(args, "log2n_total", 0))
    n_l = (1 << args.log2n_total) // world if strong else 1 << args.log2n_per_gpu
    total = n_l * world
    # Which multi-GPU path (both are products, both verified every step).  Measured on this pool
    # (profiles/r2_exchange_timeline.txt, profiles/r2_exchange_lanes.txt): the exchange path wins at every world size once
    # its copies run on ONE ordered lane from 3 ranks on (2 ranks 5.55 vs 6.24 ms, 4 ranks 6.20 vs 6.99, 8 ranks 7.0 vs 7.7-7.9);
    # THRS_MG_MODE=scatter selects the round-1 path (the partition pass stores into peer memory, then a 4-pass local sort).
    mode = os.environ.get("THRS_MG_MODE", "auto")
    if mode == "auto":
        mode = "exchange"
    keys0, vals0 = g["gen_pairs"](torch, rank * n_l, n_l, dev)
    if mode == "scatter":
        ops = CudaLocalOps("u32", 4)
        ops.reserve(int(n_l * 1.05) + 4096)
    else:
        ops = CudaExchangeOps("u32", 4, groups=int(os.environ.get("THRS_MG_GROUPS", "3" if world <= 2 else "4")))
    ds = DistributedRadixSort(dist, ops, "u32")
    keys, vals = torch.empty_like(keys0), torch.empty_like(vals0)

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
    xstats = ops.stats() if mode != "scatter" else {}
    total_ms = sum(step_ms)
    value = total * args.steps / (total_ms * 1e-3)

    # What bounds the step.  (1) per-pass device time of the dominant kernel, measured live with the sorter's
    # events on a plain local sort of this rank's slice; (2) the segmented 3-pass sort of what this rank
    # received, alone (nothing in flight): finish-stage time minus this = the part of the exchange NOT hidden.
    stream = torch.cuda.current_stream()
    srt = RadixSort([], RadixSort.Config().configureWithKeyPair("u32", 4))
    tmp = torch.empty(srt.getTemporaryBufferBytes(n_l).getTemporaryBufferBytesForSortPairs(), dtype=torch.uint8, device=dev)
    srt.profile(True)
    for _ in range(2):
        keys.copy_(keys0)
        vals.copy_(vals0)
        srt.sortPairs(keys, vals, n_l, tmp, 0, 32, stream)
        prof = srt.readProfile()
    srt.profile(False)
    seg_alone_ms = None
    del tmp
    if mode != "scatter" and cnt and n_l <= (1 << 29):  # (the int64 temporaries of this side measurement do not fit next to 2^31-item slices)
        k = (rk.to(torch.int64) & 0xFFFFFFFF) >> 24
        lo, hi = int(k[0]), int(k[-1])
        edges = torch.searchsorted(k, torch.arange(lo, hi + 2, device=dev, dtype=torch.int64)).tolist()
        del k
        sk, sv = rk.clone(), rv.clone()
        ak, av = torch.empty_like(sk), torch.empty_like(sv)
        sbytes = srt.getSegmentsScratchBytes(cnt, len(edges) - 1)
        stmp = torch.empty(sbytes, dtype=torch.uint8, device=dev)
        best = None
        for _ in range(3):  # unsorted keys inside the same bucket boundaries: the ranking work of the real step
            sk.copy_(keys0[:cnt] if cnt <= n_l else rk)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            srt.sortSegmentsOutOfPlace(sk, sv, ak, av, stmp, sbytes, edges, 0, 24, stream)
            b.record(stream)
            b.synchronize()
            best = a.elapsed_time(b) if best is None else min(best, a.elapsed_time(b))
        del ak, av
        seg_alone_ms = best
        del sk, sv, stmp
    torch.cuda.empty_cache()

    # end to end: pinned host slices in, sorted share out (H2D + sort + D2H inside the timed region)
    e2e_value, e2e_ms, e2e_note = None, None, None
    numa_node = None
    if n_l <= (1 << 29):
        # every rank's pinned slices on the NUMA node of ITS GPU where the host allows it (all ranks are started from
        # one launcher, i.e. from one node's cores: with the default first-touch policy every slice would land there)
        numa_node = _prefer_gpu_numa_node(torch, dev)
        hk0, hv0 = keys0.cpu().pin_memory(), vals0.cpu().pin_memory()
        hk_out = torch.empty(int(n_l * 1.06) + 8192, dtype=hk0.dtype).pin_memory()
        hv_out = torch.empty_like(hk_out).pin_memory()
        _default_numa_policy()
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
        e2e_value, e2e_ms = total * len(e2e) / sum(e2e), 1e3 * sum(e2e) / len(e2e)
    else:
        e2e_note = "skipped: pinned host copies of >= 2^30 pairs per rank (32+ GiB of pinned memory per process)"

    if rank == 0:
        peak, peak_src = g["_peaks"]()
        per_gpu_bytes = 2.0 * n_l * 8
        avg_pass = sum(prof["pass_ms"]) / len(prof["pass_ms"])
        achieved = per_gpu_bytes / (avg_pass * 1e-3) / 1e9
        stage_ms = {k: v / args.steps for k, v in stage_sum.items()}
        if mode == "scatter":
            send_bytes = n_l * 8 * (world - 1) / world
            x_ms = stage_ms.get("partition+exchange (peer stores)", 0.0)
            exchange = {"how": "round-1 path: partition pass storing into peer memory (P2P stores, CUDA IPC)",
                        "bytes_sent_per_gpu": send_bytes, "ms": x_ms,
                        "GBps_per_direction_per_gpu": (send_bytes / (x_ms * 1e-3) / 1e9) if x_ms else None, "nvlink_peak_GBps": 900.0}
            launches_per_step = 1 + 1 + 6
            how = "MSD splitters + partition pass storing into peer memory over NVLink + local 4-pass onesweep"
        else:
            fin = [v for k, v in stage_ms.items() if k.startswith("bucket exchange")][0]
            send_bytes = xstats["sent_remote_bytes"]
            exposed = None if seg_alone_ms is None else max(fin - seg_alone_ms, 0.0)
            exchange = {"how": "copy-engine copies of (digit, source) chunks into the owners' windows (CUDA IPC), digit-major / source-minor, "
                               f"{ops.groups} groups per rank, 4-byte flag per group, cuStreamWaitValue32 on the owner's stream",
                        "bytes_sent_per_gpu_rank0": send_bytes, "bucket_copies_per_step_rank0": xstats["bucket_copies"],
                        "finish_stage_ms": fin, "segmented_sort_alone_ms": seg_alone_ms,
                        "exchange_not_hidden_ms": exposed,
                        "overlap_fraction_of_copy_time": None if exposed is None or not send_bytes else
                        max(0.0, 1.0 - exposed / (send_bytes / 750e9 * 1e3)),
                        "GBps_per_direction_per_gpu_lower_bound": send_bytes / (fin * 1e-3) / 1e9 if fin else None,
                        "nvlink_peak_GBps": 900.0,
                        "limiter": "segmented sort (SM shared-memory pipe)" if exposed is not None and exposed < 0.25 * fin else "NVLink copies"}
            launches_per_step = xstats["kernel_launches"]
            how = ("top-digit histogram + ONE stable top-digit pass (the sort's pass 4) + copy-engine bucket exchange over NVLink "
                   "overlapped with a segmented 3-pass onesweep on bits [0,24) of the received buckets")
        line = {
            "metric": g["METRIC"], "value": value, "unit": g["UNIT"], "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "strong" if strong else "weak",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": f"C5{'' if strong and args.log2n_total == 32 else ' shape'}: {world} x {n_l} uniform-random u32 key + u32 value pairs "
                                   f"(2^{total.bit_length() - 1} total), {how}, stable, bits [0,32)",
                       "n_total": total, "n_per_gpu": n_l,
                       "parallelism": f"partitioned sort over {world} ranks, one process per GPU (torch.distributed/NCCL only passes IPC handles and the timing all_reduce)",
                       "l2": "per-GPU inputs (>= 2 GiB) are larger than L2; input restored outside the timed events",
                       "stage_ms_rank0": stage_ms, "exchange": exchange, "mode": mode},
            "verified": verified,
            "roofline": {"bound": "hbm", "kernel": "onesweep_kernel (one 8-bit pass over this rank's slice, rank 0)",
                         "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": g["_traffic"](), "peak_source": peak_src, "bytes_per_launch": per_gpu_bytes,
                         "avg_launch_ms": avg_pass},
            "e2e": {"value": e2e_value, "unit": g["UNIT"], "h2d_bytes_per_step": total * 8, "d2h_bytes_per_step": total * 8,
                    "ms_per_step": e2e_ms, "api": "DistributedRadixSort.sort_pairs (pinned host slices)", "note": e2e_note,
                    "host_memory_rank0": ("pinned, NUMA node %d preferred (the node of this rank's GPU)" % numa_node) if numa_node is not None
                    else "pinned, default NUMA policy (GPU's node unknown or set_mempolicy refused)"},
            "gpu_launches": launches_per_step * args.steps * world,
            "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    dist.barrier()
    if hasattr(ops, "close"):
        ops.close()
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

"""Host-side mirror of the reference's ``thrs::`` interface (tinyhipradixsort.hpp:638-948)
over the C ABI.  Same names, argument meaning and error behaviour, so the parity tests read
like the reference's unittest.cpp:

    config = RadixSort.Config(); config.configureWithKeyPair("u32", "u32")
    sorter = RadixSort([], config)
    tmp = Buffer(sorter.getTemporaryBufferBytes(n).getTemporaryBufferBytesForSortPairs())
    sorter.sortPairs(keys, values, n, tmp, 0, 32, stream)

Buffers may be ``Buffer`` objects, torch CUDA tensors or raw device addresses (int).
PyTorch is only plumbing here (device memory, streams); all work happens in libthrs_b200.so.
Where the reference traps (THRS_ASSERT -> __debugbreak, hpp:15) this raises ``ThrsError``.
"""
import ctypes
import enum

from . import _lib


class ThrsError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"thrs error {code}: {message}")
        self.code = code


class KeyType(enum.IntEnum):  # hpp:638-644
    U32 = 0
    U64 = 1
    F32 = 2
    F64 = 3


class ValueType(enum.IntEnum):  # hpp:645-650
    U32 = 0
    U64 = 1
    U128 = 2


class SortOrder(enum.IntEnum):  # hpp:679-683
    Ascending = 0
    Descending = 1


def bytesOf(t):  # hpp:651-678
    if isinstance(t, KeyType):
        return 4 if t in (KeyType.U32, KeyType.F32) else 8
    if isinstance(t, ValueType):
        return {ValueType.U32: 4, ValueType.U64: 8, ValueType.U128: 16}[t]
    raise ThrsError(1, "bytesOf: not a KeyType/ValueType")


_KEY_NAMES = {
    "u32": KeyType.U32, "uint32": KeyType.U32, "i32": KeyType.U32, "int32": KeyType.U32,
    "u64": KeyType.U64, "uint64": KeyType.U64, "i64": KeyType.U64, "int64": KeyType.U64,
    "f32": KeyType.F32, "float": KeyType.F32, "float32": KeyType.F32,
    "f64": KeyType.F64, "double": KeyType.F64, "float64": KeyType.F64,
}
_VALUE_BYTES = {4: ValueType.U32, 8: ValueType.U64, 16: ValueType.U128}
_VALUE_NAMES = {"u32": 4, "uint32": 4, "f32": 4, "u64": 8, "uint64": 8, "f64": 8, "u128": 16, "uint4": 16}


def _check(rc):
    if rc != 0:
        raise ThrsError(rc, _lib.load().thrs_last_error().decode())


def _ptr(x):
    """Device address of a Buffer / torch tensor / int."""
    if x is None:
        return None
    if isinstance(x, Buffer):
        return x.data()
    if isinstance(x, int):
        return x
    if hasattr(x, "data_ptr"):
        if not x.is_cuda:
            raise ThrsError(1, "expected a CUDA tensor: this library has no CPU path")
        if not x.is_contiguous():
            raise ThrsError(1, "expected a contiguous tensor")
        return x.data_ptr()
    raise ThrsError(1, f"cannot take a device pointer from {type(x)!r}")


def _stream(stream):
    """cudaStream_t handle: None -> torch's current stream (legacy default if torch has no CUDA)."""
    if stream is None:
        try:
            import torch

            if torch.cuda.is_available():
                return torch.cuda.current_stream().cuda_stream
        except ImportError:
            pass
        return None
    if isinstance(stream, int):
        return stream
    return stream.cuda_stream


def ipc_export(device_ptr):
    """64-byte CUDA IPC handle of a cudaMalloc'ed allocation (thrs_ipc_export)."""
    h = ctypes.create_string_buffer(64)
    _check(_lib.load().thrs_ipc_export(_ptr(device_ptr), h))
    return h.raw


def ipc_open(handle):
    """Map another process's allocation into this one (thrs_ipc_open); returns the device address."""
    p = ctypes.c_void_p(0)
    _check(_lib.load().thrs_ipc_open(ctypes.create_string_buffer(bytes(handle), 64), ctypes.byref(p)))
    return int(p.value)


def ipc_close(device_ptr):
    _check(_lib.load().thrs_ipc_close(int(device_ptr)))


class Buffer:
    """thrs::Buffer (hpp:501-528): RAII device allocation of max(bytes, 1), non-copyable."""

    def __init__(self, nbytes):
        self._bytes = max(int(nbytes), 1)
        p = ctypes.c_void_p()
        _check(_lib.load().thrs_buffer_alloc(self._bytes, ctypes.byref(p)))
        self._ptr = p.value

    def bytes(self):
        return self._bytes

    def data(self):
        return self._ptr

    def free(self):
        if getattr(self, "_ptr", None):
            _lib.load().thrs_buffer_free(self._ptr)
            self._ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class TemporaryBufferDef:
    """thrs::RadixSort::TemporaryBufferDef (hpp:806-832)."""

    def __init__(self, pSumBuffer, keyOutBuffer, valueOutBuffer):
        self.pSumBuffer = pSumBuffer
        self.keyOutBuffer = keyOutBuffer
        self.valueOutBuffer = valueOutBuffer

    def getTemporaryBufferBytesForSortKeys(self):
        return self.pSumBuffer + self.keyOutBuffer

    def getTemporaryBufferBytesForSortPairs(self):
        return self.pSumBuffer + self.keyOutBuffer + self.valueOutBuffer

    def getPSumBuffer(self, p):
        return p

    def getOutputKeyBuffer(self, p):
        return p + self.pSumBuffer

    def getOutputValueBuffer(self, p):
        return p + self.pSumBuffer + self.keyOutBuffer


class RadixSort:
    """thrs::RadixSort (hpp:694-948)."""

    class Config:  # hpp:697-749
        def __init__(self, keyIs16byteAligned=True, keyType=KeyType.U32, valueType=ValueType.U32,
                     sortOrder=SortOrder.Ascending):
            self.keyIs16byteAligned = keyIs16byteAligned
            self.keyType = keyType
            self.valueType = valueType
            self.sortOrder = sortOrder

        def configureWithKey(self, key):
            """``key``: a name ('u32','f32','u64','f64') or a KeyType (template <class KEY>)."""
            self.keyType = key if isinstance(key, KeyType) else _KEY_NAMES[str(key)]
            return self

        def configureWithKeyPair(self, key, value):
            """``value``: a name, a byte size (4/8/16) or a ValueType."""
            self.configureWithKey(key)
            if isinstance(value, ValueType):
                self.valueType = value
            else:
                nbytes = value if isinstance(value, int) else _VALUE_NAMES[str(value)]
                if nbytes not in _VALUE_BYTES:
                    raise ThrsError(1, "value size must be 4, 8 or 16 bytes")  # static_assert hpp:733
                self.valueType = _VALUE_BYTES[nbytes]
            return self

    def __init__(self, extraArgs=(), config=None):
        # extraArgs were NVRTC flags in the reference (hpp:751); kernels are precompiled
        # sm_100a here, so they are accepted and ignored.
        self._lib = _lib.load()
        self.m_config = config if config is not None else RadixSort.Config()
        c = _lib.thrs_config(
            int(bool(self.m_config.keyIs16byteAligned)), int(self.m_config.keyType),
            int(self.m_config.valueType), int(self.m_config.sortOrder))
        h = ctypes.c_void_p()
        _check(self._lib.thrs_create(ctypes.byref(c), ctypes.byref(h)))
        self._h = h

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            self._lib.thrs_destroy(h)
            self._h = None

    def getTemporaryBufferBytes(self, numberOfMaxInputs):
        d = _lib.thrs_temp_def()
        _check(self._lib.thrs_temp_bytes(self._h, int(numberOfMaxInputs), ctypes.byref(d)))
        return TemporaryBufferDef(d.pSumBuffer, d.keyOutBuffer, d.valueOutBuffer)

    def sortKeys(self, inputKeyBuffer, numberOfInputs, temporaryBuffer, startBits, endBits, stream=None):
        _check(self._lib.thrs_sort_keys(self._h, _ptr(inputKeyBuffer), int(numberOfInputs),
                                        _ptr(temporaryBuffer), int(startBits), int(endBits), _stream(stream)))

    def sortPairs(self, inputKeyBuffer, inputValueBuffer, numberOfInputs, temporaryBuffer, startBits, endBits,
                  stream=None):
        _check(self._lib.thrs_sort_pairs(self._h, _ptr(inputKeyBuffer), _ptr(inputValueBuffer),
                                         int(numberOfInputs), _ptr(temporaryBuffer), int(startBits),
                                         int(endBits), _stream(stream)))

    def digitHistogram(self, keyBuffer, numberOfInputs, startBit, countsOut, temporaryBuffer, stream=None):
        """256-bin histogram (uint32, device) of digit (getKeyBits(key) >> startBit) & 0xFF in sort order."""
        _check(self._lib.thrs_digit_histogram(self._h, _ptr(keyBuffer), int(numberOfInputs), int(startBit),
                                              _ptr(countsOut), _ptr(temporaryBuffer), _stream(stream)))

    # -- not in the reference: many independent sorts in one set of launches (thrs_sort_segments_*)
    def getSegmentsTemporaryBufferBytes(self, totalInputs, numberOfSegments, withValues=True):
        return int(self._lib.thrs_segments_temp_bytes(self._h, int(totalInputs), int(numberOfSegments), int(bool(withValues))))

    def sortSegments(self, keyBuffer, valueBuffer, segmentBegin, temporaryBuffer, startBits, endBits, stream=None):
        """Sort every segment [segmentBegin[i], segmentBegin[i+1]) of the arrays on its own (stable,
        in place).  segmentBegin: ascending host offsets, len = segments + 1; valueBuffer None = keys only."""
        seg = (ctypes.c_uint64 * len(segmentBegin))(*[int(x) for x in segmentBegin])
        nseg = len(segmentBegin) - 1
        if valueBuffer is None:
            _check(self._lib.thrs_sort_segments_keys(self._h, _ptr(keyBuffer), seg, nseg, _ptr(temporaryBuffer),
                                                     int(startBits), int(endBits), _stream(stream)))
        else:
            _check(self._lib.thrs_sort_segments_pairs(self._h, _ptr(keyBuffer), _ptr(valueBuffer), seg, nseg,
                                                      _ptr(temporaryBuffer), int(startBits), int(endBits), _stream(stream)))

    def getSegmentsScratchBytes(self, totalInputs, numberOfSegments):
        return int(self._lib.thrs_segments_scratch_bytes(self._h, int(totalInputs), int(numberOfSegments)))

    def sortSegmentsOutOfPlace(self, keyBuffer, valueBuffer, altKeyBuffer, altValueBuffer, scratchBuffer, scratchBytes,
                               segmentBegin, startBits, endBits, stream=None):
        """thrs_sort_segments_ex: ping-pongs between the arrays and their alternates, no copy-back;
        returns True if the sorted segments lie in the alternates."""
        seg = (ctypes.c_uint64 * len(segmentBegin))(*[int(x) for x in segmentBegin])
        in_alt = ctypes.c_int(0)
        _check(self._lib.thrs_sort_segments_ex(self._h, _ptr(keyBuffer), None if valueBuffer is None else _ptr(valueBuffer),
                                               _ptr(altKeyBuffer), None if altValueBuffer is None else _ptr(altValueBuffer),
                                               _ptr(scratchBuffer), int(scratchBytes), seg, len(segmentBegin) - 1,
                                               int(startBits), int(endBits), _stream(stream), ctypes.byref(in_alt)))
        return bool(in_alt.value)

    def sortBucketsOutOfPlace(self, keyBuffer, valueBuffer, altKeyBuffer, altValueBuffer, scratchBuffer, scratchBytes,
                              chunkBegin, chunkBucket, bucketBegin, startBits, endBits, stream=None):
        """thrs_sort_buckets_ex: pass 0 gathers the chunks (chunkBegin, chunkBucket) into the buckets
        (bucketBegin); returns True if the sorted buckets lie in the alternates."""
        cb = (ctypes.c_uint64 * len(chunkBegin))(*[int(x) for x in chunkBegin])
        cm = (ctypes.c_uint32 * len(chunkBucket))(*[int(x) for x in chunkBucket])
        bb = (ctypes.c_uint64 * len(bucketBegin))(*[int(x) for x in bucketBegin])
        in_alt = ctypes.c_int(0)
        _check(self._lib.thrs_sort_buckets_ex(self._h, _ptr(keyBuffer), None if valueBuffer is None else _ptr(valueBuffer),
                                              _ptr(altKeyBuffer), None if altValueBuffer is None else _ptr(altValueBuffer),
                                              _ptr(scratchBuffer), int(scratchBytes), cb, cm, len(chunkBucket), bb,
                                              len(bucketBegin) - 1, int(startBits), int(endBits), _stream(stream),
                                              ctypes.byref(in_alt)))
        return bool(in_alt.value)

    # -- not in the reference: what unittest.cpp:399-414 does around a sort, in one C-ABI call
    def sortHost(self, host_keys, host_values, startBits, endBits):
        """Sort numpy arrays / CPU tensors in place (H2D, sort, D2H inside the call)."""
        n = int(host_keys.shape[0])
        kp = host_keys.ctypes.data if hasattr(host_keys, "ctypes") else host_keys.data_ptr()
        vp = None
        if host_values is not None:
            vp = host_values.ctypes.data if hasattr(host_values, "ctypes") else host_values.data_ptr()
        _check(self._lib.thrs_sort_host(self._h, kp, vp, n, int(startBits), int(endBits)))

    HOST_SLOTS = 4  # THRS_HOST_SLOTS

    def sortHostSubmit(self, slot, host_keys, host_values, startBits, endBits):
        """Streaming form of sortHost: enqueue upload, sort and download of one batch in `slot` and return at once
        (thrs_sort_host_submit).  The arrays (pinned CPU tensors for real overlap) are sorted in place once
        sortHostWait(slot) has returned; up to HOST_SLOTS batches may be in flight."""
        n = int(host_keys.shape[0])
        kp = host_keys.ctypes.data if hasattr(host_keys, "ctypes") else host_keys.data_ptr()
        vp = None
        if host_values is not None:
            vp = host_values.ctypes.data if hasattr(host_values, "ctypes") else host_values.data_ptr()
        _check(self._lib.thrs_sort_host_submit(self._h, int(slot), kp, vp, n, int(startBits), int(endBits)))

    def sortHostWait(self, slot):
        """Block until the batch submitted in `slot` is back in its host arrays (thrs_sort_host_wait)."""
        _check(self._lib.thrs_sort_host_wait(self._h, int(slot)))

    def profile(self, on):
        """Record CUDA events between the launches of each sort (thrs_profile_enable)."""
        _check(self._lib.thrs_profile_enable(self._h, int(bool(on))))

    def readProfile(self):
        """Per-stage milliseconds of the last profiled sort (blocks until it finished)."""
        p = _lib.thrs_profile()
        _check(self._lib.thrs_profile_read(self._h, ctypes.byref(p)))
        return {"memset_ms": p.memset_ms, "histogram_ms": p.histogram_ms,
                "pass_ms": [p.pass_ms[i] for i in range(p.passes)], "copyback_ms": p.copyback_ms,
                "total_ms": p.total_ms, "passes": p.passes, "parts": p.parts}

    def scatterPairs(self, inputKeyBuffer, inputValueBuffer, numberOfInputs, temporaryBuffer, startBit, binFirstDigit,
                     keyDst, valDst, firstIndex, stream=None):
        """One stable pass that splits the input into len(binFirstDigit) <= 8 destination bins by the
        8-bit digit at startBit (thrs_scatter_pairs): bin b's items go, in input order, to device
        address keyDst[b] / valDst[b] (own or peer memory) from element firstIndex[b] on."""
        import numpy as np

        bf = np.ascontiguousarray(binFirstDigit, dtype=np.uint32)
        kd = np.ascontiguousarray(keyDst, dtype=np.uint64)
        vd = np.ascontiguousarray(valDst, dtype=np.uint64)
        fi = np.ascontiguousarray(firstIndex, dtype=np.uint32)
        if not (bf.ndim == 1 and bf.shape == kd.shape == vd.shape == fi.shape):
            raise ThrsError(1, "scatterPairs: the four tables have one entry per bin")
        _check(self._lib.thrs_scatter_pairs(self._h, _ptr(inputKeyBuffer), _ptr(inputValueBuffer), int(numberOfInputs),
                                            _ptr(temporaryBuffer), int(startBit), int(bf.shape[0]), bf.ctypes.data,
                                            kd.ctypes.data, vd.ctypes.data, fi.ctypes.data, _stream(stream)))

    SMALL_SORT_MAX = 4096

    def setSmallPath(self, enabled):
        """Enable/disable the one-block path for n <= SMALL_SORT_MAX (thrs_set_small_path)."""
        _check(self._lib.thrs_set_small_path(self._h, int(bool(enabled))))

    def setClusterPath(self, enabled):
        """Enable/disable the one-cluster path for SMALL_SORT_MAX < n <= 8 x SMALL_SORT_MAX (thrs_set_cluster_path)."""
        _check(self._lib.thrs_set_cluster_path(self._h, int(bool(enabled))))

    def clusterSortMax(self):
        """Largest n sorted by one cluster launch on this device (thrs_cluster_sort_max) — needs a device."""
        m = ctypes.c_uint32(0)
        _check(self._lib.thrs_cluster_sort_max(self._h, ctypes.byref(m)))
        return int(m.value)

    MID_SORT_MAX = 1 << 20

    def setMidMax(self, maxItems):
        """Largest n sorted with the quarter-size tile shape (thrs_set_mid_max); 0 = full-size tiles only."""
        _check(self._lib.thrs_set_mid_max(self._h, int(maxItems)))

    RANKING_AUTO, RANKING_MATCH_MASK, RANKING_COUNTER_RETURN = 0, 1, 2

    def setRanking(self, ranking):
        """Choose the onesweep ranking (thrs_set_ranking); AUTO lets the device probe decide."""
        _check(self._lib.thrs_set_ranking(self._h, int(ranking)))

    def getRanking(self):
        """(ranking in use, probe passed, groups the probe checked) — needs a device."""
        r, ok, g = ctypes.c_int(0), ctypes.c_int(0), ctypes.c_uint64(0)
        _check(self._lib.thrs_get_ranking(self._h, ctypes.byref(r), ctypes.byref(ok), ctypes.byref(g)))
        return r.value, bool(ok.value), int(g.value)

    def getRankingCheck(self):
        """(production check passed, items sorted under both rankings) — thrs_get_ranking_check."""
        ok, items = ctypes.c_int(0), ctypes.c_uint64(0)
        _check(self._lib.thrs_get_ranking_check(self._h, ctypes.byref(ok), ctypes.byref(items)))
        return bool(ok.value), int(items.value)

    def tileItems(self):
        return int(self._lib.thrs_tile_items(self._h))

    def lastLaunchCount(self):
        return int(self._lib.thrs_last_launch_count())

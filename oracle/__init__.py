"""oracle — ctypes access to the CPU oracle (oracle/liboracle.so) and to the reference-derived
checkers under oracle/_ref/.  TEST INFRASTRUCTURE ONLY: importable from tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, never from the
product package (tinyhipradixsort_b200 has no import of this module).

Parity is pinned (see oracle.cpp header): the restatement is checked against the reference's
own fpKey.hpp build (_ref/libref_fpkey.so), the known-answer vectors in tests/golden/, and on
the GPU box against the reference's own kernels (_ref/kernel_*.cubin via _ref/libref_gpu.so).
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(_HERE, "liboracle.so")
REF_DIR = os.path.join(_HERE, "_ref")

KEY_TYPES = {"u32": 0, "u64": 1, "f32": 2, "f64": 3}
KEY_NP = {"u32": np.uint32, "u64": np.uint64, "f32": np.uint32, "f64": np.uint64}  # raw bit patterns
KEY_BYTES = {"u32": 4, "u64": 8, "f32": 4, "f64": 8}

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(ORACLE_SO):
            raise ImportError(f"{ORACLE_SO} missing: run `make -C oracle`")
        L = ctypes.CDLL(ORACLE_SO)
        L.oracle_key_bits_u32.restype = ctypes.c_uint32
        L.oracle_key_bits_u32.argtypes = [ctypes.c_uint32]
        L.oracle_key_bits_u64.restype = ctypes.c_uint64
        L.oracle_key_bits_u64.argtypes = [ctypes.c_uint64]
        L.oracle_key_bits_f32.restype = ctypes.c_uint32
        L.oracle_key_bits_f32.argtypes = [ctypes.c_uint32]
        L.oracle_key_bits_f64.restype = ctypes.c_uint64
        L.oracle_key_bits_f64.argtypes = [ctypes.c_uint64]
        L.oracle_field.restype = ctypes.c_uint64
        L.oracle_field.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_uint64, ctypes.c_int, ctypes.c_int]
        for name in ("oracle_sort", "oracle_sort_lsd"):
            f = getattr(L, name)
            f.restype = ctypes.c_int
            f.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                          ctypes.c_uint64, ctypes.c_int, ctypes.c_int]
        L.oracle_splitmix64_next.restype = ctypes.c_uint64
        L.oracle_splitmix64_next.argtypes = [ctypes.POINTER(ctypes.c_uint64)]
        L.oracle_randomize.restype = None
        L.oracle_randomize.argtypes = [ctypes.POINTER(ctypes.c_uint64), ctypes.c_int, ctypes.c_int,
                                       ctypes.c_void_p, ctypes.c_uint64]
        L.oracle_time_stable_sort_pairs_u32.restype = ctypes.c_double
        L.oracle_time_stable_sort_pairs_u32.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64,
                                                        ctypes.c_int]
        L.oracle_max_threads.restype = ctypes.c_int
        L.oracle_check_counter_pairs_u32.restype = ctypes.c_int64
        L.oracle_check_counter_pairs_u32.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64,
                                                     ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                                     ctypes.c_int]
        _lib = L
    return _lib


class SplitMix64:
    """splitmix64 of unittest.cpp:24-35 (state lives here, stepping happens in liboracle.so)."""

    def __init__(self, x=0):
        self._s = ctypes.c_uint64(x)

    def next(self):
        return int(lib().oracle_splitmix64_next(ctypes.byref(self._s)))

    def randomize(self, key, n, mode=0):
        """randomizeValues<T> (unittest.cpp:96-116) -> numpy array of raw bit patterns.
        mode 0: floats masked finite like the reference; mode 1: raw bits (Inf/NaN occur)."""
        out = np.empty(n, dtype=KEY_NP[key])
        lib().oracle_randomize(ctypes.byref(self._s), KEY_TYPES[key], mode, out.ctypes.data, n)
        return out


def splitmix64_at(i):
    """i-th output (0-based) of splitmix64 seeded with x=0, vectorised (numpy uint64)."""
    with np.errstate(over="ignore"):
        z = (np.asarray(i, dtype=np.uint64) + np.uint64(1)) * np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def sort(key, keys, values=None, start=0, end=None, descending=False, lsd=False):
    """Stable sort on bits [start,end) of getKeyBits; returns new (keys, values) numpy arrays.
    ``keys``: raw bit patterns (uint32/uint64).  ``values``: any array whose rows are 4/8/16 bytes."""
    end = KEY_BYTES[key] * 8 if end is None else end
    k = np.ascontiguousarray(keys, dtype=KEY_NP[key]).copy()
    v = None
    vb = 0
    vp = None
    if values is not None:
        v = np.ascontiguousarray(values).copy()
        vb = v.dtype.itemsize * (int(np.prod(v.shape[1:])) if v.ndim > 1 else 1)
        vp = v.ctypes.data
    f = lib().oracle_sort_lsd if lsd else lib().oracle_sort
    rc = f(KEY_TYPES[key], int(descending), k.ctypes.data, vp, vb, k.shape[0], start, end)
    if rc != 0:
        raise ValueError(f"oracle_sort rc={rc}")
    return k, v


def key_bits(key, raw):
    L = lib()
    return {"u32": L.oracle_key_bits_u32, "u64": L.oracle_key_bits_u64,
            "f32": L.oracle_key_bits_f32, "f64": L.oracle_key_bits_f64}[key](int(raw))


def field(key, raw, start, end, descending=False):
    return int(lib().oracle_field(KEY_TYPES[key], int(descending), int(raw), start, end))


def time_stable_sort_pairs_u32(keys, values, threads=1):
    """Seconds for the reference's CPU check path (stableSortPairs, unittest.cpp:358-377)."""
    return float(lib().oracle_time_stable_sort_pairs_u32(keys.ctypes.data, values.ctypes.data, keys.shape[0],
                                                         threads))


def max_threads():
    return int(lib().oracle_max_threads())


def check_counter_pairs_u32(keys, values, key="u32", descending=False, mode=1, start=0, end=32):
    """-1 if (keys, values) is THE stable sort of key[i]=(u32)splitmix64_at(i), value[i]=i."""
    return int(lib().oracle_check_counter_pairs_u32(keys.ctypes.data, values.ctypes.data, keys.shape[0],
                                                    KEY_TYPES[key], int(descending), mode, start, end))


# ---- reference-derived checkers (oracle/_ref) --------------------------------------------
_ref_fpkey = None


def ref_fpkey():
    """The reference's own fpKey.hpp build, or None when oracle/_ref was not built."""
    global _ref_fpkey
    p = os.path.join(REF_DIR, "libref_fpkey.so")
    if _ref_fpkey is None and os.path.exists(p):
        L = ctypes.CDLL(p)
        for n, t in (("u32", ctypes.c_uint32), ("u64", ctypes.c_uint64), ("f32", ctypes.c_uint32),
                     ("f64", ctypes.c_uint64)):
            f = getattr(L, "ref_key_bits_" + n)
            f.restype = t
            f.argtypes = [t]
        _ref_fpkey = L
    return _ref_fpkey


class RefGpuSorter:
    """The reference's kernel.cu (compiled unmodified into _ref/kernel_*.cubin) driven with the
    launch sequence of tinyhipradixsort.hpp:864-943 by oracle/ref_gpu.cu.  GPU box only."""

    _so = None

    def __init__(self, key, value_bytes=4, descending=False):
        so = os.path.join(REF_DIR, "libref_gpu.so")
        cubin = os.path.join(REF_DIR, f"kernel_{key}_v{value_bytes}_{'desc' if descending else 'asc'}.cubin")
        if not (os.path.exists(so) and os.path.exists(cubin)):
            raise FileNotFoundError(cubin)
        if RefGpuSorter._so is None:
            L = ctypes.CDLL(so)
            L.ref_gpu_load.restype = ctypes.c_void_p
            L.ref_gpu_load.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_int]
            L.ref_gpu_unload.argtypes = [ctypes.c_void_p]
            L.ref_gpu_temp_bytes.restype = ctypes.c_uint64
            L.ref_gpu_temp_bytes.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int]
            L.ref_gpu_sort.restype = ctypes.c_int
            L.ref_gpu_sort.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32,
                                       ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
            RefGpuSorter._so = L
        self._h = RefGpuSorter._so.ref_gpu_load(cubin.encode(), KEY_BYTES[key], value_bytes)
        if not self._h:
            raise RuntimeError("ref_gpu_load failed for " + cubin)

    def temp_bytes(self, n, pairs):
        return int(RefGpuSorter._so.ref_gpu_temp_bytes(self._h, n, int(pairs)))

    def sort(self, keys_ptr, values_ptr, n, temp_ptr, start, end, stream=None):
        rc = RefGpuSorter._so.ref_gpu_sort(self._h, keys_ptr, values_ptr, n, temp_ptr, start, end, stream)
        if rc:
            raise RuntimeError(f"ref_gpu_sort rc={rc}")

    @staticmethod
    def available():
        return os.path.exists(os.path.join(REF_DIR, "libref_gpu.so"))

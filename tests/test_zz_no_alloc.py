"""thrs_sort_keys_noalloc (include/thrs_b200.h): the keys sort whose scratch is a static __device__ array
inside the library, for callers that cannot allocate.  No counterpart in the reference,
so the check is against the existing entry point, thrs_sort_keys, on the
same inputs — bit for bit, like every integer test here — and against the oracle.  Called straight
through the C ABI (ctypes), two calls in a row per case: the second reuses the array the first one left
dirty.  Collected last on purpose (pytest -x must not hide the other tests behind this file)."""
import ctypes

import numpy as np
import pytest
import torch

import oracle
from tinyhipradixsort_b200 import _lib

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 8 << 20  # THRS_NOALLOC_SCRATCH_BYTES
KEY_TYPE = {"u32": 0, "u64": 1, "f32": 2, "f64": 3}
KEY_NP = {"u32": np.uint32, "u64": np.uint64, "f32": np.uint32, "f64": np.uint64}


def _create(lib, key, descending=False):
    cfg = _lib.thrs_config()
    lib.thrs_config_default(ctypes.byref(cfg))
    cfg.key_type = KEY_TYPE[key]
    cfg.sort_order = 1 if descending else 0
    h = ctypes.c_void_p()
    assert lib.thrs_create(ctypes.byref(cfg), ctypes.byref(h)) == 0, lib.thrs_last_error()
    return h


def _device(a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int32 if a.dtype.itemsize == 4 else np.int64)).cuda()


def _sort_existing(lib, h, keys, start, end):
    """thrs_sort_keys with a caller-allocated temp (the entry point the new one is compared with)."""
    n = keys.shape[0]
    d = _lib.thrs_temp_def()
    assert lib.thrs_temp_bytes(h, n, ctypes.byref(d)) == 0
    dk = _device(keys)
    tmp = torch.empty(max(int(d.pSumBuffer + d.keyOutBuffer), 16), dtype=torch.uint8, device="cuda")
    assert lib.thrs_sort_keys(h, dk.data_ptr(), n, tmp.data_ptr(), start, end, None) == 0, lib.thrs_last_error()
    torch.cuda.synchronize()
    return dk.cpu().numpy().view(keys.dtype)


def _sort_noalloc(lib, h, keys, start, end, stream=None):
    dk = _device(keys)
    torch.cuda.synchronize()
    rc = lib.thrs_sort_keys_noalloc(h, dk.data_ptr(), keys.shape[0], start, end, stream)
    assert rc == 0, lib.thrs_last_error()
    torch.cuda.synchronize()
    return dk.cpu().numpy().view(keys.dtype)


@pytest.mark.parametrize("key,desc", [("u32", False), ("f32", True), ("u64", False), ("f64", True)])
def test_noalloc_equals_sort_keys_two_calls_in_a_row(key, desc):
    lib = _lib.load()
    torch.cuda.set_device(0)
    h = _create(lib, key, desc)
    try:
        bits = 32 if key in ("u32", "f32") else 64
        n_max = SCRATCH_BYTES // (2 * (bits // 8))
        rng = oracle.SplitMix64(2024)
        # the largest size first, then smaller ones over the array it left dirty: multi-block path (full and
        # quarter-size tiles), one-cluster and one-block paths, odd pass counts (copy-back from the array)
        for n, start, end in ((n_max, 0, bits), (n_max - 12345, 8, 32), (300001, 0, bits), (20000, 0, bits), (4096, 0, 8),
                              (n_max, 0, 24), (1, 0, bits)):
            keys = rng.randomize(key, n, mode=1).astype(KEY_NP[key], copy=False)
            keys[::5] = keys[0]
            want = _sort_existing(lib, h, keys, start, end)
            first = _sort_noalloc(lib, h, keys, start, end)
            second = _sort_noalloc(lib, h, keys, start, end)
            assert np.array_equal(first, want), f"{key} n={n} [{start},{end}): first call differs from thrs_sort_keys"
            assert np.array_equal(second, want), f"{key} n={n} [{start},{end}): second call differs from thrs_sort_keys"
            exp, _ = oracle.sort(key, keys, None, start, end, desc)
            assert np.array_equal(first, exp), "differs from the oracle"
    finally:
        lib.thrs_destroy(h)


def test_noalloc_back_to_back_on_one_stream_without_sync():
    """Two different sorts enqueued back to back on one stream share the array in stream order."""
    lib = _lib.load()
    torch.cuda.set_device(0)
    h = _create(lib, "u32")
    try:
        rng = oracle.SplitMix64(7)
        a = rng.randomize("u32", 1 << 20, mode=1)
        b = rng.randomize("u32", 777777, mode=1)
        da, db = _device(a), _device(b)
        st = torch.cuda.Stream()
        torch.cuda.synchronize()
        for _ in range(3):  # re-sorting sorted data on later rounds: same result
            assert lib.thrs_sort_keys_noalloc(h, da.data_ptr(), a.shape[0], 0, 32, st.cuda_stream) == 0, lib.thrs_last_error()
            assert lib.thrs_sort_keys_noalloc(h, db.data_ptr(), b.shape[0], 0, 32, st.cuda_stream) == 0, lib.thrs_last_error()
        st.synchronize()
        assert np.array_equal(da.cpu().numpy().view(np.uint32), np.sort(a, kind="stable"))
        assert np.array_equal(db.cpu().numpy().view(np.uint32), np.sort(b, kind="stable"))
    finally:
        lib.thrs_destroy(h)


def test_noalloc_allocates_nothing_and_rejects_what_does_not_fit():
    lib = _lib.load()
    torch.cuda.set_device(0)
    h = _create(lib, "u32")
    h64 = _create(lib, "u64")
    try:
        assert lib.thrs_sort_keys_noalloc(h, None, 0, 0, 32, None) == 0  # warm-up form: n = 0
        keys = oracle.SplitMix64(1).randomize("u32", 1 << 20, mode=1)
        dk = _device(keys)
        # one sort first: the driver loads a kernel's code on its first launch, which is not this call's doing
        assert lib.thrs_sort_keys_noalloc(h, dk.data_ptr(), keys.shape[0], 0, 32, None) == 0, lib.thrs_last_error()
        torch.cuda.synchronize()
        free0, _ = torch.cuda.mem_get_info()
        for _ in range(4):
            assert lib.thrs_sort_keys_noalloc(h, dk.data_ptr(), keys.shape[0], 0, 32, None) == 0, lib.thrs_last_error()
        torch.cuda.synchronize()
        free1, _ = torch.cuda.mem_get_info()
        assert free1 == free0, "device memory in use changed across thrs_sort_keys_noalloc calls"
        assert np.array_equal(dk.cpu().numpy().view(np.uint32), np.sort(keys, kind="stable"))
        # beyond the static array: a clear error, nothing enqueued, input untouched
        big = torch.zeros((1 << 20) + 1, dtype=torch.int32, device="cuda")
        assert lib.thrs_sort_keys_noalloc(h, big.data_ptr(), (1 << 20) + 1, 0, 32, None) == 1
        assert b"static scratch" in lib.thrs_last_error()
        assert lib.thrs_sort_keys_noalloc(h64, big.data_ptr(), (1 << 19) + 1, 0, 64, None) == 1
        assert lib.thrs_sort_keys_noalloc(h64, big.data_ptr(), 1 << 19, 0, 64, None) == 0
        # argument checks like thrs_sort_keys (hpp:856)
        assert lib.thrs_sort_keys_noalloc(h, dk.data_ptr(), 16, 0, 12, None) == 2
        assert lib.thrs_sort_keys_noalloc(None, dk.data_ptr(), 16, 0, 32, None) == 1
        torch.cuda.synchronize()
    finally:
        lib.thrs_destroy(h)
        lib.thrs_destroy(h64)

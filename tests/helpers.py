"""Shared helpers of the parity tests (GPU side).  Everything goes through the Python mirror of
the thrs:: interface, i.e. through the C ABI of libthrs_b200.so."""
import numpy as np
import torch

import oracle
import tinyhipradixsort_b200 as thrs

KEY_BITS = {"u32": 32, "u64": 64, "f32": 32, "f64": 64}
# torch has no uint64 arithmetic but can carry the bits
_TORCH_OF = {4: torch.int32, 8: torch.int64}


def to_device(a):
    """numpy array (any 4/8/16-byte rows) -> CUDA tensor carrying the same bytes."""
    a = np.ascontiguousarray(a)
    if a.dtype.itemsize in (4, 8) and a.ndim == 1:
        return torch.from_numpy(a.view(np.int32 if a.dtype.itemsize == 4 else np.int64)).cuda()
    return torch.from_numpy(a.view(np.uint8).reshape(a.shape[0], -1)).cuda()


def to_host(t, like):
    like = np.ascontiguousarray(like)
    out = t.cpu().numpy()
    return out.view(like.dtype).reshape(like.shape)


def make_sorter(key, value_bytes=None, descending=False):
    cfg = thrs.RadixSort.Config()
    if value_bytes is None:
        cfg.configureWithKey(key)
    else:
        cfg.configureWithKeyPair(key, value_bytes)
    cfg.sortOrder = thrs.SortOrder.Descending if descending else thrs.SortOrder.Ascending
    return thrs.RadixSort([], cfg)


def gpu_sort(sorter, keys, values=None, start=0, end=None, key="u32", temp_n=None):
    """The flow of unittest.cpp:131-153 / 379-414: upload, temp alloc, sort, sync, download."""
    n = keys.shape[0]
    end = KEY_BITS[key] if end is None else end
    dk = to_device(keys)
    dv = to_device(values) if values is not None else None
    d = sorter.getTemporaryBufferBytes(n if temp_n is None else temp_n)
    nbytes = d.getTemporaryBufferBytesForSortPairs() if values is not None else d.getTemporaryBufferBytesForSortKeys()
    tmp = torch.empty(max(nbytes, 16), dtype=torch.uint8, device="cuda")
    tmp.fill_(0xA5)  # the library must not rely on zeroed scratch
    if values is None:
        sorter.sortKeys(dk, n, tmp, start, end)
    else:
        sorter.sortPairs(dk, dv, n, tmp, start, end)
    torch.cuda.synchronize()
    return to_host(dk, keys), (to_host(dv, values) if values is not None else None)


def check_against_oracle(key, keys, values=None, start=0, end=None, descending=False, sorter=None):
    sorter = sorter or make_sorter(key, None if values is None else values[0].nbytes if values.ndim > 1 else values.dtype.itemsize, descending)
    got_k, got_v = gpu_sort(sorter, keys, values, start, end, key)
    exp_k, exp_v = oracle.sort(key, keys, values, start, end, descending)
    assert np.array_equal(got_k, exp_k), "keys differ from oracle"
    if values is not None:
        assert np.array_equal(got_v, exp_v), "values differ from oracle (stability)"
    return got_k, got_v


_cudart = None


def memcpy_d2d(dst_ptr, src_ptr, nbytes):
    """Blocking cudaMemcpy(DeviceToDevice) on raw addresses (for thrs.Buffer, which is not a tensor)."""
    import ctypes
    global _cudart
    if _cudart is None:
        _cudart = ctypes.CDLL("libcudart.so.12")
        _cudart.cudaMemcpy.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
    torch.cuda.synchronize()
    rc = _cudart.cudaMemcpy(dst_ptr, src_ptr, nbytes, 3)
    assert rc == 0, f"cudaMemcpy rc={rc}"

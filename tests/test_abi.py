"""CPU: the C-ABI library loads and exports every symbol include/thrs_b200.h declares; host-only
entry points (config, temp sizes, key bits, argument validation) behave like the reference."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

import oracle
import tinyhipradixsort_b200 as thrs
from tinyhipradixsort_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "thrs_b200.h")).read()
    return sorted(set(re.findall(r"THRS_API\s+[\w\s\*]+?\b(thrs_\w+)\s*\(", text)))


def test_header_symbols_exported_and_bound():
    names = _declared()
    assert len(names) >= 15
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in thrs_b200.h but not exported"
    assert sorted(_lib.SIGNATURES) == names, "python binding and header disagree"
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    assert set(names) <= exported
    # nothing but the C ABI leaks out of the library
    assert all(e.startswith("thrs_") for e in exported), exported


def test_library_is_sm100a_only():
    out = subprocess.run(["cuobjdump", "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\w+", out))
    assert archs == {"sm_100a"}, archs


def test_product_does_not_touch_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "tinyhipradixsort_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", "Makefile")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.lower(), (dirpath, f)
    out = subprocess.run(["ldd", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out


def test_config_defaults_and_enums():
    c = _lib.thrs_config()
    _lib.load().thrs_config_default(ctypes.byref(c))
    assert (c.key_is_16byte_aligned, c.key_type, c.value_type, c.sort_order) == (1, 0, 0, 0)  # hpp:699-705
    cfg = thrs.RadixSort.Config()
    assert cfg.keyIs16byteAligned and cfg.keyType == thrs.KeyType.U32 and cfg.valueType == thrs.ValueType.U32
    cfg.configureWithKeyPair("f64", 16)
    assert cfg.keyType == thrs.KeyType.F64 and cfg.valueType == thrs.ValueType.U128
    assert thrs.bytesOf(thrs.KeyType.F64) == 8 and thrs.bytesOf(thrs.ValueType.U128) == 16
    with pytest.raises(thrs.ThrsError):
        cfg.configureWithKeyPair("u32", 12)


def test_create_rejects_bad_enums():
    lib = _lib.load()
    h = ctypes.c_void_p()
    for bad in ((1, 9, 0, 0), (1, 0, 7, 0), (1, 0, 0, 5)):
        c = _lib.thrs_config(*bad)
        assert lib.thrs_create(ctypes.byref(c), ctypes.byref(h)) == 1
        assert lib.thrs_last_error()


@pytest.mark.parametrize("key,vbytes", [("u32", 4), ("u64", 8), ("f32", 16), ("f64", 4)])
def test_temp_bytes_contract(key, vbytes):
    cfg = thrs.RadixSort.Config().configureWithKeyPair(key, vbytes)
    s = thrs.RadixSort(["--generate-line-info"], cfg)  # extraArgs accepted and ignored
    kb = oracle.KEY_BYTES[key]
    prev = None
    for n in (0, 1, 2047, 2048, 2049, 99999, 1 << 20, (1 << 28), (1 << 31) + 100, (1 << 32) - 1):
        d = s.getTemporaryBufferBytes(n)
        for v in (d.pSumBuffer, d.keyOutBuffer, d.valueOutBuffer):
            assert v % 16 == 0
        assert d.keyOutBuffer >= kb * n and d.keyOutBuffer < kb * n + 16      # hpp:840
        assert d.valueOutBuffer >= vbytes * n and d.valueOutBuffer < vbytes * n + 16  # hpp:841
        assert d.getTemporaryBufferBytesForSortKeys() == d.pSumBuffer + d.keyOutBuffer
        assert d.getTemporaryBufferBytesForSortPairs() == d.pSumBuffer + d.keyOutBuffer + d.valueOutBuffer
        assert d.getOutputKeyBuffer(1000) == 1000 + d.pSumBuffer
        assert d.getOutputValueBuffer(1000) == 1000 + d.pSumBuffer + d.keyOutBuffer
        if prev is not None:  # monotone in n: a buffer sized for n_max serves any n <= n_max (hpp:859)
            assert d.pSumBuffer >= prev.pSumBuffer and d.keyOutBuffer >= prev.keyOutBuffer
        prev = d
    # scratch stays a small fraction of the data (reference: 0.5 B per key)
    d = s.getTemporaryBufferBytes(1 << 28)
    assert d.pSumBuffer <= (1 << 28) * kb / 2 + (1 << 20)


def test_host_key_bits_match_oracle_and_golden():
    import json

    with open(os.path.join(ROOT, "tests", "golden", "fpkey_kat.json")) as f:
        kat = json.load(f)
    for key in ("u32", "u64", "f32", "f64"):
        for raw, exp in kat[key]:
            assert thrs.key_bits_from_raw(int(raw, 16), key) == int(exp, 16)
    r = oracle.SplitMix64(5)
    for x in r.randomize("f32", 5000, mode=1):
        assert thrs.key_bits_from_raw(int(x), "f32") == oracle.key_bits("f32", x)
    for x in r.randomize("f64", 5000, mode=1):
        assert thrs.key_bits_from_raw(int(x), "f64") == oracle.key_bits("f64", x)
    assert thrs.getKeyBits(-0.0) == thrs.getKeyBits(0.0) == 0x80000000  # unittest.cpp:83


def test_argument_validation_without_gpu():
    s = thrs.RadixSort([], thrs.RadixSort.Config())
    with pytest.raises(thrs.ThrsError) as e:
        s.sortKeys(16, 10, 16, 0, 12)  # (end-start) % 8 != 0, hpp:856
    assert e.value.code == 2
    with pytest.raises(thrs.ThrsError):
        s.sortKeys(16, 10, 16, -8, 0)
    s.sortKeys(0, 0, 0, 0, 32)   # n == 0: nothing to do, nothing touched
    s.sortKeys(16, 10, 16, 8, 8)  # empty bit range: no-op like the reference loop (hpp:867)
    with pytest.raises(thrs.ThrsError):
        s.sortKeys(None, 10, 16, 0, 32)
    with pytest.raises(thrs.ThrsError) as e:
        s.sortKeys(18, 10, 16, 0, 32)  # misaligned key pointer
    assert e.value.code == 5
    with pytest.raises(thrs.ThrsError):
        s.sortKeys(np.zeros(4), 4, 16, 0, 32)  # not a device buffer


def test_no_cpu_fallback_without_device():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    s = thrs.RadixSort([], thrs.RadixSort.Config())
    with pytest.raises(thrs.ThrsError) as e:
        s.sortKeys(4096, 10, 8192, 0, 32)  # well-formed call, but no device: must fail loudly
    assert e.value.code in (3, 4)
    k = np.arange(10, dtype=np.uint32)
    with pytest.raises(thrs.ThrsError):
        s.sortHost(k, None, 0, 32)
    with pytest.raises(thrs.ThrsError):
        s.sortHostSubmit(0, k, None, 0, 32)
    with pytest.raises(thrs.ThrsError):
        s.sortHostWait(thrs.RadixSort.HOST_SLOTS)  # slot out of range is rejected without a device, too
    s.sortHostWait(0)  # nothing in flight: returns


def test_noalloc_entry_point_validates_without_a_device():
    """thrs_sort_keys_noalloc: argument errors are reported before anything touches a device, and without a
    device the call fails loudly like every other sort (no CPU fallback)."""
    import torch

    lib = _lib.load()
    cfg = _lib.thrs_config()
    lib.thrs_config_default(ctypes.byref(cfg))
    h = ctypes.c_void_p()
    assert lib.thrs_create(ctypes.byref(cfg), ctypes.byref(h)) == 0
    try:
        assert lib.thrs_sort_keys_noalloc(None, 4096, 16, 0, 32, None) == 1
        assert lib.thrs_sort_keys_noalloc(h, 4096, 16, 0, 12, None) == 2      # hpp:856: multiple of 8
        assert lib.thrs_sort_keys_noalloc(h, 4096, 16, 16, 8, None) == 2
        assert lib.thrs_sort_keys_noalloc(h, 4096, (1 << 20) + 1, 0, 32, None) == 1
        assert b"static scratch" in lib.thrs_last_error()
        if not torch.cuda.is_available():
            assert lib.thrs_sort_keys_noalloc(h, 4096, 16, 0, 32, None) in (1, 3, 4)  # never THRS_OK without a device
            assert lib.thrs_last_error()
    finally:
        lib.thrs_destroy(h)

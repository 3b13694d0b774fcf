"""CPU: the oracle against the golden vectors, the reference's own fpKey build, and itself."""
import hashlib
import json
import os

import numpy as np
import pytest

import oracle

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _kat():
    with open(os.path.join(GOLDEN, "fpkey_kat.json")) as f:
        return json.load(f)


def test_splitmix64_known_answers():
    # unittest.cpp:24-35 seeded with x=0
    r = oracle.SplitMix64(0)
    got = [r.next() for _ in range(4)]
    assert got == [0xE220A8397B1DCDAF, 0x6E789E6AA1B965F4, 0x06C45D188009454F, 0xF88BB8A8724C81EC]
    assert [hex(x) for x in got] == _kat()["splitmix64_seed0_first4"]
    # counter form used by the device-side generators
    assert [int(x) for x in oracle.splitmix64_at(np.arange(4))] == got
    # first n of testSortKeys (unittest.cpp:138)
    assert 1 + oracle.SplitMix64(0).next() % 99999 == 11258


@pytest.mark.parametrize("key", ["u32", "u64", "f32", "f64"])
def test_key_bits_golden(key):
    for raw, exp in _kat()[key]:
        assert oracle.key_bits(key, int(raw, 16)) == int(exp, 16), (key, raw)


def test_key_bits_survey_table():
    # SURVEY.md 8c probe table of the reference's fpKey.hpp
    table = {0x00000000: 0x80000000, 0x80000000: 0x80000000, 0x00000001: 0x80000001, 0x80000001: 0x7FFFFFFE,
             0x007FFFFF: 0x807FFFFF, 0x7F7FFFFF: 0xFF7FFFFF, 0x7F800000: 0xFF800000, 0xFF800000: 0x007FFFFF,
             0x7FC00000: 0xFFC00000, 0xFFC00000: 0x003FFFFF, 0x7F800001: 0xFF800001, 0x7FFFFFFF: 0xFFFFFFFF,
             0xFFFFFFFF: 0x00000000, 0xFF7FFFFF: 0x00800000, 0x3F800000: 0xBF800000, 0xBF800000: 0x407FFFFF}
    for raw, exp in table.items():
        assert oracle.key_bits("f32", raw) == exp


def test_key_bits_against_reference_build():
    ref = oracle.ref_fpkey()
    if ref is None:
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    r = oracle.SplitMix64(123)
    for x in r.randomize("f32", 20000, mode=1):
        assert oracle.key_bits("f32", x) == ref.ref_key_bits_f32(int(x))
    for x in r.randomize("f64", 20000, mode=1):
        assert oracle.key_bits("f64", x) == ref.ref_key_bits_f64(int(x))


def test_fpkeys_float_property():
    # UTEST(FPKeys, float), unittest.cpp:81-94 (shortened stream)
    import struct

    def kb(f):
        return oracle.key_bits("f32", struct.unpack("<I", struct.pack("<f", f))[0])

    assert (-0.0 < 0.0) == (kb(-0.0) < kb(0.0))
    fmax = np.finfo(np.float32).max
    assert (fmax < np.inf) == (kb(float(fmax)) < kb(float("inf")))
    r = oracle.SplitMix64(0)
    for _ in range(20000):
        a = np.float32((-1.0 if r.next() % 2 == 0 else 1.0) * r.next() * 0.1)
        b = np.float32((-1.0 if r.next() % 2 == 0 else 1.0) * r.next() * 0.1)
        assert (a < b) == (kb(float(a)) < kb(float(b)))


def test_sort_golden_streams():
    with open(os.path.join(GOLDEN, "sort_kat.json")) as f:
        cases = json.load(f)["cases"]
    streams = {}
    for c in cases:
        sid = (c["key"], c["descending"], c["pairs"], c.get("stream", "Sort"))
        if c["iteration"] == 0:
            streams[sid] = oracle.SplitMix64(0)
        rng = streams[sid]
        n = 1 + rng.next() % 99999
        assert n == c["n"]
        if c.get("stream") == "StartBits":
            s = rng.next() % 64
            assert s == c["start"]
        raw = rng.randomize(c["key"], n, mode=0)
        vals = np.arange(n, dtype=np.uint32) if c["pairs"] else None
        k, v = oracle.sort(c["key"], raw, vals, c["start"], c["end"], c["descending"])
        h = hashlib.sha256(k.tobytes())
        if c["pairs"]:
            h.update(v.tobytes())
        assert h.hexdigest() == c["sha256"], c
        assert hex(int(k[0])) == c["first_key"] and hex(int(k[-1])) == c["last_key"]


@pytest.mark.parametrize("key", ["u32", "u64", "f32", "f64"])
@pytest.mark.parametrize("descending", [False, True])
def test_field_sort_equals_lsd_passes(key, descending):
    # the one-shot field sort and the literal per-digit loop of hpp:867-934 agree
    r = oracle.SplitMix64(7)
    bits = oracle.KEY_BYTES[key] * 8
    for start, end in ((0, bits), (8, 24), (3, 19), (bits - 4, bits + 4), (0, bits + 16)):
        raw = r.randomize(key, 3000, mode=1)
        raw[::7] = raw[0]  # duplicates
        vals = np.arange(raw.shape[0], dtype=np.uint32)
        k1, v1 = oracle.sort(key, raw, vals, start, end, descending)
        k2, v2 = oracle.sort(key, raw, vals, start, end, descending, lsd=True)
        assert np.array_equal(k1, k2) and np.array_equal(v1, v2)


def test_negative_zero_and_nan_order():
    raw = np.array([0x80000000, 0x00000000, 0x7FC00000, 0xFFC00000, 0x7F800000, 0xFF800000, 0x00000001,
                    0x80000001, 0x00000000, 0x80000000], dtype=np.uint32)
    vals = np.arange(raw.shape[0], dtype=np.uint32)
    k, v = oracle.sort("f32", raw, vals)
    # -NaN < -Inf < -denorm < (-0 == +0, input order kept) < +denorm < +Inf < +NaN
    assert [hex(x) for x in k] == ["0xffc00000", "0xff800000", "0x80000001", "0x80000000", "0x0", "0x0",
                                   "0x80000000", "0x1", "0x7f800000", "0x7fc00000"]
    assert list(v[3:7]) == [0, 1, 8, 9]


def test_bad_bit_range():
    raw = np.zeros(4, dtype=np.uint32)
    with pytest.raises(ValueError):
        oracle.sort("u32", raw, None, 0, 12)  # THRS_ASSERT hpp:856


def test_counter_pair_checker():
    n = 5000
    keys = (oracle.splitmix64_at(np.arange(n)) & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    vals = np.arange(n, dtype=np.uint32)
    k, v = oracle.sort("u32", keys, vals)
    assert oracle.check_counter_pairs_u32(k, v) == -1
    k2 = k.copy()
    k2[100], k2[101] = k2[101], k2[100]
    assert oracle.check_counter_pairs_u32(k2, v) >= 0

"""Generates tests/golden/*.json.  Run in the build container (needs /root/reference through
oracle/_ref/libref_fpkey.so = the reference's own fpKey.hpp compiled by oracle/Makefile):

    make -C oracle && python tests/golden/make_golden.py

fpkey_kat.json : getKeyBits known answers computed BY THE REFERENCE'S fpKey.hpp (fpKey.hpp:15-38)
                 for special IEEE patterns and a splitmix64 stream, f32 and f64.
sort_kat.json  : the reference's own test streams (unittest.cpp:24-35,96-116,131-166,379-424:
                 seed 0, n = 1 + rng % 99999, randomizeValues) sorted by an implementation that
                 is independent of oracle/oracle.cpp: numpy stable argsort on the reference's
                 getKeyBits, recorded as sha256 of the expected output bytes.
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

F32_SPECIAL = [0x00000000, 0x80000000, 0x00000001, 0x80000001, 0x007FFFFF, 0x807FFFFF, 0x00800000, 0x80800000,
               0x7F7FFFFF, 0xFF7FFFFF, 0x7F800000, 0xFF800000, 0x7FC00000, 0xFFC00000, 0x7F800001, 0xFF800001,
               0x7FFFFFFF, 0xFFFFFFFF, 0x3F800000, 0xBF800000, 0x7FA00000, 0xFFA00000]
F64_SPECIAL = [0x0000000000000000, 0x8000000000000000, 0x0000000000000001, 0x8000000000000001,
               0x000FFFFFFFFFFFFF, 0x800FFFFFFFFFFFFF, 0x0010000000000000, 0x8010000000000000,
               0x7FEFFFFFFFFFFFFF, 0xFFEFFFFFFFFFFFFF, 0x7FF0000000000000, 0xFFF0000000000000,
               0x7FF8000000000000, 0xFFF8000000000000, 0x7FF0000000000001, 0xFFF0000000000001,
               0x7FFFFFFFFFFFFFFF, 0xFFFFFFFFFFFFFFFF, 0x3FF0000000000000, 0xBFF0000000000000]


def ref_bits_np(key, raw):
    ref = oracle.ref_fpkey()
    f = getattr(ref, "ref_key_bits_" + key)
    return np.array([f(int(x)) for x in raw], dtype=raw.dtype)


def main():
    ref = oracle.ref_fpkey()
    assert ref is not None, "build oracle/_ref first (make -C oracle)"
    # ---- fpkey KAT
    rng = oracle.SplitMix64(0)
    f32 = F32_SPECIAL + [int(x) for x in rng.randomize("f32", 200, mode=1)]
    f64 = F64_SPECIAL + [int(x) for x in rng.randomize("f64", 200, mode=1)]
    kat = {
        "source": "reference fpKey.hpp:15-38 compiled with g++ -O2 (oracle/_ref/libref_fpkey.so)",
        "splitmix64_seed0_first4": [hex(oracle.SplitMix64(0).next())] + [],
        "f32": [[hex(x), hex(ref.ref_key_bits_f32(x))] for x in f32],
        "f64": [[hex(x), hex(ref.ref_key_bits_f64(x))] for x in f64],
        "u32": [[hex(x), hex(ref.ref_key_bits_u32(x))] for x in (0, 1, 0x80000000, 0xFFFFFFFF)],
        "u64": [[hex(x), hex(ref.ref_key_bits_u64(x))] for x in (0, 1, 1 << 63, (1 << 64) - 1)],
    }
    r = oracle.SplitMix64(0)
    kat["splitmix64_seed0_first4"] = [hex(r.next()) for _ in range(4)]
    with open(os.path.join(HERE, "fpkey_kat.json"), "w") as f:
        json.dump(kat, f, indent=1)

    # ---- sort KAT: reference test streams, first 3 iterations per case
    cases = []
    for key in ("u32", "u64", "f32", "f64"):
        for descending in (False, True):
            for pairs in (False, True):
                rng = oracle.SplitMix64(0)
                for it in range(3):
                    n = 1 + rng.next() % 99999  # unittest.cpp:138
                    raw = rng.randomize(key, n, mode=0)
                    kb = ref_bits_np(key, raw)
                    if descending:
                        kb = ~kb
                    perm = np.argsort(kb, kind="stable")
                    exp_k = raw[perm]
                    h = hashlib.sha256(exp_k.tobytes())
                    if pairs:
                        h.update(perm.astype(np.uint32).tobytes())  # values = 0..n-1 (unittest.cpp:394-397)
                    cases.append({"key": key, "descending": descending, "pairs": pairs, "iteration": it,
                                  "n": int(n), "start": 0, "end": 8 * raw.dtype.itemsize,
                                  "first_key": hex(int(exp_k[0])), "last_key": hex(int(exp_k[-1])),
                                  "sha256": h.hexdigest()})
    # StartBits.u64 stream (unittest.cpp:248-355): window [s, s+8)
    for descending in (False, True):
        rng = oracle.SplitMix64(0)
        for it in range(6):
            n = 1 + rng.next() % 99999
            s = rng.next() % 64
            raw = rng.randomize("u64", n, mode=0)
            kb = ~raw if descending else raw
            fld = (kb >> np.uint64(s)) & np.uint64(0xFF)
            perm = np.argsort(fld, kind="stable")
            exp_k = raw[perm]
            h = hashlib.sha256(exp_k.tobytes())
            h.update(perm.astype(np.uint32).tobytes())
            cases.append({"key": "u64", "descending": descending, "pairs": True, "iteration": it, "n": int(n),
                          "start": int(s), "end": int(s) + 8, "first_key": hex(int(exp_k[0])),
                          "last_key": hex(int(exp_k[-1])), "sha256": h.hexdigest(), "stream": "StartBits"})
    with open(os.path.join(HERE, "sort_kat.json"), "w") as f:
        json.dump({"source": "numpy stable argsort on reference getKeyBits; streams of unittest.cpp", "cases": cases},
                  f, indent=1)
    print("wrote", len(kat["f32"]), "f32 /", len(kat["f64"]), "f64 key-bit vectors and", len(cases), "sort cases")


if __name__ == "__main__":
    main()

"""fpKey.hpp mirror (reference fpKey.hpp:15-38): the host ``getKeyBits`` family, served by the
C ABI (``thrs_key_bits_*``) so Python and C++ callers share one implementation."""
import struct

from . import _lib


def getKeyBits(x, kind=None):
    """getKeyBits(x).  ``kind`` in {'u32','u64','f32','f64'}; inferred as f32 for Python floats
    and u32/u64 for ints when omitted (the C++ overload set of fpKey.hpp)."""
    lib = _lib.load()
    if kind is None:
        kind = "f32" if isinstance(x, float) else ("u32" if int(x) < (1 << 32) else "u64")
    if kind == "u32":
        return lib.thrs_key_bits_u32(int(x))
    if kind == "u64":
        return lib.thrs_key_bits_u64(int(x))
    if kind == "f32":
        return lib.thrs_key_bits_f32(struct.unpack("<I", struct.pack("<f", x))[0])
    if kind == "f64":
        return lib.thrs_key_bits_f64(struct.unpack("<Q", struct.pack("<d", x))[0])
    raise ValueError(kind)


def key_bits_from_raw(raw_bits, kind):
    """getKeyBits on a raw IEEE bit pattern (lets tests feed NaN payloads exactly)."""
    lib = _lib.load()
    return {"u32": lib.thrs_key_bits_u32, "u64": lib.thrs_key_bits_u64,
            "f32": lib.thrs_key_bits_f32, "f64": lib.thrs_key_bits_f64}[kind](int(raw_bits))

"""tinyhipradixsort_b200 — B200-native (sm_100a) radix sort with the tinyhipradixsort API.

The product is ``libthrs_b200.so`` (C ABI: include/thrs_b200.h; C++ drop-in header:
include/tinyhipradixsort.hpp).  This package is the thin Python mirror of the reference's
``thrs::`` interface used by the tests and bench.py.
"""
from .radixsort import (Buffer, KeyType, RadixSort, SortOrder, TemporaryBufferDef, ThrsError,
                        ValueType, bytesOf)
from .fpkey import getKeyBits, key_bits_from_raw

__all__ = ["Buffer", "KeyType", "RadixSort", "SortOrder", "TemporaryBufferDef", "ThrsError",
           "ValueType", "bytesOf", "getKeyBits", "key_bits_from_raw"]

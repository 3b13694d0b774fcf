"""ctypes binding of the C ABI (include/thrs_b200.h) — libthrs_b200.so, built in-tree by
``tinyhipradixsort_b200/csrc/Makefile`` (``__graft_entry__.build()``).

There is deliberately no fallback: if the shared library is missing the import of the
sort fails loudly, and if no CUDA device is usable every sort call raises.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libthrs_b200.so")


class thrs_config(ctypes.Structure):
    _fields_ = [
        ("key_is_16byte_aligned", ctypes.c_int32),
        ("key_type", ctypes.c_int32),
        ("value_type", ctypes.c_int32),
        ("sort_order", ctypes.c_int32),
    ]


class thrs_temp_def(ctypes.Structure):
    _fields_ = [
        ("pSumBuffer", ctypes.c_uint64),
        ("keyOutBuffer", ctypes.c_uint64),
        ("valueOutBuffer", ctypes.c_uint64),
    ]


class thrs_profile(ctypes.Structure):
    _fields_ = [
        ("memset_ms", ctypes.c_float),
        ("histogram_ms", ctypes.c_float),
        ("pass_ms", ctypes.c_float * 8),
        ("copyback_ms", ctypes.c_float),
        ("total_ms", ctypes.c_float),
        ("passes", ctypes.c_int32),
        ("parts", ctypes.c_int32),
    ]


# name -> (restype, argtypes); must list every symbol include/thrs_b200.h declares
# (tests/test_abi.py checks the two against each other).
SIGNATURES = {
    "thrs_config_default": (None, [ctypes.POINTER(thrs_config)]),
    "thrs_create": (ctypes.c_int, [ctypes.POINTER(thrs_config), ctypes.POINTER(ctypes.c_void_p)]),
    "thrs_destroy": (None, [ctypes.c_void_p]),
    "thrs_temp_bytes": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_uint32, ctypes.POINTER(thrs_temp_def)]),
    "thrs_sort_keys": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p,
                                      ctypes.c_int, ctypes.c_int, ctypes.c_void_p]),
    "thrs_sort_pairs": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32,
                                       ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]),
    "thrs_sort_keys_noalloc": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int, ctypes.c_int,
                                              ctypes.c_void_p]),
    "thrs_digit_histogram": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int,
                                            ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]),
    "thrs_sort_host": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32,
                                      ctypes.c_int, ctypes.c_int]),
    "thrs_sort_host_submit": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32,
                                             ctypes.c_int, ctypes.c_int]),
    "thrs_sort_host_wait": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "thrs_key_bits_u32": (ctypes.c_uint32, [ctypes.c_uint32]),
    "thrs_key_bits_u64": (ctypes.c_uint64, [ctypes.c_uint64]),
    "thrs_key_bits_f32": (ctypes.c_uint32, [ctypes.c_uint32]),
    "thrs_key_bits_f64": (ctypes.c_uint64, [ctypes.c_uint64]),
    "thrs_buffer_alloc": (ctypes.c_int, [ctypes.c_int64, ctypes.POINTER(ctypes.c_void_p)]),
    "thrs_buffer_free": (ctypes.c_int, [ctypes.c_void_p]),
    "thrs_profile_enable": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "thrs_profile_read": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(thrs_profile)]),
    "thrs_last_error": (ctypes.c_char_p, []),
    "thrs_version": (ctypes.c_char_p, []),
    "thrs_last_launch_count": (ctypes.c_int, []),
    "thrs_tile_items": (ctypes.c_uint32, [ctypes.c_void_p]),
    "thrs_scatter_pairs": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32,
                                          ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                          ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]),
    "thrs_ipc_export": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p]),
    "thrs_ipc_open": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)]),
    "thrs_ipc_close": (ctypes.c_int, [ctypes.c_void_p]),
    "thrs_mg_create": (ctypes.c_int, [ctypes.c_int, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(thrs_config),
                                      ctypes.POINTER(ctypes.c_void_p)]),
    "thrs_mg_destroy": (None, [ctypes.c_void_p]),
    "thrs_mg_sort_pairs": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_void_p),
                                          ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_void_p),
                                          ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_uint64),
                                          ctypes.POINTER(ctypes.c_uint64)]),
    "thrs_mg_last_error": (ctypes.c_char_p, []),
    "thrs_set_small_path": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "thrs_set_mid_max": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_uint32]),
    "thrs_set_cluster_path": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "thrs_cluster_sort_max": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_uint32)]),
    "thrs_set_ranking": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "thrs_get_ranking": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int),
                                        ctypes.POINTER(ctypes.c_uint64)]),
    "thrs_segments_temp_bytes": (ctypes.c_uint64, [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_int]),
    "thrs_segments_scratch_bytes": (ctypes.c_uint64, [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint32]),
    "thrs_sort_segments_keys": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p,
                                               ctypes.c_int, ctypes.c_int, ctypes.c_void_p]),
    "thrs_sort_segments_pairs": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32,
                                                ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]),
    "thrs_sort_segments_ex": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                             ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int,
                                             ctypes.c_int, ctypes.c_void_p, ctypes.POINTER(ctypes.c_int)]),
    "thrs_partition_pairs": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p,
                                            ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                            ctypes.c_void_p]),
    "thrs_xchg_create": (ctypes.c_int, [ctypes.POINTER(thrs_config), ctypes.c_int, ctypes.c_int, ctypes.c_uint64, ctypes.c_uint64,
                                        ctypes.c_int, ctypes.POINTER(ctypes.c_void_p)]),
    "thrs_xchg_destroy": (None, [ctypes.c_void_p]),
    "thrs_xchg_window": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_uint64)]),
    "thrs_xchg_set_peers": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)]),
    "thrs_xchg_begin": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p]),
    "thrs_xchg_finish": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_void_p),
                                        ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint64)]),
    "thrs_xchg_plan_step": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_int)]),
    "thrs_xchg_send_group": (ctypes.c_int, [ctypes.c_void_p]),
    "thrs_xchg_recv_group": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p]),
    "thrs_xchg_result": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_void_p),
                                        ctypes.POINTER(ctypes.c_uint64)]),
    "thrs_xchg_stats": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint64),
                                       ctypes.POINTER(ctypes.c_uint32), ctypes.POINTER(ctypes.c_uint32)]),
    "thrs_xchg_plan": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_int32),
                                      ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32)]),
    "thrs_sort_buckets_ex": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                            ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32,
                                            ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                            ctypes.POINTER(ctypes.c_int)]),
    "thrs_xchg_last_error": (ctypes.c_char_p, []),
    "thrs_set_sm_limit": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "thrs_get_ranking_check": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_uint64)]),
}

_lib = None


def load():
    """Load libthrs_b200.so (once) and attach prototypes.  Raises if it was not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `make -C tinyhipradixsort_b200/csrc` "
            "(or __graft_entry__.build()). There is no CPU or PyTorch fallback."
        )
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export it
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib

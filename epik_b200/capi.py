"""ctypes binding of the C ABI in include/epik_b200.h (libepik_b200.so, built by __graft_entry__.build()).

This is the stub a Python caller of the reference would add; the C++ equivalent is in INTEGRATION.md.
There is no fallback: if the shared library is missing, importing this module raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("EPIK_B200_LIB") or os.path.join(_HERE, "libepik_b200.so")  # override: kernel experiments

EPK_OK, EPK_E_INVALID, EPK_E_CUDA, EPK_E_NOMEM, EPK_E_SHORT_SEQ, EPK_E_UNSUPPORTED = range(6)
STATUS_NAMES = ["EPK_OK", "EPK_E_INVALID", "EPK_E_CUDA", "EPK_E_NOMEM", "EPK_E_SHORT_SEQ", "EPK_E_UNSUPPORTED"]
UINT64_MAX = (1 << 64) - 1

# every symbol include/epik_b200.h declares (tests check that the library exports them all)
SYMBOLS = [
    "epk_abi_version", "epk_last_error", "epk_device_count", "epk_db_create", "epk_db_get_info", "epk_db_destroy",
    "epk_placer_create", "epk_placer_destroy", "epk_place", "epk_place_device", "epk_group_batch", "epk_kmers",
    "epk_placer_counters", "epk_host_alloc", "epk_host_free",
    "epk_db_create_shard", "epk_shard_of_key", "epk_partial_row_bytes", "epk_place_partial_device",
    "epk_place_finalize_device", "epk_gather_ceiling_device",
    "epk_device_alloc", "epk_device_free", "epk_ipc_export", "epk_ipc_import", "epk_ipc_close",
    "epk_gather_microbench", "epk_placer_set_trace", "epk_trace_term",
    "epk_db_get_view", "epk_placer_set_shard_views", "epk_place_finalize_seq_device", "epk_db_create_synthetic", "epk_place_many", "epk_group_batches", "epk_place_device_noalloc", "epk_db_pool", "epk_db_pool_share", "epk_db_pool_join",
    "epk_db_get_score_range", "epk_db_pool_set_score_range",
]


class EpkError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__("%s: %s" % (STATUS_NAMES[status] if 0 <= status < len(STATUS_NAMES) else status, message))
        self.status = status


class DbDesc(C.Structure):
    _fields_ = [
        ("alphabet", C.c_uint32), ("kmer_size", C.c_uint32), ("omega", C.c_float), ("n_keys", C.c_uint64),
        ("keys", C.c_void_p), ("offsets", C.c_void_p), ("entries", C.c_void_p), ("filter_value", C.c_void_p),
        ("n_nodes", C.c_uint32), ("branch_length", C.c_void_p), ("subtree_num_nodes", C.c_void_p),
        ("subtree_total_length", C.c_void_p),
    ]


class LoadFilter(C.Structure):
    _fields_ = [("mu", C.c_double), ("omega", C.c_float), ("max_entries", C.c_uint64)]


class DbInfo(C.Structure):
    _fields_ = [
        ("alphabet", C.c_uint32), ("kmer_size", C.c_uint32), ("n_nodes", C.c_uint32), ("table_kind", C.c_uint32),
        ("omega", C.c_float), ("threshold", C.c_float), ("log10_threshold", C.c_float),
        ("n_keys_loaded", C.c_uint64), ("n_entries_loaded", C.c_uint64), ("n_entries_total", C.c_uint64),
        ("table_slots", C.c_uint64), ("device_bytes", C.c_uint64),
    ]


class DbView(C.Structure):
    _fields_ = [("table", C.c_uint64), ("postings", C.c_uint64), ("hash_mask", C.c_uint64), ("hash_shift", C.c_uint32),
                ("table_kind", C.c_uint32)]


class SynthDesc(C.Structure):
    _fields_ = [("alphabet", C.c_uint32), ("kmer_size", C.c_uint32), ("omega", C.c_float), ("n_nodes", C.c_uint32),
                ("branch_length", C.c_void_p), ("subtree_num_nodes", C.c_void_p), ("subtree_total_length", C.c_void_p),
                ("seed", C.c_uint64), ("presence_threshold", C.c_uint64), ("length_table", C.c_void_p),
                ("length_table_bits", C.c_uint32), ("score_lo", C.c_double), ("key_space_limit", C.c_uint64)]


class PlacerOpts(C.Structure):
    _fields_ = [("variant", C.c_uint32), ("chunk_reads", C.c_uint32), ("reserved", C.c_uint32 * 6)]


class Result(C.Structure):
    _fields_ = [
        ("n_place", C.c_void_p), ("branch", C.c_void_p), ("score", C.c_void_p), ("lwr", C.c_void_p),
        ("count", C.c_void_p), ("distal", C.c_void_p), ("pendant", C.c_void_p),
    ]


class Counters(C.Structure):
    _fields_ = [("n_reads", C.c_uint64), ("n_windows", C.c_uint64), ("n_hits", C.c_uint64), ("n_entries", C.c_uint64),
                ("n_launches", C.c_uint64), ("n_short", C.c_uint64), ("n_unsupported", C.c_uint64)]


class Microbench(C.Structure):
    _fields_ = [("ms_per_launch", C.c_double), ("n_probes", C.c_uint64), ("n_hits", C.c_uint64),
                ("n_entries", C.c_uint64), ("bytes", C.c_uint64), ("checksum", C.c_uint64)]


_lib = None


def lib():
    """The loaded libepik_b200.so; raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "%s is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(epik_b200 has no CPU or pure-Python path)" % LIB_PATH)
    l = C.CDLL(LIB_PATH)
    l.epk_abi_version.restype = C.c_uint32
    l.epk_last_error.restype = C.c_char_p
    l.epk_device_count.restype = C.c_int
    l.epk_db_create.restype = C.c_int
    l.epk_db_create.argtypes = [C.POINTER(DbDesc), C.POINTER(LoadFilter), C.c_int, C.POINTER(C.c_void_p)]
    l.epk_db_get_info.restype = C.c_int
    l.epk_db_get_info.argtypes = [C.c_void_p, C.POINTER(DbInfo)]
    l.epk_db_destroy.argtypes = [C.c_void_p]
    l.epk_db_destroy.restype = None
    l.epk_placer_create.restype = C.c_int
    l.epk_placer_create.argtypes = [C.c_void_p, C.c_uint64, C.c_double, C.POINTER(PlacerOpts), C.POINTER(C.c_void_p)]
    l.epk_placer_destroy.argtypes = [C.c_void_p]
    l.epk_placer_destroy.restype = None
    l.epk_place.restype = C.c_int
    l.epk_place.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.POINTER(Result)]
    l.epk_place_device_noalloc.restype = C.c_int
    l.epk_place_device_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.POINTER(Result), C.c_void_p]
    l.epk_place_many.restype = C.c_int
    l.epk_place_many.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32, C.POINTER(Result)]
    l.epk_place_device.restype = C.c_int
    l.epk_place_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.POINTER(Result), C.c_void_p]
    l.epk_group_batch.restype = C.c_int
    l.epk_group_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
    l.epk_group_batches.restype = C.c_int
    l.epk_group_batches.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
    l.epk_kmers.restype = C.c_int
    l.epk_kmers.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p,
                            C.c_void_p, C.POINTER(C.c_uint64)]
    l.epk_placer_counters.restype = C.c_int
    l.epk_placer_counters.argtypes = [C.c_void_p, C.POINTER(Counters)]
    l.epk_db_create_shard.restype = C.c_int
    l.epk_db_create_shard.argtypes = [C.POINTER(DbDesc), C.POINTER(LoadFilter), C.c_int, C.c_uint32, C.c_uint32,
                                      C.POINTER(C.c_void_p)]
    l.epk_shard_of_key.restype = C.c_uint32
    l.epk_shard_of_key.argtypes = [C.c_uint32, C.c_uint32]
    l.epk_partial_row_bytes.restype = C.c_uint64
    l.epk_partial_row_bytes.argtypes = [C.c_void_p]
    l.epk_place_partial_device.restype = C.c_int
    l.epk_place_partial_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p,
                                           C.c_void_p, C.c_void_p]
    l.epk_place_finalize_device.restype = C.c_int
    l.epk_place_finalize_device.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_void_p, C.c_uint32,
                                            C.POINTER(Result), C.c_void_p]
    l.epk_gather_ceiling_device.restype = C.c_int
    l.epk_gather_ceiling_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]
    l.epk_device_alloc.restype = C.c_int
    l.epk_device_alloc.argtypes = [C.c_int, C.c_uint64, C.POINTER(C.c_void_p)]
    l.epk_device_free.restype = None
    l.epk_device_free.argtypes = [C.c_int, C.c_void_p]
    l.epk_ipc_export.restype = C.c_int
    l.epk_ipc_export.argtypes = [C.c_void_p, C.c_char_p]
    l.epk_ipc_import.restype = C.c_int
    l.epk_ipc_import.argtypes = [C.c_int, C.c_char_p, C.POINTER(C.c_void_p)]
    l.epk_ipc_close.restype = None
    l.epk_ipc_close.argtypes = [C.c_int, C.c_void_p]
    l.epk_placer_set_trace.restype = C.c_int
    l.epk_placer_set_trace.argtypes = [C.c_void_p, C.c_void_p]
    l.epk_trace_term.restype = C.c_uint32
    l.epk_trace_term.argtypes = [C.c_uint32] * 4
    l.epk_gather_microbench.restype = C.c_int
    l.epk_gather_microbench.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.POINTER(Microbench)]
    l.epk_db_create_synthetic.restype = C.c_int
    l.epk_db_create_synthetic.argtypes = [C.POINTER(SynthDesc), C.POINTER(LoadFilter), C.c_int, C.c_uint32, C.c_uint32,
                                          C.POINTER(C.c_void_p)]
    l.epk_db_pool.restype = C.c_int
    l.epk_db_pool.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
    l.epk_db_pool_share.restype = C.c_int
    l.epk_db_pool_share.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_uint64), C.c_char_p]
    l.epk_db_pool_join.restype = C.c_int
    l.epk_db_pool_join.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_char_p, C.POINTER(C.c_void_p)]
    l.epk_db_get_score_range.restype = C.c_int
    l.epk_db_get_score_range.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_uint32),
                                         C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    l.epk_db_pool_set_score_range.restype = C.c_int
    l.epk_db_pool_set_score_range.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_uint32, C.c_uint64, C.c_uint64]
    l.epk_db_get_view.restype = C.c_int
    l.epk_db_get_view.argtypes = [C.c_void_p, C.POINTER(DbView)]
    l.epk_placer_set_shard_views.restype = C.c_int
    l.epk_placer_set_shard_views.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
    l.epk_place_finalize_seq_device.restype = C.c_int
    l.epk_place_finalize_seq_device.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_void_p, C.c_void_p,
                                                C.c_uint32, C.POINTER(Result), C.c_void_p]
    l.epk_host_alloc.restype = C.c_void_p
    l.epk_host_alloc.argtypes = [C.c_size_t]
    l.epk_host_free.argtypes = [C.c_void_p]
    l.epk_host_free.restype = None
    _lib = l
    return l


def check(status: int) -> None:
    if status != EPK_OK:
        raise EpkError(status, lib().epk_last_error().decode(errors="replace"))

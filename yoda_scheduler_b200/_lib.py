"""ctypes binding of libyoda_b200.so -- one Python function per symbol of include/yoda_b200.h."""
from __future__ import annotations

import ctypes as C
import os

from ._build import SO, build, stale

_lib = None


class YodaError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"yoda_b200 error {code}: {msg}")
        self.code = code


class Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("max_nodes", C.c_uint32), ("node_base", C.c_uint32), ("shard_rank", C.c_uint32),
                ("shard_count", C.c_uint32), ("flags", C.c_uint32)]


class Timing(C.Structure):
    _fields_ = [("total_ms", C.c_float), ("prep_ms", C.c_float), ("max_ms", C.c_float), ("score_ms", C.c_float),
                ("write_ms", C.c_float), ("h2d_ms", C.c_float), ("d2h_ms", C.c_float), ("kernels", C.c_uint32),
                ("path", C.c_uint32), ("n_classes", C.c_uint32), ("n_groups", C.c_uint32)]


class TailConfig(C.Structure):
    _fields_ = [("weight", C.c_uint32), ("percentage", C.c_int32), ("seed", C.c_uint64), ("start_index", C.c_uint32), ("flags", C.c_uint32),
                ("other_scores", C.c_void_p)]


class Xchg(C.Structure):
    _fields_ = [("d_ptr", C.c_void_p), ("count", C.c_uint64), ("op", C.c_int32), ("reserved", C.c_int32)]


# every exported symbol of include/yoda_b200.h: name -> (restype, argtypes)
_P, _U32, _U64, _I = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int
SYMBOLS = {
    "yoda_create": (_I, [C.POINTER(Config), C.POINTER(_P)]),
    "yoda_destroy": (None, [_P]),
    "yoda_abi_version": (_I, []),
    "yoda_last_error": (C.c_char_p, []),
    "yoda_node_upsert": (_I, [_P, _U32, _P]),
    "yoda_node_upsert_many": (_I, [_P, _U32, _P, _P]),
    "yoda_node_set_alloc": (_I, [_P, _U32, _U64]),
    "yoda_node_set_alloc_many": (_I, [_P, _U32, _P, _P]),
    "yoda_node_remove": (_I, [_P, _U32]),
    "yoda_commit": (_I, [_P, C.POINTER(_U64)]),
    "yoda_node_count": (_I, [_P, C.POINTER(_U32)]),
    "yoda_eval_batch": (_I, [_P, _P, _U32, _U32, _P]),
    "yoda_eval_batch_device": (_I, [_P, _P, _U32, _U32, _P]),
    "yoda_fetch_row": (_I, [_P, _U32, _I, _P, C.c_size_t]),
    "yoda_fetch_rows": (_I, [_P, _U32, _U32, _P, _P, _P]),
    "yoda_sync": (_I, [_P]),
    "yoda_num_feasible_nodes_to_find": (_U32, [_U32, C.c_int32]),
    "yoda_framework_tail": (_I, [_P, C.POINTER(TailConfig), _P, C.POINTER(_U32)]),
    "yoda_last_timing": (_I, [_P, C.POINTER(Timing)]),
    "yoda_result_matrices": (_I, [_P, C.POINTER(_P), C.POINTER(_P), C.POINTER(_P), C.POINTER(_U64)]),
    "yoda_result_raw_bytes": (_I, [_P, C.POINTER(_U32)]),
    "yoda_debug_pod_order": (_I, [_P, _P, _P, _P, _U32, C.POINTER(_U32), C.POINTER(_U32)]),
    "yoda_normalize_scores": (_I, [_P, _P, _U32]),
    "yoda_eval_sequential": (_I, [_P, _P, _U32, _P, _P]),
    "yoda_queue_order": (_I, [_P, _P, _U32, _P]),
    "yoda_normalize_scores_noalloc": (_I, [_P, _P, _U32, _P]),
    "yoda_record_from_scv_json": (_I, [C.c_char_p, C.c_size_t, _P, C.c_char_p, C.c_size_t]),
    "yoda_parse_label": (_U64, [C.c_char_p, C.c_size_t]),
    "yoda_uint64_to_int64": (C.c_int64, [_U64]),
    "yoda_comm_unique_id": (_I, [_P]),
    "yoda_comm_init": (_I, [_P, _P, _U32, _U32]),
    "yoda_phase_begin": (_I, [_P, _P, _U32, _U32]),
    "yoda_phase_run": (_I, [_P, _I]),
    "yoda_phase_exchange": (_I, [_P, _I, C.POINTER(Xchg), C.POINTER(_I)]),
    "yoda_phase_finish": (_I, [_P, _P]),
    "yoda_set_stream": (_I, [_P, _P]),
}


def load(rebuild_if_stale: bool = True):
    """Load the CUDA library.  Raises if it is missing and cannot be built: no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if rebuild_if_stale and stale():
        build()
    if not os.path.exists(SO):
        raise YodaError(-2, f"{SO} is missing; run __graft_entry__.build()")
    L = C.CDLL(SO)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(L, name)  # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def check(rc: int):
    if rc != 0:
        raise YodaError(rc, load().yoda_last_error().decode("utf-8", "replace"))

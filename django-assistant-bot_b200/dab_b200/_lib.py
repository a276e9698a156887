"""ctypes binding of libdab_b200.so - the C-ABI declared in include/dab_b200.h.

The library is hand-written CUDA for sm_100a.  There is no CPU fallback and no other
backend: if the shared object is missing, or a call fails, this module raises.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_int, c_int64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdab_b200.so")

DAB_F32, DAB_BF16 = 0, 1
DAB_MASK_U8, DAB_MASK_I32, DAB_MASK_I64 = 0, 1, 2
DAB_OK, DAB_E_INVALID, DAB_E_CUDA, DAB_E_DIM, DAB_E_CAPACITY, DAB_E_INCOMPLETE, DAB_E_INEXACT = range(7)
DAB_MAX_K = 2048
DAB_MAX_QUERY_BATCH = 256

# every symbol include/dab_b200.h declares: (restype, argtypes)
SIGNATURES = {
    "dab_version": (c_char_p, []),
    "dab_last_error": (c_char_p, []),
    "dab_pool_l2norm": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int64, c_int64,
                                c_int64, c_void_p]),
    "dab_store_create": (c_int, [POINTER(c_void_p), c_int, c_int64, c_int, c_int]),
    "dab_store_destroy": (None, [c_void_p]),
    "dab_store_size": (c_int64, [c_void_p]),
    "dab_store_capacity": (c_int64, [c_void_p]),
    "dab_store_dim": (c_int, [c_void_p]),
    "dab_store_dtype": (c_int, [c_void_p]),
    "dab_store_reset": (c_int, [c_void_p]),
    "dab_store_append": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int64, c_void_p]),
    "dab_store_ingest": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_int64, c_int64, c_void_p]),
    "dab_store_delete_rows": (c_int, [c_void_p, c_void_p, c_int64, c_void_p]),
    "dab_store_compact": (c_int, [c_void_p, POINTER(c_int64), c_void_p]),
    "dab_store_export": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p]),
    "dab_store_set_id_base": (c_int, [c_void_p, c_int64]),
    "dab_store_search": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "dab_store_search_ex": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "dab_store_search_host": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "dab_store_search_status": (c_int, [c_void_p, c_void_p]),
    "dab_store_set_spare_rounds": (c_int, [c_void_p, c_int]),
    "dab_store_set_path": (c_int, [c_void_p, c_int]),
    "dab_store_set_column": (c_int, [c_void_p, c_int, c_void_p, c_int64, c_int64, c_void_p]),
    "dab_store_build_allow": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "dab_store_allow_from_rows": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p]),
    "dab_merge_topk": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "dab_doc_aggregate": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                                  c_void_p]),
    "dab_exchange_bytes": (c_int64, [c_int, c_int, c_int]),
    "dab_exchange_slot": (c_int, [c_void_p, c_int, c_int, c_int, ctypes.c_uint32, POINTER(c_void_p), POINTER(c_void_p)]),
    "dab_store_status_dev": (c_void_p, [c_void_p]),
    "dab_exchange_merge_topk": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, ctypes.c_uint32, c_void_p,
                                        c_void_p, c_void_p, c_void_p, c_void_p]),
    "dab_sharded_create": (c_int, [POINTER(c_void_p), POINTER(c_int), c_int, c_int64, c_int, c_int, c_int, c_int]),
    "dab_sharded_destroy": (None, [c_void_p]),
    "dab_sharded_num_shards": (c_int, [c_void_p]),
    "dab_sharded_shard": (c_void_p, [c_void_p, c_int, POINTER(c_int), POINTER(c_int64), POINTER(c_int64), POINTER(c_void_p)]),
    "dab_sharded_search_host": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "dab_search_noalloc": (c_int, [c_void_p, c_int, c_void_p, c_int64, c_int, c_int64, c_void_p, c_int, c_int, c_void_p,
                                   c_void_p, c_void_p, c_void_p]),
    "dab_store_describe_plan": (c_int, [c_void_p, c_int, c_int, POINTER(c_int), POINTER(c_int), POINTER(c_int),
                                        POINTER(c_int), POINTER(c_int)]),
    "dab_launch_count": (ctypes.c_uint64, []),
    "dab_store_set_profile": (c_int, [c_void_p, c_int]),
    "dab_store_get_profile": (c_int, [c_void_p, POINTER(ctypes.c_float), POINTER(c_int), POINTER(ctypes.c_float),
                                      POINTER(c_int)]),
}

DAB_MAX_COLUMNS, DAB_MAX_PREDICATES, DAB_COL_ID = 8, 8, -1
DAB_OP_EQ, DAB_OP_NE, DAB_OP_LT, DAB_OP_LE, DAB_OP_GT, DAB_OP_GE, DAB_OP_IN, DAB_OP_NOT_IN = range(8)


class Predicate(ctypes.Structure):
    """struct dab_predicate of include/dab_b200.h."""
    _fields_ = [("column", ctypes.c_int32), ("op", ctypes.c_int32), ("value", c_int64), ("set_dev", c_void_p),
                ("set_n", c_int64)]


_lib = None


class DabError(RuntimeError):
    """A libdab_b200 call failed.  Subclass of RuntimeError so the reference's handlers keep
    working: gpu_service/main.py:84-86 turns it into HTTP 500, Celery auto-retries it
    (assistant/processing/tasks.py:15-22)."""

    def __init__(self, code: int, message: str):
        super().__init__(f"libdab_b200 error {code}: {message}")
        self.code = code


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a).  There is no CPU fallback.")
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError here = header and library disagree
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc: int) -> None:
    if rc != DAB_OK:
        raise DabError(rc, lib().dab_last_error().decode("utf-8", "replace"))

"""ctypes binding of ``oracle/scan_ref.c`` (TEST INFRASTRUCTURE ONLY, see ``oracle/__init__.py``)."""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libdab_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile scan_ref.c with the committed Makefile (gcc only)."""
    src = os.path.join(_HERE, "scan_ref.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "-s"], check=True)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = ctypes.CDLL(_LIB_PATH)
        for sfx in ("_strict", "_vec"):
            f = getattr(L, "dab_oracle_distances" + sfx)
            f.restype = None
            f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_int, ctypes.c_void_p,
                          ctypes.c_void_p, ctypes.c_int]
            f = getattr(L, "dab_oracle_search" + sfx)
            f.restype = ctypes.c_int64
            f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_int, ctypes.c_void_p,
                          ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                          ctypes.c_void_p, ctypes.c_void_p]
            f = getattr(L, "dab_oracle_pool" + sfx)
            f.restype = None
            f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64,
                          ctypes.c_void_p, ctypes.c_int]
        _lib = L
    return _lib


def _rows(rows):
    rows = np.ascontiguousarray(rows)
    if rows.dtype == np.float32:
        return rows, 0
    if rows.dtype == np.uint16:  # bf16 bit patterns
        return rows, 1
    raise TypeError("rows must be float32 or uint16 (bf16 bits)")


def distances(rows, q, strict: bool = True, nthreads: int = 0) -> np.ndarray:
    rows, dt = _rows(rows)
    q = np.ascontiguousarray(q, dtype=np.float32)
    n, dim = rows.shape
    assert q.shape == (dim,)
    out = np.empty(n, dtype=np.float64)
    f = getattr(lib(), "dab_oracle_distances" + ("_strict" if strict else "_vec"))
    f(rows.ctypes.data, dt, n, dim, q.ctypes.data, out.ctypes.data, nthreads)
    return out


def search(rows, queries, k: int, allow=None, strict: bool = False, nthreads: int = 0):
    """(ids [Q,k] int64 padded with -1, distances [Q,k] float64 padded with NaN, counts [Q])."""
    rows, dt = _rows(rows)
    queries = np.ascontiguousarray(np.atleast_2d(queries), dtype=np.float32)
    n, dim = rows.shape
    nq = queries.shape[0]
    assert queries.shape[1] == dim
    ids = np.empty((nq, k), dtype=np.int64)
    dist = np.empty((nq, k), dtype=np.float64)
    cnt = np.zeros(nq, dtype=np.int64)
    a = None
    if allow is not None:
        a = np.ascontiguousarray(allow, dtype=np.uint8)
        assert a.shape == (n,)
    f = getattr(lib(), "dab_oracle_search" + ("_strict" if strict else "_vec"))
    f(rows.ctypes.data, dt, n, dim, queries.ctypes.data, nq, k, a.ctypes.data if a is not None else None,
      nthreads, ids.ctypes.data, dist.ctypes.data, cnt.ctypes.data)
    return ids, dist, cnt


def pool(hidden, mask, strict: bool = False, nthreads: int = 0) -> np.ndarray:
    hidden = np.ascontiguousarray(hidden, dtype=np.float32)
    mask = np.ascontiguousarray(mask, dtype=np.int64)
    B, T, D = hidden.shape
    out = np.empty((B, D), dtype=np.float32)
    f = getattr(lib(), "dab_oracle_pool" + ("_strict" if strict else "_vec"))
    f(hidden.ctypes.data, mask.ctypes.data, B, T, D, out.ctypes.data, nthreads)
    return out

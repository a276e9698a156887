"""CPU-side checks of the C-ABI: the library loads and exports every symbol include/dab_b200.h
declares.  No compute call is made (no GPU here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "dab_b200.h")).read()
    return sorted(set(re.findall(r"DAB_API[^;(]*?\b(dab_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_something():
    syms = _declared_symbols()
    assert "dab_store_search" in syms and "dab_pool_l2norm" in syms and len(syms) >= 15


def test_library_exports_every_declared_symbol():
    from dab_b200 import _lib
    assert os.path.exists(_lib.LIB_PATH), "build with __graft_entry__.build()"
    L = ctypes.CDLL(_lib.LIB_PATH)
    for s in _declared_symbols():
        assert hasattr(L, s), f"{s} declared in include/dab_b200.h but not exported"
    assert sorted(_lib.SIGNATURES) == _declared_symbols(), "ctypes table and header disagree"


def test_version_and_error_string():
    from dab_b200 import _lib
    L = _lib.lib()
    assert b"sm_100a" in L.dab_version()
    assert isinstance(L.dab_last_error(), bytes)


def test_argument_errors_without_gpu():
    """Validation happens before any CUDA call, so these run on the CPU box too."""
    from dab_b200 import _lib
    L = _lib.lib()
    h = ctypes.c_void_p()
    assert L.dab_store_create(ctypes.byref(h), 0, 10, 770, _lib.DAB_BF16) == _lib.DAB_E_INVALID
    assert b"multiple" in L.dab_last_error()
    assert L.dab_store_create(ctypes.byref(h), 0, 0, 768, _lib.DAB_F32) == _lib.DAB_E_INVALID
    assert L.dab_store_create(ctypes.byref(h), 0, 10, 768, 7) == _lib.DAB_E_INVALID
    assert L.dab_pool_l2norm(None, 0, None, 0, None, None, None, 1, 1, 4, None) == _lib.DAB_E_INVALID
    assert L.dab_merge_topk(None, None, 1, 1, 1, None, None, None) == _lib.DAB_E_INVALID
    assert L.dab_store_search(None, None, 1, 1, None, None, None, None) == _lib.DAB_E_INVALID


def test_ops_refuse_cpu_tensors():
    import torch
    from dab_b200 import DabError, ops
    with pytest.raises(DabError, match="CUDA tensor"):
        ops.pool_l2norm(torch.zeros(1, 2, 4))
    if not torch.cuda.is_available():
        with pytest.raises(DabError, match="no CPU fallback"):
            ops.ChunkStore(10, 8)

"""dab_search_noalloc - the entry point for callers that cannot allocate GPU memory.

Some users of this library rank embeddings from inside a process whose GPU memory is already
spoken for (the embedding service, after the encoder's allocator has reserved the device), or
from a request handler where cudaMalloc's device-wide synchronisation is unaffordable.  They hand
in their own row matrix, inverse norms and output buffers; the search scratch is a static device
array inside the library.  This file checks, through the C-ABI, that such a call returns what the
ordinary store-based search returns on the same rows, and what the CPU oracle says - twice in a row, because every call reuses
the same static scratch and must not be confused by what the previous call left there.

Named test_zz_* so pytest collects it last: a run that stops at the first failure has then
already run every other test.  It passes alone in a fresh process and after the other tests.
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _noalloc(rows, inv_norm, queries, k, allow=None, id_base=0):
    """Calls dab_search_noalloc with raw device pointers; the outputs are ordinary tensors the
    test allocates (the caller, not the library, owns them)."""
    from dab_b200 import _lib
    L = _lib.lib()
    Q = queries.shape[0]
    scores = torch.empty(Q, k, dtype=torch.float32, device=DEV)
    ids = torch.empty(Q, k, dtype=torch.int64, device=DEV)
    dt = _lib.DAB_BF16 if rows.dtype == torch.bfloat16 else _lib.DAB_F32
    rc = L.dab_search_noalloc(rows.data_ptr(), dt, inv_norm.data_ptr(), rows.shape[0], rows.shape[1], id_base,
                              queries.data_ptr(), Q, k, None if allow is None else allow.data_ptr(),
                              scores.data_ptr(), ids.data_ptr(), torch.cuda.current_stream().cuda_stream)
    return rc, scores, ids


def _same_ranking(ids_a, sc_a, ids_b, sc_b):
    """The tolerance of the other search tests: ids equal except where the two scores differ by
    less than 1e-6 (near-ties may swap); scores within 1e-5 relative."""
    ids_a, ids_b = ids_a.cpu().numpy(), ids_b.cpu().numpy()
    sc_a, sc_b = sc_a.cpu().numpy().astype(np.float64), sc_b.cpu().numpy().astype(np.float64)
    np.testing.assert_allclose(sc_a, sc_b, rtol=1e-5, atol=1e-6)
    for q in range(ids_a.shape[0]):
        for j in range(ids_a.shape[1]):
            if ids_a[q, j] != ids_b[q, j]:
                assert abs(sc_a[q, j] - sc_b[q, j]) < 1e-6, (q, j, ids_a[q, j], ids_b[q, j])


def _matches_oracle(rows, queries, ids, scores, k, id_base, allow=None):
    """Pins the answer to the CPU oracle itself (float64 cosine of the STORED values, bf16 upcast exactly), so a
    fault shared by both entry points - they run the same kernels - cannot pass unnoticed."""
    from oracle import search_ref
    S = search_ref.exact_scores(rows.float().cpu().numpy(), queries.cpu().numpy())
    got_i, got_s = ids.cpu().numpy() - id_base, scores.cpu().numpy()
    for qi in range(queries.shape[0]):
        search_ref.check_topk_parity(S[qi], got_i[qi], got_s[qi], k, None if allow is None else allow.numpy())


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_noalloc_matches_store_search_twice_in_a_row(dtype):
    from dab_b200 import _lib, ops
    g = torch.Generator().manual_seed(2026)
    N, D = 60_000, 768
    rows = torch.randn(N, D, generator=g).to(dtype).to(DEV)
    rows[123] = 0                                        # zero row: inverse norm 0, never returned
    inv_norm = torch.zeros(N + 4, device=DEV)            # readable up to a multiple of 4 past n_rows
    nrm = rows.float().norm(dim=1)
    inv_norm[:N] = torch.where(nrm > 0, 1.0 / nrm, torch.zeros_like(nrm))
    store = ops.ChunkStore(N, D, dtype, DEV, id_base=1000)
    store.set_path("simt")
    store.append(rows)

    # first call: 3 queries, k = 10
    q1 = torch.randn(3, D, generator=g).to(DEV)
    rc, s1, i1 = _noalloc(rows, inv_norm, q1, 10, id_base=1000)
    assert rc == 0, _lib.lib().dab_last_error()
    ws1, wi1 = store.search(q1, 10)
    _same_ranking(i1, s1, wi1, ws1)
    _matches_oracle(rows, q1, i1, s1, 10, 1000)
    assert 1123 not in i1.cpu().numpy()

    # second call straight after, different batch size, k and a row filter: the static scratch
    # still holds the first call's candidates and must not leak into this answer
    q2 = torch.randn(8, D, generator=g).to(DEV)
    allow = torch.rand(N, generator=g) < 0.5
    bitmap = ops.pack_allow_bitmap(allow.to(DEV))
    rc, s2, i2 = _noalloc(rows, inv_norm, q2, 100, allow=bitmap, id_base=1000)
    assert rc == 0, _lib.lib().dab_last_error()
    ws2, wi2 = store.search(q2, 100, bitmap)
    _same_ranking(i2, s2, wi2, ws2)
    _matches_oracle(rows, q2, i2, s2, 100, 1000, allow)
    assert allow.numpy()[(i2.cpu().numpy() - 1000).ravel()].all()

    # and the first question again: same answer as before
    rc, s3, i3 = _noalloc(rows, inv_norm, q1, 10, id_base=1000)
    assert rc == 0
    assert torch.equal(i3, i1) and torch.equal(s3, s1)
    store.close()


def test_noalloc_rejects_sizes_beyond_its_static_scratch():
    from dab_b200 import _lib
    L = _lib.lib()
    rows = torch.randn(64, 2048, device=DEV)
    inv = torch.ones(64, device=DEV)
    rc, _, _ = _noalloc(rows, inv, torch.randn(1, 2048, device=DEV), 5)
    assert rc == _lib.DAB_E_CAPACITY and b"DAB_NOALLOC_MAX_DIM" in L.dab_last_error()
    rows = torch.randn(64, 64, device=DEV)
    rc, _, _ = _noalloc(rows, inv, torch.randn(9, 64, device=DEV), 5)
    assert rc == _lib.DAB_E_CAPACITY and b"DAB_NOALLOC_MAX_QUERIES" in L.dab_last_error()


def test_noalloc_allocates_nothing():
    """Device memory in use is the same before and after a warmed-up call."""
    rows = torch.randn(20_000, 256, device=DEV)
    inv = torch.zeros(20_000, device=DEV)
    inv[:] = 1.0 / rows.norm(dim=1)
    q = torch.randn(2, 256, device=DEV)
    rc, s, i = _noalloc(rows, inv, q, 7)          # first call on the device: resolves the symbol address
    assert rc == 0
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    out_s = torch.empty_like(s)
    out_i = torch.empty_like(i)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()          # torch's caching allocator serves these from its pool
    from dab_b200 import _lib
    rc = _lib.lib().dab_search_noalloc(rows.data_ptr(), _lib.DAB_F32, inv.data_ptr(), 20_000, 256, 0, q.data_ptr(), 2, 7,
                                       None, out_s.data_ptr(), out_i.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    free2, _ = torch.cuda.mem_get_info()
    assert free2 == free1, "dab_search_noalloc changed the amount of free device memory"
    assert torch.equal(out_i, i) and torch.equal(out_s, s)

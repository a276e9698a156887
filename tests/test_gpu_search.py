"""Parity of the device store's exact cosine top-k with the oracle (through the C-ABI).
Bar (BASELINE.json north_star): ids bit-exact under (score desc, row asc) except where adjacent
oracle scores differ by < 1e-6; scores within 1e-5 relative."""
import os

import numpy as np
import pytest
import torch

from oracle import c_oracle, search_ref

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _stored_f32(rows: torch.Tensor, dtype) -> np.ndarray:
    return rows.to(dtype).float().numpy()


def _check(store, rows_stored, queries, k, allow=None, bitmap=None):
    scores, ids = store.search(queries.to(DEV), k, bitmap)
    S = search_ref.exact_scores(rows_stored, queries.numpy())
    ids, scores = ids.cpu().numpy(), scores.cpu().numpy()
    for qi in range(queries.shape[0]):
        search_ref.check_topk_parity(S[qi], ids[qi], scores[qi], k, allow=allow)
        n_valid = (ids[qi] >= 0).sum()
        assert np.all(np.diff(scores[qi][:n_valid]) <= 0)
    return scores, ids


def test_search_matches_reference_golden(golden_dir):
    """Device ranking vs the reference's own NumPy cosine (search_service.py:17-21,32-35) on the
    committed golden vectors, incl. duplicate rows, a scaled copy, identical and opposite queries."""
    from dab_b200 import ops
    g = np.load(os.path.join(golden_dir, "search_ref.npz"))
    rows, queries, sims = g["rows"], g["queries"], g["sims"]
    store = ops.ChunkStore(rows.shape[0], rows.shape[1], torch.float32, DEV)
    store.append(torch.from_numpy(rows).to(DEV))
    scores, ids = store.search(torch.from_numpy(queries).to(DEV), 10)
    for qi in range(queries.shape[0]):
        search_ref.check_topk_parity(sims[qi], ids[qi].cpu().numpy(), scores[qi].cpu().numpy(), 10)
    assert ids[1, 0].item() == 20 and abs(scores[1, 0].item() - 1.0) < 1e-6   # identical vector: distance 0
    s3, i3 = store.search(torch.from_numpy(rows[3]).to(DEV), 3)
    assert i3[0, :2].tolist() == [3, 10]                                      # duplicates: lower row first
    store.close()


@pytest.mark.parametrize("path", ["simt", "tcgen05"])
def test_bf16_store_matches_reference_golden_top50(golden_dir, path):
    """A bf16 store against numbers the REFERENCE computed (embeddings_similarity on the same bf16-valued rows,
    tests/golden/search_ref_bf16.npz): n = 50, a triplicated row, a zero row, an un-normalised query."""
    from dab_b200 import ops
    from oracle import make_golden
    g = np.load(os.path.join(golden_dir, "search_ref_bf16.npz"))
    rows, queries, sims = make_golden.golden_bf16_rows(), g["queries"], g["sims"]
    store = ops.ChunkStore(rows.shape[0], rows.shape[1], torch.bfloat16, DEV)
    store.set_path(path)
    store.append(torch.from_numpy(rows).to(DEV))          # bf16-representable values: the rounding is exact
    scores, ids = store.search(torch.from_numpy(queries).to(DEV), 50)
    for qi in range(queries.shape[0]):
        search_ref.check_topk_parity(sims[qi], ids[qi].cpu().numpy(), scores[qi].cpu().numpy(), 50)
    assert ids[0, :3].tolist() == [9, 150, 151]
    assert ids[1, 0].item() == 33 and abs(scores[1, 0].item() - 1.0) < 1e-6
    assert 200 not in ids.cpu().numpy()
    store.close()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("N,D,Q,k", [(10_000, 384, 1, 10), (50_000, 768, 1, 10), (50_000, 768, 4, 100),
                                     (30_011, 1024, 3, 5), (200_000, 768, 2, 250), (777, 64, 7, 1), (5, 128, 2, 10),
                                     (40_000, 2048, 2, 10)])
def test_search_parity_simt(dtype, N, D, Q, k):
    from dab_b200 import ops
    g = torch.Generator().manual_seed(N + D + Q + k)
    rows = torch.randn(N, D, generator=g)
    queries = torch.randn(Q, D, generator=g)
    store = ops.ChunkStore(N, D, dtype, DEV)
    store.set_path("simt")
    half = N // 2
    store.append(rows[:half].to(DEV))          # two appends: fp32 rows are rounded into a bf16 store
    store.append(rows[half:].to(DEV))
    assert len(store) == N
    _check(store, _stored_f32(rows, dtype), queries, k)
    store.close()


def test_search_duplicates_zero_rows_and_filter():
    from dab_b200 import ops
    g = torch.Generator().manual_seed(5)
    N, D, k = 20_000, 768, 20
    rows = torch.randn(N, D, generator=g)
    rows[100] = 0                                   # zero row: NaN distance in pgvector, never returned
    dup = torch.randint(0, N // 2, (64,), generator=g)
    rows[N // 2 + torch.arange(64)] = rows[dup]     # duplicates at higher indices
    q = rows[dup[:3]].clone()                       # queries equal to duplicated rows -> exact ties at the top
    store = ops.ChunkStore(N, D, torch.float32, DEV)
    store.append(rows.to(DEV))
    scores, ids = _check(store, rows.numpy(), q, k)
    for i in range(3):
        assert ids[i, 0] == dup[i].item() and 100 not in ids[i]
    allow = torch.rand(N, generator=g) < 0.3
    bitmap = ops.pack_allow_bitmap(allow.to(DEV))
    _check(store, rows.numpy(), q, k, allow=allow.numpy(), bitmap=bitmap)
    # filter leaving fewer than k rows: returns fewer, padded with id -1 / -inf
    few = torch.zeros(N, dtype=torch.bool)
    few[[7, 9, 100]] = True
    s, i = store.search(q[:1].to(DEV), k, ops.pack_allow_bitmap(few.to(DEV)))
    assert sorted(i[0, :2].tolist()) == [7, 9] and (i[0, 2:] == -1).all() and torch.isinf(s[0, 2:]).all()
    store.close()


def test_search_sorted_corpus():
    """Rows sorted by ascending similarity to the query: the strided level sampling still yields
    a representative threshold, so no round overflows."""
    from dab_b200 import _lib, ops
    g = torch.Generator().manual_seed(11)
    N, D, k = 300_000, 64, 10
    rows = torch.randn(N, D, generator=g)
    q = torch.randn(1, D, generator=g)
    rows = rows[torch.argsort((rows @ q[0]) / rows.norm(dim=1))].contiguous()
    store = ops.ChunkStore(N, D, torch.float32, DEV)
    store.append(rows.to(DEV))
    store.search_async(q.to(DEV), k)
    assert _lib.lib().dab_store_search_status(store._h, torch.cuda.current_stream().cuda_stream) == _lib.DAB_OK
    _check(store, rows.numpy(), q, k)
    store.close()


def test_search_clustered_corpus_overflows_and_recovers():
    """~29k near-copies of the query placed only in tiles that the sampled levels never visit (the
    plan is asked from the library): the last level admits far more than the 8192-slot candidate
    buffer.  The unchecked call must report DAB_E_INCOMPLETE with poisoned outputs; the checked
    search must retry on its own and return the exact answer."""
    from dab_b200 import _lib, ops
    g = torch.Generator().manual_seed(12)
    N, D, k = 300_000, 64, 10
    rows = torch.randn(N, D, generator=g)
    q = torch.randn(1, D, generator=g)
    store = ops.ChunkStore(N, D, torch.float32, DEV)
    store.append(rows.to(DEV))
    store.set_path("simt")          # the CUDA-core path reports the overflow as such; auto is covered below
    plan = store.describe_plan(1, k)
    assert plan["path"] == "simt" and plan["num_levels"] >= 2
    r = torch.arange(100_000, 130_000)
    r = r[(r // plan["tile_rows"]) % plan["strides"][-2] != 0]        # rows only the last level visits
    rows[r] = q + 0.05 * torch.randn(r.numel(), D, generator=g)
    store.reset()
    store.append(rows.to(DEV))
    s, i = store.search_async(q.to(DEV), k)
    rc = _lib.lib().dab_store_search_status(store._h, torch.cuda.current_stream().cuda_stream)
    assert rc == _lib.DAB_E_INCOMPLETE and (i == -1).all() and torch.isnan(s).all()
    _check(store, rows.numpy(), q, k)
    # auto path (tensor cores for this layout): the raw call may say incomplete or inexact (29k rows sit inside
    # the rescoring margin), never OK with a wrong answer; the checked search must still be exact
    store.set_path("auto")
    assert store.describe_plan(1, k)["path"] == "tcgen05"
    s, i = store.search_async(q.to(DEV), k)
    rc = _lib.lib().dab_store_search_status(store._h, torch.cuda.current_stream().cuda_stream)
    assert rc in (_lib.DAB_E_INCOMPLETE, _lib.DAB_E_INEXACT), rc
    _check(store, rows.numpy(), q, k)
    # the inexact verdict is remembered: the next 64 few-query searches go straight to the CUDA cores
    # (no wasted tensor attempt on a corpus this dense), then the tensor path is probed again
    assert store.describe_plan(1, k)["path"] == "simt" and store.describe_plan(64, k)["path"] == "tcgen05"
    for _ in range(64):
        store.search_async(q.to(DEV), k)
    assert store.describe_plan(1, k)["path"] == "tcgen05"
    store.close()


def test_search_vs_c_oracle_pgvector_semantics():
    """End to end against oracle/scan_ref.c (pgvector <=> restatement): distance = 1 - score."""
    from dab_b200 import ops
    g = torch.Generator().manual_seed(3)
    N, D, Q, k = 100_000, 768, 2, 10
    rows = torch.randn(N, D, generator=g)
    q = torch.randn(Q, D, generator=g)
    store = ops.ChunkStore(N, D, torch.float32, DEV)
    store.append(rows.to(DEV))
    scores, ids = store.search(q.to(DEV), k)
    oid, odist, _ = c_oracle.search(rows.numpy(), q.numpy(), k)
    S = search_ref.exact_scores(rows.numpy(), q.numpy())
    for qi in range(Q):
        search_ref.check_topk_parity(S[qi], ids[qi].cpu().numpy(), scores[qi].cpu().numpy(), k)
        np.testing.assert_allclose(1.0 - scores[qi].cpu().numpy(), odist[qi], atol=2e-6)
    store.close()


def test_store_errors():
    from dab_b200 import DabError, ops
    store = ops.ChunkStore(10, 16, torch.float32, DEV)
    with pytest.raises(DabError, match="different vector dimensions"):
        store.append(torch.zeros(2, 8, device=DEV))
    with pytest.raises(DabError, match="full"):
        store.append(torch.zeros(11, 16, device=DEV))
    s, i = store.search(torch.ones(1, 16, device=DEV), 3)      # empty store
    assert (i == -1).all()
    with pytest.raises(DabError, match="different vector dimensions"):
        store.search(torch.ones(1, 8, device=DEV), 3)
    with pytest.raises(DabError, match="k must be"):
        store.search(torch.ones(1, 16, device=DEV), 0)
    store.close()


def test_merge_topk():
    from dab_b200 import ops
    g = torch.Generator().manual_seed(9)
    R, Q, k = 8, 5, 100
    scores = torch.rand(R, Q, k, generator=g).sort(dim=2, descending=True).values
    ids = torch.randperm(R * Q * k, generator=g).view(R, Q, k)
    scores[3, :, 50:] = float("-inf")
    ids[3, :, 50:] = -1                                  # a shard that returned fewer rows
    scores[1, 0, 0] = scores[2, 0, 0]                    # cross-shard tie -> lower id first
    ms, mi = ops.merge_topk(scores.to(DEV), ids.to(DEV))
    for qi in range(Q):
        flat_s, flat_i = scores[:, qi].flatten().numpy(), ids[:, qi].flatten().numpy()
        keep = flat_i >= 0
        order = np.lexsort((flat_i[keep], -flat_s[keep]))[:k]
        np.testing.assert_array_equal(mi[qi].cpu().numpy(), flat_i[keep][order])
        np.testing.assert_array_equal(ms[qi].cpu().numpy(), flat_s[keep][order])

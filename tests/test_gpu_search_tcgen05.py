"""Parity of the tcgen05 (bf16 tensor-core) scoring path + exact fp32 rescoring with the oracle.
The tensor pass ranks with bf16-rounded queries; the final ids/scores must still meet the exact bar."""
import numpy as np
import pytest
import torch

from oracle import search_ref

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _run(N, D, Q, k, seed, path="tcgen05", planted=False):
    from dab_b200 import ops
    g = torch.Generator().manual_seed(seed)
    rows = torch.randn(N, D, generator=g)
    queries = torch.randn(Q, D, generator=g)
    if planted:
        # near-duplicates of the queries: a dense cluster of scores inside the rescoring margin
        for qi in range(min(Q, 4)):
            idx = torch.randint(0, N, (300,), generator=g)
            rows[idx] = queries[qi] + 0.02 * torch.randn(300, D, generator=g)
        rows[7] = 0
    store = ops.ChunkStore(N, D, torch.bfloat16, DEV)
    store.set_path(path)
    store.append(rows.to(DEV))
    stored = rows.to(torch.bfloat16).float().numpy()
    scores, ids = store.search(queries.to(DEV), k)
    S = search_ref.exact_scores(stored, queries.numpy())
    ids, scores = ids.cpu().numpy(), scores.cpu().numpy()
    for qi in range(Q):
        search_ref.check_topk_parity(S[qi], ids[qi], scores[qi], k)
    store.close()
    return ids, scores


@pytest.mark.parametrize("N,D,Q,k", [(20_000, 64, 5, 10), (50_000, 768, 16, 10), (100_000, 768, 64, 100),
                                     (60_001, 1024, 64, 100), (30_000, 384, 33, 5), (200_000, 768, 80, 10),
                                     (300, 128, 8, 10), (40_000, 768, 200, 20)])
def test_tcgen05_parity(N, D, Q, k):
    _run(N, D, Q, k, seed=N + D + Q + k)


@pytest.mark.parametrize("N,D,Q,k", [(100_000, 768, 256, 100), (50_000, 1024, 256, 10), (20_000, 768, 100, 10),
                                     (300, 768, 256, 5), (70_001, 768, 129, 50), (30_000, 384, 255, 10)])
def test_tcgen05_wide_query_blocks_cluster_kernel(N, D, Q, k):
    """More queries than fit one SM's shared memory: the 4-CTA cluster kernel (cta_group::2 pairs,
    multicast corpus tiles) scores the whole block in one corpus pass."""
    _run(N, D, Q, k, seed=N + D + Q + k)


def test_tcgen05_dense_score_cluster_and_zero_row():
    _run(80_000, 768, 8, 50, seed=77, planted=True)


def test_tcgen05_inexact_falls_back_to_exact_path():
    """5000 near-copies of a query all sit inside the bf16 rescoring margin of the k-th best: more
    than the 2048 candidates the tensor pass may keep.  The raw call must say DAB_E_INEXACT; the
    checked search must answer on the CUDA-core path and still be exact."""
    from dab_b200 import _lib, ops
    g = torch.Generator().manual_seed(21)
    N, D, Q, k = 60_000, 768, 8, 10
    rows = torch.randn(N, D, generator=g)
    queries = torch.randn(Q, D, generator=g)
    idx = torch.randperm(N, generator=g)[:5000]
    rows[idx] = queries[0] + 0.001 * torch.randn(5000, D, generator=g)
    store = ops.ChunkStore(N, D, torch.bfloat16, DEV)
    store.set_path("tcgen05")
    store.set_spare_rounds(8)
    store.append(rows.to(DEV))
    raw_s, raw_i = store.search_async(queries.to(DEV), k)
    rc = _lib.lib().dab_store_search_status(store._h, torch.cuda.current_stream().cuda_stream)
    assert rc == _lib.DAB_E_INEXACT, _lib.lib().dab_last_error()
    S = search_ref.exact_scores(rows.to(torch.bfloat16).float().numpy(), queries.numpy())
    # only the query whose margin overflowed is poisoned; the other seven answers of the raw call are final
    assert torch.isnan(raw_s[0]).all() and (raw_i[0] == -1).all()
    for qi in range(1, Q):
        search_ref.check_topk_parity(S[qi], raw_i[qi].cpu().numpy(), raw_s[qi].cpu().numpy(), k)
    store.set_spare_rounds(0)
    scores, ids = store.search(queries.to(DEV), k)           # checked: redoes query 0 alone on the CUDA cores
    for qi in range(Q):
        search_ref.check_topk_parity(S[qi], ids[qi].cpu().numpy(), scores[qi].cpu().numpy(), k)
    hs, hi = store.search_host(queries.numpy(), k)            # the C host entry splices the same way
    np.testing.assert_array_equal(hi, ids.cpu().numpy())
    np.testing.assert_array_equal(hs, scores.cpu().numpy())
    store.close()


@pytest.mark.parametrize("Q,victim", [(200, 3), (300, 280)])
def test_inexact_query_in_a_wide_block_is_redone_alone(Q, victim):
    """Many queries, one of them with 5000 near-copies in the corpus: the checked searches keep the other tensor-path
    answers and repeat only the overflowing query on the CUDA cores (not 50 CUDA-core passes for the block).  With
    300 queries the victim sits in the second 256-query chunk of the call."""
    from dab_b200 import ops
    g = torch.Generator().manual_seed(33)
    N, D, k = 60_000, 768, 10
    rows = torch.randn(N, D, generator=g)
    queries = torch.randn(Q, D, generator=g)
    idx = torch.randperm(N, generator=g)[:5000]
    rows[idx] = queries[victim] + 0.001 * torch.randn(5000, D, generator=g)
    store = ops.ChunkStore(N, D, torch.bfloat16, DEV)
    store.append(rows.to(DEV))
    S = search_ref.exact_scores(rows.to(torch.bfloat16).float().numpy(), queries.numpy())
    from dab_b200 import _lib
    n0 = _lib.lib().dab_launch_count()
    scores, ids = store.search(queries.to(DEV), k)
    n1 = _lib.lib().dab_launch_count()
    hs, hi = store.search_host(queries.numpy(), k)
    for qi in range(Q):
        search_ref.check_topk_parity(S[qi], ids[qi].cpu().numpy(), scores[qi].cpu().numpy(), k)
    np.testing.assert_array_equal(hi, ids.cpu().numpy())
    assert n1 - n0 < 80, n1 - n0       # tensor search(es) + one single-query exact search, not 50+ passes of 4
    store.close()


def test_auto_path_picks_tensor_cores_for_batches():
    ids_auto, sc_auto = _run(50_000, 768, 32, 10, seed=5, path="auto")
    ids_simt, sc_simt = _run(50_000, 768, 32, 10, seed=5, path="simt")
    np.testing.assert_array_equal(ids_auto, ids_simt)
    np.testing.assert_allclose(sc_auto, sc_simt, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("dtype,N,D,Q,k", [(torch.float32, 30_000, 100, 3, 10), (torch.float32, 50_000, 300, 16, 10),
                                           (torch.bfloat16, 40_000, 104, 5, 10), (torch.bfloat16, 60_000, 1000, 40, 20),
                                           (torch.bfloat16, 50_000, 760, 200, 10), (torch.float32, 40_000, 700, 100, 5),
                                           (torch.bfloat16, 300, 8, 2, 5), (torch.float32, 5000, 4, 1, 3)])
def test_tensor_path_on_rows_that_are_not_whole_k_blocks(dtype, N, D, Q, k):
    """Row lengths that are not a multiple of 128 bytes (GloVe's 300, 1000, ...): the tensor maps carry the true
    row length and TMA zero-fills the tail of the last k-block of rows and queries alike."""
    from dab_b200 import ops
    g = torch.Generator().manual_seed(N + D + Q + k)
    rows = torch.randn(N, D, generator=g)
    queries = torch.randn(Q, D, generator=g)
    queries[0] = rows[9]
    store = ops.ChunkStore(N, D, dtype, DEV)
    store.append(rows.to(DEV))
    assert store.describe_plan(Q, k)["path"].startswith("tcgen05")
    scores, ids = store.search(queries.to(DEV), k)
    S = search_ref.exact_scores(rows.to(dtype).float().numpy(), queries.numpy())
    ids, scores = ids.cpu().numpy(), scores.cpu().numpy()
    for qi in range(Q):
        search_ref.check_topk_parity(S[qi], ids[qi], scores[qi], k)
    assert ids[0, 0] == 9
    store.set_path("simt")
    s2, i2 = store.search(queries.to(DEV), k)
    np.testing.assert_array_equal(i2.cpu().numpy(), ids)
    store.close()


@pytest.mark.parametrize("N,D,Q,k", [(50_000, 768, 16, 10), (120_000, 768, 48, 100), (30_000, 384, 33, 5),
                                     (40_000, 768, 100, 20), (20_000, 1024, 7, 250), (300, 32, 5, 10)])
def test_tf32_tensor_path_on_fp32_stores(N, D, Q, k):
    """fp32 stores (what pgvector holds) with more than 4 queries: kind::tf32 MMAs + the same margin
    + exact fp32 rescoring; the final ranking must meet the exact bar."""
    from dab_b200 import ops
    g = torch.Generator().manual_seed(N + D + Q + k)
    rows = torch.randn(N, D, generator=g)
    queries = torch.randn(Q, D, generator=g)
    rows[N // 3] = 0
    rows[N // 2] = rows[11]
    queries[0] = rows[11]
    store = ops.ChunkStore(N, D, torch.float32, DEV)
    store.append(rows.to(DEV))
    assert store.describe_plan(Q, k)["path"].startswith("tcgen05")   # auto: tensor cores (single-CTA or cluster kernel)
    scores, ids = store.search(queries.to(DEV), k)
    S = search_ref.exact_scores(rows.numpy(), queries.numpy())
    ids, scores = ids.cpu().numpy(), scores.cpu().numpy()
    for qi in range(Q):
        search_ref.check_topk_parity(S[qi], ids[qi], scores[qi], k)
    assert ids[0, 0] == 11 and ids[0, 1] == N // 2
    store.set_path("simt")
    s2, i2 = store.search(queries.to(DEV), k)
    np.testing.assert_array_equal(i2.cpu().numpy(), ids)
    store.close()


@pytest.mark.parametrize("N,D,Q,k", [(60_000, 768, 64, 10), (60_000, 768, 96, 20), (50_000, 768, 150, 10),
                                     (40_000, 768, 256, 10), (30_000, 384, 256, 5), (70_001, 1024, 100, 50)])
def test_tf32_cluster_kernel_on_fp32_stores(N, D, Q, k):
    """fp32 stores with more queries than one SM holds (48 at dim 768): the cluster kernel with kind::tf32 -
    one pair up to 96 queries, two pairs up to 192, two corpus passes of 128 for a 256-query block."""
    from dab_b200 import ops
    g = torch.Generator().manual_seed(N + D + Q + k)
    rows = torch.randn(N, D, generator=g)
    queries = torch.randn(Q, D, generator=g)
    rows[N // 3] = 0
    queries[Q - 1] = rows[17]
    store = ops.ChunkStore(N, D, torch.float32, DEV)
    store.append(rows.to(DEV))
    assert store.describe_plan(Q, k)["path"] == "tcgen05-cluster"
    scores, ids = store.search(queries.to(DEV), k)
    S = search_ref.exact_scores(rows.numpy(), queries.numpy())
    ids, scores = ids.cpu().numpy(), scores.cpu().numpy()
    for qi in range(Q):
        search_ref.check_topk_parity(S[qi], ids[qi], scores[qi], k)
    assert ids[Q - 1, 0] == 17
    store.close()


def test_threads_on_one_store_keep_their_own_verdicts():
    """Two host threads search the SAME store on their own streams, again and again: thread A's query has 5000
    near-copies in the corpus (its tensor pass ends DAB_E_INEXACT and the raw answer is poisoned), thread B's
    queries are ordinary.  Every search keeps its own verdict slot, so A always sees INEXACT and B always sees OK
    with final answers - whatever the interleaving (the asgiref worker next to another coroutine,
    assistant/bot/services/context_service/service.py:82-83); and the checked search repeats A's query on the exact
    path without changing the store for B."""
    import threading
    from dab_b200 import _lib, ops
    L = _lib.lib()
    g = torch.Generator().manual_seed(77)
    N, D, k = 60_000, 768, 10
    rows = torch.randn(N, D, generator=g)
    qa = torch.randn(1, D, generator=g)
    qb = torch.randn(6, D, generator=g)
    idx = torch.randperm(N, generator=g)[:5000]
    rows[idx] = qa[0] + 0.001 * torch.randn(5000, D, generator=g)
    store = ops.ChunkStore(N, D, torch.bfloat16, DEV)
    store.set_path("tcgen05")
    store.append(rows.to(DEV))
    stored = rows.to(torch.bfloat16).float().numpy()
    Sa = search_ref.exact_scores(stored, qa.numpy())
    Sb = search_ref.exact_scores(stored, qb.numpy())
    torch.cuda.synchronize()
    errors = []

    def worker(name, q, S, want_rc):
        try:
            stream = torch.cuda.Stream(device=DEV)
            qd = q.to(DEV)
            torch.cuda.synchronize()
            with torch.cuda.stream(stream):
                for it in range(25):
                    s, i = store.search_async(qd, k)
                    rc = L.dab_store_search_status(store._h, stream.cuda_stream)
                    assert rc == want_rc, (name, it, rc, L.dab_last_error())
                    if want_rc == _lib.DAB_OK:
                        for qi in range(q.shape[0]):
                            search_ref.check_topk_parity(S[qi], i[qi].cpu().numpy(), s[qi].cpu().numpy(), k)
                    else:
                        assert torch.isnan(s[0]).all()
                    if it % 8 == 0:                      # the checked search answers exactly, for both threads
                        cs, ci = store.search(qd, k)
                        for qi in range(q.shape[0]):
                            search_ref.check_topk_parity(S[qi], ci[qi].cpu().numpy(), cs[qi].cpu().numpy(), k)
        except Exception as e:  # noqa: BLE001
            errors.append((name, repr(e)))

    ta = threading.Thread(target=worker, args=("A", qa, Sa, _lib.DAB_E_INEXACT))
    tb = threading.Thread(target=worker, args=("B", qb, Sb, _lib.DAB_OK))
    ta.start(); tb.start(); ta.join(); tb.join()
    assert not errors, errors
    store.close()

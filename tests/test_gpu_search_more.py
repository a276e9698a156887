"""More parity cases for the device store: explicit ids, query chunks beyond 256, filters on the
tensor-core paths, the largest k, reset/reuse, 1-D queries."""
import numpy as np
import pytest
import torch

from oracle import search_ref

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _oracle_check(stored, queries, ids, scores, k, allow=None, row_of_id=None):
    S = search_ref.exact_scores(stored, queries)
    for qi in range(queries.shape[0]):
        got = ids[qi] if row_of_id is None else np.array([row_of_id.get(int(x), -1) for x in ids[qi]])
        search_ref.check_topk_parity(S[qi], got, scores[qi], k, allow=allow)


@pytest.mark.parametrize("dtype,path", [(torch.float32, "simt"), (torch.bfloat16, "tcgen05")])
def test_explicit_ids_are_returned(dtype, path):
    """Rows appended with Django pks: results carry the pks; ties still break by row order."""
    from dab_b200 import ops
    g = torch.Generator().manual_seed(1)
    N, D, Q, k = 30_000, 768, 6, 25
    rows = torch.randn(N, D, generator=g)
    pks = (torch.arange(N) * 3 + 1000)                       # ascending with insertion order
    rows[20_000] = rows[17]
    queries = torch.randn(Q, D, generator=g)
    queries[0] = rows[17]
    store = ops.ChunkStore(N + 8, D, dtype, DEV)
    store.set_path(path)
    store.append(rows[:10_000].to(DEV), pks[:10_000].to(DEV))
    store.append(rows[10_000:].to(DEV), pks[10_000:].to(DEV))
    scores, ids = store.search(queries.to(DEV), k)
    ids, scores = ids.cpu().numpy(), scores.cpu().numpy()
    row_of = {int(p): i for i, p in enumerate(pks.tolist())}
    _oracle_check(rows.to(dtype).float().numpy(), queries.numpy(), ids, scores, k, row_of_id=row_of)
    assert ids[0, 0] == 1000 + 3 * 17 and ids[0, 1] == 1000 + 3 * 20_000
    from dab_b200 import DabError
    with pytest.raises(DabError, match="explicit and implicit ids"):
        store.append(rows[:5].to(DEV))
    store.close()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_more_than_256_queries_are_chunked(dtype):
    from dab_b200 import ops
    g = torch.Generator().manual_seed(2)
    N, D, Q, k = 20_000, 128, 300, 7
    rows = torch.randn(N, D, generator=g)
    queries = torch.randn(Q, D, generator=g)
    store = ops.ChunkStore(N, D, dtype, DEV, id_base=5_000_000_000)          # ids beyond 32 bits
    store.append(rows.to(DEV))
    scores, ids = store.search(queries.to(DEV), k)
    _oracle_check(rows.to(dtype).float().numpy(), queries.numpy(), ids.cpu().numpy() - 5_000_000_000, scores.cpu().numpy(), k)
    store.close()


@pytest.mark.parametrize("Q", [16, 200])
def test_row_filter_on_tensor_core_paths(Q):
    from dab_b200 import ops
    g = torch.Generator().manual_seed(3 + Q)
    N, D, k = 60_000, 768, 30
    rows = torch.randn(N, D, generator=g)
    queries = torch.randn(Q, D, generator=g)
    allow = torch.rand(N, generator=g) < 0.2
    store = ops.ChunkStore(N, D, torch.bfloat16, DEV)
    store.set_path("tcgen05")
    store.append(rows.to(DEV))
    scores, ids = store.search(queries.to(DEV), k, ops.pack_allow_bitmap(allow.to(DEV)))
    _oracle_check(rows.to(torch.bfloat16).float().numpy(), queries.numpy(), ids.cpu().numpy(), scores.cpu().numpy(), k,
                  allow=allow.numpy())
    store.close()


@pytest.mark.parametrize("dtype,path", [(torch.float32, "simt"), (torch.bfloat16, "tcgen05")])
def test_largest_k_and_reset_reuse(dtype, path):
    from dab_b200 import _lib, ops
    g = torch.Generator().manual_seed(4)
    N, D = 50_000, 64
    rows = torch.randn(N, D, generator=g)
    q = torch.randn(2, D, generator=g)
    store = ops.ChunkStore(N, D, dtype, DEV)
    store.set_path(path)
    store.append(rows.to(DEV))
    k = _lib.DAB_MAX_K
    scores, ids = store.search(q.to(DEV), k)
    _oracle_check(rows.to(dtype).float().numpy(), q.numpy(), ids.cpu().numpy(), scores.cpu().numpy(), k)
    # reset and refill with other rows: nothing of the old contents may leak into the answers
    store.reset()
    assert len(store) == 0
    rows2 = torch.randn(1000, D, generator=g)
    store.append(rows2.to(DEV))
    s1, i1 = store.search(q[0].to(DEV), 5)                                    # 1-D query
    assert s1.shape == (1, 5)
    _oracle_check(rows2.to(dtype).float().numpy(), q[:1].numpy(), i1.cpu().numpy(), s1.cpu().numpy(), 5)
    store.close()


def test_randomised_shapes_against_oracle():
    """A seeded sweep over odd shapes (dims that are not multiples of 64, k near N, tiny N)."""
    from dab_b200 import ops
    rng = np.random.default_rng(2026)
    for trial in range(12):
        dtype = torch.bfloat16 if trial % 2 else torch.float32
        D = int(rng.choice([8, 24, 200, 392, 520, 1000] if dtype == torch.bfloat16 else [4, 12, 100, 388, 516, 1004]))
        N = int(rng.integers(1, 6000))
        Q = int(rng.integers(1, 40))
        k = int(rng.integers(1, min(N, 300) + 1))
        g = torch.Generator().manual_seed(trial)
        rows = torch.randn(N, D, generator=g)
        queries = torch.randn(Q, D, generator=g)
        store = ops.ChunkStore(N, D, dtype, DEV)
        store.append(rows.to(DEV))
        scores, ids = store.search(queries.to(DEV), k)
        _oracle_check(rows.to(dtype).float().numpy(), queries.numpy(), ids.cpu().numpy(), scores.cpu().numpy(), k)
        store.close()


def test_search_host_matches_device_search_and_retries():
    """dab_store_search_host: NumPy in/out through one C call, same answer as the tensor API,
    including the internal retry on an adversarial corpus."""
    from dab_b200 import DabError, ops
    g = torch.Generator().manual_seed(8)
    N, D, Q, k = 40_000, 768, 300, 12                      # > 256 queries: two chunks inside the C call
    rows = torch.randn(N, D, generator=g)
    queries = torch.randn(Q, D, generator=g)
    for dtype in (torch.float32, torch.bfloat16):
        store = ops.ChunkStore(N, D, dtype, DEV)
        store.append(rows.to(DEV))
        hs, hi = store.search_host(queries.numpy(), k)
        ds, di = store.search(queries.to(DEV), k)
        np.testing.assert_array_equal(hi, di.cpu().numpy())
        np.testing.assert_array_equal(hs, ds.cpu().numpy())
        s1, i1 = store.search_host(queries[0].tolist(), 3)   # a Python list, like the reference passes
        assert s1.shape == (1, 3) and (i1[0] == hi[0, :3]).all()
        with pytest.raises(DabError, match="different vector dimensions"):
            store.search_host(np.zeros(5, np.float32), 3)
        store.close()
    # adversarial corpus (see test_gpu_search.py): the C call must retry internally
    N, D = 300_000, 64
    rows = torch.randn(N, D, generator=g)
    q = torch.randn(1, D, generator=g)
    store = ops.ChunkStore(N, D, torch.float32, DEV)
    store.append(rows.to(DEV))
    plan = store.describe_plan(1, 10)
    r = torch.arange(100_000, 130_000)
    r = r[(r // plan["tile_rows"]) % plan["strides"][-2] != 0]
    rows[r] = q + 0.05 * torch.randn(r.numel(), D, generator=g)
    store.reset()
    store.append(rows.to(DEV))
    hs, hi = store.search_host(q.numpy(), 10)
    S = search_ref.exact_scores(rows.numpy(), q.numpy())
    search_ref.check_topk_parity(S[0], hi[0], hs[0], 10)
    store.close()


def test_concurrent_searches_from_threads():
    """The reference reaches the search from an asgiref worker thread while another coroutine runs
    (context_service/service.py:82-83) and from several uvicorn workers: calls on one store are
    serialised by the library, calls on different stores are independent; answers must not mix."""
    import threading
    from dab_b200 import ops
    g = torch.Generator().manual_seed(13)
    N, D, k = 30_000, 768, 10
    rows_a = torch.randn(N, D, generator=g)
    rows_b = torch.randn(N, D, generator=g)
    qs = torch.randn(8, D, generator=g)
    store_a = ops.ChunkStore(N, D, torch.float32, DEV)
    store_b = ops.ChunkStore(N, D, torch.bfloat16, DEV)
    store_a.append(rows_a.to(DEV))
    store_b.append(rows_b.to(DEV))
    Sa = search_ref.exact_scores(rows_a.numpy(), qs.numpy())
    Sb = search_ref.exact_scores(rows_b.to(torch.bfloat16).float().numpy(), qs.numpy())
    errors = []

    def worker(store, S, qi, reps):
        try:
            stream = torch.cuda.Stream(device=DEV)
            with torch.cuda.stream(stream):
                for _ in range(reps):
                    s, i = store.search_host(qs[qi].numpy(), k)
                    search_ref.check_topk_parity(S[qi], i[0], s[0], k)
                    s2, i2 = store.search(qs[qi:qi + 1].to(DEV), k)
                    search_ref.check_topk_parity(S[qi], i2[0].cpu().numpy(), s2[0].cpu().numpy(), k)
        except Exception as e:  # pragma: no cover
            errors.append(repr(e))

    threads = [threading.Thread(target=worker, args=(store_a if t % 2 == 0 else store_b, Sa if t % 2 == 0 else Sb, t, 15))
               for t in range(8)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    store_a.close()
    store_b.close()

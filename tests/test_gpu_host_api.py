"""GPU tests of the drop-in Python surface: the reference's search_service / embedder API on the B200."""
import asyncio
import os

import numpy as np
import pytest
import torch

from oracle import pool_ref, search_ref

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture()
def table():
    from dab_b200.search_service import DeviceTable
    rng = np.random.default_rng(42)
    n, d = 3000, 768
    emb = rng.standard_normal((n, d)).astype(np.float32)
    pks = np.arange(1000, 1000 + n)                     # Django pks, ascending with insertion
    docs = rng.integers(1, 60, size=n)
    bots = docs % 3
    t = DeviceTable(n, d, torch.float32, DEV)
    t.add(pks[:1700], emb[:1700], document_ids=docs[:1700], texts=[f"q{i}" for i in range(1700)], document__wiki__bot=bots[:1700])
    t.add(pks[1700:], emb[1700:], document_ids=docs[1700:], texts=[f"q{i}" for i in range(1700, n)], document__wiki__bot=bots[1700:])
    t._test = dict(emb=emb, pks=pks, docs=docs, bots=bots)
    yield t
    t.close()


def test_objects_embedding_search_matches_oracle(table):
    """_objects_embedding_search (search_service.py:185-196): objects ordered by ascending distance,
    filters honoured, fewer than n when the queryset is small."""
    from dab_b200 import search_service as ss
    T = table._test
    q = np.random.default_rng(1).standard_normal(768).astype(np.float32)
    S = search_ref.exact_scores(T["emb"], q)[0]
    qs = ss.ChunkQuerySet(table)
    hits = asyncio.run(ss._objects_embedding_search(q.tolist(), qs, n=5))
    want, wsc = search_ref.rank_topk(S, 5)
    assert [h.id for h in hits] == list(T["pks"][want])
    np.testing.assert_allclose([h.distance for h in hits], 1 - wsc, atol=2e-6)
    assert all(hits[i].distance <= hits[i + 1].distance for i in range(4))
    assert hits[0].text == f"q{want[0]}" and hits[0].document_id == T["docs"][want[0]]
    # the query-time filter of EmbeddingsStep (steps/embeddings.py:26-29) and of MergeQuestionsStep (questions.py:124)
    fqs = qs.filter(document__wiki__bot=2).filter(document__id__lt=40)
    allow = (T["bots"] == 2) & (T["docs"] < 40)
    hits = asyncio.run(ss.embedding_search_questions(q.tolist(), fqs, n=250))
    want, wsc = search_ref.rank_topk(S, 250, allow)
    assert [h.id for h in hits] == list(T["pks"][want])
    tiny = qs.filter(pk__in=[1000, 1005, 2999])
    hits = asyncio.run(ss.embedding_search_sentences(q.tolist(), tiny, n=10))
    assert len(hits) == 3 and {h.id for h in hits} == {1000, 1005, 2999}
    with pytest.raises(Exception, match="different vector dimensions"):
        asyncio.run(ss._objects_embedding_search([0.0] * 5, qs, n=5))


def test_embedding_search_doc_level(table):
    """embedding_search (search_service.py:111-152) == oracle aggregation over the oracle ranking."""
    from dab_b200 import search_service as ss
    T = table._test
    qv = np.random.default_rng(2).standard_normal(768).astype(np.float32)

    class Emb:
        async def embeddings(self, texts):
            return [qv.tolist() for _ in texts]

    ss.configure(Emb())
    got = asyncio.run(ss.embedding_search("how do I reset my password", ss.ChunkQuerySet(table), max_scores_n=5, top_n=5))
    S = search_ref.exact_scores(T["emb"], qv)[0]
    ids, sc = search_ref.rank_topk(S, 5 * 5 * 10)
    want = search_ref.doc_level_scores(T["docs"][ids], 1 - sc, 5, 5)
    assert [d for d, _ in got] == [d for d, _ in want]
    np.testing.assert_allclose([s for _, s in got], [s for _, s in want], atol=1e-6)
    ss.configure(None)


def test_django_style_queryset_path(table):
    """A duck-typed Django queryset: pks select the allow-set, instances come back via in_bulk with .distance."""
    from dab_b200 import search_service as ss
    T = table._test

    class Obj:
        def __init__(self, pk):
            self.pk = self.id = pk

    class Manager:
        def in_bulk(self, ids):
            return {i: Obj(i) for i in ids}

    class Meta:
        label_lower = "storage.question"

    class Question:
        _meta = Meta()
        objects = Manager()

    class QS:
        model = Question

        def __init__(self, pks):
            self._pks = pks

        def values_list(self, *a, **k):
            return self._pks

    ss.register_table("storage.Question", "embedding", table)
    allowed = [int(p) for p in T["pks"][::7]]
    q = np.random.default_rng(3).standard_normal(768).astype(np.float32)
    hits = asyncio.run(ss._objects_embedding_search(q.tolist(), QS(allowed), n=10))
    allow = np.zeros(len(T["pks"]), bool)
    allow[::7] = True
    want, wsc = search_ref.rank_topk(search_ref.exact_scores(T["emb"], q)[0], 10, allow)
    assert [h.pk for h in hits] == list(T["pks"][want]) and isinstance(hits[0], Obj)
    np.testing.assert_allclose([h.distance for h in hits], 1 - wsc, atol=2e-6)


def _golden_model_and_tokenizer():
    """Rebuilds exactly what oracle/make_golden.py fed the reference: same seed, same ids."""
    from transformers import BatchEncoding, BertConfig, BertModel
    torch.manual_seed(0)
    cfg = BertConfig(vocab_size=30522, hidden_size=384, num_hidden_layers=2, num_attention_heads=12,
                     intermediate_size=1536, max_position_embeddings=512)
    model = BertModel(cfg).eval()
    lengths = [5, 40, 17, 1, 33, 24]
    rng = np.random.default_rng(20261106)
    ids_per_text = {f"text-{i}": rng.integers(1000, 30000, size=L).astype(np.int64) for i, L in enumerate(lengths)}

    def tokenizer(texts, return_tensors="pt", padding=True):
        T = max(len(ids_per_text[t]) for t in texts)
        ids = torch.zeros(len(texts), T, dtype=torch.long)
        mask = torch.zeros(len(texts), T, dtype=torch.long)
        for i, t in enumerate(texts):
            L = len(ids_per_text[t])
            ids[i, :L] = torch.from_numpy(ids_per_text[t])
            mask[i, :L] = 1
        return BatchEncoding({"input_ids": ids, "attention_mask": mask})

    return model, tokenizer, list(ids_per_text)


def test_b200_embedder_matches_reference_embedder(golden_dir):
    """B200TransformersEmbedder.embeddings (batched, fused pool) vs the reference's
    TransformersEmbedder.embeddings output recorded in tests/golden/pool_ref.npz."""
    from dab_b200.embedders import B200TransformersEmbedder
    g = np.load(os.path.join(golden_dir, "pool_ref.npz"))
    model, tok, texts = _golden_model_and_tokenizer()
    ref = g["ref_embeddings"]
    emb = B200TransformersEmbedder("random-init-minilm-shaped", model=model, tokenizer=tok, device=DEV)
    got = np.asarray(asyncio.run(emb.embeddings(texts)))
    np.testing.assert_allclose(got, ref / np.linalg.norm(ref, axis=1, keepdims=True), rtol=1e-4, atol=5e-6)
    raw = B200TransformersEmbedder("random-init-minilm-shaped", model=model, tokenizer=tok, device=DEV, unit_norm=False)
    np.testing.assert_allclose(np.asarray(asyncio.run(raw.embeddings(texts))), ref, rtol=1e-4, atol=5e-6)
    assert asyncio.run(emb.embeddings([])) == []


def test_embedder_truncation_bf16_forward_and_embed_into():
    """(f)3: over-long texts are truncated instead of failing; a bf16 encoder feeds the pool kernel's bf16 input
    path; embed_into appends the pooled rows to a store without leaving the device
    (assistant/processing/documents/steps/embeddings.py:27-39)."""
    from dab_b200 import ops
    from dab_b200.embedders import B200TransformersEmbedder
    from oracle import pool_ref
    model, tok, texts = _golden_model_and_tokenizer()
    full = B200TransformersEmbedder("m", model=model, tokenizer=tok, device=DEV)
    cut = B200TransformersEmbedder("m", model=model, tokenizer=tok, device=DEV, max_length=16)
    e_full = np.asarray(asyncio.run(full.embeddings(texts)))
    e_cut = np.asarray(asyncio.run(cut.embeddings(texts)))
    enc = tok(texts)
    ids, mask = enc["input_ids"][:, :16].to(DEV), enc["attention_mask"][:, :16].to(DEV)
    with torch.no_grad():
        hidden = model.to(DEV)(input_ids=ids, attention_mask=mask).last_hidden_state
    want, _ = pool_ref.masked_mean_pool(hidden.cpu().numpy(), mask.cpu().numpy())
    np.testing.assert_allclose(e_cut, want, rtol=1e-4, atol=5e-6)
    short = [i for i, t in enumerate(texts) if int(enc["attention_mask"][i].sum()) <= 16]
    np.testing.assert_allclose(e_cut[short], e_full[short], rtol=1e-4, atol=5e-6)      # short texts are untouched
    # bf16 forward: same direction within bf16 noise, still unit norm
    import copy
    b16 = B200TransformersEmbedder("m", model=copy.deepcopy(model), tokenizer=tok, device=DEV, dtype=torch.bfloat16)
    hidden16, _ = b16._encode_batch(texts)
    assert hidden16.dtype == torch.bfloat16
    e16 = np.asarray(asyncio.run(b16.embeddings(texts)))
    np.testing.assert_allclose(np.linalg.norm(e16, axis=1), 1.0, rtol=1e-5)
    assert ((e16 * e_full).sum(axis=1) > 0.995).all()
    # ingest: encode -> pool -> append, on the device, into an fp32 and a bf16 store
    for dtype in (torch.float32, torch.bfloat16):
        store = ops.ChunkStore(64, 384, dtype, DEV)
        assert full.embed_into(store, texts) == len(texts) and len(store) == len(texts)
        s, i = store.search(torch.from_numpy(e_full[2:3]).float().to(DEV), 2)
        assert i[0, 0].item() == 2 and abs(s[0, 0].item() - 1.0) < (1e-5 if dtype == torch.float32 else 4e-3)
        store.close()


def test_endpoint_with_b200_embedder():
    from fastapi.testclient import TestClient
    from dab_b200.embedders import B200TransformersEmbedder
    from dab_b200.gpu_service import create_app
    model, tok, texts = _golden_model_and_tokenizer()
    app = create_app(embedders={"minilm": B200TransformersEmbedder("minilm", model=model, tokenizer=tok, device=DEV)})
    with TestClient(app) as c:
        r = c.post("/embeddings/", json={"model": "MiniLM", "texts": texts[:3]})
        assert r.status_code == 200
        e = np.asarray(r.json()["embeddings"])
        assert e.shape == (3, 384)
        np.testing.assert_allclose(np.linalg.norm(e, axis=1), 1.0, rtol=1e-5)
        from dab_b200.gpu_service import BINARY_MEDIA_TYPE, decode_binary_embeddings
        rb = c.post("/embeddings/", json={"model": "minilm", "texts": texts[:3]}, headers={"Accept": BINARY_MEDIA_TYPE})
        assert rb.status_code == 200
        eb = decode_binary_embeddings(rb.content, rb.headers["x-embedding-shape"])
        assert eb.shape == (3, 384)
        np.testing.assert_allclose(eb, e, rtol=1e-6, atol=1e-7)      # same values, fewer bytes
        assert len(rb.content) * 3 < len(r.content)
        r = c.post("/embeddings/", json={"model": "minilm", "texts": ["not-in-the-stub-vocabulary"]})
        assert r.status_code == 500        # tokenizer KeyError surfaces as HTTP 500 with str(e)


@pytest.mark.parametrize("n,max_scores_n,top_n", [(250, 5, 5), (40, 1, 3), (1000, 10, 10), (64, 40, 2), (7, 2, 4)])
def test_doc_aggregate_kernel_matches_oracle(n, max_scores_n, top_n):
    """dab_doc_aggregate == oracle restatement of search_service.py:133-150 on random hit lists."""
    from dab_b200 import ops
    rng = np.random.default_rng(n * 31 + max_scores_n)
    Q, N = 6, 5000
    doc_of_row = rng.integers(0, max(3, n // 8), size=N).astype(np.int64)
    scores = np.sort(rng.random((Q, n)).astype(np.float32), axis=1)[:, ::-1].copy()
    rows = np.stack([rng.permutation(N)[:n] for _ in range(Q)]).astype(np.int64)
    rows[1, n // 2:] = -1                                   # a query with fewer hits than n
    docs, dsc = ops.doc_aggregate(torch.from_numpy(scores).to(DEV), torch.from_numpy(rows).to(DEV),
                                  torch.from_numpy(doc_of_row).to(DEV), max_scores_n, top_n)
    docs, dsc = docs.cpu().numpy(), dsc.cpu().numpy()
    for q in range(Q):
        valid = rows[q] >= 0
        want = search_ref.doc_level_scores(doc_of_row[rows[q][valid]], 1.0 - scores[q][valid].astype(np.float64),
                                           max_scores_n, top_n)
        got = [(int(d), float(s)) for d, s in zip(docs[q], dsc[q]) if d >= 0]
        assert [d for d, _ in got] == [d for d, _ in want], (q, got, want)
        np.testing.assert_allclose([s for _, s in got], [s for _, s in want], atol=2e-6)
        assert (docs[q][len(want):] == -1).all()


def test_embedding_search_many_is_the_batched_dedup_query(table):
    """One call answers the per-question top-1 searches of MergeQuestionsStep (questions.py:121-139)."""
    from dab_b200 import search_service as ss
    T = table._test
    new_q = T["emb"][[5, 700, 2500]] + 0.01 * np.random.default_rng(9).standard_normal((3, 768)).astype(np.float32)
    qs = ss.ChunkQuerySet(table).filter(document__id__lt=30)          # "older documents only"
    got = asyncio.run(ss.embedding_search_many(new_q.tolist(), qs, n=1))
    allow = T["docs"] < 30
    S = search_ref.exact_scores(T["emb"], new_q)
    for i in range(3):
        want, wsc = search_ref.rank_topk(S[i], 1, allow)
        assert len(got[i]) == 1 and got[i][0].id == T["pks"][want[0]]
        assert got[i][0].distance == pytest.approx(1 - wsc[0], abs=2e-6)
        single = asyncio.run(ss.embedding_search_questions(new_q[i].tolist(), qs, n=1))
        assert single[0].id == got[i][0].id
    assert asyncio.run(ss.embedding_search_many([], qs)) == []


def test_c1_reference_cpu_path_shape_end_to_end():
    """BASELINE config[0]: 32 chunks x 256 tokens through a MiniLM-L6-shaped 384-d encoder (random init),
    mean-pool + L2 norm, exact cosine top-10 over 10k chunks - the whole hot path on the device against
    the oracle (pool: masked mean of the same hidden states; retrieve: float64 cosine ranking)."""
    from transformers import BertConfig, BertModel
    from dab_b200 import ops
    torch.manual_seed(0)
    cfg = BertConfig(vocab_size=30522, hidden_size=384, num_hidden_layers=6, num_attention_heads=12,
                     intermediate_size=1536, max_position_embeddings=512)
    model = BertModel(cfg).eval().to(DEV)
    g = torch.Generator().manual_seed(555)
    B, T = 32, 256
    ids = torch.randint(1000, 30000, (B, T), generator=g)
    lengths = torch.randint(16, T + 1, (B,), generator=g)
    mask = (torch.arange(T)[None] < lengths[:, None]).long()
    with torch.no_grad():
        hidden = model(input_ids=ids.to(DEV), attention_mask=mask.to(DEV)).last_hidden_state
    emb = ops.pool_l2norm(hidden, mask.to(DEV))
    want, _ = pool_ref.masked_mean_pool(hidden.float().cpu().numpy(), mask.numpy())
    np.testing.assert_allclose(emb.cpu().numpy(), want, rtol=1e-5, atol=1e-6)
    # the 10k-chunk store: 32 embedded chunks + random unit vectors, queried with the chunks themselves
    N = 10_000
    corpus = torch.nn.functional.normalize(torch.randn(N, 384, generator=g), dim=1)
    corpus[:B] = emb.cpu()
    store = ops.ChunkStore(N, 384, torch.float32, DEV)
    store.append(corpus.to(DEV))
    scores, rows = store.search(emb, 10)
    S = search_ref.exact_scores(corpus.numpy(), emb.cpu().numpy())
    for q in range(B):
        search_ref.check_topk_parity(S[q], rows[q].cpu().numpy(), scores[q].cpu().numpy(), 10)
        assert rows[q, 0].item() == q and abs(scores[q, 0].item() - 1.0) < 1e-5       # each chunk finds itself first
    store.close()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_delete_and_snapshot_roundtrip(tmp_path, dtype):
    """Tombstoned rows never come back; a saved shard reloads to the same answers."""
    from dab_b200 import search_service as ss
    rng = np.random.default_rng(3)
    n, d = 4000, 768
    emb = rng.standard_normal((n, d)).astype(np.float32)
    pks = np.arange(10, 10 + n)
    docs = rng.integers(1, 40, size=n)
    t = ss.DeviceTable(n, d, dtype, DEV)
    t.add(pks, emb, document_ids=docs, texts=[f"s{i}" for i in range(n)], document__wiki__bot=docs % 2)
    stored = torch.from_numpy(emb).to(dtype).float().numpy()
    q = rng.standard_normal(d).astype(np.float32)
    S = search_ref.exact_scores(stored, q)[0]
    top, _ = search_ref.rank_topk(S, 5)
    # delete the two best hits and every row of one document (what finalisation does to older documents)
    gone = set(pks[top[:2]].tolist()) | set(pks[docs == docs[top[2]]].tolist())
    assert t.delete(gone) == len(gone) and t.delete(gone) == 0
    alive = ~np.isin(pks, list(gone))
    qs = ss.ChunkQuerySet(t)
    assert qs.count() == alive.sum()
    hits = asyncio.run(ss._objects_embedding_search(q.tolist(), qs, n=20))
    want, wsc = search_ref.rank_topk(S, 20, alive)
    assert [h.id for h in hits] == list(pks[want])
    assert qs.filter(document__wiki__bot=1).count() == (alive & (docs % 2 == 1)).sum()
    # snapshot -> reload -> same answers, deletions included
    path = str(tmp_path / "shard.npz")
    t.save(path)
    t2 = ss.DeviceTable.load(path, device=DEV)
    assert len(t2) == n and ss.ChunkQuerySet(t2).count() == alive.sum()
    hits2 = asyncio.run(ss._objects_embedding_search(q.tolist(), ss.ChunkQuerySet(t2), n=20))
    assert [h.id for h in hits2] == [h.id for h in hits] and hits2[0].text == hits[0].text
    np.testing.assert_allclose([h.distance for h in hits2], [h.distance for h in hits], atol=1e-7)
    t.close()
    t2.close()


def test_bulk_load_from_a_django_style_queryset():
    """DeviceTable.load_queryset: the device store (re)built from the database column - NULL embeddings skipped
    (the column is nullable, assistant/storage/models.py:23), rows streamed in chunks with their pk, document id,
    text, order and filter columns; then a filtered search over what was loaded."""
    from dab_b200 import search_service as ss
    rng = np.random.default_rng(8)
    n, dim = 1000, 64
    emb = rng.standard_normal((n, dim)).astype(np.float32)
    records = [dict(pk=10 + i, embedding=None if i % 9 == 0 else emb[i].tolist(), document_id=i // 5, text=f"s{i}", order=i % 5,
                    document__wiki__bot_id=i % 3) for i in range(n)]

    class QS:
        def __init__(self, recs):
            self.recs = recs

        def exclude(self, **kw):
            assert kw == {"embedding__isnull": True}
            return QS([r for r in self.recs if r["embedding"] is not None])

        def values_list(self, *names):
            self.names = names
            return self

        def iterator(self, chunk_size=2000):
            assert chunk_size == 128
            return iter([tuple(r[nm] for nm in self.names) for r in self.recs])

    t = ss.DeviceTable(n, dim, torch.float32, DEV)
    loaded = t.load_queryset(QS(records), "embedding", columns={"document__wiki__bot": "document__wiki__bot_id"}, batch=128)
    keep = np.array([i for i in range(n) if i % 9 != 0])
    assert loaded == len(keep) == len(t)
    q = emb[keep[17]] + 0.01
    hits = asyncio.run(ss._objects_embedding_search(q.tolist(), ss.ChunkQuerySet(t).filter(document__wiki__bot=keep[17] % 3), n=5))
    S = search_ref.exact_scores(emb[keep], q[None])[0]
    want, wsc = search_ref.rank_topk(S, 5, (keep % 3) == keep[17] % 3)
    assert [h.pk for h in hits] == [10 + int(keep[w]) for w in want]
    assert hits[0].text == f"s{keep[17]}" and hits[0].document_id == keep[17] // 5 and hits[0].order == keep[17] % 5
    np.testing.assert_allclose([h.distance for h in hits], 1 - wsc, atol=2e-6)
    t.close()

"""CPU tests of the host-side mirror of the reference interface (no GPU needed)."""
import asyncio

import numpy as np
import pytest

from oracle import search_ref


class _Hit:
    def __init__(self, doc, dist):
        self.document_id, self.distance = doc, dist


def test_aggregate_documents_matches_oracle():
    """dab_b200.search_service.aggregate_documents == oracle restatement of search_service.py:133-150."""
    from dab_b200.search_service import aggregate_documents
    rng = np.random.default_rng(0)
    docs = rng.integers(0, 12, size=250)
    dist = np.sort(rng.random(250))
    hits = [_Hit(int(d), float(x)) for d, x in zip(docs, dist)]
    for msn, tn in ((5, 5), (1, 3), (10, 10), (40, 2)):
        got = aggregate_documents(hits, msn, tn)
        want = search_ref.doc_level_scores(docs, dist, msn, tn)
        assert [d for d, _ in got] == [d for d, _ in want]
        np.testing.assert_allclose([s for _, s in got], [s for _, s in want], rtol=1e-12)


def test_cosine_helpers_match_reference_golden(golden_dir):
    import os
    from dab_b200.search_service import cosine_similarity, embeddings_similarity
    g = np.load(os.path.join(golden_dir, "search_ref.npz"))
    for qi in range(2):
        for r in (0, 5, 63):
            assert embeddings_similarity(g["queries"][qi].tolist(), g["rows"][r].tolist()) == pytest.approx(g["sims"][qi, r], rel=1e-14)
    assert cosine_similarity([1.0, 0.0], [0.0, 1.0]) == 0.0


class _FakeTable:
    def __init__(self, n):
        self.cols = {"pk": np.arange(n), "document_id": np.arange(n) // 3, "document__wiki__bot": np.arange(n) % 2}

    def __len__(self):
        return len(self.cols["pk"])

    def column(self, name):
        return self.cols[{"document__id": "document_id", "id": "pk"}.get(name, name)]


def test_chunk_queryset_filter_lookups():
    from dab_b200.search_service import ChunkQuerySet
    qs = ChunkQuerySet(_FakeTable(30))
    assert qs.count() == 30 and qs._mask is None
    a = qs.filter(document__wiki__bot=1)
    assert a.count() == 15 and qs.count() == 30                      # filter returns a new view
    b = a.filter(document__id__lt=4)
    assert set(np.nonzero(b._mask)[0]) == {1, 3, 5, 7, 9, 11}
    assert qs.filter(pk__in=[2, 4, 99]).count() == 2
    assert qs.exclude(document_id=0).count() == 27
    assert qs.filter(document_id__gte=9).count() == 3


class _GoodEmbedder:
    async def embeddings(self, texts):
        return [[float(len(t)), 1.0] for t in texts]


class _BadEmbedder:
    async def embeddings(self, texts):
        raise RuntimeError("device on fire")


def test_embeddings_endpoint_contract():
    """Same JSON contract and status codes as gpu_service/main.py:75-86."""
    from fastapi.testclient import TestClient
    from dab_b200.gpu_service import create_app
    app = create_app(embedders={"Some/Model": _GoodEmbedder(), "bad": _BadEmbedder()})
    with TestClient(app) as c:
        r = c.post("/embeddings/", json={"model": "SOME/model", "texts": ["ab", "abcd"]})       # lower-cased lookup
        assert r.status_code == 200 and r.json() == {"embeddings": [[2.0, 1.0], [4.0, 1.0]]}
        r = c.post("/embeddings/", json={"model": "unknown", "texts": ["x"]})
        assert r.status_code == 400 and r.json()["detail"] == "Model is not supported"
        r = c.post("/embeddings/", json={"model": "bad", "texts": ["x"]})
        assert r.status_code == 500 and r.json()["detail"] == "device on fire"
        assert c.post("/embeddings/", json={"model": "bad"}).status_code == 422


def test_embeddings_endpoint_binary_transport_is_opt_in():
    """Accept: application/octet-stream -> raw float32 + shape header; the JSON contract stays the default."""
    from fastapi.testclient import TestClient
    from dab_b200.gpu_service import BINARY_MEDIA_TYPE, create_app, decode_binary_embeddings
    app = create_app(embedders={"m": _GoodEmbedder(), "bad": _BadEmbedder()})
    with TestClient(app) as c:
        r = c.post("/embeddings/", json={"model": "m", "texts": ["ab", "abcd", ""]}, headers={"Accept": BINARY_MEDIA_TYPE})
        assert r.status_code == 200 and r.headers["content-type"] == BINARY_MEDIA_TYPE
        assert r.headers["x-embedding-shape"] == "3,2" and len(r.content) == 3 * 2 * 4
        np.testing.assert_array_equal(decode_binary_embeddings(r.content, r.headers["x-embedding-shape"]),
                                      np.asarray([[2, 1], [4, 1], [0, 1]], dtype=np.float32))
        assert c.post("/embeddings/", json={"model": "m", "texts": ["ab"]}).json() == {"embeddings": [[2.0, 1.0]]}
        r = c.post("/embeddings/", json={"model": "bad", "texts": ["x"]}, headers={"Accept": BINARY_MEDIA_TYPE})
        assert r.status_code == 500 and r.json()["detail"] == "device on fire"
        r = c.post("/embeddings/", json={"model": "nope", "texts": ["x"]}, headers={"Accept": BINARY_MEDIA_TYPE})
        assert r.status_code == 400


def test_endpoint_skips_models_that_fail_to_load():
    from fastapi.testclient import TestClient
    from dab_b200.gpu_service import create_app

    def factory(name):
        if name == "broken":
            raise OSError("no weights on disk")
        return _GoodEmbedder()

    app = create_app(embedder_models=["ok", "broken"], embedder_factory=factory)
    with TestClient(app) as c:
        assert c.post("/embeddings/", json={"model": "ok", "texts": ["a"]}).status_code == 200
        assert c.post("/embeddings/", json={"model": "broken", "texts": ["a"]}).status_code == 400


def test_get_embedding_requires_configuration():
    from dab_b200 import DabError, search_service
    search_service.configure(None)
    with pytest.raises(DabError, match="no embedder configured"):
        asyncio.run(search_service.get_embedding("x"))
    search_service.configure(_GoodEmbedder())
    assert asyncio.run(search_service.get_embedding("abc")) == [3.0, 1.0]
    search_service.configure(None)

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
    def __len__(self):
        return 30


def test_chunk_queryset_collects_predicates_lazily():
    """filter / exclude only record Django-style lookups (nothing is evaluated on the host: the device builds the
    bitmap when the search asks for it, tests/test_gpu_filters.py); each call returns a new view."""
    from dab_b200.search_service import ChunkQuerySet
    qs = ChunkQuerySet(_FakeTable())
    assert qs.count() == 30 and qs.allow_bitmap() is None
    a = qs.filter(document__wiki__bot=1)
    b = a.filter(document__id__lt=4).exclude(pk__in=(2, 4, 99))
    assert qs._preds == [] and a._preds == [("document__wiki__bot", "exact", 1, False)]
    assert b._preds == [("document__wiki__bot", "exact", 1, False), ("document__id", "lt", 4, False),
                        ("pk", "in", [2, 4, 99], True)]
    assert qs.all() is qs


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


def test_get_embedding_requires_configuration(monkeypatch):
    from dab_b200 import DabError, search_service
    search_service.configure(None)
    monkeypatch.delenv("EMBEDDING_AI_MODEL", raising=False)
    with pytest.raises(DabError, match="no embedding model named"):
        asyncio.run(search_service.get_embedding("x"))
    search_service.configure(_GoodEmbedder())
    assert asyncio.run(search_service.get_embedding("abc")) == [3.0, 1.0]
    search_service.configure(None)


def _serve(app):
    """The FastAPI app on a real socket (uvicorn in a thread), for clients that speak HTTP themselves."""
    import socket
    import threading
    import time
    import uvicorn
    sock = socket.socket()
    sock.bind(("127.0.0.1", 0))
    port = sock.getsockname()[1]
    sock.close()
    server = uvicorn.Server(uvicorn.Config(app, host="127.0.0.1", port=port, log_level="error"))
    th = threading.Thread(target=server.run, daemon=True)
    th.start()
    for _ in range(200):
        if server.started:
            break
        time.sleep(0.05)
    assert server.started
    return server, th, f"http://127.0.0.1:{port}"


def test_http_clients_and_factory_against_the_endpoint(monkeypatch):
    """The reference's kept surface end to end over HTTP: ``get_ai_embdedder('gpu_service:<model>')``
    (assistant/ai/services/ai_service.py:61-67) builds the client, the client talks to this repository's
    ``POST /embeddings/``.  Where the reference tree is present (the build container) its OWN unmodified
    ``GPUServiceEmbedder`` is driven against the app as well."""
    import asyncio
    import os
    import sys
    from dab_b200 import embedders
    from dab_b200.gpu_service import create_app
    server, th, url = _serve(create_app(embedders={"some/model": _GoodEmbedder(), "bad": _BadEmbedder()}))
    try:
        monkeypatch.setenv("GPU_SERVICE_ENDPOINT", url)
        monkeypatch.setenv("EMBEDDING_AI_MODEL", "gpu_service:Some/Model")
        e = embedders.get_ai_embdedder("gpu_service:Some/Model")
        assert isinstance(e, embedders.GPUServiceEmbedder) and isinstance(e, embedders.AIEmbedder)
        assert asyncio.run(e.embeddings(["ab", "abcd"])) == [[2.0, 1.0], [4.0, 1.0]]
        assert asyncio.run(embedders.GPUServiceEmbedder(url, "some/model", binary=True).embeddings(["abc"])) == [[3.0, 1.0]]
        with pytest.raises(Exception, match="status code 500"):
            asyncio.run(embedders.get_ai_embdedder("gpu_service:bad").embeddings(["x"]))
        with pytest.raises(Exception, match="status code 400"):
            asyncio.run(embedders.get_ai_embdedder("gpu_service:unknown").embeddings(["x"]))
        with pytest.raises(Exception):
            embedders.get_ai_embdedder("text-embedding-3-small")
        # search_service.get_embedding falls back to the factory with settings.EMBEDDING_AI_MODEL (search_service.py:155-158)
        from dab_b200 import search_service as ss
        ss.configure(embedder=None)
        assert asyncio.run(ss.get_embedding("hello")) == [5.0, 1.0]
        if os.path.isdir("/root/reference"):
            sys.path.insert(0, "/root/reference")
            try:
                from assistant.ai.embedders.gpu_service import GPUServiceEmbedder as RefClient
                assert asyncio.run(RefClient(base_url=url, model="some/model").embeddings(["ab", "abcd"])) == [[2.0, 1.0], [4.0, 1.0]]
                with pytest.raises(Exception, match="status code 400"):
                    asyncio.run(RefClient(base_url=url, model="nope").embeddings(["x"]))
            finally:
                sys.path.remove("/root/reference")
    finally:
        server.should_exit = True
        th.join(timeout=5)


def test_device_table_host_columns_without_a_gpu():
    """The host side of DeviceTable (growable NumPy columns, pk -> row index by sorted search, dictionary codes for
    non-numeric filter values) is plain NumPy and can be checked here; the device side is in tests/test_gpu_*.py."""
    from dab_b200.search_service import DeviceTable, _Grow
    g = _Grow(np.int64)
    for i in range(5):
        g.extend(np.arange(i * 700, (i + 1) * 700))
    assert g.n == 3500 and (g.view() == np.arange(3500)).all()
    g.keep(np.arange(0, 3500, 7))
    assert g.n == 500 and g.view()[3] == 21
    t = DeviceTable.__new__(DeviceTable)                      # no store: only the host-side helpers are exercised
    t._pks, t._dead, t._pk_sorted = _Grow(np.int64), _Grow(np.bool_), None
    t._codes, t._decode, t._meta = {}, {}, {}
    pks = np.array([50, 10, 40, 20, 30], dtype=np.int64)
    t._pks.extend(pks)
    t._dead.extend(np.array([False, False, True, False, False]))
    assert sorted(t._rows_of_pks([10, 20, 40, 99, 10]).tolist()) == [1, 3]      # 40 is already deleted, 99 unknown
    assert t._encode("bot", [3, 5, 3]).tolist() == [3, 5, 3] and "bot" not in t._codes
    assert t._encode("status", ["done", "new", "done"]).tolist() == [0, 1, 0]
    assert t._encode("status", ["new", "failed"]).tolist() == [1, 2] and t._decode["status"] == ["done", "new", "failed"]
    assert t._encode_operand("status", "failed") == 2 and t._encode_operand("status", "never") is None

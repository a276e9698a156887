"""Generate tests/golden/*.npz by running the REFERENCE's own Python for the hot path.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Run in the build container, where
/root/reference exists:  ``python oracle/make_golden.py``.  The GPU box has no
/root/reference; it only sees the committed fixtures.

What is executed from the reference (unmodified, imported from /root/reference):
  * ``assistant.ai.embedders.transformers.TransformersEmbedder.embeddings``
    (assistant/ai/embedders/transformers.py:15-29) - instance made with ``__new__`` and an
    injected random-init BertModel + tokenizer stub, because no weights are on disk
    (``local_files_only=True``, :11-13) and there is no network.
  * ``assistant.rag.services.search_service.cosine_similarity`` / ``embeddings_similarity``
    (assistant/rag/services/search_service.py:17-21,32-35) - the module imports django,
    asgiref and pgvector at the top (:6-12); those are absent here, so inert stand-ins
    are placed in ``sys.modules`` first.  No arithmetic is stubbed.
pgvector's ``<=>`` itself cannot be run here (no PostgreSQL): that half stays "parity unpinned".
"""
from __future__ import annotations

import asyncio
import os
import sys
import types

import numpy as np

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def _stub_modules():
    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    mod("asgiref")
    mod("asgiref.sync", sync_to_async=lambda f: f)
    mod("django")
    mod("django.conf", settings=types.SimpleNamespace(EMBEDDING_AI_MODEL="stub"))
    mod("django.db")
    mod("django.db.models", QuerySet=object)
    mod("pgvector")
    mod("pgvector.django", CosineDistance=object)
    mod("assistant.ai.services.ai_service", get_ai_embdedder=lambda *a, **k: None)
    mod("assistant.storage.models", Document=object, Sentence=object, Question=object,
        BaseEmbeddingModel=object, WikiDocument=object)


def make_pool_golden():
    import torch
    from transformers import BatchEncoding, BertConfig, BertModel

    from assistant.ai.embedders.transformers import TransformersEmbedder  # the reference class

    torch.manual_seed(0)
    cfg = BertConfig(vocab_size=30522, hidden_size=384, num_hidden_layers=2, num_attention_heads=12,
                     intermediate_size=1536, max_position_embeddings=512)
    model = BertModel(cfg).eval()

    lengths = [5, 40, 17, 1, 33, 24]
    rng = np.random.default_rng(20261106)
    ids_per_text = {f"text-{i}": rng.integers(1000, 30000, size=L).astype(np.int64) for i, L in enumerate(lengths)}

    def tokenizer(text, return_tensors="pt"):
        ids = torch.from_numpy(ids_per_text[text])[None, :]
        return BatchEncoding({"input_ids": ids, "attention_mask": torch.ones_like(ids)})

    emb = TransformersEmbedder.__new__(TransformersEmbedder)
    emb._model_name = "random-init-minilm-shaped"
    emb._tokenizer = tokenizer
    emb._device = "cpu"
    emb._model = model
    texts = list(ids_per_text.keys())
    ref = asyncio.run(emb.embeddings(texts))  # <- reference code path, transformers.py:15-29
    ref = np.asarray(ref, dtype=np.float64)

    # the same texts as one padded batch: what the build's batched endpoint feeds the pool kernel
    T = max(lengths)
    B = len(lengths)
    input_ids = torch.zeros(B, T, dtype=torch.long)
    mask = torch.zeros(B, T, dtype=torch.long)
    for i, t in enumerate(texts):
        L = lengths[i]
        input_ids[i, :L] = torch.from_numpy(ids_per_text[t])
        mask[i, :L] = 1
    with torch.no_grad():
        hidden = model(input_ids=input_ids, attention_mask=mask).last_hidden_state
        per_text = [model(input_ids=input_ids[i:i + 1, :lengths[i]]).last_hidden_state[0] for i in range(B)]
    np.savez_compressed(
        os.path.join(OUT, "pool_ref.npz"),
        hidden=hidden.numpy().astype(np.float32),
        mask=mask.numpy(),
        lengths=np.asarray(lengths),
        per_text_hidden=torch.cat(per_text, 0).numpy().astype(np.float32),
        ref_embeddings=ref,
    )
    print("pool_ref.npz: ref", ref.shape, "hidden", tuple(hidden.shape))


def make_search_golden():
    from assistant.rag.services import search_service as ss  # the reference module

    rng = np.random.default_rng(4321)
    N, D, Q = 64, 768, 4
    rows = rng.standard_normal((N, D)).astype(np.float32)
    rows[10] = rows[3]            # duplicate rows -> tie, lower index first
    rows[11] = 2.5 * rows[7]      # scaled copy -> same cosine as row 7
    queries = rng.standard_normal((Q, D)).astype(np.float32)
    queries[1] = rows[20]         # identical vector -> distance 0
    queries[2] = -rows[21]        # opposite vector -> distance 2
    sims = np.empty((Q, N), dtype=np.float64)
    for qi in range(Q):
        for r in range(N):
            # search_service.py:32-35 fed Python lists, exactly as emb_test.py:29 calls it
            sims[qi, r] = ss.embeddings_similarity(queries[qi].tolist(), rows[r].tolist())
    legacy_order = np.stack([np.argsort([ss.cosine_similarity(np.array(queries[qi].tolist()), np.array(r.tolist()))
                                          for r in rows])[::-1] for qi in range(Q)])  # :79-82
    e = np.eye(D, dtype=np.float32)
    ka = np.array([
        ss.embeddings_similarity(e[0].tolist(), e[1].tolist()),          # orthogonal -> 0
        ss.embeddings_similarity(e[0].tolist(), e[0].tolist()),          # identical  -> 1
        ss.embeddings_similarity(e[0].tolist(), (-e[0]).tolist()),       # opposite   -> -1
        ss.embeddings_similarity((3 * rows[0]).tolist(), (0.5 * rows[1]).tolist()),
        ss.embeddings_similarity(rows[0].tolist(), rows[1].tolist()),   # scale invariance pair
    ])
    np.savez_compressed(os.path.join(OUT, "search_ref.npz"), rows=rows, queries=queries, sims=sims,
                        legacy_order=legacy_order, known_answers=ka)
    print("search_ref.npz: sims", sims.shape)


def make_search_golden_bf16():
    """A second, larger retrieval fixture: the rows are bf16-representable values (what a bf16 store holds, upcast
    exactly), the reference's own ``embeddings_similarity`` (search_service.py:32-35, float64 NumPy on Python lists)
    ranks them, and n = 50 > 10.  Rows regenerate from the seed, so only queries and reference outputs are stored."""
    from assistant.rag.services import search_service as ss  # the reference module
    from oracle import search_ref

    rng = np.random.default_rng(8765)
    N, D, Q, n = 400, 768, 6, 50
    rows = search_ref.bf16_bits_to_f32(search_ref.f32_to_bf16_bits(rng.standard_normal((N, D)).astype(np.float32)))
    rows[150] = rows[9]            # duplicates inside the top of query 0's list
    rows[151] = rows[9]
    rows[200] = 0                  # zero vector: NaN similarity, never ranked
    queries = rng.standard_normal((Q, D)).astype(np.float32)
    queries[0] = rows[9] + 0.01 * queries[0]
    queries[1] = 7.0 * rows[33]    # un-normalised query, identical direction
    sims = np.empty((Q, N), dtype=np.float64)
    with np.errstate(invalid="ignore", divide="ignore"):
        for qi in range(Q):
            for r in range(N):
                sims[qi, r] = ss.embeddings_similarity(queries[qi].tolist(), rows[r].tolist())
    np.savez_compressed(os.path.join(OUT, "search_ref_bf16.npz"), seed=8765, shape=np.array([N, D, Q, n]), queries=queries,
                        sims=sims, row_checksum=np.array([float(rows.astype(np.float64).sum())]))
    print("search_ref_bf16.npz: sims", sims.shape)


def golden_bf16_rows():
    """The rows of search_ref_bf16.npz, regenerated (tests call this; it needs nothing from the reference)."""
    from oracle import search_ref
    rng = np.random.default_rng(8765)
    N, D = 400, 768
    rows = search_ref.bf16_bits_to_f32(search_ref.f32_to_bf16_bits(rng.standard_normal((N, D)).astype(np.float32)))
    rows[150] = rows[9]
    rows[151] = rows[9]
    rows[200] = 0
    return rows


if __name__ == "__main__":
    if not os.path.isdir(REF):
        sys.exit("needs /root/reference (run in the build container)")
    os.makedirs(OUT, exist_ok=True)
    sys.path.insert(0, REF)
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))   # `from oracle import ...`
    _stub_modules()
    make_pool_golden()
    make_search_golden()
    make_search_golden_bf16()

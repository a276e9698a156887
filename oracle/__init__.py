"""CPU oracle for the django-assistant-bot retrieval hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it, and only as the checker or as the timed
CPU baseline.  The shipped path (``django-assistant-bot_b200/``) never imports it
and has no CPU fallback.

Parity status
-------------
* pool half: PINNED.  ``oracle/make_golden.py`` runs the reference's own
  ``TransformersEmbedder.embeddings`` (``assistant/ai/embedders/transformers.py:15-29``)
  imported from ``/root/reference`` and records its outputs in ``tests/golden/pool_ref.npz``;
  ``tests/test_oracle.py`` checks :func:`oracle.pool_ref.masked_mean_pool` against them.
* retrieve half, NumPy statement (``assistant/rag/services/search_service.py:17-21,32-35,76-84``):
  PINNED the same way (``tests/golden/search_ref.npz``).
* retrieve half, pgvector ``<=>`` (the C extension PostgreSQL runs for
  ``search_service.py:185-196``): **parity unpinned**.  pgvector is not in
  ``/root/reference`` (server extension version unpinned, client ``pgvector==0.2.5``),
  no PostgreSQL exists in this image and the reference holds no test or golden
  vector for it.  ``oracle/scan_ref.c`` restates its published algorithm.
"""

"""dab_b200 - B200-native (sm_100a) retrieval hot path for django-assistant-bot.

Only the path BASELINE.json names: exact cosine top-k over the chunk-embedding store
(what assistant/rag delegates to pgvector ``<=>``) and the embedding post-process
(masked mean-pool + L2 normalise) that feeds it.  Python here is the host-side mirror of
the reference's interface; the arithmetic lives in libdab_b200.so (hand-written CUDA).
"""
from ._lib import DabError, LIB_PATH, lib  # noqa: F401

__all__ = ["DabError", "LIB_PATH", "lib"]

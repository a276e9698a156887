"""Embedders with the reference's interface, pooling on the B200.

Mirrors ``assistant/ai/providers/base.py:38-45`` (``AIEmbedder``) and
``assistant/ai/embedders/transformers.py:8-29`` (``TransformersEmbedder``): same constructor
arguments, same ``async embeddings(input: List[str]) -> List[List[float]]``.  What changes is
inside: the reference runs one text at a time and pools with
``outputs.last_hidden_state.mean(dim=1)`` (:17-25); here the texts go through the encoder as
padded batches and the post-process is ONE fused CUDA kernel (masked mean + L2 normalise,
``dab_pool_l2norm``).  The encoder forward itself is stock ``transformers`` and out of scope.

The reference's HTTP client ``GPUServiceEmbedder`` (assistant/ai/embedders/gpu_service.py:8-28)
needs no change: it talks to ``dab_b200.gpu_service`` through the same JSON contract
(tests/test_host_api.py drives the reference's own class against this app).  A mirror of it lives here too,
with the binary transport as an option, together with ``get_ai_embdedder`` - the factory of
assistant/ai/services/ai_service.py:51-74, same name (typo included), same ``gpu_service:`` prefix rule.
"""
from __future__ import annotations

import asyncio
import os
import threading
from abc import ABC, abstractmethod
from typing import List, Optional, Sequence

import torch

from . import ops
from ._lib import DabError


class AIEmbedder(ABC):
    """Same abstract surface as assistant/ai/providers/base.py:38-45."""

    @abstractmethod
    async def embeddings(self, input: List[str]) -> List[List[float]]:
        """Get embeddings for the given input."""


class B200TransformersEmbedder(AIEmbedder):
    """Drop-in for ``TransformersEmbedder``.

    ``model`` / ``tokenizer`` may be injected (tests, pre-loaded models); otherwise they are loaded
    exactly like the reference does (``AutoTokenizer/AutoModel.from_pretrained(model_name,
    local_files_only=...)``, transformers.py:9-13).
    ``unit_norm=True`` (default) returns L2-normalised embeddings - the documented delta, harmless
    for cosine; ``unit_norm=False`` returns the reference's raw mean (``out / inv_norm``).
    """

    def __init__(self, model_name: str, local_files_only: bool = True, *, model=None, tokenizer=None,
                 device: Optional[str] = None, max_batch: int = 256, unit_norm: bool = True,
                 max_length: Optional[int] = None, dtype: Optional[torch.dtype] = None):
        """``max_length``: texts are truncated to this many tokens (default: the tokenizer's own
        ``model_max_length``) - the reference tokenises without truncation (transformers.py:17-18), so an
        over-long chunk makes the encoder raise; here it is cut like every batched embedding service does.
        ``dtype=torch.bfloat16`` runs the encoder forward in bf16: its hidden states then enter the pool kernel
        as bf16 (half the bytes of the HBM-bound post-process); accumulation stays fp32."""
        if not torch.cuda.is_available():
            raise DabError(2, "B200TransformersEmbedder needs a CUDA device (no CPU fallback)")
        self._model_name = model_name
        self._device = torch.device(device or "cuda")
        if tokenizer is None or model is None:
            from transformers import AutoModel, AutoTokenizer
            tokenizer = tokenizer or AutoTokenizer.from_pretrained(model_name, local_files_only=local_files_only)
            model = model or AutoModel.from_pretrained(model_name, local_files_only=local_files_only)
        self._tokenizer = tokenizer
        self._model = model.to(self._device).eval()
        if dtype is not None:
            if dtype not in (torch.float32, torch.bfloat16):
                raise DabError(1, "encoder dtype must be torch.float32 or torch.bfloat16")
            self._model = self._model.to(dtype)
        self._max_batch = int(max_batch)
        self._unit_norm = unit_norm
        self._max_length = max_length
        self._forward_lock = threading.Lock()      # one encoder forward at a time per embedder

    def _encode_batch(self, texts: Sequence[str]):
        kw = {}
        limit = self._max_length if self._max_length is not None else getattr(self._tokenizer, "model_max_length", None)
        if limit is not None and 0 < int(limit) < 1_000_000:      # HF uses a huge sentinel for "no limit"
            kw = {"truncation": True, "max_length": int(limit)}
        try:
            inputs = self._tokenizer(list(texts), return_tensors="pt", padding=True, **kw)
        except TypeError:                                          # a minimal tokenizer without truncation support
            inputs = self._tokenizer(list(texts), return_tensors="pt", padding=True)
            if kw:
                inputs = {k: v[:, :kw["max_length"]] for k, v in dict(inputs).items()}
        inputs = {k: v.to(self._device) for k, v in dict(inputs).items()}
        with torch.no_grad():
            hidden = self._model(**inputs).last_hidden_state
        mask = inputs.get("attention_mask")
        return hidden, mask

    def embed_tensor(self, input: Sequence[str]) -> torch.Tensor:
        """[len(input), D] fp32 embeddings on the device (what the store appends)."""
        outs = []
        for i in range(0, len(input), self._max_batch):
            hidden, mask = self._encode_batch(input[i:i + self._max_batch])
            if self._unit_norm:
                outs.append(ops.pool_l2norm(hidden, mask))
            else:
                out, inv = ops.pool_l2norm(hidden, mask, want_inv_norm=True)
                outs.append(out / inv[:, None])
        if not outs:
            return torch.empty((0, 0), device=self._device)
        return torch.cat(outs, 0)

    async def embeddings(self, input: List[str]) -> List[List[float]]:
        """Same contract as the reference; the blocking GPU work runs in a worker thread so the event loop of the
        embedding service keeps answering while a batch is in flight (the reference computes inside ``async def``,
        gpu_service/main.py:75-86 with transformers.py:15-29, which stalls every other request of the worker)."""
        if not input:
            return []

        def work():
            with self._forward_lock:
                return self.embed_tensor(input).tolist()
        return await asyncio.to_thread(work)

    def embed_into(self, store, texts: Sequence[str], ids: Optional[torch.Tensor] = None) -> int:
        """The ingest path in one call: encode, pool + normalise, and append the rows to ``store`` (a
        ``ChunkStore``, or a ``ShardedStore`` whose local shard owns these rows) without leaving the device -
        what ``embedder.embeddings(texts)`` followed by ``bulk_update(['embedding'])`` does through PostgreSQL
        (assistant/processing/documents/steps/embeddings.py:27-39,57-69).  bf16 stores take the kernel's bf16
        output directly.  Returns the number of rows appended."""
        target = getattr(store, "local", store)
        n = 0
        for i in range(0, len(texts), self._max_batch):
            hidden, mask = self._encode_batch(texts[i:i + self._max_batch])
            if ids is None:
                n += target.ingest(hidden, mask)          # pool + append in one kernel
                continue
            if target.dtype == torch.bfloat16:
                _, rows = ops.pool_l2norm(hidden, mask, want_bf16=True)
            else:
                rows = ops.pool_l2norm(hidden, mask)
            target.append(rows, ids[i:i + rows.shape[0]])
            n += rows.shape[0]
        return n


class GPUServiceEmbedder(AIEmbedder):
    """Client side of ``POST /embeddings/`` - mirror of assistant/ai/embedders/gpu_service.py:8-28 (same
    constructor, same request, same exception on a non-200 answer).  ``binary=True`` asks for the raw float32
    transport (``Accept: application/octet-stream``), which skips ~20 bytes of JSON text per float."""

    def __init__(self, base_url: str, model=str, *, binary: bool = False):
        self._base_url = base_url
        self._model = model
        self._binary = binary

    async def embeddings(self, input: List[str]) -> List[List[float]]:
        import aiohttp
        headers = {"Accept": "application/octet-stream"} if self._binary else None
        async with aiohttp.ClientSession() as session:
            async with session.post(f"{self._base_url}/embeddings/", json={"model": self._model, "texts": input},
                                    headers=headers) as response:
                if response.status != 200:
                    raise Exception(f"Failed to get embeddings. "
                                    f"Got status code {response.status} from GPU Service with message {await response.text()}")
                if self._binary:
                    import numpy as np
                    rows, cols = (int(x) for x in response.headers["X-Embedding-Shape"].split(","))
                    return np.frombuffer(await response.read(), dtype="<f4").reshape(rows, cols).tolist()
                return (await response.json())['embeddings']


def _setting(name: str, default=None):
    """``django.conf.settings.<name>`` where Django is configured (the reference reads GPU_SERVICE_ENDPOINT and
    EMBEDDING_AI_MODEL there), else the environment variable of the same name."""
    try:
        from django.conf import settings
        if settings.configured and hasattr(settings, name):
            return getattr(settings, name)
    except Exception:
        pass
    return os.environ.get(name, default)


_local_embedders = {}


def get_ai_embdedder(model: str = None) -> AIEmbedder:
    """The factory of assistant/ai/services/ai_service.py:51-74 for the embedders on this path:

    * ``gpu_service:<name>`` -> ``GPUServiceEmbedder(base_url=settings.GPU_SERVICE_ENDPOINT, model=<name>)``,
      exactly the reference's rule (:61-67);
    * ``b200:<name>`` -> an in-process ``B200TransformersEmbedder`` (one per name, kept), for callers that
      embed inside the process that owns the GPU.

    The reference's OpenAI / Ollama embedders (:56-60, :68-73) are remote services and not part of this path."""
    if not model:
        model = _setting("EMBEDDING_AI_MODEL")
    if not model:
        raise DabError(1, "no embedding model named: pass one or set EMBEDDING_AI_MODEL")
    if model.startswith('gpu_service:'):
        endpoint = _setting("GPU_SERVICE_ENDPOINT")
        if not endpoint:
            raise DabError(1, "GPU_SERVICE_ENDPOINT is not set")
        return GPUServiceEmbedder(base_url=endpoint, model=model[len('gpu_service:'):])
    if model.startswith('b200:'):
        name = model[len('b200:'):]
        if name not in _local_embedders:
            _local_embedders[name] = B200TransformersEmbedder(name)
        return _local_embedders[name]
    raise DabError(1, f"embedding model {model!r}: only the gpu_service: and b200: prefixes are served here "
                      "(OpenAI and Ollama embedders are remote services outside this path)")

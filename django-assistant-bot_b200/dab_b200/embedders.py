"""Embedders with the reference's interface, pooling on the B200.

Mirrors ``assistant/ai/providers/base.py:38-45`` (``AIEmbedder``) and
``assistant/ai/embedders/transformers.py:8-29`` (``TransformersEmbedder``): same constructor
arguments, same ``async embeddings(input: List[str]) -> List[List[float]]``.  What changes is
inside: the reference runs one text at a time and pools with
``outputs.last_hidden_state.mean(dim=1)`` (:17-25); here the texts go through the encoder as
padded batches and the post-process is ONE fused CUDA kernel (masked mean + L2 normalise,
``dab_pool_l2norm``).  The encoder forward itself is stock ``transformers`` and out of scope.

The reference's HTTP client ``GPUServiceEmbedder`` (assistant/ai/embedders/gpu_service.py:8-28)
needs no change: it talks to ``dab_b200.gpu_service`` through the same JSON contract.
"""
from __future__ import annotations

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
                 device: Optional[str] = None, max_batch: int = 256, unit_norm: bool = True):
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
        self._max_batch = int(max_batch)
        self._unit_norm = unit_norm

    def _encode_batch(self, texts: Sequence[str]):
        inputs = self._tokenizer(list(texts), return_tensors="pt", padding=True)
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
        if not input:
            return []
        return self.embed_tensor(input).tolist()

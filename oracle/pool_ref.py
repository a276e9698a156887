"""Oracle for the embedding post-process (pool half of the hot path).

TEST INFRASTRUCTURE ONLY - see ``oracle/__init__.py``.

Follows ``assistant/ai/embedders/transformers.py:15-29`` of the reference: for one
text (batch 1, no padding) the embedding is ``last_hidden_state.mean(dim=1)`` (line 25).
The batched generalisation used by the build is the *masked* mean, which equals the
reference's plain mean for every text once padding positions are excluded, followed
by an L2 normalisation (a documented delta: the only consumer is cosine distance,
``assistant/rag/services/search_service.py:192-194``, which is scale-invariant).
"""
from __future__ import annotations

import numpy as np

EPS_NORM = 1e-12  # same floor torch.nn.functional.normalize uses


def reference_mean(hidden_1td: np.ndarray) -> np.ndarray:
    """transformers.py:25 for one un-padded text: mean over the token axis."""
    h = np.asarray(hidden_1td)
    assert h.ndim == 3 and h.shape[0] == 1
    return h.astype(np.float64).mean(axis=1)[0]


def masked_mean_pool(hidden: np.ndarray, mask: np.ndarray, normalize: bool = True):
    """e[b] = sum_t m[b,t] h[b,t,:] / max(sum_t m[b,t], 1); optionally e / max(|e|, 1e-12).

    Accumulates in float64 (the "ideal" value both the reference's fp32 torch mean and
    the CUDA kernel are compared with at 1e-5 relative) and returns
    ``(embedding float32 [B,D], inv_norm float32 [B])`` where ``inv_norm`` is
    ``1 / max(|mean|, eps)`` of the un-normalised mean.
    An all-masked row yields a zero vector (and inv_norm = 1/eps clamps to 1e12).
    """
    h = np.asarray(hidden)
    m = np.asarray(mask)
    assert h.ndim == 3 and m.shape == h.shape[:2]
    w = (m != 0).astype(np.float64)
    s = np.einsum("btd,bt->bd", h.astype(np.float64), w)
    cnt = np.maximum(w.sum(axis=1), 1.0)
    mean = s / cnt[:, None]
    nrm = np.sqrt((mean * mean).sum(axis=1))
    inv = 1.0 / np.maximum(nrm, EPS_NORM)
    out = mean * inv[:, None] if normalize else mean
    return out.astype(np.float32), inv.astype(np.float32)


def masked_mean_pool_torch_cpu(hidden, mask):
    """The CPU torch statement timed as the pool CPU baseline:
    fp32 masked mean followed by ``F.normalize`` - what the reference's line 25 becomes
    for a padded batch, executed by stock PyTorch on host cores."""
    import torch

    h = hidden.float()
    w = (mask != 0).to(h.dtype)
    s = torch.einsum("btd,bt->bd", h, w)
    mean = s / w.sum(dim=1).clamp_min(1.0)[:, None]
    return torch.nn.functional.normalize(mean, dim=1, eps=EPS_NORM)

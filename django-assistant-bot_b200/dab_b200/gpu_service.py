"""The embeddings endpoint of the GPU service, same HTTP contract as the reference.

Mirrors ``gpu_service/main.py`` of the reference for the hot path only:
``EmbeddingRequest`` (:34-36), the model registry + lifespan loader (:52-72) and
``POST /embeddings/`` (:75-86): lower-cased model lookup, 400 "Model is not supported" for an
unknown model, ``{"embeddings": [[float]]}`` on success, 500 with ``str(e)`` on any failure.
``/dialog/`` (LLM generation) is out of scope.  The embedder behind it pools on the B200
(``dab_b200.embedders.B200TransformersEmbedder``).

One addition, opt-in per request: a client that sends ``Accept: application/octet-stream`` gets the
same embeddings as raw little-endian float32 (``X-Embedding-Shape: rows,dim``) instead of JSON - at
768 floats per text the JSON text is ~6x the bytes and most of the endpoint's CPU time.  Without
that header the reference's JSON contract is unchanged.

Run like the reference (gunicorn_conf.py:9-16):
    gunicorn -k uvicorn.workers.UvicornWorker "dab_b200.gpu_service:create_app()"
"""
from __future__ import annotations

import logging
from contextlib import asynccontextmanager
from typing import Dict, Iterable, List, Optional

import numpy as np
from fastapi import FastAPI, HTTPException, Request
from fastapi.responses import Response
from pydantic import BaseModel

logger = logging.getLogger(__name__)

# gpu_service/models.py:1-3 of the reference
DEFAULT_EMBEDDER_MODELS = ["sberbank-ai/ruBert-base"]
BINARY_MEDIA_TYPE = "application/octet-stream"


def decode_binary_embeddings(content: bytes, shape_header: str) -> np.ndarray:
    """Client side of the binary transport: body + ``X-Embedding-Shape`` -> [rows, dim] float32."""
    rows, dim = (int(x) for x in shape_header.split(","))
    return np.frombuffer(content, dtype="<f4").reshape(rows, dim)


class EmbeddingRequest(BaseModel):
    model: str
    texts: List[str]


def create_app(embedder_models: Optional[Iterable[str]] = None, embedders: Optional[Dict[str, object]] = None,
               embedder_factory=None) -> FastAPI:
    """``embedders``: pre-built ``{name: AIEmbedder}`` (tests, custom loading); otherwise every name
    in ``embedder_models`` is loaded at start-up with ``embedder_factory`` (default
    ``B200TransformersEmbedder``); load failures are logged and skipped, after which that model
    answers 400 - exactly the reference's behaviour (main.py:60-63,77-79)."""
    registry: Dict[str, object] = {}
    if embedders:
        registry.update({k.lower(): v for k, v in embedders.items()})
    names = list(DEFAULT_EMBEDDER_MODELS if embedder_models is None and not embedders else (embedder_models or []))

    @asynccontextmanager
    async def lifespan(app: FastAPI):
        logger.info("App is starting.")
        factory = embedder_factory
        if factory is None and names:
            from .embedders import B200TransformersEmbedder as factory  # noqa: N813
        for model in names:
            try:
                registry[model.lower()] = factory(model)
            except Exception as e:  # same policy as the reference: log, skip, answer 400 later
                logger.exception(f"Failed to load embedder {model}: {e}")
        logger.info("App initialized.")
        yield

    app = FastAPI(lifespan=lifespan)
    app.state.embedders = registry

    @app.post("/embeddings/")
    async def get_embeddings(request: EmbeddingRequest, http: Request):
        model = request.model.lower()
        if model not in registry:
            raise HTTPException(status_code=400, detail="Model is not supported")
        try:
            if BINARY_MEDIA_TYPE in http.headers.get("accept", ""):
                if not request.texts:      # the JSON path answers {"embeddings": []}; same thing in binary
                    return Response(content=b"", media_type=BINARY_MEDIA_TYPE, headers={"X-Embedding-Shape": "0,0"})
                embedder = registry[model]
                if hasattr(embedder, "embed_tensor"):     # device tensor -> one D2H copy, no Python floats
                    arr = embedder.embed_tensor(request.texts).float().cpu().numpy()
                else:
                    arr = np.asarray(await embedder.embeddings(request.texts), dtype="<f4")
                arr = np.ascontiguousarray(arr, dtype="<f4").reshape(len(request.texts), -1)
                return Response(content=arr.tobytes(), media_type=BINARY_MEDIA_TYPE,
                                headers={"X-Embedding-Shape": f"{arr.shape[0]},{arr.shape[1]}"})
            embeddings = await registry[model].embeddings(request.texts)
            return {"embeddings": embeddings}
        except Exception as e:
            logger.exception(f"Failed to get response: {e}")
            raise HTTPException(status_code=500, detail=str(e))

    return app

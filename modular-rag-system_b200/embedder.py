"""Batched GPU embedder: the reference's embedder objects with the pooling tail on the B200.

The reference embeds one text per call (embedding_pipeline.py:582-607) through
``SentenceTransformer.encode`` (embedding_pipeline.py:424, ollama_rag_system.py:163), whose last two
modules — masked mean ``Pooling`` and ``Normalize`` — are K1 of this package (csrc/ingest.cuh).  The
transformer forward stays whatever the caller already runs: it is handed in as ``encode_tokens``.

``PooledEmbedder``            speaks the ``EmbeddingPipeline.embedders`` protocol (embedding_pipeline.py:416-434,
                              :521-538): ``.available`` and ``generate_embedding(text) -> (list, info)``;
                              ``generate_embeddings(texts)`` is the batched form (one K1 launch per batch).
``PooledEmbeddingGenerator``  speaks the vector store's protocol (ollama_rag_system.py:96-197):
                              ``.embedding_cache`` and ``generate_embedding(text) -> list``; ``generate_embeddings``
                              is what ``B200VectorStore.add_documents`` / ``search_many`` call when it exists.
"""
from __future__ import annotations

import time
from typing import Callable, Dict, List, Sequence, Tuple

import numpy as np

from .store import DeviceStore, pool_norm_pack_dev

# encode_tokens(texts) -> (token embeddings [B,S,H] CUDA fp16/bf16/fp32, attention mask [B,S] CUDA or None)
EncodeTokens = Callable[[Sequence[str]], Tuple["object", "object"]]


class PooledEmbedder:
    """Mean-pool + L2-normalise on the device for every batch of texts the pipeline hands over."""

    def __init__(self, encode_tokens: EncodeTokens, model_name: str = "b200-pooled", batch_size: int = 256):
        self.encode_tokens = encode_tokens
        self.model_name = model_name
        self.batch_size = int(batch_size)
        self.available = self._check_availability()

    def _check_availability(self) -> bool:
        try:
            import torch
            return callable(self.encode_tokens) and torch.cuda.is_available()
        except Exception:
            return False

    def pooled_dev(self, texts: Sequence[str], out_dtype: str = "fp32"):
        """[len(texts), H] unit rows on the device (torch tensor), batch_size texts per encoder call."""
        import torch
        if not self.available:
            raise RuntimeError("PooledEmbedder not available (no CUDA device or no encoder)")
        parts = []
        for i in range(0, len(texts), self.batch_size):
            tok, mask = self.encode_tokens(list(texts[i:i + self.batch_size]))
            rows, _ = pool_norm_pack_dev(tok.contiguous(), None if mask is None else mask.contiguous(), out_dtype)
            parts.append(rows)
        if not parts:
            return torch.empty((0, 0))
        return parts[0] if len(parts) == 1 else torch.cat(parts, 0)

    def generate_embeddings(self, texts: Sequence[str]) -> Tuple[np.ndarray, Dict]:
        t0 = time.time()
        rows = self.pooled_dev(texts).float().cpu().numpy()
        return rows, {"model": self.model_name, "processing_time": time.time() - t0,
                      "embedding_dim": int(rows.shape[1]) if rows.size else 0, "backend": "b200_pooled",
                      "batch": len(texts)}

    def generate_embedding(self, text: str) -> Tuple[List[float], Dict]:  # embedding_pipeline.py:416-434
        rows, info = self.generate_embeddings([text])
        info.pop("batch", None)
        return rows[0].tolist(), info

    def embed_into(self, store: DeviceStore, texts: Sequence[str]) -> int:
        """Encoder outputs go through K1 straight into the store tail (no host round trip); returns the first row."""
        first = -1
        for i in range(0, len(texts), self.batch_size):
            tok, mask = self.encode_tokens(list(texts[i:i + self.batch_size]))
            f = store.add_pooled_dev(tok.contiguous(), None if mask is None else mask.contiguous())
            first = f if first < 0 else first
        return first


class PooledEmbeddingGenerator:
    """``EmbeddingGenerator`` stand-in (ollama_rag_system.py:96-197) on top of a ``PooledEmbedder``: same cache-first
    lookup (:116-117), same return type (a list of floats)."""

    def __init__(self, embedder: PooledEmbedder):
        self.embedder = embedder
        self.embedding_cache: Dict[str, List[float]] = {}

    def generate_embeddings(self, texts: Sequence[str]) -> List[List[float]]:
        missing = [t for t in dict.fromkeys(texts) if t not in self.embedding_cache]
        if missing:
            rows, _ = self.embedder.generate_embeddings(missing)
            for t, r in zip(missing, rows):
                self.embedding_cache[t] = r.tolist()
        return [self.embedding_cache[t] for t in texts]

    def generate_embedding(self, text: str) -> List[float]:
        return self.generate_embeddings([text])[0]

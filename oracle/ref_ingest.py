"""CPU restatement of the ingest-side arithmetic (TEST INFRASTRUCTURE).

* ``normalize_f32``      <- ``RAGVectorStore._update_faiss_index``
                            (/root/reference/ollama_rag_system.py:379-380):
                            ``e = np.array(emb, float32); e = e / np.linalg.norm(e)``.
                            PINNED (fixtures produced by those two expressions).
                            One deliberate difference, stated here and in
                            DESIGN.md: the reference has no zero guard (0/0 ->
                            NaN row that FAISS would keep); the store follows
                            the *search* semantics instead (:214-215, a zero
                            norm scores 0.0), so a zero vector packs to a zero
                            row with norm 0.
* ``pool_normalize_f32`` <- the encoder tail inside
                            ``SentenceTransformer.encode`` (call sites
                            embedding_pipeline.py:424, ollama_rag_system.py:163).
                            PARITY UNPINNED: third-party sentence-transformers
                            >=2.2.0 (requirements.txt:11), not vendored, not
                            installed.  Published algorithm of its
                            ``Pooling(pooling_mode_mean_tokens)`` + ``Normalize``:
                              v = sum_s m_s h_s / max(sum_s m_s, 1e-9)
                              v = v / max(||v||_2, 1e-12)
* ``pack_f16/pack_bf16`` <- round-to-nearest-even narrowing (what
                            ``tensor.half()`` / ``.bfloat16()`` do).
"""

from __future__ import annotations

from typing import Tuple

import numpy as np


def normalize_f32(rows: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """fp32 rows -> (unit fp32 rows, fp32 norms).  ollama_rag_system.py:379-380."""
    x = np.asarray(rows, dtype=np.float32)
    x = np.atleast_2d(x)
    norms = np.linalg.norm(x, axis=1).astype(np.float32)
    out = np.zeros_like(x)
    nz = norms > 0
    out[nz] = x[nz] / norms[nz, None]
    return out, norms


def pool_normalize_f32(tok: np.ndarray, mask: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """Masked mean over tokens then L2-normalise, fp32.

    ``tok [B,S,H]`` any float dtype (widened to fp32 first, as a CPU encoder
    would hand fp32 to the pooling module), ``mask [B,S]`` 0/1.
    Returns (unit rows [B,H] fp32, norms-before-normalise [B] fp32).
    """
    h = np.asarray(tok).astype(np.float32)
    m = np.asarray(mask).astype(np.float32)
    summed = (h * m[:, :, None]).sum(axis=1, dtype=np.float32)
    cnt = np.maximum(m.sum(axis=1, dtype=np.float32), np.float32(1e-9))
    v = summed / cnt[:, None]
    norms = np.linalg.norm(v, axis=1).astype(np.float32)
    unit = v / np.maximum(norms, np.float32(1e-12))[:, None]
    return unit.astype(np.float32), norms


def pack_f16(x: np.ndarray) -> np.ndarray:
    """fp32 -> IEEE half, round-to-nearest-even (numpy's astype does RNE)."""
    return np.asarray(x, dtype=np.float32).astype(np.float16)


def pack_bf16(x: np.ndarray) -> np.ndarray:
    """fp32 -> bfloat16 bit patterns (uint16), round-to-nearest-even."""
    u = np.asarray(x, dtype=np.float32).view(np.uint32).astype(np.uint64)
    lsb = (u >> 16) & 1
    r = (u + 0x7FFF + lsb) >> 16
    nan = np.isnan(np.asarray(x, dtype=np.float32))
    r = np.where(nan, 0x7FC0, r)
    return r.astype(np.uint16)


def unpack_bf16(b: np.ndarray) -> np.ndarray:
    """bfloat16 bit patterns (uint16) -> fp32."""
    return (np.asarray(b, dtype=np.uint16).astype(np.uint32) << 16).view(np.float32)

"""PooledEmbedder / PooledEmbeddingGenerator: the reference's embedder protocols with the pooling
tail (K1) on the device.  PARITY UNPINNED for the pooling arithmetic (sentence-transformers is un-vendored): checked
against the restated Pooling(mean)+Normalize of oracle/ref_ingest.py."""
import zlib

import numpy as np
import pytest

from oracle import ref_ingest

pytestmark = pytest.mark.gpu

S, H = 48, 384


def _tokens_of(text):
    """Deterministic stand-in for the transformer forward (out of scope): token embeddings seeded by the text."""
    rng = np.random.default_rng(zlib.crc32(text.encode()))
    n = 1 + len(text) % S
    tok = rng.standard_normal((S, H)).astype(np.float16)
    mask = (np.arange(S) < n).astype(np.int64)
    return tok, mask


def _encode_tokens(texts):
    import torch
    tm = [_tokens_of(t) for t in texts]
    return (torch.from_numpy(np.stack([t for t, _ in tm])).cuda(), torch.from_numpy(np.stack([m for _, m in tm])).cuda())


def _oracle_rows(texts):
    tm = [_tokens_of(t) for t in texts]
    unit, _ = ref_ingest.pool_normalize_f32(np.stack([t for t, _ in tm]).astype(np.float32), np.stack([m for _, m in tm]))
    return unit


@pytest.mark.parametrize("out_dtype,tol", [("fp32", 2e-6), ("fp16", 2.0 ** -11 + 1e-5), ("bf16", 2.0 ** -8 + 1e-5)])
@pytest.mark.parametrize("h", [384, 100])
def test_pool_norm_pack_dev_stateless(out_dtype, tol, h):
    import torch
    from modular_rag_system_b200 import pool_norm_pack_dev
    rng = np.random.default_rng(h)
    B, s = 11, 37
    tok = rng.standard_normal((B, s, h)).astype(np.float32)
    lens = rng.integers(0, s + 1, size=B)          # includes an all-masked chunk when a length is 0
    lens[3] = 0
    mask = (np.arange(s)[None, :] < lens[:, None])
    unit, onorm = ref_ingest.pool_normalize_f32(tok, mask)
    rows, norms = pool_norm_pack_dev(torch.from_numpy(tok).cuda(), torch.from_numpy(mask).cuda(), out_dtype)
    assert rows.shape == (B, h)
    got = rows.float().cpu().numpy()
    assert np.abs(got - unit).max() <= tol
    np.testing.assert_allclose(norms.cpu().numpy(), onorm, rtol=2e-5, atol=1e-12)
    assert np.all(got[3] == 0.0) and norms[3].item() == 0.0


def test_pipeline_protocol_and_batching():
    from modular_rag_system_b200 import PooledEmbedder
    emb = PooledEmbedder(_encode_tokens, model_name="fake-encoder", batch_size=4)
    assert emb.available is True
    texts = [f"chunk number {i} " * (1 + i % 5) for i in range(10)]   # 3 encoder batches
    vec, info = emb.generate_embedding(texts[0])                      # embedding_pipeline.py:416-434
    assert isinstance(vec, list) and len(vec) == H and isinstance(vec[0], float)
    assert info["model"] == "fake-encoder" and info["embedding_dim"] == H and info["backend"] == "b200_pooled"
    assert info["processing_time"] >= 0.0
    rows, info = emb.generate_embeddings(texts)
    assert rows.shape == (10, H) and rows.dtype == np.float32 and info["batch"] == 10
    assert np.abs(rows - _oracle_rows(texts)).max() <= 2e-6
    np.testing.assert_array_equal(rows[0], np.asarray(vec, np.float32))   # batch position does not change a row


def test_embed_into_store_matches_host_route():
    from modular_rag_system_b200 import DeviceStore, PooledEmbedder
    emb = PooledEmbedder(_encode_tokens, batch_size=8)
    texts = [f"t{i}" * (i + 1) for i in range(19)]
    a, b = DeviceStore(H, "fp16", 32), DeviceStore(H, "fp16", 32)
    assert emb.embed_into(a, texts) == 0 and len(a) == 19
    rows32, _ = emb.generate_embeddings(texts)
    b.add(rows32)                                                     # host route: fp32 rows -> normalise+pack
    ra, _ = a.get_rows()
    rb, _ = b.get_rows()
    assert np.abs(ra.astype(np.float32) - rb.astype(np.float32)).max() <= 2.0 ** -10
    assert np.mean(ra == rb) > 0.97


def test_vector_store_uses_the_batched_generator(tmp_path):
    from modular_rag_system_b200 import (B200VectorStore, PooledEmbedder, PooledEmbeddingGenerator, RAGDocument,
                                         RAGQuery)
    calls = []

    def enc(texts):
        calls.append(len(texts))
        return _encode_tokens(texts)

    gen = PooledEmbeddingGenerator(PooledEmbedder(enc, batch_size=64))
    vs = B200VectorStore(tmp_path / "s", embedding_generator=gen, window=None)
    texts = [f"document about topic {i}" for i in range(40)]
    docs = [RAGDocument(id=f"d{i}", content=t, metadata={}, embedding=None, timestamp=float(i), source="s")
            for i, t in enumerate(texts)]
    assert all(vs.add_documents(docs))
    assert calls == [40]                                              # one encoder + pooling pass, not 40
    qs = [RAGQuery(query_text=texts[i], context={}, max_results=3, similarity_threshold=0.5) for i in (7, 21, 39, 0, 5)]
    res = vs.search_many(qs)
    assert calls == [40]                                              # query texts were already in the cache (:116-117)
    for i, r in zip((7, 21, 39, 0, 5), res):
        assert r.documents[0].id == f"d{i}" and abs(r.similarities[0] - 1.0) < 2e-3
    one = vs.search(RAGQuery(query_text="a text never seen", context={}, max_results=3, similarity_threshold=0.0))
    assert calls == [40, 1] and len(one.documents) == 3
    vs.close(save=False)

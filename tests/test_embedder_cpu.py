"""Host logic of the batched embedder (no GPU): protocol surface, cache-first lookup, loud failure without a device."""
import numpy as np
import pytest


class _FakePooled:
    """PooledEmbedder stand-in whose rows are a function of the text (the device kernel is covered by -m gpu)."""
    def __init__(self):
        self.calls = []

    def generate_embeddings(self, texts):
        self.calls.append(list(texts))
        return np.stack([np.full(4, float(len(t)), np.float32) for t in texts]), {}


def test_generator_is_cache_first_and_dedups():
    from modular_rag_system_b200 import PooledEmbeddingGenerator
    fake = _FakePooled()
    gen = PooledEmbeddingGenerator(fake)
    gen.embedding_cache["seen"] = [9.0, 9.0, 9.0, 9.0]            # ollama_rag_system.py:116-117
    out = gen.generate_embeddings(["seen", "ab", "abc", "ab"])
    assert fake.calls == [["ab", "abc"]]                           # cached and repeated texts are not encoded again
    assert out[0] == [9.0] * 4 and out[1] == [2.0] * 4 and out[3] == out[1] and out[2] == [3.0] * 4
    assert gen.generate_embedding("abc") == [3.0] * 4 and len(fake.calls) == 1
    assert isinstance(out[1], list) and isinstance(out[1][0], float)


def test_embedder_refuses_without_a_device():
    import torch
    from modular_rag_system_b200 import PooledEmbedder
    if torch.cuda.is_available():
        pytest.skip("CUDA device present")
    emb = PooledEmbedder(lambda texts: (None, None))
    assert emb.available is False                                  # the pipeline skips it (embedding_pipeline.py:521-538)
    with pytest.raises(RuntimeError):
        emb.generate_embedding("x")                                # no CPU fallback

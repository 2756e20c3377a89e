"""B200-native exact vector retrieval for modular-rag-system (sm_100a only).

``DeviceStore``      device-resident embedding store + exact top-k search (C ABI: include/ragb200.h)
``B200VectorStore``  drop-in for the reference's ``RAGVectorStore`` (ollama_rag_system.py:225-554)
``PooledEmbedder``   batched embedder for the reference's pipeline: encoder tail (mean-pool + normalise) on the device
``ShardedDeviceStore`` row-sharded store over several GPUs of one box, ONE process (C ABI: ragb_mstore_*)
``ShardedSearcher``  row-sharded multi-GPU search, one process per GPU (torch.distributed / NCCL)
"""
from ._lib import RagbError, build, lib  # noqa: F401
from .store import DeviceStore, ShardedDeviceStore, gather_merge_dev, kpad_of, launch_count, merge_keys_dev, pool_norm_pack_dev  # noqa: F401
from .sharded import PeerBuffers, ShardedSearcher, shard_bounds  # noqa: F401
from .embedder import PooledEmbedder, PooledEmbeddingGenerator  # noqa: F401
from .rag_store import B200VectorStore, RAGDocument, RAGQuery, RAGResult, install, uninstall  # noqa: F401

__all__ = ["DeviceStore", "ShardedDeviceStore", "RagbError", "build", "lib", "kpad_of", "launch_count", "merge_keys_dev",
           "ShardedSearcher", "PeerBuffers", "gather_merge_dev", "shard_bounds", "B200VectorStore", "RAGDocument", "RAGQuery", "RAGResult", "install", "uninstall",
           "PooledEmbedder", "PooledEmbeddingGenerator", "pool_norm_pack_dev"]

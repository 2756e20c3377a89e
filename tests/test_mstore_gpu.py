"""The row-sharded store behind the boundary (C ABI ragb_mstore_*, ShardedDeviceStore, B200VectorStore(devices=...)):
one process, rows dealt over the devices in blocks — results must equal the one-device store's, ties and policy
included.  On a 1-GPU box the shards sit on the same device (a device may be listed more than once), which covers the
layout, the global ids, the cross-shard window cut and the merge; with >= 2 GPUs the same tests also run over real
peer memory (the finish kernels store their candidates into the first device's gather buffer over NVLink)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _device_sets():
    import torch
    n = torch.cuda.device_count()
    sets = [[0, 0, 0]]
    if n >= 2:
        sets.append([0, 1])
    if n >= 4:
        sets.append([0, 1, 2, 3])
    if n >= 8:
        sets.append(list(range(8)))
    return sets


def _corpus(n, d, seed):
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, d)).astype(np.float32)
    X[2047] = X[2048] = X[7000] = X[11]     # identical rows on both sides of a block boundary and on different shards
    Q = rng.standard_normal((300, d)).astype(np.float32)
    Q[0] = X[11]
    for i in range(3, 300, 17):
        Q[i] = X[(i * 977) % n] + 0.3 * rng.standard_normal(d).astype(np.float32)
    return X, Q


@pytest.mark.parametrize("dtype", ["fp16", "bf16"])
def test_sharded_store_equals_single_store(dtype):
    from modular_rag_system_b200 import DeviceStore, ShardedDeviceStore
    from modular_rag_system_b200._lib import FORCE_SCAN
    n, d = 21_000, 128
    X, Q = _corpus(n, d, 1)
    one = DeviceStore(d, dtype, n)
    one.add(X)
    for devs in _device_sets():
        ms = ShardedDeviceStore(d, dtype, 4096, devs, block_rows=2048)   # grows; blocks of 2048 rows
        assert ms.block_rows == 2048
        for a, b in ((0, 5), (5, 2047), (2047, 2050), (2050, 9000), (9000, n)):   # appends that straddle blocks
            assert ms.add(X[a:b]) == a
        assert len(ms) == n and sum(ms.device_rows()) == n
        assert max(ms.device_rows()) - min(ms.device_rows()) <= 2048              # within one block of balanced
        r1, n1 = one.get_rows()
        r2, n2 = ms.get_rows()
        assert np.array_equal(r1.view(np.uint16), r2.view(np.uint16)) and np.array_equal(n1, n2)
        for nq, k, kw in ((1, 10, {}), (3, 5, dict(threshold=0.1, clamp01=True)), (40, 10, {}), (300, 10, {}),
                          (300, 100, {}), (130, 30, dict(threshold=0.0, clamp01=True)), (2, 10, dict(flags=FORCE_SCAN))):
            a = one.search(Q[:nq], k, "cosine", **kw)
            b = ms.search(Q[:nq], k, "cosine", **kw)
            assert np.array_equal(a[0], b[0]), (devs, nq, k, kw)
            assert np.array_equal(a[1], b[1]), (devs, nq, k, kw)
        ids, _ = ms.search(Q[:1], 4, "cosine")
        assert ids[0].tolist() == [7000, 2048, 2047, 11]                          # exact ties: larger GLOBAL row first
        valid = (np.arange(n) % 3 != 0).astype(np.uint8)
        a = one.search(Q[:50], 10, "cosine", valid=valid)
        b = ms.search(Q[:50], 10, "cosine", valid=valid)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
        ms.set(7000, -X[11])
        one2 = one.search(Q[:1], 4, "cosine")[0][0].tolist()
        assert ms.search(Q[:1], 3, "cosine")[0][0].tolist() == [2048, 2047, 11] and one2[0] == 7000
        ms.set(7000, X[11])
        ms.close()
    one.close()


def test_sharded_filter_cuts_the_window_over_global_rows():
    from modular_rag_system_b200 import DeviceStore, RagbError, ShardedDeviceStore
    rng = np.random.default_rng(3)
    n, d = 30_000, 64
    X, Q = _corpus(n, d, 2)
    live = rng.random(n) > 0.1
    has = rng.random(n) > 0.05
    src = rng.integers(0, 6, n).astype(np.uint16)
    flags = live.astype(np.uint8) | (has.astype(np.uint8) << 1)
    one = DeviceStore(d, "fp16", n)
    one.add(X)
    one.set_meta(0, flags, src)
    for devs in _device_sets():
        ms = ShardedDeviceStore(d, "fp16", n, devs, block_rows=4096)
        ms.add(X)
        ms.set_meta(0, flags, src)
        ms.set_timestamps(0, np.arange(n, dtype=np.float64))
        for allowed_codes, window in ((None, None), (None, 1000), ([1, 4], 1000), ([2], 5), ([5], 0), (None, 2048),
                                      (None, 4097), ([0, 1, 2, 3], 12_345), ([3], 100_000)):
            allowed = None
            if allowed_codes is not None:
                allowed = np.zeros(6, np.uint8); allowed[allowed_codes] = 1
            t1, t2 = one.build_filter(allowed, window), ms.build_filter(allowed, window)
            assert t1 == t2, (devs, allowed_codes, window)
            a = one.search(Q[:40], 10, "cosine", valid="filter")
            b = ms.search(Q[:40], 10, "cosine", valid="filter")
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]), (devs, allowed_codes, window)
        with pytest.raises(RagbError):                 # out-of-order timestamps need a one-device store (recency ranks)
            ms.add(X[:2])
            ms.set_timestamps(n, np.array([5.0, 4.0]))
        ms.close()
    one.close()


def test_sharded_search_dev_and_profile():
    import torch
    from modular_rag_system_b200 import DeviceStore, ShardedDeviceStore
    n, d = 40_000, 256
    X, Q = _corpus(n, d, 4)
    one = DeviceStore(d, "fp16", n)
    one.add(X)
    want = one.search(Q, 10, "ip")
    for devs in _device_sets():
        ms = ShardedDeviceStore(d, "fp16", n, devs)          # default block: 16384 rows
        ms.add_dev(torch.from_numpy(X).cuda())               # every shard's ingest kernel reads its blocks in place
        qd = torch.from_numpy(Q).cuda()
        ms.profile(True)
        for _ in range(3):                                   # back to back on the caller's stream: contexts are reused in order
            ids, sc = ms.search_dev(qd, 10, "ip")
        torch.cuda.synchronize()
        ms_ms, launches = ms.profile_read()
        assert launches == 3 and ms_ms > 0.0
        assert np.array_equal(ids.cpu().numpy(), want[0]) and np.array_equal(sc.cpu().numpy(), want[1])
        ms.close()
    one.close()


def test_vector_store_over_several_devices(tmp_path):
    """B200VectorStore(devices=[...]) — the drop-in for RAGVectorStore with the rows sharded — returns what the
    one-device drop-in returns: window, source filter, replaced ids, NULL embeddings, ties."""
    from modular_rag_system_b200 import B200VectorStore, RAGDocument, RAGQuery
    rng = np.random.default_rng(8)
    n, d = 5_000, 96
    X = rng.standard_normal((n, d))
    X[4000] = X[10]
    docs = [RAGDocument(id=f"d{i % 4900}", content=f"doc {i}", metadata={"i": i},
                        embedding=None if i % 97 == 0 else X[i].tolist(), timestamp=100.0 + i,
                        source=("a", "b", "c")[i % 3]) for i in range(n)]

    class _E:
        embedding_cache = {}

        def generate_embedding(self, text):
            return self.embedding_cache.get(text)

    qs = []
    for j in range(12):
        _E.embedding_cache[f"q{j}"] = (X[(j * 431) % n] + 0.3 * rng.standard_normal(d)).tolist()
        qs.append(RAGQuery(query_text=f"q{j}", context={}, max_results=(3, 5, 8)[j % 3], similarity_threshold=0.2,
                           source_filters=(None, ["a"], ["b", "c"])[j % 3]))
    _E.embedding_cache["dup"] = X[10].tolist()
    qs.append(RAGQuery(query_text="dup", context={}, max_results=3, similarity_threshold=0.9))
    base = B200VectorStore(tmp_path / "one", embedding_generator=_E(), window=2500)
    assert all(base.add_documents(docs))
    want = [base.search(q) for q in qs]
    for devs in _device_sets():
        vs = B200VectorStore(tmp_path / ("m%d" % len(devs)), embedding_generator=_E(), window=2500, devices=devs)
        for a in range(0, n, 1237):
            assert all(vs.add_documents(docs[a:a + 1237]))
        st = vs.get_stats()
        assert st['gpus'] == len(devs) and st['devices'] == devs and st['device_rows'] == n
        for q, w in zip(qs, want):
            r = vs.search(q)
            assert [x.id for x in r.documents] == [x.id for x in w.documents], (devs, q.query_text)
            assert r.similarities == w.similarities and r.total_documents_searched == w.total_documents_searched
        many = vs.search_many(qs)
        assert [[x.id for x in r.documents] for r in many] == [[x.id for x in w.documents] for w in want]
        assert want[-1].documents[0].metadata["i"] == 4000          # the newer of the two identical vectors
        vs.close()
        vs2 = B200VectorStore(tmp_path / ("m%d" % len(devs)), embedding_generator=_E(), window=2500, devices=devs)
        assert vs2.restored_from == "sidecar"                       # packed rows come back through ragb_mstore_put_rows
        r = vs2.search(qs[0])
        assert [x.id for x in r.documents] == [x.id for x in want[0].documents]
        vs2.close(save=False)
    base.close(save=False)

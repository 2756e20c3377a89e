"""B200VectorStore (drop-in for RAGVectorStore) against the unmodified reference's outputs (golden
fixtures made by oracle/make_golden.py): same documents, same query, same policy -> same documents back."""
import numpy as np
import pytest

from oracle import ref_search
from helpers import SEARCH_FIXTURES, load

pytestmark = pytest.mark.gpu


def _build(fx, tmp_path, **kw):
    from modular_rag_system_b200 import B200VectorStore, RAGDocument
    vs = B200VectorStore(tmp_path / "store", **kw)
    null = set(fx["null_rows"].tolist())
    docs = []
    for i in range(len(fx["X"])):
        docs.append(RAGDocument(id=str(fx["doc_ids"][i]), content=f"doc {i}", metadata={"row": i},
                                embedding=None if i in null else fx["X"][i].astype(np.float64).tolist(),
                                timestamp=float(fx["ts"][i]), source=str(fx["sources"][i])))
    return vs, docs


class _NullEmbedder:
    """generate_embedding returns nothing for documents (NULL embedding rows, :485-492); queries come from the cache."""
    def __init__(self):
        self.embedding_cache = {}

    def generate_embedding(self, text):
        return self.embedding_cache.get(text)


@pytest.mark.parametrize("name", SEARCH_FIXTURES)
def test_matches_reference_outputs(name, tmp_path):
    from modular_rag_system_b200 import RAGQuery
    fx = load(name)
    vs, docs = _build(fx, tmp_path, embedding_generator=_NullEmbedder())
    assert all(vs.add_documents(docs)) or len(docs) == 0
    vs.embedding_generator.embedding_cache["Q"] = fx["q"].astype(np.float64).tolist()
    k, thr = int(fx["k"]), float(fx["thr"])
    res = vs.search(RAGQuery(query_text="Q", context={}, max_results=k, similarity_threshold=thr,
                             source_filters=fx["source_filters"].tolist() or None))
    assert res.total_documents_searched == int(fx["ref_total"])
    assert len(res.documents) == len(res.similarities) <= k
    # map returned ids back to fixture rows (the LAST row carrying an id is the live one)
    last = {str(d): i for i, d in enumerate(fx["doc_ids"].tolist())}
    got_rows = np.array([last[d.id] for d in res.documents], dtype=np.int64)
    ref_rows, ref_sims = fx["ref_rows"], fx["ref_sims"]
    X64 = fx["X"].astype(np.float64)
    q64 = fx["q"].astype(np.float64)

    def score_of(rows):
        den = np.linalg.norm(X64[rows], axis=1) * np.linalg.norm(q64)
        s = np.divide(X64[rows] @ q64, den, out=np.zeros(len(rows)), where=den != 0)
        return np.clip(s, 0.0, 1.0)

    pad = lambda a, v: np.pad(a, (0, k - len(a)), constant_values=v)
    err = ref_search.check_topk(pad(got_rows, -1), pad(np.array(res.similarities), -np.inf), pad(ref_rows, -1),
                                pad(ref_sims, -np.inf), score_of, 1e-3, thr)
    assert err is None, err
    if "edge" in name or "planted" in name:  # separated scores or exact ties: identical order, newest first
        assert got_rows.tolist() == ref_rows.tolist()
    for d, r in zip(res.documents, got_rows):
        assert d.content == f"doc {r}" and d.metadata == {"row": int(r)} and d.access_count == 1
    assert res.query_embedding is not None and res.retrieval_time > 0
    vs.close()


def test_error_conventions_and_stats(tmp_path):
    from modular_rag_system_b200 import B200VectorStore, RAGDocument, RAGQuery
    vs = B200VectorStore(tmp_path / "s")
    r = vs.search(RAGQuery("nothing cached", {}))            # embedder fails -> empty result, no raise (:523-525)
    assert r.documents == [] and r.similarities == []
    assert vs.add_document(RAGDocument(id="a", content="x", metadata={})) is False  # cannot embed -> False (:319-321)
    rng = np.random.default_rng(0)
    X = rng.standard_normal((50, 96))
    assert all(vs.add_documents([RAGDocument(id=f"d{i}", content=str(i), metadata={}, embedding=X[i].tolist(),
                                             source="a" if i % 2 else "b") for i in range(50)]))
    assert vs.add_document(RAGDocument(id="short", content="s", metadata={}, embedding=[1.0, 2.0]))  # wrong length
    vs.embedding_generator.embedding_cache["q"] = X[7].tolist()
    r = vs.search(RAGQuery("q", {}, max_results=3, similarity_threshold=0.7))
    assert [d.id for d in r.documents] == ["d7"] and abs(r.similarities[0] - 1.0) < 1e-3
    r = vs.search(RAGQuery("q", {}, max_results=60, similarity_threshold=0.0))
    assert len(r.documents) == 51 and r.documents[0].id == "d7"   # the wrong-length row scores 0.0 and still returns
    vs.embedding_generator.embedding_cache["bad"] = [1.0, 2.0, 3.0]  # wrong-length query scores 0.0 everywhere
    r = vs.search(RAGQuery("bad", {}, max_results=4, similarity_threshold=0.0))
    assert [d.id for d in r.documents] == ["short", "d49", "d48", "d47"] and r.similarities == [0.0] * 4
    st = vs.get_stats()
    assert st["backend"] == "b200+sqlite" and st["total_documents"] == 51 and st["sources"]["a"] == 25
    assert vs.document_mapping[7] == "d7" and vs.faiss_index is None and vs.db_path.exists()
    vs.close()


def test_reopen_rebuilds_device_store(tmp_path):
    from modular_rag_system_b200 import B200VectorStore, RAGDocument, RAGQuery
    rng = np.random.default_rng(1)
    X = rng.standard_normal((300, 64))
    vs = B200VectorStore(tmp_path / "s", window=None)
    vs.add_documents([RAGDocument(id=f"d{i}", content=str(i), metadata={}, embedding=X[i].tolist(), timestamp=1000.0 + i)
                      for i in range(300)])
    vs.close()
    vs2 = B200VectorStore(tmp_path / "s", window=None)
    vs2.embedding_generator.embedding_cache["q"] = X[123].tolist()
    r = vs2.search(RAGQuery("q", {}, max_results=2, similarity_threshold=0.5))
    assert r.documents[0].id == "d123" and r.total_documents_searched == 300
    vs2.close()


def test_concurrent_writer_and_readers(tmp_path):
    """One ingest thread (the reference's worker, :652) against reader threads (rag_query_interface.py:740)."""
    import threading
    from modular_rag_system_b200 import B200VectorStore, RAGDocument, RAGQuery
    rng = np.random.default_rng(2)
    X = rng.standard_normal((400, 128))
    vs = B200VectorStore(tmp_path / "s", window=None, capacity=64)
    vs.add_documents([RAGDocument(id="d0", content="0", metadata={}, embedding=X[0].tolist())])
    vs.embedding_generator.embedding_cache["q"] = X[0].tolist()
    errors = []

    def writer():
        for i in range(1, 400, 20):
            ok = vs.add_documents([RAGDocument(id=f"d{j}", content=str(j), metadata={}, embedding=X[j].tolist())
                                   for j in range(i, min(400, i + 20))])
            if not all(ok):
                errors.append("add failed")

    def reader():
        for _ in range(30):
            r = vs.search(RAGQuery("q", {}, max_results=1, similarity_threshold=0.9))
            if [d.id for d in r.documents] != ["d0"]:
                errors.append(f"reader saw {[d.id for d in r.documents]}")

    th = [threading.Thread(target=writer)] + [threading.Thread(target=reader) for _ in range(3)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errors, errors[:3]
    assert vs.get_stats()["device_rows"] == 400
    vs.close()


def test_device_filter_equals_numpy_policy():
    """ragb_store_build_filter (source IN ..., newest-`window` matches in row order, live, non-NULL) against the
    same policy written in numpy, over random metadata; the search under valid="filter" equals the search under
    the numpy mask."""
    from modular_rag_system_b200 import DeviceStore
    rng = np.random.default_rng(5)
    n, d = 23_456, 64
    X = rng.standard_normal((n, d)).astype(np.float32)
    s = DeviceStore(d, "fp16", n)
    s.add(X)
    live = rng.random(n) > 0.1
    has = rng.random(n) > 0.05
    src = rng.integers(0, 7, n).astype(np.uint16)
    s.set_meta(0, live.astype(np.uint8) | (has.astype(np.uint8) << 1), src)
    Q = rng.standard_normal((3, d)).astype(np.float32)
    for allowed_codes, window in ((None, None), (None, 1000), ([1, 4], 1000), ([2], 5), ([6], 0), ([0, 1, 2, 3, 4, 5, 6], 20_000),
                                  (None, 2048), (None, 2049), ([3], 100_000)):
        allowed = None
        match = live.copy()
        if allowed_codes is not None:
            allowed = np.zeros(7, np.uint8); allowed[allowed_codes] = 1
            match &= np.isin(src, allowed_codes)
        total = int(match.sum())
        if window is not None and total > window:
            keep = np.flatnonzero(match)[total - window:]
            match = np.zeros(n, bool); match[keep] = True
            total = window
        got_total = s.build_filter(allowed, window)
        assert got_total == total, (allowed_codes, window)
        want = (match & has).astype(np.uint8)
        a = s.search(Q, 10, "cosine", valid="filter")
        b = s.search(Q, 10, "cosine", valid=want)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]), (allowed_codes, window)
    s.add(X[:10])  # new rows: the mask no longer covers the store
    from modular_rag_system_b200 import RagbError
    with pytest.raises(RagbError):
        s.search(Q, 10, "cosine", valid="filter")


def test_search_many_equals_looped_search(tmp_path):
    """The batched front-end returns, per query, what the single-query entry point returns (ids, order, totals)."""
    from modular_rag_system_b200 import B200VectorStore, RAGDocument, RAGQuery
    rng = np.random.default_rng(9)
    n, d = 3000, 96
    X = rng.standard_normal((n, d))
    vs = B200VectorStore(tmp_path / "many", window=None)
    docs = [RAGDocument(id=f"d{i}", content=f"doc {i}", metadata={}, embedding=X[i].tolist(), timestamp=1000.0 + i,
                        source=("a", "b", "c")[i % 3]) for i in range(n)]
    assert all(vs.add_documents(docs))
    qs = []
    for j in range(40):
        text = f"q{j}"
        vs.embedding_generator.embedding_cache[text] = (X[(j * 37) % n] + 0.4 * rng.standard_normal(d)).tolist()
        qs.append(RAGQuery(query_text=text, context={}, max_results=(3, 5, 7)[j % 3], similarity_threshold=(0.0, 0.3)[j % 2],
                           source_filters=(None, ["a"], ["b", "c"])[j % 3]))
    qs.append(RAGQuery(query_text="unknown text", context={}))  # no vector: empty result, no exception
    many = vs.search_many(qs)
    assert len(many) == len(qs)
    for qy, m in zip(qs[:-1], many[:-1]):
        one = vs.search(qy)
        assert [x.id for x in one.documents] == [x.id for x in m.documents]
        assert np.allclose(one.similarities, m.similarities, atol=1e-3)
        assert one.total_documents_searched == m.total_documents_searched
        if not qy.source_filters:
            assert m.documents[0].id == f"d{(int(qy.query_text[1:]) * 37) % n}"
    assert many[-1].documents == [] and many[-1].similarities == []


def test_sidecar_restores_bit_exact_and_yields_to_the_table(tmp_path):
    """close() writes the packed rows; a new store on the same directory restores them with one copy (bit-equal
    rows, same results); if the SQLite table changed behind the sidecar, the table is re-read instead."""
    import sqlite3, pickle, json
    from modular_rag_system_b200 import B200VectorStore, RAGDocument, RAGQuery
    rng = np.random.default_rng(21)
    n, d = 700, 128
    X = rng.standard_normal((n, d))
    vs = B200VectorStore(tmp_path / "sc", window=None)
    docs = [RAGDocument(id=f"d{i}", content=f"doc {i}", metadata={"i": i}, embedding=X[i].tolist() if i % 50 else None,
                        timestamp=10.0 + i, source=("a", "b")[i % 2]) for i in range(n)]
    vs.embedding_generator.embedding_cache.update({f"doc {i}": None for i in range(n)})
    vs.embedding_generator = type("E", (), {"embedding_cache": {}, "generate_embedding": lambda self, t: self.embedding_cache.get(t)})()
    assert all(vs.add_documents(docs))
    assert vs.add_document(RAGDocument(id="d7", content="doc 7 again", metadata={}, embedding=X[8].tolist(), timestamp=5000.0, source="a"))
    vs.embedding_generator.embedding_cache["Q"] = X[8].tolist()
    q = RAGQuery(query_text="Q", context={}, max_results=5, similarity_threshold=0.0, source_filters=["a"])
    before = vs.search(q)
    rows0, norms0 = vs._store.get_rows()
    vs.close()
    vs2 = B200VectorStore(tmp_path / "sc", window=None, embedding_generator=vs.embedding_generator)
    assert vs2.restored_from == "sidecar"
    rows1, norms1 = vs2._store.get_rows()
    assert np.array_equal(rows0.view(np.uint16), rows1.view(np.uint16)) and np.array_equal(norms0, norms1)
    after = vs2.search(q)
    assert [x.id for x in after.documents] == [x.id for x in before.documents] and after.similarities == before.similarities
    assert after.documents[0].id == "d7" and after.total_documents_searched == before.total_documents_searched
    vs2.close()
    with sqlite3.connect(tmp_path / "sc" / "rag_documents.db") as conn:  # somebody else (the reference) adds a row
        conn.execute("INSERT INTO documents VALUES (?,?,?,?,?,?,?,?)",
                     ("new", "new doc", json.dumps({}), pickle.dumps((X[8] * 3).tolist()), 9000.0, "a", 1.0, 0))
    vs3 = B200VectorStore(tmp_path / "sc", window=None, embedding_generator=vs.embedding_generator)
    assert vs3.restored_from == "sqlite"
    res = vs3.search(q)
    assert res.documents[0].id == "new" and res.documents[1].id == "d7"  # equal cosine 1.0: newer first
    vs3.close(save=False)


def test_out_of_order_timestamps_on_the_device():
    """ragb_store_set_timestamps: the newest-`window` cut and the order among equal scores follow the TIMESTAMP
    (ollama_rag_system.py:474, :512), evaluated on the device (csrc/rank.cuh: recency ranks) — against the same policy
    written in numpy.  Later row wins a timestamp tie; rows without a timestamp count as the newest, in row order."""
    from modular_rag_system_b200 import DeviceStore
    rng = np.random.default_rng(17)
    n, d = 9_000, 64
    X = rng.standard_normal((n, d)).astype(np.float32)
    X[100] = X[7000] = X[8999] = X[4]          # four identical rows: equal scores for every query
    s = DeviceStore(d, "fp16", n + 64)
    s.add(X)
    ts = rng.permutation(n).astype(np.float64)  # thoroughly out of order
    ts[200:260] = ts[200]                       # a run of equal timestamps
    ts[4], ts[100], ts[7000], ts[8999] = 5e6, 7e6, 6e6, 6e6
    s.set_timestamps(0, ts)
    live = rng.random(n) > 0.1
    has = rng.random(n) > 0.05
    live[[4, 100, 7000, 8999]] = has[[4, 100, 7000, 8999]] = True
    src = rng.integers(0, 5, n).astype(np.uint16)
    s.set_meta(0, live.astype(np.uint8) | (has.astype(np.uint8) << 1), src)
    rank = np.empty(n, np.int64)
    rank[np.lexsort((np.arange(n), ts))] = np.arange(n)
    Q = rng.standard_normal((5, d)).astype(np.float32)
    Q[0] = X[4]
    for allowed_codes, window in ((None, None), (None, 1000), ([1, 3], 500), ([2], 7), (None, 2048), (None, 5000)):
        allowed = None
        match = live.copy()
        if allowed_codes is not None:
            allowed = np.zeros(5, np.uint8); allowed[allowed_codes] = 1
            match &= np.isin(src, allowed_codes)
        total = int(match.sum())
        if window is not None and total > window:
            rows = np.flatnonzero(match)
            keep = rows[np.argsort(rank[rows])[total - window:]]
            match = np.zeros(n, bool); match[keep] = True
            total = window
        assert s.build_filter(allowed, window) == total, (allowed_codes, window)
        want = (match & has).astype(np.uint8)
        a = s.search(Q, 10, "cosine", valid="filter")
        b = s.search(Q, 10, "cosine", valid=want)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]), (allowed_codes, window)
    # equal scores: newest timestamp first, the later row among equal timestamps — on every kernel path
    from modular_rag_system_b200._lib import FORCE_GEMM, FORCE_SCAN
    for fl in (FORCE_SCAN, FORCE_GEMM, 0):
        ids, sc = s.search(np.repeat(Q[:1], 6 if fl == 0 else 1, axis=0), 4, "cosine", flags=fl)
        assert ids[0].tolist() == [100, 8999, 7000, 4], (fl, ids[0])
    ids, _ = s.search(np.repeat(Q[:1], 200, axis=0), 4, "cosine")          # SM-pair kernel + finish
    assert (ids == np.array([100, 8999, 7000, 4])).all()
    # rows appended afterwards without a timestamp are the newest
    s.add(X[4:5])
    ids, _ = s.search(Q[:1], 5, "cosine")
    assert ids[0].tolist() == [n, 100, 8999, 7000, 4]
    # back in order: ranks are dropped again (identity)
    s.set_timestamps(0, np.arange(n + 1, dtype=np.float64))
    ids, _ = s.search(Q[:1], 5, "cosine")
    assert ids[0].tolist() == [n, 8999, 7000, 100, 4]
    s.close()

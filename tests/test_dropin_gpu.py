"""The drop-in, executed through the reference's REAL classes on a GPU (north_star: "a drop-in behind the existing
Python retrieve/search and add-documents entry points in enhanced_rag_system.py, embedding_pipeline.py and
ollama_rag_system.py").

``oracle/_ref/`` holds the three reference modules byte for byte (oracle/fetch_ref.py, SHA-256 checked); they are
imported here unmodified.  Every test builds the reference's own objects — ``RAGSystem``, ``EnhancedRAGSystem``,
``EmbeddingPipeline`` — once as shipped (SQLite + per-row Python cosine, ollama_rag_system.py:454-525) and once
after ``install()`` (device store), feeds both the same documents and cached query vectors, and compares what the
reference's callers see, field by field.
"""
import contextlib
import io
import pickle
import sqlite3
import sys
import time
import zlib

import numpy as np
import pytest

import oracle
from oracle import ref_ingest

pytestmark = pytest.mark.gpu

D = 384


@pytest.fixture()
def ref_modules():
    """The unmodified reference modules, importable for the duration of one test; install() undone afterwards."""
    path = oracle.ref_path()
    if path is None:
        pytest.skip("oracle/_ref is missing (python oracle/fetch_ref.py in the authoring container builds it)")
    names = ("ollama_rag_system", "embedding_pipeline", "enhanced_rag_system")
    saved = {n: sys.modules.pop(n, None) for n in names}
    sys.path.insert(0, path)
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            import ollama_rag_system as ref
            import embedding_pipeline as pipe
            import enhanced_rag_system as enh
        assert ref.__file__.startswith(path)
        yield ref, pipe, enh
    finally:
        import modular_rag_system_b200 as pkg
        pkg.uninstall()
        sys.path.remove(path)
        for n in names:
            sys.modules.pop(n, None)
            if saved[n] is not None:
                sys.modules[n] = saved[n]


def _corpus(n, seed=0, clusters=40):
    """n documents in `clusters` tight clusters (so that several documents pass the reference's 0.7 threshold), a few
    exact duplicates (ties -> newest first), one zero vector, distinct contents."""
    rng = np.random.default_rng(seed)
    centers = rng.standard_normal((clusters, D))
    X = centers[rng.integers(0, clusters, n)] + 0.35 * rng.standard_normal((n, D))
    X[n // 2] = X[n // 2 - 7]
    X[n - 3] = X[n // 2 - 7]            # three identical vectors
    X[11] = 0.0
    texts = [f"document {i}: lorem {zlib.crc32(str(i).encode()) % 9973}" for i in range(n)]
    sources = [("keystroke", "ocr", "document")[i % 3] for i in range(n)]
    return X, texts, sources, centers


def _plant(gen, texts, X):
    for t, x in zip(texts, X):
        gen.embedding_cache[t] = [float(v) for v in x]


def _same_result(a, b, ref, k=5, thr=0.7, tol=1e-3, what=""):
    """a: the unpatched reference's result, b: the result after install().  north_star's criterion: ids identical
    excluding ties within the score tolerance (fp16 rows: two documents whose float64 cosines differ by less than the
    rounding may swap places or swap across the cut), scores within 1e-3, everything else equal."""
    assert type(b) is ref.RAGResult and type(a) is ref.RAGResult, what
    assert b.total_documents_searched == a.total_documents_searched, what
    assert len(b.documents) == len(b.similarities) and len(b.documents) <= k, what
    assert b.query_embedding == a.query_embedding and b.retrieval_time > 0.0
    sa = {d.content: float(s) for d, s in zip(a.documents, a.similarities)}
    sb = {d.content: float(s) for d, s in zip(b.documents, b.similarities)}
    assert all(x >= y - 1e-7 for x, y in zip(b.similarities, b.similarities[1:])), what
    cut = float(a.similarities[-1]) if len(a.documents) == k else thr
    for c, s in sb.items():
        if c in sa:
            assert abs(sa[c] - s) <= tol, f"{what}: {c!r} {sa[c]} vs {s}"
        else:
            assert abs(s - cut) <= 2 * tol, f"{what}: extra document {c!r} at {s}, cut {cut}"
    for c, s in sa.items():
        assert c in sb or abs(s - cut) <= 2 * tol, f"{what}: document {c!r} ({s}) missing, cut {cut}"
    for j, (x, y) in enumerate(zip(a.similarities, b.similarities)):
        assert abs(x - y) <= 2 * tol, f"{what}: position {j}: {x} vs {y}"
    da = {d.content: d for d in a.documents}
    for db in b.documents:
        assert type(db) is ref.RAGDocument
        if db.content in da:
            d0 = da[db.content]
            assert (db.source, db.metadata, db.importance_score, db.access_count) == \
                   (d0.source, d0.metadata, d0.importance_score, d0.access_count), what
            assert db.embedding == d0.embedding


def test_rag_system_query_unpatched_vs_installed(ref_modules, tmp_path):
    """RAGSystem.query (ollama_rag_system.py:683-721) over 1 100 documents: more than the newest-1000 window (:474),
    so the window, the 0.7 default threshold (:595), source_filters (:466-469) and max_results all take part."""
    ref, _, _ = ref_modules
    import modular_rag_system_b200 as pkg
    n = 1100
    X, texts, sources, centers = _corpus(n)
    with contextlib.redirect_stdout(io.StringIO()):
        cfg = ref.RAGSystem(tmp_path / "cfgprobe")._default_config()
    cfg['processing']['auto_process'] = False

    def build(root):
        rs = ref.RAGSystem(root, config=cfg)
        _plant(rs.vector_store.embedding_generator, texts, X)
        for i in range(n):
            rs.add_document(texts[i], metadata={"i": i}, source=sources[i], importance=1.0 + (i % 4))
        return rs

    with contextlib.redirect_stdout(io.StringIO()):
        plain = build(tmp_path / "plain")
        assert type(plain.vector_store).__name__ == "RAGVectorStore" and not hasattr(plain, "add_documents")
        pkg.install()
        fast = build(tmp_path / "fast")
    assert isinstance(fast.vector_store, pkg.B200VectorStore)
    assert fast.vector_store.get_stats()['device_rows'] == n
    rng = np.random.default_rng(5)
    queries = {f"q{j}": centers[j] + 0.2 * rng.standard_normal(D) for j in range(8)}
    queries["dup"] = X[n // 2 - 7]                    # exact ties among three documents: newest first (:474, :512)
    queries["far"] = rng.standard_normal(D)           # nothing reaches 0.7: empty result
    queries["old"] = X[5]                             # its best match lies outside the newest-1000 window
    for rs in (plain, fast):
        _plant(rs.vector_store.embedding_generator, list(queries), list(queries.values()))
    with contextlib.redirect_stdout(io.StringIO()):
        for name in queries:
            for kw in (dict(), dict(max_results=3), dict(max_results=12, source_filters=["ocr", "keystroke"])):
                _same_result(plain.query(name, **kw), fast.query(name, **kw), ref, k=kw.get("max_results", 5),
                             what=f"{name} {kw}")
        r = fast.query("dup", max_results=3)
        assert [d.metadata["i"] for d in r.documents] == [n - 3, n // 2, n // 2 - 7]
        far = fast.query("far")
        assert len(far.documents) == 0 and far.total_documents_searched == 1000
        assert all(d.metadata["i"] != 5 for d in fast.query("old", max_results=12).documents)
    assert fast.query_stats['total_queries'] == plain.query_stats['total_queries'] + 3
    fast.vector_store.close(save=False)


def test_add_document_worker_thread_and_add_documents(ref_modules, tmp_path):
    """RAGSystem.add_document with the default auto_process=True: documents go through the reference's own queue and
    daemon worker (:638-681) into the device store while the caller thread queries; then add_documents, the entry
    point four call sites assume (enhanced_rag_system.py:337,1194; embedding_pipeline.py:806; run_integrated_rag.py:284)."""
    ref, _, _ = ref_modules
    import modular_rag_system_b200 as pkg
    X, texts, sources, _ = _corpus(60, seed=3)
    with contextlib.redirect_stdout(io.StringIO()):
        pkg.install(window=None)
        rs = ref.RAGSystem(tmp_path / "w")
        _plant(rs.vector_store.embedding_generator, texts, X)
        ids = [rs.add_document(texts[i], {"i": i}, sources[i]) for i in range(40)]
        assert all(ids) and rs.processing_thread is not None
        deadline = time.time() + 30
        while rs.vector_store.get_stats()['device_rows'] < 40 and time.time() < deadline:
            rs.query(texts[0])                           # reader against the writer thread
            time.sleep(0.05)
        assert rs.vector_store.get_stats()['device_rows'] == 40
        docs = [ref.RAGDocument(id="", content=texts[i], metadata={"i": i}, source=sources[i],
                                embedding=[float(v) for v in X[i]]) for i in range(40, 60)]
        got = rs.add_documents(docs)
        assert got == [d.id for d in docs] and all(got)
        for i in (0, 17, 39, 40, 59):
            r = rs.query(texts[i], max_results=1)
            assert r.documents[0].metadata["i"] == i and abs(r.similarities[0] - 1.0) < 1e-3
        many = rs.query_many([texts[i] for i in range(60)], max_results=1)
        # document 11 is the zero vector (cosine 0.0 everywhere: nothing reaches 0.7, :214-215); 23, 30 and 57 are
        # identical vectors: the newest wins the tie (:474, :512)
        want = [None if i == 11 else 57 if i in (23, 30) else i for i in range(60)]
        assert [r.documents[0].metadata["i"] if r.documents else None for r in many] == want
        rs.shutdown()
        st = rs.get_system_stats()
    assert st['vector_store']['total_documents'] == 60 and st['vector_store']['backend'] == 'b200+sqlite'
    rs.vector_store.close(save=False)


def _tokens_of(text, S=48):
    """Deterministic stand-in for the transformer forward (out of scope): token embeddings seeded by the text."""
    rng = np.random.default_rng(zlib.crc32(text.encode()))
    tok = rng.standard_normal((S, D)).astype(np.float16)
    mask = (np.arange(S) < 1 + len(text) % S).astype(np.int64)
    return tok, mask


def _encode_tokens(texts):
    import torch
    tm = [_tokens_of(t) for t in texts]
    return (torch.from_numpy(np.stack([t for t, _ in tm])).cuda(), torch.from_numpy(np.stack([m for _, m in tm])).cuda())


def test_embedding_pipeline_with_pooled_embedder(ref_modules, tmp_path):
    """EmbeddingPipeline (embedding_pipeline.py:508-553) with a PooledEmbedder at embedders[0]: process_content
    (:555-635) chunks the text and every chunk's vector is what K1 produces (masked mean + L2 normalise, checked
    against the restated Pooling+Normalize); create_rag_documents -> add_documents is the ingest path of
    KeystrokeEmbeddingProcessor._process_batch (:770-812) and process_uploaded_file."""
    ref, pipe, _ = ref_modules
    import modular_rag_system_b200 as pkg
    with contextlib.redirect_stdout(io.StringIO()):
        pkg.install(window=None)
        rs = ref.RAGSystem(tmp_path / "p")
        ep = pipe.EmbeddingPipeline(tmp_path / "p")
    ep.embedders.insert(0, pkg.PooledEmbedder(_encode_tokens, model_name="fake-encoder"))
    text = "\n\n".join(f"Paragraph {i}. " + "The quick brown fox jumps over the lazy dog. " * 30 for i in range(6))
    results = ep.process_content(text, "document", {"file": "a.txt"})
    assert len(results) >= 2 and all(r.model_used == "fake-encoder" for r in results)
    for r in results:
        tok, mask = _tokens_of(r.request.content)
        unit, _ = ref_ingest.pool_normalize_f32(tok[None].astype(np.float32), mask[None])
        assert np.abs(np.asarray(r.embedding, np.float32) - unit[0]).max() <= 2e-6
    with contextlib.redirect_stdout(io.StringIO()):
        docs = ep.create_rag_documents(results)           # D3 glue: the reference's own call raises TypeError
        assert all(type(d) is ref.RAGDocument and d.id for d in docs)
        assert rs.add_documents(docs) == [d.id for d in docs]
        # the keystroke path: batch_process -> create_rag_documents -> rag_system.add_documents (:796-806)
        kp = pipe.KeystrokeEmbeddingProcessor(ep, rs)
        for j in range(5):
            kp.add_keystroke_data(f"typed text number {j} " * 8, {"app": "editor", "j": j})
    n_rows = rs.vector_store.get_stats()['device_rows']
    assert n_rows >= len(docs) + 5
    gen = rs.vector_store.embedding_generator
    q = results[1].request.content
    gen.embedding_cache[q] = results[1].embedding
    with contextlib.redirect_stdout(io.StringIO()):
        r = rs.query(q, max_results=2)
    assert r.documents[0].content == q and abs(r.similarities[0] - 1.0) < 1e-3
    assert r.documents[0].metadata["file"] == "a.txt" and r.documents[0].source == "document"
    rs.vector_store.close(save=False)


def test_enhanced_rag_system_paths(ref_modules, tmp_path):
    """EnhancedRAGSystem (enhanced_rag_system.py, the v2 definitions that win on import): process_uploaded_file
    (:1168-1203) reaches base_rag.add_documents, enhanced_query (:1205-1237) reaches base_rag.query — and returns the
    documents the unpatched RAGVectorStore returns for the same table."""
    ref, pipe, enh = ref_modules
    import modular_rag_system_b200 as pkg
    with contextlib.redirect_stdout(io.StringIO()):
        pkg.install()
        ers = enh.EnhancedRAGSystem(tmp_path / "e", upload_port=18089)
    assert isinstance(ers.base_rag.vector_store, pkg.B200VectorStore)
    f = ers.upload_folder / "txt" / "notes.txt"
    f.write_text("\n\n".join(f"Note {i}. " + "Retrieval augmented generation keeps context close. " * 25 for i in range(5)))
    with contextlib.redirect_stdout(io.StringIO()):
        for up in ers.file_manager.upload_folder_contents(ers.upload_folder):
            ers.process_uploaded_file(up)                 # FallbackEmbedder vectors (384-d), real chunker
    store = ers.base_rag.vector_store
    n_rows = store.get_stats()['device_rows']
    assert n_rows >= 2, "process_uploaded_file did not reach add_documents"
    # a query vector: the stored embedding of one chunk (the query-side generator is out of scope: cache, :116-117)
    with sqlite3.connect(store.db_path) as conn:
        rows = conn.execute("SELECT content, embedding FROM documents ORDER BY rowid").fetchall()
    content, blob = rows[len(rows) // 2]
    store.embedding_generator.embedding_cache["what do my notes say?"] = pickle.loads(blob)
    with contextlib.redirect_stdout(io.StringIO()):
        out = ers.enhanced_query("what do my notes say?", current_app="editor")
        fast = out['rag_results']
        # the same table through the reference's own store
        pkg.uninstall()
        plain = ref.RAGVectorStore(store.storage_path)
        assert type(plain).__name__ == "RAGVectorStore"
        plain.embedding_generator.embedding_cache["what do my notes say?"] = pickle.loads(blob)
        want = plain.search(ref.RAGQuery(query_text="what do my notes say?", context={}, max_results=5,
                                         similarity_threshold=0.7))
    # (the five notes differ in one digit: their FallbackEmbedder vectors tie within the fp16 rounding)
    hit = [s for d, s in zip(fast.documents, fast.similarities) if d.content == content]
    assert len(hit) == 1 and abs(hit[0] - 1.0) < 1e-3
    _same_result(want, fast, ref, k=5, what="enhanced_query")
    assert out['query'] == "what do my notes say?" and 'enhancement_insights' in out
    store.close(save=False)

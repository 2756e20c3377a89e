#!/usr/bin/env python3
"""Generate tests/golden/*.npz by running the UNMODIFIED reference (TEST INFRA).

Run in the authoring container only (needs /root/reference):

    python oracle/make_golden.py

Each search fixture drives ``RAGVectorStore`` from
/root/reference/ollama_rag_system.py exactly as SURVEY.md §8(c) describes:
documents enter through ``add_document(RAGDocument(embedding=...))`` (:307), the
query vector is planted in ``embedding_generator.embedding_cache`` (:116-117)
and ``search(RAGQuery(...))`` (:388) runs the SQLite brute-force path (:454-525).
The fixture stores the inputs (float32 values, handed to the reference as
Python floats), row timestamps, and the reference's returned ids / similarities.
"""

from __future__ import annotations

import contextlib
import io
import sys
import tempfile
from pathlib import Path

import numpy as np

REF = "/root/reference"
OUT = Path(__file__).resolve().parent.parent / "tests" / "golden"


def _load_reference():
    sys.path.insert(0, REF)
    with contextlib.redirect_stdout(io.StringIO()):
        import ollama_rag_system as R  # noqa
    return R


def run_reference(R, X, ts, doc_ids, q, k, thr, sources=None, source_filters=None, null_rows=()):
    """Returns (row indices in X, float64 similarities, total_documents_searched)."""
    with tempfile.TemporaryDirectory() as td, contextlib.redirect_stdout(io.StringIO()):
        vs = R.RAGVectorStore(Path(td))
        for i in range(len(X)):
            emb = None if i in null_rows else X[i].astype(np.float64).tolist()
            doc = R.RAGDocument(id=doc_ids[i], content=f"chunk {i}", metadata={"i": i}, embedding=emb,
                                timestamp=float(ts[i]), source=(sources[i] if sources is not None else "s"))
            if emb is None:
                # keep the row NULL: bypass add_document's auto-embed (:311-312)
                assert vs._add_document_sqlite(doc)
            else:
                assert vs.add_document(doc)
        vs.embedding_generator.embedding_cache["Q"] = np.asarray(q, dtype=np.float64).tolist()
        res = vs.search(R.RAGQuery(query_text="Q", context={}, max_results=k, similarity_threshold=thr,
                                   source_filters=source_filters))
    # last writer wins for duplicate ids (INSERT OR REPLACE, :351)
    last = {}
    for i, d in enumerate(doc_ids):
        last[d] = i
    rows = np.array([last[d.id] for d in res.documents], dtype=np.int64)
    sims = np.array([float(s) for s in res.similarities], dtype=np.float64)
    return rows, sims, res.total_documents_searched


def main():
    R = _load_reference()
    OUT.mkdir(parents=True, exist_ok=True)
    rng = np.random.default_rng(20261104)
    cases = []

    def unit(n, d):
        x = rng.standard_normal((n, d)).astype(np.float32)
        return x

    # --- random corpora, several shapes / k / thresholds ------------------- #
    for name, n, d, k, thr in [
        ("rand_n1000_d384_k5", 1000, 384, 5, 0.0),
        ("rand_n1000_d128_k10", 1000, 128, 10, 0.0),
        ("rand_n300_d768_k7", 300, 768, 7, 0.05),
        ("rand_n7_d384_k5", 7, 384, 5, 0.0),
        ("rand_n1_d384_k5", 1, 384, 5, 0.0),
        ("rand_n0_d384_k5", 0, 384, 5, 0.0),
    ]:
        X = unit(n, d)
        q = rng.standard_normal(d).astype(np.float32)
        ts = 1.0e9 + np.arange(n, dtype=np.float64)
        cases.append(dict(name=name, X=X, q=q, ts=ts, k=k, thr=thr))

    # --- planted neighbours, reference default policy (thr 0.7, k 5) ------- #
    n, d = 600, 384
    X = unit(n, d)
    q = rng.standard_normal(d).astype(np.float32)
    for j, r in enumerate([17, 123, 256, 400, 599, 3, 77]):
        noise = rng.standard_normal(d).astype(np.float32)
        X[r] = q * (1.0 + 0.1 * j) + (0.15 + 0.12 * j) * noise  # cosines from ~0.99 down to ~0.6
    cases.append(dict(name="planted_n600_d384_k5_thr07", X=X, q=q, ts=1.0e9 + np.arange(n), k=5, thr=0.7))
    cases.append(dict(name="planted_n600_d384_k10_thr07", X=X, q=q, ts=1.0e9 + np.arange(n), k=10, thr=0.7))

    # --- edge semantics: zero rows, negatives, exact ties, scale invariance - #
    d = 128
    X = unit(40, d)
    q = rng.standard_normal(d).astype(np.float32)
    X[5] = 0.0                       # zero norm -> 0.0 (:214-215)
    X[9] = -q                        # cosine -1 -> clamped 0.0 (:218)
    X[11] = 2.0 * q                  # scale invariant -> 1.0
    X[20] = q                        # exact tie with 11 and 30 (newest first)
    X[30] = q
    X[33] = X[2]                     # duplicate vector -> exact tie at a lower score
    cases.append(dict(name="edge_n40_d128_k5_thr0", X=X, q=q, ts=1.0e9 + np.arange(40), k=5, thr=0.0))
    cases.append(dict(name="edge_n40_d128_k40_thr0", X=X, q=q, ts=1.0e9 + np.arange(40), k=40, thr=0.0))
    cases.append(dict(name="edge_n40_d128_k7_thr09", X=X, q=q, ts=1.0e9 + np.arange(40), k=7, thr=0.9))
    # zero query: every row scores 0.0, newest first
    cases.append(dict(name="edge_zeroq_n40_d128_k5", X=X, q=np.zeros(d, np.float32), ts=1.0e9 + np.arange(40), k=5, thr=0.0))
    # NULL embedding rows are skipped (:485-493)
    cases.append(dict(name="edge_null_n40_d128_k5", X=X, q=q, ts=1.0e9 + np.arange(40), k=5, thr=0.0, null_rows=[11, 30]))
    # non-monotone timestamps: tie order follows timestamp, not insertion
    ts = 1.0e9 + np.arange(40, dtype=np.float64)
    ts[[11, 20, 30]] = [1.0e9 + 500, 1.0e9 + 700, 1.0e9 + 600]
    cases.append(dict(name="edge_ts_n40_d128_k5", X=X, q=q, ts=ts, k=5, thr=0.0))
    # duplicate id: the later add replaces the earlier row (:351)
    ids = [f"d{i}" for i in range(40)]
    ids[30] = "d20"
    cases.append(dict(name="edge_dupid_n40_d128_k5", X=X, q=q, ts=1.0e9 + np.arange(40), k=5, thr=0.0, doc_ids=ids))
    # source filter (:466-469)
    src = np.array(["a" if i % 3 else "b" for i in range(40)])
    cases.append(dict(name="edge_src_n40_d128_k5", X=X, q=q, ts=1.0e9 + np.arange(40), k=5, thr=0.0,
                      sources=src, source_filters=["b"]))
    # window: only the newest 1000 rows are scanned (:474)
    n, d = 1100, 64
    X = unit(n, d)
    q = rng.standard_normal(d).astype(np.float32)
    X[50] = q  # outside the newest 1000 -> must NOT be returned by the reference
    cases.append(dict(name="window_n1100_d64_k5", X=X, q=q, ts=1.0e9 + np.arange(n), k=5, thr=0.0))

    for c in cases:
        X, q, ts = c["X"], c["q"], np.asarray(c["ts"], dtype=np.float64)
        doc_ids = c.get("doc_ids") or [f"d{i}" for i in range(len(X))]
        rows, sims, total = run_reference(R, X, ts, doc_ids, q, c["k"], c["thr"], c.get("sources"),
                                          c.get("source_filters"), set(c.get("null_rows", ())))
        np.savez_compressed(
            OUT / f"search_{c['name']}.npz",
            X=X, q=q, ts=ts, k=np.int64(c["k"]), thr=np.float64(c["thr"]),
            doc_ids=np.array(doc_ids), sources=np.array(c.get("sources", np.array(["s"] * len(X)))),
            source_filters=np.array(c.get("source_filters") or [], dtype=str),
            null_rows=np.array(sorted(c.get("null_rows", [])), dtype=np.int64),
            ref_rows=rows, ref_sims=sims, ref_total=np.int64(total),
        )
        print(f"search_{c['name']}: {len(rows)} hits, total_searched={total}")

    # --- normalise (a4): the two numpy expressions of :379-380 -------------- #
    for d in (128, 384):
        x = (rng.standard_normal((64, d)) * rng.uniform(0.01, 30.0, (64, 1))).astype(np.float32)
        unit_rows = np.stack([(np.array(r.tolist(), dtype=np.float32) / np.linalg.norm(np.array(r.tolist(), dtype=np.float32))) for r in x])
        norms = np.array([np.linalg.norm(np.array(r.tolist(), dtype=np.float32)) for r in x], dtype=np.float32)
        np.savez_compressed(OUT / f"normpack_d{d}.npz", x=x, unit=unit_rows.astype(np.float32), norms=norms)
        print(f"normpack_d{d}")


if __name__ == "__main__":
    main()

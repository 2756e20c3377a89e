"""Full-size parity checker (TEST INFRASTRUCTURE): the float64 oracle against a device store of any size.

At BASELINE.json's full sizes (1M x 384, 10M x 768, 12.5M x 384) the corpus exists only in HBM.  This module pulls
the PACKED rows back in slabs (``DeviceStore.get_rows`` = C ABI ``ragb_store_get_rows``), runs
``ref_search.search_f64_blocked`` — the vectorised restatement of ollama_rag_system.py:199-222 / :480-515, pinned
against the unmodified reference by tests/test_oracle.py — on every slab for a SAMPLE of the queries, folds the
per-slab candidates with the oracle's own ``_reduce_candidates`` (score desc, newer row first, :474 / :512), and
judges the device result per query with ``ref_search.check_topk`` (north_star's criterion: ids identical excluding
ties within the tolerance, scores within tol of the float64 score of the same row).

Used by tests/test_fullsize_parity_gpu.py and by bench.py OUTSIDE its timed regions (fields ``recall_at_k``,
``max_abs_score_err``, ``parity_queries``).  Never imported by the product package.
"""
from __future__ import annotations

from concurrent.futures import ThreadPoolExecutor
from typing import Dict, Optional, Sequence

import numpy as np

from . import ref_ingest, ref_search


def _rows_f64(store, r0: int, n: int) -> np.ndarray:
    rows, _ = store.get_rows(r0, n)
    if store.dtype == "bf16":
        rows = ref_ingest.unpack_bf16(rows)
    return rows.astype(np.float64)


def oracle_topk(store, Q: np.ndarray, k: int, metric: str = "ip", score_rows: Optional[Sequence[np.ndarray]] = None,
                slab_rows: int = 131072, threads: int = 4, n_rows: Optional[int] = None):
    """float64 exact top-k of host queries Q [m, dim] over the rows held by ``store``.

    Returns (ids [m,k], scores [m,k], row_scores) where ``row_scores[i]`` maps every row in ``score_rows[i]`` (the rows
    the device returned for query i) to the oracle's float64 score, or None.  Slabs run on a small thread pool: numpy
    releases the GIL inside the conversions and the GEMM."""
    Q = np.atleast_2d(np.asarray(Q, dtype=np.float32))
    m = Q.shape[0]
    n = len(store) if n_rows is None else int(n_rows)
    want = None
    if score_rows is not None:
        want = [np.asarray(r, dtype=np.int64) for r in score_rows]
        want = [w[w >= 0] for w in want]
    q64 = Q.astype(np.float64)
    qn = np.linalg.norm(q64, axis=1)

    def slab(r0):
        r1 = min(n, r0 + slab_rows)
        X = _rows_f64(store, r0, r1 - r0)
        ids, sc = ref_search.search_f64_blocked(X, Q, k, metric, block_rows=slab_rows)
        ids = np.where(ids >= 0, ids + r0, -1)
        got = None
        if want is not None:
            got = []
            for i, w in enumerate(want):
                sel = w[(w >= r0) & (w < r1)]
                if len(sel) == 0:
                    got.append((sel, np.zeros(0)))
                    continue
                xs = X[sel - r0]
                s = xs @ q64[i]
                if metric == "cosine":
                    den = np.linalg.norm(xs, axis=1) * qn[i]
                    s = np.divide(s, den, out=np.zeros_like(s), where=den != 0)
                got.append((sel, s))
        return ids, sc, got

    best_s = np.full((m, 0), -np.inf)
    best_i = np.full((m, 0), -1, dtype=np.int64)
    row_scores: Optional[list] = None if want is None else [dict() for _ in range(m)]
    with ThreadPoolExecutor(max_workers=max(1, threads)) as pool:
        for ids, sc, got in pool.map(slab, range(0, n, slab_rows)):
            best_s = np.concatenate([best_s, sc], axis=1)
            best_i = np.concatenate([best_i, ids], axis=1)
            if best_s.shape[1] > 8 * k:
                best_s, best_i = ref_search._reduce_candidates(best_s, best_i, k)
            if got is not None:
                for i, (sel, s) in enumerate(got):
                    row_scores[i].update(zip(sel.tolist(), s.tolist()))
    best_s, best_i = ref_search._reduce_candidates(best_s, best_i, k)
    out_i = np.full((m, k), -1, dtype=np.int64)
    out_s = np.full((m, k), -np.inf)
    w = best_s.shape[1]
    out_i[:, :w] = np.where(np.isneginf(best_s), -1, best_i)
    out_s[:, :w] = best_s
    return out_i, out_s, row_scores


def judge(ids: np.ndarray, scores: np.ndarray, ref_ids: np.ndarray, ref_scores: np.ndarray, row_scores, tol: float) -> Dict:
    """Per-query ``check_topk`` plus the numbers bench.py prints.

    recall_at_k            north_star's: an oracle row that is missing counts as found when it ties with the cut within
                           2*tol (``check_topk``'s band) — 1.0 exactly when every query passes ``check_topk``'s id rule
    recall_at_k_strict     plain |returned ∩ oracle| / |oracle|, near-ties not excused
    max_abs_score_err      max |device score - float64 oracle score of the same row|
    """
    ids, scores = np.atleast_2d(ids), np.atleast_2d(scores)
    errors, worst = [], 0.0
    hit = tot = excused = 0
    for i in range(len(ids)):
        table = row_scores[i]

        def score_of(a, table=table):
            return np.array([table[int(x)] for x in a])
        err = ref_search.check_topk(ids[i], scores[i], ref_ids[i], ref_scores[i], score_of, tol)
        if err is not None:
            errors.append(f"query {i}: {err}")
        got = ids[i][ids[i] >= 0]
        if len(got):
            worst = max(worst, float(np.abs(score_of(got) - scores[i][: len(got)].astype(np.float64)).max()))
        want = ref_ids[i][ref_ids[i] >= 0]
        sg = set(got.tolist())
        cut = ref_scores[i][len(want) - 1] if len(want) else -np.inf
        for w, ws in zip(want.tolist(), ref_scores[i][: len(want)].tolist()):
            tot += 1
            if w in sg:
                hit += 1
            elif ws <= cut + 2 * tol:
                excused += 1
    return {"parity_queries": int(len(ids)), "recall_at_k": 1.0 if tot == 0 else (hit + excused) / tot,
            "recall_at_k_strict": 1.0 if tot == 0 else hit / tot, "max_abs_score_err": worst,
            "score_tol": tol, "check_topk_failures": len(errors), "errors": errors[:5]}


def sample_queries(scores_k1: np.ndarray, k: int, n: int, always: Sequence[int] = ()) -> np.ndarray:
    """Indices of n queries to judge: the given ones (planted), the queries with the smallest gap between their k-th and
    (k+1)-th score (where a wrong cut is most likely; ``scores_k1`` is a device result for k+1), the rest evenly spaced."""
    nq = len(scores_k1)
    pick = [int(i) for i in always if 0 <= int(i) < nq]
    gap = scores_k1[:, k - 1].astype(np.float64) - scores_k1[:, k].astype(np.float64)
    gap = np.where(np.isfinite(gap), gap, np.inf)
    for i in np.argsort(gap, kind="stable")[: max(0, n // 2)]:
        if int(i) not in pick:
            pick.append(int(i))
    for i in np.linspace(0, nq - 1, num=min(nq, 2 * n)).astype(int):
        if len(pick) >= min(n, nq):
            break
        if int(i) not in pick:
            pick.append(int(i))
    return np.array(sorted(pick[: min(n, nq)]), dtype=np.int64)


def check_store(store, Q: np.ndarray, ids: np.ndarray, scores: np.ndarray, k: int, metric: str, tol: float,
                sample: np.ndarray, extra=None, threads: int = 4) -> Dict:
    """Judge the device result (ids, scores over ALL queries of Q) on the sampled queries.  ``extra``: further device
    results for the same sample, {name: (ids [len(sample),k], scores)} — e.g. the scan path — judged against the same
    oracle pass."""
    sample = np.asarray(sample, dtype=np.int64)
    rows = [ids[i] for i in sample]
    if extra:
        rows = [np.unique(np.concatenate([rows[j]] + [np.asarray(e[0][j]) for e in extra.values()])) for j in range(len(sample))]
    oi, osc, tables = oracle_topk(store, Q[sample], k, metric, score_rows=rows, threads=threads)
    out = judge(ids[sample], scores[sample], oi, osc, tables, tol)
    out["sampled_queries"] = sample.tolist()
    if extra:
        out["extra"] = {name: judge(np.asarray(e[0]), np.asarray(e[1]), oi, osc, tables, tol) for name, e in extra.items()}
    return out

"""Shared test helpers: golden fixtures -> candidate sets, oracle calls."""
from pathlib import Path

import numpy as np

from oracle import ref_search

GOLDEN = Path(__file__).resolve().parent / "golden"
SEARCH_FIXTURES = sorted(p.name for p in GOLDEN.glob("search_*.npz"))


def load(name):
    return dict(np.load(GOLDEN / name, allow_pickle=False))


def policy_valid(fx, window=1000):
    """Row mask of what the reference scans: latest row per id (:351), non-NULL (:485-492),
    source filter (:466-469), newest `window` by timestamp (:474)."""
    n = len(fx["X"])
    valid = np.ones(n, dtype=bool)
    last = {}
    for i, d in enumerate(fx["doc_ids"].tolist()):
        if d in last:
            valid[last[d]] = False
        last[d] = i
    sf = fx["source_filters"].tolist()
    if sf:
        valid &= np.isin(fx["sources"], sf)
    live = np.flatnonzero(valid)
    # the window is taken over rows the SQL returns (NULL rows included), newest first
    order = live[np.argsort(-fx["ts"][live], kind="stable")][:window]
    win = np.zeros(n, dtype=bool)
    win[order] = True
    total = int(win.sum())
    win[fx["null_rows"]] = False
    return win, total


def oracle_policy_search(fx):
    """The oracle run under the reference's policy (cosine, clamp, threshold, window)."""
    valid, total = policy_valid(fx)
    rows = np.flatnonzero(valid)
    k = int(fx["k"])
    if len(rows) == 0:
        return np.zeros(0, np.int64), np.zeros(0), total
    ids, sc = ref_search.search_f64(fx["X"][rows], fx["q"], k, "cosine", float(fx["thr"]), True, recency=fx["ts"][rows])
    m = ids[0] >= 0
    return rows[ids[0][m]], sc[0][m], total

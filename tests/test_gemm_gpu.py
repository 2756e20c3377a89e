"""GPU parity of the batched tcgen05 path (K4) through the C ABI against the float64 oracle."""
import numpy as np
import pytest

from oracle import ref_search
from test_search_gpu import TOL, _check, _store

pytestmark = pytest.mark.gpu


def _flags():
    from modular_rag_system_b200._lib import FORCE_GEMM, FORCE_SCAN, NO_RESCORE
    return FORCE_GEMM, FORCE_SCAN, NO_RESCORE


@pytest.mark.parametrize("dtype,n,dim,nq,k", [
    ("fp16", 3000, 384, 128, 10),
    ("fp16", 70000, 384, 300, 10),
    ("bf16", 20000, 384, 64, 10),
    ("fp16", 20000, 768, 200, 10),
    ("fp16", 5000, 100, 17, 5),      # k-block tail (dim % 64 != 0), ragged query tile
    ("fp16", 257, 64, 1, 3),         # single query forced through the batched kernel
    ("fp16", 40000, 128, 1100, 32),  # more query tiles than fit one wave of splits, kpad 32
    ("fp16", 9000, 1024, 130, 16),
])
def test_batched_vs_oracle(dtype, n, dim, nq, k):
    FORCE_GEMM, _, _ = _flags()
    rng = np.random.default_rng(n + dim + nq)
    X = rng.standard_normal((n, dim)).astype(np.float32)
    Q = rng.standard_normal((nq, dim)).astype(np.float32)
    for i in range(0, nq, 7):  # planted neighbours
        Q[i] = X[(i * 131) % n] + 0.3 * rng.standard_normal(dim).astype(np.float32)
    s = _store(X, dtype)
    ids, sc = s.search(Q, k, "cosine", flags=FORCE_GEMM)
    rid, _ = _check(ids, sc, X, Q, k, "cosine", -np.inf, False, TOL[dtype])
    for i in range(0, nq, 7):
        assert ids[i][0] == (i * 131) % n
    if dtype == "fp16":
        assert ref_search.recall_at_k(ids, rid) > 0.995


@pytest.mark.parametrize("dtype,n,dim,nq,k", [
    ("fp16", 60000, 384, 300, 100),   # BASELINE config 5's k: candidate buffers in global memory (kpad 128)
    ("bf16", 50000, 384, 257, 100),
    ("fp16", 30000, 768, 129, 50),    # kpad 64
    ("fp16", 700, 128, 130, 128),     # k == kpad, fewer rows than fill the first prune
    ("fp16", 90, 64, 200, 100),       # fewer rows than k: short lists
])
def test_batched_large_k_vs_oracle(dtype, n, dim, nq, k):
    """k in (32, 128] runs on the SM-pair tensor-core kernel with per-query buffers of 2*kpad keys."""
    FORCE_GEMM, _, _ = _flags()
    rng = np.random.default_rng(n + dim + nq + k)
    X = rng.standard_normal((n, dim)).astype(np.float32)
    Q = rng.standard_normal((nq, dim)).astype(np.float32)
    for i in range(0, nq, 9):
        Q[i] = X[(i * 131) % n] + 0.3 * rng.standard_normal(dim).astype(np.float32)
    s = _store(X, dtype)
    import modular_rag_system_b200 as pkg
    l0 = pkg.launch_count()
    ids, sc = s.search(Q, k, "cosine")
    assert pkg.launch_count() - l0 == 3, "expected convert + pair kernel + finish (chunk merge, fp32 re-score, top-k)"
    rid, _ = _check(ids, sc, X, Q, k, "cosine", -np.inf, False, TOL[dtype])
    for i in range(0, nq, 9):
        assert ids[i][0] == (i * 131) % n
    if n < k:
        assert (ids[:, n:] == -1).all() and (ids[:, :n] >= 0).all()
    if dtype == "fp16":
        assert ref_search.recall_at_k(ids, rid) > 0.995


def test_batched_large_k_equals_scan_path():
    FORCE_GEMM, FORCE_SCAN, _ = _flags()
    rng = np.random.default_rng(13)
    X = rng.standard_normal((40000, 384)).astype(np.float32)
    X[100:140] = X[99]               # a run of exact ties straddling the candidate path
    Q = rng.standard_normal((160, 384)).astype(np.float32)
    Q[5] = X[99]
    s = _store(X, "fp16")
    valid = np.ones(40000, np.uint8); valid[::5] = 0
    for kw in (dict(), dict(threshold=0.05, clamp01=True), dict(valid=valid), dict(row_begin=777, row_end=39000)):
        a_ids, a_sc = s.search(Q, 100, "cosine", flags=FORCE_SCAN, **kw)
        b_ids, b_sc = s.search(Q, 100, "cosine", **kw)
        same = a_ids == b_ids
        assert same.mean() > 0.99, kw
        assert np.abs(a_sc[same] - b_sc[same])[np.isfinite(a_sc[same])].max() < 2e-6
    ids, _ = s.search(Q, 100, "cosine")
    assert ids[5][:41].tolist() == list(range(139, 98, -1))  # exact ties: newest (largest row) first


def test_batched_equals_scan_path():
    FORCE_GEMM, FORCE_SCAN, NO_RESCORE = _flags()
    rng = np.random.default_rng(11)
    X = rng.standard_normal((30000, 384)).astype(np.float32)
    Q = rng.standard_normal((256, 384)).astype(np.float32)
    s = _store(X, "fp16")
    a_ids, a_sc = s.search(Q, 10, "cosine", flags=FORCE_SCAN)
    b_ids, b_sc = s.search(Q, 10, "cosine", flags=FORCE_GEMM)
    # re-scored winners carry the scan's fp32 scores; id sets may differ only on near-ties
    same = (a_ids == b_ids)
    assert same.mean() > 0.99
    assert np.abs(a_sc[same] - b_sc[same]).max() < 2e-6
    c_ids, c_sc = s.search(Q, 10, "cosine", flags=FORCE_GEMM | NO_RESCORE)
    assert (c_ids == b_ids).mean() > 0.98 and np.abs(c_sc - b_sc).max() < 1e-3


def test_batched_policy_flags():
    """threshold, clamp, valid mask, row window, exact ties — same semantics as the scan."""
    FORCE_GEMM, FORCE_SCAN, _ = _flags()
    rng = np.random.default_rng(12)
    d = 128
    X = rng.standard_normal((5000, d)).astype(np.float32)
    Q = rng.standard_normal((140, d)).astype(np.float32)
    for r in (5, 1500, 4999, 17):
        X[r] = Q[3]
    X[40] = 0.0
    valid = np.ones(5000, np.uint8); valid[1500] = 0; valid[::3] = 0
    s = _store(X, "fp16")
    for kw in (dict(threshold=0.2, clamp01=True), dict(threshold=0.0, clamp01=True, valid=valid),
               dict(row_begin=1000, row_end=4000), dict(threshold=0.95)):
        a = s.search(Q, 8, "cosine", flags=FORCE_SCAN, **kw)
        b = s.search(Q, 8, "cosine", flags=FORCE_GEMM, **kw)
        same = a[0] == b[0]
        assert same.mean() > 0.99, kw
        assert (np.isneginf(a[1]) == np.isneginf(b[1])).mean() > 0.995
    ids, sc = s.search(Q, 8, "cosine", 0.0, clamp01=True, flags=FORCE_GEMM)
    assert ids[3][:4].tolist() == [4999, 1500, 17, 5]
    ids, sc = s.search(np.zeros((130, d), np.float32), 3, "cosine", 0.0, clamp01=True, flags=FORCE_GEMM)
    assert (ids == np.array([4999, 4998, 4997])).all() and (sc == 0).all()


# full-size (1M x 384, 10M x 768, 12.5M x 384) parity against the float64 oracle: tests/test_fullsize_parity_gpu.py


@pytest.mark.parametrize("nq,k,dim,dtype", [(4, 10, 384, "fp16"), (17, 5, 384, "fp16"), (128, 32, 384, "fp16"), (40, 100, 384, "fp16"),
                                            (2, 64, 384, "fp16"), (128, 10, 768, "bf16"), (77, 16, 100, "fp16"), (128, 128, 1024, "fp16")])
def test_single_tile_batches_on_every_kernel(nq, k, dim, dtype, monkeypatch):
    """B <= 128 is one query tile.  Default: the solo form of the tcgen05 kernel (one CTA per SM, cta_group::1, M=128 —
    csrc/gemm_pair.cuh kSolo).  RAGB_SOLO=0 brings back the earlier routing: below the measured crossover the SM-pair
    kernel beside an all-zero tile, above it the single-CTA row-granular kernel (RAGB_PAIR_SMALL forces either; k > 32
    exists only on the chunk-granular kernels).  All must agree with the oracle; solo and pair end in the same fp32
    re-score, so they must return the same rows and bit-identical scores."""
    import modular_rag_system_b200 as pkg
    FORCE_GEMM, _, _ = _flags()
    rng = np.random.default_rng(nq * 1000 + k)
    n = 30000
    X = rng.standard_normal((n, dim)).astype(np.float32)
    X[29990] = X[15000] = X[4321]                # exact ties across row splits
    Q = rng.standard_normal((nq, dim)).astype(np.float32)
    Q[0] = X[4321] + 0.3 * rng.standard_normal(dim).astype(np.float32)
    s = _store(X, dtype)
    got = {}
    for route, env in (("solo", {}), ("pair", {"RAGB_SOLO": "0", "RAGB_PAIR_SMALL": "1"}),
                       ("single", {"RAGB_SOLO": "0", "RAGB_PAIR_SMALL": "0"})):
        for name, v in env.items():
            monkeypatch.setenv(name, v)
        l0 = pkg.launch_count()
        ids, sc = s.search(Q, k, "cosine", flags=FORCE_GEMM)
        launches = pkg.launch_count() - l0
        _check(ids, sc, X, Q, k, "cosine", -np.inf, False, TOL[dtype])
        assert list(ids[0][:3]) == [29990, 15000, 4321]
        got[route] = (np.array(ids), np.array(sc), launches)
        for name in env:
            monkeypatch.delenv(name)
    assert got["solo"][2] == 3 and got["pair"][2] == 3      # convert + tcgen05 kernel + finish
    np.testing.assert_array_equal(got["solo"][0], got["pair"][0])
    np.testing.assert_array_equal(got["solo"][1], got["pair"][1])
    if k <= 32:
        assert got["single"][2] > 3              # convert + single-CTA kernel + merge + re-score + merge
        np.testing.assert_allclose(got["single"][1], got["solo"][1], atol=2e-3)
    s.close()


@pytest.mark.parametrize("nq,k", [(128, 10), (33, 100)])
def test_solo_masked_epilogue_window_and_threshold(nq, k):
    """The solo form under a valid mask, a row window that ends inside a tile, a score threshold and clamp — against the
    scan path (exact fp32, same re-score arithmetic)."""
    FORCE_GEMM, FORCE_SCAN, _ = _flags()
    rng = np.random.default_rng(k + nq)
    n, dim = 50_000, 256
    X = rng.standard_normal((n, dim)).astype(np.float32)
    Q = rng.standard_normal((nq, dim)).astype(np.float32)
    for i in range(0, nq, 5):
        Q[i] = X[(i * 797) % n] + 0.4 * rng.standard_normal(dim).astype(np.float32)
    valid = (rng.random(n) > 0.3).astype(np.uint8)
    s = _store(X, "fp16")
    for kw in (dict(valid=valid), dict(row_begin=1000, row_end=40_001), dict(threshold=0.05, clamp01=True),
               dict(valid=valid, row_begin=37, row_end=49_999, threshold=0.0, clamp01=True)):
        a = s.search(Q, k, "cosine", flags=FORCE_GEMM, **kw)
        b = s.search(Q, k, "cosine", flags=FORCE_SCAN, **kw)
        same = np.array(a[0]) == np.array(b[0])
        assert same.mean() > 0.999, kw                       # 16-bit-query selection: a near-tie at the cut may differ
        fin = same & np.isfinite(np.array(b[1]))
        assert np.abs(np.array(a[1])[fin] - np.array(b[1])[fin]).max() < 2e-6
    s.close()


@pytest.mark.parametrize("dim,nq,k", [(384, 300, 10), (256, 700, 100), (320, 130, 32), (100, 260, 64)])
def test_resident_query_tile_and_shared_threshold_switches(dim, nq, k, monkeypatch):
    """dim <= 384 keeps the query tile in shared memory (kQRes) and the row splits of a query share one running
    threshold (gthr) — csrc/gemm_pair.cuh.  Neither may change what is returned: every combination of the two A/B
    switches must agree with the oracle and with each other, exact ties (duplicated rows) included."""
    FORCE_GEMM, _, _ = _flags()
    rng = np.random.default_rng(dim * 7 + nq + k)
    n = 60000
    X = rng.standard_normal((n, dim)).astype(np.float32)
    X[50000:50040] = X[123]                      # 41 copies of one row: ties across chunks and across row splits
    Q = rng.standard_normal((nq, dim)).astype(np.float32)
    Q[0] = X[123] + 0.05 * rng.standard_normal(dim).astype(np.float32)
    for i in range(5, nq, 11):
        Q[i] = X[(i * 977) % n] + 0.3 * rng.standard_normal(dim).astype(np.float32)
    s = _store(X, "fp16")
    got = []
    for qres, gthr in (("1", "1"), ("0", "1"), ("1", "0"), ("0", "0")):
        monkeypatch.setenv("RAGB_PAIR_QRES", qres)
        monkeypatch.setenv("RAGB_PAIR_GTHR", gthr)
        ids, sc = s.search(Q, k, "cosine", flags=FORCE_GEMM)
        _check(ids, sc, X, Q, k, "cosine", -np.inf, False, TOL["fp16"])
        got.append((np.array(ids), np.array(sc)))
    for ids, sc in got[1:]:
        np.testing.assert_array_equal(ids, got[0][0])   # the fp32 re-score decides: same rows, same order
        np.testing.assert_array_equal(sc, got[0][1])
    top = set(int(x) for x in got[0][0][0][:min(k, 41)])
    assert top <= ({123} | set(range(50000, 50040)))
    s.close()


@pytest.mark.parametrize("dim,nq,k,dtype", [(384, 600, 100, "bf16"), (768, 513, 64, "fp16"), (100, 1100, 128, "fp16")])
def test_tile_parity_epilogue_equals_column_split(dim, nq, k, dtype, monkeypatch):
    """k > 32, two tile pairs or more: the two epilogue warps of a lane quarter split the TILES, each with its own
    candidate buffers and two lists per query and row split (csrc/gemm_pair.cuh kPar; RAGB_PAIR_PAR=0 is the column
    split).  Both end in the same fp32 re-score of the k_sel best chunks overall: same rows, bit-identical scores —
    also under a valid mask and a row window that ends inside a tile — and both agree with the oracle."""
    FORCE_GEMM, _, _ = _flags()
    rng = np.random.default_rng(dim + nq + k)
    n = 70_001
    X = rng.standard_normal((n, dim)).astype(np.float32)
    X[60_000:60_030] = X[77]                     # ties across tiles of both parities and across row splits
    Q = rng.standard_normal((nq, dim)).astype(np.float32)
    Q[0] = X[77] + 0.05 * rng.standard_normal(dim).astype(np.float32)
    for i in range(3, nq, 13):
        Q[i] = X[(i * 977) % n] + 0.3 * rng.standard_normal(dim).astype(np.float32)
    valid = (rng.random(n) > 0.25).astype(np.uint8)
    s = _store(X, dtype)
    got = {}
    for par in ("1", "0"):
        monkeypatch.setenv("RAGB_PAIR_PAR", par)
        a = s.search(Q, k, "cosine", flags=FORCE_GEMM)
        _check(a[0], a[1], X, Q, k, "cosine", -np.inf, False, TOL[dtype])
        b = s.search(Q, k, "cosine", flags=FORCE_GEMM, valid=valid, row_begin=129, row_end=n - 77, threshold=0.02, clamp01=True)
        got[par] = [np.array(x) for x in (*a, *b)]
    for x, y in zip(got["1"], got["0"]):
        np.testing.assert_array_equal(x, y)
    top = set(int(x) for x in got["1"][0][0][:31])
    assert top == ({77} | set(range(60_000, 60_030)))
    s.close()


@pytest.mark.parametrize("k,dtype", [(64, "fp16"), (100, "bf16"), (20, "fp16")])
def test_masked_epilogue_large_k_vs_oracle(k, dtype):
    """The masked epilogue (valid mask, row window ending inside a tile) of the SM-pair kernel for every candidate-buffer
    variant — shared memory (k <= 32) and global memory (kpad 64 / 128) — against the oracle and the scan."""
    FORCE_GEMM, FORCE_SCAN, _ = _flags()
    rng = np.random.default_rng(k)
    n, dim, nq = 33_333, 192, 260
    X = rng.standard_normal((n, dim)).astype(np.float32)
    Q = rng.standard_normal((nq, dim)).astype(np.float32)
    for i in range(0, nq, 13):
        Q[i] = X[(i * 211) % 30_000] + 0.3 * rng.standard_normal(dim).astype(np.float32)
    valid = (rng.random(n) > 0.3).astype(np.uint8)
    valid[4096:4608] = 0                                  # two whole tiles masked out
    s = _store(X, dtype)
    rb, re = 517, 31_991                                  # neither end on a 16-row chunk or a 256-row tile
    ids, sc = s.search(Q, k, "cosine", valid=valid, row_begin=rb, row_end=re, flags=FORCE_GEMM)
    vmask = valid.astype(bool) & (np.arange(n) >= rb) & (np.arange(n) < re)
    _check(ids, sc, X, Q, k, "cosine", -np.inf, False, TOL[dtype], valid=vmask)
    assert vmask[ids[ids >= 0]].all()
    a_ids, a_sc = s.search(Q, k, "cosine", valid=valid, row_begin=rb, row_end=re, flags=FORCE_SCAN)
    same = a_ids == ids
    assert same.mean() > 0.99 and np.abs(a_sc[same] - sc[same]).max() < 2e-6
    s.close()

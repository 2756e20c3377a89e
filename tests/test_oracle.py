"""The CPU oracle against the unmodified reference's outputs (tests/golden, made by oracle/make_golden.py)."""
import numpy as np
import pytest

from oracle import ref_ingest, ref_search
from helpers import SEARCH_FIXTURES, load, oracle_policy_search, policy_valid


@pytest.mark.parametrize("name", SEARCH_FIXTURES)
def test_search_f64_matches_reference(name):
    fx = load(name)
    rows, sims, total = oracle_policy_search(fx)
    assert total == int(fx["ref_total"])
    assert rows.tolist() == fx["ref_rows"].tolist()
    np.testing.assert_allclose(sims, fx["ref_sims"], rtol=0, atol=1e-12)


@pytest.mark.parametrize("name", [n for n in SEARCH_FIXTURES if "_n40_" in n or "_n7_" in n or "_n1_" in n])
def test_search_literal_matches_reference(name):
    fx = load(name)
    valid, _ = policy_valid(fx)
    order = np.argsort(-fx["ts"], kind="stable")
    rows = [(int(i), fx["X"][i].astype(np.float64).tolist()) for i in order if valid[i]]
    ids, sims = ref_search.search_literal(rows, fx["q"].astype(np.float64).tolist(), int(fx["k"]), float(fx["thr"]))
    assert ids == fx["ref_rows"].tolist()
    np.testing.assert_allclose(sims, fx["ref_sims"], rtol=0, atol=1e-12)


def test_blocked_equals_unblocked():
    rng = np.random.default_rng(5)
    X = rng.standard_normal((3000, 96)).astype(np.float32)
    Q = rng.standard_normal((7, 96)).astype(np.float32)
    a = ref_search.search_f64(X, Q, 10, "cosine")
    b = ref_search.search_f64_blocked(X, Q, 10, "cosine", block_rows=701)
    assert (a[0] == b[0]).all()
    np.testing.assert_allclose(a[1], b[1], atol=1e-12)


def test_length_mismatch_scores_zero():
    X = np.ones((5, 8), np.float32)
    ids, sc = ref_search.search_f64(X, np.ones((1, 4), np.float32), 3, "cosine", 0.0, True)
    assert ids[0].tolist() == [4, 3, 2] and (sc[0] == 0.0).all()


@pytest.mark.parametrize("d", [128, 384])
def test_normalize_matches_reference(d):
    fx = load(f"normpack_d{d}.npz")
    unit, norms = ref_ingest.normalize_f32(fx["x"])
    np.testing.assert_allclose(unit, fx["unit"], rtol=0, atol=2e-7)
    np.testing.assert_allclose(norms, fx["norms"], rtol=2e-7)


def test_normalize_zero_row():
    unit, norms = ref_ingest.normalize_f32(np.zeros((2, 16), np.float32))
    assert (unit == 0).all() and (norms == 0).all()


def test_pool_normalize_properties():
    rng = np.random.default_rng(2)
    tok = rng.standard_normal((6, 20, 32)).astype(np.float32)
    mask = (np.arange(20)[None, :] < rng.integers(1, 21, size=(6, 1))).astype(np.int64)
    unit, norms = ref_ingest.pool_normalize_f32(tok, mask)
    np.testing.assert_allclose(np.linalg.norm(unit, axis=1), 1.0, atol=1e-6)
    # masked tokens do not matter
    tok2 = tok.copy(); tok2[mask == 0] = 1e3
    unit2, _ = ref_ingest.pool_normalize_f32(tok2, mask)
    np.testing.assert_allclose(unit, unit2, atol=1e-6)
    # one valid token -> that token, normalised
    m1 = np.zeros((6, 20), np.int64); m1[:, 3] = 1
    u1, n1 = ref_ingest.pool_normalize_f32(tok, m1)
    np.testing.assert_allclose(u1, tok[:, 3] / np.linalg.norm(tok[:, 3], axis=1, keepdims=True), atol=1e-6)


def test_pack_bf16_roundtrip():
    x = np.array([1.0, 1.00390625, -2.5, 3.1415927, 0.0, 1e-40], np.float32)
    b = ref_ingest.pack_bf16(x)
    y = ref_ingest.unpack_bf16(b)
    assert np.all(np.abs(y - x) <= np.abs(x) * 2.0 ** -8 + 1e-38)
    import torch
    t = torch.from_numpy(x).bfloat16().view(torch.int16).numpy().view(np.uint16)
    assert (t == b).all()

"""ragb_scan_topk_static_dev (include/ragb200.h, csrc/noalloc.cu): the scan search with its scratch in a file-scope
__device__ array instead of a cudaMalloc'ed workspace — for callers that cannot allocate (CUDA graph capture, tight
serving loops).  No counterpart in the reference.

Called through the C ABI (ctypes) and compared with the existing entry point, ragb_search_dev with RAGB_FORCE_SCAN,
over the SAME packed rows (the store's own device buffer is handed to the new entry point as "rows the caller
holds"): same kernel arithmetic, so ids must be identical and scores equal within 2e-6, the band the scan-vs-scan
comparisons of tests/test_fullsize_parity_gpu.py use.  Two calls in a row share the static array: the second one must
not see what the first one left behind.

This file sorts last on purpose (a failure here must not hide the other tests); it
passes alone in a fresh process and after the other GPU tests.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SCORE_TOL = 2e-6
MAX_NQ, MAX_KEYS = 64, 1024     # RAGB_STATIC_SCAN_MAX_NQ, RAGB_STATIC_SCAN_MAX_KEYS


def _static_scan(lib, s, q, k, metric, stream, threshold=float("-inf"), flags=0, valid=None, n=None):
    import torch
    nq = q.shape[0]
    ids = torch.full((nq, k), -7, dtype=torch.int64, device=q.device)
    sc = torch.full((nq, k), 123.0, dtype=torch.float32, device=q.device)
    rc = lib.ragb_scan_topk_static_dev(
        s.device, lib.ragb_store_rows_dev(s._h), lib.ragb_store_dtype(s._h), len(s) if n is None else n, s.dim,
        lib.ragb_store_ld(s._h), q.data_ptr(), nq, k, {"ip": 0, "cosine": 1}[metric], threshold, flags,
        None if valid is None else valid.data_ptr(), sc.data_ptr(), ids.data_ptr(), stream)
    return rc, ids, sc


@pytest.mark.parametrize("dtype,n,d", [("fp16", 200_000, 384), ("bf16", 50_000, 768), ("fp32", 30_000, 100)])
def test_static_scratch_scan_equals_store_scan_twice_in_a_row(dtype, n, d):
    import torch
    from modular_rag_system_b200 import DeviceStore, lib
    from modular_rag_system_b200._lib import CLAMP01, FORCE_SCAN
    L = lib()
    g = torch.Generator(device="cuda").manual_seed(11)
    s = DeviceStore(d, dtype, n)
    X = torch.randn((n, d), generator=g, device="cuda")
    X[777] = X[n - 5] = X[3]                       # equal scores: the later row comes first
    s.add_dev(X)
    Q = torch.randn((MAX_NQ, d), generator=g, device="cuda")
    Q[0] = X[3]
    Q = Q.contiguous()
    valid = (torch.arange(n, device="cuda") % 5 != 1).to(torch.uint8)
    torch.cuda.synchronize()
    st = torch.cuda.current_stream().cuda_stream or 1          # torch's default stream by its explicit name
    assert L.ragb_scan_topk_static_dev(s.device, None, 0, 0, d, 8 * ((d + 7) // 8), None, 0, 1, 0, 0.0, 0, None, None,
                                       None, None) == 0        # nq = 0: set-up only
    # (nq, k, metric, kwargs) — consecutive calls on one stream; sizes shrink and grow so that a later call reads
    # scratch a former one wrote
    calls = [(MAX_NQ, 10, "cosine", {}), (3, 100, "ip", {}), (1, 1000, "cosine", {}), (8, 128, "ip", {}),
             (MAX_NQ, 16, "cosine", dict(threshold=0.05, flags=CLAMP01)), (5, 7, "ip", dict(valid=valid))]
    free0 = None
    for rep in range(2):                           # the whole sequence twice: every call has a predecessor's leftovers
        for nq, k, metric, kw in calls:
            q = Q[:nq].contiguous()
            want_ids, want_sc = s.search_dev(q, k, metric, threshold=kw.get("threshold", float("-inf")),
                                             clamp01=bool(kw.get("flags", 0) & CLAMP01), valid=kw.get("valid"),
                                             flags=FORCE_SCAN)
            rc, ids, sc = _static_scan(L, s, q, k, metric, st, **kw)
            assert rc == 0, L.ragb_last_error()
            rc2, ids2, sc2 = _static_scan(L, s, q, k, metric, st, **kw)      # and again, back to back
            assert rc2 == 0, L.ragb_last_error()
            torch.cuda.synchronize()
            for got_ids, got_sc in ((ids, sc), (ids2, sc2)):
                assert torch.equal(got_ids, want_ids), (dtype, nq, k, metric)
                a, b = got_sc.cpu().numpy(), want_sc.cpu().numpy()
                fin = np.isfinite(b)
                assert np.array_equal(np.isfinite(a), fin) and (np.abs(a[fin] - b[fin]) <= SCORE_TOL).all()
        if rep == 0:
            free0 = torch.cuda.mem_get_info()[0]   # everything is warm now: the second round must not allocate
    assert torch.cuda.mem_get_info()[0] == free0
    ids, _ = _static_scan(L, s, Q[:1].contiguous(), 3, "cosine", st)[1:]
    assert ids[0].tolist() == [n - 5, 777, 3]
    # empty corpus: empty slots, as the store answers an empty window
    rc, ids, sc = _static_scan(L, s, Q[:2].contiguous(), 4, "ip", st, n=0)
    assert rc == 0 and (ids == -1).all() and torch.isneginf(sc).all()
    s.close()


def test_static_scratch_limits_are_refused_with_a_clear_error():
    import torch
    from modular_rag_system_b200 import DeviceStore, lib
    from modular_rag_system_b200._lib import EUNSUPPORTED
    L = lib()
    s = DeviceStore(64, "fp16", 4096)
    s.add_dev(torch.randn((4096, 64), device="cuda"))
    Q = torch.randn((MAX_NQ + 1, 64), device="cuda")
    st = torch.cuda.current_stream().cuda_stream or 1
    for nq, k in ((MAX_NQ + 1, 1), (MAX_NQ, 17), (9, 128), (2, 1000)):   # too many queries / nq * kpad > MAX_KEYS
        assert nq > MAX_NQ or nq * L.ragb_kpad(k) > MAX_KEYS
        rc, ids, sc = _static_scan(L, s, Q[:nq].contiguous(), k, "ip", st)
        torch.cuda.synchronize()
        assert rc == EUNSUPPORTED
        msg = L.ragb_last_error().decode()
        assert "static scratch limit" in msg and "ragb_scan_topk_static_dev" in msg, msg
        assert (ids == -7).all() and (sc == 123.0).all()                 # nothing was launched
    for nq, k in ((MAX_NQ, 16), (8, 128), (1, 1024)):                    # the largest calls that fit
        rc, ids, sc = _static_scan(L, s, Q[:nq].contiguous(), k, "ip", st)
        want = s.search_dev(Q[:nq].contiguous(), k, "ip", flags=2)
        torch.cuda.synchronize()
        assert rc == 0 and torch.equal(ids, want[0])
    s.close()

"""Sweep the query-stationary kernel's knobs (env RAGB_TS_*) — development aid."""
import os, sys, itertools
import torch
sys.path.insert(0, ".")
from scripts.quick_perf import fill, timeit  # noqa
from modular_rag_system_b200._lib import NO_RESCORE

def run(n, d, nq, envs, iters=5):
    s = fill(n, d)
    g = torch.Generator(device="cuda").manual_seed(1)
    Q = torch.randn((nq, d), generator=g, device="cuda")
    out = (torch.empty((nq, 10), dtype=torch.int64, device="cuda"), torch.empty((nq, 10), dtype=torch.float32, device="cuda"))
    for env in envs:
        for k in list(os.environ):
            if k.startswith("RAGB_"): del os.environ[k]
        os.environ.update({k: str(v) for k, v in env.items()})
        try:
            ms = timeit(lambda: s.search_dev(Q, 10, "ip", flags=NO_RESCORE, out=out), iters=iters)
            print(f"{n}x{d} b{nq} {env}: {ms:8.3f} ms  {2.0*nq*n*d/ms/1e9:8.1f} TFLOP/s", flush=True)
        except Exception as e:
            print(f"{n}x{d} b{nq} {env}: FAILED {e}", flush=True)
    s.close()

if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "a"
    if which == "p":
        for n, d in ((4_000_000, 768), (1_000_000, 384)):
            run(n, d, 1024, [dict(RAGB_PAIR=1), dict(RAGB_PAIR=0)], iters=10)
        run(4_000_000, 768, 4096, [dict(RAGB_PAIR=1), dict(RAGB_PAIR=0)], iters=3)
        for b in (130, 200, 384, 640):
            run(4_000_000, 768, b, [dict(RAGB_PAIR=1), dict(RAGB_PAIR=0)], iters=5)
    if which == "e":
        run(4_000_000, 768, 1024, [dict(RAGB_TS_NT=nt, RAGB_TS_KBS=kbs, RAGB_TS_CLUSTER=1) for nt, kbs in ((128, 1), (128, 2), (128, 3), (64, 1), (64, 2), (64, 4), (64, 6))])
    if which == "d":
        run(4_000_000, 768, 1024, [dict(RAGB_TS_NT=nt, RAGB_TS_CLUSTER=c) for nt in (256, 192, 128) for c in (2, 1)])
    if which == "c":
        for n, d in ((4_000_000, 768), (2_000_000, 384)):
            run(n, d, 1024, [dict(RAGB_TS_CLUSTER=c) for c in (8, 4, 2, 1)] + [dict(RAGB_TS_NT=128, RAGB_TS_CLUSTER=8), dict(RAGB_TS=1), dict(RAGB_GEMM_SS=1)])
        run(4_000_000, 768, 4096, [dict(RAGB_TS_CLUSTER=c) for c in (8, 4)] + [dict(RAGB_GEMM_SS=1)])
    if which == "b":
        run(4_000_000, 768, 1024, [dict(RAGB_TS_NT=nt, RAGB_TS_KBS=kbs, RAGB_TS_CLUSTER=c)
                                   for nt, kbs in ((64, 4), (64, 2), (64, 6), (32, 4)) for c in (8, 4, 2, 1)])
        run(2_000_000, 384, 1024, [dict(RAGB_TS_NT=nt, RAGB_TS_KBS=kbs, RAGB_TS_CLUSTER=c)
                                   for nt, kbs in ((128, 2), (128, 3), (64, 3)) for c in (8, 1)])
    if which == "a":
        run(1_000_000, 384, 1024, [dict(RAGB_TS_NT=nt, RAGB_TS_KBS=kbs, RAGB_TS_CLUSTER=c)
                                   for nt, kbs in ((128, 2), (128, 1), (64, 4), (64, 2), (32, 4)) for c in (8, 2, 1)], iters=20)
        run(4_000_000, 768, 1024, [dict(RAGB_TS_KBS=kbs, RAGB_TS_CLUSTER=c, RAGB_TS_STAGES=st)
                                   for kbs in (4, 2) for c in (8, 4, 1) for st in (5, 3)] + [dict(RAGB_GEMM_SS=1)])

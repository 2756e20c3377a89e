"""Quick device-side timing of the search paths (development aid; bench.py is the contract)."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from modular_rag_system_b200 import DeviceStore
from modular_rag_system_b200._lib import FORCE_SCAN, FORCE_GEMM, NO_RESCORE

def fill(n, d, dtype="fp16"):
    s = DeviceStore(d, dtype, n)
    g = torch.Generator(device="cuda").manual_seed(0)
    step = 500_000
    for r0 in range(0, n, step):
        m = min(step, n - r0)
        s.add_dev(torch.randn((m, d), generator=g, device="cuda"))
    torch.cuda.synchronize()
    return s

def timeit(fn, iters=10, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters

if __name__ == "__main__":
    cases = [(1_000_000, 384), (10_000_000, 768)] if len(sys.argv) < 2 else [tuple(map(int, a.split("x"))) for a in sys.argv[1:]]
    for n, d in cases:
        s = fill(n, d)
        g = torch.Generator(device="cuda").manual_seed(1)
        for nq, flags, name in [(1, 0, "scan b1"), (8, FORCE_SCAN, "scan b8"), (1024, 0, "gemm b1024"), (1024, NO_RESCORE, "gemm b1024 norescore"), (128, 0, "gemm b128"), (4096, 0, "gemm b4096")]:
            Q = torch.randn((nq, d), generator=g, device="cuda")
            out = (torch.empty((nq, 10), dtype=torch.int64, device="cuda"), torch.empty((nq, 10), dtype=torch.float32, device="cuda"))
            ms = timeit(lambda: s.search_dev(Q, 10, "ip", flags=flags, out=out), iters=5 if n > 2_000_000 else 20)
            gb = n * d * 2 / 1e9
            tf = 2.0 * nq * n * d / 1e12
            print(f"{n}x{d} {name:24s} {ms:9.3f} ms  qps={nq/ms*1e3:12.1f}  GB/s={gb/ms*1e3*(nq if 'scan' in name else 1):9.1f}  TFLOP/s={tf/ms*1e3:8.1f}", flush=True)
        s.close(); del s
        torch.cuda.empty_cache()

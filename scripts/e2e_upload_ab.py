"""Host entry point with pageable against page-locked query buffers: python scripts/e2e_upload_ab.py [N D B]"""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from scripts.quick_perf import fill
n, d, b = (int(x) for x in sys.argv[1:4]) if len(sys.argv) > 3 else (1_000_000, 768, 1024)
s = fill(n, d, "fp16")
Q = np.random.default_rng(1).standard_normal((b, d)).astype(np.float32)
Qp = torch.from_numpy(Q).pin_memory().numpy()
for name, q in (("pageable", Q), ("page-locked", Qp), ("pageable", Q), ("page-locked", Qp)):
    for _ in range(5):
        s.search(q, 10, "ip")
    t0 = time.perf_counter()
    for _ in range(50):
        s.search(q, 10, "ip")
    print(f"{n}x{d} B={b} {name:12s}: {(time.perf_counter() - t0) / 50 * 1e3:.3f} ms per ragb_search call", flush=True)

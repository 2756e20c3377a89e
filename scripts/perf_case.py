"""One timed search case (development aid): python -m scripts.perf_case N D B K [dtype] [iters]
Prints wall (CUDA events) ms per search and the dominant kernel's ms from the C ABI's event brackets."""
import sys
import torch
sys.path.insert(0, ".")
from scripts.quick_perf import fill, timeit

n, d, b, k = (int(x) for x in sys.argv[1:5])
dtype = sys.argv[5] if len(sys.argv) > 5 else "fp16"
iters = int(sys.argv[6]) if len(sys.argv) > 6 else 5
s = fill(n, d, dtype)
Q = torch.randn((b, d), generator=torch.Generator(device="cuda").manual_seed(1), device="cuda")
out = (torch.empty((b, k), dtype=torch.int64, device="cuda"), torch.empty((b, k), dtype=torch.float32, device="cuda"))
fn = lambda: s.search_dev(Q, k, "ip", out=out)
ms = timeit(fn, iters=iters)
s.profile(True)
for _ in range(iters):
    fn()
torch.cuda.synchronize()
kms, nl = s.profile_read()
s.profile(False)
kms /= max(nl, 1)
tf = 2.0 * b * n * d / 1e12
print(f"{n}x{d} {dtype} b={b} k={k}: {ms:.3f} ms/search qps={b / ms * 1e3:.1f}  kernel {kms:.3f} ms "
      f"{tf / kms * 1e3:.1f} TFLOP/s ({n * d * 2 / kms / 1e6:.0f} GB/s of rows)", flush=True)

"""One search case for ncu: python -m scripts.prof_case N D B [iters] [k] [dtype]"""
import sys
import torch
sys.path.insert(0, ".")
from scripts.quick_perf import fill
n, d, b = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 2
k = int(sys.argv[5]) if len(sys.argv) > 5 else 10
dtype = sys.argv[6] if len(sys.argv) > 6 else "fp16"
s = fill(n, d, dtype)
Q = torch.randn((b, d), generator=torch.Generator(device="cuda").manual_seed(1), device="cuda")
for _ in range(iters):
    ids, sc = s.search_dev(Q, k, "ip")
torch.cuda.synchronize()
print("ok", ids[0, :3].tolist())

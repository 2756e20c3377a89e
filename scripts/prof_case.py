"""One search case for ncu: python -m scripts.prof_case N D B [iters]"""
import sys
import torch
sys.path.insert(0, ".")
from scripts.quick_perf import fill
n, d, b = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 2
s = fill(n, d)
Q = torch.randn((b, d), generator=torch.Generator(device="cuda").manual_seed(1), device="cuda")
for _ in range(iters):
    ids, sc = s.search_dev(Q, 10, "ip")
torch.cuda.synchronize()
print("ok", ids[0, :3].tolist())

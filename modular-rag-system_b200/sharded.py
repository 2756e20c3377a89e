"""Row-sharded exact search across the GPUs of one NVSwitch box, one process per GPU.

The reference is single-process: this is the multi-GPU form of the scan
of ``RAGVectorStore._search_sqlite`` (/root/reference/ollama_rag_system.py:480-515).  Every
rank owns a contiguous block of rows in its own ``DeviceStore``; a search is

    local top-k (scan or tcgen05 kernel) written as sortable 64-bit keys whose low word is
    the GLOBAL row id  ->  ONE ``all_gather`` of ``k`` candidates per query per rank (NCCL over
    NVLink; 8 B x kpad x nq per rank)  ->  merge kernel on every rank.

Key layout (include/ragb200.h): ``(order-preserving score bits << 32) | global_row``; a larger
key is a better candidate, so ties go to the newer (larger) row id exactly as on one GPU.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import numpy as np

NEG_INF = float("-inf")


def shard_bounds(n_rows: int, world: int) -> List[Tuple[int, int]]:
    """Contiguous, balanced row blocks: rank r owns rows [b[r][0], b[r][1])."""
    base, rem = divmod(int(n_rows), int(world))
    out, r0 = [], 0
    for r in range(world):
        m = base + (1 if r < rem else 0)
        out.append((r0, r0 + m))
        r0 += m
    return out


def encode_keys(scores: np.ndarray, ids: np.ndarray) -> np.ndarray:
    """Host mirror of make_key() in csrc/common.cuh (uint64; 0 = empty slot for id < 0)."""
    b = np.asarray(scores, dtype=np.float32).view(np.uint32).astype(np.uint64)
    neg = (b & np.uint64(0x80000000)) != 0
    o = np.where(neg, (~b) & np.uint64(0xFFFFFFFF), b | np.uint64(0x80000000))
    key = (o << np.uint64(32)) | (np.asarray(ids).astype(np.int64).astype(np.uint64) & np.uint64(0xFFFFFFFF))
    return np.where(np.asarray(ids) < 0, np.uint64(0), key)


def decode_keys(keys: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """uint64 keys -> (float32 scores, int64 ids); empty slots give (-inf, -1)."""
    keys = np.asarray(keys, dtype=np.uint64)
    o = (keys >> np.uint64(32)).astype(np.uint32)
    b = np.where((o & np.uint32(0x80000000)) != 0, o & np.uint32(0x7FFFFFFF), ~o)
    sc = b.astype(np.uint32).view(np.float32).copy()
    ids = (keys & np.uint64(0xFFFFFFFF)).astype(np.int64)
    empty = keys == 0
    sc[empty] = NEG_INF
    ids[empty] = -1
    return sc, ids


def merge_keys_host(gathered: np.ndarray, k: int) -> Tuple[np.ndarray, np.ndarray]:
    """Host statement of the merge kernel: gathered [world, nq, kpad] uint64 -> (ids, scores) [nq, k].
    Used by the gloo (CPU) tests of the exchange plumbing; the product path merges on the device."""
    w, nq, kpad = gathered.shape
    allk = np.transpose(gathered, (1, 0, 2)).reshape(nq, w * kpad)
    top = np.sort(allk.astype(np.uint64), axis=1)[:, ::-1][:, :k]
    if top.shape[1] < k:
        top = np.concatenate([top, np.zeros((nq, k - top.shape[1]), np.uint64)], axis=1)
    sc, ids = decode_keys(top)
    return ids, sc


class PeerBuffers:
    """Gather buffers + arrival words of the peer-memory candidate exchange (include/ragb200.h, ragb_exchange).

    Layout of every rank's block, in uint64 words: [16 arrival words | buffer 0 | buffer 1], each buffer
    ``world x nq_cap x kpad`` keys; consecutive exchanges alternate between the two buffers.  ``ptrs[r]`` is the
    address of rank r's block as mapped in THIS process (torch symmetric memory across processes; plain device
    tensors when several "ranks" are emulated inside one process for tests)."""
    HDR = 16

    def __init__(self, ptrs, world: int, rank: int, nq_cap: int, kpad: int):
        from ._lib import MAX_PEERS
        if world > MAX_PEERS:
            raise ValueError(f"peer exchange supports at most {MAX_PEERS} ranks, got {world}")
        self.ptrs, self.world, self.rank, self.nq_cap, self.kpad = [int(p) for p in ptrs], world, rank, nq_cap, kpad
        self.epoch = 0

    @staticmethod
    def words(world: int, nq_cap: int, kpad: int) -> int:
        return PeerBuffers.HDR + 2 * world * nq_cap * kpad

    def next_exchange(self):
        """The ragb_exchange of the next epoch (all ranks call this in lock step)."""
        from ._lib import Exchange
        self.epoch += 1
        x = Exchange()
        x.world, x.rank, x.epoch = self.world, self.rank, self.epoch
        off = 8 * (self.HDR + (self.epoch & 1) * self.world * self.nq_cap * self.kpad)
        for r in range(self.world):
            x.keys[r] = self.ptrs[r] + off
            x.flags[r] = self.ptrs[r]
        return x


class ShardedSearcher:
    """Search a corpus whose rows are split over the ranks of a torch.distributed group.

    ``store`` is this rank's DeviceStore holding rows [row_base, row_base + len(store)).
    All ranks call ``search_dev`` with the same queries; every rank gets the full result.

    exchange = "peer" (default on CUDA with world > 1): the last kernel of the local search stores its k
    candidates per query straight into every rank's gather buffer over NVLink and signals; the merge kernel
    waits for the signals — no collective launch.  exchange = "nccl": one all_gather_into_tensor + merge."""

    def __init__(self, store, row_base: int, group=None, exchange: Optional[str] = None):
        import torch.distributed as dist
        self.store, self.row_base, self.group = store, int(row_base), group
        self.dist = dist
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.exchange = exchange or ("peer" if self.world > 1 else "nccl")
        self._keys = None
        self._gathered = None
        self._peer = None      # PeerBuffers
        self._peer_mem = None  # the symmetric allocation (kept alive)
        self._status = None

    def _buffers(self, nq: int, kpad: int, device):
        import torch
        if self._keys is None or self._keys.shape != (nq, kpad):
            self._keys = torch.empty((nq, kpad), dtype=torch.int64, device=device)
            self._gathered = torch.empty((self.world * nq, kpad), dtype=torch.int64, device=device)
        return self._keys, self._gathered

    def _peer_buffers(self, nq: int, kpad: int, device):
        """(Re)allocate the symmetric gather block; collective over the group (same nq, kpad on all ranks)."""
        import torch
        import torch.distributed._symmetric_memory as symm
        if self._peer is not None and self._peer.nq_cap >= nq and self._peer.kpad == kpad:
            return self._peer
        torch.cuda.synchronize(device)
        cap = max(nq, 1024)
        mem = symm.empty(PeerBuffers.words(self.world, cap, kpad), dtype=torch.int64, device=device)
        mem.zero_()
        grp = self.group if self.group is not None else self.dist.group.WORLD
        hdl = symm.rendezvous(mem, grp)
        torch.cuda.synchronize(device)
        self.dist.barrier(group=self.group)  # every block is zeroed before anybody writes into it
        self._peer_mem = (mem, hdl)
        self._peer = PeerBuffers(list(hdl.buffer_ptrs), self.world, self.rank, cap, kpad)
        self._status = torch.zeros(1, dtype=torch.int32, device=device)
        return self._peer

    def search_dev(self, q, k: int, metric: str = "ip", threshold: float = NEG_INF, clamp01: bool = False,
                   flags: int = 0):
        """q: CUDA fp32 [nq, dim], identical on all ranks -> (ids int64 [nq,k] GLOBAL rows, scores fp32 [nq,k])."""
        from .store import gather_merge_dev, kpad_of, merge_keys_dev
        kpad = kpad_of(k)
        if self.world > 1 and self.exchange == "peer":
            pb = self._peer_buffers(q.shape[0], kpad, q.device)
            x = pb.next_exchange()
            self.store.search_scatter_dev(q, k, x, metric, threshold, clamp01, id_base=self.row_base, flags=flags)
            return gather_merge_dev(x, q.shape[0], k, self.store.device, status=self._status)
        keys, gathered = self._buffers(q.shape[0], kpad, q.device)
        self.store.search_keys_dev(q, k, metric, threshold, clamp01, id_base=self.row_base, flags=flags, out=keys)
        if self.world == 1:
            return merge_keys_dev(keys.unsqueeze(0), k, self.store.device)
        self.dist.all_gather_into_tensor(gathered, keys, group=self.group)
        return merge_keys_dev(gathered.view(self.world, q.shape[0], kpad), k, self.store.device)

    def check(self) -> None:
        """Raise if a merge ever timed out waiting for a rank (peer exchange only); synchronises."""
        if self._status is not None and int(self._status.item()) != 0:
            raise RuntimeError("peer exchange: a rank's candidates did not arrive within the kernel's timeout")

    def search(self, q: np.ndarray, k: int, metric: str = "ip", threshold: float = NEG_INF, clamp01: bool = False,
               flags: int = 0) -> Tuple[np.ndarray, np.ndarray]:
        """Host entry: numpy fp32 queries [nq, dim] (identical on all ranks) -> (ids, scores) numpy arrays.
        Queries go host->device from pinned memory and the result comes back device->host inside the call."""
        import torch
        q = np.ascontiguousarray(np.atleast_2d(q), dtype=np.float32)
        dev = torch.device("cuda", self.store.device)
        if self.world > 1 and self.exchange == "peer":
            # one C call: upload, local search + scatter, wait + merge, results in mapped pinned memory
            from .store import kpad_of
            pb = self._peer_buffers(q.shape[0], kpad_of(k), dev)
            ids, sc, timed_out = self.store.search_exchange(q, k, pb.next_exchange(), metric, threshold, clamp01,
                                                            id_base=self.row_base, flags=flags)
            if timed_out:
                raise RuntimeError("peer exchange: a rank's candidates did not arrive within the kernel's timeout")
            return ids, sc
        if getattr(self, "_qpin", None) is None or self._qpin.shape != q.shape:
            self._qpin = torch.empty(q.shape, dtype=torch.float32, pin_memory=True)
            self._qdev = torch.empty(q.shape, dtype=torch.float32, device=dev)
        if getattr(self, "_ipin", None) is None or self._ipin.shape != (q.shape[0], k):   # outputs are [nq, k]: keyed on k too
            self._ipin = torch.empty((q.shape[0], k), dtype=torch.int64, pin_memory=True)
            self._spin = torch.empty((q.shape[0], k), dtype=torch.float32, pin_memory=True)
        self._qpin.numpy()[...] = q
        with torch.cuda.device(dev):
            self._qdev.copy_(self._qpin, non_blocking=True)
            ids, sc = self.search_dev(self._qdev, k, metric, threshold, clamp01, flags)
            self._ipin.copy_(ids, non_blocking=True)
            self._spin.copy_(sc, non_blocking=True)
            torch.cuda.current_stream().synchronize()
        return self._ipin.numpy().copy(), self._spin.numpy().copy()

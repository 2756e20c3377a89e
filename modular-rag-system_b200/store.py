"""DeviceStore — the device-resident embedding store behind the C ABI.

Thin object wrapper over ``include/ragb200.h``: rows live on one B200 as a row-major
fp16 / bf16 / fp32 matrix of unit vectors plus the fp32 norm each row had before
normalising.  This is the slot the reference fills with ``faiss.IndexFlatIP``
(/root/reference/ollama_rag_system.py:298-305, :372-386) and never searches; here
``search`` is the exact scan of ``_search_sqlite`` (:480-515).

Host API: numpy arrays in, numpy arrays out.  ``*_dev`` methods take torch CUDA
tensors (only their ``data_ptr()`` crosses the ABI) and run on torch's current stream.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import numpy as np

from . import _lib
from ._lib import BF16, CLAMP01, COSINE, F16, F32, IP, check, lib

_DTYPES = {"fp16": F16, "float16": F16, "half": F16, "bf16": BF16, "bfloat16": BF16, "fp32": F32, "float32": F32}
_METRICS = {"ip": IP, "inner_product": IP, "cosine": COSINE}
_NP_MASK = {np.dtype(np.uint8): 0, np.dtype(np.bool_): 0, np.dtype(np.int32): 1, np.dtype(np.int64): 2,
            np.dtype(np.float32): 3}
NEG_INF = float("-inf")


def _np_elem(dtype_code: int):
    return {F16: np.float16, BF16: np.uint16, F32: np.float32}[dtype_code]


def _tok_code(arr) -> int:
    """Element code of a host token array; bf16 travels as uint16 bit patterns."""
    dt = np.dtype(arr.dtype)
    if dt == np.float16:
        return F16
    if dt == np.float32:
        return F32
    if dt == np.uint16:
        return BF16
    raise TypeError(f"token embeddings must be float16, float32 or uint16(bf16 bits), got {dt}")


def _torch_code(t) -> int:
    import torch
    return {torch.float16: F16, torch.bfloat16: BF16, torch.float32: F32}[t.dtype]


def _torch_mask_code(t) -> int:
    import torch
    return {torch.uint8: 0, torch.bool: 0, torch.int32: 1, torch.int64: 2, torch.float32: 3}[t.dtype]


def _cur_stream(device_index: int) -> int:
    """Handle of torch's current stream for the C ABI.  The ABI reads NULL as "no stream given: use the store's own
    stream and synchronise before returning"; torch's default stream IS handle 0, so it is passed as
    cudaStreamLegacy (0x1) — the same stream by its explicit name — and the call stays asynchronous and ordered
    with the torch work around it."""
    import torch
    h = torch.cuda.current_stream(device_index).cuda_stream
    return h if h else 1


class DeviceStore:
    """Fixed-``dim`` store of unit rows on one GPU."""

    def __init__(self, dim: int, dtype: str = "fp16", capacity: int = 1 << 16, device: int = 0):
        self._h = C.c_void_p()
        self._lib = lib()
        code = _DTYPES[dtype] if isinstance(dtype, str) else int(dtype)
        check(self._lib.ragb_store_create(int(device), int(dim), code, int(capacity), C.byref(self._h)))
        self.dim, self.dtype_code, self.device = int(dim), code, int(device)

    # -- life cycle ---------------------------------------------------------- #
    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.ragb_store_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __len__(self) -> int:
        return int(self._lib.ragb_store_size(self._h))

    @property
    def capacity(self) -> int:
        return int(self._lib.ragb_store_capacity(self._h))

    @property
    def ld(self) -> int:
        return int(self._lib.ragb_store_ld(self._h))

    @property
    def dtype(self) -> str:
        return {F16: "fp16", BF16: "bf16", F32: "fp32"}[self.dtype_code]

    def reserve(self, capacity: int) -> None:
        check(self._lib.ragb_store_reserve(self._h, int(capacity)))

    def clear(self) -> None:
        check(self._lib.ragb_store_clear(self._h))

    def _room(self, n: int) -> None:
        need = len(self) + n
        if need > self.capacity:
            self.reserve(max(need, 2 * self.capacity, 1024))

    # -- ingest ----------------------------------------------------------------- #
    def add(self, rows: np.ndarray) -> int:
        """Normalise + pack fp32 host rows (ollama_rag_system.py:379-383); returns the first row id."""
        rows = np.ascontiguousarray(np.atleast_2d(rows), dtype=np.float32)
        if rows.shape[1] != self.dim:
            raise ValueError(f"rows have dim {rows.shape[1]}, store has {self.dim}")
        self._room(rows.shape[0])
        first = C.c_int64(-1)
        check(self._lib.ragb_store_add(self._h, rows.ctypes.data, rows.shape[0], C.byref(first)))
        return first.value

    def set(self, row: int, vec: np.ndarray) -> None:
        """Replace one row in place (INSERT OR REPLACE, :351)."""
        vec = np.ascontiguousarray(vec, dtype=np.float32).reshape(-1)
        if vec.shape[0] != self.dim:
            raise ValueError("dimension mismatch")
        check(self._lib.ragb_store_set(self._h, int(row), vec.ctypes.data))

    def add_dev(self, rows) -> int:
        """torch CUDA tensor [n, dim] (fp16/bf16/fp32, row stride allowed) -> normalise + pack."""
        if rows.dim() != 2 or rows.shape[1] != self.dim or rows.stride(1) != 1:
            raise ValueError("expected a [n, dim] tensor with unit inner stride")
        self._room(rows.shape[0])
        first = C.c_int64(-1)
        check(self._lib.ragb_store_add_dev(self._h, rows.data_ptr(), _torch_code(rows), rows.shape[0], rows.stride(0),
                                           C.byref(first), _cur_stream(self.device)))
        return first.value

    def add_pooled(self, tok: np.ndarray, mask: Optional[np.ndarray]) -> int:
        """Masked mean-pool + L2-normalise + pack host encoder outputs [B,S,H] (embedding_pipeline.py:424)."""
        tok = np.ascontiguousarray(tok)
        if tok.ndim != 3 or tok.shape[2] != self.dim:
            raise ValueError(f"token embeddings must be [B,S,{self.dim}]")
        B, S, H = tok.shape
        mp, mc = None, 0
        if mask is not None:
            mask = np.ascontiguousarray(mask)
            if mask.shape != (B, S):
                raise ValueError("mask must be [B,S]")
            if mask.dtype not in _NP_MASK:
                mask = mask.astype(np.float32)
            mp, mc = mask.ctypes.data, _NP_MASK[np.dtype(mask.dtype)]
        self._room(B)
        first = C.c_int64(-1)
        check(self._lib.ragb_store_add_pooled(self._h, tok.ctypes.data, _tok_code(tok), mp, mc, B, S, H, C.byref(first)))
        return first.value

    def add_pooled_dev(self, tok, mask) -> int:
        """torch CUDA tensors: tok [B,S,H] contiguous, mask [B,S] contiguous or None."""
        if tok.dim() != 3 or tok.shape[2] != self.dim or not tok.is_contiguous():
            raise ValueError(f"token embeddings must be contiguous [B,S,{self.dim}]")
        B, S, H = tok.shape
        mp, mc = None, 0
        if mask is not None:
            if tuple(mask.shape) != (B, S) or not mask.is_contiguous():
                raise ValueError("mask must be contiguous [B,S]")
            mp, mc = mask.data_ptr(), _torch_mask_code(mask)
        self._room(B)
        first = C.c_int64(-1)
        check(self._lib.ragb_store_add_pooled_dev(self._h, tok.data_ptr(), _torch_code(tok), mp, mc, B, S, H,
                                                  C.byref(first), _cur_stream(self.device)))
        return first.value

    def get_rows(self, row0: int = 0, n: Optional[int] = None) -> Tuple[np.ndarray, np.ndarray]:
        """Packed rows (float16 / uint16 bf16 bits / float32) and their pre-normalisation norms."""
        n = len(self) - row0 if n is None else n
        rows = np.empty((n, self.dim), dtype=_np_elem(self.dtype_code))
        norms = np.empty(n, dtype=np.float32)
        check(self._lib.ragb_store_get_rows(self._h, int(row0), int(n), rows.ctypes.data, norms.ctypes.data))
        return rows, norms

    def put_rows(self, rows: np.ndarray, norms: np.ndarray) -> int:
        """Append rows that are already packed (the arrays ``get_rows`` returned), verbatim; returns the first row."""
        rows = np.ascontiguousarray(rows, dtype=_np_elem(self.dtype_code))
        norms = np.ascontiguousarray(norms, dtype=np.float32)
        if rows.ndim != 2 or rows.shape[1] != self.dim or norms.shape[0] != rows.shape[0]:
            raise ValueError("expected packed rows [n, dim] and norms [n]")
        self._room(rows.shape[0])
        first = C.c_int64(-1)
        check(self._lib.ragb_store_put_rows(self._h, rows.ctypes.data, norms.ctypes.data, rows.shape[0], C.byref(first)))
        return first.value

    # -- search ----------------------------------------------------------------- #
    def search(self, q: np.ndarray, k: int, metric: str = "ip", threshold: float = NEG_INF, clamp01: bool = False,
               row_begin: int = 0, row_end: int = -1, valid: Optional[np.ndarray] = None,
               flags: int = 0) -> Tuple[np.ndarray, np.ndarray]:
        """Exact top-k of host queries ``q [nq, dim]`` -> (ids int64 [nq,k], scores float32 [nq,k]).

        Empty slots hold id -1 / score -inf.  A query of the wrong length scores 0.0
        against every row (ollama_rag_system.py:201-202).
        """
        q = np.ascontiguousarray(np.atleast_2d(q), dtype=np.float32)
        nq = q.shape[0]
        if q.shape[1] != self.dim:
            q = np.zeros((nq, self.dim), dtype=np.float32)  # every row scores 0.0
        ids = np.empty((nq, k), dtype=np.int64)
        scores = np.empty((nq, k), dtype=np.float32)
        vp = None
        if isinstance(valid, str):
            if valid != "filter":
                raise ValueError('valid must be a byte mask or the string "filter" (the mask left by build_filter)')
            vp = _lib.STORE_FILTER
        elif valid is not None:
            valid = np.ascontiguousarray(valid, dtype=np.uint8)
            if valid.shape[0] < (len(self) if row_end < 0 else row_end):
                raise ValueError("valid mask shorter than the row range")
            vp = valid.ctypes.data
        fl = int(flags) | (CLAMP01 if clamp01 else 0)
        check(self._lib.ragb_search(self._h, q.ctypes.data, nq, int(k), _METRICS[metric], float(threshold), fl,
                                    int(row_begin), int(row_end), vp, scores.ctypes.data, ids.ctypes.data))
        return ids, scores

    def search_dev(self, q, k: int, metric: str = "ip", threshold: float = NEG_INF, clamp01: bool = False,
                   row_begin: int = 0, row_end: int = -1, valid=None, flags: int = 0, out=None):
        """torch CUDA fp32 queries [nq, dim] -> (ids int64 [nq,k], scores fp32 [nq,k]) CUDA tensors."""
        import torch
        if q.dtype != torch.float32 or q.dim() != 2 or q.shape[1] != self.dim or not q.is_contiguous():
            raise ValueError(f"queries must be contiguous fp32 [nq,{self.dim}]")
        nq = q.shape[0]
        if out is None:
            ids = torch.empty((nq, k), dtype=torch.int64, device=q.device)
            scores = torch.empty((nq, k), dtype=torch.float32, device=q.device)
        else:
            ids, scores = out
        fl = int(flags) | (CLAMP01 if clamp01 else 0)
        check(self._lib.ragb_search_dev(self._h, q.data_ptr(), nq, int(k), _METRICS[metric], float(threshold), fl,
                                        int(row_begin), int(row_end),
                                        None if valid is None else _lib.STORE_FILTER if isinstance(valid, str) else valid.data_ptr(),
                                        scores.data_ptr(), ids.data_ptr(), _cur_stream(self.device)))
        return ids, scores

    def search_keys_dev(self, q, k: int, metric: str = "ip", threshold: float = NEG_INF, clamp01: bool = False,
                        id_base: int = 0, flags: int = 0, out=None):
        """Local top-k as sortable keys [nq, kpad] (int64 tensor holding uint64 bits) for the multi-GPU merge."""
        import torch
        nq = q.shape[0]
        kpad = kpad_of(k)
        keys = out if out is not None else torch.empty((nq, kpad), dtype=torch.int64, device=q.device)
        fl = int(flags) | (CLAMP01 if clamp01 else 0)
        check(self._lib.ragb_search_keys_dev(self._h, q.data_ptr(), nq, int(k), _METRICS[metric], float(threshold), fl,
                                             0, -1, None, int(id_base), keys.data_ptr(), _cur_stream(self.device)))
        return keys


    # -- candidate-set policy on the device (csrc/filter.cuh) -------------------- #
    def set_meta(self, row0: int, flags: Optional[np.ndarray] = None, src: Optional[np.ndarray] = None) -> None:
        """Per-row metadata of rows [row0, row0+n): flags uint8 (bit 0 live, bit 1 has an embedding; default 3) and
        source codes uint16 (default 0) — what the SQL of ollama_rag_system.py:462-477 / :485-492 / :351 looks at."""
        n = len(flags) if flags is not None else len(src)
        f = None if flags is None else np.ascontiguousarray(flags, dtype=np.uint8)
        c = None if src is None else np.ascontiguousarray(src, dtype=np.uint16)
        check(self._lib.ragb_store_set_meta(self._h, int(row0), int(n), None if f is None else f.ctypes.data,
                                            None if c is None else c.ctypes.data))

    def set_timestamps(self, row0: int, ts: np.ndarray) -> None:
        """Per-row timestamps (RAGDocument.timestamp): the order "ORDER BY timestamp DESC" (ollama_rag_system.py:474)
        and the stable sort (:512) rank by.  Out-of-order timestamps make the store build recency ranks on the device
        (csrc/rank.cuh); nothing changes while they follow row order."""
        t = np.ascontiguousarray(ts, dtype=np.float64)
        check(self._lib.ragb_store_set_timestamps(self._h, int(row0), int(len(t)), t.ctypes.data))

    def build_filter(self, allowed_sources: Optional[np.ndarray] = None, window: Optional[int] = None) -> int:
        """Evaluate (source IN ..., newest `window` rows, live, non-NULL) on the device into the store's own mask;
        returns the number of rows the reference's SQL would have returned.  Search with ``valid="filter"``."""
        tot = C.c_int64(0)
        a = None if allowed_sources is None else np.ascontiguousarray(allowed_sources, dtype=np.uint8)
        check(self._lib.ragb_store_build_filter(self._h, None if a is None else a.ctypes.data, 0 if a is None else len(a),
                                                -1 if window is None else int(window), C.byref(tot)))
        return int(tot.value)

    def search_scatter_dev(self, q, k: int, exchange, metric: str = "ip", threshold: float = NEG_INF,
                           clamp01: bool = False, id_base: int = 0, flags: int = 0) -> None:
        """Local top-k whose last kernel stores the sorted keys into every rank's gather buffer (peer memory)
        and publishes the epoch; `exchange` is a _lib.Exchange.  Pair with gather_merge_dev()."""
        fl = int(flags) | (CLAMP01 if clamp01 else 0)
        check(self._lib.ragb_search_scatter_dev(self._h, q.data_ptr(), q.shape[0], int(k), _METRICS[metric], float(threshold),
                                                fl, 0, -1, None, int(id_base), C.byref(exchange), _cur_stream(self.device)))

    def search_exchange(self, q: np.ndarray, k: int, exchange, metric: str = "ip", threshold: float = NEG_INF,
                        clamp01: bool = False, id_base: int = 0, flags: int = 0):
        """Host entry of the peer-memory exchange (C ABI ragb_search_exchange): numpy queries in, numpy results out,
        upload + local search + scatter + wait + merge + read-back inside one call on one stream.
        Returns (ids, scores, timed_out)."""
        q = np.ascontiguousarray(np.atleast_2d(q), dtype=np.float32)
        nq = q.shape[0]
        ids = np.empty((nq, k), dtype=np.int64)
        scores = np.empty((nq, k), dtype=np.float32)
        status = C.c_int(0)
        fl = int(flags) | (CLAMP01 if clamp01 else 0)
        check(self._lib.ragb_search_exchange(self._h, q.ctypes.data, nq, int(k), _METRICS[metric], float(threshold), fl,
                                             int(id_base), C.byref(exchange), scores.ctypes.data, ids.ctypes.data,
                                             C.byref(status)))
        return ids, scores, bool(status.value)

    # -- measurement ------------------------------------------------------------ #
    def profile(self, enable: bool = True) -> None:
        """Bracket the dominant kernel of every later search with CUDA events (bench.py's roofline leg)."""
        check(self._lib.ragb_store_profile(self._h, 1 if enable else 0))

    def profile_read(self) -> Tuple[float, int]:
        """(summed kernel milliseconds, bracketed launches) since the last read; waits for the events."""
        ms, n = C.c_double(0.0), C.c_int64(0)
        check(self._lib.ragb_store_profile_read(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value


class ShardedDeviceStore:
    """The row-sharded store (include/ragb200.h: ragb_mstore_*): ONE process, rows dealt over several GPUs of one
    NVSwitch box in blocks, round-robin; the last kernel of every shard stores its k candidates per query into the
    first device's gather buffer over NVLink (peer memory), one merge kernel finishes.  Same host interface as
    ``DeviceStore`` for everything ``B200VectorStore`` uses; results are identical to the one-device store's."""

    def __init__(self, dim: int, dtype: str = "fp16", capacity: int = 1 << 16, devices=(0,), block_rows: int = 0):
        self._h = C.c_void_p()
        self._lib = lib()
        code = _DTYPES[dtype] if isinstance(dtype, str) else int(dtype)
        devs = (C.c_int * len(devices))(*[int(d) for d in devices])
        check(self._lib.ragb_mstore_create(int(dim), code, int(capacity), devs, len(devices), int(block_rows), C.byref(self._h)))
        self.dim, self.dtype_code, self.devices = int(dim), code, [int(d) for d in devices]
        self.device = self.devices[0]

    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.ragb_mstore_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __len__(self) -> int:
        return int(self._lib.ragb_mstore_size(self._h))

    @property
    def capacity(self) -> int:
        return int(self._lib.ragb_mstore_capacity(self._h))

    @property
    def block_rows(self) -> int:
        return int(self._lib.ragb_mstore_block_rows(self._h))

    @property
    def dtype(self) -> str:
        return {F16: "fp16", BF16: "bf16", F32: "fp32"}[self.dtype_code]

    def device_rows(self) -> list:
        return [int(self._lib.ragb_mstore_device_rows(self._h, i)) for i in range(len(self.devices))]

    def reserve(self, capacity: int) -> None:
        check(self._lib.ragb_mstore_reserve(self._h, int(capacity)))

    def clear(self) -> None:
        check(self._lib.ragb_mstore_clear(self._h))

    def _room(self, n: int) -> None:
        need = len(self) + n
        if need > self.capacity:
            self.reserve(max(need, 2 * self.capacity, 1024))

    def add(self, rows: np.ndarray) -> int:
        rows = np.ascontiguousarray(np.atleast_2d(rows), dtype=np.float32)
        if rows.shape[1] != self.dim:
            raise ValueError(f"rows have dim {rows.shape[1]}, store has {self.dim}")
        self._room(rows.shape[0])
        first = C.c_int64(-1)
        check(self._lib.ragb_mstore_add(self._h, rows.ctypes.data, rows.shape[0], C.byref(first)))
        return first.value

    def add_dev(self, rows) -> int:
        """torch CUDA tensor [n, dim] on ANY device of the box: every shard's ingest kernel reads its blocks in place."""
        if rows.dim() != 2 or rows.shape[1] != self.dim or rows.stride(1) != 1:
            raise ValueError("expected a [n, dim] tensor with unit inner stride")
        self._room(rows.shape[0])
        first = C.c_int64(-1)
        check(self._lib.ragb_mstore_add_dev(self._h, rows.data_ptr(), _torch_code(rows), rows.shape[0], rows.stride(0),
                                            C.byref(first), _cur_stream(rows.device.index or 0)))
        return first.value

    def set(self, row: int, vec: np.ndarray) -> None:
        vec = np.ascontiguousarray(vec, dtype=np.float32).reshape(-1)
        if vec.shape[0] != self.dim:
            raise ValueError("dimension mismatch")
        check(self._lib.ragb_mstore_set(self._h, int(row), vec.ctypes.data))

    def get_rows(self, row0: int = 0, n: Optional[int] = None) -> Tuple[np.ndarray, np.ndarray]:
        n = len(self) - row0 if n is None else n
        rows = np.empty((n, self.dim), dtype=_np_elem(self.dtype_code))
        norms = np.empty(n, dtype=np.float32)
        check(self._lib.ragb_mstore_get_rows(self._h, int(row0), int(n), rows.ctypes.data, norms.ctypes.data))
        return rows, norms

    def put_rows(self, rows: np.ndarray, norms: np.ndarray) -> int:
        rows = np.ascontiguousarray(rows, dtype=_np_elem(self.dtype_code))
        norms = np.ascontiguousarray(norms, dtype=np.float32)
        if rows.ndim != 2 or rows.shape[1] != self.dim or norms.shape[0] != rows.shape[0]:
            raise ValueError("expected packed rows [n, dim] and norms [n]")
        self._room(rows.shape[0])
        first = C.c_int64(-1)
        check(self._lib.ragb_mstore_put_rows(self._h, rows.ctypes.data, norms.ctypes.data, rows.shape[0], C.byref(first)))
        return first.value

    def set_meta(self, row0: int, flags: Optional[np.ndarray] = None, src: Optional[np.ndarray] = None) -> None:
        n = len(flags) if flags is not None else len(src)
        f = None if flags is None else np.ascontiguousarray(flags, dtype=np.uint8)
        c = None if src is None else np.ascontiguousarray(src, dtype=np.uint16)
        check(self._lib.ragb_mstore_set_meta(self._h, int(row0), int(n), None if f is None else f.ctypes.data,
                                             None if c is None else c.ctypes.data))

    def set_timestamps(self, row0: int, ts: np.ndarray) -> None:
        t = np.ascontiguousarray(ts, dtype=np.float64)
        check(self._lib.ragb_mstore_set_timestamps(self._h, int(row0), int(len(t)), t.ctypes.data))

    def build_filter(self, allowed_sources: Optional[np.ndarray] = None, window: Optional[int] = None) -> int:
        tot = C.c_int64(0)
        a = None if allowed_sources is None else np.ascontiguousarray(allowed_sources, dtype=np.uint8)
        check(self._lib.ragb_mstore_build_filter(self._h, None if a is None else a.ctypes.data, 0 if a is None else len(a),
                                                 -1 if window is None else int(window), C.byref(tot)))
        return int(tot.value)

    def search(self, q: np.ndarray, k: int, metric: str = "ip", threshold: float = NEG_INF, clamp01: bool = False,
               row_begin: int = 0, row_end: int = -1, valid=None, flags: int = 0) -> Tuple[np.ndarray, np.ndarray]:
        if row_begin != 0 or (row_end >= 0 and row_end != len(self)):
            raise ValueError("the sharded store searches all rows: restrict with a valid mask or build_filter")
        q = np.ascontiguousarray(np.atleast_2d(q), dtype=np.float32)
        nq = q.shape[0]
        if q.shape[1] != self.dim:
            q = np.zeros((nq, self.dim), dtype=np.float32)
        ids = np.empty((nq, k), dtype=np.int64)
        scores = np.empty((nq, k), dtype=np.float32)
        vp = None
        if isinstance(valid, str):
            if valid != "filter":
                raise ValueError('valid must be a byte mask or the string "filter"')
            vp = _lib.STORE_FILTER
        elif valid is not None:
            valid = np.ascontiguousarray(valid, dtype=np.uint8)
            if valid.shape[0] < len(self):
                raise ValueError("valid mask shorter than the store")
            vp = valid.ctypes.data
        fl = int(flags) | (CLAMP01 if clamp01 else 0)
        check(self._lib.ragb_mstore_search(self._h, q.ctypes.data, nq, int(k), _METRICS[metric], float(threshold), fl, vp,
                                           scores.ctypes.data, ids.ctypes.data))
        return ids, scores

    def search_dev(self, q, k: int, metric: str = "ip", threshold: float = NEG_INF, clamp01: bool = False, valid=None,
                   flags: int = 0, out=None):
        """torch CUDA fp32 queries [nq, dim] on the FIRST device -> (ids, scores) there, on its current stream."""
        import torch
        if q.dtype != torch.float32 or q.dim() != 2 or q.shape[1] != self.dim or not q.is_contiguous():
            raise ValueError(f"queries must be contiguous fp32 [nq,{self.dim}]")
        if (q.device.index or 0) != self.devices[0]:
            raise ValueError(f"queries must live on the store's first device (cuda:{self.devices[0]})")
        nq = q.shape[0]
        if out is None:
            ids = torch.empty((nq, k), dtype=torch.int64, device=q.device)
            scores = torch.empty((nq, k), dtype=torch.float32, device=q.device)
        else:
            ids, scores = out
        fl = int(flags) | (CLAMP01 if clamp01 else 0)
        check(self._lib.ragb_mstore_search_dev(self._h, q.data_ptr(), nq, int(k), _METRICS[metric], float(threshold), fl,
                                               None if valid is None else _lib.STORE_FILTER, scores.data_ptr(), ids.data_ptr(),
                                               _cur_stream(self.devices[0])))
        return ids, scores

    def profile(self, enable: bool = True) -> None:
        check(self._lib.ragb_mstore_profile(self._h, 1 if enable else 0))

    def profile_read(self) -> Tuple[float, int]:
        ms, n = C.c_double(0.0), C.c_int64(0)
        check(self._lib.ragb_mstore_profile_read(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value


def kpad_of(k: int) -> int:
    return int(lib().ragb_kpad(int(k)))


def merge_keys_dev(keys, k: int, device: int):
    """keys: int64 CUDA tensor [n_lists, nq, kpad] of sorted key lists -> (ids [nq,k], scores [nq,k])."""
    import torch
    n_lists, nq, kpad = keys.shape
    ids = torch.empty((nq, k), dtype=torch.int64, device=keys.device)
    scores = torch.empty((nq, k), dtype=torch.float32, device=keys.device)
    check(lib().ragb_merge_keys_dev(int(device), keys.data_ptr(), nq, n_lists, kpad, kpad, nq * kpad, int(k),
                                    scores.data_ptr(), ids.data_ptr(), _cur_stream(device)))
    return ids, scores


def gather_merge_dev(exchange, nq: int, k: int, device: int, status=None, out=None):
    """Wait until every rank's keys of `exchange.epoch` have arrived in this rank's gather buffer, merge them.
    status: optional int32 CUDA tensor [1], set to 1 by the kernel if a rank did not arrive within ~2 s."""
    import torch
    dev = torch.device("cuda", device)
    ids, scores = out if out is not None else (torch.empty((nq, k), dtype=torch.int64, device=dev),
                                               torch.empty((nq, k), dtype=torch.float32, device=dev))
    check(lib().ragb_gather_merge_dev(int(device), C.byref(exchange), int(nq), int(k), scores.data_ptr(), ids.data_ptr(),
                                      status.data_ptr() if status is not None else None, _cur_stream(device)))
    return ids, scores


def pool_norm_pack_dev(tok, mask, out_dtype: str = "fp32"):
    """Stateless K1 (csrc/ingest.cuh): masked mean over tokens -> L2 normalise -> pack, torch CUDA tensors in and out.
    tok [B,S,H] fp16/bf16/fp32 contiguous, mask [B,S] (u8/bool/i32/i64/f32) or None.  Returns (rows [B,H] of
    ``out_dtype``, norms [B] fp32 — the norm of the pooled vector before normalising; 0 marks an all-masked chunk).
    Replaces the tail of ``SentenceTransformer.encode`` (embedding_pipeline.py:424, ollama_rag_system.py:163)."""
    import torch
    if tok.dim() != 3 or not tok.is_contiguous() or not tok.is_cuda:
        raise ValueError("token embeddings must be a contiguous CUDA tensor [B,S,H]")
    B, S, H = tok.shape
    mp, mc = None, 0
    if mask is not None:
        if tuple(mask.shape) != (B, S) or not mask.is_contiguous():
            raise ValueError("mask must be contiguous [B,S]")
        mp, mc = mask.data_ptr(), _torch_mask_code(mask)
    tdt = {"fp16": torch.float16, "bf16": torch.bfloat16, "fp32": torch.float32}[out_dtype]
    ld = (H + 7) // 8 * 8
    dev = tok.device.index or 0
    out = torch.empty((B, ld), dtype=tdt, device=tok.device)
    norms = torch.empty((B,), dtype=torch.float32, device=tok.device)
    if B:
        check(lib().ragb_pool_norm_pack_dev(dev, tok.data_ptr(), _torch_code(tok), mp, mc, B, S, H, out.data_ptr(),
                                            _DTYPES[out_dtype], ld, norms.data_ptr(), _cur_stream(dev)))
    return out[:, :H], norms


def launch_count() -> int:
    return int(lib().ragb_launch_count())

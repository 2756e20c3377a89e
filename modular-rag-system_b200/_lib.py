"""ctypes binding of include/ragb200.h (libragb200.so, built in-tree for sm_100a).

This is the binding INTEGRATION.md shows a maintainer of the reference adding; there
is no CPU fallback — if the shared library is missing, loading fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

CSRC = Path(__file__).resolve().parent / "csrc"
LIB_PATH = Path(os.environ["RAGB200_LIB"]) if os.environ.get("RAGB200_LIB") else CSRC / "libragb200.so"  # override: A/B runs of build variants

ABI_VERSION = 2
F16, BF16, F32 = 0, 1, 2
MASK_U8, MASK_I32, MASK_I64, MASK_F32 = 0, 1, 2, 3
IP, COSINE = 0, 1
CLAMP01, FORCE_SCAN, FORCE_GEMM, NO_RESCORE = 1, 2, 4, 8
OK, EINVAL, ECUDA, ENOMEM, EFULL, EUNSUPPORTED = 0, -1, -2, -3, -4, -5

_vp, _i, _i64, _f = C.c_void_p, C.c_int, C.c_int64, C.c_float
_pi64 = C.POINTER(C.c_int64)

MAX_PEERS = 8
STORE_FILTER = C.c_void_p(1)   # RAGB_STORE_FILTER: "use the mask ragb_store_build_filter left in the store"


class Exchange(C.Structure):
    """struct ragb_exchange (include/ragb200.h): peer-mapped gather buffers and arrival words of one epoch."""
    _fields_ = [("world", C.c_int), ("rank", C.c_int), ("keys", C.c_void_p * MAX_PEERS),
                ("flags", C.c_void_p * MAX_PEERS), ("epoch", C.c_uint64)]


# name -> (restype, argtypes); every symbol include/ragb200.h declares
SIGNATURES = {
    "ragb_abi_version": (_i, []),
    "ragb_last_error": (C.c_char_p, []),
    "ragb_launch_count": (_i64, []),
    "ragb_device_count": (_i, []),
    "ragb_store_create": (_i, [_i, _i, _i, _i64, C.POINTER(_vp)]),
    "ragb_store_destroy": (_i, [_vp]),
    "ragb_store_reserve": (_i, [_vp, _i64]),
    "ragb_store_clear": (_i, [_vp]),
    "ragb_store_size": (_i64, [_vp]),
    "ragb_store_capacity": (_i64, [_vp]),
    "ragb_store_dim": (_i, [_vp]),
    "ragb_store_ld": (_i, [_vp]),
    "ragb_store_dtype": (_i, [_vp]),
    "ragb_store_device": (_i, [_vp]),
    "ragb_store_rows_dev": (_vp, [_vp]),
    "ragb_store_norms_dev": (_vp, [_vp]),
    "ragb_store_add": (_i, [_vp, _vp, _i64, _pi64]),
    "ragb_store_add_dev": (_i, [_vp, _vp, _i, _i64, _i64, _pi64, _vp]),
    "ragb_store_set": (_i, [_vp, _i64, _vp]),
    "ragb_store_add_pooled": (_i, [_vp, _vp, _i, _vp, _i, _i64, _i, _i, _pi64]),
    "ragb_store_add_pooled_dev": (_i, [_vp, _vp, _i, _vp, _i, _i64, _i, _i, _pi64, _vp]),
    "ragb_norm_pack_dev": (_i, [_i, _vp, _i, _i64, _i, _i64, _vp, _i, _i64, _vp, _vp]),
    "ragb_pool_norm_pack_dev": (_i, [_i, _vp, _i, _vp, _i, _i64, _i, _i, _vp, _i, _i64, _vp, _vp]),
    "ragb_store_get_rows": (_i, [_vp, _i64, _i64, _vp, _vp]),
    "ragb_store_put_rows": (_i, [_vp, _vp, _vp, _i64, _pi64]),
    "ragb_search": (_i, [_vp, _vp, _i64, _i, _i, _f, _i, _i64, _i64, _vp, _vp, _vp]),
    "ragb_search_dev": (_i, [_vp, _vp, _i64, _i, _i, _f, _i, _i64, _i64, _vp, _vp, _vp, _vp]),
    "ragb_scan_topk_static_dev": (_i, [_i, _vp, _i, _i64, _i, _i64, _vp, _i64, _i, _i, _f, _i, _vp, _vp, _vp, _vp]),
    "ragb_store_set_meta": (_i, [_vp, _i64, _i64, _vp, _vp]),
    "ragb_store_build_filter": (_i, [_vp, _vp, _i, _i64, _pi64]),
    "ragb_store_set_timestamps": (_i, [_vp, _i64, _i64, _vp]),
    "ragb_store_profile": (_i, [_vp, _i]),
    "ragb_store_profile_read": (_i, [_vp, C.POINTER(C.c_double), _pi64]),
    "ragb_kpad": (_i, [_i]),
    "ragb_search_keys_dev": (_i, [_vp, _vp, _i64, _i, _i, _f, _i, _i64, _i64, _vp, C.c_uint32, _vp, _vp]),
    "ragb_merge_keys_dev": (_i, [_i, _vp, _i64, _i, _i, _i64, _i64, _i, _vp, _vp, _vp]),
    "ragb_search_scatter_dev": (_i, [_vp, _vp, _i64, _i, _i, _f, _i, _i64, _i64, _vp, C.c_uint32, C.POINTER(Exchange), _vp]),
    "ragb_gather_merge_dev": (_i, [_i, C.POINTER(Exchange), _i64, _i, _vp, _vp, _vp, _vp]),
    "ragb_search_exchange": (_i, [_vp, _vp, _i64, _i, _i, _f, _i, C.c_uint32, C.POINTER(Exchange), _vp, _vp, C.POINTER(_i)]),
    # row-sharded store (one process, several GPUs)
    "ragb_mstore_create": (_i, [_i, _i, _i64, C.POINTER(_i), _i, _i64, C.POINTER(_vp)]),
    "ragb_mstore_destroy": (_i, [_vp]),
    "ragb_mstore_reserve": (_i, [_vp, _i64]),
    "ragb_mstore_clear": (_i, [_vp]),
    "ragb_mstore_size": (_i64, [_vp]),
    "ragb_mstore_capacity": (_i64, [_vp]),
    "ragb_mstore_block_rows": (_i64, [_vp]),
    "ragb_mstore_n_devices": (_i, [_vp]),
    "ragb_mstore_device": (_i, [_vp, _i]),
    "ragb_mstore_device_rows": (_i64, [_vp, _i]),
    "ragb_mstore_add": (_i, [_vp, _vp, _i64, _pi64]),
    "ragb_mstore_add_dev": (_i, [_vp, _vp, _i, _i64, _i64, _pi64, _vp]),
    "ragb_mstore_set": (_i, [_vp, _i64, _vp]),
    "ragb_mstore_get_rows": (_i, [_vp, _i64, _i64, _vp, _vp]),
    "ragb_mstore_put_rows": (_i, [_vp, _vp, _vp, _i64, _pi64]),
    "ragb_mstore_set_meta": (_i, [_vp, _i64, _i64, _vp, _vp]),
    "ragb_mstore_set_timestamps": (_i, [_vp, _i64, _i64, _vp]),
    "ragb_mstore_build_filter": (_i, [_vp, _vp, _i, _i64, _pi64]),
    "ragb_mstore_search": (_i, [_vp, _vp, _i64, _i, _i, _f, _i, _vp, _vp, _vp]),
    "ragb_mstore_search_dev": (_i, [_vp, _vp, _i64, _i, _i, _f, _i, _vp, _vp, _vp, _vp]),
    "ragb_mstore_profile": (_i, [_vp, _i]),
    "ragb_mstore_profile_read": (_i, [_vp, C.POINTER(C.c_double), _pi64]),
}


class RagbError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"ragb200 error {code}: {msg}")
        self.code = code


def build(force: bool = False) -> Path:
    """Compile csrc/*.cu -> csrc/libragb200.so for sm_100a (make -j: one translation unit per kernel family; nvcc
    cross-compiles without a GPU)."""
    srcs = (list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h")) + [CSRC / "Makefile"] +
            [CSRC.parent.parent / "include" / "ragb200.h"])
    if not force and LIB_PATH.exists() and all(LIB_PATH.stat().st_mtime >= s.stat().st_mtime for s in srcs):
        return LIB_PATH
    subprocess.run(["bash", str(CSRC / "build.sh")], check=True)
    return LIB_PATH


_lib = None


def lib() -> C.CDLL:
    """The loaded library.  Raises if it has not been built: no fallback exists."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise ImportError(
                f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(or modular-rag-system_b200/csrc/build.sh).  There is no CPU fallback.")
        l = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype, fn.argtypes = res, args
        if l.ragb_abi_version() != ABI_VERSION:
            raise ImportError("libragb200.so ABI version mismatch; rebuild")
        _lib = l
    return _lib


def check(code: int) -> None:
    if code != 0:
        raise RagbError(code, lib().ragb_last_error().decode("utf-8", "replace"))

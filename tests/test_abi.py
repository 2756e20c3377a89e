"""The C-ABI library loads without a GPU and exports every symbol include/ragb200.h declares."""
import ctypes
import re
from pathlib import Path

import pytest

import modular_rag_system_b200 as pkg
from modular_rag_system_b200 import _lib

ROOT = Path(__file__).resolve().parent.parent


def _declared():
    txt = (ROOT / "include" / "ragb200.h").read_text()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(ragb_[a-z0-9_]+)\s*\(", txt)))


def test_header_and_binding_agree():
    assert _declared() == sorted(_lib.SIGNATURES)


def test_library_exports_every_symbol():
    pkg.build()
    l = ctypes.CDLL(str(_lib.LIB_PATH))
    for name in _declared():
        assert hasattr(l, name), name
    assert _lib.lib().ragb_abi_version() == _lib.ABI_VERSION == 2
    assert _lib.lib().ragb_kpad(10) == 16 and _lib.lib().ragb_kpad(100) == 128


def test_no_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pkg.RagbError):
        pkg.DeviceStore(8, "fp16", 4)


def test_argument_errors_do_not_need_a_gpu():
    h = ctypes.c_void_p()
    assert _lib.lib().ragb_store_create(0, 0, 0, 4, ctypes.byref(h)) == _lib.EINVAL
    assert b"dim" in _lib.lib().ragb_last_error()
    assert _lib.lib().ragb_store_create(0, 8, 7, 4, ctypes.byref(h)) == _lib.EINVAL


def test_product_never_imports_the_oracle():
    for p in (ROOT / "modular-rag-system_b200").rglob("*.py"):
        src = p.read_text()
        assert "import oracle" not in src and "from oracle" not in src, p

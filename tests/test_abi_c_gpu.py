"""The C ABI used from plain C (gcc + include/ragb200.h + libragb200.so — the binding a non-Python maintainer of the
reference's callers would write): concurrent reader threads, no GIL in the way.  SURVEY §8(b) threading contract."""
import re
import subprocess
from pathlib import Path

import pytest

from modular_rag_system_b200 import _lib

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def test_reader_threads_scale_through_the_c_abi(tmp_path):
    exe = tmp_path / "abi_readers"
    subprocess.run(["gcc", "-O2", "-std=c11", "-D_POSIX_C_SOURCE=200809L", "-I", str(ROOT / "include"),
                    str(ROOT / "tests" / "abi_readers.c"), "-o", str(exe), str(_lib.LIB_PATH), "-lpthread", "-lm",
                    f"-Wl,-rpath,{_lib.LIB_PATH.parent}"], check=True)
    out = subprocess.run([str(exe), "1.0"], check=True, capture_output=True, text=True, timeout=120).stdout
    print(out)
    runs = {int(t): (float(q), int(b)) for t, q, b in re.findall(r"threads (\d+) qps (\d+) bad (\d+)", out)}
    assert all(bad == 0 for _, bad in runs.values()), out
    assert runs[8][0] >= 4.0 * runs[1][0], out       # >= 4x the single-thread query rate

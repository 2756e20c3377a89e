#!/usr/bin/env python3
"""Recipe for ``oracle/_ref/``: the UNMODIFIED reference modules, copied verbatim (TEST INFRASTRUCTURE).

    python oracle/fetch_ref.py            # authoring container only: needs /root/reference

The reference is pure Python: there is nothing to compile.  What the GPU box needs in order to run the drop-in
through the reference's REAL classes (``RAGSystem``, ``EnhancedRAGSystem``, ``EmbeddingPipeline``) and to time the
literal ``RAGVectorStore.search`` as the CPU baseline is the three files the north_star names as entry points.
They are copied byte for byte into the git-ignored ``oracle/_ref/`` (which travels with ``gpurun``; nothing under
it is ever committed) and checked against the SHA-256 of reference commit bf4dc62 — a file that differs is
refused, so "unmodified" is enforced, not assumed.

Users: ``tests/test_dropin_gpu.py`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs, through
``oracle.ref_path()``.  The product package never imports from here.
"""
from __future__ import annotations

import hashlib
import shutil
import sys
from pathlib import Path

REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent / "_ref"

# sha256 of the files at reference commit bf4dc62
FILES = {
    "ollama_rag_system.py": "31ed92e3111501b2d411ff1e57a42ee32dbbe32c03b0d45df5afe76ae4790d35",
    "embedding_pipeline.py": "42e76da2679543f0eff7498e0033f8feaa9f0540ec59bf956eb89fabe313bb60",
    "enhanced_rag_system.py": "79552bd859a24e4971ba61a41f1d62250f59c9fd158e7991f7e1421d26299d07",
}


def sha256(p: Path) -> str:
    return hashlib.sha256(p.read_bytes()).hexdigest()


def verify(root: Path = OUT) -> bool:
    """True when every file is present under ``root`` and is the reference's, bit for bit."""
    return all((root / n).exists() and sha256(root / n) == h for n, h in FILES.items())


def fetch() -> bool:
    """Copy the reference files into oracle/_ref/.  Returns False (and changes nothing) where /root/reference does
    not exist (the GPU box: it only uses the copy that travelled)."""
    if not REF.exists():
        return False
    OUT.mkdir(exist_ok=True)
    for name, want in FILES.items():
        got = sha256(REF / name)
        if got != want:
            raise RuntimeError(f"{REF / name}: sha256 {got} is not the surveyed reference ({want})")
        shutil.copyfile(REF / name, OUT / name)
    return verify()


if __name__ == "__main__":
    ok = fetch()
    print(f"oracle/_ref: {'ok' if ok else 'reference tree not present; nothing copied'}"
          + (f" ({', '.join(FILES)})" if ok else ""))
    sys.exit(0 if ok or not REF.exists() else 1)

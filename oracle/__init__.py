"""CPU oracle for the retrieval hot path — TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is product code.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl
reference`` legs may import it, and only as the checker or as the timed CPU
baseline.  The product package (``modular-rag-system_b200``) never imports it
and has no CPU fallback.

Parity status
-------------
* search / similarity / sort   : PINNED — ``oracle/make_golden.py`` runs the
  unmodified reference (``/root/reference/ollama_rag_system.py``) in the
  authoring container and commits its outputs as ``tests/golden/search_*.npz``;
  ``tests/test_oracle.py`` checks every function here against them.
* ingest normalise (a4)        : PINNED to the two numpy lines
  ``ollama_rag_system.py:379-380`` (fixtures ``tests/golden/normpack_*.npz``
  were produced by executing those expressions on numpy 2.3.5).
* masked mean-pool + normalise : PARITY UNPINNED — the arithmetic lives in the
  third-party ``sentence-transformers>=2.2.0`` (``requirements.txt:11``,
  unpinned, un-vendored, not installed here; call sites
  ``embedding_pipeline.py:424`` and ``ollama_rag_system.py:163``).  The
  restatement follows that library's published ``Pooling(mean)`` +
  ``Normalize`` modules.
* ``oracle/_ref/``             : the UNMODIFIED reference modules (git-ignored;
  ``oracle/fetch_ref.py`` copies them from ``/root/reference`` and refuses a
  file whose SHA-256 is not the surveyed commit's).  ``ref_path()`` gives the
  directory to the drop-in tests and to bench.py's literal CPU baseline.
"""

from .ref_search import (  # noqa: F401
    similarity_literal,
    search_literal,
    search_f64,
    search_f64_blocked,
)
from .ref_ingest import (  # noqa: F401
    normalize_f32,
    pool_normalize_f32,
    pack_f16,
    pack_bf16,
    unpack_bf16,
)


def ref_path():
    """Directory holding the unmodified reference modules (oracle/_ref, see fetch_ref.py), or None when the copy is
    absent or not bit-identical to the surveyed reference."""
    from . import fetch_ref
    return str(fetch_ref.OUT) if fetch_ref.verify() else None

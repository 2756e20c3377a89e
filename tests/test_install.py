"""install() patches the unmodified reference module (only where /root/reference exists; no GPU needed)."""
import sys
from pathlib import Path

import pytest

REF = Path("/root/reference")


@pytest.mark.skipif(not (REF / "ollama_rag_system.py").exists(), reason="reference tree not present on this box")
def test_install_patches_reference(monkeypatch, capsys):
    monkeypatch.syspath_prepend(str(REF))
    import importlib
    ref = importlib.import_module("ollama_rag_system")
    orig_store = ref.RAGVectorStore
    assert not hasattr(ref.RAGSystem, "add_documents")       # called at four sites, never defined
    import modular_rag_system_b200 as pkg
    try:
        pkg.install(window=None)
        assert issubclass(ref.RAGVectorStore, pkg.B200VectorStore)
        assert callable(ref.RAGSystem.add_documents)
        # same public surface as the class it replaces
        for name in ("add_document", "search", "get_stats"):
            assert callable(getattr(ref.RAGVectorStore, name))
    finally:
        ref.RAGVectorStore = orig_store
        if hasattr(ref.RAGSystem, "add_documents"):
            del ref.RAGSystem.add_documents
        sys.modules.pop("ollama_rag_system", None)

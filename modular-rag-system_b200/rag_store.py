"""B200VectorStore — drop-in for the reference's ``RAGVectorStore`` (ollama_rag_system.py:225-554).

Same constructor, methods, attributes, result types and error convention as the reference class;
what changes is where the vectors live and who scans them:

* content + metadata stay in the reference's own SQLite table (same schema, :275-286);
* every embedding is also normalised + packed onto the GPU by the ingest kernel — the slot the
  reference fills with ``faiss.IndexFlatIP`` and never searches (:298-305, :372-386);
* ``search`` runs the exact scan / threshold / sort / slice of ``_search_sqlite`` (:480-515) on the
  device (C ABI ``ragb_search``), then reads the winners' content rows back from SQLite by id.

Reference *policy* is kept in this shim with the reference's defaults: cosine
clamped to [0,1] (:218), ``similarity_threshold`` with ``>=`` (:495), the newest-1000-rows window
(:474), ``source_filters`` (:466-469), newest-first order among equal scores (:474, :512),
``access_count + 1`` on returned documents (:504), ``total_documents_searched`` = rows the SQL query
would have returned (:520).  ``window=None`` lifts the 1000-row cap.

Nothing here raises through ``add_document`` / ``search``: failures print and return ``False`` / an
empty ``RAGResult`` exactly as the reference does (:319-321, :523-525).  There is no CPU scoring
path: without the CUDA library the constructor fails loudly.
"""
from __future__ import annotations

import hashlib
import json
import pickle
import sqlite3
import threading
import time
from dataclasses import asdict, dataclass
from pathlib import Path
from typing import Any, Dict, Iterable, List, Optional, Sequence

import numpy as np

from .store import DeviceStore, ShardedDeviceStore

NEG_INF = float("-inf")


# --------------------------------------------------------------------------- #
# boundary types: field-for-field mirrors of ollama_rag_system.py:46-93 (duck-compatible with the
# reference's own dataclasses, which are accepted wherever these are)
# --------------------------------------------------------------------------- #
@dataclass
class RAGDocument:
    id: str
    content: str
    metadata: Dict[str, Any]
    embedding: Optional[List[float]] = None
    timestamp: float = 0.0
    source: str = "unknown"
    importance_score: float = 1.0
    access_count: int = 0

    def __post_init__(self):
        if self.timestamp == 0.0:
            self.timestamp = time.time()
        if not self.id:
            self.id = self.generate_id()

    def generate_id(self) -> str:  # :61-68
        content_hash = hashlib.md5(self.content.encode()).hexdigest()[:8]
        time_hash = hashlib.md5(str(self.timestamp).encode()).hexdigest()[:4]
        return f"{self.source}_{content_hash}_{time_hash}"

    def to_dict(self) -> Dict:
        return asdict(self)


@dataclass
class RAGQuery:
    query_text: str
    context: Dict[str, Any]
    max_results: int = 5
    similarity_threshold: float = 0.7
    time_weight: float = 0.1            # :81 — carried, never read by the reference's search either
    source_filters: Optional[List[str]] = None
    metadata_filters: Optional[Dict[str, Any]] = None


@dataclass
class RAGResult:
    documents: List[RAGDocument]
    similarities: List[float]
    query_embedding: Optional[List[float]] = None
    retrieval_time: float = 0.0
    total_documents_searched: int = 0


class CacheOnlyEmbedder:
    """Stand-in for the reference's ``EmbeddingGenerator`` (:96-197) when that module is not importable.
    The encoder forward / Ollama HTTP call is out of scope: this object only serves vectors that were
    put into ``embedding_cache`` (the reference checks the same cache first, :116-117)."""

    def __init__(self):
        self.embedding_cache: Dict[str, List[float]] = {}

    def generate_embedding(self, text: str) -> List[float]:
        if text in self.embedding_cache:
            return self.embedding_cache[text]
        raise RuntimeError("no encoder configured: put the vector into embedding_generator.embedding_cache "
                           "or pass an embedding_generator with generate_embedding(text)")


def _default_embedder():
    try:  # the reference's own generator when its module is importable (unchanged, out of scope)
        from ollama_rag_system import EmbeddingGenerator  # type: ignore
        return EmbeddingGenerator()
    except Exception:
        return CacheOnlyEmbedder()


class B200VectorStore:
    """Vector storage and retrieval for RAG with the embedding store on a B200 (see module docstring)."""

    def __init__(self, storage_path: Path, use_chromadb: bool = True, *, dtype: str = "fp16", device: int = 0,
                 devices: Optional[Sequence[int]] = None, window: Optional[int] = 1000, capacity: int = 4096,
                 embedding_generator=None, result_types=None):
        self.storage_path = Path(storage_path)
        self.storage_path.mkdir(parents=True, exist_ok=True)
        self.use_chromadb = False            # the exact path is the SQLite + device path
        self.embedding_generator = embedding_generator or _default_embedder()
        self.faiss_index = None              # the device store takes this slot (:242, :298-305)
        self.document_mapping: Dict[int, str] = {}   # device row -> doc id (:243, :383)
        self.window = window
        # devices=[0, 1, ...]: the rows are sharded over these GPUs (ShardedDeviceStore); else one GPU (`device`)
        self.devices = [int(d) for d in devices] if devices else [int(device)]
        self.dtype, self.device = dtype, self.devices[0]
        self._capacity0 = int(capacity)
        self._types = result_types or (RAGDocument, RAGResult)
        self._lock = threading.RLock()
        self._store: Optional[DeviceStore] = None     # created when the first vector fixes the dimension (D4)
        self._n = 0
        self._ts = np.zeros(0, np.float64)            # per device row
        self._src = np.zeros(0, np.int32)
        self._live = np.zeros(0, bool)                # False once the id was replaced (:351)
        self._has = np.zeros(0, bool)                 # False for NULL embeddings (:485-492)
        self._row_of: Dict[str, int] = {}
        self._src_code: Dict[str, int] = {}
        self._meta_version = 0
        self._filter_key = None                       # (meta version, sources, window) of the mask now in the store
        self._filter_total = 0
        # searches run outside the lock (the GUI spawns a thread per query, rag_query_interface.py:740-744): readers
        # of the current device mask are counted; whoever needs another mask, or appends, waits until they are out
        self._cv = threading.Condition(self._lock)
        self._readers = 0
        self._writers_waiting = 0
        self._init_sqlite_fallback()
        self.restored_from = "sqlite"                 # "sidecar" when the packed rows came back from disk
        if not self._load_sidecar():
            self._rebuild_from_sqlite()
        print(f"✅ RAG Vector Store initialized: {storage_path}")

    # -- storage (same table as the reference, :270-296) ------------------------------------ #
    def _init_sqlite_fallback(self):
        self.db_path = self.storage_path / "rag_documents.db"
        with sqlite3.connect(self.db_path) as conn:
            conn.execute("""
                CREATE TABLE IF NOT EXISTS documents (
                    id TEXT PRIMARY KEY, content TEXT NOT NULL, metadata TEXT, embedding BLOB,
                    timestamp REAL, source TEXT, importance_score REAL, access_count INTEGER)""")
            conn.execute("CREATE INDEX IF NOT EXISTS idx_timestamp ON documents(timestamp);")
            conn.execute("CREATE INDEX IF NOT EXISTS idx_source ON documents(source);")

    def _rebuild_from_sqlite(self):
        """The device store is a cache of the table: refill it on open (the reference's FAISS index is
        volatile and never rebuilt, :298-305)."""
        with sqlite3.connect(self.db_path) as conn:
            rows = conn.execute("SELECT id, embedding, timestamp, source FROM documents ORDER BY rowid").fetchall()
        batch: List[tuple] = []
        for doc_id, blob, ts, source in rows:
            emb = None
            if blob:
                try:
                    emb = pickle.loads(blob)
                except Exception:
                    emb = None
            batch.append((doc_id, emb, ts, source))
        if batch:
            self._append_rows(batch)

    # -- packed on-disk sidecar ----------------------------------------- #
    # The reference persists vectors only as pickled Python lists inside SQLite (3 472 B per 384-d vector,
    # ~20 us to unpickle, :358 / :487) and its FAISS index is volatile (:298-305).  On close the device rows go
    # to disk exactly as they sit in HBM (fp16/bf16 rows + fp32 norms + the row table); on open they come back
    # with one copy — if the table has not changed since (fingerprint), else the table wins and is re-read.
    def _fingerprint(self) -> List[float]:
        with sqlite3.connect(self.db_path) as conn:
            row = conn.execute("SELECT COUNT(*), COALESCE(MAX(rowid), 0), COALESCE(SUM(timestamp), 0.0), "
                               "COALESCE(SUM(LENGTH(embedding)), 0) FROM documents").fetchone()
        return [float(x) for x in row]

    def _sidecar_paths(self):
        return (self.storage_path / "b200_rows.npy", self.storage_path / "b200_norms.npy", self.storage_path / "b200_meta.npz")

    def save_sidecar(self) -> bool:
        """Write the packed device rows + row table next to the SQLite file; returns False (and prints) on failure."""
        try:
            with self._lock:
                if self._store is None or self._n == 0 or getattr(self, "_dim_is_placeholder", False):
                    return False
                n = self._n
                rows, norms = self._store.get_rows(0, n)
                p_rows, p_norms, p_meta = self._sidecar_paths()
                np.save(p_rows, rows)
                np.save(p_norms, norms)
                names = [None] * len(self._src_code)
                for name, code in self._src_code.items():
                    names[code] = name
                np.savez(p_meta, ids=np.array([self.document_mapping[r] for r in range(n)], dtype=np.str_),
                         ts=self._ts[:n], src=self._src[:n], live=self._live[:n], has=self._has[:n],
                         src_names=np.array(names, dtype=np.str_), dtype=np.array(self.dtype), dim=np.array(self._store.dim),
                         fingerprint=np.array(self._fingerprint()))
            return True
        except Exception as e:
            print(f"Sidecar save error: {e}")
            return False

    def _load_sidecar(self) -> bool:
        p_rows, p_norms, p_meta = self._sidecar_paths()
        if not (p_rows.exists() and p_norms.exists() and p_meta.exists()):
            return False
        try:
            meta = np.load(p_meta, allow_pickle=False)
            if str(meta["dtype"]) != self.dtype or not np.array_equal(meta["fingerprint"], np.array(self._fingerprint())):
                return False  # another dtype was asked for, or the table changed behind the sidecar: re-read the table
            rows = np.load(p_rows, mmap_mode="r")
            norms = np.load(p_norms)
            n, dim = rows.shape
            if n != len(meta["ids"]) or n != len(norms) or dim != int(meta["dim"]):
                return False
            with self._lock:
                self._store = self._new_store(dim, max(self._capacity0, n))
                for r0 in range(0, n, 262144):  # slabs: the memory-mapped file never sits in RAM as a whole
                    r1 = min(n, r0 + 262144)
                    self._store.put_rows(np.ascontiguousarray(rows[r0:r1]), norms[r0:r1])
                self._grow_meta(n)
                self._ts[:n], self._src[:n] = meta["ts"], meta["src"]
                self._live[:n], self._has[:n] = meta["live"], meta["has"]
                self._src_code = {str(name): i for i, name in enumerate(meta["src_names"].tolist())}
                ids = meta["ids"].tolist()
                self.document_mapping = dict(enumerate(ids))
                self._row_of = {d: r for r, d in enumerate(ids) if self._live[r]}
                self._n = n
                self._meta_version += 1
                self._push_meta(0, n, [])
            self.restored_from = "sidecar"
            return True
        except Exception as e:
            print(f"Sidecar load error (rebuilding from the table): {e}")
            with self._lock:
                if self._store is not None:
                    self._store.close()
                self._store, self._n = None, 0
                self.document_mapping, self._row_of, self._src_code = {}, {}, {}
            return False

    def _new_store(self, dim: int, capacity: int):
        if len(self.devices) > 1:
            return ShardedDeviceStore(dim, self.dtype, capacity, self.devices)
        return DeviceStore(dim, self.dtype, capacity, self.device)

    # -- host-side row table ------------------------------------------------------------------ #
    def _grow_meta(self, need: int):
        if need <= len(self._ts):
            return
        cap = max(need, 2 * len(self._ts), 1024)
        for name in ("_ts", "_src", "_live", "_has"):
            old = getattr(self, name)
            new = np.zeros(cap, old.dtype)
            new[: len(old)] = old
            setattr(self, name, new)

    def _append_rows(self, items: Sequence[tuple]) -> None:
        """items: (doc_id, embedding-or-None, timestamp, source).  One device row per item; NULL and
        wrong-length embeddings become zero rows (they score 0.0, :201-202 / :214-215)."""
        with self._lock:
            self._writers_waiting += 1
            try:
                while self._readers > 0:       # a search in flight reads the mask / metadata this append rewrites
                    self._cv.wait()
            finally:
                self._writers_waiting -= 1
            dim = None if self._store is None else self._store.dim
            if dim is None:
                for _, emb, _, _ in items:
                    if emb is not None and len(emb) > 0:
                        dim = len(emb)
                        break
                if dim is None:
                    dim = 1  # only NULL embeddings so far; re-created when the first real vector arrives
                self._store = self._new_store(dim, max(self._capacity0, len(items)))
                self._dim_is_placeholder = all(e is None or len(e) == 0 for _, e, _, _ in items)
            elif getattr(self, "_dim_is_placeholder", False):
                real = next((len(e) for _, e, _, _ in items if e is not None and len(e) > 0), None)
                if real is not None:  # every earlier row is a NULL row: rebuild them as zero rows
                    n_old = self._n
                    self._store.close()
                    self._store = self._new_store(real, max(self._capacity0, n_old + len(items)))
                    if n_old:
                        self._store.add(np.zeros((n_old, real), np.float32))
                        self._push_meta(0, n_old, [])
                    self._dim_is_placeholder = False
                    dim = real
            X = np.zeros((len(items), dim), np.float32)
            has = np.zeros(len(items), bool)
            for i, (_, emb, _, _) in enumerate(items):
                if emb is not None and len(emb) == dim:
                    X[i] = np.asarray(emb, dtype=np.float32)
                    has[i] = True
                elif emb is not None and len(emb) > 0:
                    has[i] = True  # wrong length: a real row that scores 0.0 against every query
            first = self._store.add(X)
            assert first == self._n
            self._grow_meta(self._n + len(items))
            replaced: List[int] = []
            for i, (doc_id, _, ts, source) in enumerate(items):
                r = first + i
                old = self._row_of.get(doc_id)
                if old is not None:
                    self._live[old] = False  # INSERT OR REPLACE (:351)
                    replaced.append(old)
                self._row_of[doc_id] = r
                self.document_mapping[r] = doc_id
                self._ts[r] = float(ts) if ts is not None else 0.0
                self._src[r] = self._src_code.setdefault(source, len(self._src_code))
                self._live[r] = True
                self._has[r] = has[i]
            self._n += len(items)
            self._meta_version += 1
            self._push_meta(first, len(items), replaced)
            self._cv.notify_all()

    def _push_meta(self, first: int, m: int, replaced: Sequence[int]) -> None:
        """Mirror live / has-embedding / source / timestamp of the new rows (and the rows they replaced) onto the
        device, where the candidate-set policy is evaluated (csrc/filter.cuh, csrc/rank.cuh): nothing about the
        candidate set is computed on the host."""
        if len(self._src_code) > 65536:
            raise ValueError("more than 65536 distinct sources: the device filter carries 16-bit source codes")
        if m == 0:
            return
        flags = self._live[first:first + m].astype(np.uint8) | (self._has[first:first + m].astype(np.uint8) << 1)
        self._store.set_meta(first, flags, self._src[first:first + m].astype(np.uint16))
        self._store.set_timestamps(first, self._ts[first:first + m])
        for r in replaced:
            if r < first:
                self._store.set_meta(r, np.array([int(self._has[r]) << 1], np.uint8))

    # -- ingest (:307-321, :346-370, :372-386) ------------------------------------------------ #
    def add_document(self, document) -> bool:
        """Add one document; returns False (and prints) on any failure, like the reference."""
        try:
            return self.add_documents([document])[0]
        except Exception as e:
            print(f"Error adding document: {e}")
            return False

    def add_documents(self, documents: Iterable) -> List[bool]:
        """Batched ingest: one SQLite transaction (``executemany``) and one normalise+pack launch for the
        whole batch instead of a commit and a FAISS append per document (:349-363, :379-383)."""
        docs = list(documents)
        ok = [False] * len(docs)
        rows, items, idx = [], [], []
        many = getattr(self.embedding_generator, "generate_embeddings", None)
        if many is not None:  # a batched embedder (embedder.py): one encoder + pooling pass for the whole batch
            need = [d for d in docs if not d.embedding]
            try:
                for d, e in zip(need, many([d.content for d in need]) if need else []):
                    d.embedding = e
            except Exception as e:
                print(f"Error generating embeddings: {e}")
        for i, d in enumerate(docs):
            try:
                if not d.embedding:  # :311-312
                    d.embedding = self.embedding_generator.generate_embedding(d.content)
                emb = list(d.embedding) if d.embedding is not None and len(d.embedding) else None
                rows.append((d.id, d.content, json.dumps(d.metadata), pickle.dumps(emb) if emb else None,
                             d.timestamp, d.source, d.importance_score, d.access_count))
                items.append((d.id, emb, d.timestamp, d.source))
                idx.append(i)
            except Exception as e:
                print(f"Error adding document: {e}")
        if not rows:
            return ok
        try:
            with self._lock:
                with sqlite3.connect(self.db_path) as conn:
                    conn.executemany("""INSERT OR REPLACE INTO documents
                        (id, content, metadata, embedding, timestamp, source, importance_score, access_count)
                        VALUES (?, ?, ?, ?, ?, ?, ?, ?)""", rows)
                self._append_rows(items)
            for i in idx:
                ok[i] = True
        except Exception as e:
            print(f"SQLite add error: {e}")
        return ok

    # -- retrieve (:388-403, :454-525) -------------------------------------------------------- #
    def search(self, query) -> Any:
        """Exact top-``max_results`` of the stored documents for ``query`` (a ``RAGQuery``)."""
        Doc, Res = self._types
        start_time = time.time()
        try:
            query_embedding = self.embedding_generator.generate_embedding(query.query_text)
            results = self._search_device(query, query_embedding)
        except Exception as e:
            print(f"Device search error: {e}")
            return Res(documents=[], similarities=[])
        results.query_embedding = query_embedding
        results.retrieval_time = time.time() - start_time
        return results

    def _device_candidates(self, source_filters) -> int:
        """The SQL's WHERE / ORDER BY timestamp DESC / LIMIT (:462-477) evaluated by kernels into the store's own mask;
        returns total_documents_searched.  Cached until rows, metadata or the filter change.  Caller holds the lock
        and has waited until no search reads the current mask."""
        srcs = tuple(sorted(set(source_filters))) if source_filters else None
        key = (self._meta_version, srcs, self.window)
        if key != self._filter_key:
            allowed = None
            if srcs is not None:
                allowed = np.zeros(max(len(self._src_code), 1), np.uint8)
                for name in srcs:
                    if name in self._src_code:
                        allowed[self._src_code[name]] = 1
            self._filter_total = self._store.build_filter(allowed, self.window)
            self._filter_key = key
        return self._filter_total

    def _search_rows(self, Q: Optional[np.ndarray], k: int, thr: float, source_filters):
        """Device search of host queries Q [nq, dim] under the reference's policy -> (ids [nq][<=k], scores, total).
        The lock is held only to pick (or rebuild) the device mask; the search itself runs outside it, so reader
        threads overlap (per-call stream + workspace inside the C ABI).  Q = None: only the candidate count."""
        srcs = tuple(sorted(set(source_filters))) if source_filters else None
        with self._lock:
            n, store = self._n, self._store
            if store is None or n == 0:
                return None, None, 0
            key = (self._meta_version, srcs, self.window)
            while self._writers_waiting > 0 or (self._filter_key != key and self._readers > 0):
                self._cv.wait()
            total = self._device_candidates(source_filters)
            if Q is None or total == 0 or k <= 0:
                return None, None, total
            self._readers += 1
        try:
            ids, sc = store.search(Q, min(k, 1024), "cosine", threshold=thr, clamp01=True, row_end=n, valid="filter")
        finally:
            with self._lock:
                self._readers -= 1
                self._cv.notify_all()
        return ids, sc, total

    def _search_device(self, query, query_embedding):
        Doc, Res = self._types
        k = int(query.max_results)
        if not query_embedding:  # :493 — nothing is scored without a query vector
            total = self._search_rows(None, k, 0.0, query.source_filters)[2]
            return Res(documents=[], similarities=[], total_documents_searched=total)
        q = np.asarray(query_embedding, dtype=np.float32)[None, :]
        ids, sc, total = self._search_rows(q, k, float(query.similarity_threshold), query.source_filters)
        if ids is None:
            return Res(documents=[], similarities=[], total_documents_searched=total)
        got = ids[0] >= 0
        ids, sc = ids[0][got], sc[0][got]
        docs, sims = self._fetch_documents([self.document_mapping[int(r)] for r in ids], sc, Doc)
        return Res(documents=docs, similarities=sims, total_documents_searched=total)

    def search_many(self, queries: Sequence, embeddings: Optional[Sequence] = None) -> List[Any]:
        """Batched ``search``: the queries of one call that share (max_results, similarity_threshold,
        source_filters) go to the device as ONE batch (the tcgen05 kernel from 4 queries up), instead of the
        caller looping over ``search`` (examples/llm_integration.py:54-59, rag_query_interface.py:287-311).
        Returns one ``RAGResult`` per query, in order; same error convention as ``search``."""
        Doc, Res = self._types
        t0 = time.time()
        out: List[Any] = [None] * len(queries)
        groups: Dict[tuple, List[int]] = {}
        embs: List[Optional[List[float]]] = []
        many = getattr(self.embedding_generator, "generate_embeddings", None)
        if embeddings is None and many is not None and len(queries):
            try:
                embeddings = many([qy.query_text for qy in queries])
            except Exception as ex:
                print(f"Device search error: {ex}")
                embeddings = None
        for i, qy in enumerate(queries):
            try:
                e = embeddings[i] if embeddings is not None else self.embedding_generator.generate_embedding(qy.query_text)
            except Exception as ex:
                print(f"Device search error: {ex}")
                e = None
            embs.append(e)
            dim = None if self._store is None else self._store.dim
            if not e or dim is None or len(e) != dim:
                # no vector / wrong length: the single-query path handles it (0.0 scores or nothing, :201-202, :493)
                try:
                    out[i] = self._search_device(qy, e) if e is not None else Res(documents=[], similarities=[])
                except Exception as ex:
                    print(f"Device search error: {ex}")
                    out[i] = Res(documents=[], similarities=[])
                continue
            key = (int(qy.max_results), float(qy.similarity_threshold), tuple(sorted(qy.source_filters)) if qy.source_filters else None)
            groups.setdefault(key, []).append(i)
        for (k, thr, srcs), idx in groups.items():
            try:
                Q = np.asarray([embs[i] for i in idx], dtype=np.float32)
                ids, sc, total = self._search_rows(Q, k, thr, list(srcs) if srcs else None)
                for j, i in enumerate(idx):
                    if ids is None:
                        out[i] = Res(documents=[], similarities=[], total_documents_searched=total)
                        continue
                    got = ids[j] >= 0
                    docs, sims = self._fetch_documents([self.document_mapping[int(r)] for r in ids[j][got]], sc[j][got], Doc)
                    out[i] = Res(documents=docs, similarities=sims, total_documents_searched=total)
            except Exception as ex:
                print(f"Device search error: {ex}")
                for i in idx:
                    out[i] = Res(documents=[], similarities=[])
        dt = time.time() - t0
        for i, r in enumerate(out):
            r.query_embedding = embs[i]
            r.retrieval_time = dt
        return out

    def _fetch_documents(self, doc_ids: List[str], scores, Doc):
        """(documents, similarities) of the winners, index-aligned as the reference returns them zipped (:511-515):
        an id that is no longer in the table (deleted behind the store) drops out of BOTH lists."""
        if not doc_ids:
            return [], []
        out: Dict[str, Any] = {}
        with sqlite3.connect(self.db_path) as conn:
            for i in range(0, len(doc_ids), 500):
                chunk = doc_ids[i:i + 500]
                cur = conn.execute("SELECT * FROM documents WHERE id IN (%s)" % ",".join("?" * len(chunk)), chunk)
                for doc_id, content, meta, blob, ts, source, imp, acc in cur.fetchall():
                    out[doc_id] = Doc(id=doc_id, content=content, metadata=json.loads(meta) if meta else {},
                                      embedding=pickle.loads(blob) if blob else None, timestamp=ts, source=source,
                                      importance_score=imp, access_count=(acc or 0) + 1)  # :495-505
        keep = [j for j, d in enumerate(doc_ids) if d in out]
        return [out[doc_ids[j]] for j in keep], [float(scores[j]) for j in keep]

    # -- stats / life cycle (:527-554) ------------------------------------------------------------ #
    def get_stats(self) -> Dict[str, Any]:
        stats = {'backend': 'b200+sqlite', 'total_documents': 0, 'sources': {},
                 'embedding_cache_size': len(getattr(self.embedding_generator, 'embedding_cache', {})),
                 'device_rows': self._n, 'dtype': self.dtype, 'gpus': len(self.devices), 'device': self.device,
                 'devices': list(self.devices),
                 'dim': None if self._store is None else self._store.dim}
        try:
            with sqlite3.connect(self.db_path) as conn:
                stats['total_documents'] = conn.execute("SELECT COUNT(*) FROM documents").fetchone()[0]
                stats['sources'] = dict(conn.execute("SELECT source, COUNT(*) FROM documents GROUP BY source").fetchall())
        except Exception as e:
            print(f"Stats error: {e}")
        return stats

    def close(self, save: bool = True) -> None:
        """Release the device store; by default the packed rows are written to the sidecar first."""
        with self._lock:
            if self._store is not None:
                if save:
                    self.save_sidecar()
                self._store.close()
                self._store = None


def install(**store_kwargs) -> None:
    """Swap the device store into the reference in place (what INTEGRATION.md describes): after
    ``install()`` ``ollama_rag_system.RAGSystem`` builds a ``B200VectorStore`` and gains the
    ``add_documents`` entry point four callers already assume (SURVEY.md §3.5 D2:
    enhanced_rag_system.py:337,1194; embedding_pipeline.py:806; run_integrated_rag.py:284)."""
    import ollama_rag_system as ref  # the unmodified reference module must be importable

    replaced = getattr(ref.RAGVectorStore, "_b200_replaces", ref.RAGVectorStore)

    class _Installed(B200VectorStore):
        _b200_replaces = replaced

        def __init__(self, storage_path, use_chromadb: bool = True):
            super().__init__(storage_path, use_chromadb, result_types=(ref.RAGDocument, ref.RAGResult), **store_kwargs)

    def add_documents(self, documents) -> List[str]:
        """Batched ingest entry point; returns the ids of the documents that were stored."""
        docs = list(documents)
        for d in docs:
            if not getattr(d, "id", None):
                d.id = d.generate_id()
        ok = self.vector_store.add_documents(docs)
        return [d.id for d, good in zip(docs, ok) if good]

    def query_many(self, query_texts, context=None, max_results=None, source_filters=None) -> List[Any]:
        """Batched ``RAGSystem.query`` (:683-721): same defaults, one device batch for all texts."""
        cfg = self.config.get('retrieval', {}) if hasattr(self, 'config') else {}
        qs = [ref.RAGQuery(query_text=t, context=context or {}, max_results=max_results or cfg.get('max_results', 5),
                           similarity_threshold=cfg.get('similarity_threshold', 0.7), source_filters=source_filters)
              for t in query_texts]
        return self.vector_store.search_many(qs)

    ref.RAGVectorStore = _Installed
    ref.RAGSystem.add_documents = add_documents
    ref.RAGSystem.query_many = query_many
    _install_pipeline_glue(ref)


def _install_pipeline_glue(ref) -> None:
    """``EmbeddingPipeline.create_rag_documents`` (embedding_pipeline.py:720-742) builds ``RAGDocument(...)`` without
    the required ``id`` and raises TypeError as shipped, so the three pipeline call sites of
    ``add_documents`` (enhanced_rag_system.py:337,1194; embedding_pipeline.py:806) are unreachable.  The adaptor keeps
    the reference's method first in line and only on that TypeError builds the same documents with ``id=""`` (which
    ``RAGDocument.__post_init__`` turns into ``generate_id()``, ollama_rag_system.py:61-62).  The reference file stays
    untouched; a reference that fixes the call upstream runs its own code."""
    try:
        import embedding_pipeline as pipe
    except Exception:
        return
    orig = pipe.EmbeddingPipeline.create_rag_documents
    if getattr(orig, "_b200_glue", False):
        return

    def create_rag_documents(self, embedding_results):
        try:
            return orig(self, embedding_results)
        except TypeError:
            return [ref.RAGDocument(id="", content=r.request.content, source=r.request.content_type,
                                    timestamp=r.request.timestamp, embedding=r.embedding,
                                    metadata={**r.request.metadata, 'embedding_model': r.model_used,
                                              'processing_time': r.processing_time, 'content_hash': r.content_hash,
                                              'chunk_info': r.chunk_info})
                    for r in embedding_results]

    create_rag_documents._b200_glue = True
    create_rag_documents._b200_orig = orig
    pipe.EmbeddingPipeline.create_rag_documents = create_rag_documents


def uninstall() -> None:
    """Undo ``install()`` (tests): the reference module gets its own classes and methods back."""
    import sys
    ref = sys.modules.get("ollama_rag_system")
    if ref is not None:
        orig = getattr(ref.RAGVectorStore, "_b200_replaces", None)
        if orig is not None:
            ref.RAGVectorStore = orig
        for name in ("add_documents", "query_many"):
            if name in ref.RAGSystem.__dict__:
                delattr(ref.RAGSystem, name)
    pipe = sys.modules.get("embedding_pipeline")
    if pipe is not None:
        cur = pipe.EmbeddingPipeline.create_rag_documents
        if getattr(cur, "_b200_glue", False):
            pipe.EmbeddingPipeline.create_rag_documents = cur._b200_orig

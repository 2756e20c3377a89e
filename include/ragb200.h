/* ragb200.h — C ABI of the B200-native embedding store + exact top-k search.
 *
 * Drop-in boundary for the retrieval hot path of modular-rag-system.  The
 * reference has no FFI of its own (it is pure Python); each entry point below
 * names the reference method whose arithmetic it replaces (file:line under the
 * reference tree).  The ctypes binding a maintainer would add is shown in
 * INTEGRATION.md and implemented in modular-rag-system_b200/_lib.py.
 *
 * Conventions
 *   - every function returns 0 on success, a negative RAGB_E* code on failure;
 *     ragb_last_error() gives a thread-local message.  Nothing throws.
 *   - "host" entry points take plain host pointers and copy (query buffers that are page-locked — cudaHostAlloc /
 *     cudaHostRegister, e.g. a torch pinned tensor — are read by the copy engine in place, pageable ones are staged
 *     through pinned memory in pieces; either way the call returns only when the buffers may be reused); "_dev" entry
 *     points take device pointers on the store's GPU plus a cudaStream_t passed
 *     as void* (NULL = the store's own stream, synchronised before return; to run on the legacy default stream
 *     asynchronously pass cudaStreamLegacy, i.e. (void*)1).
 *   - result slots: k per query, best first; unused slots hold id -1 and
 *     score -inf.  Order: score descending, then larger row id first (the
 *     reference's stable sort over "ORDER BY timestamp DESC" rows,
 *     ollama_rag_system.py:474,512).
 *   - there is no CPU fallback: without a CUDA device every compute entry point
 *     fails with RAGB_ECUDA.
 *   - threads: every entry point may be called from any thread.  Searches of different threads run side by side
 *     (each leases its own stream + workspace from the store; the reference's GUI spawns a thread per query,
 *     rag_query_interface.py:740-744, against one ingest thread, ollama_rag_system.py:652); appends are serialised
 *     among themselves and run beside searches; calls that rewrite what a search reads (set, set_meta,
 *     set_timestamps, build_filter, reserve, clear) wait for the searches in flight.
 *   - streams: work launched on a caller's stream by a "_dev" entry point is ordered for the caller: rows appended
 *     with ragb_store_add*_dev are searchable by every later call on any stream or thread (the store records an
 *     event behind the ingest kernel and later searches wait for it), and a later search that reuses the
 *     workspace of an earlier caller-stream search waits for that search first.
 */
#ifndef RAGB200_H
#define RAGB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RAGB_ABI_VERSION 2

/* element types */
#define RAGB_F16 0
#define RAGB_BF16 1
#define RAGB_F32 2
/* mask element types for ragb_store_add_pooled* */
#define RAGB_MASK_U8 0
#define RAGB_MASK_I32 1
#define RAGB_MASK_I64 2
#define RAGB_MASK_F32 3
/* metrics */
#define RAGB_IP 0      /* inner product over the stored (unit) rows            */
#define RAGB_COSINE 1  /* ollama_rag_system.py:199-222: q.x/(|q||x|), 0 if a norm is 0 */
/* search flags */
#define RAGB_CLAMP01 1      /* clamp scores to [0,1] before ranking (:218)        */
#define RAGB_FORCE_SCAN 2   /* use the bandwidth-bound scan kernel per query      */
#define RAGB_FORCE_GEMM 4   /* use the tcgen05 batched kernel even for nq == 1    */
#define RAGB_NO_RESCORE 8   /* batched path: skip the fp32 re-score of the winners */

/* error codes */
#define RAGB_OK 0
#define RAGB_EINVAL -1
#define RAGB_ECUDA -2
#define RAGB_ENOMEM -3
#define RAGB_EFULL -4
#define RAGB_EUNSUPPORTED -5

typedef struct ragb_store ragb_store;

/* ---- library ----------------------------------------------------------- */
int ragb_abi_version(void);
const char* ragb_last_error(void);
/* number of CUDA kernels this library has launched in this process */
int64_t ragb_launch_count(void);
int ragb_device_count(void);

/* ---- store life cycle --------------------------------------------------
 * Replaces the in-memory index slot of RAGVectorStore (faiss.IndexFlatIP,
 * ollama_rag_system.py:298-305): a device-resident row-major matrix of unit
 * rows (fp16 / bf16 / fp32) + the fp32 norm each row had before normalising. */
int ragb_store_create(int device, int dim, int dtype, int64_t capacity, ragb_store** out);
int ragb_store_destroy(ragb_store* s);
int ragb_store_reserve(ragb_store* s, int64_t capacity); /* grow, keeps rows */
int ragb_store_clear(ragb_store* s);
int64_t ragb_store_size(const ragb_store* s);
int64_t ragb_store_capacity(const ragb_store* s);
int ragb_store_dim(const ragb_store* s);
int ragb_store_ld(const ragb_store* s); /* row stride in elements */
int ragb_store_dtype(const ragb_store* s);
int ragb_store_device(const ragb_store* s);
void* ragb_store_rows_dev(const ragb_store* s);  /* device pointer to the packed rows */
float* ragb_store_norms_dev(const ragb_store* s);

/* ---- ingest --------------------------------------------------------------
 * ragb_store_add*: RAGVectorStore._update_faiss_index, ollama_rag_system.py:372-386
 *   (fp32 normalise :379-380, append :382, row number returned for
 *   document_mapping :383).  A zero row packs to zeros with norm 0 (search
 *   semantics :214-215) instead of the reference's NaN row.
 * ragb_store_add_pooled*: the encoder tail inside SentenceTransformer.encode
 *   (call sites embedding_pipeline.py:424, ollama_rag_system.py:163): masked
 *   mean over tokens, L2 normalise, pack. */
int ragb_store_add(ragb_store* s, const float* rows, int64_t n, int64_t* out_first_row);
int ragb_store_add_dev(ragb_store* s, const void* rows, int in_dtype, int64_t n, int64_t ld_in,
                       int64_t* out_first_row, void* stream);
int ragb_store_set(ragb_store* s, int64_t row, const float* vec); /* INSERT OR REPLACE, :351 */
int ragb_store_add_pooled(ragb_store* s, const void* tok, int tok_dtype, const void* mask,
                          int mask_dtype, int64_t batch, int seq, int hidden,
                          int64_t* out_first_row);
int ragb_store_add_pooled_dev(ragb_store* s, const void* tok, int tok_dtype, const void* mask,
                              int mask_dtype, int64_t batch, int seq, int hidden,
                              int64_t* out_first_row, void* stream);
/* stateless kernels (device pointers): pack into caller-owned rows */
int ragb_norm_pack_dev(int device, const void* in, int in_dtype, int64_t n, int dim, int64_t ld_in,
                       void* out, int out_dtype, int64_t ld_out, float* out_norms, void* stream);
int ragb_pool_norm_pack_dev(int device, const void* tok, int tok_dtype, const void* mask,
                            int mask_dtype, int64_t batch, int seq, int hidden, void* out,
                            int out_dtype, int64_t ld_out, float* out_norms, void* stream);
/* copy rows [row0,row0+n) back to the host (packed elements, ld = dim) and their norms */
int ragb_store_get_rows(ragb_store* s, int64_t row0, int64_t n, void* out_rows, float* out_norms);
/* append rows that are already packed (store dtype, ld = dim — what ragb_store_get_rows returns) with their norms,
 * verbatim: the restore half of the packed on-disk sidecar that replaces unpickling every BLOB of the
 * reference's table (ollama_rag_system.py:358, :487) on open */
int ragb_store_put_rows(ragb_store* s, const void* packed_rows, const float* norms, int64_t n,
                        int64_t* out_first_row);

/* ---- search ----------------------------------------------------------------
 * Replaces the scan / threshold / sort / slice of RAGVectorStore._search_sqlite
 * (ollama_rag_system.py:480-515) and EmbeddingGenerator.calculate_similarity
 * (:199-222) over rows [row_begin,row_end) (row_end < 0 = all rows).
 *   valid  : optional per-row uint8 (host for ragb_search, device for _dev),
 *            0 = skip the row (NULL embedding :485-492, source filter :466-469,
 *            window :474, replaced ids :351).
 *   keep a row when score >= threshold (:495); pass -INFINITY for none. */
int ragb_search(ragb_store* s, const float* q, int64_t nq, int k, int metric, float threshold,
                int flags, int64_t row_begin, int64_t row_end, const uint8_t* valid,
                float* out_scores, int64_t* out_ids);
int ragb_search_dev(ragb_store* s, const float* q, int64_t nq, int k, int metric, float threshold,
                    int flags, int64_t row_begin, int64_t row_end, const uint8_t* valid,
                    float* out_scores, int64_t* out_ids, void* stream);

/* ---- search that allocates nothing ---------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * ragb_scan_topk_static_dev is the scan search of ragb_search_dev (RAGB_FORCE_SCAN: every row read once, exact fp32
 * scores, same order among equal scores) over rows the CALLER holds on `device` — packed unit rows as written by
 * ragb_norm_pack_dev / ragb_store_rows_dev: `n` rows of `dim` elements of `dtype`, row stride `ld` elements
 * (ld * element size a multiple of 16 bytes, base 16-byte aligned, the padding elements finite) — with the kernel's
 * scratch (one list of ragb_kpad(k) candidate keys per CTA and query, one ticket word per query) in a file-scope
 * __device__ array of the library instead of a cudaMalloc'ed workspace: 2 425 088 bytes that every GPU which loads
 * the library holds until the process exits.  It calls no cudaMalloc / cudaFree / allocating library routine; all
 * buffers (q [nq][dim] fp32, valid [n] or NULL, out_scores [nq][k], out_ids [nq][k] = row numbers) are device
 * memory of the caller; the work is queued on `stream` (a cudaStream_t; NULL = the legacy default stream) and
 * the call returns without synchronising.
 *   limits  nq <= RAGB_STATIC_SCAN_MAX_NQ and nq * ragb_kpad(k) <= RAGB_STATIC_SCAN_MAX_KEYS per call (e.g. 64
 *           queries at k <= 16, 8 at k <= 128, 1 at k <= 1024); beyond that: RAGB_EUNSUPPORTED with a message that
 *           names the limit, nothing launched.  n < 2^32.
 *   set-up  the first call on a device resolves the array's device address (cudaGetSymbolAddress) — make it before a
 *           capture starts; a call with nq = 0 does only that.
 *   sharing every call on a device uses the SAME array: two calls that overlap on the GPU (different streams, or two
 *           graphs in flight) race on it and return garbage — order them on one stream or with events.  The array
 *           keeps the last call's contents; the words that must start at zero are cleared on `stream` by every call. */
#define RAGB_STATIC_SCAN_MAX_NQ 64
#define RAGB_STATIC_SCAN_MAX_KEYS 1024
int ragb_scan_topk_static_dev(int device, const void* rows, int dtype, int64_t n, int dim, int64_t ld,
                              const float* q, int64_t nq, int k, int metric, float threshold, int flags,
                              const uint8_t* valid, float* out_scores, int64_t* out_ids, void* stream);

/* ---- candidate-set policy on the device ---------------------------------------
 * The SQL of _search_sqlite (ollama_rag_system.py:462-477: WHERE source IN (...) ORDER BY timestamp DESC
 * LIMIT 1000), the NULL-embedding skip (:485-492) and INSERT OR REPLACE (:351), evaluated by kernels over
 * two per-row arrays kept next to the rows: flags (bit 0 = live, bit 1 = has an embedding; default 3) and a
 * 16-bit source code (default 0).  ragb_store_build_filter leaves a per-row mask in the store — match =
 * live and (n_codes == 0 or src_allowed[code]); only the LAST `window` matches in row order stay
 * (window < 0: all; exact for the reference when timestamps do not decrease in row order); rows without an
 * embedding are never valid — and returns the number of rows the SQL would have returned
 * (total_documents_searched, :520).  Pass RAGB_STORE_FILTER as `valid` to any search entry point to use it.
 * Adding rows or changing metadata invalidates the mask. */
#define RAGB_STORE_FILTER ((const uint8_t*)(uintptr_t)1)
int ragb_store_set_meta(ragb_store* s, int64_t row0, int64_t n, const uint8_t* flags,
                        const uint16_t* src); /* host arrays; NULL = leave unchanged */
int ragb_store_build_filter(ragb_store* s, const uint8_t* src_allowed, int n_codes, int64_t window,
                            int64_t* out_total);
/* Per-row timestamps (fp64, the reference's RAGDocument.timestamp, ollama_rag_system.py:53) for "ORDER BY timestamp DESC"
 * (:474) and for the order among equal scores (:512: the stable sort keeps the SQL order, newer first).  While
 * timestamps do not decrease in row order nothing changes (newer == larger row).  When they do, the store sorts
 * (timestamp, row) on the device (csrc/rank.cuh) and from then on every candidate key carries the row's recency rank
 * instead of its row number: the window of build_filter is the newest rows BY TIMESTAMP and equal scores come back
 * newest timestamp first, later row first among equal timestamps — no host mask, no host re-sort.  Rows that were
 * never described count as the newest so far, in row order. */
int ragb_store_set_timestamps(ragb_store* s, int64_t row0, int64_t n, const double* ts);

/* ---- measurement ------------------------------------------------------------
 * While enabled, every search call brackets its dominant kernel (the scan for the
 * bandwidth-bound path, the tcgen05 kernel for the batched path) with CUDA events on the
 * launching stream.  ragb_store_profile_read waits for them, returns the summed kernel
 * time and the number of bracketed launches, and resets the counters.  The reference has
 * only wall-clock retrieval_time (ollama_rag_system.py:390,401). */
int ragb_store_profile(ragb_store* s, int enable);
int ragb_store_profile_read(ragb_store* s, double* total_ms, int64_t* n_launches);

/* ---- multi-GPU building blocks (row-sharded corpus, one process per GPU) ----
 * local top-k as sortable 64-bit keys: (ordered score bits << 32) | (row + id_base),
 * kpad = ragb_kpad(k) keys per query, sorted best first, 0 = empty slot.
 * The keys of all ranks are all-gathered (NCCL) and merged with ragb_merge_keys_dev. */
int ragb_kpad(int k);
int ragb_search_keys_dev(ragb_store* s, const float* q, int64_t nq, int k, int metric,
                         float threshold, int flags, int64_t row_begin, int64_t row_end,
                         const uint8_t* valid, uint32_t id_base, uint64_t* out_keys, void* stream);
/* keys: n_lists lists per query; list (q,p) starts at keys + q*stride_q + p*stride_p */
int ragb_merge_keys_dev(int device, const uint64_t* keys, int64_t nq, int n_lists, int kpad,
                        int64_t stride_q, int64_t stride_p, int k, float* out_scores,
                        int64_t* out_ids, void* stream);


/* ---- candidate exchange over peer memory (NVLink P2P), no collective launch ----------------------
 * Every rank owns a gather buffer uint64 keys[world][nq][kpad] and `world` uint64 arrival words, both
 * mapped into every other rank's address space (CUDA IPC / torch symmetric memory).
 * ragb_search_scatter_dev = ragb_search_keys_dev whose LAST kernel stores this rank's sorted keys
 * straight into slot [rank] of every rank's buffer and then publishes `epoch` to word [rank] of every
 * rank (release, system scope).  ragb_gather_merge_dev waits (acquire, system scope; ~2 s timeout ->
 * *out_status = 1) until all `world` words of this rank have reached `epoch`, then merges its buffer.
 * Alternate between two buffers from call to call (keys[] of epoch e+1 must not alias those of e);
 * epochs are strictly increasing.  The reference is single-device: there is nothing this replaces. */
#define RAGB_MAX_PEERS 8
typedef struct ragb_exchange {
  int world, rank;
  uint64_t* keys[RAGB_MAX_PEERS];  /* keys[r]  = rank r's gather buffer of this epoch */
  uint64_t* flags[RAGB_MAX_PEERS]; /* flags[r] = rank r's arrival words [world] */
  uint64_t epoch;
} ragb_exchange;
int ragb_search_scatter_dev(ragb_store* s, const float* q, int64_t nq, int k, int metric,
                            float threshold, int flags, int64_t row_begin, int64_t row_end,
                            const uint8_t* valid, uint32_t id_base, const ragb_exchange* x,
                            void* stream);
int ragb_gather_merge_dev(int device, const ragb_exchange* x, int64_t nq, int k, float* out_scores,
                          int64_t* out_ids, int* out_status, void* stream);
/* The two calls above as ONE host entry point (what ragb_search is for a single store): queries from host memory
 * (identical on every rank), upload, local search + scatter into every rank's gather buffer, wait + merge, results
 * back into host memory — one stream, one synchronisation, no caller-side copies.  *out_status = 1 if a rank's
 * candidates did not arrive within the kernel's timeout. */
int ragb_search_exchange(ragb_store* s, const float* q, int64_t nq, int k, int metric, float threshold, int flags,
                         uint32_t id_base, const ragb_exchange* x, float* out_scores, int64_t* out_ids,
                         int* out_status);

/* ---- row-sharded store: ONE process, the rows dealt over several GPUs of one NVSwitch box ------------------------
 * north_star: "the embedding store ... row-sharded across GPUs", behind the same entry points as the one-device store
 * (RAGVectorStore.add_document / search, ollama_rag_system.py:372-386 / :388-403).
 *   layout  rows are dealt to the devices in blocks of block_rows (a power of two >= 2048; 0 = 16384), round-robin:
 *           global row g lives on device (g / block_rows) % n_dev at local row (g / block_rows / n_dev) * block_rows
 *           + g % block_rows, so shards stay within one block of balanced while the store grows, and global <-> local
 *           is arithmetic (no table).
 *   search  the queries go to every device, each runs its local scan / tcgen05 kernel, and the LAST kernel of every
 *           device stores its k candidates per query — keys whose low word is the GLOBAL row — straight into the gather
 *           buffer on the first device through peer-mapped memory (NVLink P2P stores, no collective, no staging
 *           copy); one merge kernel on the first device finishes.  Results equal the one-device store's, ties included.
 *   policy  set_meta / build_filter as for ragb_store (the newest-`window` cut is taken over the GLOBAL row order);
 *           timestamps must not decrease in row order (out-of-order timestamps need the recency ranks of a
 *           one-device store: RAGB_EUNSUPPORTED).
 * `valid` for the search entry points: NULL, RAGB_STORE_FILTER, or (host entry only) a mask over global rows.
 * A device may be listed more than once (several shards on one GPU; the tests cover the layout that way on one GPU). */
typedef struct ragb_mstore ragb_mstore;
int ragb_mstore_create(int dim, int dtype, int64_t capacity, const int* device_ids, int n_dev, int64_t block_rows,
                       ragb_mstore** out);
int ragb_mstore_destroy(ragb_mstore* m);
int ragb_mstore_reserve(ragb_mstore* m, int64_t capacity);
int ragb_mstore_clear(ragb_mstore* m);
int64_t ragb_mstore_size(const ragb_mstore* m);
int64_t ragb_mstore_capacity(const ragb_mstore* m);
int64_t ragb_mstore_block_rows(const ragb_mstore* m);
int ragb_mstore_n_devices(const ragb_mstore* m);
int ragb_mstore_device(const ragb_mstore* m, int i);          /* CUDA ordinal of shard i */
int64_t ragb_mstore_device_rows(const ragb_mstore* m, int i); /* rows held by shard i */
int ragb_mstore_add(ragb_mstore* m, const float* rows, int64_t n, int64_t* out_first_row);
/* rows on ANY device of the box (peer access): each shard's ingest kernel reads its blocks over NVLink; `stream` (on
 * the rows' device; NULL = none) is synchronised first */
int ragb_mstore_add_dev(ragb_mstore* m, const void* rows, int in_dtype, int64_t n, int64_t ld_in,
                        int64_t* out_first_row, void* stream);
int ragb_mstore_set(ragb_mstore* m, int64_t row, const float* vec);
int ragb_mstore_get_rows(ragb_mstore* m, int64_t row0, int64_t n, void* out_rows, float* out_norms);
int ragb_mstore_put_rows(ragb_mstore* m, const void* packed_rows, const float* norms, int64_t n,
                         int64_t* out_first_row);
int ragb_mstore_set_meta(ragb_mstore* m, int64_t row0, int64_t n, const uint8_t* flags, const uint16_t* src);
int ragb_mstore_set_timestamps(ragb_mstore* m, int64_t row0, int64_t n, const double* ts);
int ragb_mstore_build_filter(ragb_mstore* m, const uint8_t* src_allowed, int n_codes, int64_t window,
                             int64_t* out_total);
int ragb_mstore_search(ragb_mstore* m, const float* q, int64_t nq, int k, int metric, float threshold, int flags,
                       const uint8_t* valid, float* out_scores, int64_t* out_ids);
/* q, out_scores, out_ids on the FIRST device of the store; stream on that device (NULL = synchronous) */
int ragb_mstore_search_dev(ragb_mstore* m, const float* q, int64_t nq, int k, int metric, float threshold,
                           int flags, const uint8_t* valid, float* out_scores, int64_t* out_ids, void* stream);
int ragb_mstore_profile(ragb_mstore* m, int enable);
/* total_ms: the slowest shard's summed dominant-kernel time; n_launches: bracketed launches of the first shard */
int ragb_mstore_profile_read(ragb_mstore* m, double* total_ms, int64_t* n_launches);

#ifdef __cplusplus
}
#endif
#endif /* RAGB200_H */

// store.h — host-side internals shared by ragb200.cu (one-device store) and mstore.cu (row-sharded store).
// Not part of the ABI (include/ragb200.h is).
#pragma once
#include "../../include/ragb200.h"

#include <atomic>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <mutex>
#include <shared_mutex>
#include <string>
#include <vector>

#include <cuda.h>
#include <cuda_runtime.h>

#include "common.cuh"

namespace ragb {

// ---- errors, counters ------------------------------------------------------------------------------------------
int fail(int code, const char* fmt, ...);
void count_launch();
#define CU_TRY(expr)                                                                            \
  do {                                                                                          \
    cudaError_t e_ = (expr);                                                                    \
    if (e_ != cudaSuccess)                                                                      \
      return ::ragb::fail(e_ == cudaErrorMemoryAllocation ? RAGB_ENOMEM : RAGB_ECUDA, "%s: %s (%s:%d)", \
                          #expr, cudaGetErrorString(e_), __FILE__, __LINE__);                   \
  } while (0)
#define RB_TRY(expr)              \
  do {                            \
    int r_ = (expr);              \
    if (r_ != RAGB_OK) return r_; \
  } while (0)

struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
    if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
  }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
#define ON_DEVICE(dev)                                                                     \
  ::ragb::DeviceGuard guard_(dev);                                                         \
  if (!guard_.ok) return ::ragb::fail(RAGB_ECUDA, "cannot select CUDA device %d: %s", dev, \
                                      cudaGetErrorString(cudaGetLastError()))

int grow(void** p, size_t* have, size_t need, bool pinned);
// Host buffer -> device on `st`.  A page-locked source (cudaHostAlloc / cudaHostRegister, e.g. a torch pinned tensor) is
// read by the copy engine in place; a pageable one is staged through `pin` (page-locked, >= bytes) in pieces, so that
// the copy engine moves piece i while the host copies piece i + 1 (3 MB of queries: the two no longer add up).
// The call returns with copies queued; the caller synchronises the stream before the source may change.
int upload_host(void* dev, void* pin, const void* src, size_t bytes, cudaStream_t st);
inline size_t elem_size(int dtype) { return dtype == RAGB_F32 ? 4 : 2; }

// launch through a `const void*` entry point (kernels.h); argument types must be the kernel's own
template <typename... Args>
inline cudaError_t launch_k(const void* k, dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  void* a[] = {(void*)&args...};
  cudaError_t e = cudaLaunchKernel(k, grid, block, a, smem, st);
  count_launch();
  return e;
}

constexpr int64_t kScanMaxQ = 65535;  // grid y limit of one scan launch
constexpr int kMaxCtx = 8;            // searches in flight per store

// ---- one search in flight ----------------------------------------------------------------------------------------
// Everything a search writes besides its outputs lives here, so that searches of different threads (the reference's
// GUI spawns a thread per query, rag_query_interface.py:740-744) run side by side on their own streams instead of
// queueing on one workspace.
struct SearchCtx {
  cudaStream_t stream = nullptr;
  void* ws = nullptr;          // per-split key lists | candidate buffers | shared thresholds
  size_t ws_bytes = 0;
  unsigned* tickets = nullptr; // [kScanMaxQ + 1] zero between launches (scan.cuh last-CTA merge; +1: peer signal counter)
  void* qh = nullptr;          // converted queries (16-bit)
  size_t qh_bytes = 0;
  void* sync_ctr = nullptr;    // pair kernel: per-split meeting counters
  size_t sync_ctr_bytes = 0;
  void* pin = nullptr;         // host entry point: pinned q | scores | ids, and its device mirror
  size_t pin_bytes = 0;
  void* io = nullptr;
  size_t io_bytes = 0;
  cudaEvent_t done_ev = nullptr;  // last search that used these buffers on a CALLER's stream
  bool done_used = false;
  cudaEvent_t t0 = nullptr, t1 = nullptr;  // scratch events (mstore)
};

// tensor maps, driver entry point, occupancy and attribute caches of the batched kernels: per store = per device
struct GemmShared {
  std::mutex mu;
  void* encode_fn = nullptr;
  CUtensorMap map_x, map_x_ts, map_x_pair;
  const void* map_rows = nullptr; int64_t map_capacity = -1;
  const void* ts_rows = nullptr; int64_t ts_capacity = -1; int ts_nt = 0;
  const void* pair_rows = nullptr; int64_t pair_capacity = -1;
  int pair_clusters[24] = {};
  bool pair_attr[24] = {};
  bool solo_attr[3] = {};
  bool ts_attr[2] = {false, false};
  bool ss_attr = false;
  int occ_c = 0, occ_clusters = 0;
  size_t occ_smem = 0;
};

}  // namespace ragb

struct ragb_store {
  int device = 0, dim = 0, dtype = 0, n_sm = 148;
  int64_t ld = 0, capacity = 0;
  std::atomic<int64_t> n{0};
  void* rows = nullptr;
  float* norms = nullptr;
  cudaStream_t stream = nullptr;     // ingest / maintenance stream
  cudaEvent_t ingest_ev = nullptr;   // last ingest launched on a caller's stream: every search waits for it
  std::atomic<bool> ingest_ev_used{false};
  std::mutex add_mu;                 // writers (append, staging buffers)
  // readers hold it shared for the length of their host call; exclusive: anything that moves or rewrites what a
  // search reads (reserve, clear, set, metadata, filter, timestamps)
  std::shared_mutex rows_mu;
  // search contexts
  std::mutex pool_mu;
  std::condition_variable pool_cv;
  std::vector<ragb::SearchCtx*> free_ctx, all_ctx;
  // row-shard layout (mstore.cu): id of a local row (common.cuh IdMap); single store: identity
  int shard_shift = -1, shard_n = 1, shard_dev = 0;
  // candidate-set policy on the device (filter.cuh); guarded by rows_mu
  uint8_t* meta_flags = nullptr;     // [meta_cap] bit0 live, bit1 has embedding (default 3)
  uint16_t* meta_src = nullptr;      // [meta_cap] source code (default 0)
  double* ts = nullptr;              // [meta_cap] timestamps; rows >= ts_n: undescribed (newest, in row order)
  int64_t ts_n = 0;
  double ts_max = -1e300;            // max timestamp described so far
  bool ts_dirty = false;             // timestamps changed since the last order check
  uint32_t* rank = nullptr;          // [rank_cap] recency rank per row (rank.cuh) — only while timestamps are out of order
  uint32_t* row_of_rank = nullptr;
  int64_t rank_n = 0, rank_cap = 0;  // rows the ranks cover (0 = timestamps follow row order: identity)
  bool ranked = false;
  int64_t meta_cap = 0;
  uint8_t* filter = nullptr;         // [meta_cap] valid bytes built by ragb_store_build_filter
  int64_t filter_n = -1;             // rows the filter covers (-1 = none built)
  void* filter_aux = nullptr;        // counts | result | src_allowed
  size_t filter_aux_bytes = 0;
  // ingest staging (guarded by add_mu)
  void* stage_dev = nullptr;
  size_t stage_dev_bytes = 0;
  ragb::GemmShared gemm;
  // optional timing of the dominant search kernel with CUDA events on the launching stream
  std::mutex prof_mu;
  bool profile = false;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_events;
  size_t prof_used = 0;
};

namespace ragb {

// what one search asks for (device pointers)
struct SearchReq {
  const float* q = nullptr;
  int64_t nq = 0;
  int k = 0, metric = 0;
  float threshold = 0.f;
  int flags = 0;
  int64_t rb = 0, re = 0;           // resolved local row window
  const uint8_t* valid = nullptr;   // device mask over local rows, or null
  uint32_t id_base = 0;
  bool global_ids = true;           // false: keys/ids carry LOCAL rows even in a sharded store
  uint64_t* out_keys = nullptr;     // [nq][kpad] sorted keys (ids mapped, not decoded), or null
  float* out_scores = nullptr;      // [nq][k] or null
  int64_t* out_ids = nullptr;
  PeerOut peer = {};
};

struct CtxLease {
  ragb_store* s;
  SearchCtx* c = nullptr;
  explicit CtxLease(ragb_store* s_) : s(s_) {}
  int acquire();
  ~CtxLease();
};

// internals mstore.cu builds on (caller holds rows_mu shared, device selected)
int store_check_args(ragb_store* s, const void* q, int64_t nq, int k, int metric, int64_t* rb, int64_t* re);
int store_resolve_valid(ragb_store* s, const uint8_t** valid, int64_t row_end);
int store_prepare_stream(ragb_store* s, SearchCtx* c, cudaStream_t st, bool caller_stream);
void store_finish_stream(ragb_store* s, SearchCtx* c, cudaStream_t st, bool caller_stream);
int store_ensure_ranks(ragb_store* s, std::shared_lock<std::shared_mutex>& lk, int64_t re);
int search_core(ragb_store* s, SearchCtx* c, const SearchReq& r, cudaStream_t st);
int merge_launch(const uint64_t* keys, int64_t nq, int n_lists, int kpad, int64_t stride_q, int64_t stride_p, int k,
                 float* out_scores, int64_t* out_ids, uint64_t* out_keys, const uint32_t* row_of_id, cudaStream_t st,
                 bool unsorted = false, const PeerOut* po = nullptr, const uint64_t* wait_flags = nullptr, int wait_n = 0,
                 unsigned long long wait_epoch = 0, int* status = nullptr);
// the three passes of the device filter (filter.cuh), split so that mstore.cu can merge the per-CTA counts of several
// devices between pass 1 and pass 3; rows_mu held exclusively, device selected, everything on s->stream
int filter_refresh_ranks(ragb_store* s, int64_t n);
int filter_mark(ragb_store* s, const uint8_t* src_allowed_host, int n_codes, int64_t n, int* n_ctas);
int filter_counts_host(ragb_store* s, int n_ctas, std::vector<unsigned>* out);
int filter_window(ragb_store* s, int64_t n, int n_codes, int n_ctas, int64_t cut_cta, int64_t keep);

}  // namespace ragb

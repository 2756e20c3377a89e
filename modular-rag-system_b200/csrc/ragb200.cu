// ragb200.cu — C ABI (include/ragb200.h) over the sm_100a kernels.
// Host logic only: store life cycle, dispatch, workspaces.  No torch types, no CPU fallback.
#include "../../include/ragb200.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"
#include "gemm_topk.cuh"
#include "gemm_ts.cuh"
#include "gemm_pair.cuh"
#include "filter.cuh"
#include "ingest.cuh"
#include "merge.cuh"
#include "scan.cuh"

using namespace ragb;

// ---------------------------------------------------------------------------
// errors, counters
// ---------------------------------------------------------------------------
static thread_local std::string g_err;
static std::atomic<int64_t> g_launches{0};

static int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}
#define CU_TRY(expr)                                                                       \
  do {                                                                                     \
    cudaError_t e_ = (expr);                                                               \
    if (e_ != cudaSuccess)                                                                 \
      return fail(e_ == cudaErrorMemoryAllocation ? RAGB_ENOMEM : RAGB_ECUDA, "%s: %s (%s:%d)", \
                  #expr, cudaGetErrorString(e_), __FILE__, __LINE__);                      \
  } while (0)
#define RB_TRY(expr)        \
  do {                      \
    int r_ = (expr);        \
    if (r_ != RAGB_OK) return r_; \
  } while (0)
#define LAUNCHED()                   \
  do {                               \
    g_launches.fetch_add(1);         \
    CU_TRY(cudaGetLastError());      \
  } while (0)

static size_t elem_size(int dtype) { return dtype == RAGB_F32 ? 4 : 2; }

// ---------------------------------------------------------------------------
// store
// ---------------------------------------------------------------------------
constexpr int64_t kScanMaxQ = 65535;  // grid y limit of one scan launch

struct ragb_store {
  int device = 0, dim = 0, dtype = 0, n_sm = 148;
  int64_t ld = 0, capacity = 0;
  std::atomic<int64_t> n{0};
  void* rows = nullptr;
  float* norms = nullptr;
  cudaStream_t stream = nullptr;
  cudaEvent_t ingest_ev = nullptr;  // last ingest launched on a caller's stream; the store's own stream waits on it
  cudaEvent_t search_ev = nullptr;  // last search launched on a caller's stream (it uses ws / tickets): same
  bool search_ev_used = false;
  std::mutex add_mu, search_mu;
  // search workspace (guarded by search_mu)
  void* ws = nullptr;
  size_t ws_bytes = 0;
  unsigned* tickets = nullptr;  // [kScanMaxQ + 1] zero between launches: last-CTA final merge of the scan (scan.cuh);
                                // the extra word counts finished CTAs for the peer exchange signal
  PeerOut peer = {};            // set for the duration of ragb_search_scatter_dev (guarded by search_mu)
  // candidate-set policy on the device (filter.cuh); guarded by search_mu
  uint8_t* meta_flags = nullptr;   // [meta_cap] bit0 live, bit1 has embedding (default 3)
  uint16_t* meta_src = nullptr;    // [meta_cap] source code (default 0)
  int64_t meta_cap = 0;
  uint8_t* filter = nullptr;       // [meta_cap] valid bytes built by ragb_store_build_filter
  int64_t filter_n = -1;           // rows the filter covers (-1 = none built)
  void* filter_aux = nullptr;      // counts | result | src_allowed
  size_t filter_aux_bytes = 0;
  // ingest staging (guarded by add_mu)
  void* stage_dev = nullptr;
  size_t stage_dev_bytes = 0;
  void* stage_host = nullptr;  // pinned
  size_t stage_host_bytes = 0;
  // pinned buffers for the host search entry point (guarded by search_mu)
  void* pin = nullptr;
  size_t pin_bytes = 0;
  void* io = nullptr;  // device mirror of pin
  size_t io_bytes = 0;
  GemmState gemm;  // tensor maps etc. of the batched path
  // optional timing of the dominant search kernel (scan or gemm) with CUDA events on the
  // launching stream; guarded by search_mu
  bool profile = false;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_events;
  size_t prof_used = 0;
};

// brackets one kernel launch with events when profiling is on
struct ProfScope {
  ragb_store* s; cudaStream_t st; cudaEvent_t stop = nullptr;
  ProfScope(ragb_store* s_, cudaStream_t st_) : s(s_), st(st_) {
    if (!s->profile) return;
    if (s->prof_used == s->prof_events.size()) {
      cudaEvent_t a, b;
      if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) return;
      s->prof_events.emplace_back(a, b);
    }
    auto& ev = s->prof_events[s->prof_used++];
    cudaEventRecord(ev.first, st);
    stop = ev.second;
  }
  ~ProfScope() { if (stop) cudaEventRecord(stop, st); }
};

struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
    if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
  }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
#define ON_DEVICE(dev)                                                        \
  DeviceGuard guard_(dev);                                                    \
  if (!guard_.ok) return fail(RAGB_ECUDA, "cannot select CUDA device %d: %s", dev, \
                              cudaGetErrorString(cudaGetLastError()))

static int grow(void** p, size_t* have, size_t need, bool pinned) {
  if (*have >= need) return RAGB_OK;
  size_t want = need + need / 4;
  if (*p) {
    if (pinned) CU_TRY(cudaFreeHost(*p)); else CU_TRY(cudaFree(*p));
    *p = nullptr; *have = 0;
  }
  if (pinned) CU_TRY(cudaMallocHost(p, want)); else CU_TRY(cudaMalloc(p, want));
  *have = want;
  return RAGB_OK;
}

extern "C" int ragb_abi_version(void) { return RAGB_ABI_VERSION; }
extern "C" const char* ragb_last_error(void) { return g_err.c_str(); }
extern "C" int64_t ragb_launch_count(void) { return g_launches.load(); }
extern "C" int ragb_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}
extern "C" int ragb_kpad(int k) { return kpad_of(k); }

static int alloc_rows(ragb_store* s, int64_t capacity, void** rows, float** norms) {
  size_t bytes = (size_t)capacity * s->ld * elem_size(s->dtype);
  CU_TRY(cudaMalloc(rows, bytes ? bytes : 16));
  cudaError_t e = cudaMalloc((void**)norms, capacity ? (size_t)capacity * 4 : 16);
  if (e != cudaSuccess) { cudaFree(*rows); *rows = nullptr; return fail(RAGB_ENOMEM, "cudaMalloc norms: %s", cudaGetErrorString(e)); }
  return RAGB_OK;
}

extern "C" int ragb_store_create(int device, int dim, int dtype, int64_t capacity, ragb_store** out) {
  if (!out) return fail(RAGB_EINVAL, "out is NULL");
  *out = nullptr;
  if (dim <= 0 || dim > 65536) return fail(RAGB_EINVAL, "dim %d out of range [1,65536]", dim);
  if (dtype != RAGB_F16 && dtype != RAGB_BF16 && dtype != RAGB_F32) return fail(RAGB_EINVAL, "bad dtype %d", dtype);
  if (capacity < 0 || capacity >= (int64_t)1 << 31) return fail(RAGB_EINVAL, "capacity %lld out of range [0,2^31)", (long long)capacity);
  ON_DEVICE(device);
  ragb_store* s = new ragb_store();
  s->device = device; s->dim = dim; s->dtype = dtype;
  s->ld = ((int64_t)dim + 7) / 8 * 8;
  s->capacity = capacity;
  cudaDeviceProp prop;
  cudaError_t e = cudaGetDeviceProperties(&prop, device);
  if (e != cudaSuccess) { delete s; return fail(RAGB_ECUDA, "cudaGetDeviceProperties: %s", cudaGetErrorString(e)); }
  if (prop.major != 10) { delete s; return fail(RAGB_EUNSUPPORTED, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor); }
  s->n_sm = prop.multiProcessorCount;
  int r = alloc_rows(s, capacity, &s->rows, &s->norms);
  if (r != RAGB_OK) { delete s; return r; }
  e = cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) { cudaFree(s->rows); cudaFree(s->norms); delete s; return fail(RAGB_ECUDA, "cudaStreamCreate: %s", cudaGetErrorString(e)); }
  cudaEventCreateWithFlags(&s->ingest_ev, cudaEventDisableTiming);
  cudaEventCreateWithFlags(&s->search_ev, cudaEventDisableTiming);
  e = cudaMalloc(&s->tickets, (kScanMaxQ + 1) * sizeof(unsigned));
  if (e != cudaSuccess) { cudaStreamDestroy(s->stream); cudaFree(s->rows); cudaFree(s->norms); delete s; return fail(RAGB_ENOMEM, "cudaMalloc tickets: %s", cudaGetErrorString(e)); }
  cudaMemsetAsync(s->tickets, 0, (kScanMaxQ + 1) * sizeof(unsigned), s->stream);
  // padding columns must read as zero
  cudaMemsetAsync(s->rows, 0, (size_t)capacity * s->ld * elem_size(dtype), s->stream);
  cudaStreamSynchronize(s->stream);
  *out = s;
  return RAGB_OK;
}

extern "C" int ragb_store_destroy(ragb_store* s) {
  if (!s) return RAGB_OK;
  {
    DeviceGuard g(s->device);
    cudaStreamSynchronize(s->stream);
    gemm_state_free(&s->gemm);
    cudaFree(s->rows); cudaFree(s->norms); cudaFree(s->ws); cudaFree(s->stage_dev); cudaFree(s->io); cudaFree(s->tickets);
    cudaFree(s->meta_flags); cudaFree(s->meta_src); cudaFree(s->filter); cudaFree(s->filter_aux);
    if (s->stage_host) cudaFreeHost(s->stage_host);
    if (s->pin) cudaFreeHost(s->pin);
    for (auto& ev : s->prof_events) { cudaEventDestroy(ev.first); cudaEventDestroy(ev.second); }
    if (s->ingest_ev) cudaEventDestroy(s->ingest_ev);
    if (s->search_ev) cudaEventDestroy(s->search_ev);
    cudaStreamDestroy(s->stream);
  }
  delete s;
  return RAGB_OK;
}

extern "C" int ragb_store_reserve(ragb_store* s, int64_t capacity) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  if (capacity >= (int64_t)1 << 31) return fail(RAGB_EINVAL, "capacity %lld >= 2^31", (long long)capacity);
  std::lock_guard<std::mutex> la(s->add_mu);
  std::lock_guard<std::mutex> ls(s->search_mu);
  if (capacity <= s->capacity) return RAGB_OK;
  ON_DEVICE(s->device);
  void* rows = nullptr; float* norms = nullptr;
  int64_t old_cap = s->capacity;
  s->capacity = capacity;  // alloc_rows sizes from the argument, not from s
  int r = alloc_rows(s, capacity, &rows, &norms);
  if (r != RAGB_OK) { s->capacity = old_cap; return r; }
  const size_t rb = (size_t)s->ld * elem_size(s->dtype);
  const int64_t n = s->n.load();
  cudaMemcpyAsync(rows, s->rows, (size_t)n * rb, cudaMemcpyDeviceToDevice, s->stream);
  cudaMemsetAsync((char*)rows + (size_t)n * rb, 0, (size_t)(capacity - n) * rb, s->stream);
  cudaMemcpyAsync(norms, s->norms, (size_t)n * 4, cudaMemcpyDeviceToDevice, s->stream);
  cudaError_t e = cudaStreamSynchronize(s->stream);
  if (e != cudaSuccess) { cudaFree(rows); cudaFree(norms); s->capacity = old_cap; return fail(RAGB_ECUDA, "reserve copy: %s", cudaGetErrorString(e)); }
  cudaFree(s->rows); cudaFree(s->norms);
  s->rows = rows; s->norms = norms;
  gemm_state_free(&s->gemm);  // tensor maps point at the old allocation
  return RAGB_OK;
}

extern "C" int ragb_store_clear(ragb_store* s) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  std::lock_guard<std::mutex> la(s->add_mu);
  std::lock_guard<std::mutex> ls(s->search_mu);
  s->n.store(0);
  s->filter_n = -1;
  return RAGB_OK;
}
extern "C" int64_t ragb_store_size(const ragb_store* s) { return s ? s->n.load(std::memory_order_acquire) : 0; }
extern "C" int64_t ragb_store_capacity(const ragb_store* s) { return s ? s->capacity : 0; }
extern "C" int ragb_store_dim(const ragb_store* s) { return s ? s->dim : 0; }
extern "C" int ragb_store_ld(const ragb_store* s) { return s ? (int)s->ld : 0; }
extern "C" int ragb_store_dtype(const ragb_store* s) { return s ? s->dtype : -1; }
extern "C" int ragb_store_device(const ragb_store* s) { return s ? s->device : -1; }
extern "C" void* ragb_store_rows_dev(const ragb_store* s) { return s ? s->rows : nullptr; }
extern "C" float* ragb_store_norms_dev(const ragb_store* s) { return s ? s->norms : nullptr; }

// ---------------------------------------------------------------------------
// ingest: K2 / K1 dispatch
// ---------------------------------------------------------------------------
template <typename Tin, typename Tout>
static int norm_pack_launch(const void* in, int64_t n, int dim, int64_t ld_in, void* out, int64_t ld_out,
                            float* norms, cudaStream_t st) {
  const bool vec_ok = ((uintptr_t)in % 16 == 0) && ((ld_in * sizeof(Tin)) % 16 == 0);
  const int64_t chunks = std::max<int64_t>((dim + 7) / 8, ld_out / 8);
  static const bool generic_only = getenv("RAGB_NORM_PACK_GENERIC") != nullptr;  // A/B runs
#define NP_ARGS (const Tin*)in, n, dim, ld_in, (Tout*)out, ld_out, norms, vec_ok
  // measured on B200 (fp32 -> fp16, 2M x 384): R=1 5.50 TB/s, R=2 5.26; fp16 -> fp16: R=1 3.35, R=4 1.52 (128 registers,
  // 2 CTAs per SM) — occupancy beats bytes in flight per lane here, so one row per warp
  constexpr int RS = 1, RL = 1;
  auto grid_for = [&](int rows_per_warp) { return (unsigned)((n + 8 * rows_per_warp - 1) / (8 * rows_per_warp)); };
  if (generic_only || chunks > 128) norm_pack_kernel<Tin, Tout><<<grid_for(1), 256, 0, st>>>(NP_ARGS);
  else if (chunks <= 32) norm_pack_reg_kernel<Tin, Tout, 1, RS><<<grid_for(RS), 256, 0, st>>>(NP_ARGS);
  else if (chunks <= 64) norm_pack_reg_kernel<Tin, Tout, 2, RS><<<grid_for(RS), 256, 0, st>>>(NP_ARGS);
  else if (chunks <= 96) norm_pack_reg_kernel<Tin, Tout, 3, RL><<<grid_for(RL), 256, 0, st>>>(NP_ARGS);
  else norm_pack_reg_kernel<Tin, Tout, 4, RL><<<grid_for(RL), 256, 0, st>>>(NP_ARGS);
#undef NP_ARGS
  LAUNCHED();
  return RAGB_OK;
}
template <typename Tin>
static int norm_pack_out(const void* in, int64_t n, int dim, int64_t ld_in, void* out, int out_dtype,
                         int64_t ld_out, float* norms, cudaStream_t st) {
  switch (out_dtype) {
    case RAGB_F16: return norm_pack_launch<Tin, __half>(in, n, dim, ld_in, out, ld_out, norms, st);
    case RAGB_BF16: return norm_pack_launch<Tin, __nv_bfloat16>(in, n, dim, ld_in, out, ld_out, norms, st);
    case RAGB_F32: return norm_pack_launch<Tin, float>(in, n, dim, ld_in, out, ld_out, norms, st);
  }
  return fail(RAGB_EINVAL, "bad out dtype %d", out_dtype);
}
static int norm_pack_any(const void* in, int in_dtype, int64_t n, int dim, int64_t ld_in, void* out,
                         int out_dtype, int64_t ld_out, float* norms, cudaStream_t st) {
  if (n == 0) return RAGB_OK;
  if (n < 0 || n > (int64_t)8 * 0x7fffffff) return fail(RAGB_EINVAL, "bad row count %lld", (long long)n);
  if (ld_in < dim) return fail(RAGB_EINVAL, "ld_in %lld < dim %d", (long long)ld_in, dim);
  if (ld_out % 8 != 0 || ld_out < dim || (uintptr_t)out % 16 != 0)
    return fail(RAGB_EINVAL, "packed rows need a 16-byte aligned base and ld %% 8 == 0 (ld_out=%lld)", (long long)ld_out);
  switch (in_dtype) {
    case RAGB_F16: return norm_pack_out<__half>(in, n, dim, ld_in, out, out_dtype, ld_out, norms, st);
    case RAGB_BF16: return norm_pack_out<__nv_bfloat16>(in, n, dim, ld_in, out, out_dtype, ld_out, norms, st);
    case RAGB_F32: return norm_pack_out<float>(in, n, dim, ld_in, out, out_dtype, ld_out, norms, st);
  }
  return fail(RAGB_EINVAL, "bad in dtype %d", in_dtype);
}

template <typename Tin, typename Tm, typename Tout>
static int pool_launch(const void* tok, const void* mask, int64_t batch, int seq, int hidden, void* out,
                       int64_t ld_out, float* norms, cudaStream_t st) {
  const bool vec_ok = ((uintptr_t)tok % 16 == 0) && (((size_t)hidden * sizeof(Tin)) % 16 == 0);
  const int chunks = (hidden + 7) / 8;
  const int W = chunks < kPoolThreads ? chunks : kPoolThreads;
  const int groups = kPoolThreads / W;
  const size_t smem = sizeof(float) * (2 * (((size_t)seq + 3) / 4 * 4) + (size_t)groups * W * 8 + (size_t)chunks * 8 + 8 + 12);
  auto kern = pool_norm_pack_kernel<Tin, Tm, Tout>;
  if (smem > 48 * 1024) {
    if (smem > 200 * 1024) return fail(RAGB_EUNSUPPORTED, "seq %d x hidden %d needs %zu B of shared memory", seq, hidden, smem);
    CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  kern<<<(unsigned)batch, kPoolThreads, smem, st>>>((const Tin*)tok, (const Tm*)mask, seq, hidden, (Tout*)out, ld_out, norms, vec_ok);
  LAUNCHED();
  return RAGB_OK;
}
template <typename Tin, typename Tm>
static int pool_out(const void* tok, const void* mask, int64_t batch, int seq, int hidden, void* out,
                    int out_dtype, int64_t ld_out, float* norms, cudaStream_t st) {
  switch (out_dtype) {
    case RAGB_F16: return pool_launch<Tin, Tm, __half>(tok, mask, batch, seq, hidden, out, ld_out, norms, st);
    case RAGB_BF16: return pool_launch<Tin, Tm, __nv_bfloat16>(tok, mask, batch, seq, hidden, out, ld_out, norms, st);
    case RAGB_F32: return pool_launch<Tin, Tm, float>(tok, mask, batch, seq, hidden, out, ld_out, norms, st);
  }
  return fail(RAGB_EINVAL, "bad out dtype %d", out_dtype);
}
template <typename Tin>
static int pool_mask(const void* tok, const void* mask, int mask_dtype, int64_t batch, int seq, int hidden,
                     void* out, int out_dtype, int64_t ld_out, float* norms, cudaStream_t st) {
  switch (mask_dtype) {
    case RAGB_MASK_U8: return pool_out<Tin, uint8_t>(tok, mask, batch, seq, hidden, out, out_dtype, ld_out, norms, st);
    case RAGB_MASK_I32: return pool_out<Tin, int32_t>(tok, mask, batch, seq, hidden, out, out_dtype, ld_out, norms, st);
    case RAGB_MASK_I64: return pool_out<Tin, int64_t>(tok, mask, batch, seq, hidden, out, out_dtype, ld_out, norms, st);
    case RAGB_MASK_F32: return pool_out<Tin, float>(tok, mask, batch, seq, hidden, out, out_dtype, ld_out, norms, st);
  }
  return fail(RAGB_EINVAL, "bad mask dtype %d", mask_dtype);
}
static size_t mask_size(int mask_dtype) {
  return mask_dtype == RAGB_MASK_U8 ? 1 : mask_dtype == RAGB_MASK_I64 ? 8 : 4;
}
static int pool_any(const void* tok, int tok_dtype, const void* mask, int mask_dtype, int64_t batch, int seq,
                    int hidden, void* out, int out_dtype, int64_t ld_out, float* norms, cudaStream_t st) {
  if (batch == 0) return RAGB_OK;
  if (batch < 0 || batch > 0x7fffffff) return fail(RAGB_EINVAL, "bad batch %lld", (long long)batch);
  if (seq <= 0 || hidden <= 0) return fail(RAGB_EINVAL, "bad seq %d / hidden %d", seq, hidden);
  if (ld_out % 8 != 0 || ld_out < hidden || (uintptr_t)out % 16 != 0)
    return fail(RAGB_EINVAL, "packed rows need a 16-byte aligned base and ld %% 8 == 0 (ld_out=%lld)", (long long)ld_out);
  switch (tok_dtype) {
    case RAGB_F16: return pool_mask<__half>(tok, mask, mask_dtype, batch, seq, hidden, out, out_dtype, ld_out, norms, st);
    case RAGB_BF16: return pool_mask<__nv_bfloat16>(tok, mask, mask_dtype, batch, seq, hidden, out, out_dtype, ld_out, norms, st);
    case RAGB_F32: return pool_mask<float>(tok, mask, mask_dtype, batch, seq, hidden, out, out_dtype, ld_out, norms, st);
  }
  return fail(RAGB_EINVAL, "bad token dtype %d", tok_dtype);
}

extern "C" int ragb_norm_pack_dev(int device, const void* in, int in_dtype, int64_t n, int dim, int64_t ld_in,
                                  void* out, int out_dtype, int64_t ld_out, float* out_norms, void* stream) {
  if (!in || !out) return fail(RAGB_EINVAL, "NULL buffer");
  ON_DEVICE(device);
  return norm_pack_any(in, in_dtype, n, dim, ld_in, out, out_dtype, ld_out, out_norms, (cudaStream_t)stream);
}
extern "C" int ragb_pool_norm_pack_dev(int device, const void* tok, int tok_dtype, const void* mask, int mask_dtype,
                                       int64_t batch, int seq, int hidden, void* out, int out_dtype, int64_t ld_out,
                                       float* out_norms, void* stream) {
  if (!tok || !out) return fail(RAGB_EINVAL, "NULL buffer");
  ON_DEVICE(device);
  return pool_any(tok, tok_dtype, mask, mask_dtype, batch, seq, hidden, out, out_dtype, ld_out, out_norms, (cudaStream_t)stream);
}

// rows [first, first+n) are reserved under add_mu; published after the kernel has finished
static int reserve_rows(ragb_store* s, int64_t n, int64_t* first) {
  const int64_t cur = s->n.load();
  if (n < 0) return fail(RAGB_EINVAL, "negative row count");
  if (cur + n > s->capacity)
    return fail(RAGB_EFULL, "store full: %lld + %lld > capacity %lld (ragb_store_reserve grows it)",
                (long long)cur, (long long)n, (long long)s->capacity);
  *first = cur;
  return RAGB_OK;
}
static void* row_ptr(ragb_store* s, int64_t row) { return (char*)s->rows + (size_t)row * s->ld * elem_size(s->dtype); }

extern "C" int ragb_store_add_dev(ragb_store* s, const void* rows, int in_dtype, int64_t n, int64_t ld_in,
                                  int64_t* out_first_row, void* stream) {
  if (!s || (!rows && n)) return fail(RAGB_EINVAL, "NULL argument");
  std::lock_guard<std::mutex> la(s->add_mu);
  ON_DEVICE(s->device);
  int64_t first = 0;
  RB_TRY(reserve_rows(s, n, &first));
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  RB_TRY(norm_pack_any(rows, in_dtype, n, s->dim, ld_in, row_ptr(s, first), s->dtype, s->ld, s->norms + first, st));
  // the new rows become searchable once the kernel is done: our own stream is synchronised here; on a caller's
  // stream later work of that stream is ordered behind it, and the store's own stream (host entry points) is
  // made to wait for it
  if (!stream) CU_TRY(cudaStreamSynchronize(st));
  else if (s->ingest_ev) { CU_TRY(cudaEventRecord(s->ingest_ev, st)); CU_TRY(cudaStreamWaitEvent(s->stream, s->ingest_ev, 0)); }
  s->n.store(first + n, std::memory_order_release);
  if (out_first_row) *out_first_row = first;
  return RAGB_OK;
}

extern "C" int ragb_store_add(ragb_store* s, const float* rows, int64_t n, int64_t* out_first_row) {
  if (!s || (!rows && n)) return fail(RAGB_EINVAL, "NULL argument");
  std::lock_guard<std::mutex> la(s->add_mu);
  ON_DEVICE(s->device);
  int64_t first = 0;
  RB_TRY(reserve_rows(s, n, &first));
  // stream the rows through a device staging buffer in slabs of <= 64 MiB
  const size_t row_bytes = (size_t)s->dim * 4;
  const int64_t slab_rows = std::max<int64_t>(1, std::min<int64_t>(n, (int64_t)((64u << 20) / row_bytes)));
  RB_TRY(grow(&s->stage_dev, &s->stage_dev_bytes, (size_t)slab_rows * row_bytes, false));
  for (int64_t r0 = 0; r0 < n; r0 += slab_rows) {
    const int64_t m = std::min(slab_rows, n - r0);
    CU_TRY(cudaMemcpyAsync(s->stage_dev, rows + (size_t)r0 * s->dim, (size_t)m * row_bytes, cudaMemcpyHostToDevice, s->stream));
    RB_TRY(norm_pack_any(s->stage_dev, RAGB_F32, m, s->dim, s->dim, row_ptr(s, first + r0), s->dtype, s->ld,
                         s->norms + first + r0, s->stream));
  }
  CU_TRY(cudaStreamSynchronize(s->stream));
  s->n.store(first + n, std::memory_order_release);
  if (out_first_row) *out_first_row = first;
  return RAGB_OK;
}

extern "C" int ragb_store_set(ragb_store* s, int64_t row, const float* vec) {
  if (!s || !vec) return fail(RAGB_EINVAL, "NULL argument");
  std::lock_guard<std::mutex> la(s->add_mu);
  std::lock_guard<std::mutex> ls(s->search_mu);  // no search reads a half-written row
  if (row < 0 || row >= s->n.load()) return fail(RAGB_EINVAL, "row %lld out of range", (long long)row);
  ON_DEVICE(s->device);
  RB_TRY(grow(&s->stage_dev, &s->stage_dev_bytes, (size_t)s->dim * 4, false));
  CU_TRY(cudaMemcpyAsync(s->stage_dev, vec, (size_t)s->dim * 4, cudaMemcpyHostToDevice, s->stream));
  RB_TRY(norm_pack_any(s->stage_dev, RAGB_F32, 1, s->dim, s->dim, row_ptr(s, row), s->dtype, s->ld, s->norms + row, s->stream));
  CU_TRY(cudaStreamSynchronize(s->stream));
  return RAGB_OK;
}

extern "C" int ragb_store_add_pooled_dev(ragb_store* s, const void* tok, int tok_dtype, const void* mask,
                                         int mask_dtype, int64_t batch, int seq, int hidden,
                                         int64_t* out_first_row, void* stream) {
  if (!s || (!tok && batch)) return fail(RAGB_EINVAL, "NULL argument");
  if (hidden != s->dim) return fail(RAGB_EINVAL, "hidden %d != store dim %d", hidden, s->dim);
  std::lock_guard<std::mutex> la(s->add_mu);
  ON_DEVICE(s->device);
  int64_t first = 0;
  RB_TRY(reserve_rows(s, batch, &first));
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  RB_TRY(pool_any(tok, tok_dtype, mask, mask_dtype, batch, seq, hidden, row_ptr(s, first), s->dtype, s->ld, s->norms + first, st));
  if (!stream) CU_TRY(cudaStreamSynchronize(st));
  else if (s->ingest_ev) { CU_TRY(cudaEventRecord(s->ingest_ev, st)); CU_TRY(cudaStreamWaitEvent(s->stream, s->ingest_ev, 0)); }
  s->n.store(first + batch, std::memory_order_release);
  if (out_first_row) *out_first_row = first;
  return RAGB_OK;
}

extern "C" int ragb_store_add_pooled(ragb_store* s, const void* tok, int tok_dtype, const void* mask, int mask_dtype,
                                     int64_t batch, int seq, int hidden, int64_t* out_first_row) {
  if (!s || (!tok && batch)) return fail(RAGB_EINVAL, "NULL argument");
  if (hidden != s->dim) return fail(RAGB_EINVAL, "hidden %d != store dim %d", hidden, s->dim);
  if (seq <= 0) return fail(RAGB_EINVAL, "bad seq %d", seq);
  std::lock_guard<std::mutex> la(s->add_mu);
  ON_DEVICE(s->device);
  int64_t first = 0;
  RB_TRY(reserve_rows(s, batch, &first));
  const size_t tok_b = (size_t)seq * hidden * elem_size(tok_dtype);
  const size_t mask_b = mask ? (size_t)seq * mask_size(mask_dtype) : 0;
  const size_t mask_b_al = (mask_b + 15) / 16 * 16;
  const int64_t slab = std::max<int64_t>(1, std::min<int64_t>(batch, (int64_t)((128u << 20) / (tok_b + mask_b_al))));
  const size_t tok_slab = ((size_t)slab * tok_b + 15) / 16 * 16;
  RB_TRY(grow(&s->stage_dev, &s->stage_dev_bytes, tok_slab + (size_t)slab * mask_b_al + 16, false));
  char* dtok = (char*)s->stage_dev;
  char* dmask = dtok + tok_slab;
  for (int64_t b0 = 0; b0 < batch; b0 += slab) {
    const int64_t m = std::min(slab, batch - b0);
    CU_TRY(cudaMemcpyAsync(dtok, (const char*)tok + (size_t)b0 * tok_b, (size_t)m * tok_b, cudaMemcpyHostToDevice, s->stream));
    if (mask) CU_TRY(cudaMemcpyAsync(dmask, (const char*)mask + (size_t)b0 * mask_b, (size_t)m * mask_b, cudaMemcpyHostToDevice, s->stream));
    RB_TRY(pool_any(dtok, tok_dtype, mask ? dmask : nullptr, mask_dtype, m, seq, hidden, row_ptr(s, first + b0), s->dtype,
                    s->ld, s->norms + first + b0, s->stream));
  }
  CU_TRY(cudaStreamSynchronize(s->stream));
  s->n.store(first + batch, std::memory_order_release);
  if (out_first_row) *out_first_row = first;
  return RAGB_OK;
}

// append rows that are ALREADY packed (what ragb_store_get_rows returned): no arithmetic, bit-exact restore
extern "C" int ragb_store_put_rows(ragb_store* s, const void* packed_rows, const float* norms, int64_t n,
                                   int64_t* out_first_row) {
  if (!s || ((!packed_rows || !norms) && n)) return fail(RAGB_EINVAL, "NULL argument");
  std::lock_guard<std::mutex> la(s->add_mu);
  ON_DEVICE(s->device);
  int64_t first = 0;
  RB_TRY(reserve_rows(s, n, &first));
  if (n) {
    const size_t es = elem_size(s->dtype);
    CU_TRY(cudaMemcpy2DAsync(row_ptr(s, first), (size_t)s->ld * es, packed_rows, (size_t)s->dim * es, (size_t)s->dim * es,
                             (size_t)n, cudaMemcpyHostToDevice, s->stream));
    CU_TRY(cudaMemcpyAsync(s->norms + first, norms, (size_t)n * 4, cudaMemcpyHostToDevice, s->stream));
    CU_TRY(cudaStreamSynchronize(s->stream));
  }
  s->n.store(first + n, std::memory_order_release);
  if (out_first_row) *out_first_row = first;
  return RAGB_OK;
}

extern "C" int ragb_store_get_rows(ragb_store* s, int64_t row0, int64_t n, void* out_rows, float* out_norms) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  if (row0 < 0 || n < 0 || row0 + n > s->n.load()) return fail(RAGB_EINVAL, "rows [%lld,%lld) out of range", (long long)row0, (long long)(row0 + n));
  if (n == 0) return RAGB_OK;
  std::lock_guard<std::mutex> ls(s->search_mu);
  ON_DEVICE(s->device);
  const size_t es = elem_size(s->dtype);
  if (out_rows)
    CU_TRY(cudaMemcpy2DAsync(out_rows, (size_t)s->dim * es, row_ptr(s, row0), (size_t)s->ld * es, (size_t)s->dim * es,
                             (size_t)n, cudaMemcpyDeviceToHost, s->stream));
  if (out_norms) CU_TRY(cudaMemcpyAsync(out_norms, s->norms + row0, (size_t)n * 4, cudaMemcpyDeviceToHost, s->stream));
  CU_TRY(cudaStreamSynchronize(s->stream));
  return RAGB_OK;
}

// ---------------------------------------------------------------------------
// search: K3 scan (+K5 merge) dispatch
// ---------------------------------------------------------------------------
static constexpr int scan_unroll(int cpl) { return cpl <= 1 ? 8 : cpl == 2 ? 6 : cpl == 3 ? 4 : cpl == 4 ? 3 : 2; }

template <typename T, int G, int CPL>
static int scan_launch_one(const ScanParams& p, int64_t nq, int n_sm, cudaStream_t st, int* n_lists) {
  constexpr int U = scan_unroll(CPL);
  constexpr int R = (32 / G) * U;
  const int64_t units = (p.row_end - p.row_begin + R - 1) / R;
  int P = (int)std::min<int64_t>((units + kScanWarps - 1) / kScanWarps, 2 * (int64_t)n_sm);
  if (P < 1) P = 1;
  *n_lists = P;
  if (!p.out) return RAGB_OK;  // sizing pass
  const size_t smem = (size_t)kScanWarps * listbuf_len(p.kpad) * sizeof(uint64_t);
  auto kern = scan_topk_kernel<T, G, CPL, U>;
  if (smem > 48 * 1024) CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<dim3(P, (unsigned)nq), kScanThreads, smem, st>>>(p);
  LAUNCHED();
  return RAGB_OK;
}
template <typename T>
static int scan_launch_generic(const ScanParams& p, int64_t nq, int n_sm, cudaStream_t st, int* n_lists) {
  constexpr int U = 4;
  const int64_t units = (p.row_end - p.row_begin + U - 1) / U;
  int P = (int)std::min<int64_t>((units + kScanWarps - 1) / kScanWarps, 2 * (int64_t)n_sm);
  if (P < 1) P = 1;
  *n_lists = P;
  if (!p.out) return RAGB_OK;
  const size_t smem = (size_t)kScanWarps * listbuf_len(p.kpad) * sizeof(uint64_t) + (size_t)p.ld * sizeof(float);
  if (smem > 220 * 1024) return fail(RAGB_EUNSUPPORTED, "k=%d with dim=%d needs %zu B of shared memory", p.k, p.dim, smem);
  auto kern = scan_topk_generic_kernel<T, U>;
  if (smem > 48 * 1024) CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<dim3(P, (unsigned)nq), kScanThreads, smem, st>>>(p);
  LAUNCHED();
  return RAGB_OK;
}

template <typename T, int G>
static int scan_launch_g(int cpl, const ScanParams& p, int64_t nq, int n_sm, cudaStream_t st, int* nl) {
  switch (cpl) {
    case 1: return scan_launch_one<T, G, 1>(p, nq, n_sm, st, nl);
    case 2: return scan_launch_one<T, G, 2>(p, nq, n_sm, st, nl);
    case 3: return scan_launch_one<T, G, 3>(p, nq, n_sm, st, nl);
    case 4: return scan_launch_one<T, G, 4>(p, nq, n_sm, st, nl);
    case 6: return scan_launch_one<T, G, 6>(p, nq, n_sm, st, nl);
    case 8: return scan_launch_one<T, G, 8>(p, nq, n_sm, st, nl);
  }
  return fail(RAGB_EINVAL, "internal: bad cpl %d", cpl);
}
static bool cpl_ok(int c) { return c == 1 || c == 2 || c == 3 || c == 4 || c == 6 || c == 8; }

template <typename T>
static int scan_launch_t(const ScanParams& p, int64_t nq, int n_sm, cudaStream_t st, int* nl) {
  const int chunks = (int)(p.ld * sizeof(T) / 16);
  int bestG = 0;
  for (int pass = 0; pass < 2 && !bestG; ++pass)
    for (int G = 32; G >= 8; G >>= 1)
      if (chunks % G == 0 && cpl_ok(chunks / G) && (pass == 1 || chunks / G >= 2)) { bestG = G; break; }
  switch (bestG) {
    case 32: return scan_launch_g<T, 32>(chunks / 32, p, nq, n_sm, st, nl);
    case 16: return scan_launch_g<T, 16>(chunks / 16, p, nq, n_sm, st, nl);
    case 8: return scan_launch_g<T, 8>(chunks / 8, p, nq, n_sm, st, nl);
  }
  return scan_launch_generic<T>(p, nq, n_sm, st, nl);
}
static int scan_launch(int dtype, const ScanParams& p, int64_t nq, int n_sm, cudaStream_t st, int* nl) {
  switch (dtype) {
    case RAGB_F16: return scan_launch_t<__half>(p, nq, n_sm, st, nl);
    case RAGB_BF16: return scan_launch_t<__nv_bfloat16>(p, nq, n_sm, st, nl);
    case RAGB_F32: return scan_launch_t<float>(p, nq, n_sm, st, nl);
  }
  return fail(RAGB_EINVAL, "bad dtype %d", dtype);
}

struct WaitSpec { const uint64_t* flags; int n; unsigned long long epoch; int* status; };
static int merge_launch(const uint64_t* keys, int64_t nq, int n_lists, int kpad, int64_t stride_q, int64_t stride_p,
                        int k, float* out_scores, int64_t* out_ids, uint64_t* out_keys, cudaStream_t st,
                        bool unsorted = false, const PeerOut* po = nullptr, const WaitSpec* w = nullptr) {
  if (nq == 0) return RAGB_OK;
  MergeParams m = {};
  m.keys = keys; m.n_lists = n_lists; m.kpad = kpad; m.k = k; m.stride_q = stride_q; m.stride_p = stride_p;
  m.out_scores = out_scores; m.out_ids = out_ids; m.out_keys = out_keys; m.unsorted = unsorted ? 1 : 0;
  if (po) m.peers = *po;
  if (w) { m.wait_flags = w->flags; m.wait_n = w->n; m.wait_epoch = w->epoch; m.status = w->status; }
  const size_t smem = (size_t)kMergeWarps * listbuf_len(kpad) * sizeof(uint64_t);
  if (smem > 48 * 1024) CU_TRY(cudaFuncSetAttribute(merge_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  merge_topk_kernel<<<(unsigned)nq, kMergeThreads, smem, st>>>(m);
  LAUNCHED();
  return RAGB_OK;
}


// ---------------------------------------------------------------------------
// search: K4 batched tcgen05 path
// ---------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int make_map(GemmState* g, CUtensorMap* map, int dtype, const void* base, int dim, int64_t rows, int64_t ld,
                    int box_rows) {
  if (!g->encode_fn) {
    cudaDriverEntryPointQueryResult qres;
    void* fn = nullptr;
    CU_TRY(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    if (!fn || qres != cudaDriverEntryPointSuccess) return fail(RAGB_ECUDA, "cuTensorMapEncodeTiled not available from the driver");
    g->encode_fn = fn;
  }
  cuuint64_t gdim[2] = {(cuuint64_t)dim, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)kGemmKB, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = ((EncodeTiledFn)g->encode_fn)(map, dtype == RAGB_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16,
                                            2, const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(RAGB_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return RAGB_OK;
}

static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return (v && *v) ? atoi(v) : dflt;
}
static bool pair_enabled() { return env_int("RAGB_PAIR", 1) && !env_int("RAGB_GEMM_SS", 0) && !env_int("RAGB_TS", 0); }

static bool gemm_eligible(const ragb_store* s, int64_t nq, int k, int flags, int64_t n_rows) {
  if (s->dtype == RAGB_F32 || (flags & RAGB_FORCE_SCAN)) return false;
  const int kp = kpad_of(k);
  // kpad 64 / 128 exists only in the SM-pair kernel
  if (kp > kGemmMaxKpad && !(kp <= 128 && pair_enabled() && !(flags & RAGB_NO_RESCORE))) return false;
  if ((flags & RAGB_FORCE_GEMM) || nq >= 4) return true;
  // 2 or 3 queries: the scan kernel walks the rows once PER QUERY, which is free while the window sits in L2 and
  // costs a full HBM pass each beyond it (10M x 768: 4.26 / 6.38 ms against 2.23 ms for one tensor-path pass)
  return nq >= 2 && n_rows * s->ld * 2 >= (int64_t)64 << 20;
}

template <typename T>
static int rescore_launch(ragb_store* s, const float* q, int64_t nq, int metric, int flags, float threshold,
                          uint32_t id_base, int kpad, uint64_t* keys, cudaStream_t st) {
  const int64_t warps = nq * kpad;
  rescore_kernel<T><<<(unsigned)((warps + 7) / 8), 256, 0, st>>>((const T*)s->rows, s->ld, s->dim, q, nq, metric, flags,
                                                                threshold, id_base, kpad, keys);
  LAUNCHED();
  return RAGB_OK;
}

// merge the per-split lists, optionally re-score the winners in fp32, decode
static int gemm_finish(ragb_store* s, const float* q, int64_t nq, int k, int kpad, int metric, float threshold,
                       int flags, uint32_t id_base, const uint64_t* parts, int splits, uint64_t* tmp,
                       uint64_t* out_keys, float* out_scores, int64_t* out_ids, cudaStream_t st, int k_sel = 0) {
  const bool rescore = !(flags & RAGB_NO_RESCORE);
  if (!rescore)
    return merge_launch(parts, nq, splits, kpad, (int64_t)splits * kpad, kpad, k, out_scores, out_ids, out_keys, st, false, &s->peer);
  RB_TRY(merge_launch(parts, nq, splits, kpad, (int64_t)splits * kpad, kpad, k_sel > k ? k_sel : k, nullptr, nullptr, tmp, st));
  if (s->dtype == RAGB_BF16) RB_TRY(rescore_launch<__nv_bfloat16>(s, q, nq, metric, flags, threshold, id_base, kpad, tmp, st));
  else RB_TRY(rescore_launch<__half>(s, q, nq, metric, flags, threshold, id_base, kpad, tmp, st));
  return merge_launch(tmp, nq, 1, kpad, kpad, kpad, k, out_scores, out_ids, out_keys, st, /*unsorted=*/true, &s->peer);
}

// SM-pair kernel finish (merge.cuh: finish_chunks_kernel): fold the per-split CHUNK lists, score the 32 rows of every
// selected chunk in fp32, keep the k best rows — one launch, one CTA per query
static int gemm_finish_chunks(ragb_store* s, const float* q, int64_t nq, int k, int kpad, int k_sel, int metric,
                              float threshold, int flags, int64_t rb, int64_t re, const uint8_t* valid, uint32_t id_base,
                              const uint64_t* parts, int splits, uint64_t* out_keys, float* out_scores, int64_t* out_ids,
                              cudaStream_t st) {
  FinishParams f = {};
  f.parts = parts; f.n_lists = splits; f.kpad = kpad; f.k_sel = k_sel; f.k = k;
  f.X = s->rows; f.ld = s->ld; f.dim = s->dim; f.q = q; f.metric = metric; f.flags = flags; f.threshold = threshold;
  f.row_begin = rb; f.row_end = re; f.valid = valid; f.id_base = id_base;
  f.out_scores = out_scores; f.out_ids = out_ids; f.out_keys = out_keys;
  f.peers = s->peer;
  const size_t smem = finish_smem_bytes(splits, kpad);
  if (smem > 200 * 1024) return fail(RAGB_EUNSUPPORTED, "internal: %d splits x kpad %d do not fit the finish kernel", splits, kpad);
  // rows per chunk must match the epilogue that produced the keys (gemm_pair.cuh: ChunkCfg)
  typedef void (*FinishKernel)(const FinishParams);
  const bool small = kpad <= 32 || ChunkCfg<8>::rows == 16;
  static_assert(ChunkCfg<2>::rows == 16 && ChunkCfg<4>::rows == ChunkCfg<8>::rows && (ChunkCfg<8>::rows == 32 || ChunkCfg<8>::rows == 16),
                "finish kernel variants");
  FinishKernel kern = s->dtype == RAGB_BF16 ? (small ? finish_chunks_kernel<__nv_bfloat16, 16> : finish_chunks_kernel<__nv_bfloat16, 32>)
                                            : (small ? finish_chunks_kernel<__half, 16> : finish_chunks_kernel<__half, 32>);
  if (smem > 48 * 1024) CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(unsigned)nq, kMergeThreads, smem, st>>>(f);
  LAUNCHED();
  return RAGB_OK;
}

// SM-pair kernel (gemm_pair.cuh): tcgen05.mma.cta_group::2, M=256 (two query tiles) x N=256
static int gemm_pair_launch(ragb_store* s, const float* q, int64_t nq, int k, int metric, float threshold, int flags,
                            int64_t rb, int64_t re, const uint8_t* valid, uint32_t id_base, uint64_t* out_keys,
                            float* out_scores, int64_t* out_ids, cudaStream_t st);

// A batch of `pairs` query-tile pairs runs as floor(P / pairs) row splits on the P SM pairs of the device; when that
// leaves pairs idle (16 tile pairs -> 4 splits -> 64 of 74), the batch can be cut into equal groups that fill the
// device better (2 x 8 tile pairs -> 9 splits -> 72 of 74).  Measured on B200 (12.5M x 384 bf16, B=4096): top-100,
// whose epilogue is the limit, gains 8.5 % (43.6 -> 39.9 ms); top-10 LOSES 1.4 % — that kernel is power-capped, more
// busy SMs only lower the clock, and every extra group is one more pass over the corpus.  So only k > 32 is split.
static int gemm_pair_search(ragb_store* s, const float* q, int64_t nq, int k, int metric, float threshold, int flags,
                            int64_t rb, int64_t re, const uint8_t* valid, uint32_t id_base, uint64_t* out_keys,
                            float* out_scores, int64_t* out_ids, cudaStream_t st) {
  const int64_t pairs = (nq + 2 * kGemmM - 1) / (2 * kGemmM);
  const int P = s->n_sm / 2;
  int groups = 1;
  if (kpad_of(k) > 32 && pairs >= 8 && pairs <= P && !env_int("RAGB_PAIR_NOSPLIT", 0)) {
    double best = 1e30;
    for (int m = 1; m <= 4; ++m) {
      const int64_t gsz = (pairs + m - 1) / m;
      if (gsz < 4) break;
      double cost = 0;
      for (int64_t left = pairs; left > 0; left -= gsz) cost += 1.0 / (double)(P / (int)std::min(gsz, left));
      cost *= 1.0 + 0.02 * (m - 1);
      if (cost < best) { best = cost; groups = m; }
    }
  }
  if (groups == 1)
    return gemm_pair_launch(s, q, nq, k, metric, threshold, flags, rb, re, valid, id_base, out_keys, out_scores, out_ids, st);
  const int kpad = kpad_of(k);
  const int64_t step = (pairs + groups - 1) / groups * (2 * kGemmM);
  const PeerOut peer = s->peer;
  int rc = RAGB_OK;
  for (int64_t q0 = 0; q0 < nq && rc == RAGB_OK; q0 += step) {
    const int64_t m = std::min(step, nq - q0);
    if (peer.n) {  // this group's slot inside every rank's gather buffer; only the last group publishes the epoch
      PeerOut po = peer;
      for (int r = 0; r < po.n; ++r) po.keys[r] += (size_t)q0 * kpad;
      po.signal = (q0 + step >= nq) ? 1 : 0;
      s->peer = po;
    }
    rc = gemm_pair_launch(s, q + (size_t)q0 * s->dim, m, k, metric, threshold, flags, rb, re, valid, id_base,
                          out_keys ? out_keys + (size_t)q0 * kpad : nullptr, out_scores ? out_scores + (size_t)q0 * k : nullptr,
                          out_ids ? out_ids + (size_t)q0 * k : nullptr, st);
  }
  s->peer = peer;
  return rc;
}

static int gemm_pair_launch(ragb_store* s, const float* q, int64_t nq, int k, int metric, float threshold, int flags,
                            int64_t rb, int64_t re, const uint8_t* valid, uint32_t id_base, uint64_t* out_keys,
                            float* out_scores, int64_t* out_ids, cudaStream_t st) {
  GemmState* g = &s->gemm;
  const int kpad = kpad_of(k);
  const int n_kb = (s->dim + kGemmKB - 1) / kGemmKB;
  int n_qt = (int)((nq + kGemmM - 1) / kGemmM);
  // the pair kernel selects CHUNKS of 32 rows; the fp32 re-score resolves them to rows, so it cannot be skipped
  if (flags & RAGB_NO_RESCORE) return RAGB_EUNSUPPORTED;
  n_qt += n_qt & 1;  // an odd tile count (a single tile included) gets one all-zero query tile
  const int64_t nq_pad = (int64_t)n_qt * kGemmM;
  if (g->pair_rows != s->rows || g->pair_capacity != s->capacity) {
    RB_TRY(make_map(g, &g->map_x_pair, s->dtype, s->rows, s->dim, s->capacity, s->ld, kPairN / 2));
    g->pair_rows = s->rows; g->pair_capacity = s->capacity;
  }
  RB_TRY(grow(&g->qh, &g->qh_bytes, (size_t)nq_pad * s->ld * 2, false));
  if (s->dtype == RAGB_BF16)
    convert_queries_kernel<__nv_bfloat16><<<(unsigned)(nq_pad / 4), 128, 0, st>>>(q, nq, s->dim, s->ld, metric, (__nv_bfloat16*)g->qh);
  else
    convert_queries_kernel<__half><<<(unsigned)(nq_pad / 4), 128, 0, st>>>(q, nq, s->dim, s->ld, metric, (__half*)g->qh);
  LAUNCHED();
  CUtensorMap map_q;
  RB_TRY(make_map(g, &map_q, s->dtype, g->qh, s->dim, nq_pad, s->ld, kGemmM));
  // kpad <= 32: candidate buffers in shared memory, 5-stage ring; kpad 64 / 128: buffers in global memory, 6 stages
  const int E = kpad <= 32 ? 2 : kpad == 64 ? 4 : 8;
  if (kpad > 128) return RAGB_EUNSUPPORTED;
  // dim <= 384: the query tile stays in shared memory, the ring carries corpus k-blocks only (gemm_pair.cuh: kQRes)
  const bool qres = n_kb <= kPairQResMaxKB && env_int("RAGB_PAIR_QRES", 1);
  const int n_stages = qres ? (E == 2 ? kPairStagesSmallQ : kPairStagesLargeQ) : (E == 2 ? kPairStagesSmall : kPairStagesLarge);
  const size_t cand_stride = E == 2 ? (qres ? kPairCapSmallQ : kPairCapSmall) + 1 : 32 * E;
  const size_t smem = 1024 + (qres ? (size_t)n_kb * kStageBytesA + (size_t)n_stages * kStageBytesA : (size_t)n_stages * kPairStageBytes) +
                      (E == 2 ? (size_t)kGemmM * cand_stride * 8 : 0) + (2 * n_stages + 5) * 8 + 16;
  typedef void (*PairKernel)(const CUtensorMap, const CUtensorMap, const TsParams);
  PairKernel kern = qres ? (E == 2 ? gemm_topk_pair_kernel<2, kPairStagesSmallQ, true>
                          : E == 4 ? gemm_topk_pair_kernel<4, kPairStagesLargeQ, true> : gemm_topk_pair_kernel<8, kPairStagesLargeQ, true>)
                         : (E == 2 ? gemm_topk_pair_kernel<2, kPairStagesSmall>
                          : E == 4 ? gemm_topk_pair_kernel<4, kPairStagesLarge> : gemm_topk_pair_kernel<8, kPairStagesLarge>);
  const int ei = (E == 2 ? 0 : E == 4 ? 1 : 2) + (qres ? 3 : 0);
  static bool attr_set[6] = {false, false, false, false, false, false};
  if (!attr_set[ei]) {
    // the resident-query variants are sized by dim: allow the largest (6 k-blocks) once
    const size_t smem_max = smem + (qres ? (size_t)(kPairQResMaxKB - n_kb) * kStageBytesA : 0);
    CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_max));
    attr_set[ei] = true;
  }
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute attr[2];
  cfg.blockDim = dim3(kGemmThreads); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeCooperative;  // only counted in numAttrs when the pairs meet (below)
  attr[1].val.cooperative = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  if (g->pair_clusters[ei] == 0) {
    cfg.gridDim = dim3((unsigned)(2 * (s->n_sm / 2)));
    int nc = 0;
    CU_TRY(cudaOccupancyMaxActiveClusters(&nc, kern, &cfg));
    g->pair_clusters[ei] = nc > 0 ? nc : s->n_sm / 2;
  }
  const int pps = n_qt / 2;
  TsParams p = {};
  p.qh = g->qh; p.ld_q = s->ld; p.n_kb = n_kb; p.nt = kPairN; p.kbs = 1; p.stages = n_stages; p.cluster = 2;
  p.row_begin = rb; p.row_end = re; p.valid = valid; p.nq = nq; p.n_qt = n_qt;
  p.n_tiles = (re - rb + kPairN - 1) / kPairN;
  int splits = std::max(1, g->pair_clusters[ei] / pps);
  if (splits > p.n_tiles) splits = (int)p.n_tiles;
  p.n_splits = splits;
  const int k_sel = kpad;  // chunks kept per query: >= k, the slack absorbs the rounding of the 16-bit queries
  p.flags = flags; p.threshold = threshold; p.k = k_sel; p.kpad = kpad;
  p.sel_slack = s->dtype == RAGB_BF16 ? 8e-3f : 2e-3f;
  p.id_base = 0;  // chunk keys carry LOCAL start rows; the re-score adds id_base
  p.idesc = umma_idesc(s->dtype == RAGB_BF16, 2 * kGemmM, kPairN);
  // pairs of one split meet periodically (gemm_pair.cuh) — only when the whole grid is resident at once, else the
  // meeting would wait for CTAs that cannot start
  p.sync_ctr = nullptr;
  if (pps > 1 && n_qt * splits <= 2 * g->pair_clusters[ei] && env_int("RAGB_PAIR_SYNC", 1)) {
    RB_TRY(grow(&g->sync_ctr, &g->sync_ctr_bytes, 1024 * sizeof(unsigned), false));
    CU_TRY(cudaMemsetAsync(g->sync_ctr, 0, (size_t)splits * sizeof(unsigned), st));
    p.sync_ctr = (unsigned*)g->sync_ctr;
    p.sync_tiles = std::max(1, env_int("RAGB_PAIR_SYNC_TILES", kPairSyncTiles));
  }
  const size_t part_bytes = (size_t)nq_pad * splits * kpad * sizeof(uint64_t);
  const size_t cand_bytes = E == 2 ? 0 : (size_t)n_qt * splits * kGemmM * (32 * E) * sizeof(uint64_t);
  const size_t gthr_bytes = (splits > 1 && env_int("RAGB_PAIR_GTHR", 1)) ? (size_t)nq_pad * sizeof(unsigned) : 0;
  RB_TRY(grow(&s->ws, &s->ws_bytes, part_bytes + cand_bytes + gthr_bytes, false));
  p.out = (uint64_t*)s->ws;
  p.cand_ws = E == 2 ? nullptr : (uint64_t*)((char*)s->ws + part_bytes);
  p.gthr = nullptr;
  p.full_sort = env_int("RAGB_PAIR_FULLSORT", 0);
  if (gthr_bytes) {  // the row splits of a query share their running threshold (gemm_pair.cuh)
    p.gthr = (unsigned*)((char*)s->ws + part_bytes + cand_bytes);
    CU_TRY(cudaMemsetAsync(p.gthr, 0, gthr_bytes, st));
  }
  cfg.gridDim = dim3((unsigned)(n_qt * splits));
  {
    ProfScope prof(s, st);
    cudaError_t le = cudaErrorUnknown;
    if (p.sync_ctr) {
      // The meeting spins on CTAs of the same grid: it is only safe if the whole grid is resident at once.  A
      // COOPERATIVE launch makes the driver guarantee that (all CTAs start together or the launch is refused),
      // also against kernels of other streams / other stores that hold SMs at the moment.
      // (RAGB_PAIR_COOP=0 drops the attribute: Nsight Compute's kernel replay cannot measure cooperative launches)
      // Nsight Compute announces itself in the environment of the process it profiles.
      const bool under_ncu = getenv("NV_COMPUTE_PROFILER_PERFWORKS_DIR") != nullptr;
      cfg.numAttrs = env_int("RAGB_PAIR_COOP", under_ncu ? 0 : 1) ? 2 : 1;
      le = cudaLaunchKernelEx(&cfg, kern, map_q, g->map_x_pair, p);
      if (le != cudaSuccess) {  // refused (grid does not fit right now): launch without the meeting
        (void)cudaGetLastError();
        cfg.numAttrs = 1;
        p.sync_ctr = nullptr;
      }
    }
    if (le != cudaSuccess) CU_TRY(cudaLaunchKernelEx(&cfg, kern, map_q, g->map_x_pair, p));
  }
  LAUNCHED();
  return gemm_finish_chunks(s, q, nq, k, kpad, k_sel, metric, threshold, flags, rb, re, valid, id_base, p.out, splits,
                            out_keys, out_scores, out_ids, st);
}

// query-stationary tcgen05 kernel (gemm_ts.cuh); returns RAGB_EUNSUPPORTED when the shape does not fit
static int gemm_ts_search(ragb_store* s, const float* q, int64_t nq, int k, int metric, float threshold, int flags,
                          int64_t rb, int64_t re, const uint8_t* valid, uint32_t id_base, uint64_t* out_keys,
                          float* out_scores, int64_t* out_ids, cudaStream_t st) {
  GemmState* g = &s->gemm;
  const int kpad = kpad_of(k);
  const int n_kb = (s->dim + kGemmKB - 1) / kGemmKB;
  // mode: queries streamed with the corpus (nt = 256, every TMEM column an accumulator) unless
  // RAGB_TS=1 asks for the TMEM-resident query tile (nt limited by the columns left over)
  const bool ts = env_int("RAGB_TS", 0) != 0;
  int nt = 256;
  if (ts) {
    nt = n_kb <= 8 ? 128 : n_kb <= 12 ? 64 : n_kb <= 14 ? 32 : 0;
    if (!nt) return RAGB_EUNSUPPORTED;
  }
  nt = env_int("RAGB_TS_NT", nt);
  if ((ts ? n_kb * 32 : 0) + 2 * nt > 512 || nt % 32 != 0 || nt < 32) return fail(RAGB_EINVAL, "RAGB_TS_NT=%d does not fit TMEM", nt);
  int kbs = env_int("RAGB_TS_KBS", ts ? (nt == 128 ? 2 : 4) : 1);
  if (kbs > n_kb) kbs = n_kb;
  if (kbs < 1) kbs = 1;
  const size_t stage_bytes = (size_t)kbs * ((size_t)nt * 128 + (ts ? 0 : kStageBytesA));
  const size_t fixed = 1024 + (size_t)kGemmM * kTsCandStride * 8 + (2 * kTsMaxStages + 5) * 8 + 16;
  int stages = (int)((232448 - fixed) / stage_bytes);
  if (stages > kTsMaxStages) stages = kTsMaxStages;
  stages = std::min(stages, std::max(2, env_int("RAGB_TS_STAGES", stages)));
  if (stages < 2) return RAGB_EUNSUPPORTED;
  const size_t smem = fixed + (size_t)stages * stage_bytes;

  if (g->ts_rows != s->rows || g->ts_capacity != s->capacity || g->ts_nt != nt) {
    RB_TRY(make_map(g, &g->map_x_ts, s->dtype, s->rows, s->dim, s->capacity, s->ld, nt));
    g->ts_rows = s->rows; g->ts_capacity = s->capacity; g->ts_nt = nt;
  }
  const int n_qt = (int)((nq + kGemmM - 1) / kGemmM);
  const int64_t nq_pad = (int64_t)n_qt * kGemmM;
  RB_TRY(grow(&g->qh, &g->qh_bytes, (size_t)nq_pad * s->ld * 2, false));
  if (s->dtype == RAGB_BF16)
    convert_queries_kernel<__nv_bfloat16><<<(unsigned)(nq_pad / 4), 128, 0, st>>>(q, nq, s->dim, s->ld, metric, (__nv_bfloat16*)g->qh);
  else
    convert_queries_kernel<__half><<<(unsigned)(nq_pad / 4), 128, 0, st>>>(q, nq, s->dim, s->ld, metric, (__half*)g->qh);
  LAUNCHED();

  CUtensorMap map_q;
  RB_TRY(make_map(g, &map_q, s->dtype, g->qh, s->dim, nq_pad, s->ld, kGemmM));
  auto kern = ts ? gemm_topk_ts_kernel<true> : gemm_topk_ts_kernel<false>;
  static bool attr_set[2] = {false, false};
  if (!attr_set[ts]) {
    CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
    attr_set[ts] = true;
  }
  // cluster size: pairs measured best on B200 (148 SMs all busy, corpus bytes halved on the crossbar,
  // HBM traffic within 5% of one pass because sibling clusters hit in L2); RAGB_TS_CLUSTER=4/8 widen it
  int Cmax = 1;
  while (Cmax < 8 && n_qt % (2 * Cmax) == 0) Cmax *= 2;
  int C = std::min(Cmax, std::max(1, env_int("RAGB_TS_CLUSTER", 2)));
  while (C & (C - 1)) C &= C - 1;  // power of two
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute attr[1];
  cfg.blockDim = dim3(kGemmThreads); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)C; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  // co-resident clusters decide how many row splits run side by side
  int max_clusters = s->n_sm / C;
  if (C > 1) {
    if (g->occ_c != C || g->occ_smem != smem + (ts ? 1 : 0)) {
      cfg.gridDim = dim3((unsigned)(C * (s->n_sm / C)));
      int nc = 0;
      CU_TRY(cudaOccupancyMaxActiveClusters(&nc, kern, &cfg));
      g->occ_c = C; g->occ_smem = smem + (ts ? 1 : 0); g->occ_clusters = nc;
    }
    if (g->occ_clusters > 0) max_clusters = g->occ_clusters;
  }
  const int cps = n_qt / C;
  TsParams p;
  p.qh = g->qh; p.ld_q = s->ld; p.n_kb = n_kb; p.nt = nt; p.kbs = kbs; p.stages = stages; p.cluster = C;
  p.row_begin = rb; p.row_end = re; p.valid = valid; p.nq = nq; p.n_qt = n_qt;
  p.n_tiles = (re - rb + nt - 1) / nt;
  int splits = std::max(1, max_clusters / cps);
  if (splits > p.n_tiles) splits = (int)p.n_tiles;
  p.n_splits = splits;
  // with the re-score on, the tensor-core pass (16-bit queries) only SELECTS: keep all kpad >= k slots so that
  // a candidate the rounding pushed just below the k-th place is still re-scored in fp32
  const int k_sel = (flags & RAGB_NO_RESCORE) ? k : kpad;
  p.flags = flags; p.threshold = threshold; p.k = k_sel; p.kpad = kpad;
  p.id_base = id_base; p.idesc = umma_idesc(s->dtype == RAGB_BF16, kGemmM, nt);
  const size_t part_bytes = (size_t)nq_pad * splits * kpad * sizeof(uint64_t);
  const size_t tmp_bytes = (size_t)nq * kpad * sizeof(uint64_t);
  RB_TRY(grow(&s->ws, &s->ws_bytes, part_bytes + tmp_bytes, false));
  p.out = (uint64_t*)s->ws;
  uint64_t* tmp = (uint64_t*)((char*)s->ws + part_bytes);
  cfg.gridDim = dim3((unsigned)(n_qt * splits));
  {
    ProfScope prof(s, st);
    CU_TRY(cudaLaunchKernelEx(&cfg, kern, map_q, g->map_x_ts, p));
  }
  LAUNCHED();
  return gemm_finish(s, q, nq, k, kpad, metric, threshold, flags, id_base, p.out, splits, tmp, out_keys, out_scores,
                     out_ids, st, k_sel);
}

static int gemm_search(ragb_store* s, const float* q, int64_t nq, int k, int metric, float threshold, int flags,
                       int64_t rb, int64_t re, const uint8_t* valid, uint32_t id_base, uint64_t* out_keys,
                       float* out_scores, int64_t* out_ids, cudaStream_t st) {
  if (pair_enabled()) {
    // One query tile (B <= 128): the pair kernel runs it beside an all-zero tile, i.e. at half its tensor rate, but its
    // chunk-granular epilogue has no warm-up; the single-CTA kernel wastes nothing and streams at ~85 % of HBM peak, but
    // pays ~0.4 ms of row-granular insertions while thresholds are loose.  Measured crossover: N*D around 8e9
    // (1M x 384, B=128: 0.27 vs 0.64 ms; 10M x 768: 2.99 vs 2.90 ms).  RAGB_PAIR_SMALL=0/1 forces either.
    bool want = true;
    if (nq <= kGemmM && kpad_of(k) <= kGemmMaxKpad) {
      const int force = env_int("RAGB_PAIR_SMALL", -1);
      want = force >= 0 ? force != 0 : (double)(re - rb) * s->dim < 4e9;
    }
    if (want) {
      int r = gemm_pair_search(s, q, nq, k, metric, threshold, flags, rb, re, valid, id_base, out_keys, out_scores, out_ids, st);
      if (r != RAGB_EUNSUPPORTED) return r;
    }
  }
  if (!env_int("RAGB_GEMM_SS", 0)) {
    int r = gemm_ts_search(s, q, nq, k, metric, threshold, flags, rb, re, valid, id_base, out_keys, out_scores, out_ids, st);
    if (r != RAGB_EUNSUPPORTED) return r;
  }
  GemmState* g = &s->gemm;
  const int kpad = kpad_of(k);
  if (g->map_rows != s->rows || g->map_capacity != s->capacity) {
    RB_TRY(make_map(g, &g->map_x, s->dtype, s->rows, s->dim, s->capacity, s->ld, kGemmN));
    g->map_rows = s->rows; g->map_capacity = s->capacity;
  }
  const int n_qt = (int)((nq + kGemmM - 1) / kGemmM);
  const int64_t nq_pad = (int64_t)n_qt * kGemmM;
  RB_TRY(grow(&g->qh, &g->qh_bytes, (size_t)nq_pad * s->ld * 2, false));
  if (s->dtype == RAGB_BF16)
    convert_queries_kernel<__nv_bfloat16><<<(unsigned)(nq_pad / 4), 128, 0, st>>>(q, nq, s->dim, s->ld, metric, (__nv_bfloat16*)g->qh);
  else
    convert_queries_kernel<__half><<<(unsigned)(nq_pad / 4), 128, 0, st>>>(q, nq, s->dim, s->ld, metric, (__half*)g->qh);
  LAUNCHED();
  CUtensorMap map_q;
  RB_TRY(make_map(g, &map_q, s->dtype, g->qh, s->dim, nq_pad, s->ld, kGemmM));

  GemmParams p;
  p.dim = s->dim; p.row_begin = rb; p.row_end = re; p.valid = valid; p.nq = nq; p.n_qt = n_qt;
  p.n_tiles = (re - rb + kGemmN - 1) / kGemmN;
  int splits = n_qt >= s->n_sm ? 1 : s->n_sm / n_qt;
  if (splits > p.n_tiles) splits = (int)p.n_tiles;
  p.n_splits = splits;
  p.flags = flags; p.threshold = threshold; p.k = k; p.kpad = kpad; p.stages = 4;
  p.id_base = id_base; p.idesc = umma_idesc(s->dtype == RAGB_BF16, kGemmM, kGemmN);
  const size_t part_bytes = (size_t)nq_pad * splits * kpad * sizeof(uint64_t);
  const size_t tmp_bytes = (size_t)nq * kpad * sizeof(uint64_t);
  RB_TRY(grow(&s->ws, &s->ws_bytes, part_bytes + tmp_bytes, false));
  p.out = (uint64_t*)s->ws;
  uint64_t* tmp = (uint64_t*)((char*)s->ws + part_bytes);
  const size_t smem = 1024 + (size_t)p.stages * kStageBytes + (size_t)kpad * kGemmM * 8 + (2 * p.stages + 4) * 8 + 16;
  static bool attr_set = false;
  if (!attr_set) {
    CU_TRY(cudaFuncSetAttribute(gemm_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
    attr_set = true;
  }
  {
    ProfScope prof(s, st);
    gemm_topk_kernel<<<(unsigned)(n_qt * splits), kGemmThreads, smem, st>>>(map_q, g->map_x, p);
  }
  LAUNCHED();
  return gemm_finish(s, q, nq, k, kpad, metric, threshold, flags, id_base, p.out, splits, tmp, out_keys, out_scores,
                     out_ids, st);
}

static int check_search_args(ragb_store* s, const void* q, int64_t nq, int k, int metric, int64_t* rb, int64_t* re) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  if (nq < 0 || nq > 65535 * 16) return fail(RAGB_EINVAL, "nq %lld out of range", (long long)nq);
  if (nq && !q) return fail(RAGB_EINVAL, "q is NULL");
  if (k <= 0 || k > 1024) return fail(RAGB_EUNSUPPORTED, "k=%d outside [1,1024]", k);
  if (metric != RAGB_IP && metric != RAGB_COSINE) return fail(RAGB_EINVAL, "bad metric %d", metric);
  const int64_t n = s->n.load(std::memory_order_acquire);
  if (*re < 0 || *re > n) *re = n;
  if (*rb < 0) *rb = 0;
  if (*rb > *re) *rb = *re;
  return RAGB_OK;
}

static int resolve_valid(ragb_store* s, const uint8_t** valid, int64_t row_end, bool* is_store_filter);
// A search launched on a caller's stream uses the store's workspace; the store's own stream (host entry points)
// must not start its next search before that one is done.  One caller stream at a time per store is the contract.
static void mark_caller_search(ragb_store* s, cudaStream_t st) {
  if (s->search_ev && cudaEventRecord(s->search_ev, st) == cudaSuccess) s->search_ev_used = true;
}
static void own_stream_waits(ragb_store* s) {
  if (s->search_ev_used && cudaStreamWaitEvent(s->stream, s->search_ev, 0) == cudaSuccess) s->search_ev_used = false;
}

// core: device pointers, caller holds search_mu and has selected the device.
// Writes sorted keys (out_keys, kpad per query) and/or decoded (out_scores, out_ids).
static int search_core(ragb_store* s, const float* q, int64_t nq, int k, int metric, float threshold, int flags,
                       int64_t rb, int64_t re, const uint8_t* valid, uint32_t id_base, uint64_t* out_keys,
                       float* out_scores, int64_t* out_ids, cudaStream_t st) {
  if (nq == 0) return RAGB_OK;
  const int kpad = kpad_of(k);
  if (re - rb <= 0)  // empty store / empty window: RAGResult([], []) (:517-521)
    return merge_launch(nullptr, nq, 0, kpad, 0, 0, k, out_scores, out_ids, out_keys, st, false, &s->peer);
  if (gemm_eligible(s, nq, k, flags, re - rb))
    return gemm_search(s, q, nq, k, metric, threshold, flags, rb, re, valid, id_base, out_keys, out_scores, out_ids, st);
  ScanParams p;
  p.X = s->rows; p.ld = s->ld; p.dim = s->dim; p.row_begin = rb; p.row_end = re; p.valid = valid;
  p.q = q; p.ld_q = s->dim; p.metric = metric; p.flags = flags; p.threshold = threshold;
  p.k = k; p.kpad = kpad; p.id_base = id_base; p.out = nullptr;
  p.tickets = nullptr; p.out_scores = nullptr; p.out_ids = nullptr; p.out_keys = nullptr;
  p.peers = PeerOut{};
  if (s->peer.n && nq > kScanMaxQ) return fail(RAGB_EUNSUPPORTED, "peer exchange supports at most %lld queries per call on the scan path", (long long)kScanMaxQ);
  // the grid y dimension caps one launch at 65535 queries
  int n_lists = 0;
  RB_TRY(scan_launch(s->dtype, p, nq, s->n_sm, st, &n_lists));  // sizing
  const int64_t q_step = kScanMaxQ;
  RB_TRY(grow(&s->ws, &s->ws_bytes, (size_t)std::min(nq, q_step) * n_lists * kpad * sizeof(uint64_t), false));
  // one launch per <= 65535 queries: the last CTA of each query folds the per-CTA lists (RAGB_SCAN_2PASS=1
  // brings the separate merge kernel back, for A/B timing)
  const bool fused = !env_int("RAGB_SCAN_2PASS", 0);
  for (int64_t q0 = 0; q0 < nq; q0 += q_step) {
    const int64_t m = std::min(q_step, nq - q0);
    p.q = q + (size_t)q0 * s->dim;
    p.out = (uint64_t*)s->ws;
    p.tickets = fused ? s->tickets : nullptr;
    if (fused) p.peers = s->peer;
    p.out_scores = out_scores ? out_scores + q0 * k : nullptr;
    p.out_ids = out_ids ? out_ids + q0 * k : nullptr;
    p.out_keys = out_keys ? out_keys + q0 * kpad : nullptr;
    {
      ProfScope prof(s, st);
      RB_TRY(scan_launch(s->dtype, p, m, s->n_sm, st, &n_lists));
    }
    if (!fused)
      RB_TRY(merge_launch((const uint64_t*)s->ws, m, n_lists, kpad, (int64_t)n_lists * kpad, kpad, k, p.out_scores,
                          p.out_ids, p.out_keys, st, false, &s->peer));
  }
  return RAGB_OK;
}

extern "C" int ragb_search_keys_dev(ragb_store* s, const float* q, int64_t nq, int k, int metric, float threshold,
                                    int flags, int64_t row_begin, int64_t row_end, const uint8_t* valid,
                                    uint32_t id_base, uint64_t* out_keys, void* stream) {
  RB_TRY(check_search_args(s, q, nq, k, metric, &row_begin, &row_end));
  if (nq && !out_keys) return fail(RAGB_EINVAL, "out_keys is NULL");
  std::lock_guard<std::mutex> ls(s->search_mu);
  ON_DEVICE(s->device);
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  if (!stream) own_stream_waits(s);
  bool sf;
  RB_TRY(resolve_valid(s, &valid, row_end, &sf));
  RB_TRY(search_core(s, q, nq, k, metric, threshold, flags, row_begin, row_end, valid, id_base, out_keys, nullptr, nullptr, st));
  if (!stream) CU_TRY(cudaStreamSynchronize(st));
  else mark_caller_search(s, st);
  return RAGB_OK;
}

// ---- candidate-set policy on the device (filter.cuh) ----------------------------------------------------------
static int meta_reserve(ragb_store* s, int64_t need) {  // caller holds search_mu, device selected
  if (need <= s->meta_cap) return RAGB_OK;
  const int64_t cap = std::max<int64_t>(std::max<int64_t>(need, s->capacity), 2 * s->meta_cap);
  uint8_t* fl = nullptr; uint16_t* sr = nullptr; uint8_t* fi = nullptr;
  if (cudaMalloc(&fl, cap) != cudaSuccess || cudaMalloc(&sr, cap * 2) != cudaSuccess || cudaMalloc(&fi, cap) != cudaSuccess) {
    cudaFree(fl); cudaFree(sr); cudaFree(fi);
    return fail(RAGB_ENOMEM, "cudaMalloc of %lld rows of filter metadata failed", (long long)cap);
  }
  cudaMemsetAsync(fl, 3, cap, s->stream);       // rows nobody described are live and searchable
  cudaMemsetAsync(sr, 0, cap * 2, s->stream);
  if (s->meta_cap) {
    cudaMemcpyAsync(fl, s->meta_flags, s->meta_cap, cudaMemcpyDeviceToDevice, s->stream);
    cudaMemcpyAsync(sr, s->meta_src, s->meta_cap * 2, cudaMemcpyDeviceToDevice, s->stream);
  }
  CU_TRY(cudaStreamSynchronize(s->stream));
  cudaFree(s->meta_flags); cudaFree(s->meta_src); cudaFree(s->filter);
  s->meta_flags = fl; s->meta_src = sr; s->filter = fi; s->meta_cap = cap; s->filter_n = -1;
  return RAGB_OK;
}

extern "C" int ragb_store_set_meta(ragb_store* s, int64_t row0, int64_t n, const uint8_t* flags, const uint16_t* src) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  if (row0 < 0 || n < 0 || row0 + n > s->n.load()) return fail(RAGB_EINVAL, "rows [%lld,%lld) out of range", (long long)row0, (long long)(row0 + n));
  if (n == 0) return RAGB_OK;
  std::lock_guard<std::mutex> ls(s->search_mu);
  ON_DEVICE(s->device);
  RB_TRY(meta_reserve(s, row0 + n));
  if (flags) CU_TRY(cudaMemcpyAsync(s->meta_flags + row0, flags, (size_t)n, cudaMemcpyHostToDevice, s->stream));
  if (src) CU_TRY(cudaMemcpyAsync(s->meta_src + row0, src, (size_t)n * 2, cudaMemcpyHostToDevice, s->stream));
  CU_TRY(cudaStreamSynchronize(s->stream));
  s->filter_n = -1;
  return RAGB_OK;
}

extern "C" int ragb_store_build_filter(ragb_store* s, const uint8_t* src_allowed, int n_codes, int64_t window,
                                       int64_t* out_total) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  if (n_codes < 0 || n_codes > 65536 || (n_codes && !src_allowed)) return fail(RAGB_EINVAL, "bad source table");
  std::lock_guard<std::mutex> ls(s->search_mu);
  ON_DEVICE(s->device);
  const int64_t n = s->n.load(std::memory_order_acquire);
  if (out_total) *out_total = 0;
  own_stream_waits(s);  // a search still running on a caller's stream may be reading the current mask
  s->filter_n = -1;
  if (n == 0) { s->filter_n = 0; return RAGB_OK; }
  RB_TRY(meta_reserve(s, n));
  const int n_ctas = (int)((n + kFilterRowsPerCta - 1) / kFilterRowsPerCta);
  const size_t counts_b = ((size_t)n_ctas * 4 + 15) / 16 * 16;
  RB_TRY(grow(&s->filter_aux, &s->filter_aux_bytes, counts_b + 32 + 65536, false));
  FilterParams p;
  p.flags = s->meta_flags; p.src = s->meta_src; p.n = n; p.window = window; p.valid = s->filter;
  p.counts = (unsigned*)s->filter_aux;
  p.result = (int64_t*)((char*)s->filter_aux + counts_b);
  uint8_t* allowed_dev = (uint8_t*)s->filter_aux + counts_b + 32;
  p.src_allowed = n_codes ? allowed_dev : nullptr; p.n_codes = n_codes;
  if (n_codes) CU_TRY(cudaMemcpyAsync(allowed_dev, src_allowed, (size_t)n_codes, cudaMemcpyHostToDevice, s->stream));
  filter_mark_kernel<<<n_ctas, kFilterThreads, 0, s->stream>>>(p);
  LAUNCHED();
  filter_cut_kernel<<<1, 1024, 0, s->stream>>>(p, n_ctas);
  LAUNCHED();
  if (window >= 0) {
    filter_window_kernel<<<n_ctas, kFilterThreads, 0, s->stream>>>(p);
    LAUNCHED();
  }
  int64_t res[3] = {0, 0, 0};
  CU_TRY(cudaMemcpyAsync(res, p.result, sizeof(res), cudaMemcpyDeviceToHost, s->stream));
  CU_TRY(cudaStreamSynchronize(s->stream));
  if (out_total) *out_total = res[0];
  s->filter_n = n;
  return RAGB_OK;
}

// valid == RAGB_STORE_FILTER: use the mask ragb_store_build_filter left in the store
static int resolve_valid(ragb_store* s, const uint8_t** valid, int64_t row_end, bool* is_store_filter) {
  *is_store_filter = false;
  if (*valid != RAGB_STORE_FILTER) return RAGB_OK;
  if (s->filter_n < 0) return fail(RAGB_EINVAL, "RAGB_STORE_FILTER: no filter built (ragb_store_build_filter), or rows/metadata changed since");
  if (row_end > s->filter_n) return fail(RAGB_EINVAL, "RAGB_STORE_FILTER covers %lld rows, search asks for %lld: rebuild the filter", (long long)s->filter_n, (long long)row_end);
  *valid = s->filter;
  *is_store_filter = true;
  return RAGB_OK;
}

// ---- candidate exchange over peer memory -------------------------------------------------------------------
static_assert(kMaxPeers == RAGB_MAX_PEERS, "PeerOut and ragb_exchange disagree on the rank limit");
static int check_exchange(const ragb_exchange* x) {
  if (!x) return fail(RAGB_EINVAL, "exchange is NULL");
  if (x->world < 1 || x->world > RAGB_MAX_PEERS || x->rank < 0 || x->rank >= x->world)
    return fail(RAGB_EINVAL, "bad exchange: world %d rank %d (at most %d ranks)", x->world, x->rank, RAGB_MAX_PEERS);
  if (x->epoch == 0) return fail(RAGB_EINVAL, "exchange epoch must be > 0");
  for (int r = 0; r < x->world; ++r)
    if (!x->keys[r] || !x->flags[r]) return fail(RAGB_EINVAL, "exchange buffer of rank %d is NULL", r);
  return RAGB_OK;
}

extern "C" int ragb_search_scatter_dev(ragb_store* s, const float* q, int64_t nq, int k, int metric, float threshold,
                                       int flags, int64_t row_begin, int64_t row_end, const uint8_t* valid,
                                       uint32_t id_base, const ragb_exchange* x, void* stream) {
  RB_TRY(check_search_args(s, q, nq, k, metric, &row_begin, &row_end));
  RB_TRY(check_exchange(x));
  if (nq == 0) return fail(RAGB_EINVAL, "peer exchange needs at least one query");
  std::lock_guard<std::mutex> ls(s->search_mu);
  ON_DEVICE(s->device);
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  const int kpad = kpad_of(k);
  PeerOut po = {};
  po.n = x->world; po.epoch = x->epoch; po.done = s->tickets + kScanMaxQ; po.signal = 1;
  for (int r = 0; r < x->world; ++r) {
    po.keys[r] = x->keys[r] + (size_t)x->rank * nq * kpad;  // this rank's [nq][kpad] slot in rank r's buffer
    po.flag[r] = x->flags[r] + x->rank;
  }
  if (!stream) own_stream_waits(s);
  bool sf;
  RB_TRY(resolve_valid(s, &valid, row_end, &sf));
  s->peer = po;
  const int rc = search_core(s, q, nq, k, metric, threshold, flags, row_begin, row_end, valid, id_base, nullptr, nullptr, nullptr, st);
  s->peer = PeerOut{};
  if (rc != RAGB_OK) return rc;
  if (!stream) CU_TRY(cudaStreamSynchronize(st));
  else mark_caller_search(s, st);
  return RAGB_OK;
}

extern "C" int ragb_gather_merge_dev(int device, const ragb_exchange* x, int64_t nq, int k, float* out_scores,
                                     int64_t* out_ids, int* out_status, void* stream) {
  RB_TRY(check_exchange(x));
  if (nq <= 0 || k <= 0 || k > 1024) return fail(RAGB_EINVAL, "bad merge arguments");
  if (!out_scores || !out_ids) return fail(RAGB_EINVAL, "output buffer is NULL");
  ON_DEVICE(device);
  const int kpad = kpad_of(k);
  WaitSpec w = {x->flags[x->rank], x->world, x->epoch, out_status};
  return merge_launch(x->keys[x->rank], nq, x->world, kpad, kpad, (int64_t)nq * kpad, k, out_scores, out_ids, nullptr,
                      (cudaStream_t)stream, false, nullptr, &w);
}

extern "C" int ragb_search_dev(ragb_store* s, const float* q, int64_t nq, int k, int metric, float threshold, int flags,
                               int64_t row_begin, int64_t row_end, const uint8_t* valid, float* out_scores,
                               int64_t* out_ids, void* stream) {
  RB_TRY(check_search_args(s, q, nq, k, metric, &row_begin, &row_end));
  if (nq && (!out_scores || !out_ids)) return fail(RAGB_EINVAL, "output buffer is NULL");
  std::lock_guard<std::mutex> ls(s->search_mu);
  ON_DEVICE(s->device);
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  if (!stream) own_stream_waits(s);
  bool sf;
  RB_TRY(resolve_valid(s, &valid, row_end, &sf));
  RB_TRY(search_core(s, q, nq, k, metric, threshold, flags, row_begin, row_end, valid, 0, nullptr, out_scores, out_ids, st));
  if (!stream) CU_TRY(cudaStreamSynchronize(st));
  else mark_caller_search(s, st);
  return RAGB_OK;
}

extern "C" int ragb_search(ragb_store* s, const float* q, int64_t nq, int k, int metric, float threshold, int flags,
                           int64_t row_begin, int64_t row_end, const uint8_t* valid, float* out_scores, int64_t* out_ids) {
  RB_TRY(check_search_args(s, q, nq, k, metric, &row_begin, &row_end));
  if (nq && (!out_scores || !out_ids)) return fail(RAGB_EINVAL, "output buffer is NULL");
  if (nq == 0) return RAGB_OK;
  std::lock_guard<std::mutex> ls(s->search_mu);
  ON_DEVICE(s->device);
  // per-call device buffer: q | scores | ids | valid ; pinned mirror of the first three
  const size_t qb = (size_t)nq * s->dim * 4, sb = (size_t)nq * k * 4, ib = (size_t)nq * k * 8;
  bool store_filter;
  RB_TRY(resolve_valid(s, &valid, row_end, &store_filter));  // the store's own mask is already on the device
  const uint8_t* valid_dev = store_filter ? valid : nullptr;
  if (store_filter) valid = nullptr;
  const size_t vb = valid ? (size_t)row_end : 0;
  auto al = [](size_t x) { return (x + 255) / 256 * 256; };
  const size_t o_s = al(qb), o_i = o_s + al(sb), o_v = o_i + al(ib);
  RB_TRY(grow(&s->io, &s->io_bytes, o_v + al(vb), false));
  RB_TRY(grow(&s->pin, &s->pin_bytes, o_v, true));
  char* d = (char*)s->io;
  char* h = (char*)s->pin;
  memcpy(h, q, qb);
  own_stream_waits(s);
  CU_TRY(cudaMemcpyAsync(d, h, qb, cudaMemcpyHostToDevice, s->stream));
  if (valid) CU_TRY(cudaMemcpyAsync(d + o_v, valid, vb, cudaMemcpyHostToDevice, s->stream));
  RB_TRY(search_core(s, (const float*)d, nq, k, metric, threshold, flags, row_begin, row_end,
                     valid ? (const uint8_t*)(d + o_v) : valid_dev, 0, nullptr, (float*)(d + o_s), (int64_t*)(d + o_i), s->stream));
  CU_TRY(cudaMemcpyAsync(h + o_s, d + o_s, o_v - o_s, cudaMemcpyDeviceToHost, s->stream));
  CU_TRY(cudaStreamSynchronize(s->stream));
  memcpy(out_scores, h + o_s, sb);
  memcpy(out_ids, h + o_i, ib);
  return RAGB_OK;
}

extern "C" int ragb_store_profile(ragb_store* s, int enable) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  std::lock_guard<std::mutex> ls(s->search_mu);
  s->profile = enable != 0;
  s->prof_used = 0;
  return RAGB_OK;
}
extern "C" int ragb_store_profile_read(ragb_store* s, double* total_ms, int64_t* n_launches) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  std::lock_guard<std::mutex> ls(s->search_mu);
  ON_DEVICE(s->device);
  double tot = 0;
  for (size_t i = 0; i < s->prof_used; ++i) {
    float ms = 0;
    CU_TRY(cudaEventSynchronize(s->prof_events[i].second));
    CU_TRY(cudaEventElapsedTime(&ms, s->prof_events[i].first, s->prof_events[i].second));
    tot += ms;
  }
  if (total_ms) *total_ms = tot;
  if (n_launches) *n_launches = (int64_t)s->prof_used;
  s->prof_used = 0;
  return RAGB_OK;
}

extern "C" int ragb_merge_keys_dev(int device, const uint64_t* keys, int64_t nq, int n_lists, int kpad, int64_t stride_q,
                                   int64_t stride_p, int k, float* out_scores, int64_t* out_ids, void* stream) {
  if (nq < 0 || n_lists < 0 || k <= 0 || k > 1024) return fail(RAGB_EINVAL, "bad merge arguments");
  if (kpad != kpad_of(kpad) || kpad < k) return fail(RAGB_EINVAL, "kpad %d must be a power of two >= max(k,16)", kpad);
  if (nq && n_lists && !keys) return fail(RAGB_EINVAL, "keys is NULL");
  ON_DEVICE(device);
  return merge_launch(keys, nq, n_lists, kpad, stride_q, stride_p, k, out_scores, out_ids, nullptr, (cudaStream_t)stream);
}

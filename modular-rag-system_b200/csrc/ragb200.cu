// ragb200.cu — C ABI (include/ragb200.h) of the one-device store over the sm_100a kernels.
// Host logic only: store life cycle, search contexts, dispatch.  No torch types, no CPU fallback.  The kernels are
// compiled in their own translation units (k_*.cu) and reached through kernels.h.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "kernels.h"
#include "params.h"
#include "store.h"

using namespace ragb;

// ---------------------------------------------------------------------------
// errors, counters
// ---------------------------------------------------------------------------
static thread_local std::string g_err;
static std::atomic<int64_t> g_launches{0};

namespace ragb {
int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}
void count_launch() { g_launches.fetch_add(1); }

int grow(void** p, size_t* have, size_t need, bool pinned) {
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
int upload_host(void* dev, void* pin, const void* src, size_t bytes, cudaStream_t st) {
  if (bytes == 0) return RAGB_OK;
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, src) == cudaSuccess) {
    if (a.type == cudaMemoryTypeHost) {   // page-locked already: no staging copy
      CU_TRY(cudaMemcpyAsync(dev, src, bytes, cudaMemcpyHostToDevice, st));
      return RAGB_OK;
    }
  } else {
    (void)cudaGetLastError();
  }
  const size_t piece = bytes <= ((size_t)256 << 10) ? bytes : (size_t)512 << 10;
  for (size_t o = 0; o < bytes; o += piece) {
    const size_t m = std::min(piece, bytes - o);
    memcpy((char*)pin + o, (const char*)src + o, m);
    CU_TRY(cudaMemcpyAsync((char*)dev + o, (char*)pin + o, m, cudaMemcpyHostToDevice, st));
  }
  return RAGB_OK;
}
}  // namespace ragb
#define LAUNCH(...)                                                                                          \
  do {                                                                                                       \
    cudaError_t le_ = launch_k(__VA_ARGS__);                                                                 \
    if (le_ == cudaSuccess) le_ = cudaGetLastError();                                                        \
    if (le_ != cudaSuccess)                                                                                  \
      return fail(RAGB_ECUDA, "kernel launch: %s (%s:%d)", cudaGetErrorString(le_), __FILE__, __LINE__);     \
  } while (0)

extern "C" int ragb_abi_version(void) { return RAGB_ABI_VERSION; }
extern "C" const char* ragb_last_error(void) { return g_err.c_str(); }
extern "C" int64_t ragb_launch_count(void) { return g_launches.load(); }
extern "C" int ragb_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}
extern "C" int ragb_kpad(int k) { return kpad_of(k); }

// brackets one kernel launch with events when profiling is on
struct ProfScope {
  ragb_store* s; cudaStream_t st; cudaEvent_t stop = nullptr;
  ProfScope(ragb_store* s_, cudaStream_t st_) : s(s_), st(st_) {
    if (!s->profile) return;
    std::lock_guard<std::mutex> lk(s->prof_mu);
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

// ---------------------------------------------------------------------------
// search contexts
// ---------------------------------------------------------------------------
static void ctx_free(SearchCtx* c) {
  if (!c) return;
  cudaFree(c->ws); cudaFree(c->tickets); cudaFree(c->qh); cudaFree(c->sync_ctr); cudaFree(c->io);
  if (c->pin) cudaFreeHost(c->pin);
  if (c->done_ev) cudaEventDestroy(c->done_ev);
  if (c->t0) cudaEventDestroy(c->t0);
  if (c->t1) cudaEventDestroy(c->t1);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}
static SearchCtx* ctx_new() {  // device selected
  SearchCtx* c = new SearchCtx();
  bool ok = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) == cudaSuccess &&
            cudaEventCreateWithFlags(&c->done_ev, cudaEventDisableTiming) == cudaSuccess &&
            cudaEventCreateWithFlags(&c->t0, cudaEventDisableTiming) == cudaSuccess &&
            cudaEventCreateWithFlags(&c->t1, cudaEventDisableTiming) == cudaSuccess &&
            cudaMalloc(&c->tickets, (kScanMaxQ + 1) * sizeof(unsigned)) == cudaSuccess &&
            cudaMemsetAsync(c->tickets, 0, (kScanMaxQ + 1) * sizeof(unsigned), c->stream) == cudaSuccess &&
            cudaStreamSynchronize(c->stream) == cudaSuccess;
  if (!ok) { cudaGetLastError(); ctx_free(c); return nullptr; }
  return c;
}
int CtxLease::acquire() {
  std::unique_lock<std::mutex> lk(s->pool_mu);
  for (;;) {
    if (!s->free_ctx.empty()) { c = s->free_ctx.back(); s->free_ctx.pop_back(); return RAGB_OK; }
    if ((int)s->all_ctx.size() < kMaxCtx) {
      DeviceGuard g(s->device);  // the stream, events and buffers of a context belong to the store's device
      SearchCtx* n = g.ok ? ctx_new() : nullptr;
      if (!n) return fail(RAGB_ENOMEM, "cannot create a search context on device %d", s->device);
      s->all_ctx.push_back(n);
      c = n;
      return RAGB_OK;
    }
    s->pool_cv.wait(lk);
  }
}
CtxLease::~CtxLease() {
  if (!c) return;
  { std::lock_guard<std::mutex> lk(s->pool_mu); s->free_ctx.push_back(c); }
  s->pool_cv.notify_one();
}

namespace ragb {
// Before a search uses context c on stream st: an earlier search that used c's buffers on a CALLER's stream must be
// done, and so must the last ingest that was launched on a caller's stream (its rows are already published).
int store_prepare_stream(ragb_store* s, SearchCtx* c, cudaStream_t st, bool caller_stream) {
  if (c->done_used) {
    CU_TRY(cudaStreamWaitEvent(st, c->done_ev, 0));
    if (!caller_stream) c->done_used = false;  // own stream: synchronised before the context is handed back
  }
  if (s->ingest_ev_used.load(std::memory_order_acquire)) CU_TRY(cudaStreamWaitEvent(st, s->ingest_ev, 0));
  return RAGB_OK;
}
void store_finish_stream(ragb_store* s, SearchCtx* c, cudaStream_t st, bool caller_stream) {
  (void)s;
  if (caller_stream && cudaEventRecord(c->done_ev, st) == cudaSuccess) c->done_used = true;
}
}  // namespace ragb
// exclusive holders (set, reserve, filter build ...) order the store's own stream behind every caller-stream search
static void own_stream_waits(ragb_store* s) {
  std::lock_guard<std::mutex> lk(s->pool_mu);
  for (SearchCtx* c : s->all_ctx)
    if (c->done_used) cudaStreamWaitEvent(s->stream, c->done_ev, 0);
}

// ---------------------------------------------------------------------------
// store
// ---------------------------------------------------------------------------
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
    cudaDeviceSynchronize();
    for (SearchCtx* c : s->all_ctx) ctx_free(c);
    cudaFree(s->rows); cudaFree(s->norms); cudaFree(s->stage_dev);
    cudaFree(s->meta_flags); cudaFree(s->meta_src); cudaFree(s->filter); cudaFree(s->filter_aux);
    cudaFree(s->ts); cudaFree(s->rank); cudaFree(s->row_of_rank);
    for (auto& ev : s->prof_events) { cudaEventDestroy(ev.first); cudaEventDestroy(ev.second); }
    if (s->ingest_ev) cudaEventDestroy(s->ingest_ev);
    cudaStreamDestroy(s->stream);
  }
  delete s;
  return RAGB_OK;
}

static void gemm_maps_reset(ragb_store* s) {
  std::lock_guard<std::mutex> lk(s->gemm.mu);
  s->gemm.map_rows = s->gemm.ts_rows = s->gemm.pair_rows = nullptr;
  s->gemm.map_capacity = s->gemm.ts_capacity = s->gemm.pair_capacity = -1;
}

extern "C" int ragb_store_reserve(ragb_store* s, int64_t capacity) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  if (capacity >= (int64_t)1 << 31) return fail(RAGB_EINVAL, "capacity %lld >= 2^31", (long long)capacity);
  std::lock_guard<std::mutex> la(s->add_mu);
  std::unique_lock<std::shared_mutex> ls(s->rows_mu);
  if (capacity <= s->capacity) return RAGB_OK;
  ON_DEVICE(s->device);
  void* rows = nullptr; float* norms = nullptr;
  int r = alloc_rows(s, capacity, &rows, &norms);
  if (r != RAGB_OK) return r;
  const size_t rb = (size_t)s->ld * elem_size(s->dtype);
  const int64_t n = s->n.load();
  CU_TRY(cudaDeviceSynchronize());  // every search / ingest in flight reads or writes the old allocation
  cudaMemcpyAsync(rows, s->rows, (size_t)n * rb, cudaMemcpyDeviceToDevice, s->stream);
  cudaMemsetAsync((char*)rows + (size_t)n * rb, 0, (size_t)(capacity - n) * rb, s->stream);
  cudaMemcpyAsync(norms, s->norms, (size_t)n * 4, cudaMemcpyDeviceToDevice, s->stream);
  cudaError_t e = cudaStreamSynchronize(s->stream);
  if (e != cudaSuccess) { cudaFree(rows); cudaFree(norms); return fail(RAGB_ECUDA, "reserve copy: %s", cudaGetErrorString(e)); }
  cudaFree(s->rows); cudaFree(s->norms);
  s->rows = rows; s->norms = norms; s->capacity = capacity;
  gemm_maps_reset(s);  // tensor maps point at the old allocation
  return RAGB_OK;
}

extern "C" int ragb_store_clear(ragb_store* s) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  std::lock_guard<std::mutex> la(s->add_mu);
  std::unique_lock<std::shared_mutex> ls(s->rows_mu);
  s->n.store(0);
  s->filter_n = -1;
  // rows added after a clear are undescribed again: live, searchable, source 0, newest in row order
  if (s->meta_cap) {
    ON_DEVICE(s->device);
    own_stream_waits(s);
    CU_TRY(cudaMemsetAsync(s->meta_flags, 3, (size_t)s->meta_cap, s->stream));
    CU_TRY(cudaMemsetAsync(s->meta_src, 0, (size_t)s->meta_cap * 2, s->stream));
    CU_TRY(cudaStreamSynchronize(s->stream));
  }
  s->ts_n = 0; s->ts_max = -1e300; s->ts_dirty = false; s->ranked = false; s->rank_n = 0;
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

static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return (v && *v) ? atoi(v) : dflt;
}

// ---------------------------------------------------------------------------
// ingest: K2 / K1 dispatch
// ---------------------------------------------------------------------------
static int norm_pack_any(const void* in, int in_dtype, int64_t n, int dim, int64_t ld_in, void* out, int out_dtype,
                         int64_t ld_out, float* norms, cudaStream_t st) {
  if (n == 0) return RAGB_OK;
  if (n < 0 || n > (int64_t)8 * 0x7fffffff) return fail(RAGB_EINVAL, "bad row count %lld", (long long)n);
  if (ld_in < dim) return fail(RAGB_EINVAL, "ld_in %lld < dim %d", (long long)ld_in, dim);
  if (ld_out % 8 != 0 || ld_out < dim || (uintptr_t)out % 16 != 0)
    return fail(RAGB_EINVAL, "packed rows need a 16-byte aligned base and ld %% 8 == 0 (ld_out=%lld)", (long long)ld_out);
  if (in_dtype < 0 || in_dtype > 2) return fail(RAGB_EINVAL, "bad in dtype %d", in_dtype);
  if (out_dtype < 0 || out_dtype > 2) return fail(RAGB_EINVAL, "bad out dtype %d", out_dtype);
  const bool vec_ok = ((uintptr_t)in % 16 == 0) && ((ld_in * elem_size(in_dtype)) % 16 == 0);
  const int64_t chunks = std::max<int64_t>((dim + 7) / 8, ld_out / 8);
  static const bool generic_only = getenv("RAGB_NORM_PACK_GENERIC") != nullptr;  // A/B runs
  // G lanes per row, CPL chunks per lane, fetched once and kept in registers (ingest.cuh): the (G, CPL) that wastes the
  // fewest lane slots; rows longer than 1024 elements take the generic two-pass kernel
  int G = 0, cpl = 0;
  if (!generic_only && chunks <= 128) {
    int best_waste = 1 << 30;
    for (int g = 32; g >= 8; g >>= 1) {
      const int c = (int)((chunks + g - 1) / g);
      if (c > (g == 32 ? 4 : 3)) continue;
      const int waste = g * c - (int)chunks;
      if (waste < best_waste) { best_waste = waste; G = g; cpl = c; }
    }
  }
  const void* kern = k_norm_pack(in_dtype, out_dtype, G, cpl);
  if (!kern) return fail(RAGB_EINVAL, "internal: no normalise+pack kernel for G=%d cpl=%d", G, cpl);
  const int64_t rows_per_cta = 8 * (G ? 32 / G : 1);
  LAUNCH(kern, dim3((unsigned)((n + rows_per_cta - 1) / rows_per_cta)), dim3(256), 0, st, in, n, dim, ld_in, out, ld_out, norms, vec_ok);
  return RAGB_OK;
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
  if (tok_dtype < 0 || tok_dtype > 2) return fail(RAGB_EINVAL, "bad token dtype %d", tok_dtype);
  if (out_dtype < 0 || out_dtype > 2) return fail(RAGB_EINVAL, "bad out dtype %d", out_dtype);
  if (mask_dtype < 0 || mask_dtype > 3) return fail(RAGB_EINVAL, "bad mask dtype %d", mask_dtype);
  const bool vec_ok = ((uintptr_t)tok % 16 == 0) && (((size_t)hidden * elem_size(tok_dtype)) % 16 == 0);
  // persistent ring kernel (ingest.cuh: pool_ring_kernel) whenever the rows can travel by 1-D TMA and a few stages fit
  // (16-bit tokens: 5 333 -> 5 661 GB/s on ragged lengths; fp32 tokens stay on the one-chunk-per-CTA kernel, which
  // already reads at 0.97 of the copy peak and measured 5 % faster than the ring — profiles/r2_pool_ring_ab.txt)
  if (vec_ok && hidden % 8 == 0 && hidden <= 1024 && seq <= 4096 && env_int("RAGB_POOL_RING", elem_size(tok_dtype) == 2 ? 1 : 0)) {
    int dev = 0;
    CU_TRY(cudaGetDevice(&dev));
    cudaDeviceProp* prop = nullptr;
    static cudaDeviceProp props[64];
    static bool have[64] = {};
    static unsigned long long* counters[64] = {};
    static std::atomic<unsigned> next_counter{0};
    static std::mutex mu;
    {
      std::lock_guard<std::mutex> lk(mu);
      if (dev < 64 && !have[dev]) {
        CU_TRY(cudaGetDeviceProperties(&props[dev], dev));
        CU_TRY(cudaMalloc(&counters[dev], 1024 * sizeof(unsigned long long)));
        have[dev] = true;
      }
      if (dev < 64) prop = &props[dev];
    }
    if (prop) {
      const size_t stage = (size_t)kPoolRingRows * hidden * elem_size(tok_dtype);
      // two CTAs per SM with ~100 KB of ring each when at least 4 stages fit, else one CTA with the whole SM
      int ctas = 2, n_stages = 0;
      for (; ctas >= 1; --ctas) {
        const size_t budget = (ctas == 2 ? 110 : 220) * 1024;
        const size_t fixed = pool_ring_smem_bytes(tok_dtype, 0, seq, hidden);
        n_stages = fixed < budget ? (int)std::min<size_t>(16, (budget - fixed) / (stage + sizeof(PoolStageMeta) + 16)) : 0;
        if (n_stages >= (ctas == 2 ? 4 : 2)) break;
      }
      if (ctas >= 1) {
        const void* kern = k_pool_ring(tok_dtype, mask_dtype, out_dtype);
        const size_t smem = pool_ring_smem_bytes(tok_dtype, n_stages, seq, hidden);
        CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        PoolRingParams p = {};
        p.tok = tok; p.mask = mask; p.batch = batch; p.seq = seq; p.hidden = hidden; p.out = out; p.ld_out = ld_out;
        p.norms = norms; p.n_stages = n_stages;
        p.counter = counters[dev] + (next_counter.fetch_add(1) % 1024);   // one chunk counter per launch in flight
        CU_TRY(cudaMemsetAsync(p.counter, 0, sizeof(unsigned long long), st));
        const unsigned grid = (unsigned)std::min<int64_t>(batch, (int64_t)prop->multiProcessorCount * ctas);
        LAUNCH(kern, dim3(grid), dim3(kPoolRingThreads), smem, st, p);
        return RAGB_OK;
      }
    }
  }
  const int T = pool_threads();
  const int chunks = (hidden + 7) / 8;
  const int W = chunks < T ? chunks : T;
  const int groups = T / W;
  const size_t smem = sizeof(float) * (2 * (((size_t)seq + 3) / 4 * 4) + (size_t)groups * W * 8 + (size_t)chunks * 8 + 8 + 12);
  const void* kern = k_pool(tok_dtype, mask_dtype, out_dtype);
  if (smem > 48 * 1024) {
    if (smem > 200 * 1024) return fail(RAGB_EUNSUPPORTED, "seq %d x hidden %d needs %zu B of shared memory", seq, hidden, smem);
    CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  LAUNCH(kern, dim3((unsigned)batch), dim3(T), smem, st, tok, mask, seq, hidden, out, ld_out, norms, vec_ok);
  return RAGB_OK;
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

// rows [first, first+n) are reserved under add_mu; published after the kernel has been ordered before every search
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

// The new rows become searchable once the kernel is done.  Own stream: synchronised here.  Caller's stream: an event is
// recorded behind the kernel and every later search — on whatever stream — waits for it before it reads a row
// (store_prepare_stream), so publishing n right away is safe.
static int publish(ragb_store* s, int64_t first, int64_t n, cudaStream_t st, bool caller_stream) {
  if (!caller_stream) CU_TRY(cudaStreamSynchronize(st));
  else {
    CU_TRY(cudaEventRecord(s->ingest_ev, st));
    s->ingest_ev_used.store(true, std::memory_order_release);
    CU_TRY(cudaStreamWaitEvent(s->stream, s->ingest_ev, 0));
  }
  s->n.store(first + n, std::memory_order_release);
  return RAGB_OK;
}

extern "C" int ragb_store_add_dev(ragb_store* s, const void* rows, int in_dtype, int64_t n, int64_t ld_in,
                                  int64_t* out_first_row, void* stream) {
  if (!s || (!rows && n)) return fail(RAGB_EINVAL, "NULL argument");
  std::lock_guard<std::mutex> la(s->add_mu);
  std::shared_lock<std::shared_mutex> ls(s->rows_mu);  // the allocation must not move under the kernel
  ON_DEVICE(s->device);
  int64_t first = 0;
  RB_TRY(reserve_rows(s, n, &first));
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  RB_TRY(norm_pack_any(rows, in_dtype, n, s->dim, ld_in, row_ptr(s, first), s->dtype, s->ld, s->norms + first, st));
  RB_TRY(publish(s, first, n, st, stream != nullptr));
  if (out_first_row) *out_first_row = first;
  return RAGB_OK;
}

extern "C" int ragb_store_add(ragb_store* s, const float* rows, int64_t n, int64_t* out_first_row) {
  if (!s || (!rows && n)) return fail(RAGB_EINVAL, "NULL argument");
  std::lock_guard<std::mutex> la(s->add_mu);
  std::shared_lock<std::shared_mutex> ls(s->rows_mu);
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
  RB_TRY(publish(s, first, n, s->stream, false));
  if (out_first_row) *out_first_row = first;
  return RAGB_OK;
}

extern "C" int ragb_store_set(ragb_store* s, int64_t row, const float* vec) {
  if (!s || !vec) return fail(RAGB_EINVAL, "NULL argument");
  std::lock_guard<std::mutex> la(s->add_mu);
  std::unique_lock<std::shared_mutex> ls(s->rows_mu);  // no search reads a half-written row
  if (row < 0 || row >= s->n.load()) return fail(RAGB_EINVAL, "row %lld out of range", (long long)row);
  ON_DEVICE(s->device);
  own_stream_waits(s);  // ... nor one still in flight on a caller's stream
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
  std::shared_lock<std::shared_mutex> ls(s->rows_mu);
  ON_DEVICE(s->device);
  int64_t first = 0;
  RB_TRY(reserve_rows(s, batch, &first));
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  RB_TRY(pool_any(tok, tok_dtype, mask, mask_dtype, batch, seq, hidden, row_ptr(s, first), s->dtype, s->ld, s->norms + first, st));
  RB_TRY(publish(s, first, batch, st, stream != nullptr));
  if (out_first_row) *out_first_row = first;
  return RAGB_OK;
}

extern "C" int ragb_store_add_pooled(ragb_store* s, const void* tok, int tok_dtype, const void* mask, int mask_dtype,
                                     int64_t batch, int seq, int hidden, int64_t* out_first_row) {
  if (!s || (!tok && batch)) return fail(RAGB_EINVAL, "NULL argument");
  if (hidden != s->dim) return fail(RAGB_EINVAL, "hidden %d != store dim %d", hidden, s->dim);
  if (seq <= 0) return fail(RAGB_EINVAL, "bad seq %d", seq);
  if (tok_dtype < 0 || tok_dtype > 2) return fail(RAGB_EINVAL, "bad token dtype %d", tok_dtype);
  std::lock_guard<std::mutex> la(s->add_mu);
  std::shared_lock<std::shared_mutex> ls(s->rows_mu);
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
  RB_TRY(publish(s, first, batch, s->stream, false));
  if (out_first_row) *out_first_row = first;
  return RAGB_OK;
}

// append rows that are ALREADY packed (what ragb_store_get_rows returned): no arithmetic, bit-exact restore
extern "C" int ragb_store_put_rows(ragb_store* s, const void* packed_rows, const float* norms, int64_t n,
                                   int64_t* out_first_row) {
  if (!s || ((!packed_rows || !norms) && n)) return fail(RAGB_EINVAL, "NULL argument");
  std::lock_guard<std::mutex> la(s->add_mu);
  std::shared_lock<std::shared_mutex> ls(s->rows_mu);
  ON_DEVICE(s->device);
  int64_t first = 0;
  RB_TRY(reserve_rows(s, n, &first));
  if (n) {
    const size_t es = elem_size(s->dtype);
    CU_TRY(cudaMemcpy2DAsync(row_ptr(s, first), (size_t)s->ld * es, packed_rows, (size_t)s->dim * es, (size_t)s->dim * es,
                             (size_t)n, cudaMemcpyHostToDevice, s->stream));
    CU_TRY(cudaMemcpyAsync(s->norms + first, norms, (size_t)n * 4, cudaMemcpyHostToDevice, s->stream));
  }
  RB_TRY(publish(s, first, n, s->stream, false));
  if (out_first_row) *out_first_row = first;
  return RAGB_OK;
}

extern "C" int ragb_store_get_rows(ragb_store* s, int64_t row0, int64_t n, void* out_rows, float* out_norms) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  std::shared_lock<std::shared_mutex> ls(s->rows_mu);
  if (row0 < 0 || n < 0 || row0 + n > s->n.load()) return fail(RAGB_EINVAL, "rows [%lld,%lld) out of range", (long long)row0, (long long)(row0 + n));
  if (n == 0) return RAGB_OK;
  ON_DEVICE(s->device);
  CtxLease lease(s);
  RB_TRY(lease.acquire());
  cudaStream_t st = lease.c->stream;
  RB_TRY(store_prepare_stream(s, lease.c, st, false));
  const size_t es = elem_size(s->dtype);
  if (out_rows)
    CU_TRY(cudaMemcpy2DAsync(out_rows, (size_t)s->dim * es, row_ptr(s, row0), (size_t)s->ld * es, (size_t)s->dim * es,
                             (size_t)n, cudaMemcpyDeviceToHost, st));
  if (out_norms) CU_TRY(cudaMemcpyAsync(out_norms, s->norms + row0, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaStreamSynchronize(st));
  return RAGB_OK;
}

// ---------------------------------------------------------------------------
// search: K3 scan (+K5 merge) dispatch
// ---------------------------------------------------------------------------
static bool cpl_ok(int c) { return c == 1 || c == 2 || c == 3 || c == 4 || c == 6 || c == 8; }

static int scan_launch(ragb_store* s, const ScanParams& p, int64_t nq, cudaStream_t st, int* n_lists) {
  const int chunks = (int)(p.ld * elem_size(s->dtype) / 16);
  // Lanes per row: the FEWEST lanes whose share of a row is an instantiated chunk count (<= 8 x 16 B).  The loads in
  // flight per lane are the same either way (scan_unroll), but a row then costs log2(G) shuffle steps over fewer lanes:
  // 10M x 768 fp16, batch 1: G=16 x 6 chunks 6 949 GB/s against G=32 x 3 chunks 6 180; 384: G=8 x 6 7 021 against
  // G=16 x 3 6 415 (profiles/r2_scan_lanes_ab.txt).  RAGB_SCAN_G=<8|16|32> forces a width, RAGB_SCAN_WIDE=1 is the old
  // widest-first choice — both for A/B runs.
  int bestG = 0;
  static const int wide = env_int("RAGB_SCAN_WIDE", 0), force = env_int("RAGB_SCAN_G", 0);
  if (wide) {
    for (int pass = 0; pass < 2 && !bestG; ++pass)
      for (int G = 32; G >= 8; G >>= 1)
        if (chunks % G == 0 && cpl_ok(chunks / G) && (pass == 1 || chunks / G >= 2)) { bestG = G; break; }
  } else {
    for (int G = 8; G <= 32; G <<= 1)
      if (chunks % G == 0 && cpl_ok(chunks / G)) { bestG = G; break; }
  }
  if (force && chunks % force == 0 && cpl_ok(chunks / force)) bestG = force;
  int R = 0;
  const void* kern = bestG ? k_scan(s->dtype, bestG, chunks / bestG, &R) : nullptr;
  const bool generic = kern == nullptr;
  if (generic) kern = k_scan_generic(s->dtype, &R);
  if (!kern) return fail(RAGB_EINVAL, "bad dtype %d", s->dtype);
  const int64_t units = (p.row_end - p.row_begin + R - 1) / R;
  // Small windows (BASELINE configs[0]: 10k rows, 15 MB, L2 resident) are latency-bound, and most of that latency is the
  // per-CTA list reduction and the last CTA folding one list per CTA: give every warp a few units before adding CTAs
  // (RAGB_SCAN_UPW, default 4: 10k x 384 fp32 runs on 79 CTAs instead of 296).  Large windows hit the 2-CTAs-per-SM cap
  // either way.
  static const int upw = std::max(1, env_int("RAGB_SCAN_UPW", 4));
  int P = (int)std::min<int64_t>((units + kScanWarps * upw - 1) / (kScanWarps * upw), 2 * (int64_t)s->n_sm);
  if (P < 1) P = 1;
  *n_lists = P;
  if (!p.out) return RAGB_OK;  // sizing pass
  const size_t smem = (size_t)kScanWarps * listbuf_len(p.kpad) * sizeof(uint64_t) + (generic ? (size_t)p.ld * sizeof(float) : 0);
  if (smem > 220 * 1024) return fail(RAGB_EUNSUPPORTED, "k=%d with dim=%d needs %zu B of shared memory", p.k, p.dim, smem);
  if (smem > 48 * 1024) CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  LAUNCH(kern, dim3(P, (unsigned)nq), dim3(kScanThreads), smem, st, p);
  return RAGB_OK;
}

namespace ragb {
int merge_launch(const uint64_t* keys, int64_t nq, int n_lists, int kpad, int64_t stride_q, int64_t stride_p, int k,
                 float* out_scores, int64_t* out_ids, uint64_t* out_keys, const uint32_t* row_of_id, cudaStream_t st,
                 bool unsorted, const PeerOut* po, const uint64_t* wait_flags, int wait_n, unsigned long long wait_epoch,
                 int* status) {
  if (nq == 0) return RAGB_OK;
  MergeParams m = {};
  m.keys = keys; m.n_lists = n_lists; m.kpad = kpad; m.k = k; m.stride_q = stride_q; m.stride_p = stride_p;
  m.out_scores = out_scores; m.out_ids = out_ids; m.out_keys = out_keys; m.unsorted = unsorted ? 1 : 0;
  m.row_of_id = row_of_id;
  if (po) m.peers = *po;
  m.wait_flags = wait_flags; m.wait_n = wait_n; m.wait_epoch = wait_epoch; m.status = status;
  const size_t smem = (size_t)kMergeWarps * listbuf_len(kpad) * sizeof(uint64_t);
  if (smem > 48 * 1024) CU_TRY(cudaFuncSetAttribute(k_merge(), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  LAUNCH(k_merge(), dim3((unsigned)nq), dim3(kMergeThreads), smem, st, m);
  return RAGB_OK;
}
}  // namespace ragb

// ---------------------------------------------------------------------------
// search: K4 batched tcgen05 path
// ---------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// caller holds g->mu
static int make_map(GemmShared* g, CUtensorMap* map, int dtype, const void* base, int dim, int64_t rows, int64_t ld,
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

static IdMap id_map(const ragb_store* s, const SearchReq& r) {
  IdMap m;
  m.rank = (s->ranked && r.global_ids) ? s->rank : nullptr;
  m.base = r.id_base;
  m.shift = r.global_ids ? s->shard_shift : -1;
  m.n_dev = s->shard_n; m.dev = s->shard_dev;
  return m;
}
static const uint32_t* id_decode(const ragb_store* s, const SearchReq& r) {
  return (s->ranked && r.global_ids) ? s->row_of_rank : nullptr;
}

static int convert_queries(ragb_store* s, SearchCtx* c, const SearchReq& r, int64_t nq_pad, cudaStream_t st) {
  RB_TRY(grow(&c->qh, &c->qh_bytes, (size_t)nq_pad * s->ld * 2, false));
  LAUNCH(k_convert(s->dtype), dim3((unsigned)(nq_pad / 4)), dim3(128), 0, st, r.q, r.nq, s->dim, s->ld, r.metric, c->qh);
  return RAGB_OK;
}

// merge the per-split lists, optionally re-score the winners in fp32 (keys carry LOCAL rows until then), decode
static int gemm_finish(ragb_store* s, const SearchReq& r, int kpad, const uint64_t* parts, int splits, uint64_t* tmp,
                       cudaStream_t st, int k_sel) {
  const bool rescore = !(r.flags & RAGB_NO_RESCORE);
  if (!rescore)
    return merge_launch(parts, r.nq, splits, kpad, (int64_t)splits * kpad, kpad, r.k, r.out_scores, r.out_ids, r.out_keys,
                        nullptr, st, false, &r.peer);
  RB_TRY(merge_launch(parts, r.nq, splits, kpad, (int64_t)splits * kpad, kpad, k_sel > r.k ? k_sel : r.k, nullptr, nullptr, tmp,
                      nullptr, st));
  const int64_t warps = r.nq * kpad;
  LAUNCH(k_rescore(s->dtype), dim3((unsigned)((warps + 7) / 8)), dim3(256), 0, st, (const void*)s->rows, s->ld, s->dim, r.q, r.nq,
         r.metric, r.flags, r.threshold, id_map(s, r), kpad, tmp);
  return merge_launch(tmp, r.nq, 1, kpad, kpad, kpad, r.k, r.out_scores, r.out_ids, r.out_keys, id_decode(s, r), st,
                      /*unsorted=*/true, &r.peer);
}

// SM-pair kernel finish (merge.cuh: finish_chunks_kernel): fold the per-split CHUNK lists, score the rows of every
// selected chunk in fp32, keep the k best rows — one launch, one CTA per query
static int gemm_finish_chunks(ragb_store* s, const SearchReq& r, int kpad, int k_sel, const uint64_t* parts, int splits,
                              cudaStream_t st) {
  FinishParams f = {};
  f.parts = parts; f.n_lists = splits; f.kpad = kpad; f.k_sel = k_sel; f.k = r.k;
  f.X = s->rows; f.ld = s->ld; f.dim = s->dim; f.q = r.q; f.metric = r.metric; f.flags = r.flags; f.threshold = r.threshold;
  f.row_begin = r.rb; f.row_end = r.re; f.valid = r.valid; f.ids = id_map(s, r); f.row_of_id = id_decode(s, r);
  f.out_scores = r.out_scores; f.out_ids = r.out_ids; f.out_keys = r.out_keys;
  f.peers = r.peer;
  const size_t smem = finish_smem_bytes(splits, kpad);
  if (smem > 200 * 1024) return fail(RAGB_EUNSUPPORTED, "internal: %d splits x kpad %d do not fit the finish kernel", splits, kpad);
  // rows per chunk must match the epilogue that produced the keys (params.h: ChunkCfg)
  static_assert(ChunkCfg<2>::rows == 16 && ChunkCfg<4>::rows == ChunkCfg<8>::rows && (ChunkCfg<8>::rows == 32 || ChunkCfg<8>::rows == 16),
                "finish kernel variants");
  const bool small = kpad <= 32 || ChunkCfg<8>::rows == 16;
  const void* kern = k_finish(s->dtype, small ? 16 : 32);
  if (smem > 48 * 1024) CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // Programmatic dependent launch (RAGB_FINISH_PDL=1): the CTAs may be scheduled while the tensor pass drains
  // (merge.cuh, step 0).  Measured: no gain (profiles/r2_finish_ab.txt — the persistent tensor-pass CTAs all exit within
  // microseconds of each other, there is no tail to hide under), and the event behind the tensor pass then also waits
  // for the early CTAs, which inflates that kernel's bracket.  Off by default.
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.gridDim = dim3((unsigned)r.nq); cfg.blockDim = dim3(kMergeThreads); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cfg.attrs = attr; cfg.numAttrs = env_int("RAGB_FINISH_PDL", 0) ? 1 : 0;
  void* args[] = {(void*)&f};
  CU_TRY(cudaLaunchKernelExC(&cfg, kern, args));
  count_launch();
  CU_TRY(cudaGetLastError());
  return RAGB_OK;
}

// chunks the tensor pass keeps per query (<= kpad)
static int ksel_of(int k, int kpad, int E) {
  static const int spare = env_int("RAGB_KSEL_SPARE", 2);
  return (E == 2 && spare >= 0) ? std::min(kpad, k + spare) : kpad;
}

// SM-pair kernel (gemm_pair.cuh): tcgen05.mma.cta_group::2, M=256 (two query tiles) x N=256
static int gemm_pair_launch(ragb_store* s, SearchCtx* c, const SearchReq& r, cudaStream_t st) {
  GemmShared* g = &s->gemm;
  const int kpad = kpad_of(r.k);
  const int n_kb = (s->dim + kGemmKB - 1) / kGemmKB;
  int n_qt = (int)((r.nq + kGemmM - 1) / kGemmM);
  // the pair kernel selects CHUNKS of rows; the fp32 re-score resolves them to rows, so it cannot be skipped
  if (r.flags & RAGB_NO_RESCORE) return RAGB_EUNSUPPORTED;
  if (kpad > 128) return RAGB_EUNSUPPORTED;
  n_qt += n_qt & 1;  // an odd tile count (a single tile included) gets one all-zero query tile
  const int64_t nq_pad = (int64_t)n_qt * kGemmM;
  // kpad <= 32: candidate buffers in shared memory, 5-stage ring; kpad 64 / 128: buffers in global memory
  const int E = kpad <= 32 ? 2 : kpad == 64 ? 4 : 8;
  // dim <= 384: the query tile stays in shared memory, the ring carries corpus k-blocks only (gemm_pair.cuh: kQRes)
  // (not for 16 < k <= 32: next to the resident tile a buffer has 48 slots, i.e. k_sel = 32 survivors + the 16 chunks one
  // tile can add and nothing to spare — every insert forces a prune on the next tile: 1M x 384 B=1024 top-32 1.59 ms
  // against 0.55 ms for top-10.  With the queries streamed the buffers have 64 slots.)
  const bool qres = n_kb <= kPairQResMaxKB && (kpad <= 16 || kpad > 32) && env_int("RAGB_PAIR_QRES", 1);
  const int n_stages = qres ? (E == 2 ? kPairStagesSmallQ : kPairStagesLargeQ) : (E == 2 ? kPairStagesSmall : kPairStagesLarge);
  const size_t cand_stride = E == 2 ? (qres ? kPairCapSmallQ : kPairCapSmall) + 1 : 32 * E;
  const size_t smem = 1024 + (qres ? (size_t)n_kb * kStageBytesA + (size_t)n_stages * kStageBytesA : (size_t)n_stages * kPairStageBytes) +
                      (E == 2 ? (size_t)kGemmM * cand_stride * 8 : 0) + (2 * n_stages + 5) * 8 + 16 +
                      (E == 2 ? 0 : 16 + 3 * kGemmM * 4);  // kEpi8, global-memory buffers: per-query count | sorted | threshold
  // 8 epilogue warps (gemm_pair.cuh: kEpi8).  Measured on B200 against 4 (profiles/r2_epi8_ab.txt): 1M x 384 top-10
  // kernel 0.614 -> 0.545 ms (1 281 -> 1 442 TFLOP/s), top-100 1.351 -> 1.144 ms, 10M x 384 6.29 -> 5.92 ms, 12.5M x 384
  // bf16 top-100 15.64 -> 15.33 ms per launch; at dim 768, where the tensor pipe was already 97 % busy, 12.32 -> 12.30 ms.
  // RAGB_PAIR_EPI8=0 brings the 4-warp epilogue back for A/B runs.
  const bool epi8 = env_int("RAGB_PAIR_EPI8", 1) != 0;
  // k > 32: the two epilogue warps of a lane quarter split the tiles instead of the columns, each with its own candidate
  // buffers (gemm_pair.cuh: kPar) — two lists per query and row split.  RAGB_PAIR_PAR=0: column split, for A/B runs.
  // (from two tile pairs up: one pair runs as up to 74 row splits, and 148 lists of 128 keys per query is more than the
  // finish kernel should have to fold)
  // Measured (profiles/r2_topk_epilogue_ab.txt): pays on long runs at dim <= 384 (10M x 384 top-64 6.10 -> 5.83 ms,
  // 12.5M x 384 top-100 14.20 -> 14.06 ms), loses on short ones (each half builds its own top-k: 1M x 384 0.81 -> 0.97 ms)
  // and at dim 768, where the epilogue has slack anyway.  RAGB_PAIR_PAR=0/1 forces it.
  const int64_t tiles_per_cta = (r.re - r.rb) / kPairN / std::max(1, (s->n_sm / 2) / std::max(1, n_qt / 2));
  const int par_env = env_int("RAGB_PAIR_PAR", -1);
  const bool par = E != 2 && epi8 && n_qt >= 4 && (par_env >= 0 ? par_env != 0 : (qres && tiles_per_cta >= 2000));
  const void* kern = k_pair(E, qres, epi8, par);
  const int ei = (E == 2 ? 0 : E == 4 ? 1 : 2) + (qres ? 3 : 0) + (epi8 ? 6 : 0) + (par ? 12 : 0);
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute attr[2];
  cfg.blockDim = dim3(epi8 ? kPairThreads8 : kGemmThreads); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeCooperative;  // only counted in numAttrs when the pairs meet (below)
  attr[1].val.cooperative = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  CUtensorMap map_x, map_q;
  int clusters = 0;
  {
    std::lock_guard<std::mutex> lk(g->mu);  // per store, hence per device: the opt-in below is a per-device attribute
    if (g->pair_rows != s->rows || g->pair_capacity != s->capacity) {
      RB_TRY(make_map(g, &g->map_x_pair, s->dtype, s->rows, s->dim, s->capacity, s->ld, kPairN / 2));
      g->pair_rows = s->rows; g->pair_capacity = s->capacity;
    }
    map_x = g->map_x_pair;
    if (!g->pair_attr[ei]) {
      // the resident-query variants are sized by dim: allow the largest (6 k-blocks) once
      const size_t smem_max = smem + (qres ? (size_t)(kPairQResMaxKB - n_kb) * kStageBytesA : 0);
      CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_max));
      g->pair_attr[ei] = true;
    }
    if (g->pair_clusters[ei] == 0) {
      cfg.gridDim = dim3((unsigned)(2 * (s->n_sm / 2)));
      int nc = 0;
      CU_TRY(cudaOccupancyMaxActiveClusters(&nc, kern, &cfg));
      g->pair_clusters[ei] = nc > 0 ? nc : s->n_sm / 2;
    }
    clusters = g->pair_clusters[ei];
    RB_TRY(convert_queries(s, c, r, nq_pad, st));
    RB_TRY(make_map(g, &map_q, s->dtype, c->qh, s->dim, nq_pad, s->ld, kGemmM));
  }
  const int pps = n_qt / 2;
  TsParams p = {};
  p.qh = c->qh; p.ld_q = s->ld; p.n_kb = n_kb; p.nt = kPairN; p.kbs = 1; p.stages = n_stages; p.cluster = 2;
  p.row_begin = r.rb; p.row_end = r.re; p.valid = r.valid; p.nq = r.nq; p.n_qt = n_qt;
  p.n_tiles = (r.re - r.rb + kPairN - 1) / kPairN;
  int splits = std::max(1, clusters / pps);
  if (splits > p.n_tiles) splits = (int)p.n_tiles;
  p.n_splits = splits;
  p.out_lists = par ? 2 * splits : splits;
  // chunks kept per query: kpad >= k.  With k == kpad there is no spare chunk: a k-th row can lose to a (k+1)-th that
  // scores within the rounding of the 16-bit queries (measured at 1M x 384: 8 / 23 / 49 of 1024 queries for k = 16 /
  // 32 / 128, every swapped row within 3e-5 of the cut — inside the parity band, profiles/r2_parity_ksel_slack_*.json)
  // ... and k + 2 for k <= 32 since r2 (RAGB_KSEL_SPARE, -1 = kpad as before): the finish kernel re-reads 16 rows per
  // selected chunk, so top-10 with 12 chunks instead of 16 is a quarter fewer bytes there, the running threshold is the
  // 12th best chunk instead of the 16th, and two spare chunks still cover the 16-bit rounding at the cut
  const int k_sel = ksel_of(r.k, kpad, E);
  p.flags = r.flags; p.threshold = r.threshold; p.k = k_sel; p.kpad = kpad;
  p.sel_slack = s->dtype == RAGB_BF16 ? 8e-3f : 2e-3f;
  p.id_base = 0;  // chunk keys carry LOCAL start rows; the re-score maps rows to ids
  p.idesc = umma_idesc(s->dtype == RAGB_BF16, 2 * kGemmM, kPairN);
  // pairs of one split meet periodically (gemm_pair.cuh) — only when the whole grid is resident at once, else the
  // meeting would wait for CTAs that cannot start
  p.sync_ctr = nullptr;
  if (pps > 1 && n_qt * splits <= 2 * clusters && env_int("RAGB_PAIR_SYNC", 1)) {
    RB_TRY(grow(&c->sync_ctr, &c->sync_ctr_bytes, 1024 * sizeof(unsigned), false));
    CU_TRY(cudaMemsetAsync(c->sync_ctr, 0, (size_t)splits * sizeof(unsigned), st));
    p.sync_ctr = (unsigned*)c->sync_ctr;
    p.sync_tiles = std::max(1, env_int("RAGB_PAIR_SYNC_TILES", kPairSyncTiles));
  }
  const size_t part_bytes = (size_t)nq_pad * p.out_lists * kpad * sizeof(uint64_t);
  const size_t cand_bytes = E == 2 ? 0 : (size_t)n_qt * splits * (par ? 2 : 1) * kGemmM * (32 * E) * sizeof(uint64_t);
  const size_t gthr_bytes = ((splits > 1 || par) && env_int("RAGB_PAIR_GTHR", 1)) ? (size_t)nq_pad * sizeof(unsigned) : 0;
  RB_TRY(grow(&c->ws, &c->ws_bytes, part_bytes + cand_bytes + gthr_bytes, false));
  p.out = (uint64_t*)c->ws;
  p.cand_ws = E == 2 ? nullptr : (uint64_t*)((char*)c->ws + part_bytes);
  p.gthr = nullptr;
  p.full_sort = env_int("RAGB_PAIR_FULLSORT", 2);   // 2: prune by selection (gemm_pair.cuh warp_prune_select); 0 / 1: the sorts
  p.soft_prune = env_int("RAGB_PAIR_SOFT", 1);
  if (gthr_bytes) {  // the row splits of a query share their running threshold (gemm_pair.cuh)
    p.gthr = (unsigned*)((char*)c->ws + part_bytes + cand_bytes);
    CU_TRY(cudaMemsetAsync(p.gthr, 0, gthr_bytes, st));
  }
  cfg.gridDim = dim3((unsigned)(n_qt * splits));
#ifdef RAGB_PAIR_STATS
  static long long* stats_dev = nullptr;   // build variant only (scripts/pair_stats.sh): one device, one search at a time
  if (!stats_dev) CU_TRY(cudaMalloc(&stats_dev, 1024 * 14 * sizeof(long long)));
  CU_TRY(cudaMemsetAsync(stats_dev, 0, 1024 * 14 * sizeof(long long), st));
  p.stats = stats_dev;
#endif
  void* args[] = {(void*)&map_q, (void*)&map_x, (void*)&p};
  {
    ProfScope prof(s, st);
    cudaError_t le = cudaErrorUnknown;
    if (p.sync_ctr) {
      // The meeting spins on CTAs of the same grid: it is only safe if the whole grid is resident at once.  A
      // COOPERATIVE launch makes the driver guarantee that (all CTAs start together or the launch is refused),
      // also against kernels of other streams / other stores that hold SMs at the moment.
      // (RAGB_PAIR_COOP=0 drops the attribute: Nsight Compute's kernel replay cannot measure cooperative launches;
      // Nsight Compute announces itself in the environment of the process it profiles.)
      const bool under_ncu = getenv("NV_COMPUTE_PROFILER_PERFWORKS_DIR") != nullptr;
      cfg.numAttrs = env_int("RAGB_PAIR_COOP", under_ncu ? 0 : 1) ? 2 : 1;
      le = cudaLaunchKernelExC(&cfg, kern, args);
      if (le != cudaSuccess) {  // refused (grid does not fit right now): launch without the meeting
        (void)cudaGetLastError();
        cfg.numAttrs = 1;
        p.sync_ctr = nullptr;
      }
    }
    if (le != cudaSuccess) CU_TRY(cudaLaunchKernelExC(&cfg, kern, args));
  }
  count_launch();
  CU_TRY(cudaGetLastError());
#ifdef RAGB_PAIR_STATS
  if (env_int("RAGB_PAIR_STATS_PRINT", 0)) {
    static std::vector<long long> h(1024 * 14);
    CU_TRY(cudaStreamSynchronize(st));
    CU_TRY(cudaMemcpy(h.data(), stats_dev, h.size() * sizeof(long long), cudaMemcpyDeviceToHost));
    const int grid = n_qt * splits;
    double sum[14] = {}, mx[14] = {};
    for (int b = 0; b < grid; ++b)
      for (int i = 0; i < 14; ++i) {
        if ((i == 3 || i == 4) && (b & 1)) continue;   // the MMA warp works in the leader only
        sum[i] += (double)h[b * 14 + i]; mx[i] = std::max(mx[i], (double)h[b * 14 + i]);
      }
    static const char* names[14] = {"kernel", "producer: empty slot", "producer: meeting", "MMA: accumulator free", "MMA: stage full",
                                    "epilogue: accumulator full", "epilogue: prune", "epilogue: flush", "epilogue: draining tiles",
                                    "  prune: first barrier", "  prune: prunes", "  prune: second barrier", "  prunes (count, warp 2)",
                                    "  cycles in the prune function"};
    fprintf(stderr, "pair stats: grid %d, E %d, qres %d, tiles per CTA %lld\n", grid, E, (int)qres, (long long)(p.n_tiles / splits));
    for (int i = 0; i < 14; ++i) {
      const double n = (i == 3 || i == 4) ? grid / 2 : grid;
      fprintf(stderr, "  %-28s mean %12.0f cycles (%5.1f %% of the kernel)  max %12.0f\n", names[i], sum[i] / n,
              100.0 * (sum[i] / n) / (sum[0] / grid), mx[i]);
    }
  }
#endif
  return gemm_finish_chunks(s, r, kpad, k_sel, p.out, p.out_lists, st);
}

// One query tile (4 <= B <= 128) on every SM by itself (gemm_pair.cuh: kSolo): tcgen05.mma.cta_group::1, M=128 x N=256,
// every CTA walks its own row split.  The batch is HBM-bound (one pass over the rows), so the grid is one CTA per SM.
static int gemm_solo_launch(ragb_store* s, SearchCtx* c, const SearchReq& r, cudaStream_t st) {
  GemmShared* g = &s->gemm;
  const int kpad = kpad_of(r.k);
  const int n_kb = (s->dim + kGemmKB - 1) / kGemmKB;
  if ((r.flags & RAGB_NO_RESCORE) || kpad > 128 || r.nq > kGemmM) return RAGB_EUNSUPPORTED;
  const int64_t nq_pad = kGemmM;
  const int E = kpad <= 32 ? 2 : kpad == 64 ? 4 : 8;
  const int n_stages = E == 2 ? kSoloStagesSmall : kSoloStagesLarge;
  const size_t cand_stride = E == 2 ? kPairCapSmall + 1 : 32 * E;
  const size_t smem = 1024 + (size_t)n_stages * kStageBytes + (E == 2 ? (size_t)kGemmM * cand_stride * 8 : 0) +
                      (2 * n_stages + 5) * 8 + 16 + (E == 2 ? 0 : 16 + 3 * kGemmM * 4);
  const void* kern = k_solo(E);
  const int ei = E == 2 ? 0 : E == 4 ? 1 : 2;
  CUtensorMap map_x, map_q;
  {
    std::lock_guard<std::mutex> lk(g->mu);
    if (g->pair_rows != s->rows || g->pair_capacity != s->capacity) {   // the pair form's map: boxes of 128 rows
      RB_TRY(make_map(g, &g->map_x_pair, s->dtype, s->rows, s->dim, s->capacity, s->ld, kPairN / 2));
      g->pair_rows = s->rows; g->pair_capacity = s->capacity;
    }
    map_x = g->map_x_pair;
    if (!g->solo_attr[ei]) {
      CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      g->solo_attr[ei] = true;
    }
    RB_TRY(convert_queries(s, c, r, nq_pad, st));
    RB_TRY(make_map(g, &map_q, s->dtype, c->qh, s->dim, nq_pad, s->ld, kGemmM));
  }
  TsParams p = {};
  p.qh = c->qh; p.ld_q = s->ld; p.n_kb = n_kb; p.nt = kPairN; p.kbs = 1; p.stages = n_stages; p.cluster = 1;
  p.row_begin = r.rb; p.row_end = r.re; p.valid = r.valid; p.nq = r.nq; p.n_qt = 1;
  p.n_tiles = (r.re - r.rb + kPairN - 1) / kPairN;
  // a CTA's first tiles run with loose thresholds: give every CTA at least a few tiles before adding CTAs
  int splits = (int)std::min<int64_t>(s->n_sm, std::max<int64_t>(1, p.n_tiles / 4));
  p.n_splits = splits;
  p.out_lists = splits;
  const int k_sel = ksel_of(r.k, kpad, E);
  p.flags = r.flags; p.threshold = r.threshold; p.k = k_sel; p.kpad = kpad;
  p.sel_slack = s->dtype == RAGB_BF16 ? 8e-3f : 2e-3f;
  p.id_base = 0;
  p.idesc = umma_idesc(s->dtype == RAGB_BF16, kGemmM, kPairN);
  p.sync_ctr = nullptr;
  const size_t part_bytes = (size_t)nq_pad * splits * kpad * sizeof(uint64_t);
  const size_t cand_bytes = E == 2 ? 0 : (size_t)splits * kGemmM * (32 * E) * sizeof(uint64_t);
  const size_t gthr_bytes = (splits > 1 && env_int("RAGB_PAIR_GTHR", 1)) ? (size_t)nq_pad * sizeof(unsigned) : 0;
  RB_TRY(grow(&c->ws, &c->ws_bytes, part_bytes + cand_bytes + gthr_bytes, false));
  p.out = (uint64_t*)c->ws;
  p.cand_ws = E == 2 ? nullptr : (uint64_t*)((char*)c->ws + part_bytes);
  p.gthr = nullptr;
  p.full_sort = env_int("RAGB_PAIR_FULLSORT", 2);   // 2: prune by selection (gemm_pair.cuh warp_prune_select); 0 / 1: the sorts
  p.soft_prune = env_int("RAGB_PAIR_SOFT", 1);
  if (gthr_bytes) {
    p.gthr = (unsigned*)((char*)c->ws + part_bytes + cand_bytes);
    CU_TRY(cudaMemsetAsync(p.gthr, 0, gthr_bytes, st));
  }
  {
    ProfScope prof(s, st);
    LAUNCH(kern, dim3((unsigned)splits), dim3(kPairThreads8), smem, st, map_q, map_x, p);
  }
  return gemm_finish_chunks(s, r, kpad, k_sel, p.out, splits, st);
}

// A batch of `pairs` query-tile pairs runs as floor(P / pairs) row splits on the P SM pairs of the device; when that
// leaves pairs idle (16 tile pairs -> 4 splits -> 64 of 74), the batch can be cut into equal groups that fill the
// device better (2 x 8 tile pairs -> 9 splits -> 72 of 74).  Measured on B200 (12.5M x 384 bf16, B=4096): top-100,
// whose epilogue is the limit, gains 8.5 % (43.6 -> 39.9 ms); top-10 LOSES 1.4 % — that kernel is power-capped, more
// busy SMs only lower the clock, and every extra group is one more pass over the corpus.  So only k > 32 is split.
static int gemm_pair_search(ragb_store* s, SearchCtx* c, const SearchReq& r, cudaStream_t st) {
  const int64_t pairs = (r.nq + 2 * kGemmM - 1) / (2 * kGemmM);
  const int P = s->n_sm / 2;
  int groups = 1;
  if (kpad_of(r.k) > 32 && pairs >= 8 && pairs <= P && !env_int("RAGB_PAIR_NOSPLIT", 0)) {
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
  if (groups == 1) return gemm_pair_launch(s, c, r, st);
  const int kpad = kpad_of(r.k);
  const int64_t step = (pairs + groups - 1) / groups * (2 * kGemmM);
  int rc = RAGB_OK;
  for (int64_t q0 = 0; q0 < r.nq && rc == RAGB_OK; q0 += step) {
    SearchReq g = r;
    g.nq = std::min(step, r.nq - q0);
    g.q = r.q + (size_t)q0 * s->dim;
    if (r.out_keys) g.out_keys = r.out_keys + (size_t)q0 * kpad;
    if (r.out_scores) g.out_scores = r.out_scores + (size_t)q0 * r.k;
    if (r.out_ids) g.out_ids = r.out_ids + (size_t)q0 * r.k;
    if (r.peer.n) {  // this group's slot inside every rank's gather buffer; only the last group publishes the epoch
      for (int p = 0; p < g.peer.n; ++p) g.peer.keys[p] += (size_t)q0 * kpad;
      g.peer.signal = (q0 + step >= r.nq) ? r.peer.signal : 0;
    }
    rc = gemm_pair_launch(s, c, g, st);
  }
  return rc;
}

// query-stationary tcgen05 kernel (gemm_ts.cuh); returns RAGB_EUNSUPPORTED when the shape does not fit
static int gemm_ts_search(ragb_store* s, SearchCtx* c, const SearchReq& r, cudaStream_t st) {
  GemmShared* g = &s->gemm;
  const int kpad = kpad_of(r.k);
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
  const int n_qt = (int)((r.nq + kGemmM - 1) / kGemmM);
  const int64_t nq_pad = (int64_t)n_qt * kGemmM;
  const void* kern = k_ts(ts);
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
  CUtensorMap map_x, map_q;
  {
    std::lock_guard<std::mutex> lk(g->mu);
    if (g->ts_rows != s->rows || g->ts_capacity != s->capacity || g->ts_nt != nt) {
      RB_TRY(make_map(g, &g->map_x_ts, s->dtype, s->rows, s->dim, s->capacity, s->ld, nt));
      g->ts_rows = s->rows; g->ts_capacity = s->capacity; g->ts_nt = nt;
    }
    map_x = g->map_x_ts;
    if (!g->ts_attr[ts]) {
      CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
      g->ts_attr[ts] = true;
    }
    if (C > 1) {
      if (g->occ_c != C || g->occ_smem != smem + (ts ? 1 : 0)) {
        cfg.gridDim = dim3((unsigned)(C * (s->n_sm / C)));
        int nc = 0;
        CU_TRY(cudaOccupancyMaxActiveClusters(&nc, kern, &cfg));
        g->occ_c = C; g->occ_smem = smem + (ts ? 1 : 0); g->occ_clusters = nc;
      }
      if (g->occ_clusters > 0) max_clusters = g->occ_clusters;
    }
    RB_TRY(convert_queries(s, c, r, nq_pad, st));
    RB_TRY(make_map(g, &map_q, s->dtype, c->qh, s->dim, nq_pad, s->ld, kGemmM));
  }
  const bool rescore = !(r.flags & RAGB_NO_RESCORE);
  const int cps = n_qt / C;
  TsParams p = {};
  p.qh = c->qh; p.ld_q = s->ld; p.n_kb = n_kb; p.nt = nt; p.kbs = kbs; p.stages = stages; p.cluster = C;
  p.row_begin = r.rb; p.row_end = r.re; p.valid = r.valid; p.nq = r.nq; p.n_qt = n_qt;
  p.n_tiles = (r.re - r.rb + nt - 1) / nt;
  int splits = std::max(1, max_clusters / cps);
  if (splits > p.n_tiles) splits = (int)p.n_tiles;
  p.n_splits = splits;
  // with the re-score on, the tensor-core pass (16-bit queries) only SELECTS: keep all kpad >= k slots so that
  // a candidate the rounding pushed just below the k-th place is still re-scored in fp32
  const int k_sel = rescore ? kpad : r.k;
  p.flags = r.flags; p.threshold = r.threshold; p.k = k_sel; p.kpad = kpad;
  p.id_base = rescore ? 0 : r.id_base; p.idesc = umma_idesc(s->dtype == RAGB_BF16, kGemmM, nt);
  const size_t part_bytes = (size_t)nq_pad * splits * kpad * sizeof(uint64_t);
  const size_t tmp_bytes = (size_t)r.nq * kpad * sizeof(uint64_t);
  RB_TRY(grow(&c->ws, &c->ws_bytes, part_bytes + tmp_bytes, false));
  p.out = (uint64_t*)c->ws;
  uint64_t* tmp = (uint64_t*)((char*)c->ws + part_bytes);
  cfg.gridDim = dim3((unsigned)(n_qt * splits));
  void* args[] = {(void*)&map_q, (void*)&map_x, (void*)&p};
  {
    ProfScope prof(s, st);
    CU_TRY(cudaLaunchKernelExC(&cfg, kern, args));
  }
  count_launch();
  CU_TRY(cudaGetLastError());
  return gemm_finish(s, r, kpad, p.out, splits, tmp, st, k_sel);
}

static int gemm_search(ragb_store* s, SearchCtx* c, const SearchReq& r, cudaStream_t st) {
  if (pair_enabled() && r.nq <= kGemmM && env_int("RAGB_SOLO", 1)) {
    // One query tile (B <= 128): its own kernel form — one CTA per SM, M=128, nothing spent on a zero tile
    int rc = gemm_solo_launch(s, c, r, st);
    if (rc != RAGB_EUNSUPPORTED) return rc;
  }
  if (pair_enabled()) {
    // (RAGB_SOLO=0) One query tile (B <= 128): the pair kernel runs it beside an all-zero tile, i.e. at half its tensor rate, but its
    // chunk-granular epilogue has no warm-up; the single-CTA kernel wastes nothing and streams at ~85 % of HBM peak, but
    // pays ~0.4 ms of row-granular insertions while thresholds are loose.  Measured crossover: N*D around 8e9
    // (1M x 384, B=128: 0.27 vs 0.64 ms; 10M x 768: 2.99 vs 2.90 ms).  RAGB_PAIR_SMALL=0/1 forces either.
    bool want = true;
    if (r.nq <= kGemmM && kpad_of(r.k) <= kGemmMaxKpad) {
      const int force = env_int("RAGB_PAIR_SMALL", -1);
      want = force >= 0 ? force != 0 : (double)(r.re - r.rb) * s->dim < 4e9;
    }
    if (want) {
      int rc = gemm_pair_search(s, c, r, st);
      if (rc != RAGB_EUNSUPPORTED) return rc;
    }
  }
  if (!env_int("RAGB_GEMM_SS", 0)) {
    int rc = gemm_ts_search(s, c, r, st);
    if (rc != RAGB_EUNSUPPORTED) return rc;
  }
  GemmShared* g = &s->gemm;
  const int kpad = kpad_of(r.k);
  const int n_qt = (int)((r.nq + kGemmM - 1) / kGemmM);
  const int64_t nq_pad = (int64_t)n_qt * kGemmM;
  CUtensorMap map_x, map_q;
  {
    std::lock_guard<std::mutex> lk(g->mu);
    if (g->map_rows != s->rows || g->map_capacity != s->capacity) {
      RB_TRY(make_map(g, &g->map_x, s->dtype, s->rows, s->dim, s->capacity, s->ld, kGemmN));
      g->map_rows = s->rows; g->map_capacity = s->capacity;
    }
    map_x = g->map_x;
    if (!g->ss_attr) {
      CU_TRY(cudaFuncSetAttribute(k_gemm_topk(), cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
      g->ss_attr = true;
    }
    RB_TRY(convert_queries(s, c, r, nq_pad, st));
    RB_TRY(make_map(g, &map_q, s->dtype, c->qh, s->dim, nq_pad, s->ld, kGemmM));
  }
  const bool rescore = !(r.flags & RAGB_NO_RESCORE);
  GemmParams p = {};
  p.dim = s->dim; p.row_begin = r.rb; p.row_end = r.re; p.valid = r.valid; p.nq = r.nq; p.n_qt = n_qt;
  p.n_tiles = (r.re - r.rb + kGemmN - 1) / kGemmN;
  int splits = n_qt >= s->n_sm ? 1 : s->n_sm / n_qt;
  if (splits > p.n_tiles) splits = (int)p.n_tiles;
  p.n_splits = splits;
  p.flags = r.flags; p.threshold = r.threshold; p.k = r.k; p.kpad = kpad; p.stages = 4;
  p.id_base = rescore ? 0 : r.id_base; p.idesc = umma_idesc(s->dtype == RAGB_BF16, kGemmM, kGemmN);
  const size_t part_bytes = (size_t)nq_pad * splits * kpad * sizeof(uint64_t);
  const size_t tmp_bytes = (size_t)r.nq * kpad * sizeof(uint64_t);
  RB_TRY(grow(&c->ws, &c->ws_bytes, part_bytes + tmp_bytes, false));
  p.out = (uint64_t*)c->ws;
  uint64_t* tmp = (uint64_t*)((char*)c->ws + part_bytes);
  const size_t smem = 1024 + (size_t)p.stages * kStageBytes + (size_t)kpad * kGemmM * 8 + (2 * p.stages + 4) * 8 + 16;
  {
    ProfScope prof(s, st);
    LAUNCH(k_gemm_topk(), dim3((unsigned)(n_qt * splits)), dim3(kGemmThreads), smem, st, map_q, map_x, p);
  }
  return gemm_finish(s, r, kpad, p.out, splits, tmp, st, 0);
}

namespace ragb {
int store_check_args(ragb_store* s, const void* q, int64_t nq, int k, int metric, int64_t* rb, int64_t* re) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  if (nq < 0 || nq > 65535 * 16) return fail(RAGB_EINVAL, "nq %lld out of range", (long long)nq);
  if (nq && !q) return fail(RAGB_EINVAL, "q is NULL");
  if (k <= 0 || k > 1024) return fail(RAGB_EUNSUPPORTED, "k=%d outside [1,1024]", k);
  if (metric != RAGB_IP && metric != RAGB_COSINE) return fail(RAGB_EINVAL, "bad metric %d", metric);
  const int64_t n = s->n.load(std::memory_order_acquire);  // caller holds rows_mu (shared): clear / reserve are excluded
  if (*re < 0 || *re > n) *re = n;
  if (*rb < 0) *rb = 0;
  if (*rb > *re) *rb = *re;
  return RAGB_OK;
}

// core: device pointers, caller holds rows_mu (shared), has selected the device and leased c.
// Writes sorted keys (out_keys, kpad per query) and/or decoded (out_scores, out_ids).
int search_core(ragb_store* s, SearchCtx* c, const SearchReq& r, cudaStream_t st) {
  if (r.nq == 0) return RAGB_OK;
  const int kpad = kpad_of(r.k);
  if ((r.flags & RAGB_NO_RESCORE) && (s->ranked || s->shard_shift >= 0) && r.global_ids)
    return fail(RAGB_EUNSUPPORTED, "RAGB_NO_RESCORE needs plain row ids (no recency ranks, no row-sharded layout)");
  if (r.re - r.rb <= 0)  // empty store / empty window: RAGResult([], []) (:517-521)
    return merge_launch(nullptr, r.nq, 0, kpad, 0, 0, r.k, r.out_scores, r.out_ids, r.out_keys, nullptr, st, false, &r.peer);
  if (gemm_eligible(s, r.nq, r.k, r.flags, r.re - r.rb)) return gemm_search(s, c, r, st);
  ScanParams p = {};
  p.X = s->rows; p.ld = s->ld; p.dim = s->dim; p.row_begin = r.rb; p.row_end = r.re; p.valid = r.valid;
  p.q = r.q; p.ld_q = s->dim; p.metric = r.metric; p.flags = r.flags; p.threshold = r.threshold;
  p.k = r.k; p.kpad = kpad; p.ids = id_map(s, r); p.row_of_id = id_decode(s, r); p.out = nullptr;
  if (r.peer.n && r.nq > kScanMaxQ) return fail(RAGB_EUNSUPPORTED, "peer exchange supports at most %lld queries per call on the scan path", (long long)kScanMaxQ);
  // the grid y dimension caps one launch at 65535 queries
  int n_lists = 0;
  RB_TRY(scan_launch(s, p, r.nq, st, &n_lists));  // sizing
  const int64_t q_step = kScanMaxQ;
  RB_TRY(grow(&c->ws, &c->ws_bytes, (size_t)std::min(r.nq, q_step) * n_lists * kpad * sizeof(uint64_t), false));
  // one launch per <= 65535 queries: the last CTA of each query folds the per-CTA lists (RAGB_SCAN_2PASS=1
  // brings the separate merge kernel back, for A/B timing)
  const bool fused = !env_int("RAGB_SCAN_2PASS", 0);
  for (int64_t q0 = 0; q0 < r.nq; q0 += q_step) {
    const int64_t m = std::min(q_step, r.nq - q0);
    p.q = r.q + (size_t)q0 * s->dim;
    p.out = (uint64_t*)c->ws;
    p.tickets = fused ? c->tickets : nullptr;
    p.peers = PeerOut{};
    if (fused) p.peers = r.peer;
    p.out_scores = r.out_scores ? r.out_scores + q0 * r.k : nullptr;
    p.out_ids = r.out_ids ? r.out_ids + q0 * r.k : nullptr;
    p.out_keys = r.out_keys ? r.out_keys + q0 * kpad : nullptr;
    {
      ProfScope prof(s, st);
      RB_TRY(scan_launch(s, p, m, st, &n_lists));
    }
    if (!fused)
      RB_TRY(merge_launch((const uint64_t*)c->ws, m, n_lists, kpad, (int64_t)n_lists * kpad, kpad, r.k, p.out_scores,
                          p.out_ids, p.out_keys, p.row_of_id, st, false, &r.peer));
  }
  return RAGB_OK;
}
}  // namespace ragb

// ---- candidate-set policy on the device (filter.cuh, rank.cuh) ---------------------------------------------------
static int meta_reserve(ragb_store* s, int64_t need) {  // caller holds rows_mu exclusively, device selected
  if (need <= s->meta_cap) return RAGB_OK;
  const int64_t cap = std::max<int64_t>(std::max<int64_t>(need, s->capacity), 2 * s->meta_cap);
  uint8_t* fl = nullptr; uint16_t* sr = nullptr; uint8_t* fi = nullptr; double* ts = nullptr;
  if (cudaMalloc(&fl, cap) != cudaSuccess || cudaMalloc(&sr, cap * 2) != cudaSuccess || cudaMalloc(&fi, cap) != cudaSuccess ||
      cudaMalloc(&ts, cap * 8) != cudaSuccess) {
    cudaFree(fl); cudaFree(sr); cudaFree(fi); cudaFree(ts);
    cudaGetLastError();
    return fail(RAGB_ENOMEM, "cudaMalloc of %lld rows of filter metadata failed", (long long)cap);
  }
  own_stream_waits(s);
  cudaMemsetAsync(fl, 3, cap, s->stream);       // rows nobody described are live and searchable
  cudaMemsetAsync(sr, 0, cap * 2, s->stream);
  if (s->meta_cap) {
    cudaMemcpyAsync(fl, s->meta_flags, s->meta_cap, cudaMemcpyDeviceToDevice, s->stream);
    cudaMemcpyAsync(sr, s->meta_src, s->meta_cap * 2, cudaMemcpyDeviceToDevice, s->stream);
    if (s->ts_n) cudaMemcpyAsync(ts, s->ts, (size_t)s->ts_n * 8, cudaMemcpyDeviceToDevice, s->stream);
  }
  CU_TRY(cudaStreamSynchronize(s->stream));
  cudaFree(s->meta_flags); cudaFree(s->meta_src); cudaFree(s->filter); cudaFree(s->ts);
  s->meta_flags = fl; s->meta_src = sr; s->filter = fi; s->ts = ts; s->meta_cap = cap; s->filter_n = -1;
  return RAGB_OK;
}

extern "C" int ragb_store_set_meta(ragb_store* s, int64_t row0, int64_t n, const uint8_t* flags, const uint16_t* src) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  std::unique_lock<std::shared_mutex> ls(s->rows_mu);
  if (row0 < 0 || n < 0 || row0 + n > s->n.load()) return fail(RAGB_EINVAL, "rows [%lld,%lld) out of range", (long long)row0, (long long)(row0 + n));
  if (n == 0) return RAGB_OK;
  ON_DEVICE(s->device);
  RB_TRY(meta_reserve(s, row0 + n));
  own_stream_waits(s);
  if (flags) CU_TRY(cudaMemcpyAsync(s->meta_flags + row0, flags, (size_t)n, cudaMemcpyHostToDevice, s->stream));
  if (src) CU_TRY(cudaMemcpyAsync(s->meta_src + row0, src, (size_t)n * 2, cudaMemcpyHostToDevice, s->stream));
  CU_TRY(cudaStreamSynchronize(s->stream));
  s->filter_n = -1;
  return RAGB_OK;
}

static int ts_fill(ragb_store* s, int64_t r0, int64_t r1) {
  if (r1 <= r0) return RAGB_OK;
  LAUNCH(k_ts_fill(), dim3((unsigned)((r1 - r0 + kRankThreads - 1) / kRankThreads)), dim3(kRankThreads), 0, s->stream, s->ts, r0,
         r1, s->ts_max);
  return RAGB_OK;
}

extern "C" int ragb_store_set_timestamps(ragb_store* s, int64_t row0, int64_t n, const double* ts) {
  if (!s || (!ts && n)) return fail(RAGB_EINVAL, "NULL argument");
  std::unique_lock<std::shared_mutex> ls(s->rows_mu);
  if (row0 < 0 || n < 0 || row0 + n > s->n.load()) return fail(RAGB_EINVAL, "rows [%lld,%lld) out of range", (long long)row0, (long long)(row0 + n));
  if (n == 0) return RAGB_OK;
  ON_DEVICE(s->device);
  RB_TRY(meta_reserve(s, row0 + n));
  own_stream_waits(s);
  RB_TRY(ts_fill(s, s->ts_n, row0));  // rows nobody described are the newest so far, in row order
  CU_TRY(cudaMemcpyAsync(s->ts + row0, ts, (size_t)n * 8, cudaMemcpyHostToDevice, s->stream));
  CU_TRY(cudaStreamSynchronize(s->stream));
  for (int64_t i = 0; i < n; ++i) s->ts_max = std::max(s->ts_max, ts[i]);
  s->ts_n = std::max(s->ts_n, row0 + n);
  s->ts_dirty = true;
  s->filter_n = -1;
  return RAGB_OK;
}

// (Re)build the recency ranks if timestamps do not follow row order; rows_mu held exclusively, device selected.
static int ranks_refresh(ragb_store* s, int64_t n) {
  if (!s->ts || s->ts_n == 0) { s->ranked = false; s->rank_n = 0; s->ts_dirty = false; return RAGB_OK; }
  own_stream_waits(s);
  if (n > s->ts_n) {  // the tail was never described: newest, in row order
    RB_TRY(meta_reserve(s, n));
    RB_TRY(ts_fill(s, s->ts_n, n));
    s->ts_n = n;
  }
  RB_TRY(grow(&s->filter_aux, &s->filter_aux_bytes, 256, false));
  int* flag = (int*)s->filter_aux;
  CU_TRY(cudaMemsetAsync(flag, 0, 4, s->stream));
  if (n > 1) LAUNCH(k_ts_check(), dim3((unsigned)((n + kRankThreads - 1) / kRankThreads)), dim3(kRankThreads), 0, s->stream,
                    (const double*)s->ts, n, flag);
  int out_of_order = 0;
  CU_TRY(cudaMemcpyAsync(&out_of_order, flag, 4, cudaMemcpyDeviceToHost, s->stream));
  CU_TRY(cudaStreamSynchronize(s->stream));
  s->ts_dirty = false;
  if (!out_of_order) { s->ranked = false; s->rank_n = n; return RAGB_OK; }
  if (s->rank_cap < n) {
    cudaFree(s->rank); cudaFree(s->row_of_rank); s->rank = s->row_of_rank = nullptr; s->rank_cap = 0;
    const int64_t cap = std::max<int64_t>(n, s->capacity);
    CU_TRY(cudaMalloc(&s->rank, (size_t)cap * 4));
    CU_TRY(cudaMalloc(&s->row_of_rank, (size_t)cap * 4));
    s->rank_cap = cap;
  }
  int64_t npad = 1;
  while (npad < n) npad <<= 1;
  uint64_t* keys = nullptr; uint32_t* idx = nullptr;
  CU_TRY(cudaMalloc(&keys, (size_t)npad * 8));
  cudaError_t e = cudaMalloc(&idx, (size_t)npad * 4);
  if (e != cudaSuccess) { cudaFree(keys); return fail(RAGB_ENOMEM, "cudaMalloc rank scratch: %s", cudaGetErrorString(e)); }
  const unsigned gp = (unsigned)((npad + kRankThreads - 1) / kRankThreads);
  const unsigned gh = (unsigned)std::max<int64_t>(1, (npad / 2 + kRankThreads - 1) / kRankThreads);
  auto body = [&]() -> int {
    LAUNCH(k_rank_init(), dim3(gp), dim3(kRankThreads), 0, s->stream, (const double*)s->ts, n, npad, keys, idx);
    for (int64_t size = 2; size <= npad; size <<= 1)
      for (int64_t stride = size >> 1; stride > 0; stride >>= 1)
        LAUNCH(k_rank_bitonic(), dim3(gh), dim3(kRankThreads), 0, s->stream, keys, idx, npad, size, stride);
    LAUNCH(k_rank_scatter(), dim3((unsigned)((n + kRankThreads - 1) / kRankThreads)), dim3(kRankThreads), 0, s->stream,
           (const uint32_t*)idx, n, s->rank, s->row_of_rank);
    CU_TRY(cudaStreamSynchronize(s->stream));
    return RAGB_OK;
  };
  const int rc = body();
  cudaFree(keys); cudaFree(idx);
  if (rc != RAGB_OK) return rc;
  s->ranked = true; s->rank_n = n;
  return RAGB_OK;
}

// layout of filter_aux: counts [n_ctas] | result [4] | src_allowed [65536]
static FilterParams filter_params(ragb_store* s, int64_t n, int64_t window, int n_codes, int n_ctas) {
  const size_t counts_b = ((size_t)n_ctas * 4 + 15) / 16 * 16;
  FilterParams p = {};
  p.flags = s->meta_flags; p.src = s->meta_src; p.n = n; p.window = window; p.valid = s->filter;
  p.perm = s->ranked ? s->row_of_rank : nullptr;
  p.counts = (unsigned*)s->filter_aux;
  p.result = (int64_t*)((char*)s->filter_aux + counts_b);
  p.src_allowed = n_codes ? (uint8_t*)s->filter_aux + counts_b + 32 : nullptr;
  p.n_codes = n_codes;
  return p;
}

namespace ragb {
// A search over rows [.., re) needs ranks that cover them (when timestamps are out of order).  Called with rows_mu held
// shared; upgrades (drop, take exclusive, refresh, re-take shared) only when something changed since the last refresh.
int store_ensure_ranks(ragb_store* s, std::shared_lock<std::shared_mutex>& lk, int64_t re) {
  for (int tries = 0; tries < 4; ++tries) {
    if (!s->ts || (!s->ts_dirty && (!s->ranked || s->rank_n >= re))) return RAGB_OK;
    lk.unlock();
    int rc = RAGB_OK;
    {
      std::unique_lock<std::shared_mutex> ex(s->rows_mu);
      if (s->ts_dirty || (s->ranked && s->rank_n < s->n.load())) rc = ranks_refresh(s, s->n.load());
    }
    lk.lock();
    if (rc != RAGB_OK) return rc;
  }
  return RAGB_OK;
}

int filter_refresh_ranks(ragb_store* s, int64_t n) {
  if (s->ts_dirty || (s->ranked && s->rank_n < n)) return ranks_refresh(s, n);
  return RAGB_OK;
}
// pass 1: mark + per-CTA match counts
int filter_mark(ragb_store* s, const uint8_t* src_allowed, int n_codes, int64_t n, int* n_ctas_out) {
  own_stream_waits(s);  // a search still running on a caller's stream may be reading the current mask
  RB_TRY(meta_reserve(s, n));
  const int n_ctas = (int)((n + kFilterRowsPerCta - 1) / kFilterRowsPerCta);
  const size_t counts_b = ((size_t)n_ctas * 4 + 15) / 16 * 16;
  RB_TRY(grow(&s->filter_aux, &s->filter_aux_bytes, counts_b + 32 + 65536, false));
  const FilterParams p = filter_params(s, n, -1, n_codes, n_ctas);
  if (n_codes) CU_TRY(cudaMemcpyAsync(const_cast<uint8_t*>(p.src_allowed), src_allowed, (size_t)n_codes, cudaMemcpyHostToDevice, s->stream));
  LAUNCH(k_filter_mark(), dim3(n_ctas), dim3(kFilterThreads), 0, s->stream, p);
  *n_ctas_out = n_ctas;
  return RAGB_OK;
}
int filter_counts_host(ragb_store* s, int n_ctas, std::vector<unsigned>* out) {
  out->resize((size_t)n_ctas);
  CU_TRY(cudaMemcpyAsync(out->data(), s->filter_aux, (size_t)n_ctas * 4, cudaMemcpyDeviceToHost, s->stream));
  CU_TRY(cudaStreamSynchronize(s->stream));
  return RAGB_OK;
}
// pass 3 with the cut decided by the host: CTAs before cut_cta are cleared, cut_cta keeps its last `keep` matches
int filter_window(ragb_store* s, int64_t n, int n_codes, int n_ctas, int64_t cut_cta, int64_t keep) {
  const FilterParams p = filter_params(s, n, 0, n_codes, n_ctas);
  const int64_t res[3] = {0, cut_cta, keep};
  CU_TRY(cudaMemcpyAsync(p.result, res, sizeof res, cudaMemcpyHostToDevice, s->stream));
  LAUNCH(k_filter_window(), dim3(n_ctas), dim3(kFilterThreads), 0, s->stream, p);
  CU_TRY(cudaStreamSynchronize(s->stream));
  return RAGB_OK;
}
}  // namespace ragb

extern "C" int ragb_store_build_filter(ragb_store* s, const uint8_t* src_allowed, int n_codes, int64_t window,
                                       int64_t* out_total) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  if (n_codes < 0 || n_codes > 65536 || (n_codes && !src_allowed)) return fail(RAGB_EINVAL, "bad source table");
  std::unique_lock<std::shared_mutex> ls(s->rows_mu);
  ON_DEVICE(s->device);
  const int64_t n = s->n.load(std::memory_order_acquire);
  if (out_total) *out_total = 0;
  own_stream_waits(s);  // a search still running on a caller's stream may be reading the current mask
  s->filter_n = -1;
  if (n == 0) { s->filter_n = 0; return RAGB_OK; }
  RB_TRY(filter_refresh_ranks(s, n));
  int n_ctas = 0;
  RB_TRY(filter_mark(s, src_allowed, n_codes, n, &n_ctas));
  const FilterParams p = filter_params(s, n, window, n_codes, n_ctas);
  LAUNCH(k_filter_cut(), dim3(1), dim3(1024), 0, s->stream, p, n_ctas);
  if (window >= 0) LAUNCH(k_filter_window(), dim3(n_ctas), dim3(kFilterThreads), 0, s->stream, p);
  int64_t res[3] = {0, 0, 0};
  CU_TRY(cudaMemcpyAsync(res, p.result, sizeof(res), cudaMemcpyDeviceToHost, s->stream));
  CU_TRY(cudaStreamSynchronize(s->stream));
  if (out_total) *out_total = res[0];
  s->filter_n = n;
  return RAGB_OK;
}

namespace ragb {
// valid == RAGB_STORE_FILTER: use the mask ragb_store_build_filter left in the store
int store_resolve_valid(ragb_store* s, const uint8_t** valid, int64_t row_end) {
  if (*valid != RAGB_STORE_FILTER) return RAGB_OK;
  if (s->filter_n < 0) return fail(RAGB_EINVAL, "RAGB_STORE_FILTER: no filter built (ragb_store_build_filter), or rows/metadata changed since");
  if (row_end > s->filter_n) return fail(RAGB_EINVAL, "RAGB_STORE_FILTER covers %lld rows, search asks for %lld: rebuild the filter", (long long)s->filter_n, (long long)row_end);
  *valid = s->filter;
  return RAGB_OK;
}
}  // namespace ragb

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

// the common body of the device-pointer entry points
static int search_dev_common(ragb_store* s, SearchReq& r, int64_t row_begin, int64_t row_end, void* stream) {
  std::shared_lock<std::shared_mutex> ls(s->rows_mu);
  RB_TRY(store_check_args(s, r.q, r.nq, r.k, r.metric, &row_begin, &row_end));
  r.rb = row_begin; r.re = row_end;
  ON_DEVICE(s->device);
  RB_TRY(store_ensure_ranks(s, ls, row_end));
  RB_TRY(store_resolve_valid(s, &r.valid, row_end));
  CtxLease lease(s);
  RB_TRY(lease.acquire());
  cudaStream_t st = stream ? (cudaStream_t)stream : lease.c->stream;
  RB_TRY(store_prepare_stream(s, lease.c, st, stream != nullptr));
  if (r.peer.n) r.peer.done = lease.c->tickets + kScanMaxQ;
  RB_TRY(search_core(s, lease.c, r, st));
  if (!stream) CU_TRY(cudaStreamSynchronize(st));
  store_finish_stream(s, lease.c, st, stream != nullptr);
  return RAGB_OK;
}

extern "C" int ragb_search_keys_dev(ragb_store* s, const float* q, int64_t nq, int k, int metric, float threshold,
                                    int flags, int64_t row_begin, int64_t row_end, const uint8_t* valid,
                                    uint32_t id_base, uint64_t* out_keys, void* stream) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  if (nq && !out_keys) return fail(RAGB_EINVAL, "out_keys is NULL");
  SearchReq r;
  r.q = q; r.nq = nq; r.k = k; r.metric = metric; r.threshold = threshold; r.flags = flags; r.valid = valid;
  r.id_base = id_base; r.out_keys = out_keys;
  return search_dev_common(s, r, row_begin, row_end, stream);
}

extern "C" int ragb_search_scatter_dev(ragb_store* s, const float* q, int64_t nq, int k, int metric, float threshold,
                                       int flags, int64_t row_begin, int64_t row_end, const uint8_t* valid,
                                       uint32_t id_base, const ragb_exchange* x, void* stream) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  RB_TRY(check_exchange(x));
  if (nq <= 0) return fail(RAGB_EINVAL, "peer exchange needs at least one query");
  if (k <= 0 || k > 1024) return fail(RAGB_EUNSUPPORTED, "k=%d outside [1,1024]", k);
  const int kpad = kpad_of(k);
  SearchReq r;
  r.q = q; r.nq = nq; r.k = k; r.metric = metric; r.threshold = threshold; r.flags = flags; r.valid = valid;
  r.id_base = id_base;
  r.peer.n = x->world; r.peer.epoch = x->epoch; r.peer.signal = 1 | (env_int("RAGB_PEER_GPUFENCE", 1) ? 2 : 0);
  for (int p = 0; p < x->world; ++p) {
    r.peer.keys[p] = x->keys[p] + (size_t)x->rank * nq * kpad;  // this rank's [nq][kpad] slot in rank p's buffer
    r.peer.flag[p] = x->flags[p] + x->rank;
  }
  return search_dev_common(s, r, row_begin, row_end, stream);
}

extern "C" int ragb_gather_merge_dev(int device, const ragb_exchange* x, int64_t nq, int k, float* out_scores,
                                     int64_t* out_ids, int* out_status, void* stream) {
  RB_TRY(check_exchange(x));
  if (nq <= 0 || k <= 0 || k > 1024) return fail(RAGB_EINVAL, "bad merge arguments");
  if (!out_scores || !out_ids) return fail(RAGB_EINVAL, "output buffer is NULL");
  ON_DEVICE(device);
  const int kpad = kpad_of(k);
  return merge_launch(x->keys[x->rank], nq, x->world, kpad, kpad, (int64_t)nq * kpad, k, out_scores, out_ids, nullptr, nullptr,
                      (cudaStream_t)stream, false, nullptr, x->flags[x->rank], x->world, x->epoch, out_status);
}

extern "C" int ragb_search_exchange(ragb_store* s, const float* q, int64_t nq, int k, int metric, float threshold,
                                    int flags, uint32_t id_base, const ragb_exchange* x, float* out_scores,
                                    int64_t* out_ids, int* out_status) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  RB_TRY(check_exchange(x));
  if (nq <= 0 || !q || !out_scores || !out_ids) return fail(RAGB_EINVAL, "peer exchange needs at least one query and output buffers");
  std::shared_lock<std::shared_mutex> ls(s->rows_mu);
  int64_t rb = 0, re = -1;
  RB_TRY(store_check_args(s, q, nq, k, metric, &rb, &re));
  ON_DEVICE(s->device);
  RB_TRY(store_ensure_ranks(s, ls, re));
  const int kpad = kpad_of(k);
  const size_t qb = (size_t)nq * s->dim * 4, sb = (size_t)nq * k * 4, ib = (size_t)nq * k * 8;
  auto al = [](size_t v) { return (v + 255) / 256 * 256; };
  const size_t o_s = al(qb), o_i = o_s + al(sb), o_st = o_i + al(ib), total = o_st + 256;
  CtxLease lease(s);
  RB_TRY(lease.acquire());
  SearchCtx* c = lease.c;
  RB_TRY(grow(&c->io, &c->io_bytes, total, false));
  RB_TRY(grow(&c->pin, &c->pin_bytes, total, true));
  char* d = (char*)c->io;
  char* h = (char*)c->pin;
  RB_TRY(store_prepare_stream(s, c, c->stream, false));
  *(int*)(h + o_st) = 0;
  RB_TRY(upload_host(d, h, q, qb, c->stream));
  SearchReq r;
  r.q = (const float*)d; r.nq = nq; r.k = k; r.metric = metric; r.threshold = threshold; r.flags = flags;
  r.rb = rb; r.re = re; r.id_base = id_base;
  r.peer.n = x->world; r.peer.epoch = x->epoch; r.peer.signal = 1 | (env_int("RAGB_PEER_GPUFENCE", 1) ? 2 : 0); r.peer.done = c->tickets + kScanMaxQ;
  for (int p = 0; p < x->world; ++p) {
    r.peer.keys[p] = x->keys[p] + (size_t)x->rank * nq * kpad;
    r.peer.flag[p] = x->flags[p] + x->rank;
  }
  RB_TRY(search_core(s, c, r, c->stream));
  // the merge writes scores, ids and the timeout status straight into mapped pinned memory: nothing to copy back
  RB_TRY(merge_launch(x->keys[x->rank], nq, x->world, kpad, kpad, (int64_t)nq * kpad, k, (float*)(h + o_s), (int64_t*)(h + o_i),
                      nullptr, nullptr, c->stream, false, nullptr, x->flags[x->rank], x->world, x->epoch, (int*)(h + o_st)));
  CU_TRY(cudaStreamSynchronize(c->stream));
  memcpy(out_scores, h + o_s, sb);
  memcpy(out_ids, h + o_i, ib);
  if (out_status) *out_status = *(int*)(h + o_st);
  return RAGB_OK;
}

extern "C" int ragb_search_dev(ragb_store* s, const float* q, int64_t nq, int k, int metric, float threshold, int flags,
                               int64_t row_begin, int64_t row_end, const uint8_t* valid, float* out_scores,
                               int64_t* out_ids, void* stream) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  if (nq && (!out_scores || !out_ids)) return fail(RAGB_EINVAL, "output buffer is NULL");
  SearchReq r;
  r.q = q; r.nq = nq; r.k = k; r.metric = metric; r.threshold = threshold; r.flags = flags; r.valid = valid;
  r.out_scores = out_scores; r.out_ids = out_ids;
  return search_dev_common(s, r, row_begin, row_end, stream);
}

extern "C" int ragb_search(ragb_store* s, const float* q, int64_t nq, int k, int metric, float threshold, int flags,
                           int64_t row_begin, int64_t row_end, const uint8_t* valid, float* out_scores, int64_t* out_ids) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  if (nq && (!out_scores || !out_ids)) return fail(RAGB_EINVAL, "output buffer is NULL");
  std::shared_lock<std::shared_mutex> ls(s->rows_mu);
  RB_TRY(store_check_args(s, q, nq, k, metric, &row_begin, &row_end));
  if (nq == 0) return RAGB_OK;
  ON_DEVICE(s->device);
  RB_TRY(store_ensure_ranks(s, ls, row_end));
  // per-call device buffer: q | scores | ids | valid ; pinned mirror of the first three
  const size_t qb = (size_t)nq * s->dim * 4, sb = (size_t)nq * k * 4, ib = (size_t)nq * k * 8;
  const bool store_filter = valid == RAGB_STORE_FILTER;
  RB_TRY(store_resolve_valid(s, &valid, row_end));  // the store's own mask is already on the device
  const uint8_t* valid_dev = store_filter ? valid : nullptr;
  if (store_filter) valid = nullptr;
  const size_t vb = valid ? (size_t)row_end : 0;
  auto al = [](size_t x) { return (x + 255) / 256 * 256; };
  const size_t o_s = al(qb), o_i = o_s + al(sb), o_v = o_i + al(ib);
  CtxLease lease(s);
  RB_TRY(lease.acquire());
  SearchCtx* c = lease.c;
  RB_TRY(grow(&c->io, &c->io_bytes, o_v + al(vb), false));
  RB_TRY(grow(&c->pin, &c->pin_bytes, o_v, true));
  char* d = (char*)c->io;
  char* h = (char*)c->pin;
  RB_TRY(store_prepare_stream(s, c, c->stream, false));
  // Small calls are all latency (BASELINE configs[0]: one query against 10k rows): pinned host memory is mapped into the
  // device's address space, so the kernels can read the queries from it and write the k winners straight into it —
  // no copy engine round trips on either side of the launch.  RAGB_ZERO_COPY: bit 0 = results, bit 1 = queries.
  static const int zc = env_int("RAGB_ZERO_COPY", 1);
  const bool zc_out = (zc & 1) && sb + ib <= 65536;
  const bool zc_in = (zc & 2) && qb <= 16384;
  if (zc_in) memcpy(h, q, qb);
  else RB_TRY(upload_host(d, h, q, qb, c->stream));
  if (valid) CU_TRY(cudaMemcpyAsync(d + o_v, valid, vb, cudaMemcpyHostToDevice, c->stream));
  SearchReq r;
  r.q = (const float*)(zc_in ? h : d); r.nq = nq; r.k = k; r.metric = metric; r.threshold = threshold; r.flags = flags;
  r.rb = row_begin; r.re = row_end; r.valid = valid ? (const uint8_t*)(d + o_v) : valid_dev;
  r.out_scores = (float*)((zc_out ? h : d) + o_s); r.out_ids = (int64_t*)((zc_out ? h : d) + o_i);
  RB_TRY(search_core(s, c, r, c->stream));
  if (!zc_out) CU_TRY(cudaMemcpyAsync(h + o_s, d + o_s, o_v - o_s, cudaMemcpyDeviceToHost, c->stream));
  CU_TRY(cudaStreamSynchronize(c->stream));
  memcpy(out_scores, h + o_s, sb);
  memcpy(out_ids, h + o_i, ib);
  return RAGB_OK;
}

extern "C" int ragb_store_profile(ragb_store* s, int enable) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  std::lock_guard<std::mutex> lk(s->prof_mu);
  s->profile = enable != 0;
  s->prof_used = 0;
  return RAGB_OK;
}
extern "C" int ragb_store_profile_read(ragb_store* s, double* total_ms, int64_t* n_launches) {
  if (!s) return fail(RAGB_EINVAL, "store is NULL");
  std::lock_guard<std::mutex> lk(s->prof_mu);
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
  return merge_launch(keys, nq, n_lists, kpad, stride_q, stride_p, k, out_scores, out_ids, nullptr, nullptr, (cudaStream_t)stream);
}

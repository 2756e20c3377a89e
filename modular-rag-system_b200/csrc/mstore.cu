// mstore.cu — the row-sharded store (include/ragb200.h: ragb_mstore_*): one process, one ragb_store per GPU of the
// box, rows dealt round-robin in blocks, candidates stored over NVLink into the first device's gather buffer by the
// last kernel of every shard, one merge.  Host logic only; the per-device work is the one-device store's (store.h).
#include <algorithm>
#include <cstring>
#include <memory>

#include "kernels.h"
#include "params.h"
#include "store.h"

using namespace ragb;

namespace {

// gather buffer + result staging of one sharded search in flight (lives on the first device)
struct MCtx {
  void* gather = nullptr; size_t gather_bytes = 0;   // keys [n_dev][nq][kpad]
  void* io = nullptr; size_t io_bytes = 0;           // scores | ids
  void* pin = nullptr; size_t pin_bytes = 0;         // pinned q | scores | ids
  cudaEvent_t ev[RAGB_MAX_PEERS] = {};               // shard d has stored its keys
  cudaEvent_t start = nullptr, done = nullptr;
  bool done_used = false;
};

}  // namespace

struct ragb_mstore {
  int n_dev = 0, dim = 0, dtype = 0, shift = 14;
  int dev[RAGB_MAX_PEERS] = {};
  ragb_store* st[RAGB_MAX_PEERS] = {};
  int64_t block = 16384;
  std::atomic<int64_t> n{0};
  std::mutex add_mu;                 // the dealing of rows to shards is one sequence
  std::mutex pool_mu;
  std::condition_variable pool_cv;
  std::vector<MCtx*> free_ctx, all_ctx;
  double ts_last = -1e300;
  int64_t ts_n = 0;
};

// ---- layout ---------------------------------------------------------------------------------------------------
static inline int dev_of(const ragb_mstore* m, int64_t g) { return (int)((g >> m->shift) % m->n_dev); }
static inline int64_t local_of(const ragb_mstore* m, int64_t g) {
  return (((g >> m->shift) / m->n_dev) << m->shift) + (g & (m->block - 1));
}
// number of global rows < g that live on shard d  (= the local row count of d when the store holds g rows)
static int64_t local_count(const ragb_mstore* m, int64_t g, int d) {
  const int64_t full = g >> m->shift, rem = g & (m->block - 1);
  int64_t blocks = full / m->n_dev + ((full % m->n_dev) > d ? 1 : 0);
  int64_t c = blocks << m->shift;
  if ((full % m->n_dev) == d) c += rem;
  return c;
}
// walk the global range [g0, g0+n) in pieces that live contiguously on one shard
template <typename F>
static int for_pieces(const ragb_mstore* m, int64_t g0, int64_t n, F f) {
  for (int64_t g = g0; g < g0 + n;) {
    const int64_t m_ = std::min(g0 + n - g, m->block - (g & (m->block - 1)));
    RB_TRY(f(dev_of(m, g), local_of(m, g), g - g0, m_));
    g += m_;
  }
  return RAGB_OK;
}

// ---- contexts ---------------------------------------------------------------------------------------------------
static void mctx_free(MCtx* c) {
  if (!c) return;
  cudaFree(c->gather); cudaFree(c->io);
  if (c->pin) cudaFreeHost(c->pin);
  for (auto& e : c->ev) if (e) cudaEventDestroy(e);
  if (c->start) cudaEventDestroy(c->start);
  if (c->done) cudaEventDestroy(c->done);
  delete c;
}
struct MLease {
  ragb_mstore* m; MCtx* c = nullptr;
  explicit MLease(ragb_mstore* m_) : m(m_) {}
  int acquire() {
    std::unique_lock<std::mutex> lk(m->pool_mu);
    for (;;) {
      if (!m->free_ctx.empty()) { c = m->free_ctx.back(); m->free_ctx.pop_back(); return RAGB_OK; }
      if ((int)m->all_ctx.size() < kMaxCtx) {
        MCtx* n = new MCtx();
        bool ok = true;
        for (int d = 0; d < m->n_dev && ok; ++d) {   // an event belongs to the device that records it
          DeviceGuard g(m->dev[d]);
          ok = g.ok && cudaEventCreateWithFlags(&n->ev[d], cudaEventDisableTiming) == cudaSuccess;
        }
        {
          DeviceGuard g(m->dev[0]);
          ok = ok && g.ok && cudaEventCreateWithFlags(&n->start, cudaEventDisableTiming) == cudaSuccess &&
               cudaEventCreateWithFlags(&n->done, cudaEventDisableTiming) == cudaSuccess;
        }
        if (!ok) { cudaGetLastError(); mctx_free(n); return fail(RAGB_ECUDA, "cannot create the events of a sharded search"); }
        m->all_ctx.push_back(n);
        c = n;
        return RAGB_OK;
      }
      m->pool_cv.wait(lk);
    }
  }
  ~MLease() {
    if (!c) return;
    { std::lock_guard<std::mutex> lk(m->pool_mu); m->free_ctx.push_back(c); }
    m->pool_cv.notify_one();
  }
};

// ---- life cycle ---------------------------------------------------------------------------------------------------
extern "C" int ragb_mstore_create(int dim, int dtype, int64_t capacity, const int* device_ids, int n_dev,
                                  int64_t block_rows, ragb_mstore** out) {
  if (!out) return fail(RAGB_EINVAL, "out is NULL");
  *out = nullptr;
  if (!device_ids || n_dev < 1 || n_dev > RAGB_MAX_PEERS) return fail(RAGB_EINVAL, "need 1..%d devices, got %d", RAGB_MAX_PEERS, n_dev);
  if (block_rows == 0) block_rows = 16384;
  if (block_rows < kFilterRowsPerCta || (block_rows & (block_rows - 1)))
    return fail(RAGB_EINVAL, "block_rows %lld must be a power of two >= %d", (long long)block_rows, kFilterRowsPerCta);
  if (capacity < 0 || capacity >= (int64_t)1 << 32) return fail(RAGB_EINVAL, "capacity %lld out of range [0,2^32)", (long long)capacity);
  // (a device may be listed more than once: several shards on one GPU — how the tests cover the layout on a 1-GPU box)
  ragb_mstore* m = new ragb_mstore();
  m->n_dev = n_dev; m->dim = dim; m->dtype = dtype; m->block = block_rows;
  m->shift = 0;
  while (((int64_t)1 << m->shift) < block_rows) ++m->shift;
  const int64_t blocks = (capacity + block_rows - 1) / block_rows;
  const int64_t cap_local = ((blocks + n_dev - 1) / n_dev) * block_rows;
  for (int d = 0; d < n_dev; ++d) {
    m->dev[d] = device_ids[d];
    int rc = ragb_store_create(device_ids[d], dim, dtype, cap_local, &m->st[d]);
    if (rc != RAGB_OK) { ragb_mstore_destroy(m); return rc; }
    m->st[d]->shard_shift = m->shift; m->st[d]->shard_n = n_dev; m->st[d]->shard_dev = d;
  }
  // every shard stores its candidates into the first device's gather buffer, and ingest kernels read rows that sit on
  // another device of the box: peer access both ways between all pairs
  for (int a = 0; a < n_dev; ++a) {
    DeviceGuard g(m->dev[a]);
    for (int b = 0; b < n_dev; ++b) {
      if (m->dev[a] == m->dev[b]) continue;
      int can = 0;
      cudaDeviceCanAccessPeer(&can, m->dev[a], m->dev[b]);
      if (!can) { ragb_mstore_destroy(m); return fail(RAGB_EUNSUPPORTED, "device %d cannot access device %d (no peer path)", m->dev[a], m->dev[b]); }
      cudaError_t e = cudaDeviceEnablePeerAccess(m->dev[b], 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
        ragb_mstore_destroy(m);
        return fail(RAGB_ECUDA, "cudaDeviceEnablePeerAccess(%d -> %d): %s", m->dev[a], m->dev[b], cudaGetErrorString(e));
      }
      cudaGetLastError();
    }
  }
  *out = m;
  return RAGB_OK;
}

extern "C" int ragb_mstore_destroy(ragb_mstore* m) {
  if (!m) return RAGB_OK;
  for (int d = 0; d < m->n_dev; ++d)
    if (m->st[d]) { DeviceGuard g(m->dev[d]); cudaDeviceSynchronize(); }
  {
    DeviceGuard g(m->dev[0]);
    for (MCtx* c : m->all_ctx) mctx_free(c);
  }
  for (int d = 0; d < m->n_dev; ++d) ragb_store_destroy(m->st[d]);
  delete m;
  return RAGB_OK;
}

extern "C" int ragb_mstore_reserve(ragb_mstore* m, int64_t capacity) {
  if (!m) return fail(RAGB_EINVAL, "store is NULL");
  if (capacity >= (int64_t)1 << 32) return fail(RAGB_EINVAL, "capacity %lld >= 2^32", (long long)capacity);
  std::lock_guard<std::mutex> la(m->add_mu);
  const int64_t blocks = (capacity + m->block - 1) / m->block;
  const int64_t cap_local = ((blocks + m->n_dev - 1) / m->n_dev) * m->block;
  for (int d = 0; d < m->n_dev; ++d) RB_TRY(ragb_store_reserve(m->st[d], cap_local));
  return RAGB_OK;
}
extern "C" int ragb_mstore_clear(ragb_mstore* m) {
  if (!m) return fail(RAGB_EINVAL, "store is NULL");
  std::lock_guard<std::mutex> la(m->add_mu);
  for (int d = 0; d < m->n_dev; ++d) RB_TRY(ragb_store_clear(m->st[d]));
  m->n.store(0);
  m->ts_last = -1e300; m->ts_n = 0;
  return RAGB_OK;
}
extern "C" int64_t ragb_mstore_size(const ragb_mstore* m) { return m ? m->n.load(std::memory_order_acquire) : 0; }
extern "C" int64_t ragb_mstore_capacity(const ragb_mstore* m) { return m ? ragb_store_capacity(m->st[0]) * m->n_dev : 0; }
extern "C" int64_t ragb_mstore_block_rows(const ragb_mstore* m) { return m ? m->block : 0; }
extern "C" int ragb_mstore_n_devices(const ragb_mstore* m) { return m ? m->n_dev : 0; }
extern "C" int ragb_mstore_device(const ragb_mstore* m, int i) { return (m && i >= 0 && i < m->n_dev) ? m->dev[i] : -1; }
extern "C" int64_t ragb_mstore_device_rows(const ragb_mstore* m, int i) {
  return (m && i >= 0 && i < m->n_dev) ? ragb_store_size(m->st[i]) : 0;
}

// ---- ingest ---------------------------------------------------------------------------------------------------
static int check_room(ragb_mstore* m, int64_t n) {
  const int64_t cur = m->n.load();
  if (n < 0) return fail(RAGB_EINVAL, "negative row count");
  if (cur + n > ragb_mstore_capacity(m))
    return fail(RAGB_EFULL, "store full: %lld + %lld > capacity %lld (ragb_mstore_reserve grows it)", (long long)cur,
                (long long)n, (long long)ragb_mstore_capacity(m));
  return RAGB_OK;
}
template <typename F>
static int deal(ragb_mstore* m, int64_t n, int64_t* out_first_row, F add_piece) {
  std::lock_guard<std::mutex> la(m->add_mu);
  RB_TRY(check_room(m, n));
  const int64_t g0 = m->n.load();
  RB_TRY(for_pieces(m, g0, n, [&](int d, int64_t local, int64_t off, int64_t cnt) -> int {
    int64_t first = -1;
    RB_TRY(add_piece(d, off, cnt, &first));
    if (first != local) return fail(RAGB_EINVAL, "internal: shard %d appended at local row %lld, layout says %lld", d, (long long)first, (long long)local);
    return RAGB_OK;
  }));
  m->n.store(g0 + n, std::memory_order_release);
  if (out_first_row) *out_first_row = g0;
  return RAGB_OK;
}

extern "C" int ragb_mstore_add(ragb_mstore* m, const float* rows, int64_t n, int64_t* out_first_row) {
  if (!m || (!rows && n)) return fail(RAGB_EINVAL, "NULL argument");
  return deal(m, n, out_first_row, [&](int d, int64_t off, int64_t cnt, int64_t* first) {
    return ragb_store_add(m->st[d], rows + (size_t)off * m->dim, cnt, first);
  });
}
extern "C" int ragb_mstore_add_dev(ragb_mstore* m, const void* rows, int in_dtype, int64_t n, int64_t ld_in,
                                   int64_t* out_first_row, void* stream) {
  if (!m || (!rows && n)) return fail(RAGB_EINVAL, "NULL argument");
  if (in_dtype < 0 || in_dtype > 2) return fail(RAGB_EINVAL, "bad in dtype %d", in_dtype);
  if (stream) {
    cudaPointerAttributes at;
    CU_TRY(cudaPointerGetAttributes(&at, rows));
    ON_DEVICE(at.device);
    CU_TRY(cudaStreamSynchronize((cudaStream_t)stream));
  }
  const size_t es = elem_size(in_dtype);
  return deal(m, n, out_first_row, [&](int d, int64_t off, int64_t cnt, int64_t* first) {
    return ragb_store_add_dev(m->st[d], (const char*)rows + (size_t)off * ld_in * es, in_dtype, cnt, ld_in, first, nullptr);
  });
}
extern "C" int ragb_mstore_put_rows(ragb_mstore* m, const void* packed_rows, const float* norms, int64_t n,
                                    int64_t* out_first_row) {
  if (!m || ((!packed_rows || !norms) && n)) return fail(RAGB_EINVAL, "NULL argument");
  const size_t rb = (size_t)m->dim * elem_size(m->dtype);
  return deal(m, n, out_first_row, [&](int d, int64_t off, int64_t cnt, int64_t* first) {
    return ragb_store_put_rows(m->st[d], (const char*)packed_rows + (size_t)off * rb, norms + off, cnt, first);
  });
}
extern "C" int ragb_mstore_get_rows(ragb_mstore* m, int64_t row0, int64_t n, void* out_rows, float* out_norms) {
  if (!m) return fail(RAGB_EINVAL, "store is NULL");
  if (row0 < 0 || n < 0 || row0 + n > m->n.load()) return fail(RAGB_EINVAL, "rows [%lld,%lld) out of range", (long long)row0, (long long)(row0 + n));
  const size_t rb = (size_t)m->dim * elem_size(m->dtype);
  return for_pieces(m, row0, n, [&](int d, int64_t local, int64_t off, int64_t cnt) {
    return ragb_store_get_rows(m->st[d], local, cnt, out_rows ? (char*)out_rows + (size_t)off * rb : nullptr,
                               out_norms ? out_norms + off : nullptr);
  });
}
extern "C" int ragb_mstore_set(ragb_mstore* m, int64_t row, const float* vec) {
  if (!m || !vec) return fail(RAGB_EINVAL, "NULL argument");
  if (row < 0 || row >= m->n.load()) return fail(RAGB_EINVAL, "row %lld out of range", (long long)row);
  return ragb_store_set(m->st[dev_of(m, row)], local_of(m, row), vec);
}
extern "C" int ragb_mstore_set_meta(ragb_mstore* m, int64_t row0, int64_t n, const uint8_t* flags, const uint16_t* src) {
  if (!m) return fail(RAGB_EINVAL, "store is NULL");
  if (row0 < 0 || n < 0 || row0 + n > m->n.load()) return fail(RAGB_EINVAL, "rows [%lld,%lld) out of range", (long long)row0, (long long)(row0 + n));
  return for_pieces(m, row0, n, [&](int d, int64_t local, int64_t off, int64_t cnt) {
    return ragb_store_set_meta(m->st[d], local, cnt, flags ? flags + off : nullptr, src ? src + off : nullptr);
  });
}
extern "C" int ragb_mstore_set_timestamps(ragb_mstore* m, int64_t row0, int64_t n, const double* ts) {
  if (!m || (!ts && n)) return fail(RAGB_EINVAL, "NULL argument");
  if (row0 < 0 || n < 0 || row0 + n > m->n.load()) return fail(RAGB_EINVAL, "rows [%lld,%lld) out of range", (long long)row0, (long long)(row0 + n));
  std::lock_guard<std::mutex> la(m->add_mu);
  // the sharded store orders candidates by global row: timestamps must follow row order (checked, not stored)
  if (row0 < m->ts_n) return fail(RAGB_EUNSUPPORTED, "sharded store: timestamps of rows below %lld were already given", (long long)m->ts_n);
  double last = m->ts_last;
  for (int64_t i = 0; i < n; ++i) {
    if (ts[i] < last)
      return fail(RAGB_EUNSUPPORTED, "sharded store: timestamp of row %lld is older than an earlier row's; out-of-order "
                  "timestamps need the recency ranks of a one-device store", (long long)(row0 + i));
    last = ts[i];
  }
  m->ts_last = last; m->ts_n = row0 + n;
  return RAGB_OK;
}

// ---- candidate-set policy: the window is cut over the GLOBAL row order ---------------------------------------------
extern "C" int ragb_mstore_build_filter(ragb_mstore* m, const uint8_t* src_allowed, int n_codes, int64_t window,
                                        int64_t* out_total) {
  if (!m) return fail(RAGB_EINVAL, "store is NULL");
  if (n_codes < 0 || n_codes > 65536 || (n_codes && !src_allowed)) return fail(RAGB_EINVAL, "bad source table");
  if (out_total) *out_total = 0;
  const int64_t n = m->n.load(std::memory_order_acquire);
  std::vector<std::unique_lock<std::shared_mutex>> locks;
  for (int d = 0; d < m->n_dev; ++d) locks.emplace_back(m->st[d]->rows_mu);
  struct Cta { int64_t gstart; int d, j; unsigned cnt; };
  std::vector<Cta> ctas;
  std::vector<int> n_ctas(m->n_dev, 0);
  std::vector<int64_t> n_loc(m->n_dev, 0);
  int64_t total = 0;
  for (int d = 0; d < m->n_dev; ++d) {
    ragb_store* s = m->st[d];
    s->filter_n = -1;
    n_loc[d] = local_count(m, n, d);
    if (n_loc[d] == 0) { s->filter_n = 0; continue; }
    ON_DEVICE(m->dev[d]);
    RB_TRY(filter_mark(s, src_allowed, n_codes, n_loc[d], &n_ctas[d]));
    std::vector<unsigned> cnt;
    RB_TRY(filter_counts_host(s, n_ctas[d], &cnt));
    for (int j = 0; j < n_ctas[d]; ++j) {
      const int64_t l0 = (int64_t)j * kFilterRowsPerCta;  // a filter CTA never straddles a block (block % 2048 == 0)
      IdMap im; im.rank = nullptr; im.base = 0; im.shift = m->shift; im.n_dev = m->n_dev; im.dev = d;
      ctas.push_back({(int64_t)global_row(im, (uint32_t)l0), d, j, cnt[j]});
      total += cnt[j];
    }
  }
  if (window >= 0 && total > window) {
    std::sort(ctas.begin(), ctas.end(), [](const Cta& a, const Cta& b) { return a.gstart > b.gstart; });  // newest first
    int64_t acc = 0;
    size_t cut = 0;
    while (cut < ctas.size() && acc + (int64_t)ctas[cut].cnt < window) acc += ctas[cut++].cnt;
    const Cta& c = ctas[cut];   // the window starts inside this CTA: it keeps its last (window - acc) matches
    for (int d = 0; d < m->n_dev; ++d) {
      if (n_loc[d] == 0) continue;
      int64_t cut_cta, keep;
      if (d == c.d) { cut_cta = c.j; keep = window - acc; }
      else {
        // this shard's first CTA newer than the cut keeps everything; older ones are cleared
        int first_newer = n_ctas[d];
        for (const Cta& x : ctas) if (x.d == d && x.gstart > c.gstart && x.j < first_newer) first_newer = x.j;
        cut_cta = first_newer; keep = kFilterRowsPerCta;
        if (first_newer == 0) continue;   // nothing older than the cut on this shard
      }
      ON_DEVICE(m->dev[d]);
      RB_TRY(filter_window(m->st[d], n_loc[d], n_codes, n_ctas[d], cut_cta, keep));
    }
    total = window;
  } else {
    for (int d = 0; d < m->n_dev; ++d)
      if (n_loc[d]) { ON_DEVICE(m->dev[d]); CU_TRY(cudaStreamSynchronize(m->st[d]->stream)); }
  }
  for (int d = 0; d < m->n_dev; ++d) m->st[d]->filter_n = n_loc[d];
  if (out_total) *out_total = total;
  return RAGB_OK;
}

// ---- search ---------------------------------------------------------------------------------------------------
// q_host != null: host entry (queries uploaded to every shard from pinned memory, results copied back);
// q_dev  != null: queries / outputs on the first device, `stream` there (null = synchronous).
static int msearch(ragb_mstore* m, const float* q_host, const float* q_dev, int64_t nq, int k, int metric, float threshold,
                   int flags, const uint8_t* valid, float* out_scores, int64_t* out_ids, void* stream) {
  if (!m) return fail(RAGB_EINVAL, "store is NULL");
  if (nq < 0 || nq > 65535 * 16) return fail(RAGB_EINVAL, "nq %lld out of range", (long long)nq);
  if (nq && (!out_scores || !out_ids || (!q_host && !q_dev))) return fail(RAGB_EINVAL, "NULL buffer");
  if (k <= 0 || k > 1024) return fail(RAGB_EUNSUPPORTED, "k=%d outside [1,1024]", k);
  if (metric != RAGB_IP && metric != RAGB_COSINE) return fail(RAGB_EINVAL, "bad metric %d", metric);
  if (nq == 0) return RAGB_OK;
  const bool use_filter = valid == RAGB_STORE_FILTER;
  if (valid && !use_filter && !q_host) return fail(RAGB_EUNSUPPORTED, "sharded store: a mask over global rows is a host-entry argument");
  const int G = m->n_dev, kpad = kpad_of(k);
  const int64_t n = m->n.load(std::memory_order_acquire);
  const size_t qb = (size_t)nq * m->dim * 4, sb = (size_t)nq * k * 4, ib = (size_t)nq * k * 8;
  auto al = [](size_t x) { return (x + 255) / 256 * 256; };
  std::vector<std::shared_lock<std::shared_mutex>> locks;
  for (int d = 0; d < G; ++d) locks.emplace_back(m->st[d]->rows_mu);
  MLease ml(m);
  RB_TRY(ml.acquire());
  MCtx* mc = ml.c;
  std::vector<std::unique_ptr<CtxLease>> leases;
  for (int d = 0; d < G; ++d) { leases.emplace_back(new CtxLease(m->st[d])); RB_TRY(leases[d]->acquire()); }
  const bool caller = stream != nullptr;
  cudaStream_t st0 = caller ? (cudaStream_t)stream : leases[0]->c->stream;
  const void* q_src = nullptr;
  {
    ON_DEVICE(m->dev[0]);
    RB_TRY(grow(&mc->gather, &mc->gather_bytes, (size_t)G * nq * kpad * 8, false));
    if (mc->done_used) { CU_TRY(cudaStreamWaitEvent(st0, mc->done, 0)); mc->done_used = false; }
    if (q_host) {
      RB_TRY(grow(&mc->io, &mc->io_bytes, al(sb) + al(ib), false));
      RB_TRY(grow(&mc->pin, &mc->pin_bytes, al(qb) + al(sb) + al(ib), true));
      // a page-locked source is uploaded to every shard in place; a pageable one once through the pinned buffer
      cudaPointerAttributes pa;
      if (cudaPointerGetAttributes(&pa, q_host) == cudaSuccess && pa.type == cudaMemoryTypeHost) q_src = q_host;
      else { (void)cudaGetLastError(); memcpy(mc->pin, q_host, qb); q_src = mc->pin; }
    } else {
      CU_TRY(cudaEventRecord(mc->start, st0));  // the queries are ready on the first device from here on
    }
  }
  for (int d = 0; d < G; ++d) {
    ragb_store* s = m->st[d];
    SearchCtx* c = leases[d]->c;
    ON_DEVICE(m->dev[d]);
    const bool on_caller = caller && d == 0;
    cudaStream_t st = d == 0 ? st0 : c->stream;
    RB_TRY(store_prepare_stream(s, c, st, on_caller));
    const int64_t re = local_count(m, n, d);
    SearchReq r;
    r.nq = nq; r.k = k; r.metric = metric; r.threshold = threshold; r.flags = flags; r.rb = 0; r.re = re;
    r.out_keys = (uint64_t*)mc->gather + (size_t)d * nq * kpad;   // peer-mapped: the finish kernel stores over NVLink
    if (q_host) {
      const size_t vb = (valid && !use_filter) ? (size_t)re : 0;
      RB_TRY(grow(&c->io, &c->io_bytes, al(qb) + al(vb), false));
      CU_TRY(cudaMemcpyAsync(c->io, q_src, qb, cudaMemcpyHostToDevice, st));
      r.q = (const float*)c->io;
      if (vb) {  // de-interleave the global mask into this shard's local rows (block pieces)
        RB_TRY(grow(&c->pin, &c->pin_bytes, vb, true));
        uint8_t* lp = (uint8_t*)c->pin;
        for (int64_t l = 0; l < re; l += m->block) {
          IdMap im; im.rank = nullptr; im.base = 0; im.shift = m->shift; im.n_dev = G; im.dev = d;
          memcpy(lp + l, valid + global_row(im, (uint32_t)l), (size_t)std::min(m->block, re - l));
        }
        CU_TRY(cudaMemcpyAsync((char*)c->io + al(qb), lp, vb, cudaMemcpyHostToDevice, st));
        r.valid = (const uint8_t*)c->io + al(qb);
      }
    } else if (d == 0) {
      r.q = q_dev;
    } else {
      RB_TRY(grow(&c->io, &c->io_bytes, al(qb), false));
      CU_TRY(cudaStreamWaitEvent(st, mc->start, 0));
      CU_TRY(cudaMemcpyPeerAsync(c->io, m->dev[d], q_dev, m->dev[0], qb, st));
      r.q = (const float*)c->io;
    }
    if (use_filter) { const uint8_t* v = RAGB_STORE_FILTER; RB_TRY(store_resolve_valid(s, &v, re)); r.valid = v; }
    RB_TRY(search_core(s, c, r, st));
    if (d > 0) {
      CU_TRY(cudaEventRecord(mc->ev[d], st));
      if (caller) store_finish_stream(s, c, st, true);  // asynchronous: the next user of this context waits for it
    }
  }
  {
    ON_DEVICE(m->dev[0]);
    for (int d = 1; d < G; ++d) CU_TRY(cudaStreamWaitEvent(st0, mc->ev[d], 0));
    float* d_sc = q_host ? (float*)mc->io : out_scores;
    int64_t* d_id = q_host ? (int64_t*)((char*)mc->io + al(sb)) : out_ids;
    RB_TRY(merge_launch((const uint64_t*)mc->gather, nq, G, kpad, kpad, (int64_t)nq * kpad, k, d_sc, d_id, nullptr, nullptr, st0));
    if (q_host) {
      char* h = (char*)mc->pin + al(qb);
      CU_TRY(cudaMemcpyAsync(h, mc->io, al(sb) + ib, cudaMemcpyDeviceToHost, st0));
      CU_TRY(cudaStreamSynchronize(st0));
      memcpy(out_scores, h, sb);
      memcpy(out_ids, h + al(sb), ib);
    } else if (!caller) {
      CU_TRY(cudaStreamSynchronize(st0));
    } else {
      store_finish_stream(m->st[0], leases[0]->c, st0, true);
      if (cudaEventRecord(mc->done, st0) == cudaSuccess) mc->done_used = true;
    }
  }
  return RAGB_OK;
}

extern "C" int ragb_mstore_search(ragb_mstore* m, const float* q, int64_t nq, int k, int metric, float threshold, int flags,
                                  const uint8_t* valid, float* out_scores, int64_t* out_ids) {
  return msearch(m, q, nullptr, nq, k, metric, threshold, flags, valid, out_scores, out_ids, nullptr);
}
extern "C" int ragb_mstore_search_dev(ragb_mstore* m, const float* q, int64_t nq, int k, int metric, float threshold,
                                      int flags, const uint8_t* valid, float* out_scores, int64_t* out_ids, void* stream) {
  return msearch(m, nullptr, q, nq, k, metric, threshold, flags, valid, out_scores, out_ids, stream);
}

extern "C" int ragb_mstore_profile(ragb_mstore* m, int enable) {
  if (!m) return fail(RAGB_EINVAL, "store is NULL");
  for (int d = 0; d < m->n_dev; ++d) RB_TRY(ragb_store_profile(m->st[d], enable));
  return RAGB_OK;
}
extern "C" int ragb_mstore_profile_read(ragb_mstore* m, double* total_ms, int64_t* n_launches) {
  if (!m) return fail(RAGB_EINVAL, "store is NULL");
  double worst = 0;
  int64_t n0 = 0;
  for (int d = 0; d < m->n_dev; ++d) {
    double ms = 0; int64_t nl = 0;
    RB_TRY(ragb_store_profile_read(m->st[d], &ms, &nl));
    worst = std::max(worst, ms);
    if (d == 0) n0 = nl;
  }
  if (total_ms) *total_ms = worst;
  if (n_launches) *n_launches = n0;
  return RAGB_OK;
}

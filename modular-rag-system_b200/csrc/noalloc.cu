// noalloc.cu — ragb_scan_topk_static_dev (include/ragb200.h): the scan search (K3, scan.cuh) over caller-held packed
// rows for callers that may not allocate — CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
//
// The scan kernel needs scratch memory: one candidate list of kpad keys per CTA and query (read back by the last CTA of
// each query, scan.cuh fused_final_merge) and one ticket word per query.  The store entry points keep that in a
// workspace grown with cudaMalloc (store.h SearchCtx::ws / tickets).  Here it is a file-scope __device__ array of a
// fixed, modest size that every GPU which loads the library holds until the process exits.  No counterpart in the
// reference.
//
// `g_static_scratch` below is, in HOST code, only a stand-in for the symbol — not a device address.  The address a
// kernel can use comes from cudaGetSymbolAddress, asked once per device with that device current (static_setup).
#include <algorithm>
#include <mutex>

#include "kernels.h"
#include "params.h"
#include "store.h"

using namespace ragb;

namespace {

constexpr int kStaticMaxLists = 296;   // CTAs of one scan launch: 2 per SM x 148 SMs (a larger GPU is capped here)
constexpr int kStaticMaxDevices = 64;
constexpr size_t kStaticListKeys = (size_t)kStaticMaxLists * RAGB_STATIC_SCAN_MAX_KEYS;   // [nq][P][kpad] keys
constexpr size_t kStaticTicketWords = (RAGB_STATIC_SCAN_MAX_NQ * sizeof(unsigned) + 7) / 8;
constexpr size_t kStaticWords = kStaticListKeys + kStaticTicketWords;                      // 2 425 088 B

}  // namespace

// lists | tickets.  Shared by every call on this device; keeps the last call's contents.
static __device__ __align__(256) unsigned long long g_static_scratch[kStaticWords];

namespace {

struct StaticDev {
  unsigned long long* scratch = nullptr;  // device address of g_static_scratch ON THIS DEVICE
  int n_sm = 0;
};
std::mutex g_static_mu;
StaticDev g_static_dev[kStaticMaxDevices];

// once per device, with that device current
int static_setup(int device, StaticDev* out) {
  std::lock_guard<std::mutex> lk(g_static_mu);
  StaticDev& d = g_static_dev[device];
  if (!d.scratch) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      return fail(RAGB_ECUDA, "cudaGetSymbolAddress(static scan scratch) on device %d: %s", device, cudaGetErrorString(e));
    }
    int sm = 0;
    e = cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess || sm < 1) {
      cudaGetLastError();
      return fail(RAGB_ECUDA, "cudaDeviceGetAttribute(SM count) on device %d: %s", device, cudaGetErrorString(e));
    }
    d.n_sm = sm;
    d.scratch = static_cast<unsigned long long*>(p);
  }
  *out = d;
  return RAGB_OK;
}

bool static_cpl_ok(int c) { return c == 1 || c == 2 || c == 3 || c == 4 || c == 6 || c == 8; }

}  // namespace

extern "C" int ragb_scan_topk_static_dev(int device, const void* rows, int dtype, int64_t n, int dim, int64_t ld,
                                         const float* q, int64_t nq, int k, int metric, float threshold, int flags,
                                         const uint8_t* valid, float* out_scores, int64_t* out_ids, void* stream) {
  if (device < 0 || device >= kStaticMaxDevices) return fail(RAGB_EINVAL, "device %d outside [0,%d)", device, kStaticMaxDevices);
  if (dtype != RAGB_F16 && dtype != RAGB_BF16 && dtype != RAGB_F32) return fail(RAGB_EINVAL, "bad dtype %d", dtype);
  if (n < 0 || n >= (int64_t)1 << 32) return fail(RAGB_EINVAL, "row count %lld outside [0,2^32)", (long long)n);
  if (dim < 1 || ld < dim) return fail(RAGB_EINVAL, "need 1 <= dim <= ld, got dim %d ld %lld", dim, (long long)ld);
  const size_t es = elem_size(dtype);
  if ((size_t)ld * es % 16 != 0 || (uintptr_t)rows % 16 != 0)
    return fail(RAGB_EINVAL, "rows must be 16-byte aligned with a row stride that is a multiple of 16 bytes (ld %lld x %zu B)", (long long)ld, es);
  if (metric != RAGB_IP && metric != RAGB_COSINE) return fail(RAGB_EINVAL, "bad metric %d", metric);
  if (k <= 0 || k > 1024) return fail(RAGB_EUNSUPPORTED, "k=%d outside [1,1024]", k);
  if (nq < 0) return fail(RAGB_EINVAL, "negative query count");
  const int kpad = kpad_of(k);
  // the static scratch bounds one call: say so instead of overrunning it
  if (nq > RAGB_STATIC_SCAN_MAX_NQ)
    return fail(RAGB_EUNSUPPORTED, "ragb_scan_topk_static_dev: %lld queries exceed the static scratch limit of %d per call "
                "(split the batch, or use ragb_search_dev which allocates its workspace)", (long long)nq, RAGB_STATIC_SCAN_MAX_NQ);
  if (nq * kpad > RAGB_STATIC_SCAN_MAX_KEYS)
    return fail(RAGB_EUNSUPPORTED, "ragb_scan_topk_static_dev: nq %lld x ragb_kpad(k=%d) %d = %lld keys per CTA exceed the static "
                "scratch limit of %d (fewer queries per call, or a smaller k)", (long long)nq, k, kpad,
                (long long)(nq * kpad), RAGB_STATIC_SCAN_MAX_KEYS);
  if (nq && (!q || !out_scores || !out_ids)) return fail(RAGB_EINVAL, "NULL buffer");
  if (n && !rows) return fail(RAGB_EINVAL, "rows is NULL");
  ON_DEVICE(device);
  StaticDev dev;
  RB_TRY(static_setup(device, &dev));
  if (nq == 0) return RAGB_OK;   // set-up only (first use on a device: before a capture starts)
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0)  // nothing to scan: empty slots (id -1, score -inf), as the store entry points answer an empty window
    return merge_launch(nullptr, nq, 0, kpad, 0, 0, k, out_scores, out_ids, nullptr, nullptr, st);

  // kernel variant: as the store's scan dispatch (ragb200.cu scan_launch) — fewest lanes per row
  const int chunks = (int)((size_t)ld * es / 16);
  int G = 0;
  for (int g = 8; g <= 32; g <<= 1)
    if (chunks % g == 0 && static_cpl_ok(chunks / g)) { G = g; break; }
  int R = 0;
  const void* kern = G ? k_scan(dtype, G, chunks / G, &R) : nullptr;
  const bool generic = kern == nullptr;
  if (generic) kern = k_scan_generic(dtype, &R);
  if (!kern) return fail(RAGB_EINVAL, "bad dtype %d", dtype);
  const int64_t units = (n + R - 1) / R;
  const int cap = std::min(kStaticMaxLists, 2 * dev.n_sm);
  int P = (int)std::min<int64_t>((units + kScanWarps * 4 - 1) / (kScanWarps * 4), cap);
  if (P < 1) P = 1;
  const size_t smem = (size_t)kScanWarps * listbuf_len(kpad) * sizeof(uint64_t) + (generic ? (size_t)ld * sizeof(float) : 0);
  if (smem > 220 * 1024) return fail(RAGB_EUNSUPPORTED, "k=%d with ld=%lld needs %zu B of shared memory", k, (long long)ld, smem);
  if (smem > 48 * 1024) CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));

  unsigned* tickets = reinterpret_cast<unsigned*>(dev.scratch + kStaticListKeys);
  // The tickets must start at zero.  A finished launch leaves them at zero, but the array is shared and keeps whatever
  // the last call — possibly a failed or an overlapping one — left behind: clear them on the stream, every call.
  // (The lists need no clearing: every CTA writes all kpad slots of its list before any of them is read.)
  CU_TRY(cudaMemsetAsync(tickets, 0, (size_t)nq * sizeof(unsigned), st));

  ScanParams p = {};
  p.X = rows; p.ld = ld; p.dim = dim; p.row_begin = 0; p.row_end = n; p.valid = valid;
  p.q = q; p.ld_q = dim; p.metric = metric; p.flags = flags & RAGB_CLAMP01; p.threshold = threshold;
  p.k = k; p.kpad = kpad;
  p.ids = IdMap{nullptr, 0u, -1, 1, 0};   // id = row
  p.row_of_id = nullptr;
  p.out = reinterpret_cast<uint64_t*>(dev.scratch);
  p.tickets = tickets;
  p.out_scores = out_scores; p.out_ids = out_ids; p.out_keys = nullptr;
  p.peers = PeerOut{};
  cudaError_t e = launch_k(kern, dim3((unsigned)P, (unsigned)nq), dim3(kScanThreads), smem, st, p);
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) return fail(RAGB_ECUDA, "scan launch (static scratch): %s", cudaGetErrorString(e));
  return RAGB_OK;
}

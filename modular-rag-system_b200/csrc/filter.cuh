// filter.cuh — the reference's candidate-set policy evaluated on the device.
//
// RAGVectorStore._search_sqlite chooses the rows it scores with SQL (ollama_rag_system.py:462-477):
//   SELECT * FROM documents [WHERE source IN (...)] ORDER BY timestamp DESC LIMIT 1000
// and skips rows whose embedding is NULL (:485-492); INSERT OR REPLACE (:351) retires the old row of
// an id.  The store keeps two per-row arrays next to the packed rows — flags (bit 0 live, bit 1 has an
// embedding) and a 16-bit source code — and these kernels turn (allowed sources, window) into the
// per-row `valid` bytes the search kernels read, without a per-query host pass over all rows and
// without shipping a mask over PCIe.  The window is "the newest W matching rows" = the LAST W matches
// in RECENCY order: position p of that order holds row perm[p] (rank.cuh: rows sorted by (timestamp,
// row)); while timestamps do not decrease in row order perm is null and position == row.
#pragma once
#include "common.cuh"
#include "params.h"

namespace ragb {



__device__ __forceinline__ int64_t filter_row(const FilterParams& p, int64_t pos) { return p.perm ? (int64_t)p.perm[pos] : pos; }
__device__ __forceinline__ bool filter_match(const FilterParams& p, int64_t r) {
  if (!(p.flags[r] & 1)) return false;
  if (!p.src_allowed) return true;
  const int c = p.src[r];
  return c < p.n_codes && p.src_allowed[c] != 0;
}

// pass 1: valid = match && has-embedding; per-CTA match counts
__global__ void __launch_bounds__(kFilterThreads) filter_mark_kernel(const FilterParams p) {
  __shared__ unsigned warp_cnt[kFilterThreads / 32];
  const int64_t base = (int64_t)blockIdx.x * kFilterRowsPerCta;
  unsigned cnt = 0;
  for (int i = threadIdx.x; i < kFilterRowsPerCta; i += kFilterThreads) {
    if (base + i < p.n) {
      const int64_t r = filter_row(p, base + i);
      const bool m = filter_match(p, r);
      cnt += m ? 1u : 0u;
      p.valid[r] = (m && (p.flags[r] & 2)) ? 1 : 0;
    }
  }
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  if ((threadIdx.x & 31) == 0) warp_cnt[threadIdx.x >> 5] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned t = 0;
    for (int w = 0; w < kFilterThreads / 32; ++w) t += warp_cnt[w];
    p.counts[blockIdx.x] = t;
  }
}

// pass 2 (one CTA): total matches; if they exceed the window, the CTA in which the window starts and how
// many of its matches (counted from its end) belong to the window
__global__ void __launch_bounds__(1024) filter_cut_kernel(const FilterParams p, int n_ctas) {
  __shared__ unsigned long long part[1024];
  const int t = threadIdx.x;
  const int per = (n_ctas + 1023) / 1024;
  // thread t owns CTAs [hi - per, hi) counted from the END: t = 0 holds the newest rows
  const int hi = n_ctas - t * per, lo = max(hi - per, 0);
  unsigned long long s = 0;
  for (int b = hi - 1; b >= lo; --b) s += p.counts[b];
  part[t] = s;
  __syncthreads();
  if (t == 0) {
    unsigned long long total = 0;
    for (int i = 0; i < 1024; ++i) total += part[i];
    p.result[0] = (int64_t)total;
    p.result[1] = -1;
    p.result[2] = 0;
    if (p.window >= 0 && total > (unsigned long long)p.window) {
      p.result[0] = p.window;
      unsigned long long need = (unsigned long long)p.window, acc = 0;
      int owner = 0;
      while (acc + part[owner] < need) { acc += part[owner]; ++owner; }
      const int ohi = n_ctas - owner * per, olo = max(ohi - per, 0);
      int b = ohi - 1;
      for (; b >= olo; --b) {
        if (acc + p.counts[b] >= need) break;
        acc += p.counts[b];
      }
      p.result[1] = b;
      p.result[2] = (int64_t)(need - acc);
    }
  }
}

// pass 3: rows older than the window are cleared; inside the cut CTA only its last result[2] matches stay
// (result[1], result[2] come from pass 2 — or from the host, which merges the per-CTA counts of several devices when
// the rows are sharded, mstore.cu)
__global__ void __launch_bounds__(kFilterThreads) filter_window_kernel(const FilterParams p) {
  const int64_t cut = p.result[1];
  if (cut < 0 || (int64_t)blockIdx.x > cut) return;
  const int64_t base = (int64_t)blockIdx.x * kFilterRowsPerCta;
  if ((int64_t)blockIdx.x < cut) {
    for (int i = threadIdx.x; i < kFilterRowsPerCta; i += kFilterThreads)
      if (base + i < p.n) p.valid[filter_row(p, base + i)] = 0;
    return;
  }
  // the cut CTA: thread t owns rows [t*8, t*8+8); suffix counts of matches decide who stays
  __shared__ unsigned tcnt[kFilterThreads];
  constexpr int kPer = kFilterRowsPerCta / kFilterThreads;
  bool m[kPer];
  unsigned c = 0;
#pragma unroll
  for (int j = 0; j < kPer; ++j) {
    const int64_t pos = base + threadIdx.x * kPer + j;
    m[j] = pos < p.n && filter_match(p, filter_row(p, pos));
    c += m[j] ? 1u : 0u;
  }
  tcnt[threadIdx.x] = c;
  __syncthreads();
  unsigned after = 0;  // matches in rows owned by later threads
  for (int t = threadIdx.x + 1; t < kFilterThreads; ++t) after += tcnt[t];
  const unsigned keep = (unsigned)min(p.result[2], (int64_t)kFilterRowsPerCta);
#pragma unroll
  for (int j = kPer - 1; j >= 0; --j) {
    const int64_t pos = base + threadIdx.x * kPer + j;
    if (m[j]) {
      ++after;                        // this row is the `after`-th match from the end of the CTA
      if (after > keep) p.valid[filter_row(p, pos)] = 0;
    } else if (pos < p.n) {
      p.valid[filter_row(p, pos)] = 0;
    }
  }
}

}  // namespace ragb

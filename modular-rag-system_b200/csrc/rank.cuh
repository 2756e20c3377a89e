// rank.cuh — recency ranks for stores whose timestamps do not follow row order.
//
// The reference fetches candidates "ORDER BY timestamp DESC LIMIT 1000" and then sorts them stably by similarity
// (ollama_rag_system.py:474, :512): among equal scores the NEWER timestamp comes first, and the window keeps the newest
// rows by timestamp.  While timestamps do not decrease in row order, "newer" == "larger row" and nothing is needed.
// Otherwise the store keeps, next to the rows,
//     rank[row]        position of the row in ascending (timestamp, row) order  -> low word of every candidate key
//     row_of_rank[pos] the inverse                                             -> decode of the winners, filter order
// built here on the device: order-preserving 64-bit timestamp keys, a plain global bitonic sort of (key, row) pairs
// (a maintenance pass, run when a timestamp arrives out of order — log2(n)^2/2 sweeps over 12 bytes per row: about
// 25 ms at 10M rows), then a scatter.
#pragma once
#include "common.cuh"
#include "params.h"

namespace ragb {


__device__ __forceinline__ uint64_t order_bits64(double t) {
  const uint64_t b = (uint64_t)__double_as_longlong(t);
  return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}

// flag[0] = 1 if ts[i] < ts[i-1] for some i in [1, n)
__global__ void __launch_bounds__(kRankThreads) ts_check_kernel(const double* __restrict__ ts, int64_t n, int* __restrict__ flag) {
  const int64_t i = (int64_t)blockIdx.x * kRankThreads + threadIdx.x + 1;
  if (i < n && ts[i] < ts[i - 1]) *flag = 1;
}
__global__ void __launch_bounds__(kRankThreads) ts_fill_kernel(double* __restrict__ ts, int64_t r0, int64_t r1, double v) {
  const int64_t i = r0 + (int64_t)blockIdx.x * kRankThreads + threadIdx.x;
  if (i < r1) ts[i] = v;
}
__global__ void __launch_bounds__(kRankThreads) rank_init_kernel(const double* __restrict__ ts, int64_t n, int64_t npad,
                                                                 uint64_t* __restrict__ keys, uint32_t* __restrict__ idx) {
  const int64_t i = (int64_t)blockIdx.x * kRankThreads + threadIdx.x;
  if (i >= npad) return;
  keys[i] = i < n ? order_bits64(ts[i]) : ~0ull;   // padding sorts behind every real row
  idx[i] = (uint32_t)i;
}
// one compare-exchange sweep of the bitonic network, ascending in (key, idx)
__global__ void __launch_bounds__(kRankThreads) rank_bitonic_kernel(uint64_t* __restrict__ keys, uint32_t* __restrict__ idx,
                                                                    int64_t npad, int64_t size, int64_t stride) {
  const int64_t t = (int64_t)blockIdx.x * kRankThreads + threadIdx.x;
  if (t >= (npad >> 1)) return;
  const int64_t lo = 2 * t - (t & (stride - 1)), hi = lo + stride;
  const bool up = (lo & size) == 0;
  const uint64_t ka = keys[lo], kb = keys[hi];
  const uint32_t ia = idx[lo], ib = idx[hi];
  const bool a_gt_b = ka > kb || (ka == kb && ia > ib);
  if (a_gt_b == up) { keys[lo] = kb; keys[hi] = ka; idx[lo] = ib; idx[hi] = ia; }
}
__global__ void __launch_bounds__(kRankThreads) rank_scatter_kernel(const uint32_t* __restrict__ idx, int64_t n,
                                                                    uint32_t* __restrict__ rank, uint32_t* __restrict__ row_of_rank) {
  const int64_t p = (int64_t)blockIdx.x * kRankThreads + threadIdx.x;
  if (p >= n) return;
  const uint32_t r = idx[p];   // the n real rows sort before the padding
  rank[r] = (uint32_t)p;
  row_of_rank[p] = r;
}

}  // namespace ragb

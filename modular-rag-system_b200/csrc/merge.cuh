// merge.cuh — K5: fold sorted candidate lists (per CTA / per tile / per GPU) into the final k.
//
// The reference's final "sorted(...)[:max_results]" (ollama_rag_system.py:511-515).
// One CTA per query; every warp folds a strided subset of the lists with bitonic
// merges, the warps are tree-merged, warp 0 decodes keys to (score, id).
#pragma once
#include "common.cuh"

namespace ragb {

constexpr int kMergeThreads = 256;
constexpr int kMergeWarps = kMergeThreads / 32;

struct MergeParams {
  const uint64_t* keys;
  int n_lists, kpad, k;
  int unsorted;         // lists are unordered key bags (0 = empty) instead of sorted runs
  int64_t stride_q, stride_p;
  float* out_scores;    // [nq][k] or null
  int64_t* out_ids;     // [nq][k] or null
  uint64_t* out_keys;   // [nq][kpad] or null (sorted keys, for a later merge level)
};

__global__ void __launch_bounds__(kMergeThreads) merge_topk_kernel(const MergeParams p) {
  extern __shared__ __align__(16) uint64_t smem_lists[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t qi = blockIdx.x;
  WarpTopK tk;
  tk.init(smem_lists + (size_t)warp * listbuf_len(p.kpad), p.k, p.kpad, lane);
  const uint64_t* base = p.keys + qi * p.stride_q;
  for (int l = warp; l < p.n_lists; l += kMergeWarps) {
    const uint64_t* list = base + (int64_t)l * p.stride_p;
    if (p.unsorted) {
      for (int i0 = 0; i0 < p.kpad; i0 += 32) {
        const uint64_t key = (i0 + lane < p.kpad) ? list[i0 + lane] : 0;
        tk.push(key > tk.thr, key, lane);
      }
    } else {
      tk.merge_list(list, lane);
    }
  }
  tk.flush(lane);
  __syncthreads();
  for (int s = 1; s < kMergeWarps; s <<= 1) {
    if ((warp & (2 * s - 1)) == 0 && warp + s < kMergeWarps)
      tk.merge_list(smem_lists + (size_t)(warp + s) * tk.len, lane);
    __syncthreads();
  }
  if (warp == 0) {
    for (int i = lane; i < p.k; i += 32) {
      uint64_t key = tk.buf[i];
      if (p.out_scores) p.out_scores[qi * p.k + i] = key ? key_score(key) : -INFINITY;
      if (p.out_ids) p.out_ids[qi * p.k + i] = key ? (int64_t)key_id(key) : -1;
    }
    if (p.out_keys)
      for (int i = lane; i < p.kpad; i += 32) p.out_keys[qi * p.kpad + i] = i < p.k ? tk.buf[i] : 0;
  }
}

}  // namespace ragb

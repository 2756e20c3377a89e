// merge.cuh — K5: fold sorted candidate lists (per CTA / per tile / per GPU) into the final k.
//
// The reference's final "sorted(...)[:max_results]" (ollama_rag_system.py:511-515).
// One CTA per query; every warp folds a strided subset of the lists with bitonic
// merges, the warps are tree-merged, warp 0 decodes keys to (score, id).
#pragma once
#include "common.cuh"
#include "params.h"

namespace ragb {



__global__ void __launch_bounds__(kMergeThreads) merge_topk_kernel(const MergeParams p) {
  extern __shared__ __align__(16) uint64_t smem_lists[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t qi = blockIdx.x;
  if (p.wait_n > 0) {  // candidate lists arrive from the other GPUs through peer stores
    if ((int)threadIdx.x < p.wait_n && !peer_wait(p.wait_flags + threadIdx.x, p.wait_epoch) && p.status) atomicExch(p.status, 1);
    __syncthreads();
  }
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
      if (p.out_ids) p.out_ids[qi * p.k + i] = key ? decode_id(p.row_of_id, key_id(key)) : -1;
    }
    if (p.out_keys)
      for (int i = lane; i < p.kpad; i += 32) p.out_keys[qi * p.kpad + i] = i < p.k ? tk.buf[i] : 0;
    if (p.peers.n)
      for (int i = lane; i < p.kpad; i += 32) peer_emit(p.peers, qi * p.kpad + i, i < p.k ? tk.buf[i] : 0);
  }
  if (p.peers.n) peer_signal(p.peers, gridDim.x);
}

// ---------------------------------------------------------------------------------------------
// K5' finish of the SM-pair batched kernel (gemm_pair.cuh), ONE launch, one CTA per query:
//   1. fold the per-split lists of CHUNK keys (score = chunk maximum from the 16-bit-query tensor pass,
//      id = local start row of the 16-row chunk) into the best k_sel chunks;
//   2. score the 16 rows of each selected chunk exactly: fp32 query, fp32 accumulate, 8 lanes per row
//      reading consecutive 16-byte pieces (a warp instruction touches 4 cache lines, not 32), and apply
//      what the reference applies per row (ollama_rag_system.py:480-508): row window, valid mask, clamp,
//      `>= threshold`;
//   3. keep the k best rows (:511-515), decode to (score, global id).

// number of keys of a descending-sorted list (zeros at the tail) that are greater than `key`
__device__ __forceinline__ int count_greater_sorted(const uint64_t* list, int len_pow2, uint64_t key) {
  int pos = 0;
  for (int step = len_pow2 >> 1; step > 0; step >>= 1)
    if (list[pos + step - 1] > key) pos += step;
  return pos + (list[pos] > key ? 1 : 0);
}


#ifndef RAGB_FINISH_CTAS
#define RAGB_FINISH_CTAS 3  // co-resident CTAs per SM the register budget is cut for (build-time A/B)
#endif
// kRows = rows per chunk: 16 (a half-warp per chunk, two chunks per warp at a time) or 32 (a warp per chunk)
template <typename T, int kRows>
__global__ void __launch_bounds__(kMergeThreads, RAGB_FINISH_CTAS) finish_chunks_kernel(const FinishParams p) {
  constexpr int kFinishChunkRows = kRows;
  constexpr int kPerWarp = 32 / kRows;            // chunks a warp scores at a time
  constexpr int kLaneMask = kRows - 1;            // lane -> row of its chunk
  constexpr int kGroups = kRows / 8;              // 8-lane groups per chunk: 2 or 4
  extern __shared__ __align__(16) uint64_t fsm[];
  constexpr int E = Elem<T>::per16;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t qi = blockIdx.x;
  uint64_t* lists = fsm;
  uint64_t* chunk = lists + (size_t)p.n_lists * p.kpad;  // selected chunk keys, by rank
  uint64_t* cmax = chunk + p.kpad;                        // exact key of the best row of each selected chunk
  uint64_t* surv = cmax + p.kpad;                         // rows that can still be among the k best
  uint64_t* fin = surv + (size_t)p.kpad * kFinishMaxChunkRows;
  int* misc = reinterpret_cast<int*>(fin + p.kpad);       // [0] survivor count  [1..2] threshold key

  // ---- 0. what does not depend on the tensor pass: launched with programmatic stream serialization, the CTAs of
  // this kernel may start while the last CTAs of the kernel that writes p.parts are still draining (csrc/ragb200.cu,
  // gemm_finish_chunks); griddepcontrol.wait below returns when that kernel has completed and its writes are visible.
  for (int i = threadIdx.x; i < p.kpad; i += kMergeThreads) { chunk[i] = 0; cmax[i] = 0; fin[i] = 0; }
  if (threadIdx.x == 0) { misc[0] = 0; misc[1] = 0; misc[2] = 0; }
  const float* qv = p.q + qi * p.dim;
  float scale = 1.f;
  if (p.metric == 1) {
    float qq = 0.f;
    for (int i = lane; i < p.dim; i += 32) { const float a = qv[i]; qq = fmaf(a, a, qq); }
    qq = warp_sum(qq);
    scale = qq > 0.f ? rsqrtf(qq) : 0.f;  // zero query scores 0.0 everywhere, :214-215
  }
  asm volatile("griddepcontrol.wait;" ::: "memory");
  // ---- 1. best k_sel chunks: rank every key by counting (lists are sorted, keys are unique) ----
  const int n_keys = p.n_lists * p.kpad;
  const uint64_t* base = p.parts + qi * (int64_t)n_keys;
  for (int i = threadIdx.x; i < n_keys; i += kMergeThreads) lists[i] = base[i];
  __syncthreads();
  // Bound first: the first m keys of every list (m * n_lists >= k_sel of them) contain at least k_sel keys at or above
  // their k_sel-th largest, L, so the k_sel best keys overall are all >= L and only those need a rank.  With many
  // short lists (a single query tile is split 74 ways) that is a few dozen keys instead of thousands, each of which
  // would otherwise be binary-searched through every list.
  __shared__ unsigned long long s_bound;
  __shared__ int s_ncand;
  // (with a handful of long lists the binary searches below are as cheap: the bound pays from ~8 lists up)
  const bool bounded = p.n_lists > 8;
  if (threadIdx.x == 0) { s_bound = 0ull; s_ncand = bounded ? 0 : 0x40000000; }
  __syncthreads();
  if (bounded) {
    const int m = min(p.kpad, (p.k_sel + p.n_lists - 1) / p.n_lists);
    const int hs = m * p.n_lists;
    for (int i = threadIdx.x; i < hs; i += kMergeThreads) {
      const uint64_t key = lists[(i / m) * p.kpad + (i % m)];
      if (key == 0) continue;
      int rank = 0;
      for (int l = 0; l < p.n_lists; ++l)
        for (int t = 0; t < m; ++t) rank += lists[l * p.kpad + t] > key ? 1 : 0;
      if (rank == p.k_sel - 1) s_bound = key;
    }
  }
  __syncthreads();
  const uint64_t bound = s_bound;  // 0: fewer than k_sel keys among the heads — everything is a candidate
  uint64_t* cand = surv;           // free until step 2
  const int cand_cap = p.kpad * kFinishMaxChunkRows;
  for (int i = threadIdx.x; bounded && i < n_keys; i += kMergeThreads) {
    const uint64_t key = lists[i];
    if (key != 0 && key >= bound) {
      const int pos = atomicAdd(&s_ncand, 1);
      if (pos < cand_cap) cand[pos] = key;
    }
  }
  __syncthreads();
  const int n_cand1 = s_ncand;
  if (n_cand1 <= cand_cap) {
    for (int i = threadIdx.x; i < n_cand1; i += kMergeThreads) {
      const uint64_t key = cand[i];
      int rank = 0;
      for (int j = 0; j < n_cand1; ++j) rank += cand[j] > key ? 1 : 0;
      if (rank < p.k_sel) chunk[rank] = key;
    }
  } else {  // more candidates than fit: rank every key by binary search in each sorted list
    for (int i = threadIdx.x; i < n_keys; i += kMergeThreads) {
      const uint64_t key = lists[i];
      if (key == 0) continue;
      int rank = 0;
      for (int l = 0; l < p.n_lists && rank < p.k_sel; ++l) rank += count_greater_sorted(lists + l * p.kpad, p.kpad, key);
      if (rank < p.k_sel) chunk[rank] = key;
    }
  }
  __syncthreads();

  // ---- 2. exact scores of the rows of the selected chunks (registers: my_key of row start+lane) ----
  // kRows lanes per chunk, 8 lanes per row: lane group gq of the chunk scores rows start + kGroups*r + gq, r = 0..7
  const int hw = lane / kRows, gq = (lane >> 3) & (kGroups - 1), l = lane & 7;
  const int pieces = (int)(p.ld / E);     // ld is padded to 16 bytes, padding columns are zero
  const bool q_vec = (p.dim % 4 == 0) && ((reinterpret_cast<uintptr_t>(qv) & 15) == 0);
  for (int c0 = kPerWarp * warp; c0 < p.k_sel; c0 += kPerWarp * kMergeWarps) {
    const int c = c0 + hw;
    const uint64_t ckey = c < p.k_sel ? chunk[c] : 0ull;
    const bool active = ckey != 0;
    if (!__any_sync(0xffffffffu, active)) continue;
    const int64_t start = (int64_t)key_id(ckey);
    // per 16-byte piece of the query, eight independent row loads are in flight and the query piece is
    // loaded once for all eight rows
    float acc[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) acc[r] = 0.f;
    const T* x0 = static_cast<const T*>(p.X) + (size_t)(start + gq) * p.ld;
#pragma unroll 1
    for (int pc = l; pc < pieces; pc += 8) {
      float qf[E];
      const int d0 = pc * E;
      if (q_vec && d0 + E <= p.dim) {
#pragma unroll
        for (int i = 0; i < E; i += 4) {
          const float4 t = __ldg(reinterpret_cast<const float4*>(qv + d0 + i));
          qf[i] = t.x; qf[i + 1] = t.y; qf[i + 2] = t.z; qf[i + 3] = t.w;
        }
      } else {
#pragma unroll
        for (int i = 0; i < E; ++i) qf[i] = d0 + i < p.dim ? __ldg(qv + d0 + i) : 0.f;
      }
      uint4 v[8];
#pragma unroll
      for (int r = 0; r < 8; ++r)
        v[r] = (active && start + kGroups * r + gq < p.row_end)
                   ? ld_stream16(reinterpret_cast<const uint4*>(x0 + (size_t)(kGroups * r) * p.ld) + pc)
                   : make_uint4(0, 0, 0, 0);
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        float f[E];
        Elem<T>::unpack(v[r], f);
        float a = acc[r];
#pragma unroll
        for (int i = 0; i < E; ++i) a = fmaf(f[i], qf[i], a);
        acc[r] = a;
      }
    }
    float mine = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      float a = acc[r];
      a += __shfl_xor_sync(0xffffffffu, a, 4);
      a += __shfl_xor_sync(0xffffffffu, a, 2);
      a += __shfl_xor_sync(0xffffffffu, a, 1);
      // lane j of the chunk keeps the score of row start+j = start + kGroups*(j/kGroups) + j%kGroups: held by lane
      // group j%kGroups of the chunk's lanes
      const float v2 = __shfl_sync(0xffffffffu, a, (lane & ~kLaneMask) + (lane & (kGroups - 1)) * 8);
      if (((lane & kLaneMask) / kGroups) == r) mine = v2;
    }
    const int64_t row = start + (lane & kLaneMask);
    const float sc = finish_score(mine, scale, p.flags);
    bool ok = active && row >= p.row_begin && row < p.row_end && sc >= p.threshold;
    if (ok && p.valid) ok = p.valid[row] != 0;
    const uint64_t key = ok ? make_key(sc, map_id(p.ids, row)) : 0ull;
    if (active) surv[c * kFinishChunkRows + (lane & kLaneMask)] = key;  // parked; filtered below
    uint64_t m = key;
#pragma unroll
    for (int o = kRows / 2; o > 0; o >>= 1) {
      const uint64_t t = __shfl_xor_sync(0xffffffffu, m, o);
      m = t > m ? t : m;
    }
    if ((lane & kLaneMask) == 0 && active) cmax[c] = m;
  }
  __syncthreads();

  // ---- 3. the k best rows (:511-515) ----
  // the k-th best chunk-best key bounds the k-th best row from below: only rows at or above it are ranked
  if (threadIdx.x < p.k_sel) {
    const uint64_t key = cmax[threadIdx.x];
    if (key != 0) {
      int rank = 0;
      for (int j = 0; j < p.k_sel; ++j) rank += cmax[j] > key ? 1 : 0;
      if (rank == p.k - 1) { misc[1] = (int)(uint32_t)key; misc[2] = (int)(uint32_t)(key >> 32); }
    }
  }
  __syncthreads();
  const uint64_t tau = ((uint64_t)(uint32_t)misc[2] << 32) | (uint32_t)misc[1];  // 0 when fewer than k chunks hold a row
  // compact the survivors in place: slot order does not matter, ranks are computed against the whole set
  uint64_t mykeys[(128 * kFinishChunkRows) / kMergeThreads];  // k_sel <= 128 chunks
  int n_mine = 0;
  const int n_cand = p.k_sel * kFinishChunkRows;
#pragma unroll
  for (int u = 0; u < (128 * kFinishChunkRows) / kMergeThreads; ++u) {
    const int i = threadIdx.x + u * kMergeThreads;
    uint64_t key = 0;
    if (i < n_cand && chunk[i / kFinishChunkRows] != 0) key = surv[i];
    mykeys[u] = (key != 0 && key >= tau) ? key : 0ull;
  }
  __syncthreads();
#pragma unroll
  for (int u = 0; u < (128 * kFinishChunkRows) / kMergeThreads; ++u)
    if (mykeys[u] != 0) { surv[atomicAdd(&misc[0], 1)] = mykeys[u]; ++n_mine; }
  __syncthreads();
  const int n_surv = misc[0];
#pragma unroll
  for (int u = 0; u < (128 * kFinishChunkRows) / kMergeThreads; ++u) {
    const uint64_t key = mykeys[u];
    if (key == 0) continue;
    int rank = 0;
    for (int j = 0; j < n_surv && rank < p.k; ++j) rank += surv[j] > key ? 1 : 0;
    if (rank < p.k) fin[rank] = key;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < p.k; i += kMergeThreads) {
    const uint64_t key = fin[i];
    if (p.out_scores) p.out_scores[qi * p.k + i] = key ? key_score(key) : -INFINITY;
    if (p.out_ids) p.out_ids[qi * p.k + i] = key ? decode_id(p.row_of_id, key_id(key)) : -1;
  }
  if (p.out_keys)
    for (int i = threadIdx.x; i < p.kpad; i += kMergeThreads) p.out_keys[qi * p.kpad + i] = i < p.k ? fin[i] : 0;
  if (p.peers.n) {
    for (int i = threadIdx.x; i < p.kpad; i += kMergeThreads) peer_emit(p.peers, qi * p.kpad + i, i < p.k ? fin[i] : 0);
    peer_signal(p.peers, gridDim.x);
  }
}

}  // namespace ragb

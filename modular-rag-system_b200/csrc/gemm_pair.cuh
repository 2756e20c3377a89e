// gemm_pair.cuh — K4 (pair form): batched exact top-k on an SM PAIR with tcgen05.mma.cta_group::2.
//
// Two CTAs of a cluster own two different 128-query tiles and the SAME row range.  One MMA of
// shape M=256 x N=256 x K=16 spans both SMs: each SM supplies its own query tile (A, 128 x 64 per
// k-block) and HALF of the 256 corpus rows (B, 128 x 64 per k-block), and receives the scores of its
// own 128 queries against all 256 rows in its own TMEM.  Against the single-CTA kernel in
// gemm_ts.cuh this halves the corpus bytes every SM stages and reads per k-block (48 KB -> 32 KB), so
// the ring holds 5 stages instead of 3 next to the candidate buffers, and it keeps the pair in lock
// step by construction (the leader's commit releases the stage in both CTAs).
//
// Roles per CTA (320 threads by default — kEpi8 — or 192; 1 CTA/SM):
//   warp 0    producer : TMA (cp.async.bulk.tensor ... cta_group::2) of the own query k-block (or, kQRes, the whole
//                        query tile once) and the own half of the corpus k-block; bytes are accounted on the LEADER's
//                        barrier
//   warp 1    MMA      : leader CTA only; one elected lane issues tcgen05.mma.cta_group::2 and
//                        multicasts tcgen05.commit to both CTAs' barriers
//   warps 2-9 epilogue : thread == query == TMEM lane; two warps per lane quarter, splitting the accumulator's columns
//                        (shared per-query state, named barriers) or — kPar — the tiles (private state);
//                        "accumulator drained" arrives on the leader's barrier from both CTAs (mapa + mbarrier.arrive).
//                        Chunk-granular selection, prune by rank counting / bisection, flush at the end (below).
// Forms of the one template: pair (this header's title), kSolo (one CTA, cta_group::1, a batch of one query tile),
// kQRes (resident query tile, dim <= 384), kEpi8 / kPar (epilogue organisation), E (candidate buffers: shared memory for
// k <= 32, global memory for k <= 64 / 128).  -DRAGB_PAIR_STATS adds cycle counters around every wait.
//
// The reference has no batched path (callers loop RAGVectorStore.search,
// ollama_rag_system.py:388-403 / :480-515); per query the result is the same scan.
#pragma once
#include "gemm_ts.cuh"

namespace ragb {

// -DRAGB_PAIR_STATS (build variant, scripts/pair_stats.sh): where the warps of a CTA wait, in cycles, per CTA:
// [0] kernel  [1] producer: empty slot  [2] producer: meeting  [3] MMA: accumulator free  [4] MMA: stage full
// [5] epilogue warp 2: accumulator full  [6] epilogue warp 2: prune  [7] epilogue warp 2: flush
// [8] epilogue warp 2: draining a tile  [9..11] inside prune (8 warps): first barrier | prunes | second barrier
// [12] number of prunes by warp 2  [13] cycles inside the prune function itself
#ifdef RAGB_PAIR_STATS
#define STAT_T0() const long long st_t0_ = clock64()
#define STAT_ADD(i) st_acc_[i] += clock64() - st_t0_
#define STAT_DECL() long long st_acc_[14] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}; const long long st_k0_ = clock64()
#define STAT_OUT(cond, lo, hi)                                                                   \
  if ((cond) && p.stats) {                                                                       \
    for (int i_ = (lo); i_ <= (hi); ++i_) p.stats[(size_t)blockIdx.x * 14 + i_] = st_acc_[i_];   \
  }
#else
#define STAT_T0()
#define STAT_ADD(i)
#define STAT_DECL()
#define STAT_OUT(cond, lo, hi)
#endif

constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;    // shared::cluster address of the even (leader) CTA of a pair

__device__ __forceinline__ void tma_load_2d_pair(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tc_mma_f16_pair(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                                uint32_t accum) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum)
      : "memory");
}
__device__ __forceinline__ void tc_commit_pair(uint64_t* bar) {  // arrives on both CTAs' barrier at this offset
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)),
               "h"((uint16_t)3)
               : "memory");
}
// Arrive on CTA `cta`'s copy of the barrier.  RELAXED on purpose: the arrival only says "my tcgen05.ld of this
// accumulator have completed" (tcgen05.wait::ld + fence::before_thread_sync precede it); no memory is published.
// A release here costs an ERRBAR + CCTL.IVALL per tile and has to drain the candidate stores (global memory for
// k > 32) before the MMA warp may reuse the accumulator — 13 % of the kernel's stall samples at top-100.
__device__ __forceinline__ void mbar_arrive_cta(uint64_t* bar, uint32_t cta) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [ra];\n\t}" ::"r"(smem_u32(bar)),
      "r"(cta)
      : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {  // acquire at cluster scope
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAITC_%=:\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONEC_%=;\n\t"
      "bra WAITC_%=;\n\t"
      "DONEC_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

// ---- candidate buffers ----------------------------------------------------------------------
// E selects the per-query candidate buffer:
//   E == 2 : kpad <= 32, 48 slots in shared memory, pruned by warp_top32_of_64
//   E == 4 : kpad == 64, E == 8 : kpad == 128 — 32*E slots in GLOBAL memory (L2 resident; an insert is a rare
//            event, about k*ln(rows/k) per query over a CTA's lifetime), pruned by a register bitonic sort
template <int E, bool Q = false> struct CandCfg {
  static constexpr int cap = E == 2 ? (Q ? kPairCapSmallQ : kPairCapSmall) : 32 * E;
  static constexpr int stride = E == 2 ? cap + 1 : 32 * E;  // +1: conflict-free for both access patterns
};
// Resident query tile (kQRes): when the 128 x dim query tile of a CTA fits in shared memory next to the ring
// (dim <= 384: 6 k-blocks of 16 KB), it is loaded ONCE and the ring carries corpus k-blocks only.  Streaming the
// query k-block with every corpus k-block doubles the L2 -> SM bytes (measured at 12.5M x 384, B=4096, top-100:
// 154 GB per launch over the crossbar = 15.6 TB/s at full tensor rate, more than L2 delivers; the MMA warp waited
// for TMA 22 % of the time, tensor pipe 60 % active).  With the tile resident the ring moves 16 KB per k-block.
// rows behind one candidate key: 16 — half of one tcgen05.ld of 32 columns; the finish kernel then re-reads half as
// many rows (1M x 384, B=1024, top-10: 133 -> 96 us).  For k > 32 this used to lose (twice the inserts into the
// global-memory buffers, +8 % at 1M x 384 top-100 with the full-sort prune); with the sorted-prefix prune it wins or
// ties: 1M x 384 top-100 B=1024 1.96 -> 1.77 ms, 10M x 768 13.7 -> 13.4 ms, 12.5M x 384 B=4096 33.4 -> 33.6 ms.

__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}
// tcgen05.wait::ld that also redefines the registers of the load it waits for, so that no use of them
// can be scheduled above it
__device__ __forceinline__ void tc_wait_ld_dep(uint32_t* r) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                 "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]),
                 "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]),
                 "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
               :
               : "memory");
}

// bitonic sort of 32*E keys held E per lane (element index i = e*32 + lane), descending in i
template <int E>
__device__ __forceinline__ void warp_sort_desc(uint64_t (&v)[E], int lane) {
#pragma unroll
  for (int size = 2; size <= 32 * E; size <<= 1) {
#pragma unroll
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      if (stride >= 32) {
        const int se = stride >> 5;
#pragma unroll
        for (int e = 0; e < E; ++e) {
          if ((e & se) == 0) {
            const bool desc = ((e << 5) & size) == 0;
            const uint64_t a = v[e], b = v[e | se];
            const uint64_t hi = a > b ? a : b, lo = a > b ? b : a;
            v[e] = desc ? hi : lo;
            v[e | se] = desc ? lo : hi;
          }
        }
      } else {
        const bool lower = (lane & stride) == 0;
#pragma unroll
        for (int e = 0; e < E; ++e) {
          const uint64_t o = __shfl_xor_sync(0xffffffffu, v[e], stride);
          const bool desc = (((e << 5) | lane) & size) == 0;
          const bool take_max = lower == desc;
          v[e] = take_max ? (v[e] > o ? v[e] : o) : (v[e] < o ? v[e] : o);
        }
      }
    }
  }
}

// bitonic merge: 32*E keys (element i = e*32 + lane) forming a bitonic sequence -> descending in i
template <int E>
__device__ __forceinline__ void warp_merge_desc(uint64_t (&v)[E], int lane) {
#pragma unroll
  for (int stride = 16 * E; stride > 0; stride >>= 1) {
    if (stride >= 32) {
      const int se = stride >> 5;
#pragma unroll
      for (int e = 0; e < E; ++e) {
        if ((e & se) == 0) {
          const uint64_t a = v[e], b = v[e | se];
          v[e] = a > b ? a : b;
          v[e | se] = a > b ? b : a;
        }
      }
    } else {
      const bool lower = (lane & stride) == 0;
#pragma unroll
      for (int e = 0; e < E; ++e) {
        const uint64_t o = __shfl_xor_sync(0xffffffffu, v[e], stride);
        v[e] = lower ? (v[e] > o ? v[e] : o) : (v[e] < o ? v[e] : o);
      }
    }
  }
}

// A buffer of 32*E slots whose lower half (16*E keys) is sorted descending and whose upper half holds n - 16*E new
// keys in any order -> lo = the 16*E best of all of them, descending: sort only the new half, take the larger of
// lower[i] and upper[last - i] (the best half of the union, as a bitonic sequence), merge.
template <int E>
__device__ __forceinline__ void warp_best_half(const uint64_t* base, int n, int lane, uint64_t (&lo)[E / 2]) {
  constexpr int H = E / 2;
  uint64_t hi[H];
#pragma unroll
  for (int e = 0; e < H; ++e) {
    lo[e] = base[e * 32 + lane];
    hi[e] = ((e + H) * 32 + lane < n) ? base[(e + H) * 32 + lane] : 0ull;
  }
  warp_sort_desc<H>(hi, lane);
#pragma unroll
  for (int e = 0; e < H; ++e) {
    const uint64_t o = __shfl_sync(0xffffffffu, hi[H - 1 - e], 31 - lane);
    lo[e] = lo[e] > o ? lo[e] : o;
  }
  warp_merge_desc<H>(lo, lane);
}

// All 32 lanes: sort the n keys of one query's buffer, keep the best min(n,k) in place (sorted),
// return the k-th best (0 if fewer than k are held).  `sorted` = how many keys at the front of the buffer are still
// in descending order from the previous prune (the epilogue only appends behind them).
template <int E>
__device__ __forceinline__ uint64_t warp_prune(uint64_t* base, int n, int k, int lane, int* keep_out, int sorted) {
  const int keep = n < k ? n : k;
  *keep_out = keep;
  if constexpr (E >= 4) {
    // Every prune but a query's first finds the lower half of the buffer (k = 16*E keys) sorted: sort only the new
    // upper half, take the larger of lower[i] and upper[last - i] — the k best of the union, as a bitonic sequence —
    // and merge.  Half the compare-exchanges of the full sort (E=8: 144 against 288 per lane).
    if (sorted == 16 * E && k == 16 * E) {
      constexpr int H = E / 2;
      uint64_t lo[H];
      warp_best_half<E>(base, n, lane, lo);
      __syncwarp();
#pragma unroll
      for (int e = 0; e < H; ++e) base[e * 32 + lane] = lo[e];
      return __shfl_sync(0xffffffffu, lo[H - 1], 31);
    }
  }
  if constexpr (E == 2) {
    const uint64_t v = warp_top32_of_64(base, n, lane);
    __syncwarp();
    if (lane < keep) base[lane] = v;
    return __shfl_sync(0xffffffffu, v, k - 1);
  } else {
    uint64_t v[E];
#pragma unroll
    for (int e = 0; e < E; ++e) v[e] = (e * 32 + lane < n) ? base[e * 32 + lane] : 0ull;
    warp_sort_desc<E>(v, lane);
    __syncwarp();
    uint64_t kv = 0;
#pragma unroll
    for (int e = 0; e < E; ++e) {
      if (e * 32 + lane < keep) base[e * 32 + lane] = v[e];
      if (e == ((k - 1) >> 5)) kv = v[e];
    }
    return __shfl_sync(0xffffffffu, kv, (k - 1) & 31);
  }
}

// Prune by SELECTION instead of sorting (p.full_sort == 2): the epilogue only needs the k best keys of a
// buffer and the k-th of them as its new threshold — not their order (the flush sorts once, at the end).  The k-th
// largest of the n <= 32*E keys is found by bisection on the 64-bit key value: every lane holds E keys in registers,
// one step is E compares and one redux.sync; keys are unique (low word = chunk start row), and the search stops as soon
// as a pivot separates exactly k keys — typically after 12-16 steps, since the keys of a buffer share their sign and
// exponent bits.  Survivors are compacted to the front with ballots, unsorted.  About a tenth of the instructions of
// the 256-key bitonic sort + merge (profiles/r2_pair_wait_stats.txt: the sort-based prune was 14 % / 50 % of an epilogue
// warp's time at 12.5M x 384 / 1M x 384 top-100).
template <int E>
__device__ __forceinline__ uint64_t warp_prune_select(uint64_t* base, int n, int k, int lane, int* keep_out) {
  uint64_t v[E];
#pragma unroll
  for (int e = 0; e < E; ++e) v[e] = (e * 32 + lane < n) ? base[e * 32 + lane] : 0ull;
  if (n <= k) {   // nothing to drop; the k-th best exists only when n == k
    *keep_out = n;
    if (n < k) return 0ull;
    uint64_t m = ~0ull;
#pragma unroll
    for (int e = 0; e < E; ++e) m = (v[e] != 0ull && v[e] < m) ? v[e] : m;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { const uint64_t t = __shfl_xor_sync(0xffffffffu, m, o); m = t < m ? t : m; }
    return m;
  }
  // bounds from the score words: lo = smallest present key's score word, hi = largest
  unsigned hmax = 0u, hmin = 0xffffffffu;
#pragma unroll
  for (int e = 0; e < E; ++e) {
    const unsigned h = (unsigned)(v[e] >> 32);
    hmax = h > hmax ? h : hmax;
    hmin = (v[e] != 0ull && h < hmin) ? h : hmin;
  }
  hmax = __reduce_max_sync(0xffffffffu, hmax);
  hmin = __reduce_min_sync(0xffffffffu, hmin);
  uint64_t lo = (uint64_t)hmin << 32, hi = ((uint64_t)hmax << 32) | 0xffffffffull;
  // invariant: count(v >= lo) >= k (all n > k keys are >= lo), count(v > hi) == 0
  uint64_t cut = lo;
  while (lo < hi) {
    const uint64_t mid = lo + ((hi - lo) >> 1) + 1ull;
    int c = 0;
#pragma unroll
    for (int e = 0; e < E; ++e) c += v[e] >= mid ? 1 : 0;
    c = __reduce_add_sync(0xffffffffu, c);
    if (c >= k) { lo = mid; cut = mid; if (c == k) break; }
    else hi = mid - 1ull;
  }
  if (lo >= hi) cut = lo;   // converged without an exact split: lo is the k-th largest key itself
  // compact the survivors (exactly k: keys are unique) to the front; every key is in registers already
  __syncwarp();
  int off = 0;
  uint64_t m = ~0ull;
#pragma unroll
  for (int e = 0; e < E; ++e) {
    const bool keep = v[e] >= cut;
    const unsigned b = __ballot_sync(0xffffffffu, keep);
    if (keep) { base[off + __popc(b & ((1u << lane) - 1u))] = v[e]; m = v[e] < m ? v[e] : m; }
    off += __popc(b);
  }
  *keep_out = off;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { const uint64_t t = __shfl_xor_sync(0xffffffffu, m, o); m = t < m ? t : m; }
  return m;   // the k-th best key
}

// Prune of a SHARED-MEMORY buffer (k <= 32, n <= 64 keys) by rank counting: a lane holds up to two keys and counts, for
// each, the keys of the buffer above it — n broadcast loads, every step independent of the last (the bisection above is
// a chain of ~15 redux round trips, ~1 400 cycles per prune measured) — and writes the keys of rank < k to slot `rank`:
// the survivors come out sorted.
__device__ __forceinline__ uint64_t warp_prune_rank(uint64_t* base, int n, int k, int lane, int* keep_out) {
  const uint64_t v0 = lane < n ? base[lane] : 0ull, v1 = lane + 32 < n ? base[lane + 32] : 0ull;
  int r0 = 0, r1 = 0;
#pragma unroll 4
  for (int j = 0; j < n; ++j) {
    const uint64_t kj = base[j];
    r0 += kj > v0 ? 1 : 0;
    r1 += kj > v1 ? 1 : 0;
  }
  const int keep = n < k ? n : k;
  *keep_out = keep;
  __syncwarp();
  if (lane < n && r0 < keep) base[r0] = v0;
  if (lane + 32 < n && r1 < keep) base[r1] = v1;
  // the k-th best key (0 if fewer than k are held)
  const unsigned b0 = __ballot_sync(0xffffffffu, lane < n && r0 == k - 1);
  const unsigned b1 = __ballot_sync(0xffffffffu, lane + 32 < n && r1 == k - 1);
  uint64_t kth = 0ull;
  if (b0) kth = __shfl_sync(0xffffffffu, v0, __ffs(b0) - 1);
  else if (b1) kth = __shfl_sync(0xffffffffu, v1, __ffs(b1) - 1);
  return kth;
}

// ---- the fused top-k epilogue: CHUNK-granular selection -----------------------------------------
// Thread == query holds the scores of its query against 32 consecutive rows (two chunks of 16).  The epilogue
// keeps, per query, the best k_sel CHUNKS (16 consecutive rows) ranked by (chunk maximum, chunk start
// row): a 3-input max tree and one compare per chunk, and — rarely, about k*ln(chunks/k) times per query — one store.
// Why that is exact: the k best rows of a query lie inside its k best chunks.  (A chunk outside them
// is beaten by k chunks whose maximum rows each beat every row of it — by score, or on equal score by
// the larger row id, since a later chunk holds later rows.)  The k-th best chunk maximum is therefore
// also a valid running threshold.  rescore_chunks_kernel then scores the 32 rows of every selected
// chunk in fp32 and keeps the k best rows; row range, valid mask and the user threshold are
// applied there, exactly.  Here rows outside the range / masked out are only kept from raising a
// chunk's maximum (kMasked), so the threshold stays a lower bound.
// Shared running threshold (p.gthr): the row splits of one query each see 1/n_splits of the rows, so each split's own
// k-th best chunk maximum tightens n_splits times more slowly than it could.  Any split's k-th best chunk maximum is
// a lower bound of the query's k-th best over ALL rows, so it may cut every other split as well: after each prune a
// lane publishes its threshold with atomicMax (order-preserving bits, 0 = none yet) and every thread picks up the
// query's word once per tile (the load is issued before the wait for the accumulator).  Which chunks a split keeps
// then depends on timing; the k best chunks overall — all the finish kernel uses — do not.
__device__ __forceinline__ unsigned ld_relaxed_u32(const unsigned* ptr) {
  unsigned v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ptr) : "memory");
  return v;
}
struct EpiCtx {
  unsigned* gthr;       // this query's word of the shared threshold, or null
  uint64_t* my;         // this query's candidate buffer
  uint64_t* warp_cand;  // buffer of lane 0 of this warp; lane j's is warp_cand + j*stride
  int cnt;
  int sorted;           // keys at the front of the buffer still sorted from the last prune
  float user_fast, fthr;
  // kEpi8 (two epilogue warps per TMEM lane quarter, each reading half of the accumulator's columns): the two threads
  // of a query share its buffer, so the count / sorted prefix / threshold live in shared memory, [128] each
  int* cnt_s;           // this query's entry
  int* sorted_s;
  float* thr_s;
};
// append a candidate key to the query's buffer
template <bool kEpi8>
__device__ __forceinline__ void epi_push(EpiCtx& e, uint64_t key) {
  if constexpr (kEpi8) e.my[atomicAdd(e.cnt_s, 1)] = key;   // rare: about k*ln(chunks/k) times per query and CTA
  else e.my[e.cnt++] = key;
}
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
// max of 16 consecutive columns: five 3-input maxima, then 6 -> 2 -> 1
__device__ __forceinline__ float max16(const uint32_t* r) {
  float t[5];
#pragma unroll
  for (int i = 0; i < 5; ++i)
    t[i] = fmax3(__uint_as_float(r[3 * i]), __uint_as_float(r[3 * i + 1]), __uint_as_float(r[3 * i + 2]));
  return fmaxf(fmax3(t[0], t[1], t[2]), fmax3(t[3], t[4], __uint_as_float(r[15])));
}
// one tcgen05.ld of 32 columns = two chunks of 16 rows (kRows == 16) or one chunk of 32
template <bool kMasked, int kRows, bool kEpi8 = false>
__device__ __forceinline__ void epi_chunk(EpiCtx& e, uint32_t (&r)[32], int64_t row_c, const TsParams& p, int lane) {
  bool any_a = true, any_b = true;
  if (kMasked) {
    const int64_t row = row_c + lane;
    bool v = row < p.row_end;
    if (v && p.valid) v = p.valid[row] != 0;
    const unsigned vb = __ballot_sync(0xffffffffu, v);
    any_a = (vb & 0xffffu) != 0;
    any_b = (vb >> 16) != 0;
    if (vb != 0xffffffffu) {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (!((vb >> j) & 1u)) r[j] = 0xff800000u;  // -inf: cannot raise the chunk maximum
    }
  }
  const float ma = max16(&r[0]), mb = max16(&r[16]);
  if constexpr (kRows == 16) {
    if (ma >= e.fthr && any_a) epi_push<kEpi8>(e, make_key(finish_score(ma, 1.0f, p.flags), (uint32_t)row_c));
    if (mb >= e.fthr && any_b) epi_push<kEpi8>(e, make_key(finish_score(mb, 1.0f, p.flags), (uint32_t)row_c + 16));
  } else {
    const float mx = fmaxf(ma, mb);
    if (mx >= e.fthr && (any_a || any_b)) epi_push<kEpi8>(e, make_key(finish_score(mx, 1.0f, p.flags), (uint32_t)row_c));
  }
}
// once per tile (a query gains at most kPairN / rows-per-chunk = 16 or 8 chunks per tile): prune the buffers that could overflow next tile
template <int E, bool Q>
__device__ __forceinline__ void epi_prune(EpiCtx& e, const TsParams& p, int lane) {
  constexpr int kCap = CandCfg<E, Q>::cap, kStride = CandCfg<E, Q>::stride;
  unsigned need = __ballot_sync(0xffffffffu, e.cnt > kCap - kPairN / ChunkCfg<E>::rows);
  while (need) {
    const int j = __ffs(need) - 1;
    need &= need - 1;
    const int n = __shfl_sync(0xffffffffu, e.cnt, j);
    __syncwarp();
    int keep;
    const int sorted = p.full_sort ? 0 : __shfl_sync(0xffffffffu, e.sorted, j);
    uint64_t kth;
    if (E == 2 && p.full_sort == 2) kth = warp_prune_rank(e.warp_cand + (size_t)j * kStride, n, p.k, lane, &keep);
    else if (p.full_sort == 2) kth = warp_prune_select<E>(e.warp_cand + (size_t)j * kStride, n, p.k, lane, &keep);
    else kth = warp_prune<E>(e.warp_cand + (size_t)j * kStride, n, p.k, lane, &keep, sorted);
    if (lane == j) {
      e.cnt = keep;
      e.sorted = keep;
      if (n >= p.k) {
        e.fthr = fmaxf(e.fthr, fmaxf(e.user_fast, fast_bound(key_score(kth), p.flags)));
        if (e.gthr) atomicMax(e.gthr, order_bits(e.fthr));
      }
    }
    __syncwarp();
  }
}
// kEpi8: the two warps of a lane quarter meet (named barrier: every append of this tile is in the buffers), share the
// prunes — warp half h takes the queries of its parity — and meet again before the next tile picks up the thresholds
// non-blocking: has this phase of the barrier completed?
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
// `next_full` / `next_parity`: the barrier the NEXT accumulator completes.  Besides the prunes that must happen now (a
// buffer that could overflow during the next tile), a warp prunes buffers that are past half of their spare room for
// as long as that accumulator has not arrived — the epilogue warps wait for accumulators 40-60 % of the time, while a
// forced prune in the wrong tile (1 400 cycles and up, against 3 072 tensor cycles per tile at dim 384) is what makes
// the MMA warp wait for an accumulator (profiles/r2_pair_wait_stats.txt).  RAGB_PAIR_SOFT=0 (p.soft_prune) turns it off.
template <int E, bool Q>
__device__ __forceinline__ void epi_prune8(EpiCtx& e, const TsParams& p, int lane, int quarter, int half, uint64_t* next_full,
                                           uint32_t next_parity, long long* st_acc_ = nullptr) {
  constexpr int kCap = CandCfg<E, Q>::cap, kStride = CandCfg<E, Q>::stride;
  constexpr int kHard = kCap - kPairN / ChunkCfg<E>::rows;
#ifdef RAGB_PAIR_STATS
  long long st_a_ = clock64();
#endif
  named_bar_sync(1 + quarter, 64);
#ifdef RAGB_PAIR_STATS
  { const long long t_ = clock64(); st_acc_[9] += t_ - st_a_; st_a_ = t_; }
#endif
  const int my_cnt = *e.cnt_s;
  const unsigned mine = half ? 0xAAAAAAAAu : 0x55555555u;
  unsigned need = __ballot_sync(0xffffffffu, my_cnt > kHard) & mine;
  unsigned soft = p.soft_prune ? (__ballot_sync(0xffffffffu, my_cnt > (p.k + kHard) / 2) & mine & ~need) : 0u;
  for (;;) {
    int j;
    if (need) { j = __ffs(need) - 1; need &= need - 1; }
    else if (soft && !__shfl_sync(0xffffffffu, (int)mbar_test(next_full, next_parity), 0)) { j = __ffs(soft) - 1; soft &= soft - 1; }
    else break;
    const int n = __shfl_sync(0xffffffffu, my_cnt, j);
    const int sorted = p.full_sort ? 0 : __shfl_sync(0xffffffffu, *e.sorted_s, j);
    __syncwarp();
    int keep;
    uint64_t kth;
#ifdef RAGB_PAIR_STATS
    const long long st_p0_ = clock64();
#endif
    if (E == 2 && p.full_sort == 2) kth = warp_prune_rank(e.warp_cand + (size_t)j * kStride, n, p.k, lane, &keep);
    else if (p.full_sort == 2) kth = warp_prune_select<E>(e.warp_cand + (size_t)j * kStride, n, p.k, lane, &keep);
    else kth = warp_prune<E>(e.warp_cand + (size_t)j * kStride, n, p.k, lane, &keep, sorted);
#ifdef RAGB_PAIR_STATS
    st_acc_[12] += 1; st_acc_[13] += clock64() - st_p0_;
#endif
    if (lane == j) {
      *e.cnt_s = keep;
      *e.sorted_s = keep;
      if (n >= p.k) {
        const float t = fmaxf(*e.thr_s, fmaxf(e.user_fast, fast_bound(key_score(kth), p.flags)));
        *e.thr_s = t;
        if (e.gthr) atomicMax(e.gthr, order_bits(t));
      }
    }
    __syncwarp();
  }
#ifdef RAGB_PAIR_STATS
  { const long long t_ = clock64(); st_acc_[10] += t_ - st_a_; st_a_ = t_; }
#endif
  named_bar_sync(1 + quarter, 64);
  e.fthr = fmaxf(e.fthr, *e.thr_s);
#ifdef RAGB_PAIR_STATS
  st_acc_[11] += clock64() - st_a_;
#endif
}

// sort every query's survivors and write kpad keys
template <int E, bool Q>
__device__ __forceinline__ void epi_flush(EpiCtx& e, int64_t qrow0, int split, const TsParams& p, int lane, int j0 = 0,
                                          int jstep = 1, bool shared_state = false) {
  constexpr int kStride = CandCfg<E, Q>::stride;
  __syncwarp();
  const int my_cnt = shared_state ? *e.cnt_s : e.cnt;
  const int my_sorted = shared_state ? *e.sorted_s : e.sorted;
  for (int j = j0; j < 32; j += jstep) {
    const int n = __shfl_sync(0xffffffffu, my_cnt, j);
    uint64_t* base = e.warp_cand + (size_t)j * kStride;
    uint64_t* out = p.out + ((size_t)(qrow0 + j) * p.out_lists + split) * p.kpad;
    const int keep = n < p.k ? n : p.k;
    if constexpr (E == 2) {
      uint64_t v;
      if (n <= 32) {   // the usual case after the last tile's prune: one key per lane, sorted in registers
        uint64_t v1[1] = {lane < n ? base[lane] : 0ull};
        warp_sort_desc<1>(v1, lane);
        v = v1[0];
      } else {
        v = warp_top32_of_64(base, n, lane);
      }
      if (lane < p.kpad) out[lane] = lane < keep ? v : 0ull;
    } else {
      const int sorted = p.full_sort ? 0 : __shfl_sync(0xffffffffu, my_sorted, j);
      if (sorted == 16 * E && p.k == 16 * E) {  // as in warp_prune: the lower half is in order, keep == k == 16*E
        uint64_t lo[E / 2];
        warp_best_half<E>(base, n, lane, lo);
#pragma unroll
        for (int x = 0; x < E / 2; ++x) out[x * 32 + lane] = lo[x];
      } else {
        uint64_t v[E];
#pragma unroll
        for (int x = 0; x < E; ++x) v[x] = (x * 32 + lane < n) ? base[x * 32 + lane] : 0ull;
        warp_sort_desc<E>(v, lane);
#pragma unroll
        for (int x = 0; x < E / 2; ++x) out[x * 32 + lane] = (x * 32 + lane < keep) ? v[x] : 0ull;
      }
    }
  }
}

// kEpi8: 8 epilogue warps (320 threads).  A warp may only touch the TMEM lanes of its quarter (warp % 4), so warps
// 2-5 and 6-9 pair up per quarter and split the 256 accumulator columns of a tile between them.  Why: one epilogue warp
// per scheduler issues a dependent tcgen05.ld -> wait -> max tree -> compare chain (an instruction every ~12 cycles,
// profiles/r1d_pair_top100_waits.txt); at dim <= 384 a tile's MMA takes 3 072 cycles, the epilogue 2 500 - 3 000, and
// any prune lands on the critical path (tensor pipe 74 % active at 1M x 384).  Two warps per scheduler overlap those
// chains, which leaves the TMEM read rate (64 B/cycle: 2 048 cycles per tile) as the epilogue's floor.
//
// kSolo: ONE CTA per SM, no cluster — tcgen05.mma.cta_group::1 of shape M=128 x N=256 x K=16, the CTA stages its query
// k-block and all 256 corpus rows of a k-block itself (48 KB stages).  For a batch of ONE query tile (4 <= B <= 128): the
// pair form would run it beside an all-zero query tile, i.e. spend half of every MMA and half of the TMEM reads on
// nothing.  Such a batch is HBM-bound (128 queries x 2 flop per corpus byte pair), so what matters is corpus bytes in
// flight per SM — every SM walks its own row split, 32 KB of corpus per stage.  Same epilogue, same keys out.
//
// kPar (k > 32, with kEpi8): the two epilogue warps of a lane quarter split the TILES, not the columns — warp half h
// drains accumulator h, i.e. every second tile, all 256 columns — and each keeps its OWN candidate buffers (global
// memory), count and threshold in registers: no shared per-query state, no appends by atomic, no named barriers.
// Why (cycle counters of the column-split form at 12.5M x 384 bf16 top-100, profiles/r2_pair_wait_stats.txt): an
// epilogue warp waits for accumulators 58 % of the time on average, yet the MMA warp waits 21 % of its time for an
// accumulator to be handed back — a 256-key prune takes longer than the one tile of slack the two accumulators give,
// and the per-tile barriers pass every delay on to the partner warp.  With the tile split a warp has two tile times per
// accumulator.  The two halves see disjoint tiles, so a query gets two lists per row split (out_lists = 2 * n_splits);
// they share their running threshold through p.gthr like the row splits do.
template <int E, int kPairStages, bool kQRes = false, bool kEpi8 = false, bool kSolo = false, bool kPar = false>
__global__ void __launch_bounds__(kEpi8 ? kPairThreads8 : kGemmThreads, 1)
gemm_topk_pair_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_x,
                      const TsParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve: [resident query k-blocks] | stages | candidate buffers | barriers | tmem base   (identical offsets in both CTAs)
  typedef CandCfg<E, kQRes> Cand;
  static_assert(!(kSolo && kQRes), "the solo form streams its query k-blocks");
  static_assert(!kPar || (kEpi8 && E >= 4), "tile-parity epilogue: 8 epilogue warps, candidate buffers in global memory");
  constexpr bool kShared = kEpi8 && !kPar;   // the two threads of a query share its buffer (column split)
  constexpr int kStageB = kSolo ? kStageBytes : kQRes ? kStageBytesA : kPairStageBytes;  // bytes per ring stage
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* q_base = smem;
  uint8_t* stage_base = smem + (kQRes ? (size_t)p.n_kb * kStageBytesA : 0);
  uint64_t* cand_smem = reinterpret_cast<uint64_t*>(stage_base + (size_t)kPairStages * kStageB);
  uint64_t* bars = cand_smem + (E == 2 ? (size_t)kGemmM * Cand::stride : 0);
  uint64_t* cand = E == 2 ? cand_smem : p.cand_ws + (size_t)blockIdx.x * kGemmM * Cand::stride;
  uint64_t* full_bar = bars;                     // leader's copy is the one in use
  uint64_t* empty_bar = bars + kPairStages;      // one per CTA
  uint64_t* tfull_bar = bars + 2 * kPairStages;  // one per CTA
  uint64_t* tempty_bar = tfull_bar + 2;          // leader's copy is the one in use
  uint64_t* q_bar = tempty_bar + 2;              // kQRes: the query tiles of both CTAs have landed (leader's copy)
  uint32_t* tmem_base_smem = reinterpret_cast<uint32_t*>(q_bar + 1);
  int* qs_cnt = reinterpret_cast<int*>(tmem_base_smem + 4);  // kEpi8: per-query count | sorted prefix | threshold
  int* qs_sorted = qs_cnt + kGemmM;
  float* qs_thr = reinterpret_cast<float*>(qs_sorted + kGemmM);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = kSolo ? 0 : (int)(blockIdx.x & 1);   // cluster dims (2,1,1): rank == %cluster_ctarank
  const int pair = kSolo ? (int)blockIdx.x : (int)(blockIdx.x >> 1);
  const int pps = kSolo ? p.n_qt : p.n_qt >> 1;  // pairs (solo: CTAs) per row split
  const int split = pair / pps, qt = kSolo ? pair % pps : (pair % pps) * 2 + rank;
  const int64_t t0 = p.n_tiles * split / p.n_splits, t1 = p.n_tiles * (split + 1) / p.n_splits;
  const int n_kb = p.n_kb;
  STAT_DECL();

  if (threadIdx.x == 0) {
    for (int i = 0; i < kPairStages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], (kShared ? 16 : 8) / (kSolo ? 2 : 1)); }
    mbar_init(q_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // the same warp of BOTH CTAs allocates for the pair
    if constexpr (kSolo) {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_base_smem)), "r"(512));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    } else {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_base_smem)), "r"(512));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
    }
  }
  tc_fence_before();
  __syncthreads();
  if constexpr (!kSolo) cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_base_smem;

  if (warp == 0) {
    // ===== producer (both CTAs) =====
    const int q_row = qt * kGemmM;
    int s = 0;
    uint32_t ph = 0;
    if constexpr (kQRes) {
      if (t0 < t1 && elect_one()) {
        if (rank == 0) mbar_expect_tx(q_bar, 2u * (uint32_t)n_kb * kStageBytesA);  // both CTAs' tiles land on the leader
        for (int kb = 0; kb < n_kb; ++kb)
          tma_load_2d_pair(q_base + (size_t)kb * kStageBytesA, &map_q, q_bar, kb * kGemmKB, q_row);
      }
      __syncwarp();
    }
    for (int64_t t = t0; t < t1; ++t) {
      const int row0 = (int)(p.row_begin + t * kPairN) + (kSolo ? 0 : rank * (kPairN / 2));
      // The pps pairs of a row split read the same rows; nothing else keeps them together, and once they drift
      // further apart than L2 holds, every pair fetches the split from HBM by itself (measured: 22 - 40 GB read for
      // 15.4 GB of corpus).  Every kPairSyncTiles tiles the leaders' producers meet at a counter in global memory;
      // the ring is full at that point, so the tensor pipe keeps running while they wait.  Only enabled by the
      // host when every CTA of the grid is resident at once.
      if (p.sync_ctr != nullptr && rank == 0 && t > t0 && ((t - t0) % p.sync_tiles) == 0) {
        { STAT_T0();
        if (elect_one()) {
          const unsigned target = (unsigned)pps * (unsigned)((t - t0) / p.sync_tiles);
          atomicAdd(&p.sync_ctr[split], 1u);
          while (*reinterpret_cast<volatile unsigned*>(&p.sync_ctr[split]) < target) __nanosleep(64);
        }
        __syncwarp();
        STAT_ADD(2); }
      }
      for (int kb = 0; kb < n_kb; ++kb) {
        { STAT_T0(); mbar_wait(&empty_bar[s], ph ^ 1); STAT_ADD(1); }  // the leader's commit released this slot in both CTAs
        if (elect_one()) {
          uint8_t* d = stage_base + (size_t)s * kStageB;
          if constexpr (kSolo) {  // query k-block | 256 corpus rows as two boxes of 128 (the pair form's tensor map)
            mbar_expect_tx(&full_bar[s], (uint32_t)kStageB);
            tma_load_2d(d, &map_q, &full_bar[s], kb * kGemmKB, q_row);
            tma_load_2d(d + kStageBytesA, &map_x, &full_bar[s], kb * kGemmKB, row0);
            tma_load_2d(d + 2 * kStageBytesA, &map_x, &full_bar[s], kb * kGemmKB, row0 + kPairN / 2);
          } else if constexpr (kQRes) {
            if (rank == 0) mbar_expect_tx(&full_bar[s], 2u * kStageB);  // both CTAs' bytes land on the leader
            tma_load_2d_pair(d, &map_x, &full_bar[s], kb * kGemmKB, row0);
          } else {
            if (rank == 0) mbar_expect_tx(&full_bar[s], 2u * kStageB);
            tma_load_2d_pair(d, &map_q, &full_bar[s], kb * kGemmKB, q_row);
            tma_load_2d_pair(d + kStageBytesA, &map_x, &full_bar[s], kb * kGemmKB, row0);
          }
        }
        if (++s == kPairStages) { s = 0; ph ^= 1; }
      }
    }
    STAT_OUT(lane == 0, 1, 2);
  } else if (warp == 1) {
    // ===== MMA issuer (leader only): M=256 across the pair =====
    if (rank == 0) {
      const uint32_t idesc = p.idesc;
      const uint64_t desc0 = umma_desc_sw128(smem_u32(stage_base));
      const uint64_t descq = umma_desc_sw128(smem_u32(q_base));
      int s = 0;
      uint32_t ph = 0, acc = 0, aph = 0;
      if constexpr (kQRes) {
        if (t0 < t1) { mbar_wait(q_bar, 0); tc_fence_after(); }
      }
      for (int64_t t = t0; t < t1; ++t) {
        { STAT_T0();
        if constexpr (kSolo) mbar_wait(&tempty_bar[acc], aph ^ 1);
        else mbar_wait_cluster(&tempty_bar[acc], aph ^ 1);  // both CTAs' epilogues drained this accumulator
        STAT_ADD(3); }
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * (uint32_t)kPairN;
        for (int kb = 0; kb < n_kb; ++kb) {
          { STAT_T0(); mbar_wait(&full_bar[s], ph); STAT_ADD(4); }
          tc_fence_after();
          if (elect_one()) {
            const uint64_t sdesc = desc0 + (uint64_t)((uint32_t)s * (kStageB >> 4));
            const uint64_t adesc = kQRes ? descq + (uint64_t)((uint32_t)kb * (kStageBytesA >> 4)) : sdesc;
            const uint64_t bdesc = kQRes ? sdesc : sdesc + (kStageBytesA >> 4);
            if constexpr (kSolo) {
#pragma unroll
              for (int k4 = 0; k4 < kGemmKB / 16; ++k4)
                tc_mma_f16(d_tmem, adesc + (uint64_t)(k4 * 2), bdesc + (uint64_t)(k4 * 2), idesc, (kb | k4) != 0);
              tc_commit(&empty_bar[s]);
              if (kb == n_kb - 1) tc_commit(&tfull_bar[acc]);
            } else {
#pragma unroll
              for (int k4 = 0; k4 < kGemmKB / 16; ++k4)
                tc_mma_f16_pair(d_tmem, adesc + (uint64_t)(k4 * 2), bdesc + (uint64_t)(k4 * 2), idesc, (kb | k4) != 0);
              tc_commit_pair(&empty_bar[s]);
              if (kb == n_kb - 1) tc_commit_pair(&tfull_bar[acc]);
            }
          }
          if (++s == kPairStages) { s = 0; ph ^= 1; }
        }
        if ((acc ^= 1) == 0) aph ^= 1;
      }
      STAT_OUT(lane == 0, 3, 4);
    }
  } else {
    // ===== epilogue (both CTAs): thread == query (kEpi8: two threads per query, half of the columns each) =====
    const int quarter = warp & 3;
    const int half = kEpi8 ? ((warp - 2) >> 2) : 0;
    const int tid = quarter * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(quarter * 32) << 16);
    EpiCtx e;
    if constexpr (kPar) cand = p.cand_ws + ((size_t)blockIdx.x * 2 + half) * kGemmM * Cand::stride;
    e.my = cand + (size_t)tid * Cand::stride;
    e.warp_cand = cand + (size_t)(quarter * 32) * Cand::stride;
    e.cnt = 0;
    e.sorted = 0;
    e.gthr = p.gthr ? p.gthr + (size_t)qt * kGemmM + tid : nullptr;
    // rows of the query tile past nq are zero padding: they must never take the candidate path
    // the 16-bit-query scores only SELECT: the user threshold is applied with slack here, exactly by the re-score
    e.user_fast = ((int64_t)qt * kGemmM + tid < p.nq) ? fast_bound(p.threshold - p.sel_slack, p.flags) : INFINITY;
    e.fthr = e.user_fast;
    if constexpr (E == 2) {
      // the shared-memory variants are full to the last kilobyte: the per-query words sit in the padding slot each
      // buffer carries for bank spread (slot `cap` of `stride` = cap + 1); no sorted prefix is kept for E == 2
      int* pad = reinterpret_cast<int*>(e.my + Cand::cap);
      e.cnt_s = pad; e.sorted_s = pad; e.thr_s = reinterpret_cast<float*>(pad + 1);
    } else {
      e.cnt_s = qs_cnt + tid; e.sorted_s = qs_sorted + tid; e.thr_s = qs_thr + tid;
    }
    if constexpr (kShared) {
      if (half == 0) { *e.cnt_s = 0; *e.sorted_s = 0; *e.thr_s = e.user_fast; }
      named_bar_sync(1 + quarter, 64);
    }
    constexpr int kBlocks = kPairN / 32;                     // tcgen05.ld blocks of 32 columns per tile
    const int c_lo = kShared ? half * (kBlocks / 2) : 0, c_hi = kShared ? c_lo + kBlocks / 2 : kBlocks;
    uint32_t acc = kPar ? (uint32_t)half : 0u, aph = 0;
    for (int64_t t = t0 + (kPar ? half : 0); t < t1; t += (kPar ? 2 : 1)) {
      const int64_t row0 = p.row_begin + t * kPairN;
      const unsigned gv = e.gthr ? ld_relaxed_u32(e.gthr) : 0u;
      { STAT_T0(); mbar_wait(&tfull_bar[acc], aph); STAT_ADD(5); }
#ifdef RAGB_PAIR_STATS
      const long long st_d0_ = clock64();
#endif
      tc_fence_after();
      if (gv) e.fthr = fmaxf(e.fthr, unorder_bits(gv));
      const uint32_t taddr = lane_base + acc * (uint32_t)kPairN;
      // two register sets: the TMEM load of the next 32 columns is in flight while this one is examined
      uint32_t ra[32], rb[32];
      tc_ld32(taddr + c_lo * 32, ra);
      if (p.valid == nullptr && row0 + kPairN <= p.row_end) {
#pragma unroll 1
        for (int c = c_lo; c < c_hi; c += 2) {
          tc_wait_ld_dep(ra);
          tc_ld32(taddr + (c + 1) * 32, rb);
          epi_chunk<false, ChunkCfg<E>::rows, kShared>(e, ra, row0 + c * 32, p, lane);
          tc_wait_ld_dep(rb);
          if (c + 2 < c_hi) tc_ld32(taddr + (c + 2) * 32, ra);
          epi_chunk<false, ChunkCfg<E>::rows, kShared>(e, rb, row0 + (c + 1) * 32, p, lane);
        }
      } else {  // valid mask given, or the tile reaches past row_end
#pragma unroll 1
        for (int c = c_lo; c < c_hi; c += 2) {
          tc_wait_ld_dep(ra);
          tc_ld32(taddr + (c + 1) * 32, rb);
          epi_chunk<true, ChunkCfg<E>::rows, kShared>(e, ra, row0 + c * 32, p, lane);
          tc_wait_ld_dep(rb);
          if (c + 2 < c_hi) tc_ld32(taddr + (c + 2) * 32, ra);
          epi_chunk<true, ChunkCfg<E>::rows, kShared>(e, rb, row0 + (c + 1) * 32, p, lane);
        }
      }
      tc_fence_before();
      __syncwarp();
#ifdef RAGB_PAIR_STATS
      st_acc_[8] += clock64() - st_d0_;
#endif
      if (lane == 0) {
        if constexpr (kSolo) mbar_arrive(&tempty_bar[acc]);
        else mbar_arrive_cta(&tempty_bar[acc], 0);
      }
      if constexpr (kPar) aph ^= 1;             // this warp half always drains accumulator `half`
      else if ((acc ^= 1) == 0) aph ^= 1;
      // after the accumulator is handed back: off the MMA's critical path
      { STAT_T0();
#ifdef RAGB_PAIR_STATS
      if constexpr (kShared) epi_prune8<E, kQRes>(e, p, lane, quarter, half, &tfull_bar[acc], aph, st_acc_);
#else
      if constexpr (kShared) epi_prune8<E, kQRes>(e, p, lane, quarter, half, &tfull_bar[acc], aph);
#endif
      else epi_prune<E, kQRes>(e, p, lane);
      STAT_ADD(6); }
    }
    { STAT_T0();
    if constexpr (kShared) epi_flush<E, kQRes>(e, (int64_t)qt * kGemmM + quarter * 32, split, p, lane, half, 2, true);
    else if constexpr (kPar) epi_flush<E, kQRes>(e, (int64_t)qt * kGemmM + quarter * 32, split * 2 + half, p, lane);
    else epi_flush<E, kQRes>(e, (int64_t)qt * kGemmM + quarter * 32, split, p, lane);
    STAT_ADD(7); }
    STAT_OUT(warp == 2 && lane == 0, 5, 13);
  }

#ifdef RAGB_PAIR_STATS
  if (threadIdx.x == 0 && p.stats) p.stats[(size_t)blockIdx.x * 14] = clock64() - st_k0_;
#endif
  tc_fence_before();
  __syncthreads();
  if constexpr (!kSolo) cluster_sync_all();  // neither CTA leaves (or frees TMEM) while the other may still use the pair
  if (warp == 1) {
    tc_fence_after();
    if constexpr (kSolo) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
    else asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
  }
}

}  // namespace ragb

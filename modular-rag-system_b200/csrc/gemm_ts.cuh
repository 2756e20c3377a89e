// gemm_ts.cuh — K4 (query-stationary form): batched exact top-k with the 128 queries of a CTA held
// in TENSOR MEMORY as the A operand of tcgen05.mma, the corpus streamed through shared memory by
// TMA (multicast across a thread-block cluster), fp32 accumulators double-buffered in TMEM and the
// top-k fused into the TMEM epilogue.  sm_100a only.
//
// Why this shape (numbers from profiles/r1_gemm_ss.md): the operand-streaming kernel in
// gemm_topk.cuh re-fetches its 128-query tile for every row tile and lets the 8 CTAs that share a
// row range drift apart, so a 10M x 768 scan moved 184 GB over the L2->SM crossbar and 75 GB from
// HBM for 15.4 GB of corpus.  Here
//   * the query tile is converted once, written to TMEM with tcgen05.st (lane == query, two
//     16-bit elements per 32-bit column) and never re-read: only corpus rows cross the crossbar;
//   * the C CTAs of a cluster own C different query tiles and the SAME row range; each stage of
//     the ring is fetched from L2 by ONE of them and multicast into all C shared memories, so a
//     corpus byte leaves HBM/L2 once per cluster, and the cluster moves in lock step by
//     construction (a stage is refilled only after all C consumers have released it).
//
// The reference has no batched path (callers loop RAGVectorStore.search,
// ollama_rag_system.py:388-403 / :480-515); per query the result is the same scan.
//
// One CTA = 192 threads, 1 CTA per SM:
//   warp 0    producer : arms the stage's mbarrier, and on its turn issues the multicast TMA
//   warp 1    MMA      : tcgen05.mma kind::f16, A from TMEM, B (NT corpus rows x 64) from smem
//   warps 2-5 epilogue : first write the query tile into TMEM; then thread == query: tcgen05.ld
//                        32 columns, hierarchical max against the query's running threshold, rare
//                        survivors appended to a per-query buffer that the warp prunes together
//                        with a shuffle bitonic sort.
#pragma once
#include "gemm_topk.cuh"

namespace ragb {



__device__ __forceinline__ void tma_load_2d_mc(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1,
                                               uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(mask)
      : "memory");
}
__device__ __forceinline__ void tc_commit_mc(uint64_t* bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)),
               "h"(mask)
               : "memory");
}
__device__ __forceinline__ void tc_mma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc,
                                              uint32_t accum) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accum)
      : "memory");
}
__device__ __forceinline__ void tc_st4(uint32_t taddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "r"(a), "r"(b), "r"(c),
               "r"(d)
               : "memory");
}
__device__ __forceinline__ void tc_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// ---- warp-cooperative prune of one query's candidate buffer -----------------------------------
// All 32 lanes call it.  `base` = that query's buffer, n = candidates held (<= 64).  Returns, in
// every lane, the lane-th best key (descending; 0 = none).
__device__ __forceinline__ uint64_t warp_top32_of_64(const uint64_t* base, int n, int lane) {
  uint64_t a = lane < n ? base[lane] : 0ull;
  uint64_t b = lane + 32 < n ? base[lane + 32] : 0ull;
#pragma unroll
  for (int size = 2; size <= 32; size <<= 1) {
#pragma unroll
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      const uint64_t oa = __shfl_xor_sync(0xffffffffu, a, stride);
      const uint64_t ob = __shfl_xor_sync(0xffffffffu, b, stride);
      const bool lower = (lane & stride) == 0;
      const bool seg = (lane & size) == 0;
      // a: descending overall, b: ascending overall
      const bool a_max = (lower == seg), b_max = (lower != seg);
      a = a_max ? (a > oa ? a : oa) : (a < oa ? a : oa);
      b = b_max ? (b > ob ? b : ob) : (b < ob ? b : ob);
    }
  }
  a = a > b ? a : b;  // bitonic: the 32 largest of the 64
#pragma unroll
  for (int stride = 16; stride > 0; stride >>= 1) {
    const uint64_t oa = __shfl_xor_sync(0xffffffffu, a, stride);
    const bool lower = (lane & stride) == 0;
    a = lower ? (a > oa ? a : oa) : (a < oa ? a : oa);
  }
  return a;
}

// kTS = true : A (queries) resident in TMEM, stage = kbs x [nt rows x 64] of the corpus
// kTS = false: A streamed with the corpus, stage = kbs x ([128 queries x 64] own + [nt rows x 64] multicast);
//              all 512 TMEM columns are accumulators, so nt can be 256 (the MMA shape with the best issue
//              efficiency: ~32 fixed cycles per instruction against nt/2 of streaming)
template <bool kTS>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_topk_ts_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_x,
                    const TsParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve: stages | candidate buffers | barriers | tmem base
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t xb_bytes = (uint32_t)p.nt * 128u;       // corpus part of one k-block: nt rows x 128 B
  const uint32_t qb_bytes = kTS ? 0u : (uint32_t)kStageBytesA;  // query part: 128 rows x 128 B
  const uint32_t kb_bytes = xb_bytes + qb_bytes;
  const uint32_t stage_bytes = (uint32_t)p.kbs * kb_bytes;
  uint8_t* stage_base = smem;
  uint64_t* cand = reinterpret_cast<uint64_t*>(smem + (size_t)p.stages * stage_bytes);
  uint64_t* bars = cand + (size_t)kGemmM * kTsCandStride;
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kTsMaxStages;
  uint64_t* tfull_bar = bars + 2 * kTsMaxStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* qfull_bar = tempty_bar + 2;
  uint32_t* tmem_base_smem = reinterpret_cast<uint32_t*>(qfull_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int C = p.cluster;
  const int ci = blockIdx.x / C, rank = blockIdx.x % C;  // 1-D clusters along x: rank == %cluster_ctarank
  const int cps = p.n_qt / C;                              // clusters per row split
  const int split = ci / cps, qt = (ci % cps) * C + rank;
  const int64_t t0 = p.n_tiles * split / p.n_splits, t1 = p.n_tiles * (split + 1) / p.n_splits;
  const int spt = (p.n_kb + p.kbs - 1) / p.kbs;            // stages per tile
  const uint32_t a_cols = kTS ? (uint32_t)p.n_kb * 32u : 0u;  // TMEM columns holding the query tile
  const uint16_t mc_mask = (uint16_t)((1u << C) - 1u);

  if (threadIdx.x == 0) {
    for (int i = 0; i < p.stages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], (uint32_t)C); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 4); }
    mbar_init(qfull_bar, 4);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_base_smem)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  if (C > 1) cluster_sync_all();  // every CTA's barriers exist before any remote arrive / multicast
  tc_fence_after();
  const uint32_t tmem_base = *tmem_base_smem;

  if (warp == 0) {
    // ===== producer: the whole warp walks the ring (no divisions, state kept incrementally) =====
    const int stages = p.stages, kbs = p.kbs, n_kb = p.n_kb, nt = p.nt;
    const int q_row = qt * kGemmM;
    int s = 0, turn = 0;          // ring slot, and whose turn it is to multicast
    uint32_t ph = 0;
    for (int64_t t = t0; t < t1; ++t) {
      const int row0 = (int)(p.row_begin + t * nt);
      for (int kb0 = 0; kb0 < n_kb; kb0 += kbs) {
        mbar_wait(&empty_bar[s], ph ^ 1);  // released by all C consumers of the cluster
        const int nkb = min(kbs, n_kb - kb0);
        uint8_t* dst = stage_base + (size_t)s * stage_bytes;
        const bool mine = turn == rank;
        if (elect_one()) {
          mbar_expect_tx(&full_bar[s], (uint32_t)nkb * kb_bytes);
          for (int j = 0; j < nkb; ++j) {
            uint8_t* d = dst + (size_t)j * kb_bytes;
            if (!kTS) tma_load_2d(d, &map_q, &full_bar[s], (kb0 + j) * kGemmKB, q_row);
            if (C == 1) tma_load_2d(d + qb_bytes, &map_x, &full_bar[s], (kb0 + j) * kGemmKB, row0);
            else if (mine) tma_load_2d_mc(d + qb_bytes, &map_x, &full_bar[s], (kb0 + j) * kGemmKB, row0, mc_mask);
          }
        }
        if (++s == stages) { s = 0; ph ^= 1; }
        if (++turn == C) turn = 0;
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the whole warp walks the loop, one elected lane issues =====
    if (kTS) {
      mbar_wait(qfull_bar, 0);  // the query tile is in TMEM
      tc_fence_after();
    }
    const int stages = p.stages, kbs = p.kbs, n_kb = p.n_kb;
    const uint32_t idesc = p.idesc;
    const uint64_t desc0 = umma_desc_sw128(smem_u32(stage_base));  // +bytes/16 moves the start address
    const uint32_t stage_step = stage_bytes >> 4, kb_step = kb_bytes >> 4, q_step = qb_bytes >> 4;
    int s = 0;
    uint32_t ph = 0, acc = 0, aph = 0;
    for (int64_t t = t0; t < t1; ++t) {
      mbar_wait(&tempty_bar[acc], aph ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + a_cols + acc * (uint32_t)p.nt;
      for (int kb0 = 0; kb0 < n_kb; kb0 += kbs) {
        mbar_wait(&full_bar[s], ph);
        tc_fence_after();
        const int nkb = min(kbs, n_kb - kb0);
        if (elect_one()) {
          uint64_t sdesc = desc0 + (uint64_t)((uint32_t)s * stage_step);
          for (int j = 0; j < nkb; ++j, sdesc += kb_step) {
            const uint64_t bdesc = sdesc + q_step;
            if (kTS) {
              const uint32_t a_t = tmem_base + (uint32_t)(kb0 + j) * 32u;
#pragma unroll
              for (int k4 = 0; k4 < kGemmKB / 16; ++k4)  // K=16 per MMA: 8 TMEM columns of A, 32 bytes of B
                tc_mma_f16_ts(d_tmem, a_t + (uint32_t)k4 * 8u, bdesc + (uint64_t)(k4 * 2), idesc, ((kb0 + j) | k4) != 0);
            } else {
#pragma unroll
              for (int k4 = 0; k4 < kGemmKB / 16; ++k4)
                tc_mma_f16(d_tmem, sdesc + (uint64_t)(k4 * 2), bdesc + (uint64_t)(k4 * 2), idesc, ((kb0 + j) | k4) != 0);
            }
          }
          if (C == 1) tc_commit(&empty_bar[s]); else tc_commit_mc(&empty_bar[s], mc_mask);
          if (kb0 + kbs >= n_kb) tc_commit(&tfull_bar[acc]);  // last stage of the tile: accumulator complete
        }
        if (++s == stages) { s = 0; ph ^= 1; }
      }
      if ((acc ^= 1) == 0) aph ^= 1;
    }
  } else {
    // ===== epilogue: thread == query =====
    const int quarter = warp & 3;            // TMEM lanes this warp may touch
    const int tid = quarter * 32 + lane;     // query within the tile == TMEM lane
    const uint32_t lane_base = tmem_base + ((uint32_t)(quarter * 32) << 16);
    if (kTS) {  // query tile -> TMEM (A operand): 8 elements = 4 columns per step
      const uint4* src = reinterpret_cast<const uint4*>(static_cast<const uint16_t*>(p.qh) + ((size_t)qt * kGemmM + tid) * p.ld_q);
      const int have = (int)(p.ld_q / 8), steps = p.n_kb * 8;
      for (int c = 0; c < steps; ++c) {
        uint4 v = c < have ? src[c] : make_uint4(0, 0, 0, 0);
        tc_st4(lane_base + (uint32_t)c * 4u, v.x, v.y, v.z, v.w);
      }
      tc_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(qfull_bar);
    }
    uint64_t* my = cand + (size_t)tid * kTsCandStride;
    uint64_t* warp_cand = cand + (size_t)(quarter * 32) * kTsCandStride;
    int cnt = 0;
    // rows of the query tile past nq are zero padding: they must never take the candidate path
    const float user_fast = ((int64_t)qt * kGemmM + tid < p.nq) ? fast_bound(p.threshold, p.flags) : INFINITY;
    float fthr = user_fast;
    uint32_t tt = 0;
    for (int64_t t = t0; t < t1; ++t, ++tt) {
      const uint32_t acc = tt & 1, aph = (tt >> 1) & 1;
      const int64_t row0 = p.row_begin + t * p.nt;
      mbar_wait(&tfull_bar[acc], aph);
      tc_fence_after();
      const uint32_t taddr = lane_base + a_cols + acc * (uint32_t)p.nt;
#pragma unroll 1
      for (int c = 0; c < p.nt / 32; ++c) {
        uint32_t r[32];
        tc_ld32(taddr + c * 32, r);
        tc_wait_ld();
        float m[4];
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          float a = fmaxf(__uint_as_float(r[8 * g]), __uint_as_float(r[8 * g + 1]));
          float b = fmaxf(__uint_as_float(r[8 * g + 2]), __uint_as_float(r[8 * g + 3]));
          float cc = fmaxf(__uint_as_float(r[8 * g + 4]), __uint_as_float(r[8 * g + 5]));
          float d = fmaxf(__uint_as_float(r[8 * g + 6]), __uint_as_float(r[8 * g + 7]));
          m[g] = fmaxf(fmaxf(a, b), fmaxf(cc, d));
        }
        const float mx = fmaxf(fmaxf(m[0], m[1]), fmaxf(m[2], m[3]));
        if (__any_sync(0xffffffffu, mx >= fthr)) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            if (__any_sync(0xffffffffu, m[g] >= fthr)) {
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float raw = __uint_as_float(r[8 * g + j]);
                if (raw >= fthr) {
                  const int64_t row = row0 + c * 32 + g * 8 + j;
                  const float sc = finish_score(raw, 1.0f, p.flags);
                  bool ok = row < p.row_end && sc >= p.threshold;
                  if (ok && p.valid) ok = p.valid[row] != 0;
                  if (ok) my[cnt++] = make_key(sc, (uint32_t)row + p.id_base);
                }
              }
            }
          }
          // a chunk adds at most 32: pruning whenever more than 32 are held keeps cnt <= 64
          unsigned need = __ballot_sync(0xffffffffu, cnt > kTsCap - 32);
          while (need) {
            const int j = __ffs(need) - 1;
            need &= need - 1;
            const int n = __shfl_sync(0xffffffffu, cnt, j);
            __syncwarp();
            const uint64_t v = warp_top32_of_64(warp_cand + (size_t)j * kTsCandStride, n, lane);
            const int keep = n < p.k ? n : p.k;
            __syncwarp();
            if (lane < keep) warp_cand[(size_t)j * kTsCandStride + lane] = v;
            const uint64_t kth = __shfl_sync(0xffffffffu, v, p.k - 1);
            if (lane == j) {
              cnt = keep;
              if (n >= p.k) fthr = fmaxf(user_fast, fast_bound(key_score(kth), p.flags));
            }
            __syncwarp();
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[acc]);
    }
    // final: sort every query's survivors and write kpad keys (kpad <= 32)
    __syncwarp();
    for (int j = 0; j < 32; ++j) {
      const int n = __shfl_sync(0xffffffffu, cnt, j);
      const uint64_t v = warp_top32_of_64(warp_cand + (size_t)j * kTsCandStride, n, lane);
      const int keep = n < p.k ? n : p.k;
      const int64_t qrow = (int64_t)qt * kGemmM + quarter * 32 + j;
      if (lane < p.kpad)
        p.out[((size_t)qrow * p.n_splits + split) * p.kpad + lane] = lane < keep ? v : 0ull;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (C > 1) cluster_sync_all();  // no CTA leaves while a peer may still multicast into it
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
  }
}

}  // namespace ragb

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
// Roles per CTA (192 threads, 1 CTA/SM):
//   warp 0    producer : TMA (cp.async.bulk.tensor ... cta_group::2) of the own query k-block and the
//                        own half of the corpus k-block; bytes are accounted on the LEADER's barrier
//   warp 1    MMA      : leader CTA only; one elected lane issues tcgen05.mma.cta_group::2 and
//                        multicasts tcgen05.commit to both CTAs' barriers
//   warps 2-5 epilogue : as in gemm_ts.cuh (thread == query == TMEM lane); "accumulator drained"
//                        arrives on the leader's barrier from both CTAs (mapa + mbarrier.arrive)
//
// The reference has no batched path (callers loop RAGVectorStore.search,
// ollama_rag_system.py:388-403 / :480-515); per query the result is the same scan.
#pragma once
#include "gemm_ts.cuh"

namespace ragb {

constexpr int kPairStages = 5;
constexpr int kPairN = 256;                       // corpus rows per accumulator tile (both halves)
constexpr int kPairStageBytes = 2 * kStageBytesA; // 16 KB queries + 16 KB half corpus tile
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
__device__ __forceinline__ void mbar_arrive_cta(uint64_t* bar, uint32_t cta) {  // arrive on CTA `cta`'s copy
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t}" ::"r"(smem_u32(bar)),
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

// ---- the fused top-k epilogue on one 32-column chunk (shared with nothing else yet) -----------
struct EpiCtx {
  uint64_t* my;         // this query's candidate buffer
  uint64_t* warp_cand;  // buffer of lane 0 of this warp
  int cnt;
  float user_fast, fthr;
};
__device__ __forceinline__ void epi_chunk(EpiCtx& e, const uint32_t (&r)[32], int64_t row_c, const TsParams& p, int lane) {
  float m[4];
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    float a = fmaxf(__uint_as_float(r[8 * g]), __uint_as_float(r[8 * g + 1]));
    float b = fmaxf(__uint_as_float(r[8 * g + 2]), __uint_as_float(r[8 * g + 3]));
    float c = fmaxf(__uint_as_float(r[8 * g + 4]), __uint_as_float(r[8 * g + 5]));
    float d = fmaxf(__uint_as_float(r[8 * g + 6]), __uint_as_float(r[8 * g + 7]));
    m[g] = fmaxf(fmaxf(a, b), fmaxf(c, d));
  }
  const float mx = fmaxf(fmaxf(m[0], m[1]), fmaxf(m[2], m[3]));
  if (!__any_sync(0xffffffffu, mx >= e.fthr)) return;
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    if (__any_sync(0xffffffffu, m[g] >= e.fthr)) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float raw = __uint_as_float(r[8 * g + j]);
        if (raw >= e.fthr) {
          const int64_t row = row_c + g * 8 + j;
          const float sc = finish_score(raw, 1.0f, p.flags);
          bool ok = row < p.row_end && sc >= p.threshold;
          if (ok && p.valid) ok = p.valid[row] != 0;
          if (ok) e.my[e.cnt++] = make_key(sc, (uint32_t)row + p.id_base);
        }
      }
    }
  }
  // a chunk adds at most 32: pruning whenever more than 32 are held keeps cnt <= 64
  unsigned need = __ballot_sync(0xffffffffu, e.cnt > kTsCap - 32);
  while (need) {
    const int j = __ffs(need) - 1;
    need &= need - 1;
    const int n = __shfl_sync(0xffffffffu, e.cnt, j);
    __syncwarp();
    const uint64_t v = warp_top32_of_64(e.warp_cand + (size_t)j * kTsCandStride, n, lane);
    const int keep = n < p.k ? n : p.k;
    __syncwarp();
    if (lane < keep) e.warp_cand[(size_t)j * kTsCandStride + lane] = v;
    const uint64_t kth = __shfl_sync(0xffffffffu, v, p.k - 1);
    if (lane == j) {
      e.cnt = keep;
      if (n >= p.k) e.fthr = fmaxf(e.user_fast, fast_bound(key_score(kth), p.flags));
    }
    __syncwarp();
  }
}
// sort every query's survivors and write kpad keys (kpad <= 32)
__device__ __forceinline__ void epi_flush(EpiCtx& e, int64_t qrow0, int split, const TsParams& p, int lane) {
  __syncwarp();
  for (int j = 0; j < 32; ++j) {
    const int n = __shfl_sync(0xffffffffu, e.cnt, j);
    const uint64_t v = warp_top32_of_64(e.warp_cand + (size_t)j * kTsCandStride, n, lane);
    const int keep = n < p.k ? n : p.k;
    if (lane < p.kpad) p.out[((size_t)(qrow0 + j) * p.n_splits + split) * p.kpad + lane] = lane < keep ? v : 0ull;
  }
}

__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_topk_pair_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_x,
                      const TsParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve: stages | candidate buffers | barriers | tmem base     (identical offsets in both CTAs)
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* stage_base = smem;
  uint64_t* cand = reinterpret_cast<uint64_t*>(smem + (size_t)kPairStages * kPairStageBytes);
  uint64_t* bars = cand + (size_t)kGemmM * kTsCandStride;
  uint64_t* full_bar = bars;                     // leader's copy is the one in use
  uint64_t* empty_bar = bars + kPairStages;      // one per CTA
  uint64_t* tfull_bar = bars + 2 * kPairStages;  // one per CTA
  uint64_t* tempty_bar = tfull_bar + 2;          // leader's copy is the one in use
  uint32_t* tmem_base_smem = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = blockIdx.x & 1;               // cluster dims (2,1,1): rank == %cluster_ctarank
  const int pair = blockIdx.x >> 1;
  const int pps = p.n_qt >> 1;                   // pairs per row split
  const int split = pair / pps, qt = (pair % pps) * 2 + rank;
  const int64_t t0 = p.n_tiles * split / p.n_splits, t1 = p.n_tiles * (split + 1) / p.n_splits;
  const int n_kb = p.n_kb;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kPairStages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // the same warp of BOTH CTAs allocates for the pair
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_base_smem)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_base_smem;

  if (warp == 0) {
    // ===== producer (both CTAs) =====
    const int q_row = qt * kGemmM;
    int s = 0;
    uint32_t ph = 0;
    for (int64_t t = t0; t < t1; ++t) {
      const int row0 = (int)(p.row_begin + t * kPairN) + rank * (kPairN / 2);
      for (int kb = 0; kb < n_kb; ++kb) {
        mbar_wait(&empty_bar[s], ph ^ 1);  // the leader's commit released this slot in both CTAs
        if (elect_one()) {
          if (rank == 0) mbar_expect_tx(&full_bar[s], 2u * kPairStageBytes);  // both CTAs' bytes land on the leader
          uint8_t* d = stage_base + (size_t)s * kPairStageBytes;
          tma_load_2d_pair(d, &map_q, &full_bar[s], kb * kGemmKB, q_row);
          tma_load_2d_pair(d + kStageBytesA, &map_x, &full_bar[s], kb * kGemmKB, row0);
        }
        if (++s == kPairStages) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (leader only): M=256 across the pair =====
    if (rank == 0) {
      const uint32_t idesc = p.idesc;
      const uint64_t desc0 = umma_desc_sw128(smem_u32(stage_base));
      int s = 0;
      uint32_t ph = 0, acc = 0, aph = 0;
      for (int64_t t = t0; t < t1; ++t) {
        mbar_wait_cluster(&tempty_bar[acc], aph ^ 1);  // both CTAs' epilogues drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * (uint32_t)kPairN;
        for (int kb = 0; kb < n_kb; ++kb) {
          mbar_wait(&full_bar[s], ph);
          tc_fence_after();
          if (elect_one()) {
            const uint64_t adesc = desc0 + (uint64_t)((uint32_t)s * (kPairStageBytes >> 4));
            const uint64_t bdesc = adesc + (kStageBytesA >> 4);
#pragma unroll
            for (int k4 = 0; k4 < kGemmKB / 16; ++k4)
              tc_mma_f16_pair(d_tmem, adesc + (uint64_t)(k4 * 2), bdesc + (uint64_t)(k4 * 2), idesc, (kb | k4) != 0);
            tc_commit_pair(&empty_bar[s]);
            if (kb == n_kb - 1) tc_commit_pair(&tfull_bar[acc]);
          }
          if (++s == kPairStages) { s = 0; ph ^= 1; }
        }
        if ((acc ^= 1) == 0) aph ^= 1;
      }
    }
  } else {
    // ===== epilogue (both CTAs): thread == query =====
    const int quarter = warp & 3;
    const int tid = quarter * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(quarter * 32) << 16);
    EpiCtx e;
    e.my = cand + (size_t)tid * kTsCandStride;
    e.warp_cand = cand + (size_t)(quarter * 32) * kTsCandStride;
    e.cnt = 0;
    // rows of the query tile past nq are zero padding: they must never take the candidate path
    e.user_fast = ((int64_t)qt * kGemmM + tid < p.nq) ? fast_bound(p.threshold, p.flags) : INFINITY;
    e.fthr = e.user_fast;
    uint32_t acc = 0, aph = 0;
    for (int64_t t = t0; t < t1; ++t) {
      const int64_t row0 = p.row_begin + t * kPairN;
      mbar_wait(&tfull_bar[acc], aph);
      tc_fence_after();
      const uint32_t taddr = lane_base + acc * (uint32_t)kPairN;
#pragma unroll 1
      for (int c = 0; c < kPairN / 32; ++c) {
        uint32_t r[32];
        tc_ld32(taddr + c * 32, r);
        tc_wait_ld();
        epi_chunk(e, r, row0 + c * 32, p, lane);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cta(&tempty_bar[acc], 0);
      if ((acc ^= 1) == 0) aph ^= 1;
    }
    epi_flush(e, (int64_t)qt * kGemmM + quarter * 32, split, p, lane);
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // neither CTA leaves (or frees TMEM) while the other may still use the pair
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
  }
}

}  // namespace ragb

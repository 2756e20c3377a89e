// gemm_topk.cuh — K4: batched exact top-k = tcgen05 GEMM (scores = Q.X^T) with the top-k
// fused into the TMEM epilogue, so the score matrix never reaches HBM (sm_100a only).
//
// The reference has no batched path (callers loop RAGVectorStore.search,
// ollama_rag_system.py:388-403 / :480-515); per query the result is the same scan.
//
// Shape of one CTA (192 threads, 1 CTA per SM):
//   warp 0    TMA producer : per k-block (64 elements) one 128x64 tile of Q (A operand) and
//                            one 256x64 tile of X (B operand), 128-byte swizzled, 4-stage ring
//   warp 1    MMA issuer   : tcgen05.mma cta_group::1 kind::f16, M=128 (queries) x N=256 (rows)
//                            x K=16, fp32 accumulators in TMEM, two accumulators (512 columns)
//   warps 2-5 epilogue     : thread == query (TMEM lane).  tcgen05.ld 32 columns at a time,
//                            max-reduce, one compare against the query's running threshold;
//                            the rare survivors go to a per-thread list in shared memory.
// A CTA owns one tile of 128 queries and a contiguous range of row tiles; CTAs that share a
// row range (different query tiles) are adjacent in the grid so X is fetched from HBM once
// and served to the others from L2.
#pragma once
#include <cuda.h>

#include "common.cuh"
#include "params.h"

namespace ragb {


// ---- PTX wrappers ---------------------------------------------------------------------
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
// one lane of a converged warp; the surrounding code stays warp-uniform so that descriptors and
// addresses live in uniform registers (no per-instruction R2UR/ELECT waterfall around UTCHMMA)
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_mma_f16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum)
      : "memory");
}
__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major, 128-byte swizzled operand tile: rows of 128 bytes, 8-row groups 1024 bytes apart
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3ffff) >> 4);  // start address
  d |= (uint64_t)1 << 16;                       // leading byte offset (unused for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;             // stride byte offset
  d |= (uint64_t)1 << 46;                       // descriptor version (sm_100)
  d |= (uint64_t)2 << 61;                       // SWIZZLE_128B
  return d;
}

// ---- query conversion: fp32 -> fp16/bf16, pre-scaled by 1/|q| for the cosine metric --------
template <typename T>
__global__ void __launch_bounds__(128) convert_queries_kernel(const float* __restrict__ q, int64_t nq, int dim,
                                                              int64_t ld_out, int metric, T* __restrict__ out) {
  const int64_t row = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  T* dst = out + row * ld_out;
  if (row >= nq) {  // padding rows of the last query tile
    for (int i = lane; i < ld_out; i += 32) dst[i] = Elem<T>::from_float(0.f);
    return;
  }
  const float* src = q + row * dim;
  float scale = 1.f;
  if (metric == 1) {
    float s = 0.f;
    for (int i = lane; i < dim; i += 32) s = fmaf(src[i], src[i], s);
    s = warp_sum(s);
    scale = s > 0.f ? rsqrtf(s) : 0.f;
  }
  for (int i = lane; i < ld_out; i += 32) dst[i] = Elem<T>::from_float(i < dim ? src[i] * scale : 0.f);
}


// ---- per-thread (per-query) candidate list in shared memory: list[j*128 + tid] ---------------
struct EpiList {
  uint64_t* slot;     // &list[tid]
  int k, cnt, thr_pos;
  uint64_t thr_key;   // smallest held key once k are held, else 0
};
static __device__ __noinline__ void epi_insert(EpiList& L, uint64_t key) {
  if (L.cnt < L.k) {
    L.slot[L.cnt * kGemmM] = key;
    if (++L.cnt < L.k) return;
  } else {
    if (key <= L.thr_key) return;
    L.slot[L.thr_pos * kGemmM] = key;
  }
  uint64_t m = ~0ull;
  int pos = 0;
  for (int j = 0; j < L.k; ++j) {
    uint64_t x = L.slot[j * kGemmM];
    if (x < m) { m = x; pos = j; }
  }
  L.thr_key = m; L.thr_pos = pos;
}
// raw score a candidate must reach before the exact (slow) test is worth running
__device__ __forceinline__ float fast_bound(float t, int flags) { return ((flags & 1) && t <= 0.f) ? -INFINITY : t; }

// (a template only so that the kernel is instantiated in ONE translation unit, k_gemm.cu, although several include this header)
template <int kUnused>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_topk_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_x, const GemmParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve: stages | lists | barriers | tmem base
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* stage_base = smem;
  uint64_t* lists = reinterpret_cast<uint64_t*>(smem + (size_t)p.stages * kStageBytes);
  uint64_t* bars = lists + (size_t)p.kpad * kGemmM;
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + p.stages;
  uint64_t* tfull_bar = bars + 2 * p.stages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_base_smem = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qt = blockIdx.x % p.n_qt, split = blockIdx.x / p.n_qt;
  const int64_t t0 = p.n_tiles * split / p.n_splits, t1 = p.n_tiles * (split + 1) / p.n_splits;
  const int n_kb = (p.dim + kGemmKB - 1) / kGemmKB;

  if (threadIdx.x == 0) {
    for (int i = 0; i < p.stages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_base_smem)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_base_smem;

  if (warp == 0) {
    // ===== TMA producer =====
    uint32_t it = 0;
    for (int64_t t = t0; t < t1; ++t) {
      const int row0 = (int)(p.row_begin + t * kGemmN);
      for (int kb = 0; kb < n_kb; ++kb, ++it) {
        const int s = it % p.stages;
        const uint32_t ph = (it / p.stages) & 1;
        mbar_wait(&empty_bar[s], ph ^ 1);
        if (elect_one()) {
          mbar_expect_tx(&full_bar[s], kStageBytes);
          uint8_t* a = stage_base + (size_t)s * kStageBytes;
          tma_load_2d(a, &map_q, &full_bar[s], kb * kGemmKB, qt * kGemmM);
          tma_load_2d(a + kStageBytesA, &map_x, &full_bar[s], kb * kGemmKB, row0);
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    uint32_t it = 0, tt = 0;
    for (int64_t t = t0; t < t1; ++t, ++tt) {
      const uint32_t acc = tt & 1, aph = (tt >> 1) & 1;
      mbar_wait(&tempty_bar[acc], aph ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * kGemmN;
      for (int kb = 0; kb < n_kb; ++kb, ++it) {
        const int s = it % p.stages;
        const uint32_t ph = (it / p.stages) & 1;
        mbar_wait(&full_bar[s], ph);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(stage_base + (size_t)s * kStageBytes);
        const uint64_t adesc = umma_desc_sw128(a_addr);
        const uint64_t bdesc = umma_desc_sw128(a_addr + kStageBytesA);
        if (elect_one()) {
#pragma unroll
          for (int k4 = 0; k4 < kGemmKB / 16; ++k4)  // +32 bytes per K=16 step inside the swizzle row
            tc_mma_f16(d_tmem, adesc + (uint64_t)(k4 * 2), bdesc + (uint64_t)(k4 * 2), p.idesc, (kb | k4) != 0);
          tc_commit(&empty_bar[s]);  // frees the stage when these MMAs have read it
        }
        __syncwarp();
      }
      if (elect_one()) tc_commit(&tfull_bar[acc]);  // accumulator complete
      __syncwarp();
    }
  } else {
    // ===== epilogue: thread == query =====
    const int quarter = warp & 3;              // TMEM lanes this warp may read
    const int tid = quarter * 32 + lane;       // query within the tile == TMEM lane
    EpiList L;
    L.slot = lists + tid; L.k = p.k; L.cnt = 0; L.thr_pos = 0; L.thr_key = 0;
    for (int j = 0; j < p.kpad; ++j) L.slot[j * kGemmM] = 0;
    const float user_fast = fast_bound(p.threshold, p.flags);
    float fthr = user_fast;
    uint32_t tt = 0;
    for (int64_t t = t0; t < t1; ++t, ++tt) {
      const uint32_t acc = tt & 1, aph = (tt >> 1) & 1;
      const int64_t row0 = p.row_begin + t * kGemmN;
      mbar_wait(&tfull_bar[acc], aph);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * kGemmN;
#pragma unroll 1
      for (int c = 0; c < kGemmN / 32; ++c) {
        uint32_t r[32];
        tc_ld32(taddr + c * 32, r);
        tc_wait_ld();
        float mx = __uint_as_float(r[0]);
#pragma unroll
        for (int j = 1; j < 32; ++j) mx = fmaxf(mx, __uint_as_float(r[j]));
        if (mx >= fthr) {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float raw = __uint_as_float(r[j]);
            if (raw >= fthr) {
              const int64_t row = row0 + c * 32 + j;
              const float sc = finish_score(raw, 1.0f, p.flags);
              bool ok = row < p.row_end && sc >= p.threshold;
              if (ok && p.valid) ok = p.valid[row] != 0;
              if (ok) {
                epi_insert(L, make_key(sc, (uint32_t)row + p.id_base));
                if (L.cnt == L.k) fthr = fmaxf(user_fast, fast_bound(key_score(L.thr_key), p.flags));
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[acc]);
    }
    // sort this query's list (descending) and write it out
    for (int i = 1; i < L.cnt; ++i) {
      uint64_t x = L.slot[i * kGemmM];
      int j = i - 1;
      while (j >= 0 && L.slot[j * kGemmM] < x) { L.slot[(j + 1) * kGemmM] = L.slot[j * kGemmM]; --j; }
      L.slot[(j + 1) * kGemmM] = x;
    }
    uint64_t* dst = p.out + (((size_t)qt * kGemmM + tid) * p.n_splits + split) * p.kpad;
    for (int j = 0; j < p.kpad; ++j) dst[j] = L.slot[j * kGemmM];
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
  }
}

// ---- fp32 re-score of the winners: exact fp32 q . x for each (query, candidate) ---------------
template <typename T>
__global__ void __launch_bounds__(256) rescore_kernel(const T* __restrict__ X, int64_t ld, int dim,
                                                      const float* __restrict__ q, int64_t nq, int metric, int flags,
                                                      float threshold, IdMap ids, int kpad,
                                                      uint64_t* __restrict__ keys /* [nq][kpad]: LOCAL rows in, ids out */) {
  const int64_t w = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (w >= nq * kpad) return;
  const uint64_t key = keys[w];
  if (key == 0) return;
  const int64_t qi = w / kpad;
  const float* qv = q + qi * dim;
  const uint32_t id = key_id(key);
  const T* row = X + (size_t)id * ld;
  float dot = 0.f, qq = 0.f;
  for (int i = lane; i < dim; i += 32) {
    const float a = qv[i];
    dot = fmaf(a, Elem<T>::to_float(row[i]), dot);
    qq = fmaf(a, a, qq);
  }
  dot = warp_sum(dot);
  qq = warp_sum(qq);
  const float scale = metric == 1 ? (qq > 0.f ? rsqrtf(qq) : 0.f) : 1.f;
  const float sc = finish_score(dot, scale, flags);
  // a winner whose exact score falls just below the threshold is dropped, as the scan would
  if (lane == 0) keys[w] = sc >= threshold ? make_key(sc, map_id(ids, id)) : 0;
}

}  // namespace ragb

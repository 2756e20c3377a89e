// ingest.cuh — K2 norm_pack and K1 pool_norm_pack (sm_100a, HBM-bound).
//
// K2: RAGVectorStore._update_faiss_index, ollama_rag_system.py:379-380 —
//     e = float32(x); e = e / ||e||  -> packed fp16/bf16/fp32 row + fp32 norm.
//     A zero row packs to zeros (norm 0): search semantics :214-215.
// K1: encoder tail of SentenceTransformer.encode (call sites
//     embedding_pipeline.py:424, ollama_rag_system.py:163; sentence-transformers
//     Pooling(mean) + Normalize):  v = sum_s m_s h_s / max(sum_s m_s, 1e-9);
//     v / max(||v||, 1e-12) -> packed row + fp32 norm.  Tokens with m_s == 0 are
//     never fetched from HBM.
#pragma once
#include "common.cuh"
#include "params.h"

namespace ragb {

// 8 consecutive elements [8c, 8c+8) of a row of `dim` elements as fp32, zero filled
template <typename T>
__device__ __forceinline__ void load8(const T* row, int c, int dim, bool vec_ok, float* f) {
  const int e0 = c * 8;
  if (vec_ok && e0 + 8 <= dim) {
    if constexpr (sizeof(T) == 2) {
      uint4 v = *reinterpret_cast<const uint4*>(row + e0);
      Elem<T>::unpack(v, f);
    } else {
      float4 a = *reinterpret_cast<const float4*>(row + e0);
      float4 b = *reinterpret_cast<const float4*>(row + e0 + 4);
      f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
    }
  } else {
#pragma unroll
    for (int e = 0; e < 8; ++e) f[e] = (e0 + e < dim) ? Elem<T>::to_float(row[e0 + e]) : 0.f;
  }
}
// streaming variant for data touched once
template <typename T>
__device__ __forceinline__ void load8_stream(const T* row, int c, int dim, bool vec_ok, float* f) {
  const int e0 = c * 8;
  if (vec_ok && e0 + 8 <= dim) {
    if constexpr (sizeof(T) == 2) {
      uint4 v = ld_stream16(row + e0);
      Elem<T>::unpack(v, f);
    } else {
      uint4 a = ld_stream16(row + e0), b = ld_stream16(row + e0 + 4);
      Elem<float>::unpack(a, f); Elem<float>::unpack(b, f + 4);
    }
  } else {
#pragma unroll
    for (int e = 0; e < 8; ++e) f[e] = (e0 + e < dim) ? Elem<T>::to_float(row[e0 + e]) : 0.f;
  }
}
// 8 elements to a 16-byte aligned packed row (ld is a multiple of 8)
template <typename T>
__device__ __forceinline__ void store8(T* row, int c, const float* f) {
  if constexpr (sizeof(T) == 2) {
    union { uint4 v; T h[8]; } u;
#pragma unroll
    for (int e = 0; e < 8; ++e) u.h[e] = Elem<T>::from_float(f[e]);
    *reinterpret_cast<uint4*>(row + c * 8) = u.v;
  } else {
    *reinterpret_cast<float4*>(row + c * 8) = make_float4(f[0], f[1], f[2], f[3]);
    *reinterpret_cast<float4*>(row + c * 8 + 4) = make_float4(f[4], f[5], f[6], f[7]);
  }
}

// ---- K2 -------------------------------------------------------------------
template <typename Tin, typename Tout>
__global__ void __launch_bounds__(256) norm_pack_kernel(const Tin* __restrict__ in, int64_t n, int dim,
                                                        int64_t ld_in, Tout* __restrict__ out,
                                                        int64_t ld_out, float* __restrict__ norms,
                                                        bool vec_ok) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= n) return;
  const Tin* src = in + row * ld_in;
  Tout* dst = out + row * ld_out;
  const int chunks_in = (dim + 7) / 8, chunks_out = (int)(ld_out / 8);
  float ss = 0.f;
  for (int c = lane; c < chunks_in; c += 32) {
    float f[8];
    load8(src, c, dim, vec_ok, f);
#pragma unroll
    for (int e = 0; e < 8; ++e) ss = fmaf(f[e], f[e], ss);
  }
  ss = warp_sum(ss);
  const float nrm = sqrtf(ss);
  for (int c = lane; c < chunks_out; c += 32) {
    float f[8];
    if (c < chunks_in) load8(src, c, dim, vec_ok, f);
    else {
#pragma unroll
      for (int e = 0; e < 8; ++e) f[e] = 0.f;
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) f[e] = nrm > 0.f ? __fdiv_rn(f[e], nrm) : 0.f;
    store8(dst, c, f);
  }
  if (lane == 0 && norms) norms[row] = nrm;
}

// 8 floats -> 8 packed elements with the paired converts (cvt.rn.f16x2.f32 / cvt.rn.bf16x2.f32): half the convert
// instructions of the element-wise path
__device__ __forceinline__ void store8_packed(__half* row, int c, const float* f) {
  union { uint4 v; __half2 h[4]; } u;
#pragma unroll
  for (int e = 0; e < 4; ++e) u.h[e] = __floats2half2_rn(f[2 * e], f[2 * e + 1]);
  *reinterpret_cast<uint4*>(row + c * 8) = u.v;
}
__device__ __forceinline__ void store8_packed(__nv_bfloat16* row, int c, const float* f) {
  union { uint4 v; __nv_bfloat162 h[4]; } u;
#pragma unroll
  for (int e = 0; e < 4; ++e) u.h[e] = __floats2bfloat162_rn(f[2 * e], f[2 * e + 1]);
  *reinterpret_cast<uint4*>(row + c * 8) = u.v;
}
__device__ __forceinline__ void store8_packed(float* row, int c, const float* f) { store8(row, c, f); }

// K2, rows of at most 32*CPL chunks of 8 elements: G lanes own a row (32/G rows per warp), each lane fetches its CPL
// chunks ONCE (streaming 128-bit loads, all issued before the first is used) and keeps them in registers between the
// sum of squares and the scaling.  G is picked by the host so that G*CPL wastes as few lane slots as possible: a
// 384-element row is 48 chunks = 16 lanes x 3, two rows per warp with every lane busy — against 32 lanes x 2 with a
// quarter of the load slots idle, which together with an IEEE divide per element kept the fp16 -> fp16 case at 0.52 of
// the HBM peak (instruction-bound: half the bytes of the fp32-input case for the same instruction count).  16-bit
// outputs are scaled by the reciprocal of the norm (one divide per row; the product differs from the quotient by at
// most one fp32 ulp, which survives the rounding to 16 bits about once in 2^13 elements); fp32 outputs keep the
// divide, so that they stay the quotient numpy computes (ollama_rag_system.py:380).
template <typename Tin, typename Tout, int G, int CPL>
__global__ void __launch_bounds__(256) norm_pack_sub_kernel(const Tin* __restrict__ in, int64_t n, int dim,
                                                            int64_t ld_in, Tout* __restrict__ out,
                                                            int64_t ld_out, float* __restrict__ norms,
                                                            bool vec_ok) {
  constexpr int RPW = 32 / G;
  const int lane = threadIdx.x & 31, gl = lane % G, gr = lane / G;
  const int64_t row = ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * RPW + gr;
  const bool live = row < n;
  const int chunks_in = (dim + 7) / 8, chunks_out = (int)(ld_out / 8);
  const Tin* src = in + (live ? row : 0) * ld_in;
  float f[CPL][8];
#pragma unroll
  for (int j = 0; j < CPL; ++j) {
    const int c = gl + G * j;
    if (live && c < chunks_in) load8_stream(src, c, dim, vec_ok, f[j]);
    else {
#pragma unroll
      for (int e = 0; e < 8; ++e) f[j][e] = 0.f;
    }
  }
  float ss = 0.f;
#pragma unroll
  for (int j = 0; j < CPL; ++j)
#pragma unroll
    for (int e = 0; e < 8; ++e) ss = fmaf(f[j][e], f[j][e], ss);
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
  const float nrm = sqrtf(ss);
  if (!live) return;
  Tout* dst = out + row * ld_out;
  if constexpr (sizeof(Tout) == 2) {
    const float inv = nrm > 0.f ? __fdiv_rn(1.0f, nrm) : 0.f;
#pragma unroll
    for (int j = 0; j < CPL; ++j) {
      const int c = gl + G * j;
      if (c < chunks_out) {
#pragma unroll
        for (int e = 0; e < 8; ++e) f[j][e] *= inv;
        store8_packed(dst, c, f[j]);
      }
    }
  } else {
#pragma unroll
    for (int j = 0; j < CPL; ++j) {
      const int c = gl + G * j;
      if (c < chunks_out) {
#pragma unroll
        for (int e = 0; e < 8; ++e) f[j][e] = nrm > 0.f ? __fdiv_rn(f[j][e], nrm) : 0.f;
        store8(dst, c, f[j]);
      }
    }
  }
  if (gl == 0 && norms) norms[row] = nrm;
}

// ---- K1 -------------------------------------------------------------------
template <typename Tm> __device__ __forceinline__ float mask_weight(Tm m) { return (float)m; }
__device__ __forceinline__ void named_bar_sync_pool(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// block-wide sum, deterministic; red = 8 floats of shared memory
__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = 0.f;
  for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
  return t;
}

constexpr int kPoolThreads = 256;
// co-resident CTAs per SM and token rows in flight per thread, measured on B200 (S=256, H=384, 16 384 chunks):
//   16-bit tokens: (2 CTAs, 16 rows) 5.2 TB/s on full-length chunks, (3, 12) 5.15, (4, 8) 5.9, (5, 6) 6.04, (6, 4) 5.95;
//   fp32 tokens (two 16-byte loads per row): (3, 8) 6.4 TB/s, (4, 8) 6.25.
// More, smaller CTAs win because a CTA's phases (mask, loads, reduce, normalise, store) do not overlap within it.
template <typename Tin> struct PoolCfg {
  static constexpr int ctas = sizeof(Tin) == 2 ? 5 : 3;
  static constexpr int unroll = sizeof(Tin) == 2 ? 6 : 8;
};

// dynamic shared memory: float wt[seq] | int idx[seq] | float acc[groups][W][8] | float v[chunks*8] | float red[8] | int cnt[9]
template <typename Tin, typename Tm, typename Tout>
__global__ void __launch_bounds__(kPoolThreads, PoolCfg<Tin>::ctas) pool_norm_pack_kernel(
    const Tin* __restrict__ tok, const Tm* __restrict__ mask, int seq, int hidden,
    Tout* __restrict__ out, int64_t ld_out, float* __restrict__ norms, bool vec_ok) {
  extern __shared__ __align__(16) float sm[];
  // token rows in flight per thread
  constexpr int kPoolUnroll = PoolCfg<Tin>::unroll;
  const int chunks = (hidden + 7) / 8;
  const int W = chunks < kPoolThreads ? chunks : kPoolThreads;
  const int groups = kPoolThreads / W;
  const int seq4 = (seq + 3) & ~3;
  float* wt = sm;                                     // weights of the tokens that count, compacted
  int* idx = reinterpret_cast<int*>(wt + seq4);       // their positions
  float* accs = reinterpret_cast<float*>(idx + seq4);
  float* v = accs + (size_t)groups * W * 8;
  float* red = v + (size_t)chunks * 8;
  int* wcnt = reinterpret_cast<int*>(red + 8);        // per-warp counts of one compaction round, [8] = running total
  const int64_t b = blockIdx.x;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;

  // ---- mask: total weight, and the list of tokens with a non-zero weight (the others are never fetched) ----
  float wsum = 0.f;
  if (t == 0) wcnt[8] = 0;
  for (int s0 = 0; s0 < seq; s0 += kPoolThreads) {
    const int s = s0 + t;
    const float m = s < seq ? (mask ? mask_weight(mask[b * seq + s]) : 1.0f) : 0.f;
    wsum += m;
    const unsigned bal = __ballot_sync(0xffffffffu, m != 0.f);
    if (lane == 0) wcnt[warp] = __popc(bal);
    __syncthreads();
    int off = wcnt[8];
    for (int w2 = 0; w2 < warp; ++w2) off += wcnt[w2];
    if (m != 0.f) {
      const int pos = off + __popc(bal & ((1u << lane) - 1u));
      wt[pos] = m;
      idx[pos] = s;
    }
    __syncthreads();
    if (t == 0) {
      int tot = wcnt[8];
      for (int w2 = 0; w2 < kPoolThreads / 32; ++w2) tot += wcnt[w2];
      wcnt[8] = tot;
    }
    __syncthreads();
  }
  wsum = block_sum(wsum, red);  // includes the barriers that publish wt[], idx[] and the count
  const int nv = wcnt[8];
  const float inv_cnt = 1.0f / fmaxf(wsum, 1e-9f);

  const int g = t / W, col = t - g * W;
  const bool active = g < groups;
  const Tin* base = tok + (size_t)b * seq * hidden;
  for (int c0 = 0; c0 < chunks; c0 += W) {
    const int c = c0 + col;
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
    if (active && c < chunks) {
      constexpr int NV = sizeof(Tin) == 2 ? 1 : 2;  // 16-byte loads per 8 elements
      const bool fast = vec_ok && (c * 8 + 8 <= hidden);
      const Tin* colbase = base + c * 8;
      for (int j0 = g; j0 < nv; j0 += groups * kPoolUnroll) {
        float ws[kPoolUnroll];
        if (fast) {
          // phase 1: every load of the batch is in flight before anything waits on one
          uint4 raw[kPoolUnroll][NV];
#pragma unroll
          for (int u = 0; u < kPoolUnroll; ++u) {
            const int j = j0 + u * groups;
            const bool in = j < nv;
            ws[u] = in ? wt[j] : 0.f;
            const Tin* src = colbase + (size_t)(in ? idx[j] : 0) * hidden;
#pragma unroll
            for (int q = 0; q < NV; ++q)
              raw[u][q] = in ? ld_stream16(reinterpret_cast<const char*>(src) + 16 * q) : make_uint4(0, 0, 0, 0);
          }
          // phase 2: convert and accumulate
#pragma unroll
          for (int u = 0; u < kPoolUnroll; ++u) {
            float f[8];
            if constexpr (NV == 1) Elem<Tin>::unpack(raw[u][0], f);
            else { Elem<float>::unpack(raw[u][0], f); Elem<float>::unpack(raw[u][NV - 1], f + 4); }
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] = fmaf(ws[u], f[e], acc[e]);
          }
        } else {
#pragma unroll
          for (int u = 0; u < kPoolUnroll; ++u) {
            const int j = j0 + u * groups;
            if (j < nv) {
              float f[8];
              load8_stream(base + (size_t)idx[j] * hidden, c, hidden, false, f);
              const float wj = wt[j];
#pragma unroll
              for (int e = 0; e < 8; ++e) acc[e] = fmaf(wj, f[e], acc[e]);
            }
          }
        }
      }
    }
    __syncthreads();
    if (active) {
#pragma unroll
      for (int e = 0; e < 8; ++e) accs[((size_t)g * W + col) * 8 + e] = acc[e];
    }
    __syncthreads();
    if (t < W && c0 + t < chunks) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        float s = 0.f;
        for (int gg = 0; gg < groups; ++gg) s += accs[((size_t)gg * W + t) * 8 + e];
        v[(size_t)(c0 + t) * 8 + e] = s * inv_cnt;
      }
    }
  }
  __syncthreads();
  float ss = 0.f;
  for (int i = t; i < chunks * 8; i += kPoolThreads) ss = fmaf(v[i], v[i], ss);
  ss = block_sum(ss, red);
  const float nrm = sqrtf(ss);
  const float den = fmaxf(nrm, 1e-12f);
  Tout* dst = out + b * ld_out;
  for (int c = t; c < (int)(ld_out / 8); c += kPoolThreads) {
    float f[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) f[e] = c < chunks ? __fdiv_rn(v[(size_t)c * 8 + e], den) : 0.f;
    store8(dst, c, f);
  }
  if (t == 0 && norms) norms[b] = nrm;
}

// ---- K1, persistent form -------------------------------------------------------------------------------------------
// The kernel above gives a CTA one chunk and runs its phases one after the other (mask compaction, loads, block
// reduction, normalise, store); with ragged lengths U{16..256} the short chunks are mostly fixed cost and nothing is
// in flight while a CTA is between phases: 0.81 of the HBM peak on the bytes actually read (ncu: DRAM 63 % busy).
// Here the CTAs are persistent and warp-specialised:
//   warp 8 (producer)  draws chunks from a global counter (ragged lengths balance themselves), compacts the chunk's
//                      mask into (token, weight) pairs, and streams the VALID token rows — kPoolRingRows at a time —
//                      into a ring of shared-memory stages with 1-D TMA (cp.async.bulk, one copy per row issued by
//                      one lane each, or a single copy when the rows are adjacent), mbarrier per stage;
//   warps 0-7          wait for a stage, accumulate weight x row into registers straight from shared memory (a
//                      thread owns 8 consecutive elements; `groups` thread groups take alternate rows), hand the
//                      stage back, and at a chunk's last stage fold the groups, normalise and store the packed row.
// The producer runs ahead through the ring across chunk boundaries, so HBM stays busy while the consumers finish a
// chunk.  Same arithmetic as above (masked mean with max(sum m, 1e-9), v / max(|v|, 1e-12)).
template <typename Tin, typename Tm, typename Tout>
__global__ void __launch_bounds__(kPoolRingThreads) pool_ring_kernel(const PoolRingParams p) {
  extern __shared__ __align__(128) unsigned char psm[];
  const int C = p.hidden / 8;                       // 8-element columns per row
  const int groups = kPoolRingConsumers / C;
  const size_t rb = (size_t)p.hidden * sizeof(Tin);
  const size_t stage_bytes = (size_t)kPoolRingRows * rb;
  unsigned char* stages = psm;
  PoolStageMeta* meta = reinterpret_cast<PoolStageMeta*>(stages + (size_t)p.n_stages * stage_bytes);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(meta + p.n_stages);
  uint64_t* empty_bar = full_bar + p.n_stages;
  const int seq4 = (p.seq + 3) & ~3;
  int* p_idx = reinterpret_cast<int*>(empty_bar + p.n_stages);
  float* p_w = reinterpret_cast<float*>(p_idx + seq4);
  float* accs = p_w + seq4;
  float* red = accs + (size_t)2 * groups * C * 8;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;

  if (t == 0) {
    for (int i = 0; i < p.n_stages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], kPoolRingConsumers / 32); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (warp == kPoolRingConsumers / 32) {
    // ===== producer =====
    const Tin* tok = static_cast<const Tin*>(p.tok);
    const Tm* mask = static_cast<const Tm*>(p.mask);
    int s = 0;
    uint32_t ph = 0;
    for (;;) {
      long long b = 0;
      if (lane == 0) b = (long long)atomicAdd(p.counter, 1ull);
      b = __shfl_sync(0xffffffffu, b, 0);
      const bool done = b >= p.batch;
      int nv = 0;
      float wsum = 0.f;
      if (!done) {  // compact the mask: tokens with a non-zero weight, in order
        for (int s0 = 0; s0 < p.seq; s0 += 32) {
          const int tk = s0 + lane;
          const float m = tk < p.seq ? (mask ? (float)mask[b * p.seq + tk] : 1.0f) : 0.f;
          wsum += m;
          const unsigned bal = __ballot_sync(0xffffffffu, m != 0.f);
          if (m != 0.f) {
            const int pos = nv + __popc(bal & ((1u << lane) - 1u));
            p_idx[pos] = tk;
            p_w[pos] = m;
          }
          nv += __popc(bal);
        }
        wsum = warp_sum(wsum);
        __syncwarp();
      }
      const int n_st = done ? 1 : (nv == 0 ? 1 : (nv + kPoolRingRows - 1) / kPoolRingRows);
      for (int g = 0; g < n_st; ++g) {
        mbar_wait(&empty_bar[s], ph ^ 1);
        const int r0 = g * kPoolRingRows;
        const int n = done ? -1 : min(kPoolRingRows, nv - r0);
        PoolStageMeta* mt = &meta[s];
        if (lane == 0) { mt->n_rows = n; mt->last = (g == n_st - 1) ? 1 : 0; mt->chunk = b; mt->wsum = wsum; }
        if (n > 0 && lane < n) mt->w[lane] = p_w[r0 + lane];
        __syncwarp();
        if (n > 0) {
          const Tin* base = tok + (size_t)b * p.seq * p.hidden;
          unsigned char* dst = stages + (size_t)s * stage_bytes;
          const bool adjacent = p_idx[r0 + n - 1] - p_idx[r0] == n - 1;
          if (lane == 0) mbar_expect_tx(&full_bar[s], (uint32_t)(n * rb));
          __syncwarp();
          if (adjacent) {
            if (lane == 0) bulk_load_1d(dst, base + (size_t)p_idx[r0] * p.hidden, (uint32_t)(n * rb), &full_bar[s]);
          } else if (lane < n) {
            bulk_load_1d(dst + (size_t)lane * rb, base + (size_t)p_idx[r0 + lane] * p.hidden, (uint32_t)rb, &full_bar[s]);
          }
        } else if (lane == 0) {
          mbar_arrive(&full_bar[s]);   // an empty stage: all-masked chunk, or the end marker
        }
        if (++s == p.n_stages) { s = 0; ph ^= 1; }
      }
      if (done) break;
    }
  } else {
    // ===== consumers =====
    const int g = t / C, col = t - g * C;
    const bool active = g < groups;
    const int n_tail_warps = (C + 31) / 32;          // warps that hold the C finishing threads
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
    int s = 0, buf = 0;
    uint32_t ph = 0;
    Tout* out = static_cast<Tout*>(p.out);
    for (;;) {
      mbar_wait(&full_bar[s], ph);
      const PoolStageMeta* mt = &meta[s];
      const int n = mt->n_rows;
      if (n < 0) break;
      const int last = mt->last;
      const long long chunk = mt->chunk;
      const float wsum = mt->wsum;
      if (active) {
        const unsigned char* rows = stages + (size_t)s * stage_bytes + (size_t)col * (8 * sizeof(Tin));
        for (int r = g; r < n; r += groups) {
          const float w = mt->w[r];
          float f[8];
          if constexpr (sizeof(Tin) == 2) {
            const uint4 v = *reinterpret_cast<const uint4*>(rows + (size_t)r * rb);
            Elem<Tin>::unpack(v, f);
          } else {
            const uint4 a = *reinterpret_cast<const uint4*>(rows + (size_t)r * rb);
            const uint4 b2 = *reinterpret_cast<const uint4*>(rows + (size_t)r * rb + 16);
            Elem<float>::unpack(a, f); Elem<float>::unpack(b2, f + 4);
          }
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[e] = fmaf(w, f[e], acc[e]);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[s]);     // this warp is done with the stage
      if (++s == p.n_stages) { s = 0; ph ^= 1; }
      if (last) {
        float* ab = accs + (size_t)buf * groups * C * 8;
        if (active) {
#pragma unroll
          for (int e = 0; e < 8; ++e) ab[((size_t)g * C + col) * 8 + e] = acc[e];
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = 0.f;
        named_bar_sync_pool(1, kPoolRingConsumers);  // every group's partial sums are in shared memory
        if (warp < n_tail_warps) {
          float v[8];
          float ss = 0.f;
          const float inv_cnt = 1.0f / fmaxf(wsum, 1e-9f);
          if (t < C) {
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              float sum = 0.f;
              for (int gg = 0; gg < groups; ++gg) sum += ab[((size_t)gg * C + t) * 8 + e];
              v[e] = sum * inv_cnt;
              ss = fmaf(v[e], v[e], ss);
            }
          }
          ss = warp_sum(ss);
          if (lane == 0) red[buf * 4 + warp] = ss;
          named_bar_sync_pool(2, n_tail_warps * 32);
          float tot = 0.f;
          for (int w2 = 0; w2 < n_tail_warps; ++w2) tot += red[buf * 4 + w2];
          const float nrm = sqrtf(tot);
          const float den = fmaxf(nrm, 1e-12f);
          if (t < C) {
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = __fdiv_rn(v[e], den);
            store8_packed(out + chunk * p.ld_out, t, v);
          }
          if (t == 0 && p.norms) p.norms[chunk] = nrm;
        }
        buf ^= 1;
      }
    }
  }
}

}  // namespace ragb

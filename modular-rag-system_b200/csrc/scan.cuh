// scan.cuh — K3: bandwidth-bound exact top-k scan for a single query (sm_100a).
//
// Replaces the per-row loop of RAGVectorStore._search_sqlite
// (ollama_rag_system.py:480-515) + calculate_similarity (:199-222) for one query:
// every stored row is read exactly once from HBM with 128-bit streaming loads,
// the dot product is accumulated in fp32 against the fp32 query held in
// registers, a group of G lanes reduces one row with warp shuffles, and each warp
// keeps its own running top-k (WarpTopK) behind a threshold so that almost every
// row costs one compare.  Per-CTA results go to a small partial buffer that K5
// (merge.cuh) folds into the final k.
#pragma once
#include "common.cuh"
#include "params.h"

namespace ragb {



__device__ __forceinline__ float query_scale(const float* q, int dim, int metric, int lane) {
  if (metric != 1) return 1.0f;
  float s = 0.f;
  for (int i = lane; i < dim; i += 32) { float v = q[i]; s += v * v; }
  s = warp_sum(s);
  return s > 0.f ? rsqrtf(s) : 0.f;  // zero query scores 0.0 everywhere, :214-215
}

// fold the 8 warps' lists into warp 0's and write it out
__device__ __forceinline__ void cta_reduce_and_store(WarpTopK& tk, uint64_t* smem_lists, int warp,
                                                     int lane, uint64_t* out) {
  tk.flush(lane);
  __syncthreads();
  for (int s = 1; s < kScanWarps; s <<= 1) {
    if ((warp & (2 * s - 1)) == 0 && warp + s < kScanWarps)
      tk.merge_list(smem_lists + (size_t)(warp + s) * tk.len, lane);
    __syncthreads();
  }
  if (warp == 0) {
    for (int i = lane; i < tk.kpad; i += 32) out[i] = tk.buf[i];
    __threadfence();  // the list must be visible before this CTA takes its ticket
  }
}

// The reference's final "sorted(...)[:max_results]" (ollama_rag_system.py:511-515) without a second
// launch: every CTA takes a ticket after its list is in global memory; the CTA that draws the last
// ticket of its query reads all gridDim.x lists back (L2), folds them and decodes the k winners.
__device__ __forceinline__ void fused_final_merge(WarpTopK& tk, uint64_t* smem_lists, const ScanParams& p, int warp,
                                                  int lane) {
  __shared__ int is_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned t = atomicAdd(&p.tickets[blockIdx.y], 1u);
    is_last = (t == gridDim.x - 1);
    if (is_last) p.tickets[blockIdx.y] = 0;  // everybody else has already drawn: ready for the next launch
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  const uint64_t* base = p.out + (size_t)blockIdx.y * gridDim.x * p.kpad;
  const int P = (int)gridDim.x;
  const int64_t qi = blockIdx.y;
  const int smem_keys = kScanWarps * listbuf_len(p.kpad);
  if (P + (p.k + 1) * p.kpad <= smem_keys) {
    // Rank counting instead of merge networks.  Every list is sorted, so the k-th largest list HEAD bounds the
    // k-th best key from below (k distinct rows reach it): only the <= k lists whose head reaches it can hold a
    // winner, i.e. at most k*kpad survivors, each ranked by counting the survivors above it.
    uint64_t* heads = smem_lists;
    uint64_t* fin = heads + P;
    uint64_t* surv = fin + p.kpad;
    __shared__ unsigned long long tau_s;
    __shared__ int n_surv_s;
    for (int i = threadIdx.x; i < P; i += kScanThreads) heads[i] = __ldcg(base + (size_t)i * p.kpad);
    for (int i = threadIdx.x; i < p.kpad; i += kScanThreads) fin[i] = 0;
    if (threadIdx.x == 0) { tau_s = 0ull; n_surv_s = 0; }
    __syncthreads();
    for (int i = threadIdx.x; i < P; i += kScanThreads) {
      const uint64_t h = heads[i];
      if (h == 0) continue;
      int rank = 0;
      for (int j = 0; j < P; ++j) rank += heads[j] > h ? 1 : 0;
      if (rank == p.k - 1) tau_s = h;
    }
    __syncthreads();
    const uint64_t tau = tau_s;  // 0 when fewer than k lists hold anything: everything survives
    for (int i = threadIdx.x; i < P * p.kpad; i += kScanThreads) {
      const int l = i / p.kpad;
      if (heads[l] == 0 || heads[l] < tau) continue;
      const uint64_t key = __ldcg(base + i);
      if (key != 0 && key >= tau) surv[atomicAdd(&n_surv_s, 1)] = key;
    }
    __syncthreads();
    const int n_surv = n_surv_s;
    for (int i = threadIdx.x; i < n_surv; i += kScanThreads) {
      const uint64_t key = surv[i];
      int rank = 0;
      for (int j = 0; j < n_surv && rank < p.k; ++j) rank += surv[j] > key ? 1 : 0;
      if (rank < p.k) fin[rank] = key;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < p.kpad; i += kScanThreads) {
      const uint64_t key = fin[i];
      if (i < p.k) {
        if (p.out_scores) p.out_scores[qi * p.k + i] = key ? key_score(key) : -INFINITY;
        if (p.out_ids) p.out_ids[qi * p.k + i] = key ? decode_id(p.row_of_id, key_id(key)) : -1;
      }
      if (p.out_keys) p.out_keys[qi * p.kpad + i] = key;
      if (p.peers.n) peer_emit(p.peers, qi * p.kpad + i, key);
    }
  } else {
    tk.init(smem_lists + (size_t)warp * listbuf_len(p.kpad), p.k, p.kpad, lane);
    const int total = P * p.kpad;
    for (int i0 = warp * 32; i0 < total; i0 += kScanWarps * 32) {
      const uint64_t key = (i0 + lane < total) ? __ldcg(base + i0 + lane) : 0ull;
      tk.push(key > tk.thr, key, lane);
    }
    tk.flush(lane);
    __syncthreads();
    for (int s = 1; s < kScanWarps; s <<= 1) {
      if ((warp & (2 * s - 1)) == 0 && warp + s < kScanWarps)
        tk.merge_list(smem_lists + (size_t)(warp + s) * tk.len, lane);
      __syncthreads();
    }
    if (warp == 0) {
      for (int i = lane; i < p.k; i += 32) {
        const uint64_t key = tk.buf[i];
        if (p.out_scores) p.out_scores[qi * p.k + i] = key ? key_score(key) : -INFINITY;
        if (p.out_ids) p.out_ids[qi * p.k + i] = key ? decode_id(p.row_of_id, key_id(key)) : -1;
      }
      if (p.out_keys)
        for (int i = lane; i < p.kpad; i += 32) p.out_keys[qi * p.kpad + i] = i < p.k ? tk.buf[i] : 0;
      if (p.peers.n)
        for (int i = lane; i < p.kpad; i += 32) peer_emit(p.peers, qi * p.kpad + i, i < p.k ? tk.buf[i] : 0);
    }
  }
  if (p.peers.n) peer_signal(p.peers, gridDim.y);  // one merging CTA per query
}

// G lanes per row, CPL 16-byte chunks per lane (row = G*CPL chunks), U row-slots in flight
template <typename T, int G, int CPL, int U>
__global__ void __launch_bounds__(kScanThreads, 2) scan_topk_kernel(const ScanParams p) {
  extern __shared__ __align__(16) uint64_t smem_lists[];
  constexpr int E = Elem<T>::per16;
  constexpr int RPW = 32 / G;   // rows per warp per slot
  constexpr int R = RPW * U;    // rows per warp per iteration
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int gl = lane % G, gr = lane / G;
  const float* q = p.q + (size_t)blockIdx.y * p.ld_q;

  WarpTopK tk;
  tk.init(smem_lists + (size_t)warp * listbuf_len(p.kpad), p.k, p.kpad, lane);

  float qr[CPL][E];
#pragma unroll
  for (int j = 0; j < CPL; ++j)
#pragma unroll
    for (int e = 0; e < E; ++e) {
      int d = (gl + j * G) * E + e;
      qr[j][e] = d < p.dim ? q[d] : 0.f;
    }
  const float scale = query_scale(q, p.dim, p.metric, lane);

  const int64_t n = p.row_end - p.row_begin;
  const int64_t units = (n + R - 1) / R;
  const int64_t per_cta = (units + gridDim.x - 1) / gridDim.x;
  const int64_t u0 = (int64_t)blockIdx.x * per_cta;
  const int64_t u1 = min(units, u0 + per_cta);
  const size_t row_bytes = (size_t)p.ld * sizeof(T);
  const char* base = static_cast<const char*>(p.X) + (size_t)gl * 16;

  for (int64_t unit = u0 + warp; unit < u1; unit += kScanWarps) {
    const int64_t r0 = p.row_begin + unit * R + gr;
    uint4 v[U][CPL];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t r = r0 + u * RPW;
      if (r < p.row_end) {
        const char* rp = base + (size_t)r * row_bytes;
#pragma unroll
        for (int j = 0; j < CPL; ++j) v[u][j] = ld_stream16(rp + (size_t)j * G * 16);
      } else {
#pragma unroll
        for (int j = 0; j < CPL; ++j) v[u][j] = make_uint4(0, 0, 0, 0);
      }
    }
    float acc[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float a0 = 0.f, a1 = 0.f;
#pragma unroll
      for (int j = 0; j < CPL; ++j) {
        float f[E];
        Elem<T>::unpack(v[u][j], f);
#pragma unroll
        for (int e = 0; e < E; e += 2) { a0 = fmaf(f[e], qr[j][e], a0); a1 = fmaf(f[e + 1], qr[j][e + 1], a1); }
      }
      acc[u] = a0 + a1;
    }
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int o = G / 2; o > 0; o >>= 1) acc[u] += __shfl_xor_sync(0xffffffffu, acc[u], o);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t r = r0 + u * RPW;
      float s = finish_score(acc[u], scale, p.flags);
      bool ok = (gl == 0) && (r < p.row_end) && (s >= p.threshold);
      const uint64_t key = ok ? make_key(s, map_id(p.ids, r)) : 0ull;
      ok = ok && (key > tk.thr);
      if (ok && p.valid) ok = p.valid[r] != 0;
      tk.push(ok, key, lane);
    }
  }
  cta_reduce_and_store(tk, smem_lists, warp, lane,
                       p.out + ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * p.kpad);
  if (p.tickets) fused_final_merge(tk, smem_lists, p, warp, lane);
}

// any row length: one warp per row, chunks strided over the lanes, query in shared memory
template <typename T, int U>
__global__ void __launch_bounds__(kScanThreads, 2) scan_topk_generic_kernel(const ScanParams p) {
  extern __shared__ __align__(16) uint64_t smem_lists[];
  constexpr int E = Elem<T>::per16;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float* q = p.q + (size_t)blockIdx.y * p.ld_q;
  const int chunks = (int)((p.ld * sizeof(T)) / 16);
  float* qs = reinterpret_cast<float*>(smem_lists + (size_t)kScanWarps * listbuf_len(p.kpad));
  for (int d = threadIdx.x; d < chunks * E; d += kScanThreads) qs[d] = d < p.dim ? q[d] : 0.f;

  WarpTopK tk;
  tk.init(smem_lists + (size_t)warp * listbuf_len(p.kpad), p.k, p.kpad, lane);
  const float scale = query_scale(q, p.dim, p.metric, lane);
  __syncthreads();

  const int64_t n = p.row_end - p.row_begin;
  const int64_t units = (n + U - 1) / U;
  const int64_t per_cta = (units + gridDim.x - 1) / gridDim.x;
  const int64_t u0 = (int64_t)blockIdx.x * per_cta;
  const int64_t u1 = min(units, u0 + per_cta);
  const size_t row_bytes = (size_t)p.ld * sizeof(T);

  for (int64_t unit = u0 + warp; unit < u1; unit += kScanWarps) {
    const int64_t r0 = p.row_begin + unit * U;
    float acc[U];
#pragma unroll
    for (int u = 0; u < U; ++u) acc[u] = 0.f;
    for (int c = lane; c < chunks; c += 32) {
      uint4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        v[u] = (r0 + u < p.row_end)
                   ? ld_stream16(static_cast<const char*>(p.X) + (size_t)(r0 + u) * row_bytes + (size_t)c * 16)
                   : make_uint4(0, 0, 0, 0);
      float qv[E];
#pragma unroll
      for (int e = 0; e < E; ++e) qv[e] = qs[c * E + e];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        float f[E];
        Elem<T>::unpack(v[u], f);
#pragma unroll
        for (int e = 0; e < E; ++e) acc[u] = fmaf(f[e], qv[e], acc[u]);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) acc[u] = warp_sum(acc[u]);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t r = r0 + u;
      float s = finish_score(acc[u], scale, p.flags);
      bool ok = (lane == 0) && (r < p.row_end) && (s >= p.threshold);
      const uint64_t key = ok ? make_key(s, map_id(p.ids, r)) : 0ull;
      ok = ok && (key > tk.thr);
      if (ok && p.valid) ok = p.valid[r] != 0;
      tk.push(ok, key, lane);
    }
  }
  cta_reduce_and_store(tk, smem_lists, warp, lane,
                       p.out + ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * p.kpad);
  if (p.tickets) fused_final_merge(tk, smem_lists, p, warp, lane);
}

}  // namespace ragb

// common.cuh — sortable keys, warp-level top-k lists, load helpers (sm_100a).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace ragb {

constexpr int kStage = 64;  // per-warp staging slots for candidates awaiting a merge

// ---------------------------------------------------------------------------
// 64-bit candidate key: (order-preserving score bits << 32) | row id.
// Larger key == better candidate: higher score first, then larger (newer) row
// id (ollama_rag_system.py:474,512).  0 is "empty" and below every real key.
// ---------------------------------------------------------------------------
__host__ __device__ __forceinline__ uint32_t order_bits(float s) {
  uint32_t b;
#ifdef __CUDA_ARCH__
  b = __float_as_uint(s);
#else
  union { float f; uint32_t u; } c; c.f = s; b = c.u;
#endif
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ __forceinline__ float unorder_bits(uint32_t o) {
  uint32_t b = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
#ifdef __CUDA_ARCH__
  return __uint_as_float(b);
#else
  union { float f; uint32_t u; } c; c.u = b; return c.f;
#endif
}
__device__ __forceinline__ uint64_t make_key(float s, uint32_t id) {
  if (s != s) s = -INFINITY;  // NaN ranks last
  return ((uint64_t)order_bits(s) << 32) | id;
}
__device__ __forceinline__ float key_score(uint64_t k) { return unorder_bits((uint32_t)(k >> 32)); }
__device__ __forceinline__ uint32_t key_id(uint64_t k) { return (uint32_t)k; }

// ---------------------------------------------------------------------------
// The id carried in the low word of a key.  One device, timestamps in row order: the row itself (+ base).
// Row-sharded store (mstore.cu): rows are dealt to the devices in blocks of 2^shift rows, round-robin, so the GLOBAL row
// of a local row is arithmetic.  Non-monotone timestamps: a per-row recency RANK (position of the row in
// (timestamp, row) order — what "ORDER BY timestamp DESC" + the stable sort of ollama_rag_system.py:474,512 rank equal
// scores by), decoded back to the row through the inverse permutation when the winners are written out.
// ---------------------------------------------------------------------------
struct IdMap {
  const uint32_t* rank;  // [rows] id of every local row, or null
  uint32_t base;         // rank == null: id = global_row(row) + base
  int shift;             // >= 0: block-cyclic shard, block = 2^shift rows; < 0: global_row = row
  int n_dev, dev;
};
__host__ __device__ __forceinline__ uint32_t global_row(const IdMap& m, uint32_t r) {
  if (m.shift < 0) return r;
  return ((((r >> m.shift) * (uint32_t)m.n_dev + (uint32_t)m.dev) << m.shift) | (r & ((1u << m.shift) - 1u)));
}
__device__ __forceinline__ uint32_t map_id(const IdMap& m, int64_t row) {
  if (m.rank) return m.rank[row];
  return global_row(m, (uint32_t)row) + m.base;
}
// id -> what the caller sees (row_of_id = inverse of the rank permutation, or null)
__device__ __forceinline__ int64_t decode_id(const uint32_t* row_of_id, uint32_t id) {
  return row_of_id ? (int64_t)row_of_id[id] : (int64_t)id;
}

// ---------------------------------------------------------------------------
// Candidate exchange over peer memory (row-sharded corpus, one process per GPU).
// The kernel that produces a rank's final keys stores them straight into EVERY rank's gather buffer
// (its own included) through peer-mapped pointers — NVLink P2P stores, no collective launch — and the
// last CTA to finish publishes an epoch word to every rank with release semantics at system scope.
// The merge kernel of each rank waits on its `world` epoch words (acquire, system scope) and folds.
// ---------------------------------------------------------------------------
constexpr int kMaxPeers = 8;
struct PeerOut {
  uint64_t* keys[kMaxPeers];        // rank p's gather buffer, already offset to THIS rank's [nq][kpad] slot
  uint64_t* flag[kMaxPeers];        // rank p's arrival word of THIS rank
  unsigned* done;                   // CTA completion counter of this launch (zero between launches)
  unsigned long long epoch;
  int n;                            // ranks (0 = exchange off)
  int signal;                       // 0: store keys only (an earlier launch of a split batch; the last one signals)
};
__device__ __forceinline__ void peer_emit(const PeerOut& po, int64_t idx, uint64_t key) {
  for (int r = 0; r < po.n; ++r) po.keys[r][idx] = key;
}
// every thread of the CTA calls it after its peer_emit()s; n_ctas = CTAs that will call it in this launch
__device__ __forceinline__ void peer_signal(const PeerOut& po, unsigned n_ctas) {
  if (!po.signal) return;
  // ONE system-scope fence per CTA, by the thread that takes the ticket: the barrier orders every thread's peer stores
  // before it and fences are cumulative (the pattern of cooperative-groups grid sync).  r1 fenced in all 256 threads of
  // every finishing CTA, which made the peer exchange slower than an NCCL all-gather at 1 024 queries.
  __syncthreads();
  if (threadIdx.x == 0) {
    // signal & 2: the per-CTA fence at GPU scope only.  Causality is transitive across scopes: this CTA's peer stores
    // are ordered before its ticket at GPU scope, the last CTA observes every ticket and fences at SYSTEM scope before
    // it publishes the epoch — so a rank that acquires the epoch sees every CTA's keys — and no CTA but the last waits
    // for its NVLink stores to be acknowledged.
    if (po.signal & 2) __threadfence(); else __threadfence_system();
    const unsigned t = atomicAdd(po.done, 1u);
    if (t == n_ctas - 1) {
      *po.done = 0;  // everybody else has already arrived: ready for the next launch
      __threadfence_system();
      for (int r = 0; r < po.n; ++r)
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(po.flag[r]), "l"(po.epoch) : "memory");
    }
  }
}
// thread r < n waits until rank r's epoch word has reached `epoch`; returns false on timeout (~2 s)
__device__ __forceinline__ bool peer_wait(const uint64_t* flag, unsigned long long epoch) {
  const long long t0 = clock64();
  for (;;) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flag) : "memory");
    if (v >= epoch) return true;
    if (clock64() - t0 > 4000000000ll) return false;
    __nanosleep(200);
  }
}

// score post-processing shared by every search kernel
__device__ __forceinline__ float finish_score(float dot, float scale, int flags) {
  float s = dot * scale;
  if (flags & 1) s = fminf(fmaxf(s, 0.0f), 1.0f);  // RAGB_CLAMP01, :218
  return s;
}

// smallest power of two >= k, at least 16
__host__ __device__ inline int kpad_of(int k) {
  int p = 16;
  while (p < k) p <<= 1;
  return p;
}
// per-warp list buffer length: the sorted list [0,kpad) plus room for the
// staging area / a second list at the tail
__host__ __device__ inline int listbuf_len(int kpad) { return kpad >= kStage ? 2 * kpad : 2 * kStage; }

// ---------------------------------------------------------------------------
// Warp-level bitonic primitives on a shared-memory array of 64-bit keys.
// ---------------------------------------------------------------------------
// full bitonic sort of buf[0,n) (n power of two), ascending or descending
__device__ __forceinline__ void warp_bitonic_sort(uint64_t* buf, int n, bool descending, int lane) {
  for (int size = 2; size <= n; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      __syncwarp();
      for (int t = lane; t < (n >> 1); t += 32) {
        int lo = 2 * t - (t & (stride - 1));
        int hi = lo + stride;
        bool up = ((lo & size) == 0);  // direction of this sub-sequence
        if (descending) up = !up;
        uint64_t a = buf[lo], b = buf[hi];
        if ((a > b) == up) { buf[lo] = b; buf[hi] = a; }
      }
    }
  }
  __syncwarp();
}
// buf[0,n) is bitonic (descending run then ascending run); sort it descending
__device__ __forceinline__ void warp_bitonic_merge_desc(uint64_t* buf, int n, int lane) {
  for (int stride = n >> 1; stride > 0; stride >>= 1) {
    __syncwarp();
    for (int t = lane; t < (n >> 1); t += 32) {
      int lo = 2 * t - (t & (stride - 1));
      int hi = lo + stride;
      uint64_t a = buf[lo], b = buf[hi];
      if (a < b) { buf[lo] = b; buf[hi] = a; }
    }
  }
  __syncwarp();
}

// A warp's running top-k: buf[0,kpad) sorted descending, staging area at
// buf[len-kStage,len) filled from its start, `cnt` staged keys (warp-uniform).
struct WarpTopK {
  uint64_t* buf;
  int kpad, len, k;
  int cnt;
  uint64_t thr;  // key of the k-th best so far (0 while fewer than k are held)

  __device__ __forceinline__ void init(uint64_t* b, int k_, int kpad_, int lane) {
    buf = b; k = k_; kpad = kpad_; len = listbuf_len(kpad_); cnt = 0; thr = 0;
    for (int i = lane; i < len; i += 32) buf[i] = 0;
    __syncwarp();
  }
  // merge the staged keys into the list and refresh the threshold
  __device__ __forceinline__ void flush(int lane) {
    if (cnt == 0) return;
    uint64_t* st = buf + (len - kStage);
    __syncwarp();
    for (int i = cnt + lane; i < kStage; i += 32) st[i] = 0;
    warp_bitonic_sort(st, kStage, /*descending=*/false, lane);
    // the middle (between the list and the staging area) is already zero
    warp_bitonic_merge_desc(buf, len, lane);
    for (int i = kpad + lane; i < len; i += 32) buf[i] = 0;
    __syncwarp();
    thr = buf[k - 1];
    cnt = 0;
  }
  // warp-collective: lanes with pred push `key`; call with all 32 lanes
  __device__ __forceinline__ void push(bool pred, uint64_t key, int lane) {
    unsigned m = __ballot_sync(0xffffffffu, pred);
    if (m == 0) return;
    int n = __popc(m);
    if (cnt + n > kStage) flush(lane);  // cnt <= 32 afterwards, n <= 32
    if (pred) {
      // staged keys may be stale against a threshold tightened by the flush: harmless
      int pos = cnt + __popc(m & ((1u << lane) - 1));
      buf[len - kStage + pos] = key;
    }
    cnt += n;
    __syncwarp();
    if (cnt >= 32) flush(lane);
  }
  // fold another sorted-descending list of kpad keys (shared or global) into this one
  __device__ __forceinline__ void merge_list(const uint64_t* other, int lane) {
    flush(lane);
    for (int i = lane; i < kpad; i += 32) buf[len - 1 - i] = other[i];
    warp_bitonic_merge_desc(buf, len, lane);
    for (int i = kpad + lane; i < len; i += 32) buf[i] = 0;
    __syncwarp();
    thr = buf[k - 1];
  }
};

// ---------------------------------------------------------------------------
// element helpers
// ---------------------------------------------------------------------------
template <typename T> struct Elem;
template <> struct Elem<__half> {
  static constexpr int per16 = 8;
  __device__ static __forceinline__ void unpack(const uint4& v, float* f) {
    const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) { float2 t = __half22float2(h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
  }
  __device__ static __forceinline__ __half from_float(float x) { return __float2half_rn(x); }
  __device__ static __forceinline__ float to_float(__half x) { return __half2float(x); }
};
template <> struct Elem<__nv_bfloat16> {
  static constexpr int per16 = 8;
  __device__ static __forceinline__ void unpack(const uint4& v, float* f) {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      f[2 * i] = __uint_as_float(w[i] << 16);
      f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  __device__ static __forceinline__ __nv_bfloat16 from_float(float x) { return __float2bfloat16_rn(x); }
  __device__ static __forceinline__ float to_float(__nv_bfloat16 x) { return __bfloat162float(x); }
};
template <> struct Elem<float> {
  static constexpr int per16 = 4;
  __device__ static __forceinline__ void unpack(const uint4& v, float* f) {
    f[0] = __uint_as_float(v.x); f[1] = __uint_as_float(v.y);
    f[2] = __uint_as_float(v.z); f[3] = __uint_as_float(v.w);
  }
  __device__ static __forceinline__ float from_float(float x) { return x; }
  __device__ static __forceinline__ float to_float(float x) { return x; }
};

// streaming 128-bit load: read-only path, do not allocate in L1
__device__ __forceinline__ uint4 ld_stream16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

// ---- mbarrier (shared::cta) and 1-D bulk copy wrappers ------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// 1-D TMA: `bytes` (multiple of 16) from global to shared memory, completion counted on the mbarrier
__device__ __forceinline__ void bulk_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace ragb

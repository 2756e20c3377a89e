// params.h — kernel parameter blocks and shape constants: everything the host side needs to know about a kernel in
// order to launch it through its `const void*` entry point (kernels.h).  No kernel code here; the kernels themselves
// are in the .cuh file named above each block.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace ragb {

// ---- gemm_topk.cuh ---------------------------------------------------------------------------------------
constexpr int kGemmM = 128;       // queries per CTA
constexpr int kGemmN = 256;       // rows per MMA tile
constexpr int kGemmKB = 64;       // elements per k-block (128 bytes = one swizzle row)
constexpr int kGemmThreads = 192;
constexpr int kStageBytesA = kGemmM * kGemmKB * 2;
constexpr int kStageBytesB = kGemmN * kGemmKB * 2;
constexpr int kStageBytes = kStageBytesA + kStageBytesB;
constexpr int kGemmMaxKpad = 32;  // per-thread lists live in shared memory

// kind::f16 instruction descriptor: fp32 accumulate, A and B K-major, dense
__host__ __device__ inline uint32_t umma_idesc(bool bf16, int M, int N) {
  uint32_t d = 0;
  d |= 1u << 4;                     // D format: F32
  d |= (bf16 ? 1u : 0u) << 7;       // A format
  d |= (bf16 ? 1u : 0u) << 10;      // B format
  d |= (uint32_t)(N >> 3) << 17;
  d |= (uint32_t)(M >> 4) << 24;
  return d;
}

struct GemmParams {
  int dim;
  int64_t row_begin, row_end;  // absolute rows
  const uint8_t* valid;
  int64_t nq;
  int n_qt, n_splits;          // grid = n_qt * n_splits, block b: qt = b % n_qt, split = b / n_qt
  int64_t n_tiles;             // row tiles of kGemmN over [row_begin,row_end)
  int flags;
  float threshold;
  int k, kpad, stages;
  uint32_t id_base, idesc;
  uint64_t* out;               // [n_qt*128][n_splits][kpad] sorted keys
};

// ---- gemm_ts.cuh -----------------------------------------------------------------------------------------
constexpr int kTsCap = 64;               // candidate slots per query
constexpr int kTsCandStride = kTsCap + 1;  // +1: conflict-free for both access patterns
constexpr int kTsMaxStages = 8;

struct TsParams {
  const void* qh;       // converted queries [n_qt*128][ld_q], 16-bit elements, zero padded
  int64_t ld_q;
  int n_kb;             // 64-element k-blocks (ceil(dim/64))
  int nt;               // corpus rows per accumulator tile: 32 / 64 / 128
  int kbs;              // k-blocks per stage
  int stages;
  int cluster;          // CTAs per cluster: 1, 2, 4, 8
  int64_t row_begin, row_end;
  const uint8_t* valid;
  int64_t nq;
  int n_qt, n_splits;
  int64_t n_tiles;      // tiles of nt rows over [row_begin,row_end)
  int flags;
  float threshold;
  int k, kpad;
  uint32_t id_base, idesc;
  uint64_t* out;        // [n_qt*128][n_splits][kpad] sorted keys
  uint64_t* cand_ws;    // pair kernel, kpad > 32: [grid][128][2*kpad] candidate buffers in global memory
  float sel_slack;      // pair kernel: slack under the user threshold for the 16-bit-query selection pass
  int sync_tiles;       // pair kernel: tiles between two meetings
  int out_lists;        // pair kernel: lists per query in `out` (n_splits, or 2 * n_splits with the tile-parity epilogue)
  int soft_prune;       // pair kernel, 8 epilogue warps: prune half-full buffers while the next accumulator is not there yet
  int full_sort;        // pair kernel, k > 32: 1 = every prune sorts the whole buffer (A/B switch RAGB_PAIR_FULLSORT)
  unsigned* gthr;       // pair kernel: [n_qt*128] zeroed per launch, or null — running threshold shared by the row splits
#ifdef RAGB_PAIR_STATS
  long long* stats;     // [grid][8] cycle counters (build variant -DRAGB_PAIR_STATS only; scripts/pair_stats.sh)
#endif
  unsigned* sync_ctr;   // pair kernel: [n_splits] zeroed per launch, or null — the SM pairs that stream the same row
                        // split meet every kPairSyncTiles tiles so that they keep hitting each other's rows in L2
};

// ---- gemm_pair.cuh ---------------------------------------------------------------------------------------
constexpr int kPairStagesSmall = 5;              // ring depth next to the shared-memory candidate buffers (kpad <= 32)
constexpr int kPairStagesLarge = 6;              // kpad 64 / 128: candidates live in global memory
constexpr int kPairThreads8 = 320;                // producer + MMA + 8 epilogue warps (gemm_pair.cuh: kEpi8)
constexpr int kPairN = 256;                       // corpus rows per accumulator tile (both halves)
constexpr int kPairStageBytes = 2 * kStageBytesA; // 16 KB queries + 16 KB half corpus tile
constexpr int kPairSyncTiles = 8;                 // default tiles between two meetings of the pairs that share a row split

constexpr int kPairCapSmall = 64;
constexpr int kPairCapSmallQ = 48;  // next to a resident query tile: k <= 32 survivors + the 16 chunks one tile can add

// solo form (one CTA, cta_group::1, a batch of one query tile): 48 KB stages = 16 KB query k-block + 32 KB corpus k-block
constexpr int kSoloStagesSmall = 3;    // next to the 65 KB of shared-memory candidate buffers (kpad <= 32)
constexpr int kSoloStagesLarge = 4;    // kpad 64 / 128: candidates live in global memory

constexpr int kPairQResMaxKB = 6;
constexpr int kPairStagesSmallQ = 5;   // x 16 KB, next to 96 KB of queries and the 49 KB of candidate buffers
constexpr int kPairStagesLargeQ = 8;   // x 16 KB, next to 96 KB of queries

#ifndef RAGB_CHUNK_ROWS_LARGE
#define RAGB_CHUNK_ROWS_LARGE 16  // build-time A/B: -DRAGB_CHUNK_ROWS_LARGE=32
#endif
template <int E> struct ChunkCfg { static constexpr int rows = E == 2 ? 16 : RAGB_CHUNK_ROWS_LARGE; };

// ---- scan.cuh --------------------------------------------------------------------------------------------
constexpr int kScanThreads = 256;
constexpr int kScanWarps = kScanThreads / 32;

struct ScanParams {
  const void* X;         // packed rows
  int64_t ld;            // row stride, elements
  int dim;
  int64_t row_begin, row_end;
  const uint8_t* valid;  // optional, indexed by absolute row
  const float* q;        // [nq][ld_q] fp32
  int64_t ld_q;
  int metric, flags;
  float threshold;
  int k, kpad;
  IdMap ids;             // row -> id in the low word of a key
  const uint32_t* row_of_id;  // decode of the winners' ids (out_ids), or null
  uint64_t* out;         // [nq][gridDim.x][kpad]
  // fused final merge (the last CTA of a query to finish folds all per-CTA lists): null tickets = off
  unsigned* tickets;     // [nq], zero before the launch; the merging CTA resets its entry
  float* out_scores;     // [nq][k] or null
  int64_t* out_ids;      // [nq][k] or null
  uint64_t* out_keys;    // [nq][kpad] or null (sorted keys, for the cross-GPU merge)
  PeerOut peers;         // peers.n > 0 (needs tickets): the merging CTA also stores the keys into every rank's buffer
};

// ---- merge.cuh -------------------------------------------------------------------------------------------
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
  const uint32_t* row_of_id;  // decode of the winners' ids (out_ids), or null
  PeerOut peers;        // peers.n > 0: the kpad sorted keys also go to every rank's gather buffer
  const uint64_t* wait_flags;  // wait_n > 0: first wait until wait_flags[0..wait_n) have reached wait_epoch
  int wait_n;
  unsigned long long wait_epoch;
  int* status;          // set to 1 when the wait timed out (a rank never arrived)
};

struct FinishParams {
  const uint64_t* parts;  // [nq][n_lists][kpad] sorted chunk keys
  int n_lists, kpad, k_sel, k;
  const void* X;          // packed rows
  int64_t ld;
  int dim;
  const float* q;         // [nq][dim] fp32
  int metric, flags;
  float threshold;
  int64_t row_begin, row_end;
  const uint8_t* valid;
  IdMap ids;              // row -> id in the low word of a key
  const uint32_t* row_of_id;  // decode of the winners' ids (out_ids), or null
  float* out_scores;      // [nq][k] or null
  int64_t* out_ids;       // [nq][k] or null
  uint64_t* out_keys;     // [nq][kpad] or null
  PeerOut peers;          // peers.n > 0: the kpad sorted keys also go to every rank's gather buffer
};
constexpr int kFinishMaxChunkRows = 32;

// shared memory: lists [n_lists*kpad] | chunk [kpad] | cmax [kpad] | surv [kpad*32] | final [kpad] | misc
__host__ __device__ inline size_t finish_smem_bytes(int n_lists, int kpad) {
  return ((size_t)n_lists * kpad + (size_t)kpad * (3 + kFinishMaxChunkRows)) * sizeof(uint64_t) + 16;
}

// ---- filter.cuh ------------------------------------------------------------------------------------------
constexpr int kFilterThreads = 256;
constexpr int kFilterRowsPerCta = 2048;  // 8 rows per thread

struct FilterParams {
  const uint8_t* flags;        // [n] bit0 live, bit1 has embedding
  const uint16_t* src;         // [n] source code
  const uint32_t* perm;        // [n] recency position -> row, or null (identity)
  const uint8_t* src_allowed;  // [n_codes] or null = every source
  int n_codes;
  int64_t n;
  int64_t window;              // < 0: no window
  uint8_t* valid;              // [n] out
  unsigned* counts;            // [ceil(n / kFilterRowsPerCta)] matches per CTA
  int64_t* result;             // [0] total rows the SQL would return, [1] cut CTA, [2] matches to keep in the cut CTA
};

// ---- rank.cuh --------------------------------------------------------------------------------------------
constexpr int kRankThreads = 256;

// ---- ingest.cuh (pool_ring_kernel) ----------------------------------------------------------------------
#ifndef RAGB_POOL_RING_ROWS
#define RAGB_POOL_RING_ROWS 16  // token rows per ring stage (build-time A/B)
#endif
constexpr int kPoolRingRows = RAGB_POOL_RING_ROWS;
constexpr int kPoolRingConsumers = 256;
constexpr int kPoolRingThreads = kPoolRingConsumers + 32;
struct PoolStageMeta {
  int n_rows;      // < 0: no more chunks
  int last;        // last stage of its chunk
  long long chunk;
  float wsum;      // valid in the last stage: sum of ALL mask weights of the chunk
  float w[kPoolRingRows];
};
struct PoolRingParams {
  const void* tok;
  const void* mask;
  long long batch;
  int seq, hidden;
  void* out;
  long long ld_out;
  float* norms;
  unsigned long long* counter;  // zero before the launch
  int n_stages;
};
// dynamic shared memory: stages [n_stages][16][row bytes] | meta [n_stages] | full/empty barriers | producer list
// (int idx[seq] | float w[seq]) | accs [2][groups][C][8] | red [8]
template <typename Tin>
__host__ __device__ inline size_t pool_ring_smem(int n_stages, int seq, int hidden) {
  const size_t rb = (size_t)hidden * sizeof(Tin);
  const int C = hidden / 8, groups = kPoolRingConsumers / C;
  return (size_t)n_stages * kPoolRingRows * rb + (size_t)n_stages * sizeof(PoolStageMeta) + (size_t)n_stages * 16 +
         (size_t)((seq + 3) / 4 * 4) * 8 + (size_t)2 * groups * C * 8 * 4 + 64 + 128;
}


}  // namespace ragb

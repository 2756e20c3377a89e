// gemm_topk.cuh — K4: batched exact top-k (tcgen05 GEMM + fused top-k epilogue).  Placeholder
// state until the kernel lands; ragb200.cu falls back to one scan pass per query.
#pragma once
#include "common.cuh"

namespace ragb {
struct GemmState {
  bool ready = false;
};
inline void gemm_state_free(GemmState* g) { g->ready = false; }
}  // namespace ragb

// k_pair_e4.cu — instantiations of gemm_topk_pair_kernel<4, ...> (gemm_pair.cuh)
#include "gemm_pair.cuh"
#include "kernels.h"
namespace ragb {
const void* k_pair_e4(bool qres, bool epi8, bool par) {
  if (par)  // tile-parity epilogue (8 warps)
    return qres ? (const void*)gemm_topk_pair_kernel<4, kPairStagesLargeQ, true, true, false, true>
                : (const void*)gemm_topk_pair_kernel<4, kPairStagesLarge, false, true, false, true>;
  if (epi8)
    return qres ? (const void*)gemm_topk_pair_kernel<4, kPairStagesLargeQ, true, true> : (const void*)gemm_topk_pair_kernel<4, kPairStagesLarge, false, true>;
  return qres ? (const void*)gemm_topk_pair_kernel<4, kPairStagesLargeQ, true, false> : (const void*)gemm_topk_pair_kernel<4, kPairStagesLarge, false, false>;
}
}  // namespace ragb

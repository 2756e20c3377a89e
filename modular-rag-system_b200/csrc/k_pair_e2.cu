// k_pair_e2.cu — instantiations of gemm_topk_pair_kernel<2, ...> (gemm_pair.cuh)
#include "gemm_pair.cuh"
#include "kernels.h"
namespace ragb {
const void* k_pair_e2(bool qres, bool epi8) {
  if (epi8)
    return qres ? (const void*)gemm_topk_pair_kernel<2, kPairStagesSmallQ, true, true> : (const void*)gemm_topk_pair_kernel<2, kPairStagesSmall, false, true>;
  return qres ? (const void*)gemm_topk_pair_kernel<2, kPairStagesSmallQ, true, false> : (const void*)gemm_topk_pair_kernel<2, kPairStagesSmall, false, false>;
}
}  // namespace ragb

// k_pair_e4.cu — instantiations of gemm_topk_pair_kernel<4, ...> (gemm_pair.cuh)
#include "gemm_pair.cuh"
#include "kernels.h"
namespace ragb {
const void* k_pair_e4(bool qres) {
  return qres ? (const void*)gemm_topk_pair_kernel<4, kPairStagesLargeQ, true> : (const void*)gemm_topk_pair_kernel<4, kPairStagesLarge, false>;
}
}  // namespace ragb

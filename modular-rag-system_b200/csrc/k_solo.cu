// k_solo.cu — instantiations of gemm_topk_pair_kernel<E, stages, false, true, kSolo = true> (gemm_pair.cuh): one CTA per
// SM, tcgen05.mma.cta_group::1 M=128 x N=256, for batches of one query tile (4 <= B <= 128)
#include "gemm_pair.cuh"
#include "kernels.h"
namespace ragb {
const void* k_solo(int E) {
  switch (E) {
    case 2: return (const void*)gemm_topk_pair_kernel<2, kSoloStagesSmall, false, true, true>;
    case 4: return (const void*)gemm_topk_pair_kernel<4, kSoloStagesLarge, false, true, true>;
    case 8: return (const void*)gemm_topk_pair_kernel<8, kSoloStagesLarge, false, true, true>;
  }
  return nullptr;
}
}  // namespace ragb

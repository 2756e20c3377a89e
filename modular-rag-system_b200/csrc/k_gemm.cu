// k_gemm.cu — single-CTA tcgen05 fallbacks (gemm_ts.cuh, gemm_topk.cuh), query conversion, fp32 re-score
#include "../../include/ragb200.h"
#include "gemm_ts.cuh"
#include "kernels.h"
namespace ragb {
const void* k_ts(bool ts) { return ts ? (const void*)gemm_topk_ts_kernel<true> : (const void*)gemm_topk_ts_kernel<false>; }
const void* k_gemm_topk() { return (const void*)gemm_topk_kernel<0>; }
const void* k_convert(int dtype) {
  return dtype == RAGB_BF16 ? (const void*)convert_queries_kernel<__nv_bfloat16> : (const void*)convert_queries_kernel<__half>;
}
const void* k_rescore(int dtype) {
  return dtype == RAGB_BF16 ? (const void*)rescore_kernel<__nv_bfloat16> : (const void*)rescore_kernel<__half>;
}
}  // namespace ragb

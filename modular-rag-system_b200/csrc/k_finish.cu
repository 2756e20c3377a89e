// k_finish.cu — merge.cuh (K5 merge, K5' finish of the pair kernel), filter.cuh, rank.cuh
#include "../../include/ragb200.h"
#include "filter.cuh"
#include "merge.cuh"
#include "rank.cuh"
#include "kernels.h"
namespace ragb {
const void* k_merge() { return (const void*)merge_topk_kernel; }
const void* k_finish(int dtype, int rows) {
  if (dtype == RAGB_BF16) return rows == 16 ? (const void*)finish_chunks_kernel<__nv_bfloat16, 16> : (const void*)finish_chunks_kernel<__nv_bfloat16, 32>;
  return rows == 16 ? (const void*)finish_chunks_kernel<__half, 16> : (const void*)finish_chunks_kernel<__half, 32>;
}
const void* k_filter_mark() { return (const void*)filter_mark_kernel; }
const void* k_filter_cut() { return (const void*)filter_cut_kernel; }
const void* k_filter_window() { return (const void*)filter_window_kernel; }
const void* k_ts_check() { return (const void*)ts_check_kernel; }
const void* k_ts_fill() { return (const void*)ts_fill_kernel; }
const void* k_rank_init() { return (const void*)rank_init_kernel; }
const void* k_rank_bitonic() { return (const void*)rank_bitonic_kernel; }
const void* k_rank_scatter() { return (const void*)rank_scatter_kernel; }
}  // namespace ragb

// kernels.h — the kernels of this library as plain function pointers.
//
// Every kernel template is instantiated in exactly one translation unit (k_*.cu), which are compiled in parallel
// (Makefile) and cost minutes of ptxas time between them; the host side (ragb200.cu, mstore.cu) only sees
// `const void*` entry points and launches them with cudaLaunchKernel / cudaLaunchKernelExC, so editing host logic
// never recompiles a kernel.  The argument lists are the kernels' own (see the .cuh each getter names).
#pragma once
#include <cstddef>

namespace ragb {

// gemm_pair.cuh: gemm_topk_pair_kernel<E, stages, qres, epi8>(CUtensorMap map_q, CUtensorMap map_x, TsParams)
const void* k_pair_e2(bool qres, bool epi8);
const void* k_pair_e4(bool qres, bool epi8, bool par);   // par: tile-parity epilogue (kPar; implies 8 epilogue warps)
const void* k_pair_e8(bool qres, bool epi8, bool par);
inline const void* k_pair(int E, bool qres, bool epi8, bool par) {
  return E == 2 ? k_pair_e2(qres, epi8) : E == 4 ? k_pair_e4(qres, epi8, par) : k_pair_e8(qres, epi8, par);
}
// gemm_pair.cuh, kSolo: gemm_topk_pair_kernel<E, stages, false, true, true> — one CTA, M=128, for a batch of one query tile
const void* k_solo(int E);
// gemm_ts.cuh / gemm_topk.cuh (single-CTA fallbacks), query conversion, fp32 re-score
const void* k_ts(bool ts);                  // gemm_topk_ts_kernel<ts>(map_q, map_x, TsParams)
const void* k_gemm_topk();                  // gemm_topk_kernel(map_q, map_x, GemmParams)
const void* k_convert(int dtype);           // convert_queries_kernel<T>(q, nq, dim, ld_out, metric, out)
const void* k_rescore(int dtype);           // rescore_kernel<T>(X, ld, dim, q, nq, metric, flags, threshold, IdMap, kpad, keys)
// merge.cuh, filter.cuh, rank.cuh
const void* k_merge();                      // merge_topk_kernel(MergeParams)
const void* k_finish(int dtype, int rows);  // finish_chunks_kernel<T, rows>(FinishParams)
const void* k_filter_mark();                // filter_*_kernel(FilterParams[, n_ctas])
const void* k_filter_cut();
const void* k_filter_window();
const void* k_ts_check();                   // rank.cuh
const void* k_ts_fill();
const void* k_rank_init();
const void* k_rank_bitonic();
const void* k_rank_scatter();
// scan.cuh: scan_topk_kernel<T, G, CPL, U>(ScanParams); null when (G, cpl) has no instantiation
const void* k_scan(int dtype, int G, int cpl, int* rows_per_warp_iter);
const void* k_scan_generic(int dtype, int* rows_per_warp_iter);
// ingest.cuh
const void* k_norm_pack(int tin, int tout, int G, int cpl);   // norm_pack_sub_kernel<Tin,Tout,G,cpl>; G == 0: norm_pack_kernel (any length)
const void* k_pool(int tin, int tm, int tout);         // pool_norm_pack_kernel<Tin,Tm,Tout>
const void* k_pool_ring(int tin, int tm, int tout);    // pool_ring_kernel<Tin,Tm,Tout>(PoolRingParams)
size_t pool_ring_smem_bytes(int tin, int n_stages, int seq, int hidden);
int pool_threads();

}  // namespace ragb

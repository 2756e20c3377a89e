// k_scan.cu — scan.cuh (K3) instantiations
#include "../../include/ragb200.h"
#include "scan.cuh"
#include "kernels.h"
namespace ragb {
static constexpr int scan_unroll(int cpl) { return cpl <= 1 ? 8 : cpl == 2 ? 6 : cpl == 3 ? 4 : cpl == 4 ? 3 : 2; }
template <typename T, int G, int CPL>
static const void* one(int* rows) {
  constexpr int U = scan_unroll(CPL);
  *rows = (32 / G) * U;
  return (const void*)scan_topk_kernel<T, G, CPL, U>;
}
template <typename T, int G>
static const void* by_cpl(int cpl, int* rows) {
  switch (cpl) {
    case 1: return one<T, G, 1>(rows);
    case 2: return one<T, G, 2>(rows);
    case 3: return one<T, G, 3>(rows);
    case 4: return one<T, G, 4>(rows);
    case 6: return one<T, G, 6>(rows);
    case 8: return one<T, G, 8>(rows);
  }
  return nullptr;
}
template <typename T>
static const void* by_g(int G, int cpl, int* rows) {
  switch (G) {
    case 32: return by_cpl<T, 32>(cpl, rows);
    case 16: return by_cpl<T, 16>(cpl, rows);
    case 8: return by_cpl<T, 8>(cpl, rows);
  }
  return nullptr;
}
const void* k_scan(int dtype, int G, int cpl, int* rows) {
  switch (dtype) {
    case RAGB_F16: return by_g<__half>(G, cpl, rows);
    case RAGB_BF16: return by_g<__nv_bfloat16>(G, cpl, rows);
    case RAGB_F32: return by_g<float>(G, cpl, rows);
  }
  return nullptr;
}
const void* k_scan_generic(int dtype, int* rows) {
  *rows = 4;
  switch (dtype) {
    case RAGB_F16: return (const void*)scan_topk_generic_kernel<__half, 4>;
    case RAGB_BF16: return (const void*)scan_topk_generic_kernel<__nv_bfloat16, 4>;
    case RAGB_F32: return (const void*)scan_topk_generic_kernel<float, 4>;
  }
  return nullptr;
}
}  // namespace ragb

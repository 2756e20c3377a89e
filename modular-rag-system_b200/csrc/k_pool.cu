// k_pool.cu — ingest.cuh K1 (pool + normalise + pack) instantiations
#include "../../include/ragb200.h"
#include "ingest.cuh"
#include "kernels.h"
namespace ragb {
template <typename Tin, typename Tm>
static const void* pl2(int tout) {
  switch (tout) {
    case RAGB_F16: return (const void*)pool_norm_pack_kernel<Tin, Tm, __half>;
    case RAGB_BF16: return (const void*)pool_norm_pack_kernel<Tin, Tm, __nv_bfloat16>;
    case RAGB_F32: return (const void*)pool_norm_pack_kernel<Tin, Tm, float>;
  }
  return nullptr;
}
template <typename Tin>
static const void* pl1(int tm, int tout) {
  switch (tm) {
    case RAGB_MASK_U8: return pl2<Tin, uint8_t>(tout);
    case RAGB_MASK_I32: return pl2<Tin, int32_t>(tout);
    case RAGB_MASK_I64: return pl2<Tin, int64_t>(tout);
    case RAGB_MASK_F32: return pl2<Tin, float>(tout);
  }
  return nullptr;
}
const void* k_pool(int tin, int tm, int tout) {
  switch (tin) {
    case RAGB_F16: return pl1<__half>(tm, tout);
    case RAGB_BF16: return pl1<__nv_bfloat16>(tm, tout);
    case RAGB_F32: return pl1<float>(tm, tout);
  }
  return nullptr;
}
template <typename Tin, typename Tm>
static const void* pr2(int tout) {
  switch (tout) {
    case RAGB_F16: return (const void*)pool_ring_kernel<Tin, Tm, __half>;
    case RAGB_BF16: return (const void*)pool_ring_kernel<Tin, Tm, __nv_bfloat16>;
    case RAGB_F32: return (const void*)pool_ring_kernel<Tin, Tm, float>;
  }
  return nullptr;
}
template <typename Tin>
static const void* pr1(int tm, int tout) {
  switch (tm) {
    case RAGB_MASK_U8: return pr2<Tin, uint8_t>(tout);
    case RAGB_MASK_I32: return pr2<Tin, int32_t>(tout);
    case RAGB_MASK_I64: return pr2<Tin, int64_t>(tout);
    case RAGB_MASK_F32: return pr2<Tin, float>(tout);
  }
  return nullptr;
}
const void* k_pool_ring(int tin, int tm, int tout) {
  switch (tin) {
    case RAGB_F16: return pr1<__half>(tm, tout);
    case RAGB_BF16: return pr1<__nv_bfloat16>(tm, tout);
    case RAGB_F32: return pr1<float>(tm, tout);
  }
  return nullptr;
}
size_t pool_ring_smem_bytes(int tin, int n_stages, int seq, int hidden) {
  return tin == RAGB_F32 ? pool_ring_smem<float>(n_stages, seq, hidden) : pool_ring_smem<__half>(n_stages, seq, hidden);
}
int pool_threads() { return kPoolThreads; }
}  // namespace ragb

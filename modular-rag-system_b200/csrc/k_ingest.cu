// k_ingest.cu — ingest.cuh (K2 normalise+pack, K1 pool+normalise+pack) instantiations
#include "../../include/ragb200.h"
#include "ingest.cuh"
#include "kernels.h"
namespace ragb {
template <typename Tin, typename Tout>
static const void* np2(int cpl) {
  switch (cpl) {
    case 0: return (const void*)norm_pack_kernel<Tin, Tout>;
    case 1: return (const void*)norm_pack_reg_kernel<Tin, Tout, 1, 1>;
    case 2: return (const void*)norm_pack_reg_kernel<Tin, Tout, 2, 1>;
    case 3: return (const void*)norm_pack_reg_kernel<Tin, Tout, 3, 1>;
    case 4: return (const void*)norm_pack_reg_kernel<Tin, Tout, 4, 1>;
  }
  return nullptr;
}
template <typename Tin>
static const void* np1(int tout, int cpl) {
  switch (tout) {
    case RAGB_F16: return np2<Tin, __half>(cpl);
    case RAGB_BF16: return np2<Tin, __nv_bfloat16>(cpl);
    case RAGB_F32: return np2<Tin, float>(cpl);
  }
  return nullptr;
}
const void* k_norm_pack(int tin, int tout, int cpl) {
  switch (tin) {
    case RAGB_F16: return np1<__half>(tout, cpl);
    case RAGB_BF16: return np1<__nv_bfloat16>(tout, cpl);
    case RAGB_F32: return np1<float>(tout, cpl);
  }
  return nullptr;
}
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
int pool_threads() { return kPoolThreads; }
}  // namespace ragb

// k_ingest.cu — ingest.cuh K2 (normalise + pack) instantiations
#include "../../include/ragb200.h"
#include "ingest.cuh"
#include "kernels.h"
namespace ragb {
template <typename Tin, typename Tout>
static const void* np2(int G, int cpl) {
  if (G == 0) return (const void*)norm_pack_kernel<Tin, Tout>;
#define NP(g, c) if (G == g && cpl == c) return (const void*)norm_pack_sub_kernel<Tin, Tout, g, c>
  NP(32, 1); NP(32, 2); NP(32, 3); NP(32, 4);
  NP(16, 1); NP(16, 2); NP(16, 3);
  NP(8, 1); NP(8, 2); NP(8, 3);
#undef NP
  return nullptr;
}
template <typename Tin>
static const void* np1(int tout, int G, int cpl) {
  switch (tout) {
    case RAGB_F16: return np2<Tin, __half>(G, cpl);
    case RAGB_BF16: return np2<Tin, __nv_bfloat16>(G, cpl);
    case RAGB_F32: return np2<Tin, float>(G, cpl);
  }
  return nullptr;
}
const void* k_norm_pack(int tin, int tout, int G, int cpl) {
  switch (tin) {
    case RAGB_F16: return np1<__half>(tout, G, cpl);
    case RAGB_BF16: return np1<__nv_bfloat16>(tout, G, cpl);
    case RAGB_F32: return np1<float>(tout, G, cpl);
  }
  return nullptr;
}
}  // namespace ragb

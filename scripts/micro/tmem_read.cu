// TMEM read rate microbenchmark (sm_100a): W warps per lane quarter issue tcgen05.ld.32x32b.x32 in a loop, `depth`
// loads in flight per warp.  Prints bytes per cycle per SM.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
// -o /tmp/tmem_read scripts/micro/tmem_read.cu ; run on one SM (grid 1) and on all SMs (grid 148).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tc_wait_ld_dep(uint32_t* r) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                 "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]),
                 "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]),
                 "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
               :: "memory");
}

template <int kDepth>
__global__ void __launch_bounds__(512, 1) tmem_read_kernel(int iters, unsigned* sink, long long* cycles) {
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&tmem_base_s)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = tmem_base_s + ((uint32_t)((warp & 3) * 32) << 16);
  uint32_t ra[32], rb[32];
  unsigned acc = 0;
  __syncthreads();
  const long long t0 = clock64();
  if (kDepth == 1) {
    for (int i = 0; i < iters; ++i) {
      tc_ld32(base + ((i * 32) & 511), ra);
      tc_wait_ld_dep(ra);
      acc ^= ra[0] ^ ra[31];
    }
  } else {
    tc_ld32(base, ra);
    for (int i = 0; i < iters; i += 2) {
      tc_wait_ld_dep(ra);
      tc_ld32(base + (((i + 1) * 32) & 511), rb);
      acc ^= ra[0] ^ ra[31];
      tc_wait_ld_dep(rb);
      tc_ld32(base + (((i + 2) * 32) & 511), ra);
      acc ^= rb[0] ^ rb[31];
    }
    tc_wait_ld_dep(ra);
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
  if (acc == 0x12345u) sink[threadIdx.x] = acc;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base_s), "r"(512));
}

int main() {
  unsigned* sink; long long* cyc;
  cudaMalloc(&sink, 4096); cudaMallocManaged(&cyc, 8);
  const int iters = 1 << 14;
  for (int grid : {1, 148})
    for (int warps : {4, 8, 16})
      for (int depth : {1, 2}) {
        for (int rep = 0; rep < 2; ++rep) {
          if (depth == 1) tmem_read_kernel<1><<<grid, warps * 32>>>(iters, sink, cyc);
          else tmem_read_kernel<2><<<grid, warps * 32>>>(iters, sink, cyc);
          cudaError_t e = cudaDeviceSynchronize();
          if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
        }
        const double bytes = (double)warps * iters * 32 * 32 * 4;
        printf("grid %3d  warps %2d (%d per lane quarter)  loads in flight %d : %lld cycles, %.1f B/cycle/SM, %.0f cycles per x32 load per warp\n",
               grid, warps, warps / 4, depth, *cyc, bytes / (double)*cyc, (double)*cyc / iters);
      }
  return 0;
}

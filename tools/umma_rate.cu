// Microbenchmark: issue rate of tcgen05.mma (M=128, N, K=16, bf16) for different shared-memory operand layouts.
// Data content is irrelevant (zeros); only the descriptor geometry matters.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o gpurun_out/umma_rate tools/umma_rate.cu && gpurun_out/umma_rate
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "../gnn_uds_b200/csrc/umma.cuh"

using namespace gu;

struct RateP { int variant, iters, N, ksteps; long long* out; };

__device__ __forceinline__ uint64_t desc_raw(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout << 61;
  return d;
}

__global__ void __launch_bounds__(128, 1) rate_kernel(RateP p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x;
  for (int i = tid; i < 160 * 1024 / 16; i += 128) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  if (tid < 32) tc::tmem_alloc(&slot, 512);
  if (tid == 32) { tc::mbar_init(&bar, 1); tc::mbar_fence_init(); }
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = slot;
  if (tid == 0) {
    const uint32_t a = tc::smem_u32(smem), b = a + 80 * 1024;
    const int v = p.variant;
    const uint32_t idesc = tc::idesc_bf16(128, p.N, 0, (v == 3) ? 1 : 0);
    if (v >= 4) {            // descriptors precomputed, loop unrolled: nothing but the MMA instructions in the timed loop
      uint64_t ad[8], bd[8];
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
        ad[ks] = desc_raw(a + 2u * ks * 2048u, 2048u, 128u, 0);
        bd[ks] = desc_raw(b + 2u * ks * (uint32_t)p.N * 16u, (uint32_t)p.N * 16u, 128u, 0);
      }
      const long long t0 = clock64();
      for (int it = 0; it < p.iters; ++it) {
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          if (v == 4) {
            tc::umma(tmem, ad[ks], bd[ks], idesc, 1u);
          } else {
            asm volatile("tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, 1;" ::"r"(tmem), "l"(ad[ks]), "l"(bd[ks]), "r"(idesc) : "memory");
          }
        }
      }
      const long long t1 = clock64();
      tc::umma_commit(&bar);
      tc::mbar_wait(&bar, 0);
      const long long t2 = clock64();
      if (blockIdx.x == 0) { p.out[0] = t1 - t0; p.out[1] = t2 - t0; }
    } else {
    const long long t0 = clock64();
    for (int it = 0; it < p.iters; ++it) {
      for (int ks = 0; ks < p.ksteps; ++ks) {
        uint64_t ad, bd;
        if (v == 0) {          // no swizzle, panel stride multiple of 128 B
          ad = desc_raw(a + 2u * ks * 2048u, 2048u, 128u, 0);
          bd = desc_raw(b + 2u * ks * (uint32_t)p.N * 16u, (uint32_t)p.N * 16u, 128u, 0);
        } else if (v == 1) {   // no swizzle, panel stride = 16 mod 128
          ad = desc_raw(a + 2u * ks * 2064u, 2064u, 128u, 0);
          bd = desc_raw(b + 2u * ks * ((uint32_t)p.N * 16u + 16u), (uint32_t)p.N * 16u + 16u, 128u, 0);
        } else if (v == 2) {   // 128-byte swizzle, K-major: rows of 128 B, 8-row atoms of 1024 B, K step = 32 B inside the row
          ad = desc_raw(a + (uint32_t)(ks >> 2) * 128u * 128u + (uint32_t)(ks & 3) * 32u, 16u, 1024u, 2);
          bd = desc_raw(b + (uint32_t)(ks >> 2) * (uint32_t)p.N * 128u + (uint32_t)(ks & 3) * 32u, 16u, 1024u, 2);
        } else {               // A no swizzle K-major (stride mult of 128), B MN-major no swizzle (aggregation form)
          ad = desc_raw(a + 2u * ks * 2048u, 2048u, 128u, 0);
          bd = desc_raw(b + (uint32_t)ks * 256u, 128u, 2304u, 0);
        }
        tc::umma(tmem, ad, bd, idesc, (it | ks) != 0);
      }
    }
    const long long t1 = clock64();
    tc::umma_commit(&bar);
    tc::mbar_wait(&bar, 0);
    const long long t2 = clock64();
    if (blockIdx.x == 0) { p.out[0] = t1 - t0; p.out[1] = t2 - t0; }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (tid < 32) tc::tmem_free(tmem, 512);
}

int main() {
  long long* d;
  cudaMalloc(&d, 16);
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const char* names[6] = {"noswz pstride%128==0", "noswz pstride%128==16", "swizzle128 K-major", "noswz A K-major x B MN-major", "noswz precomputed desc", "noswz precomputed, imm predicate"};
  for (int grid : {1})
    for (int N : {64, 128, 256})
      for (int v : {0, 4, 5}) {
        RateP p{v, 200, N, 8, d};
        rate_kernel<<<grid, 128, 200 * 1024>>>(p);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[2];
        cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        const double n = 200.0 * 8;
        printf("grid %3d N %3d %-32s issue %.1f cyc/mma  complete %.1f cyc/mma  (nominal %d)  %s\n", grid, N, names[v], h[0] / n,
               h[1] / n, N / 2, e == cudaSuccess ? "" : cudaGetErrorString(e));
      }
  return 0;
}

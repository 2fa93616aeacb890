// Microbenchmark 3: why are the aggregation MMAs of sp_agg_kernel at ~120 cycles per K step instead of 64?
//   (a) panel stride of the SWIZZLE_NONE K-major A operand (1920 / 2048 / 2176 bytes),
//   (b) bursts: `nk` MMAs, one commit, wait on the mbarrier, repeat (the kernel's pattern) against a continuous stream.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o gpurun_out/umma_rate3 tools/umma_rate3.cu && gpurun_out/umma_rate3
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "../gnn_uds_b200/csrc/umma.cuh"

using namespace gu;

struct P3 { int psA, lb, N, nk, burst, iters; long long* out; };

__device__ __forceinline__ uint64_t desc_raw(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout << 61;
  return d;
}

__global__ void __launch_bounds__(128, 1) k3(P3 p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  for (int i = tid; i < 200 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u + (uint32_t)i % 7u;   // fp16 ~1.0
  if (warp == 0) tc::tmem_alloc(&slot, 512);
  if (tid == 32) { tc::mbar_init(&bar, 1); tc::mbar_fence_init(); }
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = slot;
  if (warp == 1) {
    const uint32_t a = tc::smem_u32(smem), b = a + 96 * 1024;
    const uint32_t idesc = tc::idesc_bf16(128, p.N, 0, 0);
    const uint32_t N = (uint32_t)p.N;
    long long t0 = clock64();
    uint32_t ph = 0;
    for (int it = 0; it < p.iters; ++it) {
      if (tc::elect_one()) {
        for (int ks = 0; ks < p.nk; ++ks) {
          const uint64_t ad = desc_raw(a + 2u * ks * (uint32_t)p.psA, (uint32_t)p.psA, 128u, 0);
          const uint64_t bd = p.lb == 0 ? desc_raw(b + 2u * ks * N * 16u, N * 16u, 128u, 0)
                                        : desc_raw(b + (uint32_t)(ks >> 2) * N * 128u + (uint32_t)(ks & 3) * 32u, 16u, 1024u, 2);
          tc::umma(tmem, ad, bd, idesc, 1u);
        }
        if (p.burst) tc::umma_commit(&bar);
      }
      __syncwarp();
      if (p.burst) { tc::mbar_wait(&bar, ph); ph ^= 1u; tc::tc_fence_after(); }
    }
    if (!p.burst) { if (tc::elect_one()) tc::umma_commit(&bar); __syncwarp(); tc::mbar_wait(&bar, 0); }
    long long t1 = clock64();
    if (blockIdx.x == 0 && (tid & 31) == 0) p.out[0] = t1 - t0;
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_free(tmem, 512);
}

int main() {
  long long* d;
  cudaMalloc(&d, 64);
  cudaFuncSetAttribute(k3, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int N : {64, 128, 144})
    for (int ps : {2048, 1920, 2176})
      for (int lb : {0, 1})
        for (int burst : {0, 1})
          for (int nk : {8, 12}) {
            if (lb == 1 && N == 144) continue;
            P3 p{ps, lb, N, nk, burst, 200, d};
            k3<<<1, 128, 200 * 1024>>>(p);
            cudaError_t e = cudaDeviceSynchronize();
            long long h[1];
            cudaMemcpy(h, d, 8, cudaMemcpyDeviceToHost);
            printf("N %3d A none-K pstride %4d  B %-7s %s nk %2d: %.1f cyc per MMA, %.0f per %s %s\n", N, ps, lb ? "sw128-K" : "none-K",
                   burst ? "burst " : "stream", nk, h[0] / (200.0 * nk), h[0] / 200.0, burst ? "burst (incl. commit + wait)" : "group",
                   e == cudaSuccess ? "" : cudaGetErrorString(e));
            if (e != cudaSuccess) return 1;
          }
  return 0;
}

// Microbenchmark 2: tcgen05.mma rate by operand layout (SWIZZLE_NONE vs SWIZZLE_128B, K-major vs MN-major B, one or two
// accumulators) and tcgen05.ld drain throughput by warp count.  Data content is irrelevant (zeros).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o gpurun_out/umma_rate2 tools/umma_rate2.cu && gpurun_out/umma_rate2
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "../gnn_uds_b200/csrc/umma.cuh"

using namespace gu;

struct RateP { int la, lb, N, M, two_acc, iters; long long* out; };

__device__ __forceinline__ uint64_t desc_raw(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout << 61;
  return d;
}

// la / lb: 0 = SWIZZLE_NONE K-major panels, 1 = SWIZZLE_128B K-major, 2 = SWIZZLE_128B MN-major (B only)
__global__ void __launch_bounds__(128, 1) rate_kernel(RateP p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x;
  for (int i = tid; i < 200 * 1024 / 16; i += 128) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  if (tid < 32) tc::tmem_alloc(&slot, 512);
  if (tid == 32) { tc::mbar_init(&bar, 1); tc::mbar_fence_init(); }
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = slot;
  if (tid == 0) {
    const uint32_t a = tc::smem_u32(smem), b = a + 64 * 1024;
    const uint32_t idesc = tc::idesc_bf16(p.M, p.N, 0, p.lb == 2 ? 1 : 0);
    uint64_t ad[8], bd[8];
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
      // A: 128 rows x 128 K
      if (p.la == 0) ad[ks] = desc_raw(a + 2u * ks * 2048u, 2048u, 128u, 0);
      else ad[ks] = desc_raw(a + (uint32_t)(ks >> 2) * 16384u + (uint32_t)(ks & 3) * 32u, 16u, 1024u, 2);
      // B: N rows x 128 K
      const uint32_t N = (uint32_t)p.N;
      if (p.lb == 0) bd[ks] = desc_raw(b + 2u * ks * N * 16u, N * 16u, 128u, 0);
      else if (p.lb == 1) bd[ks] = desc_raw(b + (uint32_t)(ks >> 2) * N * 128u + (uint32_t)(ks & 3) * 32u, 16u, 1024u, 2);
      else bd[ks] = desc_raw(b + (uint32_t)ks * 2048u, 128u * 128u /* next 64-column atom: a [128 K x 128 B] tile */, 1024u, 2);
    }
    const long long t0 = clock64();
    for (int it = 0; it < p.iters; ++it) {
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) tc::umma(tmem + ((p.two_acc && (ks & 1)) ? 256u : 0u), ad[ks], bd[ks], idesc, 1u);
    }
    const long long t1 = clock64();
    tc::umma_commit(&bar);
    tc::mbar_wait(&bar, 0);
    const long long t2 = clock64();
    if (blockIdx.x == 0) { p.out[0] = t1 - t0; p.out[1] = t2 - t0; }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (tid < 32) tc::tmem_free(tmem, 512);
}

// drain throughput: nw warps, warp w reads lane quarter w&3, column block (w>>2)*ncol; x32 loads
__global__ void __launch_bounds__(512, 1) drain_kernel(int nw, int ncol, int iters, int pack_store, long long* out) {
  __shared__ uint32_t slot;
  __shared__ __align__(16) uint8_t sink[512 * 64];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (warp == 0) tc::tmem_alloc(&slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = slot;
  long long t0 = 0, t1 = 0;
  float keep = 0.f;
  if (warp < nw) {
    const uint32_t ta = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((warp >> 2) * ncol);
    t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      for (int c0 = 0; c0 < ncol; c0 += 32) {
        float v[32];
        tc::tmem_ld32(ta + (uint32_t)c0, v);
        tc::tmem_ld_wait();
        if (pack_store) {
#pragma unroll
          for (int j = 0; j < 4; ++j) *reinterpret_cast<uint4*>(sink + tid * 64 + j * 16) = tc::pack8_f16(v + 8 * j);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) keep += v[j];
        }
      }
    }
    t1 = clock64();
  }
  if (keep == 123.456f) out[7] = 1;
  if (tid == 0) { out[0] = t1 - t0; }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_free(tmem, 512);
}

int main() {
  long long* d;
  cudaMalloc(&d, 64);
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const char* ln[3] = {"none-K", "sw128-K", "sw128-MN"};
  for (int M : {128, 64})
    for (int N : {64, 128, 136, 256})
      for (int la : {0, 1})
        for (int lb : {0, 1, 2})
          for (int two : {0, 1}) {
            if (lb == 2 && N > 128) continue;
            if (M == 64 && (la == 0 || two)) continue;
            RateP p{la, lb, N, M, two, 200, d};
            rate_kernel<<<1, 128, 200 * 1024>>>(p);
            cudaError_t e = cudaDeviceSynchronize();
            long long h[2];
            cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
            printf("M %3d N %3d A %-8s B %-8s acc %d  issue %.1f  complete %.1f cyc/mma (floor %d) %s\n", M, N, ln[la], ln[lb], two + 1,
                   h[0] / 1600.0, h[1] / 1600.0, N / 2, e == cudaSuccess ? "" : cudaGetErrorString(e));
            if (e != cudaSuccess) return 1;
          }
  for (int ps : {0, 1})
    for (int nw : {1, 4, 8, 16})
      for (int ncol : {128}) {
        drain_kernel<<<1, 512>>>(nw, ncol, 100, ps, d);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[1];
        cudaMemcpy(h, d, 8, cudaMemcpyDeviceToHost);
        const double bytes = (double)nw * 32 * ncol * 4 * 100;
        printf("drain warps %2d cols %3d store %d: %.0f cyc per pass, %.1f B/clk %s\n", nw, ncol, ps, h[0] / 100.0, bytes / h[0],
               e == cudaSuccess ? "" : cudaGetErrorString(e));
      }
  return 0;
}

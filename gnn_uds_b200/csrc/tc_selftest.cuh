// Single-CTA self-test of the tcgen05 building blocks (umma.cuh): panel layout written by
// threads, weight image fetched with a bulk copy, K-major and MN-major operand views, shifted
// start addresses, TMEM epilogue.  Driven by gnnuds_debug_umma_selftest (tests/test_umma_gpu.py).
#pragma once
#include "umma.cuh"

namespace gu {

struct SelfTestP {
  const float* act;        // (Ra, Ca) fp32 row-major
  const uint16_t* img;     // bf16 K-major image of mat (Rm rows, Km cols), panel stride Rm*16 bytes
  float* out;              // (128, Nout)
  int Ra, Ca, Rm, Km, shift, swap, mode, Nout;
};

__global__ void __launch_bounds__(256) umma_selftest_kernel(SelfTestP p) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar_w, bar_mma;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t ps_act = (uint32_t)(p.Ra + 1) * 16u;
  const uint32_t ps_img = (uint32_t)p.Rm * 16u;
  uint8_t* act_s = smem;
  const uint32_t act_bytes = ((uint32_t)(p.Ca / 8) * ps_act + 127u) & ~127u;
  uint8_t* img_s = smem + act_bytes;
  const uint32_t img_bytes = (uint32_t)(p.Km / 8) * ps_img;

  if (warp == 0) tc::tmem_alloc(&tmem_slot, 256);
  if (tid == 32) {
    tc::mbar_init(&bar_w, 1);
    tc::mbar_init(&bar_mma, 1);
    tc::mbar_fence_init();
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = tmem_slot;

  if (tid == 0) {
    tc::mbar_expect_tx(&bar_w, img_bytes);
    tc::bulk_g2s(img_s, p.img, img_bytes, &bar_w);
  }
  // activations -> bf16 panels (row-parallel 16-byte stores)
  for (int idx = tid; idx < p.Ra * (p.Ca / 8); idx += blockDim.x) {
    const int r = idx % p.Ra, c = idx / p.Ra;
    float v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = p.act[(size_t)r * p.Ca + c * 8 + i];
    *reinterpret_cast<uint4*>(act_s + (size_t)c * ps_act + (size_t)r * 16) = tc::pack8_bf16(v);
  }
  tc::fence_async_smem();
  __syncthreads();

  if (tid == 0) {
    tc::mbar_wait(&bar_w, 0);
    tc::tc_fence_after();
    const uint32_t a_act = tc::smem_u32(act_s), a_img = tc::smem_u32(img_s);
    auto mk = [&](uint32_t addr, uint32_t lbo, uint32_t sbo) {
      return p.swap ? tc::smem_desc(addr, sbo, lbo) : tc::smem_desc(addr, lbo, sbo);
    };
    if (p.mode == 0) {
      const uint32_t idesc = tc::idesc_bf16(128, p.Nout, 0, 0);
      for (int ks = 0; ks < p.Ca / 16; ++ks) {
        const uint64_t ad = mk(a_act + (uint32_t)p.shift * 16u + (uint32_t)(2 * ks) * ps_act, ps_act, 128u);
        const uint64_t bd = mk(a_img + (uint32_t)(2 * ks) * ps_img, ps_img, 128u);
        tc::umma(tmem, ad, bd, idesc, ks > 0);
      }
    } else {
      const uint32_t idesc = tc::idesc_bf16(128, p.Nout, 0, 1);
      for (int ks = 0; ks < p.Km / 16; ++ks) {
        const uint64_t ad = mk(a_img + (uint32_t)(2 * ks) * ps_img, ps_img, 128u);
        const uint64_t bd = mk(a_act + (uint32_t)ks * 256u, 128u, ps_act);
        tc::umma(tmem, ad, bd, idesc, ks > 0);
      }
    }
    tc::umma_commit(&bar_mma);
  }
  __syncwarp();
  tc::mbar_wait(&bar_mma, 0);
  tc::tc_fence_after();
  {
    const int q = warp & 3, half = warp >> 2;
    const int row = q * 32 + lane;
    for (int c0 = half * 32; c0 < p.Nout; c0 += 64) {
      float v[32];
      tc::tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
      tc::tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (c0 + i < p.Nout) p.out[(size_t)row * p.Nout + c0 + i] = v[i];
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_free(tmem, 256);
}

}  // namespace gu

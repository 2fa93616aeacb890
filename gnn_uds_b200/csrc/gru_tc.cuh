// GRU recurrence on tensor cores (sm_100a), K6 of SURVEY 8(a): keras GRU v2, reset_after=True, gate order z,r,h
//
//   mh_t = h_{t-1} @ R + b1          tcgen05.mma, M = 128 sequences, N = 3H, K = H, bf16 operands, fp32 in TMEM
//   z = sigmoid(mx_z + mh_z)   r = sigmoid(mx_r + mh_r)   hh = tanh(mx_h + r * mh_h)   h_t = z*h_{t-1} + (1-z)*hh
//
// mx = x_t @ K + b0 comes from tc_rows_kernel (bf16 rows (b,t,n) x 3H).  A CTA owns 128 consecutive sequences
// q = b*Nn + n for all Ts steps; the fp32 state h lives in registers and enters the MMA as TWO bf16 operands
// (hi = bf16(h), lo = bf16(h - hi), K-major panels in shared memory, both multiplied with the same R image), so the
// recurrent product sees ~16 mantissa bits of the state and the rounding does not build up over the time steps.
// R^T is resident in shared memory as two K-major weight images (bf16(R) and the bf16 residual R - bf16(R));
// the step's product is h_hi*R_hi + h_lo*R_hi + h_hi*R_lo, 3*H/16 MMAs into one accumulator.
//   warps 0..15 gate group: thread = (sequence row = TMEM lane, quarter of the H units).  Prefetches the step's mx
//               slices, waits for the MMAs, reads its 3 x H/4 accumulator columns, does the gate math, writes the new
//               state (bf16) into the A panels and the output row, then signals the MMA issuer.
//   warp  16    MMA issuer (one thread): H/16 MMAs per step.
// One CTA per SM (register budget of the gate math); persistent over tiles of 128 sequences.
#pragma once
#include "kernels.cuh"
#include "tc_gemm.cuh"
#include "umma.cuh"

namespace gu {

constexpr int GRU_GATE_WARPS = 16;
constexpr int GRU_THREADS = (GRU_GATE_WARPS + 1) * 32;   // 544

struct GruTcP {
  const bf16* mx;        // (Bc, Ts, Nn, 3H)
  const uint8_t* Rimg;   // two bf16 K-major images of R^T (high part, residual): element (n, k) at (k/8)*G*16 + n*16 + (k%8)*2
  const float* b1;       // (3H) recurrent bias
  bf16* out;             // (Bc, Td, Nn, H) last Td steps
  long long nseq;        // Bc*Nn
  int Nn, Ts, Td, H;
  unsigned n_tiles;
};

template <int H>
__global__ void __launch_bounds__(GRU_THREADS, 1) gru_tc_kernel(const __grid_constant__ GruTcP p) {
  constexpr int G = 3 * H, HH = H / 4;               // a gate thread owns HH units
  constexpr uint32_t R_BYTES = (uint32_t)H * G * 2u;  // weight image
  constexpr uint32_t A_BYTES = 128u * H * 2u;         // state panels: H/8 panels of 128 rows x 16 bytes
  constexpr uint32_t PSTRIDE_A = 128u * 16u, PSTRIDE_W = (uint32_t)G * 16u;
  static_assert(H % 64 == 0 && G <= 256, "hidden size must be a multiple of 64 with 3H <= 256");
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* Rs = smem;
  uint8_t* As = smem + 2 * R_BYTES;
  uint8_t* Al = As + A_BYTES;                         // low-order part of the state
  float* b1s = reinterpret_cast<float*>(Al + A_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(b1s + G);
  uint64_t* h_ready = bars;        // gate group wrote the A panels (and drained the accumulators)
  uint64_t* mma_done = bars + 1;   // the step's MMAs completed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;   // warp-uniform to the compiler

  if (warp == 0) tc::tmem_alloc(tmem_slot, 256);
  if (tid == 32) {
    tc::mbar_init(h_ready, GRU_GATE_WARPS * 32);
    tc::mbar_init(mma_done, 1);
    tc::mbar_fence_init();
  }
  for (uint32_t i = tid * 16u; i < 2 * R_BYTES; i += GRU_THREADS * 16u)
    *reinterpret_cast<uint4*>(Rs + i) = *reinterpret_cast<const uint4*>(p.Rimg + i);
  for (int i = tid; i < G; i += GRU_THREADS) b1s[i] = p.b1[i];
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  tc::pdl_sync();
  const uint32_t tmem = *tmem_slot;
  uint32_t ph_h = 0, ph_m = 0;     // phases of h_ready (MMA thread) / mma_done (gate threads)

  if (warp == GRU_GATE_WARPS) {
    // ================================ MMA issuer ================================
    {
      const uint32_t idesc = tc::idesc_bf16(128, G, 0, 0);
      const uint64_t a0 = tc::smem_desc(tc::smem_u32(As), PSTRIDE_A, 128u), w0 = tc::smem_desc(tc::smem_u32(Rs), PSTRIDE_W, 128u);
      const uint64_t l0 = tc::smem_desc(tc::smem_u32(Al), PSTRIDE_A, 128u), w1 = tc::smem_desc(tc::smem_u32(Rs + R_BYTES), PSTRIDE_W, 128u);
      const uint64_t a_step = (uint64_t)((2u * PSTRIDE_A) >> 4), w_step = (uint64_t)((2u * PSTRIDE_W) >> 4);
      for (unsigned tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        for (int t = 0; t < p.Ts; ++t) {
          tc::mbar_wait(h_ready, ph_h); ph_h ^= 1u;
          tc::tc_fence_after();
          if (tc::elect_one()) {      // warp-uniform issue code, one elected lane (tc::elect_one)
#pragma unroll
            for (int ks = 0; ks < H / 16; ++ks) tc::umma(tmem, a0 + (uint64_t)ks * a_step, w0 + (uint64_t)ks * w_step, idesc, ks != 0);
#pragma unroll
            for (int ks = 0; ks < H / 16; ++ks) tc::umma(tmem, l0 + (uint64_t)ks * a_step, w0 + (uint64_t)ks * w_step, idesc, 1u);
#pragma unroll
            for (int ks = 0; ks < H / 16; ++ks) tc::umma(tmem, a0 + (uint64_t)ks * a_step, w1 + (uint64_t)ks * w_step, idesc, 1u);
            tc::umma_commit(mma_done);
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ================================ gate group ================================
    const int q4 = warp & 3, half = warp >> 2;
    const int row = q4 * 32 + lane;                  // TMEM lane = sequence of the tile
    const int u0 = half * HH;
    const uint32_t ta = tmem + ((uint32_t)(q4 * 32) << 16);
    uint8_t* arow = As + (size_t)row * 16;
    uint8_t* lrow = Al + (size_t)row * 16;
    for (unsigned tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      const long long q = (long long)tile * 128 + row;
      const bool live = q < p.nseq;
      const long long b = live ? q / p.Nn : 0;
      const int n = live ? (int)(q % p.Nn) : 0;
      float h[HH];
#pragma unroll
      for (int k = 0; k < HH; ++k) h[k] = 0.f;
      // zero initial state -> A panels
#pragma unroll
      for (int c = 0; c < HH / 8; ++c) {
        *reinterpret_cast<uint4*>(arow + (size_t)(u0 / 8 + c) * PSTRIDE_A) = make_uint4(0u, 0u, 0u, 0u);
        *reinterpret_cast<uint4*>(lrow + (size_t)(u0 / 8 + c) * PSTRIDE_A) = make_uint4(0u, 0u, 0u, 0u);
      }
      tc::fence_async_smem();
      mbar_arrive(h_ready);
      for (int t = 0; t < p.Ts; ++t) {
        // this step's mx slices (z, r, h gates of the thread's units) before the wait: 3 x HH bf16
        uint4 mz[HH / 8], mr[HH / 8], mc[HH / 8];
        const bf16* mxr = p.mx + ((size_t)(b * p.Ts + t) * p.Nn + n) * G + u0;
#pragma unroll
        for (int c = 0; c < HH / 8; ++c) {
          mz[c] = live ? *reinterpret_cast<const uint4*>(mxr + c * 8) : make_uint4(0u, 0u, 0u, 0u);
          mr[c] = live ? *reinterpret_cast<const uint4*>(mxr + H + c * 8) : make_uint4(0u, 0u, 0u, 0u);
          mc[c] = live ? *reinterpret_cast<const uint4*>(mxr + 2 * H + c * 8) : make_uint4(0u, 0u, 0u, 0u);
        }
        tc::mbar_wait(mma_done, ph_m); ph_m ^= 1u;
        tc::tc_fence_after();
        const int td = t - (p.Ts - p.Td);
        bf16* orow = p.out + ((size_t)(b * p.Td + (td > 0 ? td : 0)) * p.Nn + n) * H + u0;
#pragma unroll
        for (int c0 = 0; c0 < HH; c0 += 8) {         // 8 units at a time: 3 x 8 accumulator columns (register budget)
          float az[8], ar[8], ac[8];
          tc::tmem_ld8(ta + (uint32_t)(u0 + c0), az);
          tc::tmem_ld8(ta + (uint32_t)(H + u0 + c0), ar);
          tc::tmem_ld8(ta + (uint32_t)(2 * H + u0 + c0), ac);
          tc::tmem_ld_wait();
          float xz[8], xr[8], xc[8];
          tc::unpack8_bf16(mz[c0 / 8], xz);
          tc::unpack8_bf16(mr[c0 / 8], xr);
          tc::unpack8_bf16(mc[c0 / 8], xc);
          float hn[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int u = u0 + c0 + k;
            const float z = __fdividef(1.f, 1.f + __expf(-(xz[k] + az[k] + b1s[u])));
            const float r = __fdividef(1.f, 1.f + __expf(-(xr[k] + ar[k] + b1s[H + u])));
            const float pre = xc[k] + r * (ac[k] + b1s[2 * H + u]);
            const float e2 = __expf(-2.f * fabsf(pre));          // tanh(pre) = sign(pre) * (1 - e2) / (1 + e2)
            const float hh = copysignf(__fdividef(1.f - e2, 1.f + e2), pre);
            hn[k] = z * h[c0 + k] + (1.f - z) * hh;
            h[c0 + k] = hn[k];
          }
          const uint4 p0 = tc::pack8_bf16(hn);
          *reinterpret_cast<uint4*>(arow + (size_t)((u0 + c0) / 8) * PSTRIDE_A) = p0;
          float hi[8];
          tc::unpack8_bf16(p0, hi);
#pragma unroll
          for (int k = 0; k < 8; ++k) hi[k] = hn[k] - hi[k];
          *reinterpret_cast<uint4*>(lrow + (size_t)((u0 + c0) / 8) * PSTRIDE_A) = tc::pack8_bf16(hi);
          if (live && td >= 0) *reinterpret_cast<uint4*>(orow + c0) = p0;
        }
        tc::tc_fence_before();
        tc::fence_async_smem();
        if (t + 1 < p.Ts) mbar_arrive(h_ready);      // the next tile's zero state does the arrival after the last step
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_free(tmem, 256);
}

// ---------------------------------------------------------------------------------------------------------------
// LSTM recurrence on tensor cores: keras LSTM v2, gate order i,f,c,o (emulator.py:160-161)
//   z_t = mx_t + h_{t-1} @ R     (mx = x_t @ K + b from tc_rows_kernel, 4H wide; R^T resident as hi + residual images)
//   i,f,o = sigmoid(z_i,z_f,z_o)   c_t = f*c_{t-1} + i*tanh(z_c)   h_t = o*tanh(c_t)
// Same roles and barriers as gru_tc_kernel; N = 4H = 256 accumulator columns, fp32 cell and hidden state in
// registers, the hidden state enters the MMA as hi + lo 16-bit operands.  8 units per accumulator read to stay
// inside the register budget (4 gates x 8 columns live).
struct LstmTcP {
  const bf16* mx;        // (Bc, Ts, Nn, 4H)
  const uint8_t* Rimg;   // two K-major images of R^T (high part, residual), as GruTcP::Rimg with G = 4H
  bf16* out;             // (Bc, Td, Nn, H)
  long long nseq;
  int Nn, Ts, Td, H;
  unsigned n_tiles;
};

__device__ __forceinline__ float fast_sigmoid(float v) { return __fdividef(1.f, 1.f + __expf(-v)); }
__device__ __forceinline__ float fast_tanh(float v) {
  const float e2 = __expf(-2.f * fabsf(v));
  return copysignf(__fdividef(1.f - e2, 1.f + e2), v);
}

template <int H>
__global__ void __launch_bounds__(GRU_THREADS, 1) lstm_tc_kernel(const __grid_constant__ LstmTcP p) {
  constexpr int G = 4 * H, HH = H / 4;
  constexpr uint32_t R_BYTES = (uint32_t)H * G * 2u;
  constexpr uint32_t A_BYTES = 128u * H * 2u;
  constexpr uint32_t PSTRIDE_A = 128u * 16u, PSTRIDE_W = (uint32_t)G * 16u;
  static_assert(H % 64 == 0 && G <= 256, "hidden size must be a multiple of 64 with 4H <= 256");
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* Rs = smem;
  uint8_t* As = smem + 2 * R_BYTES;
  uint8_t* Al = As + A_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(Al + A_BYTES);
  uint64_t* h_ready = bars;
  uint64_t* mma_done = bars + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;   // warp-uniform to the compiler

  if (warp == 0) tc::tmem_alloc(tmem_slot, 256);
  if (tid == 32) {
    tc::mbar_init(h_ready, GRU_GATE_WARPS * 32);
    tc::mbar_init(mma_done, 1);
    tc::mbar_fence_init();
  }
  for (uint32_t i = tid * 16u; i < 2 * R_BYTES; i += GRU_THREADS * 16u)
    *reinterpret_cast<uint4*>(Rs + i) = *reinterpret_cast<const uint4*>(p.Rimg + i);
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  tc::pdl_sync();
  const uint32_t tmem = *tmem_slot;
  uint32_t ph_h = 0, ph_m = 0;

  if (warp == GRU_GATE_WARPS) {
    {
      const uint32_t idesc = tc::idesc_bf16(128, G, 0, 0);
      const uint64_t a0 = tc::smem_desc(tc::smem_u32(As), PSTRIDE_A, 128u), w0 = tc::smem_desc(tc::smem_u32(Rs), PSTRIDE_W, 128u);
      const uint64_t l0 = tc::smem_desc(tc::smem_u32(Al), PSTRIDE_A, 128u), w1 = tc::smem_desc(tc::smem_u32(Rs + R_BYTES), PSTRIDE_W, 128u);
      const uint64_t a_step = (uint64_t)((2u * PSTRIDE_A) >> 4), w_step = (uint64_t)((2u * PSTRIDE_W) >> 4);
      for (unsigned tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        for (int t = 0; t < p.Ts; ++t) {
          tc::mbar_wait(h_ready, ph_h); ph_h ^= 1u;
          tc::tc_fence_after();
          if (tc::elect_one()) {      // warp-uniform issue code, one elected lane (tc::elect_one)
#pragma unroll
            for (int ks = 0; ks < H / 16; ++ks) tc::umma(tmem, a0 + (uint64_t)ks * a_step, w0 + (uint64_t)ks * w_step, idesc, ks != 0);
#pragma unroll
            for (int ks = 0; ks < H / 16; ++ks) tc::umma(tmem, l0 + (uint64_t)ks * a_step, w0 + (uint64_t)ks * w_step, idesc, 1u);
#pragma unroll
            for (int ks = 0; ks < H / 16; ++ks) tc::umma(tmem, a0 + (uint64_t)ks * a_step, w1 + (uint64_t)ks * w_step, idesc, 1u);
            tc::umma_commit(mma_done);
          }
          __syncwarp();
        }
      }
    }
  } else {
    const int q4 = warp & 3, part = warp >> 2;
    const int row = q4 * 32 + lane;
    const int u0 = part * HH;
    const uint32_t ta = tmem + ((uint32_t)(q4 * 32) << 16);
    uint8_t* arow = As + (size_t)row * 16;
    uint8_t* lrow = Al + (size_t)row * 16;
    for (unsigned tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      const long long q = (long long)tile * 128 + row;
      const bool live = q < p.nseq;
      const long long b = live ? q / p.Nn : 0;
      const int n = live ? (int)(q % p.Nn) : 0;
      float cst[HH];
#pragma unroll
      for (int k = 0; k < HH; ++k) cst[k] = 0.f;
#pragma unroll
      for (int c = 0; c < HH / 8; ++c) {
        *reinterpret_cast<uint4*>(arow + (size_t)(u0 / 8 + c) * PSTRIDE_A) = make_uint4(0u, 0u, 0u, 0u);
        *reinterpret_cast<uint4*>(lrow + (size_t)(u0 / 8 + c) * PSTRIDE_A) = make_uint4(0u, 0u, 0u, 0u);
      }
      tc::fence_async_smem();
      mbar_arrive(h_ready);
      for (int t = 0; t < p.Ts; ++t) {
        uint4 mg[4][HH / 8];                         // this step's mx slices of the 4 gates, loaded before the wait
        const bf16* mxr = p.mx + ((size_t)(b * p.Ts + t) * p.Nn + n) * G + u0;
#pragma unroll
        for (int g = 0; g < 4; ++g)
#pragma unroll
          for (int c = 0; c < HH / 8; ++c)
            mg[g][c] = live ? *reinterpret_cast<const uint4*>(mxr + g * H + c * 8) : make_uint4(0u, 0u, 0u, 0u);
        tc::mbar_wait(mma_done, ph_m); ph_m ^= 1u;
        tc::tc_fence_after();
        const int td = t - (p.Ts - p.Td);
        bf16* orow = p.out + ((size_t)(b * p.Td + (td > 0 ? td : 0)) * p.Nn + n) * H + u0;
#pragma unroll
        for (int c0 = 0; c0 < HH; c0 += 8) {
          float ai[8], af[8], ac[8], ao[8];
          tc::tmem_ld8(ta + (uint32_t)(u0 + c0), ai);
          tc::tmem_ld8(ta + (uint32_t)(H + u0 + c0), af);
          tc::tmem_ld8(ta + (uint32_t)(2 * H + u0 + c0), ac);
          tc::tmem_ld8(ta + (uint32_t)(3 * H + u0 + c0), ao);
          tc::tmem_ld_wait();
          float xi[8], xf[8], xc[8], xo[8];
          tc::unpack8_bf16(mg[0][c0 / 8], xi); tc::unpack8_bf16(mg[1][c0 / 8], xf);
          tc::unpack8_bf16(mg[2][c0 / 8], xc); tc::unpack8_bf16(mg[3][c0 / 8], xo);
          float hn[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const float ig = fast_sigmoid(xi[k] + ai[k]);
            const float fg = fast_sigmoid(xf[k] + af[k]);
            const float cg = fast_tanh(xc[k] + ac[k]);
            const float og = fast_sigmoid(xo[k] + ao[k]);
            const float cn = fg * cst[c0 + k] + ig * cg;
            cst[c0 + k] = cn;
            hn[k] = og * fast_tanh(cn);
          }
          const uint4 p0 = tc::pack8_bf16(hn);
          *reinterpret_cast<uint4*>(arow + (size_t)((u0 + c0) / 8) * PSTRIDE_A) = p0;
          float hi[8];
          tc::unpack8_bf16(p0, hi);
#pragma unroll
          for (int k = 0; k < 8; ++k) hi[k] = hn[k] - hi[k];
          *reinterpret_cast<uint4*>(lrow + (size_t)((u0 + c0) / 8) * PSTRIDE_A) = tc::pack8_bf16(hi);
          if (live && td >= 0) *reinterpret_cast<uint4*>(orow + c0) = p0;
        }
        tc::tc_fence_before();
        tc::fence_async_smem();
        if (t + 1 < p.Ts) mbar_arrive(h_ready);
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_free(tmem, 256);
}

}  // namespace gu

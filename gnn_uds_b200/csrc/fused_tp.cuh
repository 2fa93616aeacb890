// Fused temporal stack (sm_100a): the P causal dilated Conv1D layers of one temporal net (emulator.py:154-163, applied
// 243-257 / 298-310) in ONE kernel.  The layer-by-layer schedule (one tc_rows_kernel launch per layer) writes every
// intermediate (B, T, Nn, H) tensor to HBM and reads every input row k times through L2; here only the stack's input is
// read and only its output is written.
//
// A tile = G whole sequences (b, n0 .. n0+G-1), G = floor(128 / Ts), rows interleaved as r = t*G + g, so that a time
// shift of s steps is a shift of s*G rows for EVERY sequence of the tile and the rows before t = 0 are one block:
//   layer 1   A operand = k TMA boxes {64 columns, G nodes, Ts steps} of the (B, Ts, Nn, C) input taken at t0 = -shift (rows
//             before the sequence start fall outside the tensor and arrive as zeros = the causal padding), SWIZZLE_128B
//             K-major stages in a ring;  D[128 x H] in TMEM
//   layer l>1 A operand = the previous layer's output, drained to fp16 SWIZZLE_NONE K-major panels in shared memory with
//             `pad` zero rows in front: tap shift s is the same panel buffer with the start address moved back by s*G*16
//             bytes.  Taps whose shift reaches Ts see only padding and are skipped.
//   last      TMEM -> bias, act -> fp16 -> HBM rows (b, t, n) (the block-1 net keeps the last Td steps, emulator.py:249)
// Weights of all layers stay resident in shared memory (K-major images, as tc_rows_kernel uses them).
//
// Warps 0..7 drain (TMEM lanes 32*(w%4).., 32 of the H=64 columns each), warp 8 issues the MMAs (warp-uniform code, one
// elected lane), warp 9 is the TMA producer.  With P <= 3 the pipe sees L2[i], L1[i+1], L3[i] per tile and the drains run
// L2d[i], L1d[i+1], L3d[i]: the layer-1 MMAs of the next tile cover the drains of this one.
#pragma once
#include <cuda.h>

#include "fused_conv.cuh"
#include "kernels.cuh"
#include "tc_gemm.cuh"
#include "umma.cuh"

namespace gu {

constexpr int TP_DRAIN_WARPS = 8;
constexpr int TP_THREADS = (TP_DRAIN_WARPS + 2) * 32;   // 320
constexpr int TP_MAXP = 6;       // Conv1D layers + the optional Dense tail

struct TpConvP {
  CUtensorMap tm_in;          // {C, Nn, Ts, B}, box {64, G, Ts, 1}, SWIZZLE_128B
  const uint8_t* Wimg[TP_MAXP]; uint32_t w_bytes[TP_MAXP], w_off[TP_MAXP];   // K-major images of W^T, panel stride H*16
  const float* bias[TP_MAXP];
  bf16* out;                  // (B, Td, Nn, Nlast)
  // P layers in all.  The first `Pc` are the Conv1D stack (k taps, dilation 2^l, H outputs, activation `act`); an optional
  // last layer (P == Pc + 1) is the Dense that follows the block-2 temporal net (emulator.py:313 dense_resx / :317):
  // one tap, Nlast outputs, activation `act_last`.
  int P, Pc, k, cin, H, Nlast, act_last, G, Ts, Td, Nn, act, tpb, interleave, NS;
  unsigned n_tiles;
  uint32_t pad_rows, pstrideP, off_stage, off_pan[TP_MAXP], off_bias, off_bar;
};

__global__ void __launch_bounds__(TP_THREADS, 1) tp_conv_kernel(const __grid_constant__ TpConvP p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.off_bar);
  const int NS = p.NS, P = p.P;
  uint64_t* full = bars;                       // [NS] TMA complete_tx
  uint64_t* empty = bars + NS;                 // [NS] the stage's MMAs completed (tcgen05.commit)
  uint64_t* l_done = bars + 2 * NS;            // [P]  layer l MMAs of a tile completed
  uint64_t* pan_ready = l_done + TP_MAXP;      // [P]  layer l output drained to its panels (l < P-1) / to HBM (l = P-1)
  uint64_t* wbar = pan_ready + TP_MAXP;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wbar + 1);
  float* bias_s = reinterpret_cast<float*>(smem + p.off_bias);     // [P][128]

  if (warp == 0) tc::tmem_alloc(tmem_slot, 512);
  if (tid == 32) {
    for (int i = 0; i < NS; ++i) { tc::mbar_init(&full[i], 1); tc::mbar_init(&empty[i], 1); }
    for (int l = 0; l < P; ++l) { tc::mbar_init(&l_done[l], 1); tc::mbar_init(&pan_ready[l], TP_DRAIN_WARPS * 32); }
    tc::mbar_init(wbar, 1);
    tc::mbar_fence_init();
  }
  for (int i = tid; i < P * 128; i += TP_THREADS) { const int l = i >> 7, c = i & 127, nl = l < p.Pc ? p.H : p.Nlast; bias_s[i] = (c < nl && p.bias[l]) ? p.bias[l][c] : 0.f; }
  for (int b = 0; b < P - 1; ++b) {            // panels: the pad rows in front must be zero, the rest finite
    if (b >= 2 && !p.interleave) break;
    uint4* z = reinterpret_cast<uint4*>(smem + p.off_pan[b]);
    const int n16 = (int)((p.H / 8) * p.pstrideP / 16);
    for (int i = tid; i < n16; i += TP_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  tc::pdl_sync();
  const uint32_t tmem = *tmem_slot;
  const int n_it = p.n_tiles > blockIdx.x ? (int)((p.n_tiles - 1 - blockIdx.x) / gridDim.x) + 1 : 0;
  const int nchunk = p.cin >> 6, nst = p.k * nchunk;        // layer-1 stages per tile
  const uint32_t stage_bytes = 16384u;

  if (warp == TP_DRAIN_WARPS + 1) {
    // ================================ TMA producer ================================
    if (lane == 0 && n_it > 0) {
      tc::tma_prefetch_desc(&p.tm_in);
      int stage = 0; uint32_t eph = 1;          // parity to wait for on empty[]: passes at once on the first lap
      for (int i = 0; i < n_it; ++i) {
        const unsigned tile = blockIdx.x + (unsigned)i * gridDim.x;
        const int b = (int)(tile / (unsigned)p.tpb), n0 = (int)(tile % (unsigned)p.tpb) * p.G;
        for (int j = 0; j < p.k; ++j) {
          const int shift = p.k - 1 - j;        // layer 1: dilation 1
          for (int c = 0; c < nchunk; ++c) {
            tc::mbar_wait(&empty[stage], eph);
            tc::mbar_expect_tx(&full[stage], (uint32_t)(p.G * p.Ts) * 128u);
            tc::tma_load_4d(smem + p.off_stage + (size_t)stage * stage_bytes, &p.tm_in, 64 * c, n0, -shift, b, &full[stage]);
            if (++stage == NS) { stage = 0; eph ^= 1u; }
          }
        }
      }
    }
  } else if (warp == TP_DRAIN_WARPS) {
    // ================================ MMA issuer (warp-uniform, one elected lane) ================================
    if (n_it > 0) {
      if (tc::elect_one()) {
        uint32_t tot = 0;
        for (int l = 0; l < P; ++l) tot += p.w_bytes[l];
        tc::mbar_expect_tx(wbar, tot);
        for (int l = 0; l < P; ++l)
          for (uint32_t o = 0; o < p.w_bytes[l]; o += 16384) tc::bulk_g2s(smem + p.w_off[l] + o, p.Wimg[l] + o, min(p.w_bytes[l] - o, 16384u), wbar);
      }
      __syncwarp();
      tc::mbar_wait(wbar, 0);
      const int H = p.H, G = p.G, Ts = p.Ts, kk = p.k;
      const uint32_t idesc = tc::idesc_bf16(128, H, 0, 0);
      const uint32_t pstrideW = (uint32_t)H * 16u;
      const uint64_t w_step = (uint64_t)((2u * pstrideW) >> 4), p_step = (uint64_t)((2u * p.pstrideP) >> 4);
      const uint32_t st_addr = tc::smem_u32(smem + p.off_stage);
      int stage = 0; uint32_t fph = 0;
      auto issue_l1 = [&](int i) {
        // accumulators of layer 1: free once the layer-1 drain of the previous tile is done
        if (i >= 1) tc::mbar_wait(&pan_ready[0], (i - 1) & 1);
        tc::tc_fence_after();
        const uint64_t w0 = tc::smem_desc(tc::smem_u32(smem + p.w_off[0]), pstrideW, 128u);
        for (int s = 0; s < nst; ++s) {
          tc::mbar_wait(&full[stage], fph);
          tc::tc_fence_after();
          const uint64_t ad0 = tc::smem_desc_sw128(st_addr + (uint32_t)stage * stage_bytes), wd0 = w0 + (uint64_t)(4 * s) * w_step;
          if (tc::elect_one()) {
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) tc::umma(tmem, ad0 + (uint64_t)(2 * ks), wd0 + (uint64_t)ks * w_step, idesc, (s | ks) != 0);
            tc::umma_commit(&empty[stage]);
          }
          __syncwarp();
          if (++stage == NS) { stage = 0; fph ^= 1u; }
        }
        if (tc::elect_one()) tc::umma_commit(&l_done[0]);
        __syncwarp();
      };
      auto issue_l = [&](int l, int i) {        // layer l >= 1 (0-based) from the panels of layer l-1
        tc::mbar_wait(&pan_ready[l - 1], i & 1);                          // input panels written
        if (i >= 1) tc::mbar_wait(&pan_ready[l], (i - 1) & 1);            // this layer's accumulators drained (previous tile)
        tc::tc_fence_after();
        const uint32_t pan = tc::smem_u32(smem + p.off_pan[p.interleave ? l - 1 : (l - 1) & 1]);
        const bool tail = l >= p.Pc;            // the Dense tail: one tap, Nlast outputs
        const int nl = tail ? p.Nlast : H, kl = tail ? 1 : kk;
        const uint32_t pstrideWl = (uint32_t)nl * 16u;
        const uint64_t wl = tc::smem_desc(tc::smem_u32(smem + p.w_off[l]), pstrideWl, 128u), wl_step = (uint64_t)((2u * pstrideWl) >> 4);
        const uint32_t idesc_l = tc::idesc_bf16(128, nl, 0, 0);
        const int d = 1 << l;
        bool first = true;
        for (int j = 0; j < kl; ++j) {
          const int shift = (kl - 1 - j) * d;
          if (shift >= Ts) continue;            // the tap sees only the causal padding
          const uint64_t ad0 = tc::smem_desc(pan + (p.pad_rows - (uint32_t)(shift * G)) * 16u, p.pstrideP, 128u);
          const uint64_t wd0 = wl + (uint64_t)(j * (H >> 4)) * wl_step;
          if (tc::elect_one()) {
            for (int ks = 0; ks < (H >> 4); ++ks) tc::umma(tmem + (uint32_t)(l * H), ad0 + (uint64_t)ks * p_step, wd0 + (uint64_t)ks * wl_step, idesc_l, !(first && ks == 0));
          }
          __syncwarp();
          first = false;
        }
        if (tc::elect_one()) tc::umma_commit(&l_done[l]);
        __syncwarp();
      };
      issue_l1(0);
      for (int i = 0; i < n_it; ++i) {
        issue_l(1, i);
        if (p.interleave && i + 1 < n_it) issue_l1(i + 1);
        for (int l = 2; l < P; ++l) issue_l(l, i);
        if (!p.interleave && i + 1 < n_it) issue_l1(i + 1);
      }
    }
  } else {
    // ================================ drain warps ================================
    const int q = warp & 3, cs = warp >> 2;     // TMEM lanes 32q.., columns 32cs.. of the H = 64
    const int H = p.H, G = p.G, rows = p.G * p.Ts, r = q * 32 + lane;
    auto drain = [&](int l, int i) {
      tc::mbar_wait(&l_done[l], i & 1);
      tc::tc_fence_after();
      const bool tail = l >= p.Pc;
      const int nl = tail ? p.Nlast : H, act_l = tail ? p.act_last : p.act;
      const unsigned tile = blockIdx.x + (unsigned)i * gridDim.x;
      const int b = (int)(tile / (unsigned)p.tpb), n0 = (int)(tile % (unsigned)p.tpb) * G;
      const int t = r / G, g = r - t * G, n = n0 + g, t_out = t - (p.Ts - p.Td);
      // this warp's columns: 32 of the 64, or 64 of the 128 of the Dense tail, in passes of 32
      for (int c0 = cs * (nl >> 1); c0 < (cs + 1) * (nl >> 1); c0 += 32) {
        float v[32];
        tc::tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(l * H + c0), v);
        float4 b4[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) b4[j] = *reinterpret_cast<const float4*>(bias_s + l * 128 + c0 + 4 * j);
        tc::tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 8; ++j) { v[4 * j] += b4[j].x; v[4 * j + 1] += b4[j].y; v[4 * j + 2] += b4[j].z; v[4 * j + 3] += b4[j].w; }
        act_tile<32>(v, act_l);
        if (l < p.P - 1) {
          if (r < rows) {
            uint8_t* dst = smem + p.off_pan[p.interleave ? l : l & 1] + (size_t)(c0 >> 3) * p.pstrideP + (size_t)(p.pad_rows + r) * 16;
#pragma unroll
            for (int j = 0; j < 4; ++j) *reinterpret_cast<uint4*>(dst + (size_t)j * p.pstrideP) = tc::pack8_bf16(v + 8 * j);
          }
        } else if (r < rows && n < p.Nn && t_out >= 0) {
          bf16* orow = p.out + (((size_t)b * p.Td + t_out) * p.Nn + n) * nl + c0;
          st_global_v8(orow, tc::pack8_bf16(v), tc::pack8_bf16(v + 8));
          st_global_v8(orow + 16, tc::pack8_bf16(v + 16), tc::pack8_bf16(v + 24));
        }
      }
      if (l < p.P - 1) tc::fence_async_smem();  // panels -> tensor-core reads
      tc::tc_fence_before();
      mbar_arrive(&pan_ready[l]);
    };
    if (n_it > 0) drain(0, 0);
    for (int i = 0; i < n_it; ++i) {
      drain(1, i);
      if (p.interleave && i + 1 < n_it) drain(0, i + 1);
      for (int l = 2; l < P; ++l) drain(l, i);
      if (!p.interleave && i + 1 < n_it) drain(0, i + 1);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_free(tmem, 512);
}

}  // namespace gu

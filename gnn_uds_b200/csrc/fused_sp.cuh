// Fused spatial-layer kernel (sm_100a), one stream side per launch:
//
//   out[s] = act( Agg( [ self[s] | Mat @ other_dense[s] ] @ Wconv ) + bias )          (emulator.py:227-230)
//
// i.e. NodeEdge (K2), the concat, the GCN/GAT linear part and the CSR neighbour aggregation /
// GAT edge softmax (K4/K5) of one (b,t) sample without any intermediate leaving the SM.
//   self[s]        (R x Kself bf16)   node (or edge) activations of the sample, 1..2 column segments
//   other_dense[s] (Ro x F bf16)      Dense(E/2)(other stream) = x_e (or e_x), produced by tc_rows_kernel
//   Mat            (R x Ro)           folded NodeEdge matrix  w*|inci| + b, resident in smem (bf16 K-major image)
//   Wconv          (Kself+F x Nout)   conv kernel, resident in smem (bf16 K-major image of W^T)
//
// Persistent, one CTA per SM, one sample per pass, inputs double-buffered by TMA:
//   warps 0..3   "drain" group : TMEM -> smem.  NodeEdge result -> bf16 panels (A operand of the conv GEMM);
//                                conv result h -> bf16 row tile + GAT scores from the fp32 accumulators
//   warps 4..11  "graph" group : edge softmax + weighted neighbour sum over the h tile -> bias/act -> HBM
//   warp  12     MMA issuer    : tcgen05.mma  NE: D[128 x F] = Mat(K-major) x other_dense(MN-major)
//                                             conv: D[128 x Nout] = [self | NE](K-major) x Wconv(K-major)
//   warp  13     TMA producer  : cp.async.bulk.tensor loads of sample i+2 once the graph group released the buffer
// The h tile of a sample aliases that sample's own input buffer (dead once the conv MMAs committed), so
// the graph group works on sample i while MMA + drain already process sample i+1 in the other buffer.
#pragma once
#include <cuda.h>

#include "kernels.cuh"
#include "tc_gemm.cuh"
#include "umma.cuh"

namespace gu {

constexpr int SP_DRAIN_WARPS = 4;
constexpr int SP_GRAPH_WARPS = 8;
constexpr int SP_THREADS = (SP_DRAIN_WARPS + SP_GRAPH_WARPS + 2) * 32;   // 448

struct SpConvP {
  CUtensorMap tm_self[2];     // {8, S*R rows, w/8 panels}, box {8, Rbox, w/8}
  CUtensorMap tm_other;       // {8, S*Ro rows, F/8 panels}, box {8, Kp, F/8}
  int nseg, segw[2];
  const uint8_t* Wimg; uint32_t w_bytes;      // conv weight image, panel stride Nout*16
  const uint8_t* Mimg; uint32_t m_bytes;      // NodeEdge matrix image, panel stride pstrideM
  const float* bias; const float* a_self; const float* a_neigh;
  const int* row_ptr; const int* col_idx; const float* val; int nnz;
  bf16* out;                  // (S*R, Nout)
  int S, R, Ro, Rbox, Kp, MT, Kself, F, Nout, act, gat;
  uint32_t pstrideA, pstrideM, pstrideX;      // bytes: self/NE panels (Rbox*16), Mat panels, other_dense panels (Kp*16)
  uint32_t off_w, off_m, off_in, in_stride, xe_off, off_nx, off_sc, off_bar, hsb;
  uint32_t self_bytes, xe_bytes;
};

__device__ __forceinline__ void graph_bar_sync() { asm volatile("bar.sync 2, %0;" ::"n"(SP_GRAPH_WARPS * 32) : "memory"); }

__global__ void __launch_bounds__(SP_THREADS, 1) sp_conv_kernel(const __grid_constant__ SpConvP p) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.off_bar);
  uint64_t* in_full = bars;          // [2] TMA complete_tx
  uint64_t* in_empty = bars + 2;     // [2] graph group released the buffer (h tile consumed)
  uint64_t* h_full = bars + 4;       // [2] drain group wrote the h tile
  uint64_t* ne_done = bars + 6;      // NodeEdge MMAs committed
  uint64_t* nx_ready = bars + 7;     // NE panels written (drain group)
  uint64_t* conv_done = bars + 8;    // conv MMAs committed
  uint64_t* wbar = bars + 9;         // resident weights landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);
  // scratch: bias[Nout] asf[Nout] anf[Nout] | ss[2][RP] sn[2][RP] | inv[RP] | alpha[nnz] | rp[R+1] ci[nnz]
  const int RP = 128 * p.MT;
  float* bias_s = reinterpret_cast<float*>(smem + p.off_sc);
  float* asf = bias_s + p.Nout;
  float* anf = asf + p.Nout;
  float* ss = anf + p.Nout;
  float* sn = ss + 2 * RP;
  float* inv_s = sn + 2 * RP;
  float* alpha = inv_s + RP;
  int* rp_s = reinterpret_cast<int*>(alpha + p.nnz);
  int* ci_s = rp_s + p.R + 1;

  if (warp == 0) tc::tmem_alloc(tmem_slot, 512);
  if (tid == 32) {
    for (int i = 0; i < 2; ++i) {
      tc::mbar_init(&in_full[i], 1);
      tc::mbar_init(&in_empty[i], 1);
      tc::mbar_init(&h_full[i], SP_DRAIN_WARPS * 32);
    }
    tc::mbar_init(ne_done, 1);
    tc::mbar_init(nx_ready, SP_DRAIN_WARPS * 32);
    tc::mbar_init(conv_done, 1);
    tc::mbar_init(wbar, 1);
    tc::mbar_fence_init();
  }
  // constants into scratch (all threads), zero the NE panels once (rows >= R stay zero/finite)
  for (int i = tid; i < p.Nout; i += SP_THREADS) {
    bias_s[i] = p.bias ? p.bias[i] : 0.f;
    asf[i] = p.gat ? p.a_self[i] : 0.f;
    anf[i] = p.gat ? p.a_neigh[i] : 0.f;
  }
  for (int i = tid; i <= p.R; i += SP_THREADS) rp_s[i] = p.row_ptr[i];
  for (int i = tid; i < p.nnz; i += SP_THREADS) { ci_s[i] = p.col_idx[i]; if (!p.gat) alpha[i] = p.val[i]; }
  {
    uint4* z = reinterpret_cast<uint4*>(smem + p.off_nx);
    const int n16 = (int)((p.F / 8) * p.pstrideA / 16);
    for (int i = tid; i < n16; i += SP_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t acc_conv = tmem;              // MT * Nout columns
  const uint32_t acc_ne = tmem + 256u;         // MT * F columns
  const int n_it = p.S > (int)blockIdx.x ? (p.S - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

  if (warp == SP_DRAIN_WARPS + SP_GRAPH_WARPS + 1) {
    // ================================ TMA producer ================================
    if (lane == 0) {
      for (int s = 0; s < p.nseg; ++s) tc::tma_prefetch_desc(&p.tm_self[s]);
      tc::tma_prefetch_desc(&p.tm_other);
      for (int i = 0; i < n_it; ++i) {
        const int buf = i & 1;
        const int smp = (int)blockIdx.x + i * (int)gridDim.x;
        if (i >= 2) tc::mbar_wait(&in_empty[buf], ((i >> 1) & 1) ^ 1);
        uint8_t* dst = smem + p.off_in + (size_t)buf * p.in_stride;
        tc::mbar_expect_tx(&in_full[buf], p.self_bytes + p.xe_bytes);
        uint32_t o = 0;
        for (int s = 0; s < p.nseg; ++s) {
          tc::tma_load_3d(dst + o, &p.tm_self[s], 0, smp * p.R, 0, &in_full[buf]);
          o += (uint32_t)(p.segw[s] >> 3) * p.pstrideA;
        }
        tc::tma_load_3d(dst + p.xe_off, &p.tm_other, 0, smp * p.Ro, 0, &in_full[buf]);
      }
    }
  } else if (warp == SP_DRAIN_WARPS + SP_GRAPH_WARPS) {
    // ================================ MMA issuer ================================
    if (lane == 0) {
      tc::mbar_expect_tx(wbar, p.w_bytes + p.m_bytes);
      for (uint32_t o = 0; o < p.w_bytes; o += 16384) tc::bulk_g2s(smem + p.off_w + o, p.Wimg + o, min(p.w_bytes - o, 16384u), wbar);
      for (uint32_t o = 0; o < p.m_bytes; o += 16384) tc::bulk_g2s(smem + p.off_m + o, p.Mimg + o, min(p.m_bytes - o, 16384u), wbar);
      tc::mbar_wait(wbar, 0);
      const uint32_t idesc_ne = tc::idesc_bf16(128, p.F, 0, 1);        // B = other_dense, MN-major
      const uint32_t idesc_cv = tc::idesc_bf16(128, p.Nout, 0, 0);
      const uint32_t w_addr = tc::smem_u32(smem + p.off_w), m_addr = tc::smem_u32(smem + p.off_m);
      const uint32_t nx_addr = tc::smem_u32(smem + p.off_nx);
      const uint32_t pstrideW = (uint32_t)p.Nout * 16u;
      const int ks_self = p.Kself >> 4, ks_ne = p.F >> 4;
      for (int i = 0; i < n_it; ++i) {
        const int buf = i & 1;
        const uint32_t in_addr = tc::smem_u32(smem + p.off_in + (size_t)buf * p.in_stride);
        tc::mbar_wait(&in_full[buf], (i >> 1) & 1);
        tc::tc_fence_after();
        // NodeEdge: D[row, f] = sum_k Mat[row, k] * other_dense[k, f]
        for (int ks = 0; ks < (p.Kp >> 4); ++ks) {
          const uint64_t bd = tc::smem_desc(in_addr + p.xe_off + (uint32_t)ks * 256u, 128u, p.pstrideX);
          for (int mt = 0; mt < p.MT; ++mt) {
            const uint64_t ad = tc::smem_desc(m_addr + (uint32_t)mt * 2048u + 2u * ks * p.pstrideM, p.pstrideM, 128u);
            tc::umma(acc_ne + (uint32_t)(mt * p.F), ad, bd, idesc_ne, ks != 0);
          }
        }
        tc::umma_commit(ne_done);
        // conv over the self columns first (they are ready), then over the NodeEdge columns.  The conv
        // accumulators of the previous sample must have been drained (h_full of sample i-1).
        if (i >= 1) { tc::mbar_wait(&h_full[(i - 1) & 1], ((i - 1) >> 1) & 1); tc::tc_fence_after(); }
        for (int ks = 0; ks < ks_self; ++ks) {
          const uint64_t bd = tc::smem_desc(w_addr + 2u * ks * pstrideW, pstrideW, 128u);
          for (int mt = 0; mt < p.MT; ++mt) {
            const uint64_t ad = tc::smem_desc(in_addr + (uint32_t)mt * 2048u + 2u * ks * p.pstrideA, p.pstrideA, 128u);
            tc::umma(acc_conv + (uint32_t)(mt * p.Nout), ad, bd, idesc_cv, ks != 0);
          }
        }
        tc::mbar_wait(nx_ready, i & 1);
        tc::tc_fence_after();
        for (int ks = 0; ks < ks_ne; ++ks) {
          const uint64_t bd = tc::smem_desc(w_addr + 2u * (ks_self + ks) * pstrideW, pstrideW, 128u);
          for (int mt = 0; mt < p.MT; ++mt) {
            const uint64_t ad = tc::smem_desc(nx_addr + (uint32_t)mt * 2048u + 2u * ks * p.pstrideA, p.pstrideA, 128u);
            tc::umma(acc_conv + (uint32_t)(mt * p.Nout), ad, bd, idesc_cv, 1u);
          }
        }
        tc::umma_commit(conv_done);
      }
    }
  } else if (warp < SP_DRAIN_WARPS) {
    // ================================ drain group ================================
    const int q = warp;                               // TMEM lanes 32q .. 32q+31
    for (int i = 0; i < n_it; ++i) {
      const int buf = i & 1;
      // ---- NodeEdge accumulators -> bf16 panels (A operand of the conv GEMM) ----
      tc::mbar_wait(ne_done, i & 1);
      tc::tc_fence_after();
      for (int mt = 0; mt < p.MT; ++mt) {
        if (mt * 128 + q * 32 >= p.R) continue;       // warp-uniform
        const int row = mt * 128 + q * 32 + lane;
        const uint32_t ta = acc_ne + ((uint32_t)(q * 32) << 16) + (uint32_t)(mt * p.F);
        uint8_t* dst = smem + p.off_nx + (size_t)row * 16;
        for (int c0 = 0; c0 < p.F; c0 += 32) {
          float v[32];
          tc::tmem_ld32(ta + (uint32_t)c0, v);
          tc::tmem_ld_wait();
          if (row < p.R) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              *reinterpret_cast<uint4*>(dst + (size_t)((c0 >> 3) + j) * p.pstrideA) = tc::pack8_bf16(v + 8 * j);
          }
        }
      }
      tc::fence_async_smem();                         // generic-proxy panel writes -> tensor-core reads
      tc::tc_fence_before();
      mbar_arrive(nx_ready);
      // ---- conv accumulators -> h tile (bf16 rows) + GAT scores from the fp32 values ----
      tc::mbar_wait(conv_done, i & 1);
      tc::tc_fence_after();
      uint8_t* hbase = smem + p.off_in + (size_t)buf * p.in_stride;
      for (int mt = 0; mt < p.MT; ++mt) {
        if (mt * 128 + q * 32 >= p.R) continue;
        const int row = mt * 128 + q * 32 + lane;
        const uint32_t ta = acc_conv + ((uint32_t)(q * 32) << 16) + (uint32_t)(mt * p.Nout);
        uint8_t* hrow = hbase + (size_t)row * p.hsb;
        float ps = 0.f, pn = 0.f;
        for (int c0 = 0; c0 < p.Nout; c0 += 32) {
          float v[32];
          tc::tmem_ld32(ta + (uint32_t)c0, v);
          tc::tmem_ld_wait();
          if (row < p.R) {
#pragma unroll
            for (int j = 0; j < 4; ++j) *reinterpret_cast<uint4*>(hrow + c0 * 2 + j * 16) = tc::pack8_bf16(v + 8 * j);
            if (p.gat) {
#pragma unroll
              for (int j = 0; j < 32; ++j) { ps = fmaf(v[j], asf[c0 + j], ps); pn = fmaf(v[j], anf[c0 + j], pn); }
            }
          }
        }
        if (p.gat && row < p.R) { ss[buf * RP + row] = ps; sn[buf * RP + row] = pn; }
      }
      tc::tc_fence_before();
      mbar_arrive(&h_full[buf]);
    }
  } else {
    // ================================ graph group ================================
    const int gt = tid - SP_DRAIN_WARPS * 32;         // 0..255
    const int gw = gt >> 5;
    const int lg = lane >> 3, gl = lane & 7;
    const int R = p.R;
    for (int i = 0; i < n_it; ++i) {
      const int buf = i & 1;
      const int smp = (int)blockIdx.x + i * (int)gridDim.x;
      tc::mbar_wait(&h_full[buf], (i >> 1) & 1);
      const uint8_t* hbase = smem + p.off_in + (size_t)buf * p.in_stride;
      if (p.gat) {
        // edge softmax over the CSR pattern: one thread per row
        const float* ssb = ss + buf * RP;
        const float* snb = sn + buf * RP;
        for (int r = gt; r < R; r += SP_GRAPH_WARPS * 32) {
          const int beg = rp_s[r], end = rp_s[r + 1];
          const float si = ssb[r];
          float mx = -INFINITY;
          for (int qq = beg; qq < end; ++qq) {
            float sc = si + snb[ci_s[qq]];
            sc = sc > 0.f ? sc : 0.2f * sc;
            mx = fmaxf(mx, sc);
          }
          float den = 0.f;
          for (int qq = beg; qq < end; ++qq) {
            float sc = si + snb[ci_s[qq]];
            sc = sc > 0.f ? sc : 0.2f * sc;
            const float w = __expf(sc - mx);
            alpha[qq] = w;
            den += w;
          }
          inv_s[r] = 1.f / den;
        }
        graph_bar_sync();
      }
      // weighted neighbour sum: 8 lanes per output row (8 features each per 64-feature block), 4 rows per warp pass
      for (int r = gw * 4 + lg; r < R; r += SP_GRAPH_WARPS * 4) {
        const int beg = rp_s[r], end = rp_s[r + 1];
        const float scale = p.gat ? inv_s[r] : 1.f;
        bf16* orow = p.out + ((size_t)smp * R + r) * p.Nout + gl * 8;
        for (int e0 = 0; e0 < p.Nout; e0 += 64) {
          float acc[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] = 0.f;
          const uint8_t* hb = hbase + (e0 + gl * 8) * 2;
          for (int qq = beg; qq < end; ++qq) {
            const float w = alpha[qq];
            const uint4 u = *reinterpret_cast<const uint4*>(hb + (size_t)ci_s[qq] * p.hsb);
            float hv[8];
            tc::unpack8_bf16(u, hv);
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] = fmaf(w, hv[k], acc[k]);
          }
          const float* bp = bias_s + e0 + gl * 8;
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] = fmaf(acc[k], scale, bp[k]);
          act_tile<8>(acc, p.act);
          *reinterpret_cast<uint4*>(orow + e0) = tc::pack8_bf16(acc);
        }
      }
      graph_bar_sync();                               // everyone done with h / alpha of this sample
      if (gt == 0) mbar_arrive(&in_empty[buf]);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_free(tmem, 512);
}

}  // namespace gu

// Fused graph-convolution kernel (sm_100a): spektral GCNConv / GATConv (K4/K5) of one (b,t) sample per pass,
// every contraction on tcgen05, nothing but the input and the output touching HBM:
//
//   h      = [seg0 | seg1 | seg2][s] @ W                       (R x Ktot) x (Ktot x Nout)   conv MMA
//   alpha  = softmax_j( leaky_relu(h_i.a_self + h_j.a_neigh) ) over the CSR pattern (GAT) | filter values (GCN)
//   out[s] = act( alpha @ h + bias )                           (R x R) x (R x Nout)          aggregation MMA
//
// The attention matrix of a 30..256-row drainage graph is tiny, so it is materialised per sample as a dense
// bf16 K-major A operand in shared memory (zeros off the pattern) and the neighbour sum runs on the tensor
// cores with h as the MN-major B operand: a 128x128x128 MMA takes 512 cycles where the SIMT gather over the
// same CSR took ~6000.
//
// Persistent, one CTA per SM, one sample per pass, inputs double-buffered by TMA:
//   warps 0..3   conv drain  : conv accumulators -> bf16 h panels + GAT scores (fp32)
//   warps 4..7   out drain   : aggregation accumulators -> bias/act -> bf16 -> HBM (overlaps the next conv drain)
//   warps 8..11  alpha group : zero + scatter the attention matrix (softmax over each row's neighbours)
//   warp  12     MMA issuer  : conv MMAs of sample i+1 are issued before the aggregation MMAs of sample i
//   warp  13     TMA producer
// alpha of a sample is built in that sample's own input buffer (dead once its conv MMAs committed).
#pragma once
#include <cuda.h>

#include "kernels.cuh"
#include "tc_gemm.cuh"
#include "umma.cuh"

namespace gu {

constexpr int CA_DRAIN_WARPS = 4;                                          // per drain group (two groups)
constexpr int CA_ALPHA_WARPS = 4;
constexpr int CA_THREADS = (2 * CA_DRAIN_WARPS + CA_ALPHA_WARPS + 2) * 32;   // 448

struct ConvAggP {
  CUtensorMap tm[3];          // per segment {8, S*R rows, w/8 panels}, box {8, Rbox, w/8}
  int nseg, segw[3];
  const uint8_t* Wimg; uint32_t w_bytes;      // K-major image of W^T, panel stride Nout*16
  const float* bias; const float* a_self; const float* a_neigh;
  const int* row_ptr; const int* col_idx; const float* val; int nnz;
  bf16* out;                  // (S*R, Nout)
  int S, R, Rbox, KpJ, MT, Ktot, Nout, act, gat;
  uint32_t pstrideA;          // input / alpha panels: Rbox*16 bytes
  uint32_t pstrideH;          // h panels: KpJ*16 bytes
  uint32_t off_w, off_in, in_stride, off_h, off_sc, off_bar, in_bytes, alpha_bytes;
  int trace_mode;             // 0: phase stamps, 1: stamps inside the conv drain
  long long* trace;           // debug: clock64 stamps of CTA 0, 8 per sample (GNNUDS_TRACE_FILE), else null
};

#define CA_STAMP(i, k) do { if (p.trace && p.trace_mode == 0 && blockIdx.x == 0 && (i) < 64) p.trace[(i) * 8 + (k)] = clock64(); } while (0)
#define CA_STAMP1(i, k) do { if (p.trace && p.trace_mode == 1 && blockIdx.x == 0 && tid == 0 && (i) < 64) p.trace[(i) * 8 + (k)] = clock64(); } while (0)

__device__ __forceinline__ void st_global_v8(void* p, const uint4& a, const uint4& b) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b.x),
               "r"(b.y), "r"(b.z), "r"(b.w)
               : "memory");
}
__device__ __forceinline__ void alpha_bar_sync() { asm volatile("bar.sync 3, %0;" ::"n"(CA_ALPHA_WARPS * 32) : "memory"); }

__global__ void __launch_bounds__(CA_THREADS, 1) conv_agg_kernel(const __grid_constant__ ConvAggP p) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.off_bar);
  uint64_t* in_full = bars;           // [2] TMA complete_tx
  uint64_t* in_empty = bars + 2;      // [2] aggregation MMAs that read alpha from this buffer committed
  uint64_t* conv_done = bars + 4;     // conv MMAs committed
  uint64_t* h_ready = bars + 5;       // drain group wrote h panels + scores (and drained the conv accumulators)
  uint64_t* alpha_ready = bars + 6;   // alpha group wrote the attention operand
  uint64_t* agg_done = bars + 7;      // aggregation MMAs committed
  uint64_t* out_done = bars + 8;      // drain group drained the aggregation accumulators
  uint64_t* wbar = bars + 9;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);
  // scratch: bias[Nout] asf[Nout] anf[Nout] | ss[RP] sn[RP] | rp[R+1] ci[nnz] | val[nnz] (GCN)
  const int RP = 128 * p.MT;
  float* bias_s = reinterpret_cast<float*>(smem + p.off_sc);
  float* asf = bias_s + p.Nout;
  float* anf = asf + p.Nout;
  float* ss = anf + p.Nout;
  float* sn = ss + RP;
  int* rp_s = reinterpret_cast<int*>(sn + RP);
  int* ci_s = rp_s + p.R + 1;
  float* val_s = reinterpret_cast<float*>(ci_s + p.nnz);

  if (warp == 0) tc::tmem_alloc(tmem_slot, 512);
  if (tid == 32) {
    for (int i = 0; i < 2; ++i) { tc::mbar_init(&in_full[i], 1); tc::mbar_init(&in_empty[i], 1); }
    tc::mbar_init(conv_done, 1);
    tc::mbar_init(h_ready, CA_DRAIN_WARPS * 32);
    tc::mbar_init(alpha_ready, CA_ALPHA_WARPS * 32);
    tc::mbar_init(agg_done, 1);
    tc::mbar_init(out_done, CA_DRAIN_WARPS * 32);
    tc::mbar_init(wbar, 1);
    tc::mbar_fence_init();
  }
  for (int i = tid; i < p.Nout; i += CA_THREADS) {
    bias_s[i] = p.bias ? p.bias[i] : 0.f;
    asf[i] = p.gat ? p.a_self[i] : 0.f;
    anf[i] = p.gat ? p.a_neigh[i] : 0.f;
  }
  for (int i = tid; i <= p.R; i += CA_THREADS) rp_s[i] = p.row_ptr[i];
  for (int i = tid; i < p.nnz; i += CA_THREADS) { ci_s[i] = p.col_idx[i]; if (!p.gat) val_s[i] = p.val[i]; }
  {  // h panels: rows >= R are read by the aggregation MMA against zero alpha columns, keep them finite
    uint4* z = reinterpret_cast<uint4*>(smem + p.off_h);
    const int n16 = (int)((p.Nout / 8) * p.pstrideH / 16);
    for (int i = tid; i < n16; i += CA_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t acc_conv = tmem;              // MT * Nout columns (<= 256)
  const uint32_t acc_agg = tmem + 256u;        // MT * Nout columns
  const int n_it = p.S > (int)blockIdx.x ? (p.S - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  // kernel parameters used inside the role loops, copied to registers once (the mbarrier / tcgen05 asm
  // statements are memory clobbers, so reads of `p` inside the loops would be re-issued as constant loads)
  const int kMT = p.MT, kNout = p.Nout, kR = p.R, kKtot = p.Ktot, kKpJ = p.KpJ, kgat = p.gat, kact = p.act, knseg = p.nseg;
  const uint32_t kpstrideA = p.pstrideA, kpstrideH = p.pstrideH, kin_stride = p.in_stride, koff_in = p.off_in, koff_h = p.off_h, kalpha_bytes = p.alpha_bytes, kin_bytes = p.in_bytes, koff_w = p.off_w, kw_bytes = p.w_bytes;

  if (warp == 2 * CA_DRAIN_WARPS + CA_ALPHA_WARPS + 1) {
    // ================================ TMA producer ================================
    if (lane == 0) {
      for (int s = 0; s < knseg; ++s) tc::tma_prefetch_desc(&p.tm[s]);
      for (int i = 0; i < n_it; ++i) {
        const int buf = i & 1;
        const int smp = (int)blockIdx.x + i * (int)gridDim.x;
        if (i >= 2) tc::mbar_wait(&in_empty[buf], ((i >> 1) & 1) ^ 1);
        uint8_t* dst = smem + koff_in + (size_t)buf * kin_stride;
        tc::mbar_expect_tx(&in_full[buf], kin_bytes);
        uint32_t o = 0;
        for (int s = 0; s < knseg; ++s) {
          tc::tma_load_3d(dst + o, &p.tm[s], 0, smp * kR, 0, &in_full[buf]);
          o += (uint32_t)(p.segw[s] >> 3) * kpstrideA;
        }
      }
    }
  } else if (warp == 2 * CA_DRAIN_WARPS + CA_ALPHA_WARPS) {
    // ================================ MMA issuer ================================
    if (lane == 0) {
      tc::mbar_expect_tx(wbar, kw_bytes);
      for (uint32_t o = 0; o < kw_bytes; o += 16384) tc::bulk_g2s(smem + koff_w + o, p.Wimg + o, min(kw_bytes - o, 16384u), wbar);
      tc::mbar_wait(wbar, 0);
      const uint32_t idesc_cv = tc::idesc_bf16(128, kNout, 0, 0);
      const uint32_t idesc_ag = tc::idesc_bf16(128, kNout, 0, 1);       // B = h panels, MN-major
      const uint32_t w_addr = tc::smem_u32(smem + koff_w), h_addr = tc::smem_u32(smem + koff_h);
      const uint32_t pstrideW = (uint32_t)kNout * 16u;
      const uint64_t w_desc0 = tc::smem_desc(w_addr, pstrideW, 128u), h_desc0 = tc::smem_desc(h_addr, 128u, kpstrideH);
      const uint64_t a_step = (uint64_t)((2u * kpstrideA) >> 4), w_step = (uint64_t)((2u * pstrideW) >> 4);   // one K=16 step
      auto issue_conv = [&](int i) {
        const int buf = i & 1;
        const uint32_t in_addr = tc::smem_u32(smem + koff_in + (size_t)buf * kin_stride);
        tc::mbar_wait(&in_full[buf], (i >> 1) & 1);
        if (i >= 1) tc::mbar_wait(h_ready, (i - 1) & 1);                 // conv accumulators of sample i-1 drained
        tc::tc_fence_after();
        CA_STAMP(i, 0);
        // descriptors advance linearly with the start address: one 64-bit add per MMA in the issue loop
        uint64_t ad = tc::smem_desc(in_addr, kpstrideA, 128u), bd = w_desc0;
        const int nks = kKtot >> 4;
        if (kMT == 1) {
#pragma unroll 4
          for (int ks = 0; ks < nks; ++ks, ad += a_step, bd += w_step) tc::umma(acc_conv, ad, bd, idesc_cv, ks != 0);
        } else {
#pragma unroll 2
          for (int ks = 0; ks < nks; ++ks, ad += a_step, bd += w_step) {
            tc::umma(acc_conv, ad, bd, idesc_cv, ks != 0);
            tc::umma(acc_conv + (uint32_t)kNout, ad + 128u, bd, idesc_cv, ks != 0);
          }
        }
        tc::umma_commit(conv_done);
        CA_STAMP(i, 1);
      };
      auto issue_agg = [&](int i) {
        const int buf = i & 1;
        const uint32_t al_addr = tc::smem_u32(smem + koff_in + (size_t)buf * kin_stride);
        tc::mbar_wait(alpha_ready, i & 1);
        if (i >= 1) tc::mbar_wait(out_done, (i - 1) & 1);                // aggregation accumulators of sample i-1 drained
        tc::tc_fence_after();
        uint64_t ad = tc::smem_desc(al_addr, kpstrideA, 128u), bd = h_desc0;
        const int nks = kKpJ >> 4;
        if (kMT == 1) {
#pragma unroll 4
          for (int ks = 0; ks < nks; ++ks, ad += a_step, bd += 16u) tc::umma(acc_agg, ad, bd, idesc_ag, ks != 0);
        } else {
#pragma unroll 2
          for (int ks = 0; ks < nks; ++ks, ad += a_step, bd += 16u) {
            tc::umma(acc_agg, ad, bd, idesc_ag, ks != 0);
            tc::umma(acc_agg + (uint32_t)kNout, ad + 128u, bd, idesc_ag, ks != 0);
          }
        }
        tc::umma_commit(agg_done);
        tc::umma_commit(&in_empty[buf]);                                 // alpha consumed: the buffer may be reloaded
        CA_STAMP(i, 5);
      };
      if (n_it > 0) issue_conv(0);
      for (int i = 0; i < n_it; ++i) {
        if (i + 1 < n_it) issue_conv(i + 1);
        issue_agg(i);
      }
    }
  } else if (warp < CA_DRAIN_WARPS) {
    // ================================ conv drain ================================
    const int q = warp;
    for (int i = 0; i < n_it; ++i) {
      tc::mbar_wait(conv_done, i & 1);
      if (i >= 1) tc::mbar_wait(agg_done, (i - 1) & 1);                  // h panels of sample i-1 consumed
      tc::tc_fence_after();
      if (tid == 0) CA_STAMP(i, 2);
      CA_STAMP1(i, 0);
      for (int mt = 0; mt < kMT; ++mt) {
        if (mt * 128 + q * 32 >= kR) continue;                           // warp-uniform
        const int row = mt * 128 + q * 32 + lane;
        const uint32_t ta = acc_conv + ((uint32_t)(q * 32) << 16) + (uint32_t)(mt * kNout);
        uint8_t* hrow = smem + koff_h + (size_t)row * 16;
        float ps[4] = {0.f, 0.f, 0.f, 0.f}, pn[4] = {0.f, 0.f, 0.f, 0.f};
        for (int c0 = 0; c0 < kNout; c0 += 32) {
          float v[32];
          tc::tmem_ld32(ta + (uint32_t)c0, v);
          // attention vectors of this column chunk: 16-byte broadcast loads issued while the TMEM load is in flight
          float4 as4[8], an4[8];
          if (kgat) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              as4[j] = *reinterpret_cast<const float4*>(asf + c0 + 4 * j);
              an4[j] = *reinterpret_cast<const float4*>(anf + c0 + 4 * j);
            }
          }
          tc::tmem_ld_wait();
          if (mt == 0 && c0 == 0) CA_STAMP1(i, 1);
          if (mt == 0 && c0 == 32) CA_STAMP1(i, 3);
          if (row < kR) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              *reinterpret_cast<uint4*>(hrow + (size_t)((c0 >> 3) + j) * kpstrideH) = tc::pack8_bf16(v + 8 * j);
            if (kgat) {
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                ps[0] = fmaf(v[4 * j], as4[j].x, ps[0]); ps[1] = fmaf(v[4 * j + 1], as4[j].y, ps[1]);
                ps[2] = fmaf(v[4 * j + 2], as4[j].z, ps[2]); ps[3] = fmaf(v[4 * j + 3], as4[j].w, ps[3]);
                pn[0] = fmaf(v[4 * j], an4[j].x, pn[0]); pn[1] = fmaf(v[4 * j + 1], an4[j].y, pn[1]);
                pn[2] = fmaf(v[4 * j + 2], an4[j].z, pn[2]); pn[3] = fmaf(v[4 * j + 3], an4[j].w, pn[3]);
              }
            }
          }
          if (mt == 0 && c0 == 0) CA_STAMP1(i, 2);
          if (mt == 0 && c0 == 32) CA_STAMP1(i, 4);
        }
        if (kgat && row < kR) { ss[row] = (ps[0] + ps[1]) + (ps[2] + ps[3]); sn[row] = (pn[0] + pn[1]) + (pn[2] + pn[3]); }
      }
      CA_STAMP1(i, 5);
      tc::fence_async_smem();                                            // h panels -> tensor-core reads
      CA_STAMP1(i, 6);
      tc::tc_fence_before();
      mbar_arrive(h_ready);
      CA_STAMP1(i, 7);
      if (tid == 0) CA_STAMP(i, 3);
    }
  } else if (warp < 2 * CA_DRAIN_WARPS) {
    // ================================ out drain ================================
    const int q = warp - CA_DRAIN_WARPS;
    for (int i = 0; i < n_it; ++i) {
      const int smp = (int)blockIdx.x + i * (int)gridDim.x;
      tc::mbar_wait(agg_done, i & 1);
      tc::tc_fence_after();
      if (tid == CA_DRAIN_WARPS * 32) CA_STAMP(i, 6);
      for (int mt = 0; mt < kMT; ++mt) {
        if (mt * 128 + q * 32 >= kR) continue;
        const int row = mt * 128 + q * 32 + lane;
        const uint32_t ta = acc_agg + ((uint32_t)(q * 32) << 16) + (uint32_t)(mt * kNout);
        bf16* orow = p.out + ((size_t)smp * kR + row) * kNout;
        for (int c0 = 0; c0 < kNout; c0 += 32) {
          float v[32];
          tc::tmem_ld32(ta + (uint32_t)c0, v);
          float4 b4[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) b4[j] = *reinterpret_cast<const float4*>(bias_s + c0 + 4 * j);
          tc::tmem_ld_wait();
          if (row < kR) {
#pragma unroll
            for (int j = 0; j < 8; ++j) { v[4 * j] += b4[j].x; v[4 * j + 1] += b4[j].y; v[4 * j + 2] += b4[j].z; v[4 * j + 3] += b4[j].w; }
            act_tile<32>(v, kact);
            // 32-byte stores: a lane's 64 contiguous bytes leave as two full sectors
            const uint4 u0 = tc::pack8_bf16(v), u1 = tc::pack8_bf16(v + 8), u2 = tc::pack8_bf16(v + 16), u3 = tc::pack8_bf16(v + 24);
            st_global_v8(orow + c0, u0, u1);
            st_global_v8(orow + c0 + 16, u2, u3);
          }
        }
      }
      tc::tc_fence_before();
      mbar_arrive(out_done);
      if (tid == CA_DRAIN_WARPS * 32) CA_STAMP(i, 7);
    }
  } else {
    // ================================ alpha group ================================
    const int gt = tid - 2 * CA_DRAIN_WARPS * 32;      // 0..127, one thread per row
    const int R = kR;
    for (int i = 0; i < n_it; ++i) {
      const int buf = i & 1;
      tc::mbar_wait(h_ready, i & 1);
      uint8_t* al = smem + koff_in + (size_t)buf * kin_stride;
      {
        uint4* z = reinterpret_cast<uint4*>(al);
        const int n16 = (int)(kalpha_bytes >> 4);
        for (int k = gt; k < n16; k += CA_ALPHA_WARPS * 32) z[k] = make_uint4(0u, 0u, 0u, 0u);
      }
      alpha_bar_sync();
      for (int r = gt; r < R; r += CA_ALPHA_WARPS * 32) {
        const int beg = rp_s[r], end = rp_s[r + 1];
        uint8_t* arow = al + (size_t)r * 16;
        if (kgat) {
          const float si = ss[r];
          if (end - beg <= 16) {
            // neighbours cached in registers: all shared-memory reads independent
            int js[16]; float sc[16];
            float mx = -INFINITY;
#pragma unroll
            for (int k = 0; k < 16; ++k) {
              const bool ok = beg + k < end;
              js[k] = ok ? ci_s[beg + k] : 0;
              float t = si + sn[js[k]];
              t = t > 0.f ? t : 0.2f * t;
              sc[k] = ok ? t : -INFINITY;
              mx = fmaxf(mx, sc[k]);
            }
            float den = 0.f;
#pragma unroll
            for (int k = 0; k < 16; ++k) { sc[k] = __expf(sc[k] - mx); den += sc[k]; }
            const float inv = 1.f / den;
#pragma unroll
            for (int k = 0; k < 16; ++k)
              if (beg + k < end)
                st1(reinterpret_cast<bf16*>(arow + (size_t)(js[k] >> 3) * kpstrideA + (js[k] & 7) * 2), sc[k] * inv);
          } else {
            float mx = -INFINITY;
            for (int qq = beg; qq < end; ++qq) {
              float t = si + sn[ci_s[qq]];
              t = t > 0.f ? t : 0.2f * t;
              mx = fmaxf(mx, t);
            }
            float den = 0.f;
            for (int qq = beg; qq < end; ++qq) {
              float t = si + sn[ci_s[qq]];
              t = t > 0.f ? t : 0.2f * t;
              den += __expf(t - mx);
            }
            const float inv = 1.f / den;
            for (int qq = beg; qq < end; ++qq) {
              const int j = ci_s[qq];
              float t = si + sn[j];
              t = t > 0.f ? t : 0.2f * t;
              st1(reinterpret_cast<bf16*>(arow + (size_t)(j >> 3) * kpstrideA + (j & 7) * 2), __expf(t - mx) * inv);
            }
          }
        } else {
          for (int qq = beg; qq < end; ++qq) {
            const int j = ci_s[qq];
            st1(reinterpret_cast<bf16*>(arow + (size_t)(j >> 3) * kpstrideA + (j & 7) * 2), val_s[qq]);
          }
        }
      }
      tc::fence_async_smem();                                            // alpha -> tensor-core reads
      mbar_arrive(alpha_ready);
      if (gt == 0) CA_STAMP(i, 4);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_free(tmem, 512);
}

}  // namespace gu

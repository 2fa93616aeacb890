// Fused spatial-layer kernel, second generation (sm_100a), one stream side per launch:
//
//   out[s] = act( alpha[s] @ ( [ self[s] | Mat @ other_dense[s] ] @ Wconv ) + bias )          (emulator.py:227-230)
//
// Same contract as sp_conv_kernel (fused_sp.cuh) - NodeEdge (K2), the concat, the GCN/GAT linear part and the graph
// aggregation (K4/K5) of one (b,t) sample with no intermediate leaving the SM - but EVERY contraction runs on tcgen05,
// the neighbour sum included: the attention (GAT) / filter (GCN) matrix of a 30..256-row drainage graph is tiny, so it
// is kept as a dense fp16 K-major A operand in shared memory (zeros off the CSR pattern, the pattern's entries
// rewritten per sample by the edge softmax) and `alpha @ h` is one more MMA series with the h tile as MN-major B
// operand.  The SIMT gather of sp_conv_kernel (550 instructions per warp and sample, 6.4 k of its 8.4 k cycle period)
// becomes <= 18 MMAs + one TMEM drain.
//
// GAT scores also come from the tensor core: the conv weight image carries 16 extra output columns,
// [hi(W a_self) lo(W a_self) hi(W a_neigh) lo(W a_neigh) 0...], so h.a_self and h.a_neigh arrive as accumulator
// columns Nout..Nout+3 (hi + lo = fp16 pair split of the fp32 product W a, ~22 mantissa bits) and the conv drain is a
// pure TMEM -> fp16 copy.
//
// Persistent, one CTA per SM, one sample per pass:
//   warps 0..15  workers: every TMEM drain and the softmax, in phases (a warp owns TMEM lanes 32*(w%4).. and the
//                column slice w/4):  A conv drain -> h tile (fp16, MN-major panels) + scores
//                                     B edge softmax -> alpha entries (GAT only)
//                                     C NodeEdge drain of the NEXT sample -> fp16 panels (A operand of its conv GEMM)
//                                     D aggregation drain -> 1/den, bias, act -> fp16 -> HBM
//   warp  16     MMA issuer (one thread): conv[i], NodeEdge[i+1], agg[i] in that order, so the tensor pipe works on
//                conv/NodeEdge while the workers run the softmax, and on agg while they drain NodeEdge
//   warp  17     TMA producer: other_dense tile (single buffer, free again once the NodeEdge MMAs read it) and the
//                self tiles (nb_self buffers; the h tile of a sample aliases its own self tiles)
// MT == 2 (R > 128, the 131-link side of shunqing): one self/h buffer and the aggregation accumulators alias the conv
// accumulators (shared memory and TMEM are both full), so conv[i+1] waits for the aggregation drain of sample i.
#pragma once
#include <cuda.h>

#include "fused_conv.cuh"
#include "kernels.cuh"
#include "tc_gemm.cuh"
#include "umma.cuh"

namespace gu {

constexpr int SA_WORKERS = 16;
constexpr int SA_THREADS = (SA_WORKERS + 2) * 32;   // 576

struct SpAggP {
  CUtensorMap tm_self[2];     // {w, S*R rows}, box {64, Rbox}, SWIZZLE_128B (K-major tiles of 64 columns)
  CUtensorMap tm_other;       // {F, S*Ro rows}, box {64, Kp}, SWIZZLE_128B (one MN-major tile, F == 64)
  int nseg, segw[2];
  const uint8_t* Wimg; uint32_t w_bytes;      // conv weight image, NW rows per panel (panel stride NW*16)
  const uint8_t* Mimg; uint32_t m_bytes;      // NodeEdge matrix image, panel stride pstrideM
  const float* bias;
  const int* rp;              // CSR row pointers (R+1), block-diagonal replica for sample groups
  const uint32_t* cj;         // per CSR entry: (byte offset of alpha[i][j] in the alpha operand) << 8 | j
  const float* val;           // GCN filter values per CSR entry (null for GAT)
  int nnz;
  bf16* out; long long rows_total;
  int S, R, Ro, Rbox, Kp, KpJ, MT, Kself, F, Nout, NW, act, gat, nb_self, alias_acc, max_len;
  uint32_t pstrideP, pstrideM, pstrideH, pstrideAl, chunk_bytes, self_bytes, xe_bytes, al_bytes, nx_bytes;
  uint32_t off_w, off_m, off_self, self_stride, off_xe, off_nx, off_al, off_sc, off_bar;
  long long* trace;           // debug: clock64 stamps of CTA 0, SA_NSTAMP per sample (GNNUDS_TRACE_FILE), else null
};
constexpr int SA_NSTAMP = 8;
#define SA_STAMP(i, k) do { if (p.trace && blockIdx.x == 0 && tid == 0 && (i) < 64) p.trace[(i) * SA_NSTAMP + (k)] = clock64(); } while (0)

__device__ __forceinline__ void workers_bar_sync() { asm volatile("bar.sync 4, %0;" ::"n"(SA_WORKERS * 32) : "memory"); }

// Edge softmax of one CSR row by a quad of lanes (lane `sub` takes entries sub, sub+4, ...), at most 4*KM entries.
// The weight is rounded to fp16 and stored straight into the dense alpha operand; the denominator sums the ROUNDED
// weights, so alpha/den stays a convex combination (the division happens in the aggregation drain).
template <int KM>
__device__ __forceinline__ void sa_softmax_row(int r, bool live, int sub, const int* rp_s, const uint32_t* cj_s, const float* ss,
                                               const float* sn, uint8_t* al, float* inv) {
  const int beg = live ? rp_s[r] : 0, end = live ? rp_s[r + 1] : 0;
  const float si = live ? ss[r] : 0.f;
  uint32_t ee[KM];
#pragma unroll
  for (int k = 0; k < KM; ++k) { const int qq = beg + sub + 4 * k; ee[k] = qq < end ? cj_s[qq] : 0u; }
  float sc[KM];
#pragma unroll
  for (int k = 0; k < KM; ++k) {
    float v = si + sn[ee[k] & 255u];
    v = v > 0.f ? v : 0.2f * v;
    sc[k] = beg + sub + 4 * k < end ? v : -INFINITY;
  }
  float mx = sc[0];
#pragma unroll
  for (int k = 1; k < KM; ++k) mx = fmaxf(mx, sc[k]);
  mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
  mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
  float den = 0.f;
#pragma unroll
  for (int k = 0; k < KM; ++k) {
    const bool on = beg + sub + 4 * k < end;
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"((sc[k] - mx) * 1.4426950408889634f));
    const __half wh = __float2half_rn(on ? e : 0.f);
    den += __half2float(wh);
    if (on) *reinterpret_cast<__half*>(al + (ee[k] >> 8)) = wh;
  }
  den += __shfl_xor_sync(0xffffffffu, den, 1);
  den += __shfl_xor_sync(0xffffffffu, den, 2);
  if (live && sub == 0) inv[r] = 1.f / den;
}

__global__ void __launch_bounds__(SA_THREADS, 1) sp_agg_kernel(const __grid_constant__ SpAggP p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.off_bar);
  uint64_t* self_full = bars;        // [2] TMA complete_tx, self tiles
  uint64_t* self_empty = bars + 2;   // [2] aggregation MMAs that read the h tile completed (tcgen05.commit)
  uint64_t* xe_full = bars + 4;      // TMA complete_tx, other_dense tile
  uint64_t* xe_empty = bars + 5;     // NodeEdge MMAs that read the other_dense tile completed
  uint64_t* ne_done = bars + 6;      // NodeEdge MMAs committed
  uint64_t* nx_ready = bars + 7;     // NodeEdge panels written, NodeEdge accumulators drained (workers)
  uint64_t* conv_done = bars + 8;    // conv MMAs committed
  uint64_t* h_ready = bars + 9;      // h tile + scores written, conv accumulators drained (workers)
  uint64_t* alpha_ready = bars + 10; // alpha entries written (workers)
  uint64_t* agg_done = bars + 11;    // aggregation MMAs committed
  uint64_t* out_done = bars + 12;    // aggregation accumulators drained (workers)
  uint64_t* wbar = bars + 13;        // resident weights landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 14);
  // scratch: bias[Nout] | ss[RP] sn[RP] inv[RP] | rp[R+1] | cj[nnz]
  const int RP = 128 * p.MT;
  float* bias_s = reinterpret_cast<float*>(smem + p.off_sc);
  float* ss = bias_s + p.Nout;
  float* sn = ss + RP;
  float* inv_s = sn + RP;
  int* rp_s = reinterpret_cast<int*>(inv_s + RP);
  uint32_t* cj_s = reinterpret_cast<uint32_t*>(rp_s + p.R + 1);

  if (warp == 0) tc::tmem_alloc(tmem_slot, 512);
  if (tid == 32) {
    for (int i = 0; i < 2; ++i) { tc::mbar_init(&self_full[i], 1); tc::mbar_init(&self_empty[i], 1); }
    tc::mbar_init(xe_full, 1);
    tc::mbar_init(xe_empty, 1);
    tc::mbar_init(ne_done, 1);
    tc::mbar_init(nx_ready, SA_WORKERS * 32);
    tc::mbar_init(conv_done, 1);
    tc::mbar_init(h_ready, SA_WORKERS * 32);
    tc::mbar_init(alpha_ready, SA_WORKERS * 32);
    tc::mbar_init(agg_done, 1);
    tc::mbar_init(out_done, SA_WORKERS * 32);
    tc::mbar_init(wbar, 1);
    tc::mbar_fence_init();
  }
  for (int i = tid; i < p.Nout; i += SA_THREADS) bias_s[i] = p.bias ? p.bias[i] : 0.f;
  for (int i = tid; i < RP; i += SA_THREADS) { inv_s[i] = 1.f; ss[i] = 0.f; sn[i] = 0.f; }
  for (int i = tid; i <= p.R; i += SA_THREADS) rp_s[i] = p.rp[i];
  for (int i = tid; i < p.nnz; i += SA_THREADS) cj_s[i] = p.cj[i];
  {  // zero the alpha operand and the NodeEdge panels once: everything off the CSR pattern / beyond row R stays zero
    uint4* z = reinterpret_cast<uint4*>(smem + p.off_al);
    for (int i = tid; i < (int)(p.al_bytes >> 4); i += SA_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
    z = reinterpret_cast<uint4*>(smem + p.off_nx);
    for (int i = tid; i < (int)(p.nx_bytes >> 4); i += SA_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  __syncthreads();
  if (!p.gat) {   // GCN: the filter values are the aggregation operand, written once
    for (int i = tid; i < p.nnz; i += SA_THREADS) *reinterpret_cast<__half*>(smem + p.off_al + (p.cj[i] >> 8)) = __float2half_rn(p.val[i]);
  }
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t acc_conv = tmem;                                  // MT * NW columns (<= 288)
  const uint32_t acc_agg = p.alias_acc ? tmem : tmem + 160u;       // MT * Nout columns
  const uint32_t acc_ne = tmem + 320u;                             // MT * F columns (<= 128)
  const int n_it = p.S > (int)blockIdx.x ? (p.S - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  const int NB = p.nb_self;

  if (warp == SA_WORKERS + 1) {
    // ================================ TMA producer ================================
    if (lane == 0 && n_it > 0) {
      for (int s = 0; s < p.nseg; ++s) tc::tma_prefetch_desc(&p.tm_self[s]);
      tc::tma_prefetch_desc(&p.tm_other);
      auto issue_xe = [&](int i) {
        const int smp = (int)blockIdx.x + i * (int)gridDim.x;
        if (i >= 1) tc::mbar_wait(xe_empty, (i - 1) & 1);
        tc::mbar_expect_tx(xe_full, p.xe_bytes);
        tc::tma_load_2d(smem + p.off_xe, &p.tm_other, 0, smp * p.Ro, xe_full);
      };
      auto issue_self = [&](int i) {
        const int buf = i % NB, use = i / NB;
        const int smp = (int)blockIdx.x + i * (int)gridDim.x;
        if (use >= 1) tc::mbar_wait(&self_empty[buf], (use - 1) & 1);
        uint8_t* dst = smem + p.off_self + (size_t)buf * p.self_stride;
        tc::mbar_expect_tx(&self_full[buf], p.self_bytes);
        uint32_t o = 0;
        for (int s = 0; s < p.nseg; ++s)
          for (int k0 = 0; k0 < p.segw[s]; k0 += 64, o += p.chunk_bytes) tc::tma_load_2d(dst + o, &p.tm_self[s], k0, smp * p.R, &self_full[buf]);
      };
      issue_xe(0);
      issue_self(0);
      for (int i = 0; i + 1 < n_it; ++i) { issue_xe(i + 1); issue_self(i + 1); }
    }
  } else if (warp == SA_WORKERS) {
    // ================================ MMA issuer ================================
    if (lane == 0 && n_it > 0) {
      tc::mbar_expect_tx(wbar, p.w_bytes + p.m_bytes);
      for (uint32_t o = 0; o < p.w_bytes; o += 16384) tc::bulk_g2s(smem + p.off_w + o, p.Wimg + o, min(p.w_bytes - o, 16384u), wbar);
      for (uint32_t o = 0; o < p.m_bytes; o += 16384) tc::bulk_g2s(smem + p.off_m + o, p.Mimg + o, min(p.m_bytes - o, 16384u), wbar);
      tc::mbar_wait(wbar, 0);
      const int F = p.F, Nout = p.Nout, NW = p.NW;
      const uint32_t idesc_ne = tc::idesc_bf16(128, F, 0, 1);          // B = other_dense, MN-major
      const uint32_t idesc_cv = tc::idesc_bf16(128, NW, 0, 0);
      const uint32_t idesc_ag = tc::idesc_bf16(128, Nout, 0, 1);       // B = h tile, MN-major
      const uint32_t pstrideW = (uint32_t)NW * 16u;
      const uint64_t w_desc0 = tc::smem_desc(tc::smem_u32(smem + p.off_w), pstrideW, 128u);
      const uint64_t m_desc0 = tc::smem_desc(tc::smem_u32(smem + p.off_m), p.pstrideM, 128u);
      const uint64_t nx_desc0 = tc::smem_desc(tc::smem_u32(smem + p.off_nx), p.pstrideP, 128u);
      const uint64_t al_desc0 = tc::smem_desc(tc::smem_u32(smem + p.off_al), p.pstrideAl, 128u);
      const uint64_t xe_desc0 = tc::smem_desc_sw128(tc::smem_u32(smem + p.off_xe));
      const uint64_t p_step = (uint64_t)((2u * p.pstrideP) >> 4), w_step = (uint64_t)((2u * pstrideW) >> 4);
      const uint64_t m_step = (uint64_t)((2u * p.pstrideM) >> 4), al_step = (uint64_t)((2u * p.pstrideAl) >> 4);
      const int ks_self = p.Kself >> 4, ks_ne = F >> 4, ks_kp = p.Kp >> 4, ks_j = p.KpJ >> 4;
      const bool MT2 = p.MT > 1, alias = p.alias_acc != 0;
      const uint32_t chunk_bytes = p.chunk_bytes, pstrideH = p.pstrideH;
      auto issue_ne = [&](int i) {
        tc::mbar_wait(xe_full, i & 1);
        tc::tc_fence_after();
        uint64_t ad = m_desc0, bd = xe_desc0;
        for (int ks = 0; ks < ks_kp; ++ks, ad += m_step, bd += 128u) {
          tc::umma(acc_ne, ad, bd, idesc_ne, ks != 0);
          if (MT2) tc::umma(acc_ne + (uint32_t)F, ad + 128u, bd, idesc_ne, ks != 0);
        }
        tc::umma_commit(ne_done);
        tc::umma_commit(xe_empty);
      };
      issue_ne(0);
      for (int i = 0; i < n_it; ++i) {
        const int buf = i % NB, use = i / NB;
        const uint32_t in_addr = tc::smem_u32(smem + p.off_self + (size_t)buf * p.self_stride);
        // ---- conv[i]: self columns first (ready early), then the NodeEdge columns ----
        tc::mbar_wait(&self_full[buf], use & 1);
        if (i >= 1) tc::mbar_wait(h_ready, (i - 1) & 1);                  // conv accumulators of sample i-1 drained
        if (alias && i >= 1) tc::mbar_wait(out_done, (i - 1) & 1);        // ... and the aggregation accumulators they alias
        tc::tc_fence_after();
        uint64_t wd = w_desc0;
        for (int c = 0; c < (ks_self >> 2); ++c) {
          uint64_t ad = tc::smem_desc_sw128(in_addr + (uint32_t)c * chunk_bytes);
#pragma unroll
          for (int j = 0; j < 4; ++j, ad += 2u, wd += w_step) {
            tc::umma(acc_conv, ad, wd, idesc_cv, (c | j) != 0);
            if (MT2) tc::umma(acc_conv + (uint32_t)NW, ad + 1024u, wd, idesc_cv, (c | j) != 0);
          }
        }
        tc::mbar_wait(nx_ready, i & 1);
        tc::tc_fence_after();
        {
          uint64_t ad = nx_desc0;
          for (int ks = 0; ks < ks_ne; ++ks, ad += p_step, wd += w_step) {
            tc::umma(acc_conv, ad, wd, idesc_cv, 1u);
            if (MT2) tc::umma(acc_conv + (uint32_t)NW, ad + 128u, wd, idesc_cv, 1u);
          }
        }
        tc::umma_commit(conv_done);
        // ---- NodeEdge[i+1] (its accumulators were drained before nx_ready[i]) ----
        if (i + 1 < n_it) issue_ne(i + 1);
        // ---- agg[i] ----
        tc::mbar_wait(alpha_ready, i & 1);
        if (!alias && i >= 1) tc::mbar_wait(out_done, (i - 1) & 1);       // aggregation accumulators of sample i-1 drained
        tc::tc_fence_after();
        {
          uint64_t ad = al_desc0, bd = tc::smem_desc(in_addr, 128u, pstrideH);
          for (int ks = 0; ks < ks_j; ++ks, ad += al_step, bd += 16u) {
            tc::umma(acc_agg, ad, bd, idesc_ag, ks != 0);
            if (MT2) tc::umma(acc_agg + (uint32_t)Nout, ad + 128u, bd, idesc_ag, ks != 0);
          }
        }
        tc::umma_commit(agg_done);
        tc::umma_commit(&self_empty[buf]);
      }
    }
  } else {
    // ================================ workers ================================
    const int q = warp & 3, cs = warp >> 2;            // TMEM lanes 32q .. 32q+31, column slice cs of 4
    const int R = p.R, MT = p.MT, NW = p.NW, Nout = p.Nout, F = p.F, KpJ = p.KpJ, gat = p.gat, act = p.act;
    const uint32_t pstrideH = p.pstrideH, pstrideP = p.pstrideP;
    uint8_t* al = smem + p.off_al;
    const int km = (p.max_len + 3) >> 2;
    auto drain_ne = [&](int i) {
      tc::mbar_wait(ne_done, i & 1);
      tc::tc_fence_after();
      for (int mt = 0; mt < MT; ++mt) {
        if (mt * 128 + q * 32 >= R) continue;          // warp-uniform
        const int row = mt * 128 + q * 32 + lane;
        float v[16];
        tc::tmem_ld16(acc_ne + ((uint32_t)(q * 32) << 16) + (uint32_t)(mt * F + cs * 16), v);
        tc::tmem_ld_wait();
        if (row < R) {
          uint8_t* dst = smem + p.off_nx + (size_t)row * 16 + (size_t)(cs * 2) * pstrideP;
          *reinterpret_cast<uint4*>(dst) = tc::pack8_bf16(v);
          *reinterpret_cast<uint4*>(dst + pstrideP) = tc::pack8_bf16(v + 8);
        }
      }
      tc::fence_async_smem();                          // generic-proxy panel writes -> tensor-core reads
      tc::tc_fence_before();
      mbar_arrive(nx_ready);
    };
    if (n_it > 0) drain_ne(0);
    for (int i = 0; i < n_it; ++i) {
      const int buf = i % NB;
      const int smp = (int)blockIdx.x + i * (int)gridDim.x;
      uint8_t* hbase = smem + p.off_self + (size_t)buf * p.self_stride;
      // ---- A: conv accumulators -> h tile (fp16 MN-major panels: feature panel c, row j at c*pstrideH + j*16) + scores ----
      tc::mbar_wait(conv_done, i & 1);
      tc::tc_fence_after();
      SA_STAMP(i, 0);
      for (int mt = 0; mt < MT; ++mt) {
        if (mt * 128 + q * 32 >= KpJ) continue;        // warp-uniform
        const int row = mt * 128 + q * 32 + lane;
        const uint32_t ta = acc_conv + ((uint32_t)(q * 32) << 16) + (uint32_t)(mt * NW);
        float v[32];
        tc::tmem_ld32(ta + (uint32_t)(cs * 32), v);
        float sv[8];
        if (gat && cs == 0) tc::tmem_ld8(ta + (uint32_t)Nout, sv);
        tc::tmem_ld_wait();
        if (row < KpJ) {
          const bool live = row < R;                   // rows R..KpJ-1 meet zero alpha columns: keep them finite (zero)
          uint8_t* hrow = hbase + (size_t)row * 16 + (size_t)(cs * 4) * pstrideH;
#pragma unroll
          for (int j = 0; j < 4; ++j)
            *reinterpret_cast<uint4*>(hrow + (size_t)j * pstrideH) = live ? tc::pack8_f16(v + 8 * j) : make_uint4(0u, 0u, 0u, 0u);
          if (gat && cs == 0 && live) { ss[row] = sv[0] + sv[1]; sn[row] = sv[2] + sv[3]; }
        }
      }
      tc::fence_async_smem();                          // h tile -> tensor-core reads
      tc::tc_fence_before();
      mbar_arrive(h_ready);
      SA_STAMP(i, 1);
      // ---- B: edge softmax over the CSR pattern -> alpha entries ----
      if (gat) {
        workers_bar_sync();                            // scores of every row visible
        for (int r0 = 0; r0 < R; r0 += SA_WORKERS * 8) {
          if (r0 + warp * 8 >= R) continue;            // warp-uniform (R = 131: only warp 0 runs the second pass)
          const int r = r0 + (tid >> 2);
          if (km <= 3) sa_softmax_row<3>(r, r < R, tid & 3, rp_s, cj_s, ss, sn, al, inv_s);
          else if (km == 4) sa_softmax_row<4>(r, r < R, tid & 3, rp_s, cj_s, ss, sn, al, inv_s);
          else sa_softmax_row<5>(r, r < R, tid & 3, rp_s, cj_s, ss, sn, al, inv_s);
        }
        tc::fence_async_smem();                        // alpha -> tensor-core reads
      }
      mbar_arrive(alpha_ready);
      SA_STAMP(i, 2);
      // ---- C: NodeEdge accumulators of the next sample -> fp16 panels ----
      if (i + 1 < n_it) drain_ne(i + 1);
      SA_STAMP(i, 3);
      // ---- D: aggregation accumulators -> 1/den, bias, act -> fp16 -> HBM ----
      tc::mbar_wait(agg_done, i & 1);
      tc::tc_fence_after();
      SA_STAMP(i, 4);
      for (int mt = 0; mt < MT; ++mt) {
        if (mt * 128 + q * 32 >= R) continue;          // warp-uniform
        const int row = mt * 128 + q * 32 + lane;
        float v[32];
        tc::tmem_ld32(acc_agg + ((uint32_t)(q * 32) << 16) + (uint32_t)(mt * Nout + cs * 32), v);
        float4 b4[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) b4[j] = *reinterpret_cast<const float4*>(bias_s + cs * 32 + 4 * j);
        const float sc = row < R ? inv_s[row] : 0.f;
        tc::tmem_ld_wait();
        if (row < R && (long long)smp * R + row < p.rows_total) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            v[4 * j] = fmaf(v[4 * j], sc, b4[j].x); v[4 * j + 1] = fmaf(v[4 * j + 1], sc, b4[j].y);
            v[4 * j + 2] = fmaf(v[4 * j + 2], sc, b4[j].z); v[4 * j + 3] = fmaf(v[4 * j + 3], sc, b4[j].w);
          }
          act_tile<32>(v, act);
          bf16* orow = p.out + ((size_t)smp * R + row) * Nout + cs * 32;
          const uint4 u0 = tc::pack8_bf16(v), u1 = tc::pack8_bf16(v + 8), u2 = tc::pack8_bf16(v + 16), u3 = tc::pack8_bf16(v + 24);
          st_global_v8(orow, u0, u1);
          st_global_v8(orow + 16, u2, u3);
        }
      }
      tc::tc_fence_before();
      mbar_arrive(out_done);
      SA_STAMP(i, 5);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_free(tmem, 512);
}

}  // namespace gu

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
//   warps 4..19  "graph" group : edge softmax + weighted neighbour sum over the h tile -> bias/act -> HBM
//   warp  20     MMA issuer    : tcgen05.mma  NE: D[128 x F] = Mat(K-major) x other_dense(MN-major)
//                                             conv: D[128 x Nout] = [self | NE](K-major) x Wconv(K-major)
//   warp  21     TMA producer  : cp.async.bulk.tensor loads of sample i+2 once the graph group released the buffer
// The h tile of a sample aliases that sample's own input buffer (dead once the conv MMAs committed), so
// the graph group works on sample i while MMA + drain already process sample i+1 in the other buffer.
#pragma once
#include <cuda.h>

#include "kernels.cuh"
#include "tc_gemm.cuh"
#include "umma.cuh"

namespace gu {

constexpr int SP_DRAIN_WARPS = 4;
constexpr int SP_GRAPH_WARPS = 16;
constexpr int SP_THREADS = (SP_DRAIN_WARPS + SP_GRAPH_WARPS + 2) * 32;   // 704
constexpr int SP_BINS = SP_GRAPH_WARPS * 4;                               // 8-lane groups of the graph warps

struct SpConvP {
  CUtensorMap tm_self[2];     // {8, S*R rows, w/8 panels}, box {8, Rbox, w/8};  sw128: {w, S*R rows}, box {64, Rbox}, SWIZZLE_128B
  int sw_x;                   // other_dense as one SWIZZLE_128B MN-major tile [Kp x 64] (F == 64): map {F, S*Ro rows}, box {64, Kp}
  int sw128; uint32_t chunk_bytes;   // self operand as SWIZZLE_128B K-major tiles of Rbox*128 bytes per 64 columns
  CUtensorMap tm_other;       // {8, S*Ro rows, F/8 panels}, box {8, Kp, F/8}
  int nseg, segw[2];
  const uint8_t* Wimg; uint32_t w_bytes;      // conv weight image, panel stride Nout*16
  const uint8_t* Mimg; uint32_t m_bytes;      // NodeEdge matrix image, panel stride pstrideM
  const float* bias; const float* a_self; const float* a_neigh;
  const int* row_ptr; const int* col_idx; const float* val; int nnz;
  const int* bins;            // gather schedule: bin_ptr[SP_BINS+1] | bin_rows[R]  (gnnuds.cu::upload_bins)
  bf16* out;                  // (rows_total, Nout)
  long long rows_total;       // true row count of `out` (a partial last sample group stores fewer than R rows)
  int S, R, Ro, Rbox, Kp, MT, Kself, F, Nout, act, gat;
  uint32_t pstrideA, pstrideM, pstrideX;      // bytes: self/NE panels (Rbox*16), Mat panels, other_dense panels (Kp*16)
  uint32_t off_w, off_m, off_in, in_stride, xe_off, off_nx, off_sc, off_bar, hsb;
  uint32_t self_bytes, xe_bytes;
  int max_len;                // longest CSR row (true entries): <= 20 selects the register softmax
  int dbg;                    // GNNUDS_SP_DBG timing experiments (results are wrong when set): 1 no h loads, 2 no neighbour blocks, 4 no softmax
  long long* trace;           // debug: clock64 stamps of CTA 0, SP_NSTAMP per sample (GNNUDS_TRACE_FILE), else null
};
constexpr int SP_NSTAMP = 12;
#define SP_STAMP(i, k) do { if (p.trace && blockIdx.x == 0 && (i) < 64) p.trace[(i) * SP_NSTAMP + (k)] = clock64(); } while (0)

// 16-byte shared load that is not issued when `on` is false (zeros instead): the pads of a neighbour block carry
// weight 0, and a quarter-warp whose lanes are all predicated off costs no shared-memory wavefront
template <int OFS>
__device__ __forceinline__ uint4 lds128_if(uint32_t saddr, bool on) {
  uint4 v = make_uint4(0u, 0u, 0u, 0u);
  asm("{\n\t.reg .pred q;\n\tsetp.ne.u32 q, %5, 0;\n\t@q ld.shared.v4.u32 {%0, %1, %2, %3}, [%4+%6];\n\t}"
      : "+r"(v.x), "+r"(v.y), "+r"(v.z), "+r"(v.w)
      : "r"(saddr), "r"((uint32_t)on), "n"(OFS));
  return v;
}

// Edge softmax of one CSR row by a quad of lanes (lane `sub` takes the neighbours sub, sub+4, ...), at most 4*KM
// neighbours: scores stay in registers, so the row costs two rounds of independent shared loads.  The weight is stored
// as an fp16 pair and the denominator sums the ROUNDED weights, so the attention stays a convex combination; the pads
// of the last block of 4 get weight 0.  ci holds byte offsets into the h tile (index << hshift | swizzle bits).
template <int KM>
__device__ __forceinline__ void softmax_row(int r, int R, int sub, int hshift, const int* rp_s, const int* rlen_s, const int* ci_s,
                                            const float* ssb, const float* snb, float* al, float* inv) {
  const bool live = r < R;
  const int beg = live ? rp_s[r] : 0, pend = live ? rp_s[r + 1] : 0, end = live ? beg + rlen_s[r] : 0;
  const float si = live ? ssb[r] : 0.f;
  int jj[KM];
#pragma unroll
  for (int k = 0; k < KM; ++k) { const int qq = beg + sub + 4 * k; jj[k] = qq < end ? (ci_s[qq] >> hshift) : 0; }
  float sc[KM];
#pragma unroll
  for (int k = 0; k < KM; ++k) {
    float v = si + snb[jj[k]];
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
    const int qq = beg + sub + 4 * k;
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"((sc[k] - mx) * 1.4426950408889634f));
    const __half wh = __float2half_rn(qq < end ? e : 0.f);
    den += __half2float(wh);
    if (qq < pend) al[qq] = __uint_as_float((uint32_t)__half_as_ushort(wh) * 0x10001u);
  }
  den += __shfl_xor_sync(0xffffffffu, den, 1);
  den += __shfl_xor_sync(0xffffffffu, den, 2);
  if (live && sub == 0) inv[r] = 1.f / den;
}

__device__ __forceinline__ void graph_bar_sync() { asm volatile("bar.sync 2, %0;" ::"n"(SP_GRAPH_WARPS * 32) : "memory"); }

__global__ void __launch_bounds__(SP_THREADS, 1) sp_conv_kernel(const __grid_constant__ SpConvP p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.off_bar);
  uint64_t* xe_full = bars;          // [2] TMA complete_tx, other_dense tile
  uint64_t* self_full = bars + 2;    // [2] TMA complete_tx, self tiles
  uint64_t* xe_empty = bars + 4;     // [2] NodeEdge MMAs that read the other_dense tile completed (tcgen05.commit)
  uint64_t* self_empty = bars + 6;   // [2] graph group released the self region (h tile consumed)
  uint64_t* h_full = bars + 8;       // [2] drain group wrote the h tile
  uint64_t* ne_done = bars + 10;     // NodeEdge MMAs committed
  uint64_t* nx_ready = bars + 11;    // NE panels written (drain group)
  uint64_t* conv_done = bars + 12;   // conv MMAs committed
  uint64_t* wbar = bars + 13;        // resident weights landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 14);
  // scratch: bias[Nout] asf[Nout] anf[Nout] | ss[2][RP] sn[2][RP] | inv[2][RP] | alpha[2][nnz] | ci[nnz] | rp[R+1] | bins
  // (nnz counts the PADDED CSR: rows start on multiples of 4 entries, so alpha / ci blocks are 16-byte aligned)
  const int RP = 128 * p.MT;
  float* bias_s = reinterpret_cast<float*>(smem + p.off_sc);
  float* asf = bias_s + p.Nout;
  float* anf = asf + p.Nout;
  float* ss = anf + p.Nout;
  float* sn = ss + 2 * RP;
  float* inv_s = sn + 2 * RP;
  float* alpha = inv_s + 2 * RP;
  int* ci_s = reinterpret_cast<int*>(alpha + 2 * p.nnz);
  int* rp_s = ci_s + p.nnz;
  int* bin_s = rp_s + p.R + 1;       // bin_ptr[SP_BINS+1] | bin_rows[R] | row_len[R]
  const int* rlen_s = bin_s + SP_BINS + 1 + p.R;

  if (warp == 0) tc::tmem_alloc(tmem_slot, 512);
  if (tid == 32) {
    for (int i = 0; i < 2; ++i) {
      tc::mbar_init(&xe_full[i], 1);
      tc::mbar_init(&self_full[i], 1);
      tc::mbar_init(&xe_empty[i], 1);
      tc::mbar_init(&self_empty[i], SP_GRAPH_WARPS * 32);
      tc::mbar_init(&h_full[i], SP_DRAIN_WARPS * 32);
    }
    tc::mbar_init(ne_done, 1);
    tc::mbar_init(nx_ready, SP_DRAIN_WARPS * 32);
    tc::mbar_init(conv_done, 1);
    tc::mbar_init(wbar, 1);
    tc::mbar_fence_init();
  }
  // constants into scratch (all threads), zero the NE panels once (rows >= R stay zero/finite)
  // fp16 build, relu / linear activation: the aggregation epilogue stays in half2 (scale, bias, relu, store) and reads
  // the bias as fp16 from the same scratch slot
  const bool fast_epi = GU_F16 && (p.act == ACT_RELU || p.act == ACT_LINEAR);
  for (int i = tid; i < p.Nout; i += SP_THREADS) {
    if (fast_epi) reinterpret_cast<__half*>(bias_s)[i] = __float2half_rn(p.bias ? p.bias[i] : 0.f);
    else bias_s[i] = p.bias ? p.bias[i] : 0.f;
    asf[i] = p.gat ? p.a_self[i] : 0.f;
    anf[i] = p.gat ? p.a_neigh[i] : 0.f;
  }
  for (int i = tid; i <= p.R; i += SP_THREADS) rp_s[i] = p.row_ptr[i];
  for (int i = tid; i < SP_BINS + 1 + 2 * p.R; i += SP_THREADS) bin_s[i] = p.bins[i];
  for (int i = tid; i < p.nnz; i += SP_THREADS) { const int j = p.col_idx[i]; ci_s[i] = j * (int)p.hsb + ((j & 7) << 4); if (!p.gat) { const float v = __uint_as_float(tc::pack_f16(p.val[i], p.val[i])); alpha[i] = v; alpha[p.nnz + i] = v; } }
  {
    uint4* z = reinterpret_cast<uint4*>(smem + p.off_nx);
    const int n16 = (int)((p.F / 8) * p.pstrideA / 16);
    for (int i = tid; i < n16; i += SP_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  tc::pdl_sync();
  const uint32_t tmem = *tmem_slot;
  const uint32_t acc_conv = tmem;              // MT * Nout columns
  const uint32_t acc_ne = tmem + 256u;         // MT * F columns
  const int n_it = p.S > (int)blockIdx.x ? (p.S - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

  if (warp == SP_DRAIN_WARPS + SP_GRAPH_WARPS + 1) {
    // ================================ TMA producer ================================
    if (lane == 0) {
      for (int s = 0; s < p.nseg; ++s) tc::tma_prefetch_desc(&p.tm_self[s]);
      tc::tma_prefetch_desc(&p.tm_other);
      // The two halves of an input buffer are released at different times: the other_dense tile as soon as the
      // NodeEdge MMAs have read it, the self tiles (whose bytes then hold the h tile) only when the graph group is
      // done with the sample.  So NodeEdge + its drain of sample i+2 run while the graph group still works on
      // sample i, and only [self load -> conv -> drain] remains between two samples of the graph group.
      auto issue_xe = [&](int i) {
        const int buf = i & 1;
        const int smp = (int)blockIdx.x + i * (int)gridDim.x;
        if (i >= 2) tc::mbar_wait(&xe_empty[buf], ((i >> 1) & 1) ^ 1);
        uint8_t* dst = smem + p.off_in + (size_t)buf * p.in_stride + p.xe_off;
        tc::mbar_expect_tx(&xe_full[buf], p.xe_bytes);
        if (p.sw_x) tc::tma_load_2d(dst, &p.tm_other, 0, smp * p.Ro, &xe_full[buf]);
        else tc::tma_load_3d(dst, &p.tm_other, 0, smp * p.Ro, 0, &xe_full[buf]);
      };
      if (n_it > 0) issue_xe(0);
      for (int i = 0; i < n_it; ++i) {
        const int buf = i & 1;
        const int smp = (int)blockIdx.x + i * (int)gridDim.x;
        if (i >= 2) tc::mbar_wait_sleep(&self_empty[buf], ((i >> 1) & 1) ^ 1, 64);
        SP_STAMP(i, 0);
        uint8_t* dst = smem + p.off_in + (size_t)buf * p.in_stride;
        tc::mbar_expect_tx(&self_full[buf], p.self_bytes);
        uint32_t o = 0;
        for (int s = 0; s < p.nseg; ++s) {
          if (p.sw128) {        // one SWIZZLE_128B tile [Rbox x 64] per 64 columns (128-byte TMA requests)
            for (int k0 = 0; k0 < p.segw[s]; k0 += 64, o += p.chunk_bytes) tc::tma_load_2d(dst + o, &p.tm_self[s], k0, smp * p.R, &self_full[buf]);
          } else {
            tc::tma_load_3d(dst + o, &p.tm_self[s], 0, smp * p.R, 0, &self_full[buf]);
            o += (uint32_t)(p.segw[s] >> 3) * p.pstrideA;
          }
        }
        if (i + 1 < n_it) issue_xe(i + 1);
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
      const int ks_self = p.Kself >> 4, ks_ne = p.F >> 4, ks_kp = p.Kp >> 4, F = p.F, Nout = p.Nout;
      const bool MT2 = p.MT > 1;
      const uint32_t pstrideA = p.pstrideA, pstrideX = p.pstrideX, xe_off = p.xe_off, off_in = p.off_in, in_stride = p.in_stride;
      const uint64_t w_desc0 = tc::smem_desc(w_addr, pstrideW, 128u), m_desc0 = tc::smem_desc(m_addr, p.pstrideM, 128u);
      const uint64_t nx_desc0 = tc::smem_desc(nx_addr, pstrideA, 128u);
      const uint64_t a_step = (uint64_t)((2u * pstrideA) >> 4), w_step = (uint64_t)((2u * pstrideW) >> 4), m_step = (uint64_t)((2u * p.pstrideM) >> 4);
      for (int i = 0; i < n_it; ++i) {
        const int buf = i & 1;
        const uint32_t in_addr = tc::smem_u32(smem + off_in + (size_t)buf * in_stride);
        tc::mbar_wait(&xe_full[buf], (i >> 1) & 1);
        tc::tc_fence_after();
        SP_STAMP(i, 1);
        // NodeEdge: D[row, f] = sum_k Mat[row, k] * other_dense[k, f]
        // descriptors advance linearly with the start address: one 64-bit add per MMA in the issue loops
        {
          // B = other_dense[s], MN-major: 16 k-rows per step = 256 bytes of a panel, or two 1024-byte swizzle atoms
          uint64_t ad = m_desc0, bd = p.sw_x ? tc::smem_desc_sw128(in_addr + xe_off) : tc::smem_desc(in_addr + xe_off, 128u, pstrideX);
          const uint64_t b_step = p.sw_x ? 128u : 16u;
          for (int ks = 0; ks < ks_kp; ++ks, ad += m_step, bd += b_step) {
            tc::umma(acc_ne, ad, bd, idesc_ne, ks != 0);
            if (MT2) tc::umma(acc_ne + (uint32_t)F, ad + 128u, bd, idesc_ne, ks != 0);
          }
        }
        tc::umma_commit(ne_done);
        tc::umma_commit(&xe_empty[buf]);
        SP_STAMP(i, 2);
        // conv over the self columns first (they are ready), then over the NodeEdge columns.  The conv
        // accumulators of the previous sample must have been drained (h_full of sample i-1).
        if (i >= 1) { tc::mbar_wait(&h_full[(i - 1) & 1], ((i - 1) >> 1) & 1); tc::tc_fence_after(); }
        tc::mbar_wait(&self_full[buf], (i >> 1) & 1);
        tc::tc_fence_after();
        uint64_t wd = w_desc0;
        {
          if (p.sw128) {
            for (int c = 0; c < (ks_self >> 2); ++c) {
              uint64_t ad = tc::smem_desc_sw128(in_addr + (uint32_t)c * p.chunk_bytes);
#pragma unroll
              for (int j = 0; j < 4; ++j, ad += 2u, wd += w_step) {
                tc::umma(acc_conv, ad, wd, idesc_cv, (c | j) != 0);
                if (MT2) tc::umma(acc_conv + (uint32_t)Nout, ad + 1024u, wd, idesc_cv, (c | j) != 0);
              }
            }
          } else {
            uint64_t ad = tc::smem_desc(in_addr, pstrideA, 128u);
#pragma unroll 4
            for (int ks = 0; ks < ks_self; ++ks, ad += a_step, wd += w_step) {
              tc::umma(acc_conv, ad, wd, idesc_cv, ks != 0);
              if (MT2) tc::umma(acc_conv + (uint32_t)Nout, ad + 128u, wd, idesc_cv, ks != 0);
            }
          }
        }
        tc::mbar_wait(nx_ready, i & 1);
        tc::tc_fence_after();
        {
          uint64_t ad = nx_desc0;
          for (int ks = 0; ks < ks_ne; ++ks, ad += a_step, wd += w_step) {
            tc::umma(acc_conv, ad, wd, idesc_cv, 1u);
            if (MT2) tc::umma(acc_conv + (uint32_t)Nout, ad + 128u, wd, idesc_cv, 1u);
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
      if (tid == 0) SP_STAMP(i, 3);
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
      if (tid == 0) SP_STAMP(i, 4);
      // ---- conv accumulators -> h tile (bf16 rows) + GAT scores from the fp32 values ----
      tc::mbar_wait(conv_done, i & 1);
      tc::tc_fence_after();
      if (tid == 0) SP_STAMP(i, 5);
      uint8_t* hbase = smem + p.off_in + (size_t)buf * p.in_stride;
      for (int mt = 0; mt < p.MT; ++mt) {
        if (mt * 128 + q * 32 >= p.R) continue;
        const int row = mt * 128 + q * 32 + lane;
        const uint32_t ta = acc_conv + ((uint32_t)(q * 32) << 16) + (uint32_t)(mt * p.Nout);
        uint8_t* hrow = hbase + (size_t)row * p.hsb;       // 16-byte chunk c of a row lives at chunk c ^ (row & 7)
        const int hsw = row & 7;
        // scores from the fp32 accumulators: two independent chains per vector, 16-byte broadcast loads of the
        // attention vectors (8 columns at a time keeps the 704-thread register budget)
        float ps0 = 0.f, ps1 = 0.f, pn0 = 0.f, pn1 = 0.f;
        for (int c0 = 0; c0 < p.Nout; c0 += 32) {
          float v[32];
          tc::tmem_ld32(ta + (uint32_t)c0, v);
          tc::tmem_ld_wait();
          if (row < p.R) {
#pragma unroll
            for (int j = 0; j < 4; ++j) *reinterpret_cast<uint4*>(hrow + ((((c0 >> 3) + j) ^ hsw) << 4)) = tc::pack8_f16(v + 8 * j);
            if (p.gat) {
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float4 a = *reinterpret_cast<const float4*>(asf + c0 + 4 * j);
                const float4 b = *reinterpret_cast<const float4*>(anf + c0 + 4 * j);
                ps0 = fmaf(v[4 * j], a.x, ps0); ps1 = fmaf(v[4 * j + 1], a.y, ps1);
                ps0 = fmaf(v[4 * j + 2], a.z, ps0); ps1 = fmaf(v[4 * j + 3], a.w, ps1);
                pn0 = fmaf(v[4 * j], b.x, pn0); pn1 = fmaf(v[4 * j + 1], b.y, pn1);
                pn0 = fmaf(v[4 * j + 2], b.z, pn0); pn1 = fmaf(v[4 * j + 3], b.w, pn1);
              }
            }
          }
        }
        if (p.gat && row < p.R) { ss[buf * RP + row] = ps0 + ps1; sn[buf * RP + row] = pn0 + pn1; }
      }
      tc::tc_fence_before();
      mbar_arrive(&h_full[buf]);
      if (tid == 0) SP_STAMP(i, 6);
    }
  } else {
    // ================================ graph group ================================
    const int gt = tid - SP_DRAIN_WARPS * 32;         // 0..511
    const int gw = gt >> 5;
    const int lg = lane >> 3, gl = lane & 7;
    const int R = p.R;
    const bool two = p.Nout > 64;                     // a lane owns 8 features of each 64-feature half
    const uint32_t glsh = (uint32_t)gl << 4;
    for (int i = 0; i < n_it; ++i) {
      const int buf = i & 1;
      const int smp = (int)blockIdx.x + i * (int)gridDim.x;
      tc::mbar_wait_sleep(&h_full[buf], (i >> 1) & 1, 64);
      if (gt == 0) SP_STAMP(i, 7);
      const uint8_t* hbase = smem + p.off_in + (size_t)buf * p.in_stride;
      float* al = alpha + buf * p.nnz;
      float* inv = inv_s + buf * RP;
      const int hshift = 31 - __clz((int)p.hsb);         // h rows are 128 or 256 bytes
      if (p.gat && !(p.dbg & 4) && p.max_len <= 20) {
        // one row per quad and pass; a warp whose 8 rows are all beyond R skips the pass (R = 131: only warp 0 runs the
        // second one).  KM is sized to the longest row of the graph.
        const float* ssb = ss + buf * RP;
        const float* snb = sn + buf * RP;
        const int km = (p.max_len + 3) >> 2;
        for (int r0 = 0; r0 < R; r0 += SP_GRAPH_WARPS * 8) {
          // the short tail pass goes to the LAST warps, which carry the lightest gather lists
          const int gwp = r0 ? SP_GRAPH_WARPS - 1 - gw : gw;
          if (r0 + gwp * 8 >= R) continue;
          const int r = r0 + gwp * 8 + (lane >> 2);
          if (km <= 3) softmax_row<3>(r, R, gt & 3, hshift, rp_s, rlen_s, ci_s, ssb, snb, al, inv);
          else if (km == 4) softmax_row<4>(r, R, gt & 3, hshift, rp_s, rlen_s, ci_s, ssb, snb, al, inv);
          else softmax_row<5>(r, R, gt & 3, hshift, rp_s, rlen_s, ci_s, ssb, snb, al, inv);
        }
        graph_bar_sync();
      } else if (p.gat && !(p.dbg & 4)) {
        // edge softmax over the CSR pattern: 4 lanes per row, neighbours strided over the lanes
        const float* ssb = ss + buf * RP;
        const float* snb = sn + buf * RP;
        const int sub = gt & 3;
        for (int r0 = 0; r0 < R; r0 += SP_GRAPH_WARPS * 8) {
          const int r = r0 + (gt >> 2);
          const bool live = r < R;
          const int beg = live ? rp_s[r] : 0, pend = live ? rp_s[r + 1] : 0, end = live ? beg + rlen_s[r] : 0;
          const float si = live ? ssb[r] : 0.f;
          float mx = -INFINITY;
          for (int qq = end + ((sub - end) & 3); qq < pend; qq += 4) al[qq] = 0.f;     // pads of the last block: weight 0
          for (int qq = beg + sub; qq < end; qq += 4) {
            float sc = si + snb[ci_s[qq] >> hshift];
            sc = sc > 0.f ? sc : 0.2f * sc;
            al[qq] = sc;
            mx = fmaxf(mx, sc);
          }
          mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
          mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
          float den = 0.f;
          for (int qq = beg + sub; qq < end; qq += 4) {
            // the weight is stored as an fp16 pair; the denominator sums the rounded weights so that the
            // attention stays a convex combination
            const __half wh = __float2half_rn(__expf(al[qq] - mx));
            al[qq] = __uint_as_float(tc::pack_f16(0.f, 0.f) | ((uint32_t)__half_as_ushort(wh) * 0x10001u));
            den += __half2float(wh);
          }
          den += __shfl_xor_sync(0xffffffffu, den, 1);
          den += __shfl_xor_sync(0xffffffffu, den, 2);
          if (live && sub == 0) inv[r] = 1.f / den;
        }
        graph_bar_sync();
      }
      if (gt == 0) SP_STAMP(i, 8);
      // weighted neighbour sum: 8 lanes per output row, a lane owns features [8gl, 8gl+8) and [64+8gl, 64+8gl+8).
      // Each 8-lane group walks its own row list of the balanced schedule; two neighbours per step keep four
      // 16-byte shared loads in flight.
      const int bin = gw * 4 + lg;
      for (int bi = bin_s[bin]; bi < bin_s[bin + 1]; ++bi) {
        const int r = bin_s[SP_BINS + 1 + bi];
        const int beg = rp_s[r], end = (p.dbg & 2) ? beg : rp_s[r + 1], tend = beg + rlen_s[r];
        const float scale = p.gat ? inv[r] : 1.f;
        const bool store = (long long)smp * R + r < p.rows_total;
        bf16* orow = p.out + ((size_t)smp * R + r) * p.Nout + gl * 8;
        for (int e0 = 0; e0 < p.Nout; e0 += 128) {
          // fp16 h tile, fp16 weights, HFMA2 accumulation (two features per instruction, no unpacking); the few
          // (<= ~20) terms per row keep the fp16 partial sums within ~1e-3 relative, below the bf16 output rounding
          __half2 h0[4], h1[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) { h0[k] = __float2half2_rn(0.f); h1[k] = h0[k]; }
          const uint32_t hb32 = tc::smem_u32(hbase) + (uint32_t)e0 * 2u;
          // one block = 4 neighbours: their indices and weights arrive as two 16-byte loads, then eight 16-byte
          // h loads are in flight before the first HFMA2 (pads carry weight 0)
          for (int qq = beg; qq < end; qq += 4) {
            const int4 jj = *reinterpret_cast<const int4*>(ci_s + qq);
            const uint4 ww = *reinterpret_cast<const uint4*>(al + qq);
            const uint32_t a0 = hb32 + ((uint32_t)jj.x ^ glsh), a1 = hb32 + ((uint32_t)jj.y ^ glsh);
            const uint32_t a2 = hb32 + ((uint32_t)jj.z ^ glsh), a3 = hb32 + ((uint32_t)jj.w ^ glsh);
            const bool nl = p.dbg & 1;
            const bool k1 = !nl && qq + 1 < tend, k2 = !nl && qq + 2 < tend, k3 = !nl && qq + 3 < tend;
            const uint4 u0 = lds128_if<0>(a0, !nl), u1 = lds128_if<0>(a1, k1);
            const uint4 u2 = lds128_if<0>(a2, k2), u3 = lds128_if<0>(a3, k3);
            if (two) {
              const uint4 v0 = lds128_if<128>(a0, !nl), v1 = lds128_if<128>(a1, k1);
              const uint4 v2 = lds128_if<128>(a2, k2), v3 = lds128_if<128>(a3, k3);
              tc::hfma8(h1, v0, tc::as_half2(ww.x)); tc::hfma8(h1, v1, tc::as_half2(ww.y));
              tc::hfma8(h1, v2, tc::as_half2(ww.z)); tc::hfma8(h1, v3, tc::as_half2(ww.w));
            }
            tc::hfma8(h0, u0, tc::as_half2(ww.x)); tc::hfma8(h0, u1, tc::as_half2(ww.y));
            tc::hfma8(h0, u2, tc::as_half2(ww.z)); tc::hfma8(h0, u3, tc::as_half2(ww.w));
          }
#if GU_F16
          if (fast_epi) {
            const __half2 sc2 = __float2half2_rn(scale), z2 = __float2half2_rn(0.f);
            const bool relu = p.act == ACT_RELU;
            const __half* bh = reinterpret_cast<const __half*>(bias_s) + e0 + gl * 8;
            {
              const uint4 bq = *reinterpret_cast<const uint4*>(bh);
              const __half2* b2 = reinterpret_cast<const __half2*>(&bq);
              uint4 o;
              __half2* o2 = reinterpret_cast<__half2*>(&o);
#pragma unroll
              for (int k = 0; k < 4; ++k) { const __half2 v = __hfma2(h0[k], sc2, b2[k]); o2[k] = relu ? __hmax2(v, z2) : v; }
              if (store) *reinterpret_cast<uint4*>(orow + e0) = o;
            }
            if (two) {
              const uint4 bq = *reinterpret_cast<const uint4*>(bh + 64);
              const __half2* b2 = reinterpret_cast<const __half2*>(&bq);
              uint4 o;
              __half2* o2 = reinterpret_cast<__half2*>(&o);
#pragma unroll
              for (int k = 0; k < 4; ++k) { const __half2 v = __hfma2(h1[k], sc2, b2[k]); o2[k] = relu ? __hmax2(v, z2) : v; }
              if (store) *reinterpret_cast<uint4*>(orow + e0 + 64) = o;
            }
            continue;
          }
#endif
          float a0[8], a1[8];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const float2 f0 = __half22float2(h0[k]), f1 = __half22float2(h1[k]);
            a0[2 * k] = f0.x; a0[2 * k + 1] = f0.y; a1[2 * k] = f1.x; a1[2 * k + 1] = f1.y;
          }
          const float* bp = bias_s + e0 + gl * 8;
#pragma unroll
          for (int k = 0; k < 8; ++k) a0[k] = fmaf(a0[k], scale, bp[k]);
          act_tile<8>(a0, p.act);
          if (store) *reinterpret_cast<uint4*>(orow + e0) = tc::pack8_bf16(a0);
          if (two) {
#pragma unroll
            for (int k = 0; k < 8; ++k) a1[k] = fmaf(a1[k], scale, bp[64 + k]);
            act_tile<8>(a1, p.act);
            if (store) *reinterpret_cast<uint4*>(orow + e0 + 64) = tc::pack8_bf16(a1);
          }
        }
      }
      if (gt == 0) SP_STAMP(i, 9);
      if (p.trace && blockIdx.x == 0 && i < 64 && lane == 0) {     // slowest / per-warp-0..3 spread of the gather
        atomicMax(reinterpret_cast<unsigned long long*>(p.trace + i * SP_NSTAMP + 10), (unsigned long long)clock64());
        if (gw == 15) p.trace[i * SP_NSTAMP + 11] = clock64();
      }
      mbar_arrive(&self_empty[buf]);                  // this thread is done with the h tile of the sample
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_free(tmem, 512);
}

}  // namespace gu

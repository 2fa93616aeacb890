// Fused spatial-layer kernel, second generation (sm_100a), one stream side per launch:
//
//   out[s] = act( alpha[s] @ ( [ self[s] | Mat @ other_dense[s] ] @ Wconv ) + bias )          (emulator.py:227-230)
//
// Same contract as sp_conv_kernel (fused_sp.cuh) - NodeEdge (K2), the concat, the GCN/GAT linear part and the graph
// aggregation (K4/K5) of one (b,t) sample with no intermediate leaving the SM - but EVERY contraction runs on tcgen05,
// the neighbour sum included: the attention (GAT) / filter (GCN) matrix of a 30..256-row drainage graph is tiny, so it
// is kept as a dense fp16 K-major A operand in shared memory (zeros off the CSR pattern, the pattern's entries
// rewritten per sample by the edge softmax) and `alpha @ h` is one more MMA series.  The SIMT gather of sp_conv_kernel
// (550 instructions per warp and sample, 6.4 k of its 8.4 k cycle period) becomes <= 18 MMAs + one TMEM drain.
//
// GAT scores also come from the tensor core: the conv weight image carries 16 extra output columns,
// [hi(W a_self) lo(W a_self) hi(W a_neigh) lo(W a_neigh) 0...], so h.a_self and h.a_neigh arrive as accumulator
// columns Nout..Nout+3 (hi + lo = fp16 pair split of the fp32 product W a, ~22 mantissa bits) and the conv drain is a
// pure TMEM -> fp16 copy.
//
// Persistent, one CTA per SM, one sample per pass:
//   warps 0..15  workers: every TMEM drain and the softmax, in phases (a warp owns TMEM lanes 32*(w%4).. and the
//                column slice w/4):
//                  A  conv drain -> h^T (fp16 K-major SWIZZLE_128B tiles [features x 64 neighbours], written transposed
//                     through tcgen05.ld.16x256b fragments + stmatrix.trans) + scores
//                  C  NodeEdge drain -> fp16 panels (A operand of a later conv GEMM)
//                  B  edge softmax -> alpha entries (GAT) / filter values (GCN)
//                  D  aggregation drain -> 1/den, bias, act -> fp16, staged in the alpha region as SWIZZLE_128B tiles
//                Order A C B D with NodeEdge one sample ahead (MT == 2), A B C D with NodeEdge TWO samples ahead and C in
//                the window where the workers would wait for the aggregation MMAs (MT == 1).
//   warp  16     MMA issuer (warp-uniform code, one elected lane): per sample NE[i+2], conv[i+1], agg[i] (MT == 1) or
//                agg[i], conv[i+1], NE[i+2] (MT == 2)
//   warp  17     TMA producer: other_dense tile (single buffer, free again once the NodeEdge MMAs read it) and the
//                self tiles (nb_self buffers; h^T of a sample aliases its own self tiles)
//   warp  18     store warp: TMA store of the staged output tile, then a bulk copy of zeros re-arms the alpha region
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
constexpr int SA_THREADS = (SA_WORKERS + 3) * 32;   // 608: workers, MMA issuer, TMA producer, store warp

struct SpAggP {
  CUtensorMap tm_self[2];     // {w, S*R rows}, box {64, Rbox}, SWIZZLE_128B (K-major tiles of 64 columns)
  CUtensorMap tm_other;       // {F, S*Ro rows}, box {64, Kp}, SWIZZLE_128B (one MN-major tile, F == 64)
  CUtensorMap tm_out;         // {Nout, S*R rows}, box {64, R}, SWIZZLE_128B: TMA store of the staged output tile
  int nseg, segw[2];
  const uint8_t* Wimg; uint32_t w_bytes;      // conv weight image, NW rows per panel (panel stride NW*16)
  const uint8_t* Mimg; uint32_t m_bytes;      // NodeEdge matrix image, panel stride pstrideM
  const float* bias;
  const int* rp;              // CSR row pointers (R+1), block-diagonal replica for sample groups
  const uint32_t* cj;         // per CSR entry: (byte offset of alpha[i][j] in the alpha operand) << 8 | j
  const float* val;           // GCN filter values per CSR entry (null for GAT)
  const uint8_t* zeros;       // >= al_zero zero bytes in global memory: the alpha region is re-zeroed by a bulk copy
  int nnz;
  bf16* out; long long rows_total;
  int S, R, Ro, Rbox, Kp, KpJ, MT, Kself, F, Nout, NW, act, gat, nb_self, alias_acc, max_len;
  uint32_t pstrideP, pstrideM, pstrideH, pstrideAl, chunk_bytes, self_bytes, xe_bytes, al_bytes, nx_bytes;
  uint32_t off_w, off_m, off_self, self_stride, off_xe, off_nx, off_al, off_sc, off_bar;
  uint32_t stage_half;        // bytes between the two 64-column halves of the staged output tile (in the alpha region)
  uint32_t al_zero;           // bytes of the alpha region to re-zero per sample (operand or staging, whichever is larger)
  int agg_n128;               // 1: one N=128 aggregation MMA per K step (LBO = distance of the two h tiles), 0: two N=64 MMAs
  long long* trace;           // debug: clock64 stamps of CTA 0, SA_NSTAMP per sample (GNNUDS_TRACE_FILE), else null
};
constexpr int SA_NSTAMP = 24;
#define SA_STAMP(i, k) do { if (p.trace && blockIdx.x == 0 && tid == 0 && (i) < 64) p.trace[(i) * SA_NSTAMP + (k)] = clock64(); } while (0)
#define SA_STAMPM(i, k) do { if (p.trace && blockIdx.x == 0 && (i) < 64) p.trace[(i) * SA_NSTAMP + (k)] = clock64(); } while (0)

__device__ __forceinline__ void tma_store_2d(const void* tmap, int c0, int c1, const void* src_smem) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(tmap), "r"(c0), "r"(c1),
               "r"(tc::smem_u32(src_smem))
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// SWIZZLE_128B MN-major B operand spanning TWO 64-column tiles: LBO = byte distance between the tiles
__device__ __forceinline__ uint64_t smem_desc_sw128_mn2(uint32_t addr, uint32_t tile_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)((tile_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)(1024u >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// 16 TMEM lanes x 32 fp32 columns in the mma accumulator fragment layout: thread t holds, for column block n (8 columns),
// v[4n], v[4n+1] = (lane t/4, columns 8n + 2(t%4), +1) and v[4n+2], v[4n+3] = (lane t/4 + 8, same columns)
__device__ __forceinline__ void tmem_ld_16x256b_x4(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.16x256b.x4.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
// four 8x8 b16 matrices, stored TRANSPOSED: lane 8i + r supplies the address of the 16-byte memory row that receives
// column r of matrix i
__device__ __forceinline__ void stmatrix_x4_trans(uint32_t saddr, uint32_t r0, uint32_t r1, uint32_t r2, uint32_t r3) {
  asm volatile("stmatrix.sync.aligned.m8n8.x4.trans.shared.b16 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "r"(r0), "r"(r1), "r"(r2), "r"(r3)
               : "memory");
}
__device__ __forceinline__ void workers_bar_sync() { asm volatile("bar.sync 4, %0;" ::"n"(SA_WORKERS * 32) : "memory"); }

// Edge softmax of one CSR row by LPR lanes (lane `sub` takes entries sub, sub+LPR, ...), at most LPR*KM entries.
// LPR = 4: 8 rows per warp (128 per pass of the 16 workers); LPR = 3: 10 rows per warp (160 per pass, lanes 30/31 idle) so
// that a 131-row side needs ONE pass.  The weight is rounded to fp16 and stored straight into the dense alpha operand;
// the denominator sums the ROUNDED weights, so alpha/den stays a convex combination (divided in the aggregation drain).
template <int KM, int LPR>
__device__ __forceinline__ void sa_softmax_row(int r, bool live, int sub, int lane, const int* rp_s, const uint32_t* cj_s,
                                               const float* ss, const float* sn, uint8_t* al, float* inv) {
  const int beg = live ? rp_s[r] : 0, end = live ? rp_s[r + 1] : 0;
  const float si = live ? ss[r] : 0.f;
  uint32_t ee[KM];
#pragma unroll
  for (int k = 0; k < KM; ++k) { const int qq = beg + sub + LPR * k; ee[k] = qq < end ? cj_s[qq] : 0u; }
  float sc[KM];
#pragma unroll
  for (int k = 0; k < KM; ++k) {
    float v = si + sn[ee[k] & 255u];
    v = v > 0.f ? v : 0.2f * v;
    sc[k] = beg + sub + LPR * k < end ? v : -INFINITY;
  }
  float mx = sc[0];
#pragma unroll
  for (int k = 1; k < KM; ++k) mx = fmaxf(mx, sc[k]);
  // the other lanes of the row group
  const int l1 = LPR == 4 ? (lane ^ 1) : lane - sub + (sub + 1) % 3, l2 = LPR == 4 ? (lane ^ 2) : lane - sub + (sub + 2) % 3;
  if (LPR == 4) {
    mx = fmaxf(mx, __shfl_sync(0xffffffffu, mx, l1));
    mx = fmaxf(mx, __shfl_sync(0xffffffffu, mx, l2));
  } else {
    const float m1 = __shfl_sync(0xffffffffu, mx, l1 & 31), m2 = __shfl_sync(0xffffffffu, mx, l2 & 31);
    mx = fmaxf(mx, fmaxf(m1, m2));
  }
  float den = 0.f;
#pragma unroll
  for (int k = 0; k < KM; ++k) {
    const bool on = beg + sub + LPR * k < end;
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"((sc[k] - mx) * 1.4426950408889634f));
    const __half wh = __float2half_rn(on ? e : 0.f);
    den += __half2float(wh);
    if (on) *reinterpret_cast<__half*>(al + (ee[k] >> 8)) = wh;
  }
  if (LPR == 4) {
    den += __shfl_sync(0xffffffffu, den, l1);
    den += __shfl_sync(0xffffffffu, den, l2);
  } else {
    const float d1 = __shfl_sync(0xffffffffu, den, l1 & 31), d2 = __shfl_sync(0xffffffffu, den, l2 & 31);
    den += d1 + d2;
  }
  if (live && sub == 0) inv[r] = 1.f / den;
}

__global__ void __launch_bounds__(SA_THREADS, 1) sp_agg_kernel(const __grid_constant__ SpAggP p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  // warp index through a shuffle: the compiler then knows it is warp-uniform and keeps the role code on the uniform datapath
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
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
  uint64_t* zbar = bars + 14;        // alpha region zeroed again (bulk copy complete_tx) after it served as the output staging tile
  uint64_t* staged = bars + 15;      // output tile staged in shared memory (workers)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);
  // scratch: bias[Nout] | ss[RP] sn[RP] inv[RP] | rp[R+1] | cj[nnz] | val[nnz] (GCN)
  const int RP = 128 * p.MT;
  float* bias_s = reinterpret_cast<float*>(smem + p.off_sc);
  float* ss = bias_s + p.Nout;
  float* sn = ss + RP;
  float* inv_s = sn + RP;
  int* rp_s = reinterpret_cast<int*>(inv_s + RP);
  uint32_t* cj_s = reinterpret_cast<uint32_t*>(rp_s + p.R + 1);
  float* val_s = reinterpret_cast<float*>(cj_s + p.nnz);

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
    tc::mbar_init(zbar, 1);
    tc::mbar_init(staged, SA_WORKERS * 32);
    tc::mbar_fence_init();
  }
  for (int i = tid; i < p.Nout; i += SA_THREADS) bias_s[i] = p.bias ? p.bias[i] : 0.f;
  for (int i = tid; i < RP; i += SA_THREADS) { inv_s[i] = 1.f; ss[i] = 0.f; sn[i] = 0.f; }
  for (int i = tid; i <= p.R; i += SA_THREADS) rp_s[i] = p.rp[i];
  for (int i = tid; i < p.nnz; i += SA_THREADS) { cj_s[i] = p.cj[i]; if (!p.gat) val_s[i] = p.val[i]; }
  {  // zero the alpha operand and the NodeEdge panels once: everything off the CSR pattern / beyond row R stays zero
    uint4* z = reinterpret_cast<uint4*>(smem + p.off_al);
    for (int i = tid; i < (int)(p.al_zero >> 4); i += SA_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
    z = reinterpret_cast<uint4*>(smem + p.off_nx);
    for (int i = tid; i < (int)(p.nx_bytes >> 4); i += SA_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  tc::pdl_sync();
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
      // the other_dense tile of sample k+2 is free to load (NodeEdge[k+1] committed) long before the self buffer of sample
      // k+1 (aggregation of sample k-1 / k committed): keep that order, or the NodeEdge MMAs wait for a late tile
      issue_xe(0);
      issue_self(0);
      if (n_it > 1) issue_xe(1);
      for (int k = 0; k + 1 < n_it; ++k) { if (k + 2 < n_it) issue_xe(k + 2); issue_self(k + 1); }
    }
  } else if (warp == SA_WORKERS + 2) {
    // ================================ store warp ================================
    // The staged output tile of sample i leaves through two TMA stores; once they have read shared memory, the alpha
    // region the tile occupied is zeroed again by a bulk copy from a zero buffer - all asynchronous to the workers, who
    // only wait for `zbar` before the next softmax writes the pattern's entries.
    if (lane == 0) {
      for (int i = 0; i < n_it; ++i) {
        const int smp = (int)blockIdx.x + i * (int)gridDim.x;
        tc::mbar_wait(staged, i & 1);
        tma_store_2d(&p.tm_out, 0, smp * p.R, smem + p.off_al);
        tma_store_2d(&p.tm_out, 64, smp * p.R, smem + p.off_al + p.stage_half);
        bulk_commit();
        bulk_wait_read();
        tc::mbar_expect_tx(zbar, p.al_zero);
        for (uint32_t o = 0; o < p.al_zero; o += 16384) tc::bulk_g2s(smem + p.off_al + o, p.zeros + o, min(p.al_zero - o, 16384u), zbar);
      }
      bulk_wait_all();
    }
  } else if (warp == SA_WORKERS) {
    // ================================ MMA issuer ================================
    // warp-uniform code: every lane tracks the barriers and the descriptor arithmetic (uniform datapath), one elected
    // lane issues the tcgen05 instructions (see tc::elect_one)
    if (n_it > 0) {
      if (tc::elect_one()) {
        tc::mbar_expect_tx(wbar, p.w_bytes + p.m_bytes);
        for (uint32_t o = 0; o < p.w_bytes; o += 16384) tc::bulk_g2s(smem + p.off_w + o, p.Wimg + o, min(p.w_bytes - o, 16384u), wbar);
        for (uint32_t o = 0; o < p.m_bytes; o += 16384) tc::bulk_g2s(smem + p.off_m + o, p.Mimg + o, min(p.m_bytes - o, 16384u), wbar);
      }
      __syncwarp();
      tc::mbar_wait(wbar, 0);
      const int F = p.F, Nout = p.Nout, NW = p.NW;
      const uint32_t idesc_ne = tc::idesc_bf16(128, F, 0, 1);          // B = other_dense, MN-major
      const uint32_t idesc_cv = tc::idesc_bf16(128, NW, 0, 0);
      const uint32_t idesc_ag = tc::idesc_bf16(128, Nout, 0, 0);       // B = h^T tiles, K-major
      const uint32_t h_tile = (uint32_t)Nout * 128u;                   // one [Nout x 64 neighbours] tile
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
      const uint32_t chunk_bytes = p.chunk_bytes;
      auto issue_ne = [&](int i) {
        tc::mbar_wait(xe_full, i & 1);
        if (i >= 1) tc::mbar_wait(nx_ready, (i - 1) & 1);                 // NodeEdge accumulators of sample i-1 drained
        tc::tc_fence_after();
        if (lane == 0) SA_STAMPM(i, 16);
        // groups of 4 K steps behind ONE election, every descriptor computed from the base (no dependent add chain on the
        // uniform datapath between two MMAs)
        for (int k0 = 0; k0 < ks_kp; k0 += 4) {
          if (tc::elect_one()) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int ks = k0 + j;
              if (ks < ks_kp) {
                const uint64_t ad = m_desc0 + (uint64_t)ks * m_step, bd = xe_desc0 + (uint64_t)ks * 128u;
                tc::umma(acc_ne, ad, bd, idesc_ne, ks != 0);
                if (MT2) tc::umma(acc_ne + (uint32_t)F, ad + 128u, bd, idesc_ne, ks != 0);
              }
            }
          }
          __syncwarp();
        }
        if (tc::elect_one()) { tc::umma_commit(ne_done); tc::umma_commit(xe_empty); }
        __syncwarp();
        if (lane == 0) SA_STAMPM(i, 17);
      };
      auto issue_conv = [&](int i) {
        const int buf = i % NB, use = i / NB;
        const uint32_t in_addr = tc::smem_u32(smem + p.off_self + (size_t)buf * p.self_stride);
        // self columns first (ready early), then the NodeEdge columns
        tc::mbar_wait(&self_full[buf], use & 1);
        if (i >= 1) tc::mbar_wait(h_ready, (i - 1) & 1);                  // conv accumulators of sample i-1 drained
        if (alias && i >= 1) tc::mbar_wait(out_done, (i - 1) & 1);        // ... and the aggregation accumulators they alias
        tc::tc_fence_after();
        if (lane == 0) SA_STAMPM(i, 18);
        for (int c = 0; c < (ks_self >> 2); ++c) {
          const uint64_t ad0 = tc::smem_desc_sw128(in_addr + (uint32_t)c * chunk_bytes), wd0 = w_desc0 + (uint64_t)(4 * c) * w_step;
          if (tc::elect_one()) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint64_t ad = ad0 + (uint64_t)(2 * j), wd = wd0 + (uint64_t)j * w_step;
              tc::umma(acc_conv, ad, wd, idesc_cv, (c | j) != 0);
              if (MT2) tc::umma(acc_conv + (uint32_t)NW, ad + 1024u, wd, idesc_cv, (c | j) != 0);
            }
          }
          __syncwarp();
        }
        tc::mbar_wait(nx_ready, i & 1);
        tc::tc_fence_after();
        for (int k0 = 0; k0 < ks_ne; k0 += 4) {
          if (tc::elect_one()) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int ks = k0 + j;
              if (ks < ks_ne) {
                const uint64_t ad = nx_desc0 + (uint64_t)ks * p_step, wd = w_desc0 + (uint64_t)(ks_self + ks) * w_step;
                tc::umma(acc_conv, ad, wd, idesc_cv, 1u);
                if (MT2) tc::umma(acc_conv + (uint32_t)NW, ad + 128u, wd, idesc_cv, 1u);
              }
            }
          }
          __syncwarp();
        }
        if (tc::elect_one()) tc::umma_commit(conv_done);
        __syncwarp();
        if (lane == 0) SA_STAMPM(i, 13);
      };
      // pipe order: conv[i+1] goes in BEFORE agg[i] when the accumulators are separate (MT == 1): it runs while the workers
      // do the softmax of sample i.  With aliased accumulators conv[i+1] has to wait for the aggregation
      // drain of sample i.
      // Separate accumulators (MT == 1): NodeEdge runs TWO samples ahead - per sample the pipe sees NE[i+2], conv[i+1],
      // agg[i]; the workers drain NE[i+2] while they would otherwise wait for agg[i], and conv[i+1] has its NodeEdge panels
      // from the previous sample's wait window, so it starts as soon as the conv accumulators of sample i are drained.
      if (!alias) { issue_ne(0); if (n_it > 1) issue_ne(1); issue_conv(0); }
      else { issue_ne(0); issue_conv(0); if (n_it > 1) issue_ne(1); }
      for (int i = 0; i < n_it; ++i) {
        const int buf = i % NB;
        const uint32_t in_addr = tc::smem_u32(smem + p.off_self + (size_t)buf * p.self_stride);
        if (!alias) { if (i + 2 < n_it) issue_ne(i + 2); if (i + 1 < n_it) issue_conv(i + 1); }
        // ---- agg[i] ----
        tc::mbar_wait(alpha_ready, i & 1);
        if (!alias && i >= 1) tc::mbar_wait(out_done, (i - 1) & 1);       // aggregation accumulators of sample i-1 drained
        tc::tc_fence_after();
        if (lane == 0) SA_STAMPM(i, 14);
        {
          // A = alpha (K-major panels), B = h^T (K-major SWIZZLE_128B tiles [Nout features x 64 neighbours], written
          // transposed by the conv drain): 4 K steps of 32 bytes per 64-neighbour tile
          for (int k0 = 0; k0 < ks_j; k0 += 4) {
            const uint64_t bd0 = tc::smem_desc_sw128(in_addr + (uint32_t)(k0 >> 2) * h_tile);
            if (tc::elect_one()) {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const int ks = k0 + j;
                if (ks < ks_j) {
                  const uint64_t ad = al_desc0 + (uint64_t)ks * al_step, bd = bd0 + (uint64_t)(2 * j);
                  tc::umma(acc_agg, ad, bd, idesc_ag, ks != 0);
                  if (MT2) tc::umma(acc_agg + (uint32_t)Nout, ad + 128u, bd, idesc_ag, ks != 0);
                }
              }
            }
            __syncwarp();
          }
        }
        if (tc::elect_one()) { tc::umma_commit(agg_done); tc::umma_commit(&self_empty[buf]); }
        __syncwarp();
        if (lane == 0) SA_STAMPM(i, 15);
        // NodeEdge[i+2] goes in behind agg[i] (its MN-major B operand makes it slow, ~100 cycles per K step, and the
        // aggregation drain is waiting for agg[i]); its result is not needed before the conv drain of sample i+1 is over
        if (alias) { if (i + 1 < n_it) issue_conv(i + 1); if (i + 2 < n_it) issue_ne(i + 2); }
      }

    }
  } else {
    // ================================ workers ================================
    const int q = warp & 3, cs = warp >> 2;            // TMEM lanes 32q .. 32q+31, column slice cs of 4
    const int R = p.R, MT = p.MT, NW = p.NW, Nout = p.Nout, F = p.F, KpJ = p.KpJ, gat = p.gat, act = p.act;
    const uint32_t pstrideP = p.pstrideP, h_tile = (uint32_t)p.Nout * 128u;
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
    const bool alias_w = p.alias_acc != 0;
    if (n_it > 0) drain_ne(0);
    if (!alias_w && n_it > 1) { tc::mbar_wait(conv_done, 0); drain_ne(1); }   // panels[1] once conv[0] has read panels[0]
    for (int i = 0; i < n_it; ++i) {
      const int buf = i % NB;
      const int smp = (int)blockIdx.x + i * (int)gridDim.x;
      uint8_t* hbase = smem + p.off_self + (size_t)buf * p.self_stride;
      // ---- A: conv accumulators -> h tile (fp16, MN-major B operand of the aggregation MMA) + scores ----
      tc::mbar_wait(conv_done, i & 1);
      tc::tc_fence_after();
      SA_STAMP(i, 0);
      for (int mt = 0; mt < MT; ++mt) {
        if (mt * 128 + q * 32 >= KpJ) continue;        // warp-uniform
        const uint32_t ta = acc_conv + ((uint32_t)(q * 32) << 16) + (uint32_t)(mt * NW);
        float v[32];
        // accumulator-fragment loads: v[0..15] = TMEM lanes 32q..32q+15, v[16..31] = lanes 32q+16..32q+31, columns 32cs..32cs+31
        tmem_ld_16x256b_x4(ta + (uint32_t)(cs * 32), v);
        tmem_ld_16x256b_x4(ta + (16u << 16) + (uint32_t)(cs * 32), v + 16);
        float sv[8];
        if (gat && cs == 0) tc::tmem_ld8(ta + (uint32_t)Nout, sv);
        tc::tmem_ld_wait();
        if (mt == 0) SA_STAMP(i, 6);
        // h^T as K-major SWIZZLE_128B tiles [Nout features x 64 neighbours]: an 8 (neighbours) x 8 (features) block of the
        // fragment goes out through stmatrix.trans as eight 16-byte chunks, one per feature row.  Neighbour rows >= R
        // meet zero alpha columns and must stay finite: zeroed.
        const int jrow = mt * 128 + q * 32 + (lane >> 2);          // fragment rows jrow (+8) (+16) (+24)
        const int mi = lane >> 3, mr = lane & 7;                    // stmatrix: this lane addresses row mr of matrix mi
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          const bool ok0 = jrow + 16 * hh < R, ok1 = jrow + 16 * hh + 8 < R;
          const int jb = mt * 128 + q * 32 + 16 * hh + 8 * (mi & 1);   // 8-aligned neighbour base of this lane's matrix
#pragma unroll
          for (int nn = 0; nn < 4; nn += 2) {
            const float* w = v + 16 * hh + 4 * nn;
            const uint32_t r0 = ok0 ? tc::pack_f16(w[0], w[1]) : 0u, r1 = ok1 ? tc::pack_f16(w[2], w[3]) : 0u;
            const uint32_t r2 = ok0 ? tc::pack_f16(w[4], w[5]) : 0u, r3 = ok1 ? tc::pack_f16(w[6], w[7]) : 0u;
            const int f = cs * 32 + 8 * (nn + (mi >> 1)) + mr;
            const uint32_t sa = tc::smem_u32(hbase) + (uint32_t)(jb >> 6) * h_tile + (uint32_t)f * 128u +
                                (uint32_t)((((jb & 63) >> 3) ^ (f & 7)) << 4);
            stmatrix_x4_trans(sa, r0, r1, r2, r3);
          }
        }
        if (gat && cs == 0) { const int row = mt * 128 + q * 32 + lane; if (row < R) { ss[row] = sv[0] + sv[1]; sn[row] = sv[2] + sv[3]; } }
      }
      SA_STAMP(i, 7);
      tc::fence_async_smem();                          // h tile -> tensor-core reads
      tc::tc_fence_before();
      mbar_arrive(h_ready);
      SA_STAMP(i, 1);
      // ---- C: NodeEdge accumulators of the next sample -> fp16 panels (back to back with the conv drain: the softmax
      //         then runs with no TMEM reads in flight, which slow concurrent MMAs ~2x) ----
      if (alias_w && i + 1 < n_it) drain_ne(i + 1);
      if (alias_w) SA_STAMP(i, 3);
      // ---- B: edge softmax over the CSR pattern -> alpha entries ----
      // the alpha region doubled as the output staging tile of the previous sample; the MMA warp stored it and had it
      // zeroed again (handle_store): rewrite the pattern's entries
      if (i >= 1) tc::mbar_wait(zbar, (i - 1) & 1);    // the bulk copy that zeroed the region after sample i-1's store
      SA_STAMP(i, 8);
      workers_bar_sync();                              // scores of every row visible
      SA_STAMP(i, 9);
      if (!gat) {
        for (int k = tid; k < p.nnz; k += SA_WORKERS * 32) *reinterpret_cast<__half*>(al + (cj_s[k] >> 8)) = __float2half_rn(val_s[k]);
        tc::fence_async_smem();
      } else {
        if (R <= SA_WORKERS * 8) {                     // 4 lanes per row, one pass
          const int r = tid >> 2;
          if (km <= 3) sa_softmax_row<3, 4>(r, r < R, tid & 3, lane, rp_s, cj_s, ss, sn, al, inv_s);
          else if (km == 4) sa_softmax_row<4, 4>(r, r < R, tid & 3, lane, rp_s, cj_s, ss, sn, al, inv_s);
          else sa_softmax_row<5, 4>(r, r < R, tid & 3, lane, rp_s, cj_s, ss, sn, al, inv_s);
        } else {                                       // 3 lanes per row, 10 rows per warp: up to 160 rows in one pass
          for (int r0 = 0; r0 < R; r0 += SA_WORKERS * 10) {
            if (r0 + warp * 10 >= R) continue;         // warp-uniform
            const int r = r0 + warp * 10 + lane / 3, sub = lane % 3;
            const bool live = r < R && lane < 30;
            if (p.max_len <= 12) sa_softmax_row<4, 3>(r, live, sub, lane, rp_s, cj_s, ss, sn, al, inv_s);
            else if (p.max_len <= 18) sa_softmax_row<6, 3>(r, live, sub, lane, rp_s, cj_s, ss, sn, al, inv_s);
            else sa_softmax_row<7, 3>(r, live, sub, lane, rp_s, cj_s, ss, sn, al, inv_s);
          }
        }
        tc::fence_async_smem();                        // alpha -> tensor-core reads
      }
      mbar_arrive(alpha_ready);
      SA_STAMP(i, 2);
      // ---- C' (separate accumulators): NodeEdge accumulators of sample i+2 -> fp16 panels while agg[i] runs; conv[i+1]
      //         must have read the panels of sample i+1 ----
      if (!alias_w) {
        if (i + 2 < n_it) { tc::mbar_wait(conv_done, (i + 1) & 1); drain_ne(i + 2); }
        SA_STAMP(i, 3);
      }
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
        if (mt == 0) SA_STAMP(i, 10);
        if (row < R) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            v[4 * j] = fmaf(v[4 * j], sc, b4[j].x); v[4 * j + 1] = fmaf(v[4 * j + 1], sc, b4[j].y);
            v[4 * j + 2] = fmaf(v[4 * j + 2], sc, b4[j].z); v[4 * j + 3] = fmaf(v[4 * j + 3], sc, b4[j].w);
          }
          act_tile<32>(v, act);
          // staged as two SWIZZLE_128B tiles [R x 64 features] in the (now dead) alpha region; one TMA store per tile
          // writes full 128-byte lines (per-lane row stores touch 32 lines per instruction)
          uint8_t* srow = al + (size_t)(cs >> 1) * p.stage_half + (size_t)row * 128;
          const int cb = (cs & 1) * 4, sw = row & 7;
#pragma unroll
          for (int j = 0; j < 4; ++j) *reinterpret_cast<uint4*>(srow + (((cb + j) ^ sw) << 4)) = tc::pack8_bf16(v + 8 * j);
        }
      }
      SA_STAMP(i, 11);
      tc::tc_fence_before();
      mbar_arrive(out_done);
      tc::fence_async_smem();                          // staged tile -> TMA store reads
      mbar_arrive(staged);
      SA_STAMP(i, 12);
      SA_STAMP(i, 5);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_free(tmem, 512);
}

}  // namespace gu

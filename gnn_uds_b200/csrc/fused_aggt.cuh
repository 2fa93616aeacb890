// Fused spatial-layer kernel for stream sides with 128 < R <= 192 rows (the 131-link side of shunqing), TRANSPOSED:
//
//   h^T   [E features x R rows]  =  Wconv^T [E x K]  @  [ self | Mat @ other_dense ]^T [K x R]
//   out^T [E features x R rows]  =  h^T [E x R]      @  alpha^T [R x R]                              (emulator.py:227-230)
//
// sp_agg_kernel (fused_agg.cuh) keeps the rows of a sample on the M axis of the MMAs: 131 rows are two 128-row M tiles,
// the second with 3 real rows - twice the MMA time, a second pass of every TMEM drain on one warp quadrant only, and
// conv / aggregation accumulators that no longer fit TMEM side by side.  With the FEATURES on M (E = 128 = exactly one
// tile) and the rows on N (any multiple of 16: 144), conv is 12 MMAs instead of 24 and the aggregation 9 instead of 18,
// every drain is one balanced pass over all 128 TMEM lanes, and conv (144 columns), aggregation (144) and NodeEdge (128)
// accumulators are separate.  The operands are the ones sp_agg_kernel already keeps in shared memory, used the other way
// round: the K-major image of Wconv^T is the A operand, the self tiles (SWIZZLE_128B, rows = N index) and the NodeEdge
// panels are B operands; the conv drain reads lane = feature and writes h^T row-wise as K-major SWIZZLE_128B tiles (the A
// operand of the aggregation) with plain 16-byte stores; alpha[i][j] (j contiguous) is the K-major B operand; the
// aggregation drain (lane = feature, column = row i) goes through tcgen05.ld.16x256b + stmatrix.trans into the row-major
// staging tile of the TMA store.  NodeEdge stays as it is (rows on M, two tiles): its result is a B operand here.
// GAT scores h.a_self / h.a_neigh = [self | NE] . (W a): sp_agg_kernel reads them as extra N columns of the conv GEMM; here
// they would be extra M rows (a second M tile).  Instead a second, tiny A operand - 8 rows [hi / lo of W a_self, hi / lo of
// W a_neigh, 0 0 0 0], stride-byte-offset 0 so that all sixteen 8-row groups of the M tile alias it - multiplies the same B
// operands into the aggregation accumulators (free until agg[i] is issued; TMEM has no 144 columns left): lanes 8g+0..3 of
// every lane group hold the scores, each worker warp reads some columns from its own quadrant.  Any SIMT version of this product (plain FMAs, mma.sync + ldmatrix) took 2.5-4 k cycles next to the running MMAs:
// shared-memory loads issue every ~250 cycles while the tensor pipe streams operands.
//
// Shared memory is what bounds the pipeline (W^T image 48 KB + NodeEdge image 34 KB + self tiles 34 KB + h^T 36..48 KB + x_e
// 16 KB + NodeEdge panels 17 KB + alpha / staging 38 KB > 227 KB): with h^T written over the self tiles the load of the next
// sample could only start when agg[i] had completed, and TMA latency + conv[i+1] sat on the critical path behind the
// aggregation drain.  So the persistent A operand of conv^T lives in TENSOR MEMORY (96 columns, written once per CTA with
// tcgen05.st; tcgen05.mma takes A from TMEM - which also removes the 4 KB/MMA shared-memory read of A), h^T gets its own
// region (two SWIZZLE_128B tiles + plain panels for the rows 128..143), and so does the staging tile of the TMA store: the
// alpha operand is zeroed once per CTA, every sample's softmax rewrites the same CSR positions.
// TMEM: conv 144 | aggregation + scores 144 | W^T 96 | NodeEdge 128 = 512 columns.
//
// Roles: 16 workers (A conv drain -> h^T, score read, B softmax, C NodeEdge drain of the next sample while agg[i] runs,
// D aggregation drain), MMA issuer (agg[i], conv[i+1], scores[i+1], NE[i+2]), TMA producer, store warp.
#pragma once
#include <cuda.h>

#include "fused_agg.cuh"

namespace gu {

__device__ __forceinline__ void tmem_ld_16x256b_x2(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
}

// extra parameters of the transposed kernel (the rest is SpAggP)
struct SpAggTP {
  SpAggP b;
  const uint8_t* wa;          // GAT score operand: 2 images x K/8 panels x [8 rows x 8 K] 16-bit (set_weights), K = Kself + F
  int Nr;                     // rows of a sample rounded up to 16: N of the transposed MMAs
  uint32_t off_wa, off_dump;  // smem: the score operand images; 256 bytes that swallow stmatrix rows beyond R
  uint32_t off_h;             // smem: h^T, two SWIZZLE_128B tiles [Nout x 64 rows] + (Nr - 128) / 8 plain panels [Nout x 8 rows]
  uint32_t off_stage;         // smem: SWIZZLE_128B staging tile of the TMA store (two 64-column halves, stage_half apart)
};

__global__ void __launch_bounds__(SA_THREADS, 1) sp_aggt_kernel(const __grid_constant__ SpAggTP pp) {
  const SpAggP& p = pp.b;
  extern __shared__ __align__(1024) uint8_t smem[];
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.off_bar);
  uint64_t* self_full = bars;        // TMA complete_tx, self tiles (ONE buffer)
  uint64_t* self_empty = bars + 2;   // aggregation MMAs that read h^T completed (tcgen05.commit)
  uint64_t* xe_full = bars + 4;
  uint64_t* xe_empty = bars + 5;
  uint64_t* ne_done = bars + 6;
  uint64_t* nx_ready = bars + 7;
  uint64_t* conv_done = bars + 8;
  uint64_t* h_ready = bars + 9;
  uint64_t* alpha_ready = bars + 10;
  uint64_t* agg_done = bars + 11;
  uint64_t* agg_read = bars + 12;    // aggregation accumulators of this sample are in registers (workers)
  uint64_t* wbar = bars + 13;
  uint64_t* st_free = bars + 14;     // TMA store of this sample has read the staging tile (store warp)
  uint64_t* staged = bars + 15;
  uint64_t* sc_done = bars + 16;     // score MMAs of this sample completed (tcgen05.commit)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 17);
  const int RP = 256;
  float* bias_s = reinterpret_cast<float*>(smem + p.off_sc);
  float* ss = bias_s + p.Nout;       // [h.a_self 256 | h.a_neigh 256]
  float* inv_s = ss + 2 * RP;
  int* rp_s = reinterpret_cast<int*>(inv_s + RP);
  uint32_t* cj_s = reinterpret_cast<uint32_t*>(rp_s + p.R + 1);
  float* val_s = reinterpret_cast<float*>(cj_s + p.nnz);
  const int K = p.Kself + p.F;
  const uint32_t wa_bytes = (uint32_t)(K >> 3) * 128u;

  if (warp == 0) tc::tmem_alloc(tmem_slot, 512);
  if (tid == 32) {
    for (int i = 0; i < 2; ++i) { tc::mbar_init(&self_full[i], 1); tc::mbar_init(&self_empty[i], 1); }
    tc::mbar_init(xe_full, 1); tc::mbar_init(xe_empty, 1); tc::mbar_init(ne_done, 1);
    // worker barriers: one arrival per warp (mbar_arrive_warp)
    tc::mbar_init(nx_ready, SA_WORKERS);
    tc::mbar_init(conv_done, 1);
    tc::mbar_init(h_ready, SA_WORKERS);
    tc::mbar_init(alpha_ready, SA_WORKERS);
    tc::mbar_init(agg_done, 1);
    tc::mbar_init(agg_read, SA_WORKERS);
    tc::mbar_init(wbar, 1); tc::mbar_init(st_free, 1);
    tc::mbar_init(staged, SA_WORKERS);
    tc::mbar_init(sc_done, 1);
    tc::mbar_fence_init();
  }
  for (int i = tid; i < p.Nout; i += SA_THREADS) bias_s[i] = p.bias ? p.bias[i] : 0.f;
  for (int i = tid; i < RP; i += SA_THREADS) inv_s[i] = 1.f;
  for (int i = tid; i < 2 * RP; i += SA_THREADS) ss[i] = 0.f;
  for (int i = tid; i <= p.R; i += SA_THREADS) rp_s[i] = p.rp[i];
  for (int i = tid; i < p.nnz; i += SA_THREADS) { cj_s[i] = p.cj[i]; if (!p.gat) val_s[i] = p.val[i]; }
  if (p.gat) for (int i = tid; i < (int)(wa_bytes >> 4); i += SA_THREADS) reinterpret_cast<uint4*>(smem + pp.off_wa)[i] = reinterpret_cast<const uint4*>(pp.wa)[i];
  {
    uint4* z = reinterpret_cast<uint4*>(smem + p.off_al);
    for (int i = tid; i < (int)(p.al_zero >> 4); i += SA_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
    z = reinterpret_cast<uint4*>(smem + p.off_nx);
    for (int i = tid; i < (int)(p.nx_bytes >> 4); i += SA_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t acc_conv = tmem;               // Nr columns (<= 144)
  const uint32_t acc_agg = tmem + 144u;         // Nr columns
  const uint32_t acc_sc = acc_agg;              // GAT scores of the next sample, lanes 8g+0..3, between the aggregations
  const uint32_t a_w = tmem + 288u;             // W^T as A operand: lane = feature, K/2 columns (<= 96)
  const uint32_t acc_ne = tmem + 384u;          // 2 x F columns
  if (warp < SA_WORKERS) {
    // W^T image (global, K-major panels [K/8][Nout][8]) -> tensor memory: 16 bytes of a panel row = 4 columns of lane f
    const int q = warp & 3, cs = warp >> 2, f = q * 32 + lane;
    for (int pn = cs; pn < (K >> 3); pn += 4) {
      const uint4 w = f < p.Nout ? *reinterpret_cast<const uint4*>(p.Wimg + ((size_t)pn * p.Nout + f) * 16) : make_uint4(0u, 0u, 0u, 0u);
      tc::tmem_st4(a_w + ((uint32_t)(q * 32) << 16) + (uint32_t)(4 * pn), w);
    }
    tc::tmem_st_wait();
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  tc::pdl_sync();
  const int n_it = p.S > (int)blockIdx.x ? (p.S - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  const int Nr = pp.Nr;
  const uint32_t h_tile = (uint32_t)p.Nout * 128u;            // one h^T tile [Nout features x 64 rows]

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
        const int smp = (int)blockIdx.x + i * (int)gridDim.x;
        if (i >= 1) tc::mbar_wait(&self_empty[0], (i - 1) & 1);
        tc::mbar_expect_tx(&self_full[0], p.self_bytes);
        uint32_t o = 0;
        for (int s = 0; s < p.nseg; ++s)
          for (int k0 = 0; k0 < p.segw[s]; k0 += 64, o += p.chunk_bytes) tc::tma_load_2d(smem + p.off_self + o, &p.tm_self[s], k0, smp * p.R, &self_full[0]);
        // ONE self buffer: the load of sample i+1 starts when conv[i] and scores[i] have read it - bring its lines into L2 a
        // whole sample early, so that it is an L2 hit
        if (i + 1 < n_it)
          for (int s = 0; s < p.nseg; ++s)
            for (int k0 = 0; k0 < p.segw[s]; k0 += 64) tc::tma_prefetch_2d(&p.tm_self[s], k0, (smp + (int)gridDim.x) * p.R);
      };
      issue_xe(0);
      issue_self(0);
      if (n_it > 1) issue_xe(1);
      for (int k = 0; k + 1 < n_it; ++k) { issue_self(k + 1); if (k + 2 < n_it) issue_xe(k + 2); }   // NE[k+2] is issued late in sample k: its x_e slot must not hold up the self tiles
    }
  } else if (warp == SA_WORKERS + 2) {
    // ================================ store warp ================================
    if (lane == 0) {
      for (int i = 0; i < n_it; ++i) {
        const int smp = (int)blockIdx.x + i * (int)gridDim.x;
        tc::mbar_wait(staged, i & 1);
        tma_store_2d(&p.tm_out, 0, smp * p.R, smem + pp.off_stage);
        tma_store_2d(&p.tm_out, 64, smp * p.R, smem + pp.off_stage + p.stage_half);
        bulk_commit();
        bulk_wait_read();
        mbar_arrive(st_free);
      }
      bulk_wait_all();
    }
  } else if (warp == SA_WORKERS) {
#include "fused_aggt_mma.inc"
  } else {
#include "fused_aggt_workers.inc"
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_free(tmem, 512);
}

}  // namespace gu

// Tensor-core rows kernel (sm_100a): every dense contraction of the Emulator forward on
// tcgen05 with bf16 operands and fp32 accumulation in TMEM.
//
//   out[r,:] = epilogue( sum_seg A_seg[src(r,seg), :] @ W_seg + bias )
//
// Same logical-row addressing as rows_gemm_f32 (kernels.cuh): rows are (b,t,n); a segment
// is a [rows][C] bf16 activation buffer read with a time shift (zero for t<0), so one kernel
// covers Dense, the GCN/GAT linear part over the concatenated [x | NodeEdge(x_e)] input,
// the causal dilated Conv1D (k shifted segments) and the GRU/LSTM input projections.
//
// Persistent, warp specialised, one CTA per SM:
//   warps 0..7   epilogue   TMEM -> registers -> bias/act -> bf16 -> HBM, or (agg != 0) the
//                           spektral GCN/GAT neighbour aggregation on the tile's h = x@W kept
//                           in shared memory (h never goes to HBM)
//   warps 8..11  producers  cp.async gathers of 128*MT rows x kc columns into the panel layout
//                           of umma.cuh, NS-stage ring, completion signalled by cp.async.mbarrier.arrive
//   warp  12     MMA issuer one thread: tcgen05.mma (M=128, N=Nout, K=16), weights resident in
//                           shared memory as a K-major bf16 image fetched once by a bulk copy;
//                           accumulators double-buffered in TMEM so tile i's epilogue overlaps
//                           tile i+1's MMAs and loads.
// A tile is rpt logical rows in MT M-tiles: 128 arbitrary rows (plain), or g whole (b,t)
// samples of R rows when the epilogue aggregates over the graph.
#pragma once
#include <cuda.h>

#include "kernels.cuh"
#include "umma.cuh"

namespace gu {

constexpr int TC_EPI_WARPS = 8;
constexpr int TC_PROD_WARPS = 4;
constexpr int TC_THREADS = (TC_EPI_WARPS + TC_PROD_WARPS + 1) * 32;   // 416

struct TcSeg { const bf16* p; int ld, width, shift, pitch, off; };

struct TcGemmP {
  // one tensor map per segment whose rows are contiguous in the tile order (no time shift, no
  // pitch): {8 elements, rows, C/8 panels} with box {8, boxrows, kc/8}, so a single TMA instruction
  // lands a whole pipeline stage in the panel layout.  Other segments use the cp.async producers.
  CUtensorMap tmap[4];
  int use_tma[4];
  int boxrows;               // rows per panel in a stage (multiple of 8, >= rows of a tile, <= 256)
  TcSeg seg[4]; int nseg;
  const uint8_t* Wimg;       // bf16 K-major image of W^T: element (n,k) at (k/8)*Nout*16 + n*16 + (k%8)*2
  const float* bias;         // (Nout) or null
  bf16* out; int ldo;
  int Nout, Ktot, kc, act;
  unsigned rows; int Nn, Td, Ts, pitch_d, off_d;
  int rpt, MT; unsigned n_tiles; int NS;
  // scenario-aligned tiling (tpb > 0; time-shifted segments, i.e. the causal Conv1D): a tile is 128 rows of ONE
  // scenario b (tpb tiles per scenario, the last one partial), every segment is a rank-4 tensor map
  // {8, Ts*Nn rows of a scenario, Bc, C/8}, and tap `shift` is the same box moved up by shift*Nn rows - rows before
  // t = 0 fall outside the scenario's extent and arrive as zeros (the causal padding), no gather needed.
  int tpb; unsigned rows_per_b;
  // sw128 (all segments by TMA, kc == 64): a stage is one SWIZZLE_128B K-major tile [boxrows x 64] written by a TMA box
  // with a 128-byte inner extent (8x fewer TMA requests than the 16-byte panel boxes); maps are {C, rows} or {C, rows, Bc}
  int sw128;
  uint32_t w_bytes, stage_bytes, pstrideA, stage_off, h_off, sc_off, bar_off;
  // aggregation epilogue (agg: 0 none, 1 GCN, 2 GAT) over samples of R rows
  int agg, R, hsb, nnz;      // hsb = row stride in bytes of the bf16 h tile ((Nout + 8) * 2)
  const int* row_ptr; const int* col_idx; const float* val;
  const float* a_self; const float* a_neigh;
  // narrow-head epilogue (agg == 0): out_head[row*head_ldo + head_col] = head_act( act(acc + bias) . head_w + head_b );
  // the Nout-wide row itself is not written (K8 flood head, emulator.py:326-333)
  const float* head_w; const float* head_b; float* head_out; int head_ldo, head_col, head_act;
};

__device__ __forceinline__ void cp_async16(void* dst_smem, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(tc::smem_u32(dst_smem)), "l"(src), "r"(src_bytes)
               : "memory");
}
__device__ __forceinline__ void cp_async16_full(void* dst_smem, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(tc::smem_u32(dst_smem)), "l"(src) : "memory");
}
// one arrival on `bar` when all cp.async issued so far by this thread have landed (no wait, no extra count)
__device__ __forceinline__ void cp_async_arrive(uint64_t* bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(tc::smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tc::smem_u32(bar)) : "memory");
}
// One arrival per WARP instead of 32 arrivals on the same barrier word.  The lanes' own fences (tcgen05 / async proxy)
// come before; __syncwarp orders their writes before lane 0's release.
__device__ __forceinline__ void mbar_arrive_warp(uint64_t* bar, int lane) {
  __syncwarp();
  if (lane == 0) mbar_arrive(bar);
}
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(TC_EPI_WARPS * 32) : "memory"); }

// activation over a register tile; the branch is warp-uniform and sits outside the element loop
template <int N>
__device__ __forceinline__ void act_tile(float* v, int act) {
  if (act == ACT_RELU) {
#pragma unroll
    for (int i = 0; i < N; ++i) v[i] = fmaxf(v[i], 0.f);
  } else if (act != ACT_LINEAR) {
#pragma unroll
    for (int i = 0; i < N; ++i) v[i] = apply_act(v[i], act);
  }
}

// NC (multiple of 16) consecutive fp32 accumulator columns of this thread's TMEM lane; all loads in flight, one wait
template <int NC>
__device__ __forceinline__ void tmem_load_cols(uint32_t taddr, float* v) {
#pragma unroll
  for (int i = 0; i < NC; i += 16) tc::tmem_ld16(taddr + (uint32_t)i, v + i);
  tc::tmem_ld_wait();
}

// plain epilogue of NC columns: bias + activation -> bf16 -> HBM
template <int NC>
__device__ __forceinline__ void epi_store_cols(uint32_t taddr, int c0, bool valid, bf16* orow, const float* bias_s, int act) {
  float v[NC];
  tmem_load_cols<NC>(taddr + (uint32_t)c0, v);
  if (valid) {
#pragma unroll
    for (int i = 0; i < NC; i += 4) {
      const float4 bb = *reinterpret_cast<const float4*>(bias_s + c0 + i);
      v[i] += bb.x; v[i + 1] += bb.y; v[i + 2] += bb.z; v[i + 3] += bb.w;
    }
    act_tile<NC>(v, act);
    // 32-byte stores (st.global.v8.b32): a lane's row slice goes out in whole sectors
#pragma unroll
    for (int i = 0; i < NC; i += 16)
      asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(orow + c0 + i), "r"(pack2_bf16(v[i], v[i + 1])),
                   "r"(pack2_bf16(v[i + 2], v[i + 3])), "r"(pack2_bf16(v[i + 4], v[i + 5])), "r"(pack2_bf16(v[i + 6], v[i + 7])),
                   "r"(pack2_bf16(v[i + 8], v[i + 9])), "r"(pack2_bf16(v[i + 10], v[i + 11])), "r"(pack2_bf16(v[i + 12], v[i + 13])),
                   "r"(pack2_bf16(v[i + 14], v[i + 15]))
                   : "memory");
  }
}

// narrow-head epilogue of NC columns: bias + activation, then the partial dot product with the head weights
template <int NC>
__device__ __forceinline__ float epi_head_cols(uint32_t taddr, int c0, const float* bias_s, const float* hw_s, int act) {
  float v[NC];
  tmem_load_cols<NC>(taddr + (uint32_t)c0, v);
#pragma unroll
  for (int i = 0; i < NC; i += 4) {
    const float4 bb = *reinterpret_cast<const float4*>(bias_s + c0 + i);
    v[i] += bb.x; v[i + 1] += bb.y; v[i + 2] += bb.z; v[i + 3] += bb.w;
  }
  act_tile<NC>(v, act);
  float d0 = 0.f, d1 = 0.f;
#pragma unroll
  for (int i = 0; i < NC; i += 4) {
    const float4 w = *reinterpret_cast<const float4*>(hw_s + c0 + i);
    d0 = fmaf(v[i], w.x, d0); d1 = fmaf(v[i + 1], w.y, d1); d0 = fmaf(v[i + 2], w.z, d0); d1 = fmaf(v[i + 3], w.w, d1);
  }
  return d0 + d1;
}

__device__ __forceinline__ void tile_rows(const TcGemmP& p, unsigned tile, unsigned& row_base, int& nvalid, unsigned& b, unsigned& rb) {
  if (p.tpb) {
    b = tile / (unsigned)p.tpb;
    rb = (tile - b * (unsigned)p.tpb) * 128u;
    row_base = b * p.rows_per_b + rb;
    nvalid = (int)min(128u, p.rows_per_b - rb);
  } else {
    b = 0; rb = 0;
    row_base = tile * (unsigned)p.rpt;
    nvalid = (int)min((unsigned)p.rpt, p.rows - row_base);
  }
}

__global__ void __launch_bounds__(TC_THREADS, 1) tc_rows_kernel(const __grid_constant__ TcGemmP p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  // warp index through a shuffle: known warp-uniform to the compiler, so the role code stays on the uniform datapath
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  uint8_t* Wsm = smem;
  uint8_t* stages = smem + p.stage_off;
  float* hbuf = reinterpret_cast<float*>(smem + p.h_off);
  float* scbuf = reinterpret_cast<float*>(smem + p.sc_off);     // GAT: ss[2][RP], sn[2][RP], a_self[Nout], a_neigh[Nout]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.bar_off);
  const int NS = p.NS;
  uint64_t* full = bars;
  uint64_t* empty = bars + NS;
  uint64_t* tfull = bars + 2 * NS;
  uint64_t* tempty = bars + 2 * NS + 2;
  uint64_t* wbar = bars + 2 * NS + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * NS + 5);

  if (warp == 0) tc::tmem_alloc(tmem_slot, 512);
  if (tid == 32) {
    for (int i = 0; i < NS; ++i) { tc::mbar_init(&full[i], TC_PROD_WARPS * 32); tc::mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { tc::mbar_init(&tfull[i], 1); tc::mbar_init(&tempty[i], TC_EPI_WARPS * 32); }
    tc::mbar_init(wbar, 1);
    tc::mbar_fence_init();
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  tc::pdl_sync();
  const uint32_t tmem = *tmem_slot;
  const int nchunk = p.Ktot / p.kc;           // K chunks (= pipeline stages) per tile

  if (warp >= TC_EPI_WARPS && warp < TC_EPI_WARPS + TC_PROD_WARPS) {
    // ================================ producers ================================
    // one producer thread owns row (ptid + 128*m) of every M-tile: per stage it computes one source
    // address per row and issues kc/8 cp.async of 16 bytes (a warp instruction writes 512 contiguous
    // bytes of one panel); the row's (b,t,n) is worked out once per tile
    const int ptid = tid - TC_EPI_WARPS * 32;
    const int cpr = p.kc >> 3;                 // 16-byte chunks per row per stage (4 or 8)
    if (ptid == 0)
      for (int s = 0; s < p.nseg; ++s) if (p.use_tma[s]) tc::tma_prefetch_desc(&p.tmap[s]);
    int stage = 0;
    uint32_t ephase = 1;                       // parity to wait for on empty[]: passes at once on the first lap
    uint32_t tq[2], tn[2];                     // per owned row: b*Ts + t + (Ts-Td) (~0 = no such row), (t << 20) | n
    for (unsigned tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      unsigned row_base, tb_b, tb_rb; int nv_unused;
      tile_rows(p, tile, row_base, nv_unused, tb_b, tb_rb);
      if (p.tpb) {
        // every segment is a rank-4 TMA box; one elected thread issues, the others only arrive
        for (int s = 0; s < p.nseg; ++s) {
          const int row0 = (int)tb_rb + (p.Ts - p.Td - p.seg[s].shift) * p.Nn;
          for (int k0 = 0; k0 < p.seg[s].width; k0 += p.kc) {
            tc::mbar_wait(&empty[stage], ephase);
            if (ptid == 0) {
              tc::mbar_expect_tx(&full[stage], (uint32_t)p.boxrows * 16u * (uint32_t)cpr);
              if (p.sw128) tc::tma_load_3d(stages + (size_t)stage * p.stage_bytes, &p.tmap[s], k0, row0, (int)tb_b, &full[stage]);
              else tc::tma_load_4d(stages + (size_t)stage * p.stage_bytes, &p.tmap[s], 0, row0, (int)tb_b, k0 >> 3, &full[stage]);
            } else {
              mbar_arrive(&full[stage]);
            }
            if (++stage == NS) { stage = 0; ephase ^= 1; }
          }
        }
        continue;
      }
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        const int rloc = ptid + 128 * m;
        const unsigned rl = row_base + rloc;
        tq[m] = 0xFFFFFFFFu; tn[m] = 0;
        if (m < p.MT && rloc < p.rpt && rl < p.rows) {
          const unsigned n = rl % p.Nn, bt = rl / p.Nn;
          const unsigned t = bt % p.Td, b = bt / p.Td;
          tq[m] = b * p.Ts + t + (unsigned)(p.Ts - p.Td);
          tn[m] = (t << 20) | n;
        }
      }
      for (int s = 0; s < p.nseg; ++s) {
        const TcSeg sg = p.seg[s];
        const int tofs = (p.Ts - p.Td) - sg.shift;
        const bf16* rp[2];
#pragma unroll
        for (int m = 0; m < 2; ++m) {
          const bool ok = tq[m] != 0xFFFFFFFFu && (int)(tn[m] >> 20) + tofs >= 0;
          const size_t row = (size_t)(tq[m] - (unsigned)sg.shift) * (unsigned)sg.pitch + (unsigned)sg.off + (tn[m] & 0xFFFFFu);
          rp[m] = ok ? sg.p + row * (unsigned)sg.ld : nullptr;
        }
        for (int k0 = 0; k0 < sg.width; k0 += p.kc) {
          tc::mbar_wait(&empty[stage], ephase);
          if (p.use_tma[s]) {
            if (ptid == 0) {
              tc::mbar_expect_tx(&full[stage], (uint32_t)p.boxrows * 16u * (uint32_t)cpr);
              if (p.sw128) tc::tma_load_2d(stages + (size_t)stage * p.stage_bytes, &p.tmap[s], k0, (int)row_base, &full[stage]);
              else tc::tma_load_3d(stages + (size_t)stage * p.stage_bytes, &p.tmap[s], 0, (int)row_base, k0 >> 3, &full[stage]);
            } else {
              mbar_arrive(&full[stage]);
            }
            if (++stage == NS) { stage = 0; ephase ^= 1; }
            continue;
          }
          uint8_t* sb = stages + (size_t)stage * p.stage_bytes + ptid * 16;
#pragma unroll
          for (int m = 0; m < 2; ++m) {
            if (m < p.MT && ptid + 128 * m < p.boxrows) {
              if (rp[m]) {
                const bf16* src = rp[m] + k0;
#pragma unroll
                for (int j = 0; j < 8; ++j)
                  if (j < cpr) cp_async16_full(sb + m * 2048 + j * p.pstrideA, src + j * 8);
              } else {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                  if (j < cpr) cp_async16(sb + m * 2048 + j * p.pstrideA, sg.p, 0u);
              }
            }
          }
          cp_async_arrive(&full[stage]);
          if (++stage == NS) { stage = 0; ephase ^= 1; }
        }
      }
    }
  } else if (warp == TC_EPI_WARPS + TC_PROD_WARPS) {
    // ================================ MMA issuer ================================
    // warp-uniform code, one elected lane issues (tc::elect_one): no ELECT / R2UR waterfall around the tcgen05 instructions
    {
      if (tc::elect_one()) {
        tc::mbar_expect_tx(wbar, p.w_bytes);
        for (uint32_t o = 0; o < p.w_bytes; o += 16384) {
          const uint32_t n = p.w_bytes - o < 16384u ? p.w_bytes - o : 16384u;
          tc::bulk_g2s(Wsm + o, p.Wimg + o, n, wbar);
        }
      }
      __syncwarp();
      tc::mbar_wait(wbar, 0);
      const uint32_t idesc = tc::idesc_bf16(128, p.Nout, 0, 0);
      bool any_cp = false;
      for (int s = 0; s < p.nseg; ++s) any_cp |= !p.use_tma[s];
      const uint32_t w_addr = tc::smem_u32(Wsm), st_addr = tc::smem_u32(stages);
      const uint32_t pstrideW = (uint32_t)p.Nout * 16u;
      const uint32_t stage_bytes = p.stage_bytes;
      const int kpc = p.kc / 16, Nout = p.Nout;
      const bool MT2 = p.MT > 1;
      const uint64_t a_desc0 = p.sw128 ? tc::smem_desc_sw128(st_addr) : tc::smem_desc(st_addr, p.pstrideA, 128u);
      const uint64_t w_desc0 = tc::smem_desc(w_addr, pstrideW, 128u);
      const uint64_t a_step = p.sw128 ? 2u : (uint64_t)((2u * p.pstrideA) >> 4), w_step = (uint64_t)((2u * pstrideW) >> 4);
      const uint64_t a_mt2 = p.sw128 ? 1024u : 128u;        // second M-tile: +128 rows
      unsigned tile_it = 0;
      int stage = 0;
      uint32_t fphase = 0;
      for (unsigned tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++tile_it) {
        const int as = tile_it & 1;
        if (tile_it >= 2) tc::mbar_wait(&tempty[as], ((tile_it >> 1) & 1) ^ 1);
        tc::tc_fence_after();
        const uint32_t acc = tmem + (uint32_t)as * 256u;
        uint64_t bd = w_desc0;
        for (int ch = 0; ch < nchunk; ++ch) {
          tc::mbar_wait(&full[stage], fphase);
          if (any_cp) tc::fence_async_smem();  // cp.async (generic proxy) writes -> tensor-core (async proxy) reads
          tc::tc_fence_after();
          // descriptors advance linearly with the start address: one 64-bit add per MMA in the issue loop
          uint64_t ad = a_desc0 + (uint64_t)(((uint32_t)stage * stage_bytes) >> 4);
#pragma unroll 4
          for (int ks = 0; ks < kpc; ++ks, ad += a_step, bd += w_step) {
            if (tc::elect_one()) {
              tc::umma(acc, ad, bd, idesc, (ch | ks) != 0);
              if (MT2) tc::umma(acc + (uint32_t)Nout, ad + a_mt2, bd, idesc, (ch | ks) != 0);
            }
          }
          if (tc::elect_one()) tc::umma_commit(&empty[stage]);
          __syncwarp();
          if (++stage == NS) { stage = 0; fphase ^= 1; }
        }
        if (tc::elect_one()) tc::umma_commit(&tfull[as]);
        __syncwarp();
      }
    }
  } else {
    // ================================ epilogue ================================
    const int q = warp & 3, hf = warp >> 2;
    const int half = p.Nout >> 1;
    const int RP = 128 * p.MT;
    // score/graph scratch (agg): ssp[2][RP] snp[2][RP] asf[Nout] anf[Nout] inv[RP] alpha[g*nnz] | rp_s[R+1] ci_s[nnz]
    float* bias_s = scbuf;              // [Nout], zeros when the layer has no bias
    float* ssp = scbuf + p.Nout;
    float* snp = ssp + 2 * RP;
    float* asf = ssp + 4 * RP;
    float* anf = asf + p.Nout;
    float* inv_s = anf + p.Nout;
    float* alpha = inv_s + RP;
    const int gsamp = p.agg ? p.rpt / p.R : 0;
    int* rp_s = reinterpret_cast<int*>(alpha + (p.agg == 2 ? gsamp * p.nnz : p.nnz));
    int* ci_s = rp_s + p.R + 1;
    for (int i = tid; i < p.Nout; i += TC_EPI_WARPS * 32) bias_s[i] = p.bias ? p.bias[i] : 0.f;
    float* hw_s = scbuf + p.Nout;       // narrow head (agg == 0 only): head_w[Nout], part[2][2][128]
    float* part = hw_s + p.Nout;
    if (p.head_w)
      for (int i = tid; i < p.Nout; i += TC_EPI_WARPS * 32) hw_s[i] = p.head_w[i];
    if (p.agg) {
      for (int i = tid; i <= p.R; i += TC_EPI_WARPS * 32) rp_s[i] = p.row_ptr[i];
      for (int i = tid; i < p.nnz; i += TC_EPI_WARPS * 32) { ci_s[i] = p.col_idx[i]; if (p.agg == 1) alpha[i] = p.val[i]; }
      if (p.agg == 2)
        for (int i = tid; i < p.Nout; i += TC_EPI_WARPS * 32) { asf[i] = p.a_self[i]; anf[i] = p.a_neigh[i]; }
    }
    epi_bar_sync();
    unsigned tile_it = 0;
    for (unsigned tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++tile_it) {
      const int as = tile_it & 1;
      unsigned row_base, tb_b, tb_rb; int nvalid;
      tile_rows(p, tile, row_base, nvalid, tb_b, tb_rb);
      tc::mbar_wait(&tfull[as], (tile_it >> 1) & 1);
      tc::tc_fence_after();
      for (int mt = 0; mt < p.MT; ++mt) {
        if (mt * 128 + q * 32 >= nvalid) continue;                 // warp-uniform
        const int rloc = mt * 128 + q * 32 + lane;
        const bool valid = rloc < nvalid;
        const uint32_t ta = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * 256 + mt * p.Nout);
        if (p.agg == 0 && p.head_w) {
          const int c0 = hf * half;
          float d = 0.f;
          switch (half) {                                          // warp-uniform
            case 16: d = epi_head_cols<16>(ta, c0, bias_s, hw_s, p.act); break;
            case 32: d = epi_head_cols<32>(ta, c0, bias_s, hw_s, p.act); break;
            case 64: d = epi_head_cols<64>(ta, c0, bias_s, hw_s, p.act); break;
            default:
              for (int c = c0; c < c0 + half; c += 16) d += epi_head_cols<16>(ta, c, bias_s, hw_s, p.act);
          }
          part[(as * 2 + hf) * 128 + rloc] = d;
        } else if (p.agg == 0) {
          const unsigned rl = row_base + rloc;
          const size_t rr = (size_t)(rl / p.Nn) * p.pitch_d + p.off_d + rl % p.Nn;
          bf16* orow = p.out + rr * p.ldo;
          const int c0 = hf * half;
          switch (half) {                                          // warp-uniform
            case 16: epi_store_cols<16>(ta, c0, valid, orow, bias_s, p.act); break;
            case 32: epi_store_cols<32>(ta, c0, valid, orow, bias_s, p.act); break;
            case 64: epi_store_cols<64>(ta, c0, valid, orow, bias_s, p.act); break;
            case 96: epi_store_cols<48>(ta, c0, valid, orow, bias_s, p.act); epi_store_cols<48>(ta, c0 + 48, valid, orow, bias_s, p.act); break;
            case 128: epi_store_cols<64>(ta, c0, valid, orow, bias_s, p.act); epi_store_cols<64>(ta, c0 + 64, valid, orow, bias_s, p.act); break;
            default:
              for (int c = c0; c < c0 + half; c += 16) epi_store_cols<16>(ta, c, valid, orow, bias_s, p.act);
          }
        } else {
          // h tile -> shared memory as bf16 rows; attention scores from the fp32 accumulators
          float ps = 0.f, pn = 0.f;
          uint8_t* hrow = reinterpret_cast<uint8_t*>(hbuf) + (size_t)rloc * p.hsb;
          for (int c0 = hf * half; c0 < (hf + 1) * half; c0 += 16) {
            float v[16];
            tc::tmem_ld16(ta + (uint32_t)c0, v);
            tc::tmem_ld_wait();
            if (valid) {
              uint32_t w[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) w[i] = pack2_bf16(v[2 * i], v[2 * i + 1]);
              *reinterpret_cast<uint4*>(hrow + c0 * 2) = make_uint4(w[0], w[1], w[2], w[3]);
              *reinterpret_cast<uint4*>(hrow + c0 * 2 + 16) = make_uint4(w[4], w[5], w[6], w[7]);
              if (p.agg == 2) {
#pragma unroll
                for (int i = 0; i < 16; ++i) { ps = fmaf(v[i], asf[c0 + i], ps); pn = fmaf(v[i], anf[c0 + i], pn); }
              }
            }
          }
          if (p.agg == 2 && valid) { ssp[hf * RP + rloc] = ps; snp[hf * RP + rloc] = pn; }
        }
      }
      tc::tc_fence_before();
      mbar_arrive(&tempty[as]);
      if (p.head_w && !p.agg) {
        epi_bar_sync();
        if (tid < nvalid) {                                        // MT == 1: tid is the tile row
          const unsigned rl = row_base + (unsigned)tid;
          const size_t rr = (size_t)(rl / p.Nn) * p.pitch_d + p.off_d + rl % p.Nn;
          p.head_out[rr * p.head_ldo + p.head_col] = apply_act(part[as * 256 + tid] + part[as * 256 + 128 + tid] + p.head_b[0], p.head_act);
        }
      }
      if (p.agg) {
        epi_bar_sync();
        const int R = p.R;
        if (p.agg == 2) {
          // edge softmax: one thread per row over its <= few dozen neighbours
          for (int rloc = tid; rloc < nvalid; rloc += TC_EPI_WARPS * 32) {
            int s = 0, i = rloc;
            while (i >= R) { i -= R; ++s; }
            const int beg = rp_s[i], end = rp_s[i + 1];
            const float si = ssp[rloc] + ssp[RP + rloc];
            const float* sn0 = snp + s * R;
            float mx = -INFINITY;
            for (int qq = beg; qq < end; ++qq) {
              const int j = ci_s[qq];
              float sc = si + sn0[j] + sn0[RP + j];
              sc = sc > 0.f ? sc : 0.2f * sc;
              mx = fmaxf(mx, sc);
            }
            float den = 0.f;
            float* al = alpha + s * p.nnz;
            for (int qq = beg; qq < end; ++qq) {
              const int j = ci_s[qq];
              float sc = si + sn0[j] + sn0[RP + j];
              sc = sc > 0.f ? sc : 0.2f * sc;
              const float w = __expf(sc - mx);
              al[qq] = w;
              den += w;
            }
            inv_s[rloc] = 1.f / den;
          }
          epi_bar_sync();
        }
        // weighted neighbour sum: 8 lanes per output row (E/8 features each), 4 rows per warp pass
        const int lg = lane >> 3, gl = lane & 7;
        for (int rloc = warp * 4 + lg; rloc < nvalid; rloc += TC_EPI_WARPS * 4) {
          int s = 0, i = rloc;
          while (i >= R) { i -= R; ++s; }
          const int beg = rp_s[i], end = rp_s[i + 1];
          const uint8_t* hb = reinterpret_cast<const uint8_t*>(hbuf) + (size_t)s * R * p.hsb + gl * 16;
          const float* al = alpha + (p.agg == 2 ? s * p.nnz : 0);
          const float scale = p.agg == 2 ? inv_s[rloc] : 1.f;
          bf16* orow = p.out + (size_t)(row_base + rloc) * p.ldo + gl * 8;
          const float* bp = bias_s + gl * 8;
          // lane gl owns features [64k + 8gl, 64k + 8gl + 8): a lane group reads 128 contiguous bytes
          for (int e0 = 0; e0 < p.Nout; e0 += 64) {
            float acc[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] = 0.f;
            for (int qq = beg; qq < end; ++qq) {
              const float w = al[qq];
              const uint4 u = *reinterpret_cast<const uint4*>(hb + (size_t)ci_s[qq] * p.hsb + e0 * 2);
              float hv[8];
              tc::unpack8_bf16(u, hv);
#pragma unroll
              for (int k = 0; k < 8; ++k) acc[k] = fmaf(w, hv[k], acc[k]);
            }
            const float4 b0 = *reinterpret_cast<const float4*>(bp + e0), b1 = *reinterpret_cast<const float4*>(bp + e0 + 4);
            acc[0] = fmaf(acc[0], scale, b0.x); acc[1] = fmaf(acc[1], scale, b0.y);
            acc[2] = fmaf(acc[2], scale, b0.z); acc[3] = fmaf(acc[3], scale, b0.w);
            acc[4] = fmaf(acc[4], scale, b1.x); acc[5] = fmaf(acc[5], scale, b1.y);
            acc[6] = fmaf(acc[6], scale, b1.z); acc[7] = fmaf(acc[7], scale, b1.w);
            act_tile<8>(acc, p.act);
            *reinterpret_cast<uint4*>(orow + e0) = make_uint4(pack2_bf16(acc[0], acc[1]), pack2_bf16(acc[2], acc[3]),
                                                              pack2_bf16(acc[4], acc[5]), pack2_bf16(acc[6], acc[7]));
          }
        }
        epi_bar_sync();
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_free(tmem, 512);
}

}  // namespace gu

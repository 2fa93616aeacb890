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
//                           of umma.cuh, NS-stage ring, fence.proxy.async + mbarrier hand-over
//   warp  12     MMA issuer one thread: tcgen05.mma (M=128, N=Nout, K=16), weights resident in
//                           shared memory as a K-major bf16 image fetched once by a bulk copy;
//                           accumulators double-buffered in TMEM so tile i's epilogue overlaps
//                           tile i+1's MMAs and loads.
// A tile is rpt logical rows in MT M-tiles: 128 arbitrary rows (plain), or g whole (b,t)
// samples of R rows when the epilogue aggregates over the graph.
#pragma once
#include "kernels.cuh"
#include "umma.cuh"

namespace gu {

constexpr int TC_EPI_WARPS = 8;
constexpr int TC_PROD_WARPS = 4;
constexpr int TC_THREADS = (TC_EPI_WARPS + TC_PROD_WARPS + 1) * 32;   // 416
constexpr int TC_LAG = 2;                                             // cp.async groups in flight per producer thread

struct TcSeg { const bf16* p; int ld, width, shift, pitch, off; };

struct TcGemmP {
  TcSeg seg[4]; int nseg;
  const uint8_t* Wimg;       // bf16 K-major image of W^T: element (n,k) at (k/8)*Nout*16 + n*16 + (k%8)*2
  const float* bias;         // (Nout) or null
  bf16* out; int ldo;
  int Nout, Ktot, kc, act;
  unsigned rows; int Nn, Td, Ts, pitch_d, off_d;
  int rpt, MT; unsigned n_tiles; int NS;
  uint32_t w_bytes, stage_bytes, pstrideA, stage_off, h_off, sc_off, bar_off;
  // aggregation epilogue (agg: 0 none, 1 GCN, 2 GAT) over samples of R rows
  int agg, R, hs;            // hs = row stride of the fp32 h tile in floats (Nout + 4)
  const int* row_ptr; const int* col_idx; const float* val;
  const float* a_self; const float* a_neigh;
};

__device__ __forceinline__ void cp_async16(void* dst_smem, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(tc::smem_u32(dst_smem)), "l"(src), "r"(src_bytes)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tc::smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(TC_EPI_WARPS * 32) : "memory"); }

__global__ void __launch_bounds__(TC_THREADS, 1) tc_rows_kernel(const __grid_constant__ TcGemmP p) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
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
  const uint32_t tmem = *tmem_slot;
  const int nchunk = p.Ktot / p.kc;           // K chunks (= pipeline stages) per tile

  if (warp >= TC_EPI_WARPS && warp < TC_EPI_WARPS + TC_PROD_WARPS) {
    // ================================ producers ================================
    const int ptid = tid - TC_EPI_WARPS * 32;
    const int cpr = p.kc >> 3;                 // 16-byte chunks per row per stage (4 or 8)
    const int rstep = (TC_PROD_WARPS * 32) / cpr;
    const int c = ptid % cpr, r0 = ptid / cpr;
    const int npass = (128 * p.MT) / rstep;
    unsigned it = 0;
    for (unsigned tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      const unsigned row_base = tile * (unsigned)p.rpt;
      for (int s = 0; s < p.nseg; ++s) {
        const TcSeg sg = p.seg[s];
        for (int k0 = 0; k0 < sg.width; k0 += p.kc, ++it) {
          const int stage = it % NS;
          if (it >= (unsigned)NS) tc::mbar_wait(&empty[stage], ((it / NS) & 1) ^ 1);
          uint8_t* sb = stages + (size_t)stage * p.stage_bytes + (size_t)c * p.pstrideA;
          for (int i = 0; i < npass; ++i) {
            const int rloc = r0 + i * rstep;
            const unsigned rl = row_base + rloc;
            const bf16* src = sg.p;
            uint32_t nb = 0;
            if (rloc < p.rpt && rl < p.rows) {
              const unsigned n = rl % p.Nn, bt = rl / p.Nn;
              const int t = bt % p.Td;
              const unsigned b = bt / p.Td;
              const int ts = t + (p.Ts - p.Td) - sg.shift;
              if (ts >= 0) {
                src = sg.p + ((size_t)(b * p.Ts + ts) * sg.pitch + sg.off + n) * sg.ld + k0 + c * 8;
                nb = 16;
              }
            }
            cp_async16(sb + rloc * 16, src, nb);
          }
          cp_async_commit();
          if (it >= TC_LAG) {
            cp_async_wait<TC_LAG>();
            tc::fence_async_smem();
            mbar_arrive(&full[(it - TC_LAG) % NS]);
          }
        }
      }
    }
    cp_async_wait<0>();
    tc::fence_async_smem();
    for (unsigned j = it > TC_LAG ? it - TC_LAG : 0; j < it; ++j) mbar_arrive(&full[j % NS]);
  } else if (warp == TC_EPI_WARPS + TC_PROD_WARPS) {
    // ================================ MMA issuer ================================
    if (lane == 0) {
      tc::mbar_expect_tx(wbar, p.w_bytes);
      for (uint32_t o = 0; o < p.w_bytes; o += 16384) {
        const uint32_t n = p.w_bytes - o < 16384u ? p.w_bytes - o : 16384u;
        tc::bulk_g2s(Wsm + o, p.Wimg + o, n, wbar);
      }
      tc::mbar_wait(wbar, 0);
      const uint32_t idesc = tc::idesc_bf16(128, p.Nout, 0, 0);
      const uint32_t w_addr = tc::smem_u32(Wsm), st_addr = tc::smem_u32(stages);
      const uint32_t pstrideW = (uint32_t)p.Nout * 16u;
      unsigned it = 0, tile_it = 0;
      for (unsigned tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++tile_it) {
        const int as = tile_it & 1;
        if (tile_it >= 2) tc::mbar_wait(&tempty[as], ((tile_it >> 1) & 1) ^ 1);
        tc::tc_fence_after();
        const uint32_t acc = tmem + (uint32_t)as * 256u;
        for (int ch = 0; ch < nchunk; ++ch, ++it) {
          const int stage = it % NS;
          tc::mbar_wait(&full[stage], (it / NS) & 1);
          tc::tc_fence_after();
          const uint32_t sa = st_addr + (uint32_t)stage * p.stage_bytes;
          for (int ks = 0; ks < p.kc / 16; ++ks) {
            const uint32_t kg = (uint32_t)(ch * (p.kc / 16) + ks);
            const uint64_t bd = tc::smem_desc(w_addr + 2u * kg * pstrideW, pstrideW, 128u);
            for (int mt = 0; mt < p.MT; ++mt) {
              const uint64_t ad = tc::smem_desc(sa + (uint32_t)mt * 2048u + 2u * ks * p.pstrideA, p.pstrideA, 128u);
              tc::umma(acc + (uint32_t)(mt * p.Nout), ad, bd, idesc, (ch | ks) != 0);
            }
          }
          tc::umma_commit(&empty[stage]);
        }
        tc::umma_commit(&tfull[as]);
      }
    }
  } else {
    // ================================ epilogue ================================
    const int q = warp & 3, hf = warp >> 2;
    const int half = p.Nout >> 1;
    const int RP = 128 * p.MT;
    float* ssp = scbuf;                 // [2][RP]
    float* snp = scbuf + 2 * RP;        // [2][RP]
    float* asf = scbuf + 4 * RP;        // [Nout]
    float* anf = asf + p.Nout;          // [Nout]
    if (p.agg == 2) {
      for (int i = tid; i < p.Nout; i += TC_EPI_WARPS * 32) { asf[i] = p.a_self[i]; anf[i] = p.a_neigh[i]; }
      epi_bar_sync();
    }
    unsigned tile_it = 0;
    for (unsigned tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++tile_it) {
      const int as = tile_it & 1;
      const unsigned row_base = tile * (unsigned)p.rpt;
      const int nvalid = (int)min((unsigned)p.rpt, p.rows - row_base);
      tc::mbar_wait(&tfull[as], (tile_it >> 1) & 1);
      tc::tc_fence_after();
      for (int mt = 0; mt < p.MT; ++mt) {
        if (mt * 128 + q * 32 >= nvalid) continue;                 // warp-uniform
        const int rloc = mt * 128 + q * 32 + lane;
        const bool valid = rloc < nvalid;
        const uint32_t ta = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * 256 + mt * p.Nout);
        if (p.agg == 0) {
          const unsigned rl = row_base + rloc;
          const size_t rr = (size_t)(rl / p.Nn) * p.pitch_d + p.off_d + rl % p.Nn;
          bf16* orow = p.out + rr * p.ldo;
          for (int c0 = hf * half; c0 < (hf + 1) * half; c0 += 16) {
            float v[16];
            tc::tmem_ld16(ta + (uint32_t)c0, v);
            tc::tmem_ld_wait();
            if (valid) {
              uint32_t w[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                float a = v[2 * i], b = v[2 * i + 1];
                if (p.bias) { a += p.bias[c0 + 2 * i]; b += p.bias[c0 + 2 * i + 1]; }
                w[i] = pack2_bf16(apply_act(a, p.act), apply_act(b, p.act));
              }
              *reinterpret_cast<uint4*>(orow + c0) = make_uint4(w[0], w[1], w[2], w[3]);
              *reinterpret_cast<uint4*>(orow + c0 + 8) = make_uint4(w[4], w[5], w[6], w[7]);
            }
          }
        } else {
          float ps = 0.f, pn = 0.f;
          for (int c0 = hf * half; c0 < (hf + 1) * half; c0 += 16) {
            float v[16];
            tc::tmem_ld16(ta + (uint32_t)c0, v);
            tc::tmem_ld_wait();
            if (valid) {
              float* hr = hbuf + (size_t)rloc * p.hs + c0;
#pragma unroll
              for (int i = 0; i < 4; ++i)
                *reinterpret_cast<float4*>(hr + 4 * i) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
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
      if (p.agg) {
        epi_bar_sync();
        // neighbour aggregation: one warp per output row, lanes over features (float4)
        const int E = p.Nout;
        for (int rloc = warp; rloc < nvalid; rloc += TC_EPI_WARPS) {
          const int s = rloc / p.R, i = rloc - s * p.R;
          const int beg = p.row_ptr[i], end = p.row_ptr[i + 1];
          const float* hb = hbuf + (size_t)s * p.R * p.hs;
          float mx = -INFINITY, si = 0.f;
          if (p.agg == 2) {
            si = ssp[rloc] + ssp[RP + rloc];
            for (int qq = beg + lane; qq < end; qq += 32) {
              const int j = s * p.R + p.col_idx[qq];
              float sc = si + snp[j] + snp[RP + j];
              sc = sc > 0.f ? sc : 0.2f * sc;
              mx = fmaxf(mx, sc);
            }
            mx = warp_max(mx);
          }
          const unsigned rl = row_base + rloc;
          const size_t rr = (size_t)(rl / p.Nn) * p.pitch_d + p.off_d + rl % p.Nn;
          for (int c0 = 0; c0 < E; c0 += 128) {
            const int c = c0 + lane * 4;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            float den = 0.f;
            for (int q0 = beg; q0 < end; q0 += 32) {
              const int qq = q0 + lane;
              int cj = 0; float wj = 0.f;
              if (qq < end) {
                cj = p.col_idx[qq];
                if (p.agg == 2) {
                  const int j = s * p.R + cj;
                  float sc = si + snp[j] + snp[RP + j];
                  sc = sc > 0.f ? sc : 0.2f * sc;
                  wj = expf(sc - mx);
                } else {
                  wj = p.val[qq];
                }
              }
              if (p.agg == 2) den += wj;
              const int cnt = min(32, end - q0);
              for (int u = 0; u < cnt; ++u) {
                const float w = __shfl_sync(0xffffffffu, wj, u);
                const int j = __shfl_sync(0xffffffffu, cj, u);
                if (c < E) {
                  const float4 hv = *reinterpret_cast<const float4*>(hb + (size_t)j * p.hs + c);
                  acc.x = fmaf(w, hv.x, acc.x); acc.y = fmaf(w, hv.y, acc.y);
                  acc.z = fmaf(w, hv.z, acc.z); acc.w = fmaf(w, hv.w, acc.w);
                }
              }
            }
            if (p.agg == 2) {
              den = warp_sum(den);
              const float inv = 1.f / den;
              acc.x *= inv; acc.y *= inv; acc.z *= inv; acc.w *= inv;
            }
            if (c < E) {
              const float4 bb = *reinterpret_cast<const float4*>(p.bias + c);
              float4 v;
              v.x = apply_act(acc.x + bb.x, p.act); v.y = apply_act(acc.y + bb.y, p.act);
              v.z = apply_act(acc.z + bb.z, p.act); v.w = apply_act(acc.w + bb.w, p.act);
              st4(p.out + rr * p.ldo + c, v);
            }
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

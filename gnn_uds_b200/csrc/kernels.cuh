// Device kernels of the Emulator forward (sm_100a).  fp32 SIMT parity path.
//
// Activation layout everywhere: rows = (scenario b, time t, element n) flattened,
// features contiguous -> [B][T][Nn][C].  In this layout the reference's
// transpose/reshape pairs around the temporal nets (emulator.py:244-257,299-310)
// disappear: a causal dilated Conv1D tap is the same GEMM reading rows shifted
// by shift*Nn, and the GRU walks rows with stride Nn.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace gu {

// Activation storage: float (fp32 parity path) or a 16-bit type (tensor-core path).  Every kernel below
// that touches an activation tensor is a template over the storage type and computes in fp32.
// The 16-bit type - historically named `bf16` throughout - is IEEE fp16 by default (GU_F16 = 1: 11 significand bits,
// operands of tcgen05 kind::f16 as F16 x F16 -> F32) and bfloat16 when built with -DGNNUDS_TC_BF16 (8 bits, wider range).
// fp16 cuts the tensor-core path's deviation from the fp32 path by ~8x; the model's activations are O(1..100), far
// from fp16's 65504.
#ifdef GNNUDS_TC_BF16
#define GU_F16 0
typedef __nv_bfloat16 bf16;
#else
#define GU_F16 1
typedef __half bf16;
#endif
// two packed 16-bit storage values <-> fp32
__device__ __forceinline__ float2 unpack2_16(uint32_t u) {
#if GU_F16
  return __half22float2(*reinterpret_cast<const __half2*>(&u));
#else
  return make_float2(__uint_as_float(u << 16), __uint_as_float(u & 0xFFFF0000u));
#endif
}
__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ float4 ld4(const bf16* p) {
  const uint2 u = *reinterpret_cast<const uint2*>(p);
  const float2 a = unpack2_16(u.x), b = unpack2_16(u.y);
  return make_float4(a.x, a.y, b.x, b.y);
}
// fp16 storage saturates at +-65504 instead of overflowing to inf (a trained model with large activations must not
// turn into inf -> NaN behind the bounded heads); NaN stays NaN
__device__ __forceinline__ uint32_t cvt_f16x2_sat(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ uint32_t pack2_bf16(float lo, float hi) {
#if GU_F16
  return cvt_f16x2_sat(lo, hi);
#else
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
#endif
}
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ void st4(bf16* p, float4 v) {
  *reinterpret_cast<uint2*>(p) = make_uint2(pack2_bf16(v.x, v.y), pack2_bf16(v.z, v.w));
}
__device__ __forceinline__ float ld1(const float* p) { return *p; }
#if GU_F16
__device__ __forceinline__ float ld1(const bf16* p) { return __half2float(*p); }
#else
__device__ __forceinline__ float ld1(const bf16* p) { return __bfloat162float(*p); }
#endif
__device__ __forceinline__ void st1(float* p, float v) { *p = v; }
#if GU_F16
__device__ __forceinline__ void st1(bf16* p, float v) { *reinterpret_cast<unsigned short*>(p) = (unsigned short)(cvt_f16x2_sat(v, 0.f) & 0xFFFFu); }
#else
__device__ __forceinline__ void st1(bf16* p, float v) { *p = __float2bfloat16_rn(v); }
#endif

enum { ACT_LINEAR = 0, ACT_RELU = 1, ACT_TANH = 2, ACT_SIGMOID = 3, ACT_HARD_SIGMOID = 4 };

__device__ __forceinline__ float apply_act(float v, int act) {
  switch (act) {
    case ACT_RELU: return fmaxf(v, 0.f);
    case ACT_TANH: return tanhf(v);
    case ACT_SIGMOID: return 1.f / (1.f + expf(-v));
    case ACT_HARD_SIGMOID: return fminf(fmaxf(0.2f * v + 0.5f, 0.f), 1.f);   // keras 2.x hard_sigmoid
    default: return v;
  }
}

__device__ __forceinline__ void act_tile4(float* v, int act) {
  if (act == ACT_RELU) {
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = fmaxf(v[i], 0.f);
  } else if (act != ACT_LINEAR) {
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = apply_act(v[i], act);
  }
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---------------------------------------------------------------------------
// K1 embedding (emulator.py:195-213): Dense over a tiny channel count fused with
// the min-max normalisation (emulator.py:803-810) and, for the control input,
// with get_edge_action's scatter (emulator.py:385-398).
// ---------------------------------------------------------------------------
struct EmbedP {
  const float* in;        // (Bc, T, Nn, K) raw physical values; mode 2: a (Bc, T, n_act)
  const float* nmax;      // (Nn, cn) or null (no normalisation)
  const float* nmin;
  int cn;
  const int* gather;      // mode 2: act_of_edge (Nn) column of a or -1
  int n_act;
  const float* W;         // (K, Eo)
  const float* bias;      // (Eo)
  void* out;              // (rows, Eo) activated, storage type T
  float* res;             // (Bc*Nn, Eo) pre-activation of the last time step, or null
  long long rows;
  int T, Nn, K, Eo, act, mode;   // mode 0: plain, 2: gathered control input (K == 1)
};

template <typename T>
__global__ void __launch_bounds__(256) embed_kernel(EmbedP p) {
  extern __shared__ float sm[];
  float* Ws = sm;                      // K*Eo
  float* bs = sm + p.K * p.Eo;         // Eo
  for (int i = threadIdx.x; i < p.K * p.Eo; i += blockDim.x) Ws[i] = p.W[i];
  for (int i = threadIdx.x; i < p.Eo; i += blockDim.x) bs[i] = p.bias[i];
  __syncthreads();
  const int g4 = p.Eo >> 2;                       // float4 groups per row
  const int rows_per_blk = blockDim.x / g4;
  const int lr = threadIdx.x / g4, c4 = threadIdx.x % g4;
  if (lr >= rows_per_blk) return;
  if (g4 <= 32 && g4 >= p.K && (g4 & (g4 - 1)) == 0 && p.rows < (1ll << 31)) {
    // fast path: the g4 lanes of a row split the K input channels (one normalisation each) and exchange them
    // by shuffle; 32-bit row arithmetic.  Same operation order per output as the generic path below.
    const unsigned rows = (unsigned)p.rows, Nn = (unsigned)p.Nn, Tn = (unsigned)p.T;
    const float4 bb = *reinterpret_cast<const float4*>(bs + c4 * 4);
    float mx = 1.f, mn = 0.f;
    const unsigned step = gridDim.x * (unsigned)rows_per_blk;
    const unsigned r_first = blockIdx.x * (unsigned)rows_per_blk + (unsigned)lr;
    const unsigned n_iter = (rows + step - 1) / step;          // uniform trip count: shuffles stay convergent
    for (unsigned it = 0, r = r_first; it < n_iter; ++it, r += step) {
      const bool live = r < rows;
      const unsigned rr = live ? r : 0;
      const unsigned n = rr % Nn, bt = rr / Nn;
      float mine = 0.f;
      if (p.mode == 2) {
        if (c4 == 0) { const int k = p.gather[n]; mine = (k >= 0) ? p.in[(size_t)bt * p.n_act + k] : 1.f; }
      } else if (c4 < p.K) {
        mine = p.in[(size_t)rr * p.K + c4];
        if (p.nmax) {
          mx = p.nmax[n * p.cn + c4]; mn = p.nmin[n * p.cn + c4];
          mine = (mine - mn) / (mx - mn);
        }
      }
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int c = 0; c < p.K; ++c) {
        const float xc = __shfl_sync(0xffffffffu, mine, c, g4);
        const float4 w = *reinterpret_cast<const float4*>(Ws + c * p.Eo + c4 * 4);
        s.x = fmaf(xc, w.x, s.x); s.y = fmaf(xc, w.y, s.y); s.z = fmaf(xc, w.z, s.z); s.w = fmaf(xc, w.w, s.w);
      }
      float4 acc = make_float4(bb.x + s.x, bb.y + s.y, bb.z + s.z, bb.w + s.w);
      if (!live) continue;
      if (p.res) {
        const unsigned t = bt % Tn;
        if (t == Tn - 1) *reinterpret_cast<float4*>(p.res + ((size_t)(bt / Tn) * Nn + n) * p.Eo + c4 * 4) = acc;
      }
      acc.x = apply_act(acc.x, p.act); acc.y = apply_act(acc.y, p.act);
      acc.z = apply_act(acc.z, p.act); acc.w = apply_act(acc.w, p.act);
      st4(static_cast<T*>(p.out) + (size_t)rr * p.Eo + c4 * 4, acc);
    }
    return;
  }
  for (long long r = (long long)blockIdx.x * rows_per_blk + lr; r < p.rows;
       r += (long long)gridDim.x * rows_per_blk) {
    const int n = (int)(r % p.Nn);
    const long long bt = r / p.Nn;
    float xin[8];
    if (p.mode == 2) {
      const int k = p.gather[n];
      xin[0] = (k >= 0) ? p.in[bt * p.n_act + k] : 1.f;
    } else {
      for (int c = 0; c < p.K; ++c) {
        float v = p.in[r * p.K + c];
        if (p.nmax) {
          const float mx = p.nmax[n * p.cn + c], mn = p.nmin[n * p.cn + c];
          v = (v - mn) / (mx - mn);
        }
        xin[c] = v;
      }
    }
    float4 acc = *reinterpret_cast<const float4*>(bs + c4 * 4);
    // keras Dense: x @ W + b, accumulate over input channels in order
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int c = 0; c < p.K; ++c) {
      const float4 w = *reinterpret_cast<const float4*>(Ws + c * p.Eo + c4 * 4);
      s.x = fmaf(xin[c], w.x, s.x); s.y = fmaf(xin[c], w.y, s.y);
      s.z = fmaf(xin[c], w.z, s.z); s.w = fmaf(xin[c], w.w, s.w);
    }
    acc.x += s.x; acc.y += s.y; acc.z += s.z; acc.w += s.w;
    if (p.res) {
      const int t = (int)(bt % p.T);
      if (t == p.T - 1) {
        const long long b = bt / p.T;
        *reinterpret_cast<float4*>(p.res + (b * p.Nn + n) * p.Eo + c4 * 4) = acc;
      }
    }
    acc.x = apply_act(acc.x, p.act); acc.y = apply_act(acc.y, p.act);
    acc.z = apply_act(acc.z, p.act); acc.w = apply_act(acc.w, p.act);
    st4(static_cast<T*>(p.out) + r * p.Eo + c4 * 4, acc);
  }
}

// Tensor-core path variant of the embedding (bf16 rows out, mode 0, K <= 8, Eo = 8*LPR): a thread owns 8
// output columns of a row (one 16-byte store), reads its K x 8 weight slice from shared memory, and walks
// rows with incremental (n, t) bookkeeping instead of a division per row.  The lanes c < K of a row's lane
// group each load and normalise one input channel and broadcast it by shuffle.  Two rows per iteration keep
// two independent global loads in flight per thread.
struct EmbedRow { unsigned r, n, bt, t; };

template <int LPR, int KT>                          // KT: compile-time channel count (full unrolling), 0 = runtime p.K
__global__ void __launch_bounds__(256, 4) embed_bf16_kernel(EmbedP p) {
  constexpr int RPB = 256 / LPR;                   // rows per CTA pass
  constexpr int Eo = LPR * 8;
  __shared__ __align__(16) float Ws[8 * LPR * 8];
  __shared__ __align__(16) float bs[LPR * 8];
  for (int i = threadIdx.x; i < p.K * p.Eo; i += 256) Ws[i] = p.W[i];
  for (int i = threadIdx.x; i < p.Eo; i += 256) bs[i] = p.bias[i];
  __syncthreads();
  const int lr = threadIdx.x / LPR, c8 = threadIdx.x % LPR;
  const unsigned rows = (unsigned)p.rows, Nn = (unsigned)p.Nn, Tn = (unsigned)p.T;
  const unsigned step = gridDim.x * (unsigned)RPB;
  const unsigned step_n = step % Nn, step_bt = step / Nn, step_t = step_bt % Tn;
  EmbedRow cur;
  cur.r = blockIdx.x * (unsigned)RPB + (unsigned)lr;
  cur.n = cur.r % Nn; cur.bt = cur.r / Nn; cur.t = cur.bt % Tn;
  auto advance = [&](EmbedRow& q) {
    q.r += step; q.n += step_n; q.bt += step_bt; q.t += step_t;
    if (q.n >= Nn) { q.n -= Nn; ++q.bt; ++q.t; }
    if (q.t >= Tn) q.t -= Tn;
  };
  const unsigned n_iter = (rows + step - 1) / step;   // uniform trip count: shuffles stay convergent
  const int K = KT ? KT : p.K;
  const float* wcol = Ws + c8 * 8;
  // a thread keeps its column slice for all its rows: with a compile-time channel count the K x 8 weights and the bias
  // live in registers (re-reading them from shared memory cost 12 LDS.128 per thread and row = 24 clk per row and SM,
  // twice what the 16-byte stores of the row need at the HBM rate)
  float wr[KT ? KT : 1][8], br[8];
#pragma unroll
  for (int c = 0; c < (KT ? KT : 1); ++c)
#pragma unroll
    for (int j = 0; j < 8; ++j) wr[c][j] = KT ? wcol[c * Eo + j] : 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) br[j] = bs[c8 * 8 + j];
  for (unsigned it = 0; it < n_iter; it += 2) {
    EmbedRow q[2];
    q[0] = cur; advance(cur); q[1] = cur; advance(cur);
    float mine[2] = {0.f, 0.f};
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (q[u].r < rows && c8 < K) {
        mine[u] = __ldg(p.in + (size_t)q[u].r * K + c8);
        if (p.nmax) {
          const float mx = __ldg(p.nmax + q[u].n * p.cn + c8), mn = __ldg(p.nmin + q[u].n * p.cn + c8);
          mine[u] = __fdividef(mine[u] - mn, mx - mn);      // 2 ulp: far below the 16-bit rounding of this path's output
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      float s[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] = 0.f;
#pragma unroll
      for (int c = 0; c < (KT ? KT : 8); ++c) {      // warp-uniform trip count
        if (!KT && c >= K) break;
        const float xc = __shfl_sync(0xffffffffu, mine[u], c, LPR);
        if (KT) {
#pragma unroll
          for (int j = 0; j < 8; ++j) s[j] = fmaf(xc, wr[KT ? c : 0][j], s[j]);
        } else {
          const float4 w0 = *reinterpret_cast<const float4*>(wcol + c * Eo), w1 = *reinterpret_cast<const float4*>(wcol + c * Eo + 4);
          s[0] = fmaf(xc, w0.x, s[0]); s[1] = fmaf(xc, w0.y, s[1]); s[2] = fmaf(xc, w0.z, s[2]); s[3] = fmaf(xc, w0.w, s[3]);
          s[4] = fmaf(xc, w1.x, s[4]); s[5] = fmaf(xc, w1.y, s[5]); s[6] = fmaf(xc, w1.z, s[6]); s[7] = fmaf(xc, w1.w, s[7]);
        }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] += br[j];
      if (q[u].r < rows) {
        if (p.res && q[u].t == Tn - 1) {
          float* rp = p.res + ((size_t)(q[u].bt / Tn) * Nn + q[u].n) * Eo + c8 * 8;
          *reinterpret_cast<float4*>(rp) = make_float4(s[0], s[1], s[2], s[3]);
          *reinterpret_cast<float4*>(rp + 4) = make_float4(s[4], s[5], s[6], s[7]);
        }
        if (p.act == ACT_RELU) {
#pragma unroll
          for (int j = 0; j < 8; ++j) s[j] = fmaxf(s[j], 0.f);
        } else if (p.act != ACT_LINEAR) {
#pragma unroll
          for (int j = 0; j < 8; ++j) s[j] = apply_act(s[j], p.act);
        }
        *reinterpret_cast<uint4*>(static_cast<bf16*>(p.out) + (size_t)q[u].r * Eo + c8 * 8) =
            make_uint4(pack2_bf16(s[0], s[1]), pack2_bf16(s[2], s[3]), pack2_bf16(s[4], s[5]), pack2_bf16(s[6], s[7]));
      }
    }
  }
}

// ---------------------------------------------------------------------------
// rows GEMM (fp32 SIMT): out[r,:] = act( sum_seg A_seg[src(r,seg), :] @ W_seg + bias ).
// Covers every Dense, the GCN/GAT linear part on the concatenated [x | NodeEdge(x_e)]
// input (two segments, emulator.py:227-230), the causal dilated Conv1D (k shifted
// segments, emulator.py:156-157) and the GRU/LSTM input projections.
// ---------------------------------------------------------------------------
// Logical rows are (b, t, n), n < Nn.  A segment's physical row is
// (b*Ts + t + (Ts-Td) - shift) * pitch + off + n (zero when the time index is negative);
// the destination row is (b*Td + t) * pitch_d + off_d + n.  pitch/off let node and edge
// rows share one joint buffer (graph_base != 0).
struct Seg { const float* p; int ld; int width; int shift; int pitch; int off; };
struct GemmP {
  Seg seg[4]; int nseg;
  const float* W;        // (sum width, Nout) row-major
  const float* bias;     // (Nout) or null
  float* out; int ldo; int Nout; int act;
  long long rows;        // logical rows = Bc*Td*Nn
  int Nn, Td, Ts;        // destination / source time lengths (Ts >= Td keeps the last Td steps)
  int pitch_d, off_d;
};

template <int BN>
__global__ void __launch_bounds__(256) rows_gemm_f32(GemmP p) {
  constexpr int BM = 128, BK = 16, TN = BN / 16;
  __shared__ __align__(16) float As[2][BK][BM];
  __shared__ __align__(16) float Ws[2][BK][BN];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const long long row0 = (long long)blockIdx.x * BM;
  const int col0 = blockIdx.y * BN;

  // loader mapping: one A row per thread-pair
  const int lrow = tid & 127, lhalf = tid >> 7;
  const long long r = row0 + lrow;
  long long srow[4];
  {
    const int n = (int)(r % p.Nn);
    const long long bt = r / p.Nn;
    const int t = (int)(bt % p.Td);
    const long long b = bt / p.Td;
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      srow[s] = -1;
      if (s < p.nseg && r < p.rows) {
        const int ts = t + (p.Ts - p.Td) - p.seg[s].shift;
        if (ts >= 0) srow[s] = (b * p.Ts + ts) * p.seg[s].pitch + p.seg[s].off + n;
      }
    }
  }
  float acc[8][TN];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  int nk = 0;
  for (int s = 0; s < p.nseg; ++s) nk += p.seg[s].width / BK;

  constexpr int WLD = BN * BK / 4 / 256;   // float4 loads of W per thread (1 for BN=64, 2 for BN=128)
  float4 ra0, ra1;
  float4 rw[WLD];

  int seg_i = 0, seg_k = 0, kglob = 0;
  auto load_tile = [&]() {
    const Seg& sg = p.seg[seg_i];
    const long long sr = seg_i == 0 ? srow[0] : seg_i == 1 ? srow[1] : seg_i == 2 ? srow[2] : srow[3];
    if (sr >= 0) {
      const float* ap = sg.p + sr * sg.ld + seg_k + lhalf * 8;
      ra0 = *reinterpret_cast<const float4*>(ap);
      ra1 = *reinterpret_cast<const float4*>(ap + 4);
    } else {
      ra0 = make_float4(0.f, 0.f, 0.f, 0.f); ra1 = ra0;
    }
#pragma unroll
    for (int i = 0; i < WLD; ++i) {
      const int idx = tid + i * 256;
      const int kk = idx / (BN / 4), c4 = idx % (BN / 4);
      rw[i] = *reinterpret_cast<const float4*>(p.W + (long long)(kglob + kk) * p.Nout + col0 + c4 * 4);
    }
    kglob += BK; seg_k += BK;
    if (seg_k >= sg.width) { seg_k = 0; ++seg_i; }
  };
  auto store_tile = [&](int buf) {
    const int k0 = lhalf * 8;
    As[buf][k0 + 0][lrow] = ra0.x; As[buf][k0 + 1][lrow] = ra0.y;
    As[buf][k0 + 2][lrow] = ra0.z; As[buf][k0 + 3][lrow] = ra0.w;
    As[buf][k0 + 4][lrow] = ra1.x; As[buf][k0 + 5][lrow] = ra1.y;
    As[buf][k0 + 6][lrow] = ra1.z; As[buf][k0 + 7][lrow] = ra1.w;
#pragma unroll
    for (int i = 0; i < WLD; ++i) {
      const int idx = tid + i * 256;
      const int kk = idx / (BN / 4), c4 = idx % (BN / 4);
      *reinterpret_cast<float4*>(&Ws[buf][kk][c4 * 4]) = rw[i];
    }
  };

  load_tile();
  store_tile(0);
  __syncthreads();
  int cur = 0;
  for (int it = 0; it < nk; ++it) {
    if (it + 1 < nk) load_tile();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      const float4 a0 = *reinterpret_cast<const float4*>(&As[cur][kk][ty * 8]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[cur][kk][ty * 8 + 4]);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      float b[TN];
#pragma unroll
      for (int j4 = 0; j4 < TN / 4; ++j4) {
        const float4 bv = *reinterpret_cast<const float4*>(&Ws[cur][kk][j4 * 64 + tx * 4]);
        b[j4 * 4 + 0] = bv.x; b[j4 * 4 + 1] = bv.y; b[j4 * 4 + 2] = bv.z; b[j4 * 4 + 3] = bv.w;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (it + 1 < nk) store_tile(cur ^ 1);
    __syncthreads();
    cur ^= 1;
  }

#pragma unroll
  for (int j4 = 0; j4 < TN / 4; ++j4) {
    const int col = col0 + j4 * 64 + tx * 4;
    float4 bb = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p.bias) bb = *reinterpret_cast<const float4*>(p.bias + col);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const long long rl = row0 + ty * 8 + i;
      if (rl < p.rows) {
        const long long rr = (rl / p.Nn) * p.pitch_d + p.off_d + (rl % p.Nn);
        float4 v;
        v.x = apply_act(acc[i][j4 * 4 + 0] + bb.x, p.act);
        v.y = apply_act(acc[i][j4 * 4 + 1] + bb.y, p.act);
        v.z = apply_act(acc[i][j4 * 4 + 2] + bb.z, p.act);
        v.w = apply_act(acc[i][j4 * 4 + 3] + bb.w, p.act);
        *reinterpret_cast<float4*>(p.out + rr * p.ldo + col) = v;
      }
    }
  }
}

// Small row counts (one scenario of a 5-step chunk is 150 rows: simulate per event, the gradient MPC's single decision
// vector): the 128-row kernel above runs 2 CTAs through Ktot/16 dependent load -> compute rounds, ~1 us each.  Here a CTA
// owns 16 rows x 32 columns, brings its whole A tile (16 x Ktot) and W tile (Ktot x 32) into shared memory with ONE
// round of cp.async, and 128 threads (row x 4 columns) accumulate over k in ascending order with one fmaf chain per
// output - the same order as rows_gemm_f32, so the results are bitwise the same.  Dynamic smem: (16 (Ktot + 4) + 32 Ktot) floats.
__device__ __forceinline__ void cp_async_f4(float* dst_smem, const float* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
__global__ void __launch_bounds__(128) rows_gemm_f32_skinny(GemmP p, int Ktot) {
  extern __shared__ __align__(16) float skinny_sm[];
  const int LDA = Ktot + 4;
  float* As = skinny_sm;                 // [16][LDA]
  float* Ws = skinny_sm + 16 * LDA;      // [Ktot][32]
  const int tid = threadIdx.x;
  const long long row0 = (long long)blockIdx.x * 16;
  const int col0 = blockIdx.y * 32;
  const int kq_n = Ktot >> 2;
  for (int idx = tid; idx < Ktot * 8; idx += 128) {
    const int k = idx >> 3, c4 = idx & 7;
    cp_async_f4(Ws + k * 32 + c4 * 4, p.W + (long long)k * p.Nout + col0 + c4 * 4);
  }
  // Programmatic dependent launch (launch_gemm sets the attribute): the weight tile above does not depend on the kernel
  // before this one in the stream; everything below does.  Without the launch attribute the wait returns at once.
  asm volatile("griddepcontrol.wait;" ::: "memory");
  for (int idx = tid; idx < 16 * kq_n; idx += 128) {
    const int lr = idx / kq_n, k = (idx - lr * kq_n) * 4;
    const long long r = row0 + lr;
    const float* src = nullptr;
    if (r < p.rows) {
      const int n = (int)(r % p.Nn);
      const long long bt = r / p.Nn;
      const int t = (int)(bt % p.Td);
      const long long b = bt / p.Td;
      int k0 = 0;
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        if (s < p.nseg) {
          if (k >= k0 && k < k0 + p.seg[s].width) {
            const int ts = t + (p.Ts - p.Td) - p.seg[s].shift;
            if (ts >= 0) src = p.seg[s].p + ((b * p.Ts + ts) * p.seg[s].pitch + p.seg[s].off + n) * p.seg[s].ld + (k - k0);
          }
          k0 += p.seg[s].width;
        }
      }
    }
    float* dst = As + lr * LDA + k;
    if (src) cp_async_f4(dst, src);
    else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");      // a dependent small-tile GEMM may start its weight prefetch
  const int lr = tid >> 3, cg = tid & 7;
  const float* ar = As + lr * LDA;
  const float* wr = Ws + cg * 4;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 2
  for (int k = 0; k < Ktot; k += 4) {
    const float4 a = *reinterpret_cast<const float4*>(ar + k);
    const float4 w0 = *reinterpret_cast<const float4*>(wr + (k + 0) * 32);
    const float4 w1 = *reinterpret_cast<const float4*>(wr + (k + 1) * 32);
    const float4 w2 = *reinterpret_cast<const float4*>(wr + (k + 2) * 32);
    const float4 w3 = *reinterpret_cast<const float4*>(wr + (k + 3) * 32);
    acc.x = fmaf(a.x, w0.x, acc.x); acc.y = fmaf(a.x, w0.y, acc.y); acc.z = fmaf(a.x, w0.z, acc.z); acc.w = fmaf(a.x, w0.w, acc.w);
    acc.x = fmaf(a.y, w1.x, acc.x); acc.y = fmaf(a.y, w1.y, acc.y); acc.z = fmaf(a.y, w1.z, acc.z); acc.w = fmaf(a.y, w1.w, acc.w);
    acc.x = fmaf(a.z, w2.x, acc.x); acc.y = fmaf(a.z, w2.y, acc.y); acc.z = fmaf(a.z, w2.z, acc.z); acc.w = fmaf(a.z, w2.w, acc.w);
    acc.x = fmaf(a.w, w3.x, acc.x); acc.y = fmaf(a.w, w3.y, acc.y); acc.z = fmaf(a.w, w3.z, acc.z); acc.w = fmaf(a.w, w3.w, acc.w);
  }
  const long long rl = row0 + lr;
  if (rl < p.rows) {
    const int col = col0 + cg * 4;
    float4 bb = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p.bias) bb = *reinterpret_cast<const float4*>(p.bias + col);
    const long long rr = (rl / p.Nn) * p.pitch_d + p.off_d + (rl % p.Nn);
    float4 v;
    v.x = apply_act(acc.x + bb.x, p.act); v.y = apply_act(acc.y + bb.y, p.act);
    v.z = apply_act(acc.z + bb.z, p.act); v.w = apply_act(acc.w + bb.w, p.act);
    *reinterpret_cast<float4*>(p.out + rr * p.ldo + col) = v;
  }
}

// Narrow-output variant (Nout a multiple of 16 or 32 but not 64: E/2 or H of small models).
// Same row addressing as rows_gemm_f32; 128 rows x BN columns per CTA, BN/8 rows x 4 columns
// per thread.
template <int BN>
__global__ void __launch_bounds__(256) rows_gemm_f32_narrow(GemmP p) {
  constexpr int BM = 128, BK = 16, TX = BN / 4, TY = 256 / TX, RM = BM / TY;
  __shared__ __align__(16) float As[BK][BM + 4];
  __shared__ __align__(16) float Ws[BK][BN];
  const int tid = threadIdx.x;
  const int tx = tid % TX, ty = tid / TX;
  const long long row0 = (long long)blockIdx.x * BM;
  const int col0 = blockIdx.y * BN;
  const int lrow = tid & 127, lhalf = tid >> 7;
  const long long r = row0 + lrow;
  long long srow[4];
  {
    const int n = (int)(r % p.Nn);
    const long long bt = r / p.Nn;
    const int t = (int)(bt % p.Td);
    const long long b = bt / p.Td;
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      srow[s] = -1;
      if (s < p.nseg && r < p.rows) {
        const int ts = t + (p.Ts - p.Td) - p.seg[s].shift;
        if (ts >= 0) srow[s] = (b * p.Ts + ts) * p.seg[s].pitch + p.seg[s].off + n;
      }
    }
  }
  float acc[RM][4];
#pragma unroll
  for (int i = 0; i < RM; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  int kglob = 0;
  for (int s = 0; s < p.nseg; ++s) {
    const Seg& sg = p.seg[s];
    const long long sr = s == 0 ? srow[0] : s == 1 ? srow[1] : s == 2 ? srow[2] : srow[3];
    for (int k0 = 0; k0 < sg.width; k0 += BK, kglob += BK) {
      float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
      if (sr >= 0) {
        const float* ap = sg.p + sr * sg.ld + k0 + lhalf * 8;
        a0 = *reinterpret_cast<const float4*>(ap);
        a1 = *reinterpret_cast<const float4*>(ap + 4);
      }
      __syncthreads();
      const int kb = lhalf * 8;
      As[kb + 0][lrow] = a0.x; As[kb + 1][lrow] = a0.y; As[kb + 2][lrow] = a0.z; As[kb + 3][lrow] = a0.w;
      As[kb + 4][lrow] = a1.x; As[kb + 5][lrow] = a1.y; As[kb + 6][lrow] = a1.z; As[kb + 7][lrow] = a1.w;
      if (tid < BK * BN / 4) {
        const int kk = tid / (BN / 4), c4 = tid % (BN / 4);
        *reinterpret_cast<float4*>(&Ws[kk][c4 * 4]) =
            *reinterpret_cast<const float4*>(p.W + (long long)(kglob + kk) * p.Nout + col0 + c4 * 4);
      }
      __syncthreads();
#pragma unroll
      for (int kk = 0; kk < BK; ++kk) {
        const float4 bv = *reinterpret_cast<const float4*>(&Ws[kk][tx * 4]);
#pragma unroll
        for (int i = 0; i < RM; ++i) {
          const float a = As[kk][ty * RM + i];
          acc[i][0] = fmaf(a, bv.x, acc[i][0]); acc[i][1] = fmaf(a, bv.y, acc[i][1]);
          acc[i][2] = fmaf(a, bv.z, acc[i][2]); acc[i][3] = fmaf(a, bv.w, acc[i][3]);
        }
      }
    }
  }
  const int col = col0 + tx * 4;
  float4 bb = make_float4(0.f, 0.f, 0.f, 0.f);
  if (p.bias) bb = *reinterpret_cast<const float4*>(p.bias + col);
#pragma unroll
  for (int i = 0; i < RM; ++i) {
    const long long rl = row0 + ty * RM + i;
    if (rl < p.rows) {
      const long long rr = (rl / p.Nn) * p.pitch_d + p.off_d + (rl % p.Nn);
      float4 v;
      v.x = apply_act(acc[i][0] + bb.x, p.act); v.y = apply_act(acc[i][1] + bb.y, p.act);
      v.z = apply_act(acc[i][2] + bb.z, p.act); v.w = apply_act(acc[i][3] + bb.w, p.act);
      *reinterpret_cast<float4*>(p.out + rr * p.ldo + col) = v;
    }
  }
}

// ---------------------------------------------------------------------------
// K2 NodeEdge (emulator.py:27-45): out[s] = (w*|inci| + b) @ in[s], the folded
// dense (R_out x R_in) matrix resident in shared memory (transposed), SPB samples
// per CTA pass, persistent over samples.
// ---------------------------------------------------------------------------
struct NodeEdgeP {
  const float* matT;     // (R_in, ldm) transposed folded matrix, zero padded to ldm
  const void* in;        // (S, R_in, F) storage type T
  void* out;             // (S, R_out, F)
  long long S;
  int R_in, R_out, F, ldm;
};

template <int TI, typename T>
__global__ void __launch_bounds__(512) nodeedge_kernel(NodeEdgeP p) {
  extern __shared__ __align__(16) float sm[];
  float* Ms = sm;                                   // R_in * ldm
  float* Xs = sm + (size_t)p.R_in * p.ldm;          // 2 * R_in * F
  const int tid = threadIdx.x;
  for (int i = tid * 4; i < p.R_in * p.ldm; i += blockDim.x * 4)
    *reinterpret_cast<float4*>(Ms + i) = *reinterpret_cast<const float4*>(p.matT + i);
  const int half = tid >> 8, t = tid & 255;
  const int FG = p.F >> 2;                          // float4 feature groups
  const int fg = t % FG, ig = t / FG;
  const int IG = 256 / FG;
  const int i0 = ig * TI;
  float* Xh = Xs + (size_t)half * p.R_in * p.F;
  for (long long s0 = (long long)blockIdx.x * 2; s0 < p.S; s0 += (long long)gridDim.x * 2) {
    const long long s = s0 + half;
    __syncthreads();
    if (s < p.S) {
      const T* src = static_cast<const T*>(p.in) + s * p.R_in * p.F;
      for (int i = t * 4; i < p.R_in * p.F; i += 256 * 4)
        *reinterpret_cast<float4*>(Xh + i) = ld4(src + i);
    }
    __syncthreads();
    if (s >= p.S || ig >= IG || i0 >= p.R_out) continue;
    float4 acc[TI];
#pragma unroll
    for (int i = 0; i < TI; ++i) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int j = 0; j < p.R_in; ++j) {
      const float4 x = *reinterpret_cast<const float4*>(Xh + j * p.F + fg * 4);
      const float* mrow = Ms + (size_t)j * p.ldm + i0;
#pragma unroll
      for (int i4 = 0; i4 < TI / 4; ++i4) {
        const float4 m = *reinterpret_cast<const float4*>(mrow + i4 * 4);
        const float mm[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          float4& a = acc[i4 * 4 + q];
          a.x = fmaf(mm[q], x.x, a.x); a.y = fmaf(mm[q], x.y, a.y);
          a.z = fmaf(mm[q], x.z, a.z); a.w = fmaf(mm[q], x.w, a.w);
        }
      }
    }
    T* dst = static_cast<T*>(p.out) + s * p.R_out * p.F;
#pragma unroll
    for (int i = 0; i < TI; ++i)
      if (i0 + i < p.R_out) st4(dst + (size_t)(i0 + i) * p.F + fg * 4, acc[i]);
  }
}

// ---------------------------------------------------------------------------
// K4/K5 neighbour aggregation on CSR (spektral GCNConv / GATConv dense mode restated
// sparse): one CTA per (b,t) sample, the sample's h = x@W tile staged in shared memory.
//   GCN: out_i = act( sum_j val_ij h_j + bias )
//   GAT: score_ij = leaky_relu(h_i.a_self + h_j.a_neigh, 0.2) over the nonzero pattern
//        (diagonal forced), alpha = softmax_j, out_i = act( sum_j alpha_ij h_j + bias ).
// The reference's -1e10 mask makes masked entries exactly 0 after exp, so the CSR softmax
// over the pattern is the same function.
// ---------------------------------------------------------------------------
struct AggP {
  const void* h;         // (S, R, E) storage type T
  void* out;             // (S, R, E)
  const int* row_ptr; const int* col_idx; const float* val;   // val null for GAT
  const float* a_self; const float* a_neigh;                  // (E) GAT only
  const float* bias;     // (E)
  long long S; int R, E, act, gat;
  // use_adj (emulator.py:268-271, 343-362), GAT: per-sample attention mask.  Entry q of the pattern belongs to controlled
  // link act_entry[q]-1 (0 = none): it is kept iff trunc(setting) != 0 (`adj.astype(int)` of the adjacency with the
  // setting written into it); other entries are kept iff trunc(adjacency value) != 0 (keep[q]); the diagonal always.
  const int* act_entry; const unsigned char* keep; const float* a_dyn; int n_act;   // a_dyn (S, n_act) or null
};

__device__ __forceinline__ bool agg_kept(const AggP& p, long long s, int q, int i, int j) {
  if (!p.a_dyn || i == j) return true;
  const int k = p.act_entry[q];
  return k ? truncf(p.a_dyn[s * p.n_act + k - 1]) != 0.f : p.keep[q] != 0;
}

template <typename T>
__global__ void __launch_bounds__(1024) agg_small_kernel(AggP p) {      // 256 threads per sample, 1024 when there are few samples (same results: a row is one warp's)
  extern __shared__ __align__(16) float sm[];
  float* hs = sm;                          // R*E
  float* ss = sm + (size_t)p.R * p.E;      // R
  float* sn = ss + p.R;                    // R
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
  const int E = p.E;
  for (long long s = blockIdx.x; s < p.S; s += gridDim.x) {
    const T* src = static_cast<const T*>(p.h) + s * p.R * E;
    __syncthreads();
    for (int i = tid * 4; i < p.R * E; i += blockDim.x * 4)
      *reinterpret_cast<float4*>(hs + i) = ld4(src + i);
    __syncthreads();
    if (p.gat) {
      for (int i = warp; i < p.R; i += nwarp) {
        float d1 = 0.f, d2 = 0.f;
        for (int c = lane * 4; c < E; c += 128) {
          const float4 hv = *reinterpret_cast<const float4*>(hs + (size_t)i * E + c);
          const float4 a1 = *reinterpret_cast<const float4*>(p.a_self + c);
          const float4 a2 = *reinterpret_cast<const float4*>(p.a_neigh + c);
          d1 += hv.x * a1.x + hv.y * a1.y + hv.z * a1.z + hv.w * a1.w;
          d2 += hv.x * a2.x + hv.y * a2.y + hv.z * a2.z + hv.w * a2.w;
        }
        d1 = warp_sum(d1); d2 = warp_sum(d2);
        if (lane == 0) { ss[i] = d1; sn[i] = d2; }
      }
      __syncthreads();
    }
    T* dst = static_cast<T*>(p.out) + s * p.R * E;
    for (int i = warp; i < p.R; i += nwarp) {
      const int beg = p.row_ptr[i], end = p.row_ptr[i + 1];
      float mx = -INFINITY;
      if (p.gat) {
        const float si = ss[i];
        for (int q = beg + lane; q < end; q += 32) {
          const int cj = p.col_idx[q];
          float sc = si + sn[cj];
          sc = sc > 0.f ? sc : 0.2f * sc;
          if (agg_kept(p, s, q, i, cj)) mx = fmaxf(mx, sc);
        }
        mx = warp_max(mx);
      }
      for (int c0 = 0; c0 < E; c0 += 128) {
        const int c = c0 + lane * 4;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        float den = 0.f;
        for (int q0 = beg; q0 < end; q0 += 32) {
          const int q = q0 + lane;
          int cj = 0; float wj = 0.f;
          if (q < end) {
            cj = p.col_idx[q];
            if (p.gat) {
              float sc = ss[i] + sn[cj];
              sc = sc > 0.f ? sc : 0.2f * sc;
              wj = agg_kept(p, s, q, i, cj) ? expf(sc - mx) : 0.f;
            } else {
              wj = p.val[q];
            }
          }
          if (p.gat) den += wj;
          const int cnt = min(32, end - q0);
          for (int u = 0; u < cnt; ++u) {
            const float w = __shfl_sync(0xffffffffu, wj, u);
            const int j = __shfl_sync(0xffffffffu, cj, u);
            if (c < E) {
              const float4 hv = *reinterpret_cast<const float4*>(hs + (size_t)j * E + c);
              acc.x = fmaf(w, hv.x, acc.x); acc.y = fmaf(w, hv.y, acc.y);
              acc.z = fmaf(w, hv.z, acc.z); acc.w = fmaf(w, hv.w, acc.w);
            }
          }
        }
        if (p.gat) {
          den = warp_sum(den);
          const float inv = 1.f / den;
          acc.x *= inv; acc.y *= inv; acc.z *= inv; acc.w *= inv;
        }
        if (c < E) {
          const float4 bb = *reinterpret_cast<const float4*>(p.bias + c);
          float4 v;
          v.x = apply_act(acc.x + bb.x, p.act); v.y = apply_act(acc.y + bb.y, p.act);
          v.z = apply_act(acc.z + bb.z, p.act); v.w = apply_act(acc.w + bb.w, p.act);
          st4(dst + (size_t)i * E + c, v);
        }
      }
    }
  }
}

// Large-graph path: h does not fit on chip; one warp per (sample,row),
// 128-bit gathers of neighbour rows straight from HBM/L2.  Scores for GAT come from
// row_scores_kernel.
struct AggLargeP {
  const void* h; void* out;   // storage type T
  const int* row_ptr; const int* col_idx; const float* val;
  const float* ss; const float* sn;     // (S*R) GAT scores or null
  const float* bias;
  long long S; int R, E, act, gat;
};

template <typename T>
__global__ void __launch_bounds__(256) row_scores_kernel(const T* __restrict__ h, const float* __restrict__ a_self,
                                                         const float* __restrict__ a_neigh, float* __restrict__ ss,
                                                         float* __restrict__ sn, long long rows, int E) {
  const int lane = threadIdx.x & 31;
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nw = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long r = w; r < rows; r += nw) {
    float d1 = 0.f, d2 = 0.f;
    for (int c = lane * 4; c < E; c += 128) {
      const float4 hv = ld4(h + r * E + c);
      const float4 a1 = *reinterpret_cast<const float4*>(a_self + c);
      const float4 a2 = *reinterpret_cast<const float4*>(a_neigh + c);
      d1 += hv.x * a1.x + hv.y * a1.y + hv.z * a1.z + hv.w * a1.w;
      d2 += hv.x * a2.x + hv.y * a2.y + hv.z * a2.z + hv.w * a2.w;
    }
    d1 = warp_sum(d1); d2 = warp_sum(d2);
    if (lane == 0) { ss[r] = d1; sn[r] = d2; }
  }
}

template <typename T>
__global__ void __launch_bounds__(256) agg_large_kernel(AggLargeP p) {
  const int lane = threadIdx.x & 31;
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nw = ((long long)gridDim.x * blockDim.x) >> 5;
  const long long total = p.S * p.R;
  const int E = p.E;
  for (long long g = w; g < total; g += nw) {
    const long long s = g / p.R;
    const int i = (int)(g % p.R);
    const T* hb = static_cast<const T*>(p.h) + s * p.R * E;
    const int beg = p.row_ptr[i], end = p.row_ptr[i + 1];
    float mx = -INFINITY, si = 0.f;
    if (p.gat) {
      si = p.ss[s * p.R + i];
      for (int q = beg + lane; q < end; q += 32) {
        float sc = si + p.sn[s * p.R + p.col_idx[q]];
        sc = sc > 0.f ? sc : 0.2f * sc;
        mx = fmaxf(mx, sc);
      }
      mx = warp_max(mx);
    }
    for (int c0 = 0; c0 < E; c0 += 128) {
      const int c = c0 + lane * 4;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      float den = 0.f;
      for (int q0 = beg; q0 < end; q0 += 32) {
        const int q = q0 + lane;
        int cj = 0; float wj = 0.f;
        if (q < end) {
          cj = p.col_idx[q];
          if (p.gat) {
            float sc = si + p.sn[s * p.R + cj];
            sc = sc > 0.f ? sc : 0.2f * sc;
            wj = expf(sc - mx);
          } else wj = p.val[q];
        }
        if (p.gat) den += wj;
        const int cnt = min(32, end - q0);
        for (int u = 0; u < cnt; ++u) {
          const float wv = __shfl_sync(0xffffffffu, wj, u);
          const int j = __shfl_sync(0xffffffffu, cj, u);
          if (c < E) {
            const float4 hv = ld4(hb + (size_t)j * E + c);
            acc.x = fmaf(wv, hv.x, acc.x); acc.y = fmaf(wv, hv.y, acc.y);
            acc.z = fmaf(wv, hv.z, acc.z); acc.w = fmaf(wv, hv.w, acc.w);
          }
        }
      }
      if (p.gat) {
        den = warp_sum(den);
        const float inv = 1.f / den;
        acc.x *= inv; acc.y *= inv; acc.z *= inv; acc.w *= inv;
      }
      if (c < E) {
        const float4 bb = *reinterpret_cast<const float4*>(p.bias + c);
        float4 v;
        v.x = apply_act(acc.x + bb.x, p.act); v.y = apply_act(acc.y + bb.y, p.act);
        v.z = apply_act(acc.z + bb.z, p.act); v.w = apply_act(acc.w + bb.w, p.act);
        st4(static_cast<T*>(p.out) + (s * p.R + i) * E + c, v);
      }
    }
  }
}

// 16-bit storage, E == 128 (the C5 shape).  LPR lanes per row, a lane owns 8 channels (one 16-byte load per neighbour row
// and lane; the one-warp-per-row kernel above moves 8 bytes per lane), the neighbour loads of a row are issued four at a
// time before their FMAs.  LPR = 16: the whole 256-byte row per pass.  LPR = 8: TWO passes over the grid, one per
// 128-byte half of the rows, so that the gather's working set per sample is half the h matrix (25.6 MB at C5 instead of
// 51 MB - ncu showed 9.5 % L2 hits with the full rows).  Same order of the per-channel FMAs as agg_large_kernel (bitwise
// equal for GCN).
template <int LPR, bool GAT, int MINB>
__global__ void __launch_bounds__(256, MINB) agg_large16_kernel(AggLargeP p) {
  constexpr int E = 128, RPW = 32 / LPR, NPASS = 16 / LPR;
  const int lane = threadIdx.x & 31, hl = lane % LPR, sub = lane / LPR;
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nw = ((long long)gridDim.x * blockDim.x) >> 5;
  const long long total = p.S * p.R;
  for (int pass = 0; pass < NPASS; ++pass) {
    const int c0 = pass * LPR * 8 + hl * 8;                   // this lane's 8 channels
    const float4 b0 = *reinterpret_cast<const float4*>(p.bias + c0), b1 = *reinterpret_cast<const float4*>(p.bias + c0 + 4);
    for (long long gw = w; gw * RPW < total; gw += nw) {        // warp-uniform trip count
      const long long g = gw * RPW + sub;
      const bool live = g < total;
      const long long s = live ? g / p.R : 0;
      const int i = live ? (int)(g % p.R) : 0;
      const bf16* hb = static_cast<const bf16*>(p.h) + s * p.R * E;
      const int beg = live ? __ldg(p.row_ptr + i) : 0, end = live ? __ldg(p.row_ptr + i + 1) : 0;
      float mx = -INFINITY, si = 0.f;
      if (GAT) {
        if (live) si = p.ss[s * p.R + i];
        for (int q = beg + hl; q < end; q += LPR) {
          float sc = si + p.sn[s * p.R + p.col_idx[q]];
          sc = sc > 0.f ? sc : 0.2f * sc;
          mx = fmaxf(mx, sc);
        }
#pragma unroll
        for (int o = LPR / 2; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o, LPR));
      }
      float acc[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] = 0.f;
      float den = 0.f;
      const int len = end - beg;
      int len_w = len;                                          // every row group of the warp walks the longest row
#pragma unroll
      for (int o = LPR; o < 32; o <<= 1) len_w = max(len_w, __shfl_xor_sync(0xffffffffu, len_w, o));
      for (int q0 = 0; q0 < len_w; q0 += LPR) {
        const int q = beg + q0 + hl;
        int cj = 0; float wj = 0.f;
        if (q < end) {
          cj = __ldg(p.col_idx + q);
          if (GAT) {
            float sc = si + p.sn[s * p.R + cj];
            sc = sc > 0.f ? sc : 0.2f * sc;
            wj = expf(sc - mx);
          } else wj = __ldg(p.val + q);
        }
        if (GAT) den += wj;
        const int cnt_w = min(LPR, len_w - q0);
        for (int u0 = 0; u0 < cnt_w; u0 += 4) {
          uint4 raw[4]; float wv[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int src = min(u0 + u, LPR - 1);
            wv[u] = __shfl_sync(0xffffffffu, wj, src, LPR);
            const int j = __shfl_sync(0xffffffffu, cj, src, LPR);
            const bool ok = u0 + u < LPR && q0 + u0 + u < len;
            raw[u] = ok ? *reinterpret_cast<const uint4*>(hb + (size_t)j * E + c0) : make_uint4(0u, 0u, 0u, 0u);
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            if (u0 + u < LPR && q0 + u0 + u < len) {           // keep the FMA sequence of agg_large_kernel exactly
              const float2 f0 = unpack2_16(raw[u].x), f1 = unpack2_16(raw[u].y), f2 = unpack2_16(raw[u].z), f3 = unpack2_16(raw[u].w);
              acc[0] = fmaf(wv[u], f0.x, acc[0]); acc[1] = fmaf(wv[u], f0.y, acc[1]); acc[2] = fmaf(wv[u], f1.x, acc[2]); acc[3] = fmaf(wv[u], f1.y, acc[3]);
              acc[4] = fmaf(wv[u], f2.x, acc[4]); acc[5] = fmaf(wv[u], f2.y, acc[5]); acc[6] = fmaf(wv[u], f3.x, acc[6]); acc[7] = fmaf(wv[u], f3.y, acc[7]);
            }
          }
        }
      }
      if (GAT) {
#pragma unroll
        for (int o = LPR / 2; o > 0; o >>= 1) den += __shfl_xor_sync(0xffffffffu, den, o, LPR);
        const float inv = 1.f / den;
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] *= inv;
      }
      if (live) {
        float v[8] = {acc[0] + b0.x, acc[1] + b0.y, acc[2] + b0.z, acc[3] + b0.w, acc[4] + b1.x, acc[5] + b1.y, acc[6] + b1.z, acc[7] + b1.w};
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = apply_act(v[k], p.act);
        const uint4 o4 = make_uint4(pack2_bf16(v[0], v[1]), pack2_bf16(v[2], v[3]), pack2_bf16(v[4], v[5]), pack2_bf16(v[6], v[7]));
        bf16* dst = static_cast<bf16*>(p.out) + (s * p.R + i) * E + c0;
        *reinterpret_cast<uint4*>(dst) = o4;
      }
    }
  }
}

// ---------------------------------------------------------------------------
// K6 GRU recurrence (keras GRU v2, reset_after=True): mx = x_t@K + b[0] comes from
// rows_gemm; here h_t = z*h + (1-z)*tanh(mx_h + r*(h@R + b[1])_h) walks t in-kernel.
// One CTA owns 32 (b,n) sequences; R (H x 3H) lives in shared memory.
// ---------------------------------------------------------------------------
struct GruP {
  const void* mx;        // (Bc, Ts, Nn, 3H) storage type T
  const float* R;        // (H, 3H)
  const float* b1;       // (3H) recurrent bias
  void* out;             // (Bc, Td, Nn, H) last Td steps
  long long nseq;        // Bc*Nn
  int Nn, Ts, Td, H;
  float* hall;           // reverse-pass tape: the hidden state of EVERY step (Bc, Ts, Nn, H), or null
};

template <typename T>
__global__ void __launch_bounds__(256) gru_kernel(GruP p) {
  extern __shared__ __align__(16) float sm[];
  const int H = p.H, G = 3 * H;
  float* Rs = sm;                    // H*G
  float* hs = Rs + (size_t)H * G;    // 32*H
  float* mh = hs + 32 * H;           // 32*G
  const int tid = threadIdx.x;
  for (int i = tid * 4; i < H * G; i += 256 * 4)
    *reinterpret_cast<float4*>(Rs + i) = *reinterpret_cast<const float4*>(p.R + i);
  for (int i = tid; i < 32 * H; i += 256) hs[i] = 0.f;
  const int rr = tid >> 3, gg = tid & 7;             // 32 sequences x 8 column groups
  const long long q = (long long)blockIdx.x * 32 + rr;
  const bool live = q < p.nseq;
  const long long b = live ? q / p.Nn : 0;
  const int n = live ? (int)(q % p.Nn) : 0;
  const int ncol4 = G / 32;                          // float4 column groups per thread (G=192 -> 6)
  __syncthreads();
  for (int t = 0; t < p.Ts; ++t) {
    // mh[rr][:] = hs[rr][:] @ R + b1
    for (int qq = 0; qq < ncol4; ++qq) {
      const int col = qq * 32 + gg * 4;
      float4 acc = *reinterpret_cast<const float4*>(p.b1 + col);
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int k = 0; k < H; ++k) {
        const float hv = hs[rr * H + k];
        const float4 rv = *reinterpret_cast<const float4*>(Rs + (size_t)k * G + col);
        s.x = fmaf(hv, rv.x, s.x); s.y = fmaf(hv, rv.y, s.y);
        s.z = fmaf(hv, rv.z, s.z); s.w = fmaf(hv, rv.w, s.w);
      }
      acc.x += s.x; acc.y += s.y; acc.z += s.z; acc.w += s.w;
      *reinterpret_cast<float4*>(mh + rr * G + col) = acc;
    }
    __syncthreads();
    if (live) {
      const T* mxr = static_cast<const T*>(p.mx) + ((b * p.Ts + t) * p.Nn + n) * G;
      for (int u = gg; u < H; u += 8) {
        const float z = 1.f / (1.f + expf(-(ld1(mxr + u) + mh[rr * G + u])));
        const float r = 1.f / (1.f + expf(-(ld1(mxr + H + u) + mh[rr * G + H + u])));
        const float hh = tanhf(ld1(mxr + 2 * H + u) + r * mh[rr * G + 2 * H + u]);
        const float hn = z * hs[rr * H + u] + (1.f - z) * hh;
        hs[rr * H + u] = hn;
        if (p.hall) p.hall[((b * p.Ts + t) * p.Nn + n) * H + u] = hn;
        const int td = t - (p.Ts - p.Td);
        if (td >= 0) st1(static_cast<T*>(p.out) + ((b * p.Td + td) * p.Nn + n) * H + u, hn);
      }
    }
    __syncthreads();
  }
}

// K6 LSTM recurrence (keras LSTM v2; gate order i,f,c,o): zx = x_t@K + b from rows_gemm.
struct LstmP {
  const void* zx; const float* R; void* out; long long nseq; int Nn, Ts, Td, H;   // zx/out storage type T
  float* hall; float* call;   // reverse-pass tape: hidden and cell state of EVERY step (Bc, Ts, Nn, H), or null
};

template <typename T>
__global__ void __launch_bounds__(256) lstm_kernel(LstmP p) {
  extern __shared__ __align__(16) float sm[];
  const int H = p.H, G = 4 * H;
  float* Rs = sm;                    // H*G
  float* hs = Rs + (size_t)H * G;    // 32*H
  float* cs = hs + 32 * H;           // 32*H
  float* mh = cs + 32 * H;           // 32*G
  const int tid = threadIdx.x;
  for (int i = tid * 4; i < H * G; i += 256 * 4)
    *reinterpret_cast<float4*>(Rs + i) = *reinterpret_cast<const float4*>(p.R + i);
  for (int i = tid; i < 32 * H; i += 256) { hs[i] = 0.f; cs[i] = 0.f; }
  const int rr = tid >> 3, gg = tid & 7;
  const long long q = (long long)blockIdx.x * 32 + rr;
  const bool live = q < p.nseq;
  const long long b = live ? q / p.Nn : 0;
  const int n = live ? (int)(q % p.Nn) : 0;
  const int ncol4 = G / 32;
  __syncthreads();
  for (int t = 0; t < p.Ts; ++t) {
    for (int qq = 0; qq < ncol4; ++qq) {
      const int col = qq * 32 + gg * 4;
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int k = 0; k < H; ++k) {
        const float hv = hs[rr * H + k];
        const float4 rv = *reinterpret_cast<const float4*>(Rs + (size_t)k * G + col);
        s.x = fmaf(hv, rv.x, s.x); s.y = fmaf(hv, rv.y, s.y);
        s.z = fmaf(hv, rv.z, s.z); s.w = fmaf(hv, rv.w, s.w);
      }
      *reinterpret_cast<float4*>(mh + rr * G + col) = s;
    }
    __syncthreads();
    if (live) {
      const T* zr = static_cast<const T*>(p.zx) + ((b * p.Ts + t) * p.Nn + n) * G;
      for (int u = gg; u < H; u += 8) {
        const float ig = 1.f / (1.f + expf(-(ld1(zr + u) + mh[rr * G + u])));
        const float fg = 1.f / (1.f + expf(-(ld1(zr + H + u) + mh[rr * G + H + u])));
        const float cg = tanhf(ld1(zr + 2 * H + u) + mh[rr * G + 2 * H + u]);
        const float og = 1.f / (1.f + expf(-(ld1(zr + 3 * H + u) + mh[rr * G + 3 * H + u])));
        const float cn = fg * cs[rr * H + u] + ig * cg;
        const float hn = og * tanhf(cn);
        cs[rr * H + u] = cn; hs[rr * H + u] = hn;
        if (p.hall) { p.hall[((b * p.Ts + t) * p.Nn + n) * H + u] = hn; p.call[((b * p.Ts + t) * p.Nn + n) * H + u] = cn; }
        const int td = t - (p.Ts - p.Td);
        if (td >= 0) st1(static_cast<T*>(p.out) + ((b * p.Td + td) * p.Nn + n) * H + u, hn);
      }
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------
// K8 resnet: x = act(cumsum_t(x_out) + res) in place (emulator.py:315-320).
// ---------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) cumsum_res_act_kernel(T* __restrict__ x, const float* __restrict__ res,
                                                             long long Bc, int Tn, int Nn, int E, int act) {
  const int e4 = E >> 2;
  const long long total = Bc * Nn * e4;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < total;
       g += (long long)gridDim.x * blockDim.x) {
    const int c4 = (int)(g % e4);
    const long long bn = g / e4;
    const int n = (int)(bn % Nn);
    const long long b = bn / Nn;
    const float4 r = *reinterpret_cast<const float4*>(res + bn * E + c4 * 4);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int t = 0; t < Tn; ++t) {
      T* ptr = x + ((b * Tn + t) * Nn + n) * E + c4 * 4;
      const float4 v = ld4(ptr);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      float4 o;
      o.x = apply_act(acc.x + r.x, act); o.y = apply_act(acc.y + r.y, act);
      o.z = apply_act(acc.z + r.z, act); o.w = apply_act(acc.w + r.w, act);
      st4(ptr, o);
    }
  }
}

// K8 resnet + narrow head fused (tensor-core path, E == 128): one warp per (b,n) series, a lane owns 4 channels.
//   x_t = act(cumsum_t(x_out) + res)          (written back as bf16 only when write_x)
//   out[(b,t,n), col0 + o] = hact( x_t . W[:, o] + bias[o] ),  o < Nout <= NOUT
// The head consumes the bf16-rounded x_t, i.e. the same values a separate head kernel would read back.
// TB time steps are loaded ahead; their TB*NOUT partial dot products are reduced across the warp by
// recursive halving (NV-1+log terms instead of 5 shuffles per value).
struct CumHeadP {
  bf16* x; const float* res; long long Bc; int Tn, Nn, act, write_x;
  const float* W; const float* bias; float* out; int ldo, col0, Nout, hact;
};

// NV (power of two, <= 32) per-lane values -> lane l holds the warp total of value l / (32 / NV) in v[0]
template <int NV>
__device__ __forceinline__ void warp_reduce_spread(float* v, int lane) {
  int n = NV;
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    if (n > 1) {
      n >>= 1;
      const bool up = (lane & s) != 0;
#pragma unroll
      for (int i = 0; i < NV / 2; ++i) {
        if (i < n) {
          const float keep = up ? v[i + n] : v[i];
          const float send = up ? v[i] : v[i + n];
          v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
        }
      }
    } else {
      v[0] += __shfl_xor_sync(0xffffffffu, v[0], s);
    }
  }
}

template <int NOUT, int TB>
__global__ void __launch_bounds__(256, 4) cumsum_head_bf16_kernel(CumHeadP p) {
  constexpr int E = 128, NV = NOUT * TB, DUP = 32 / NV;
  static_assert(NV <= 32 && (NV & (NV - 1)) == 0, "NOUT*TB must be a power of two <= 32");
  const int lane = threadIdx.x & 31;
  const long long gw = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nw = ((long long)gridDim.x * blockDim.x) >> 5;
  float w[4][NOUT];
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int o = 0; o < NOUT; ++o) w[j][o] = o < p.Nout ? p.W[(lane * 4 + j) * p.Nout + o] : 0.f;
  const int my = lane / DUP, my_u = my / NOUT, my_o = my % NOUT;      // the (time step, column) this lane finalises
  const bool writer = (lane % DUP) == 0 && my_o < p.Nout;
  const float my_b = my_o < p.Nout ? p.bias[my_o] : 0.f;
  const size_t tstride = (size_t)p.Nn * E;
  for (long long bn = gw; bn < p.Bc * p.Nn; bn += nw) {
    const int n = (int)(bn % p.Nn);
    const long long b = bn / p.Nn;
    const float4 r = *reinterpret_cast<const float4*>(p.res + bn * E + lane * 4);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    bf16* base = p.x + ((size_t)b * p.Tn * p.Nn + n) * E + lane * 4;
    float* obase = p.out + ((size_t)b * p.Tn * p.Nn + n) * p.ldo + p.col0;
    for (int t0 = 0; t0 < p.Tn; t0 += TB) {
      uint2 raw[TB];
#pragma unroll
      for (int u = 0; u < TB; ++u)
        raw[u] = t0 + u < p.Tn ? *reinterpret_cast<const uint2*>(base + (size_t)(t0 + u) * tstride) : make_uint2(0u, 0u);
      float d[NV];
#pragma unroll
      for (int u = 0; u < TB; ++u) {
        { const float2 f0 = unpack2_16(raw[u].x), f1 = unpack2_16(raw[u].y); acc.x += f0.x; acc.y += f0.y; acc.z += f1.x; acc.w += f1.y; }
        float v[4] = {acc.x + r.x, acc.y + r.y, acc.z + r.z, acc.w + r.w};
        act_tile4(v, p.act);
        const uint2 pk = make_uint2(pack2_bf16(v[0], v[1]), pack2_bf16(v[2], v[3]));
        if (p.write_x && t0 + u < p.Tn) *reinterpret_cast<uint2*>(base + (size_t)(t0 + u) * tstride) = pk;
        const float2 qa = unpack2_16(pk.x), qb = unpack2_16(pk.y);
        const float q0 = qa.x, q1 = qa.y, q2 = qb.x, q3 = qb.y;
#pragma unroll
        for (int o = 0; o < NOUT; ++o) d[u * NOUT + o] = fmaf(q0, w[0][o], fmaf(q1, w[1][o], fmaf(q2, w[2][o], q3 * w[3][o])));
      }
      warp_reduce_spread<NV>(d, lane);
      if (writer && t0 + my_u < p.Tn) obase[(size_t)(t0 + my_u) * p.Nn * p.ldo + my_o] = apply_act(d[0] + my_b, p.hact);
    }
  }
}

// K8 narrow heads (n_out<=4 columns): one warp per row.
struct HeadP {
  const void* in; int ld; int K;   // storage type T
  const float* W;       // (K, Nout)
  const float* bias;
  float* out; int ldo; int col0; int Nout; int act;
  long long rows;
};

template <typename T>
__global__ void __launch_bounds__(256) head_kernel(HeadP p) {
  extern __shared__ float sm[];
  float* Ws = sm;
  for (int i = threadIdx.x; i < p.K * p.Nout; i += blockDim.x) Ws[i] = p.W[i];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nw = ((long long)gridDim.x * blockDim.x) >> 5;
  if (p.K % 128 == 0 && p.K <= 256 && p.ld % 4 == 0) {
    // fast path: a lane owns 4 consecutive input columns per 128-column block (one 8/16-byte load per row),
    // its weight rows live in registers, two rows per iteration; one butterfly reduction per output.
    const int nb = p.K >> 7;
    float wreg[2][4][4];
#pragma unroll
    for (int b = 0; b < 2; ++b)
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int o = 0; o < 4; ++o) wreg[b][j][o] = (b < nb && o < p.Nout) ? Ws[(b * 128 + lane * 4 + j) * p.Nout + o] : 0.f;
    float bo[4];
#pragma unroll
    for (int o = 0; o < 4; ++o) bo[o] = o < p.Nout ? p.bias[o] : 0.f;
    for (long long r = w * 2; r < p.rows; r += nw * 2) {
      float acc[2][4];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
#pragma unroll
        for (int o = 0; o < 4; ++o) acc[u][o] = 0.f;
        if (r + u < p.rows) {
#pragma unroll
          for (int b = 0; b < 2; ++b) {
            if (b < nb) {
              const float4 v = ld4(static_cast<const T*>(p.in) + (r + u) * p.ld + b * 128 + lane * 4);
#pragma unroll
              for (int o = 0; o < 4; ++o)
                acc[u][o] = fmaf(v.x, wreg[b][0][o], fmaf(v.y, wreg[b][1][o], fmaf(v.z, wreg[b][2][o], fmaf(v.w, wreg[b][3][o], acc[u][o]))));
            }
          }
        }
      }
#pragma unroll
      for (int u = 0; u < 2; ++u)
#pragma unroll
        for (int o = 0; o < 4; ++o)
          if (o < p.Nout) acc[u][o] = warp_sum(acc[u][o]);
      if (lane < 2 && r + lane < p.rows) {
#pragma unroll
        for (int o = 0; o < 4; ++o)
          if (o < p.Nout) p.out[(r + lane) * p.ldo + p.col0 + o] = apply_act((lane ? acc[1][o] : acc[0][o]) + bo[o], p.act);
      }
    }
    return;
  }
  for (long long r = w; r < p.rows; r += nw) {
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int k = lane; k < p.K; k += 32) {
      const float v = ld1(static_cast<const T*>(p.in) + r * p.ld + k);
      for (int o = 0; o < p.Nout; ++o) acc[o] = fmaf(v, Ws[k * p.Nout + o], acc[o]);
    }
    for (int o = 0; o < p.Nout; ++o) {
      const float s = warp_sum(acc[o]);
      if (lane == 0) p.out[r * p.ldo + p.col0 + o] = apply_act(s + p.bias[o], p.act);
    }
  }
}

// ---------------------------------------------------------------------------
// P1..P4 fused epilogue: post_proc (emulator.py:643-678), de-normalisation + depth clip
// (emulator.py:586-588) and constrain (emulator.py:727-748) for pump-free, offset-free
// networks.  One CTA per (b,t) sample.
// ---------------------------------------------------------------------------
struct PostP {
  const float* out_n;     // (S, N, ldn): [head cols..., flood prob] normalised
  int ldn;
  const float* out_e;     // (S, M, lde) tanh head
  int lde;
  const float* b_raw;     // (S, N, b_in) physical boundary
  const float* a;         // (S, n_act) or null
  int n_act;
  const int* act_of_edge; const int* act_out_of_node; const int* act_in_of_node;
  const int* inc_ptr; const int* inc_edge; const int* inc_sign;   // node -> incident edges
  const float* ymax; const float* ymin; int cy;
  const float* emax; const float* emin; int ce;
  const float* bmax; const float* bmin; int cb;
  const float* hmin; const float* hmax; const float* is_outfall; const float* ehmax;
  float* y; float* ey;    // (S, N, ny), (S, M, e_out)
  long long S;
  int N, M, b_in, n_out, e_out, ny;
  int if_flood, edge_fusion, act, tide, normalised;
  float epsilon;
  float* efl_g;           // (S, M) scratch for the physical edge flows when M floats exceed shared memory, else null
  // pumped / offset networks (all null on pump-free, offset-free ones)
  const int* edge_from;   // (M) FromNode of a link: np.matmul(v, clip(node_edge,0,1))[m] = v[from(m)]
  const float* pump; const float* offset;        // (M)
  const float* pump_in; const float* pump_out;   // (N)
  int tf_rules;           // post_proc_tf (emulator.py:680-725): node-side pump gates use h > 0 instead of h > 0.01; the host
                          // passes pump == null unless every link is a pump (emulator.py:698)
  float* hn_g;            // (S, N) scratch for the normalised node depth when the edge pass needs it and N floats exceed smem
};

__global__ void __launch_bounds__(512) postproc_kernel(PostP p) {
  extern __shared__ float efl_s[];   // M physical edge flows [+ N normalised node depths on pumped/offset networks]
  const bool need_h = p.pump || p.offset;
  for (long long s = blockIdx.x; s < p.S; s += gridDim.x) {
    float* efl = p.efl_g ? p.efl_g + s * p.M : efl_s;
    float* hn = p.hn_g ? p.hn_g + s * p.N : efl_s + p.M;
    __syncthreads();
    if (need_h) {                                                // node depth (tide applied) read by the edge pass
      for (int n = threadIdx.x; n < p.N; n += blockDim.x) {
        float hN = p.out_n[(s * p.N + n) * p.ldn];
        if (p.tide) {
          const float braw = p.b_raw[(s * p.N + n) * p.b_in + p.b_in - 1];
          const float mx = p.bmax[n * p.cb + p.b_in - 1], mn = p.bmin[n * p.cb + p.b_in - 1];
          hN = hN * (1.f - p.is_outfall[n]) + (braw - mn) / (mx - mn);
        }
        hn[n] = hN;
      }
      __syncthreads();
    }
    for (int m = threadIdx.x; m < p.M; m += blockDim.x) {
      const float* src = p.out_e + (s * p.M + m) * p.lde;
      float v[4];
      for (int c = 0; c < p.e_out; ++c) v[c] = src[c];
      if (p.offset && p.offset[m] > 0.f) {                       // emulator.py:649-653: inlet offset gate on positive flow
        const int fn = p.edge_from[m];
        const float mx = p.ymax[fn * p.cy], mn = p.ymin[fn * p.cy];
        const float inoff = (hn[fn] * (mx - mn) + mn) - p.hmin[fn];
        float& f = v[p.e_out - 1];
        if (f > 0.f && !(inoff > p.offset[m])) f = 0.f;
      }
      if (p.act) {
        if (p.pump && p.pump[m] != 0.f) {                        // emulator.py:656-659: rated flow while the inlet tank holds water
          const float mx = p.emax[m * p.ce + p.e_out - 1];
          const float fl = p.pump[m] * (hn[p.edge_from[m]] > 0.01f ? 1.f : 0.f) * (mx > 1e-3f ? 1.f / mx : 0.f);
          if (fl != 0.f) v[p.e_out - 1] = fl;
        }
        const int k = p.act_of_edge[m];
        const float ae = (k >= 0) ? p.a[s * p.n_act + k] : 1.f;
        v[p.e_out - 1] *= ae;                                    // emulator.py:660
      }
      float phys[4];
      for (int c = 0; c < p.e_out; ++c) {
        const float mx = p.emax[m * p.ce + c], mn = p.emin[m * p.ce + c];
        phys[c] = v[c] * (mx - mn) + mn;
      }
      efl[m] = phys[p.e_out - 1];
      float* dst = p.ey + (s * p.M + m) * p.e_out;
      if (p.normalised) {
        for (int c = 0; c < p.e_out; ++c) dst[c] = v[c];
      } else {
        phys[0] = fminf(fmaxf(phys[0], 0.f), p.ehmax[m]);        // emulator.py:588
        for (int c = 0; c < p.e_out; ++c) dst[c] = phys[c];
      }
    }
    __syncthreads();
    for (int n = threadIdx.x; n < p.N; n += blockDim.x) {
      const float* src = p.out_n + (s * p.N + n) * p.ldn;
      float yn[4];
      float hN = src[0];
      if (p.tide) {                                              // emulator.py:645-647
        const float braw = p.b_raw[(s * p.N + n) * p.b_in + p.b_in - 1];
        const float mx = p.bmax[n * p.cb + p.b_in - 1], mn = p.bmin[n * p.cb + p.b_in - 1];
        hN = hN * (1.f - p.is_outfall[n]) + (braw - mn) / (mx - mn);
      }
      yn[0] = hN;
      if (p.edge_fusion) {                                       // emulator.py:671-677
        float op = 0.f, on = 0.f, ip = 0.f, in_ = 0.f;
        for (int q = p.inc_ptr[n]; q < p.inc_ptr[n + 1]; ++q) {
          const float f = efl[p.inc_edge[q]];
          const float pos = fmaxf(f, 0.f), neg = fmaxf(-f, 0.f);
          if (p.inc_sign[q] > 0) { op += pos; in_ += neg; } else { ip += pos; on += neg; }
        }
        const float m2 = p.ymax[n * p.cy + 2], m1 = p.ymax[n * p.cy + 1];
        yn[2] = (op + on) * (m2 > 1e-3f ? 1.f / m2 : 0.f);
        yn[1] = (ip + in_) * (m1 > 1e-3f ? 1.f / m1 : 0.f);
      } else {
        yn[1] = src[1]; yn[2] = src[2];
        if (p.act) {                                             // emulator.py:661-670
          if (p.pump_in) {
            const float on = hN > (p.tf_rules ? 0.f : 0.01f) ? 1.f : 0.f;
            const float fli = p.pump_in[n] * on / p.ymax[n * p.cy + 1], flo = p.pump_out[n] * on / p.ymax[n * p.cy + 2];
            if (fli != 0.f) yn[1] = fli;
            if (flo != 0.f) yn[2] = flo;
          }
          const int ko = p.act_out_of_node[n], ki = p.act_in_of_node[n];
          if (ko >= 0) yn[2] *= p.a[s * p.n_act + ko];
          if (ki >= 0) yn[1] *= p.a[s * p.n_act + ki];
        }
      }
      const float prob = p.if_flood ? src[p.n_out] : 0.f;
      float* dst = p.y + (s * p.N + n) * p.ny;
      if (p.normalised) {
        dst[0] = yn[0]; dst[1] = yn[1]; dst[2] = yn[2];
        if (p.if_flood) dst[3] = prob;
        continue;
      }
      float yp[3];
      for (int c = 0; c < 3; ++c) {
        const float mx = p.ymax[n * p.cy + c], mn = p.ymin[n * p.cy + c];
        yp[c] = yn[c] * (mx - mn) + mn;
      }
      float pr = prob;
      if (p.if_flood) {
        const float mx = p.ymax[n * p.cy + 3], mn = p.ymin[n * p.cy + 3];
        pr = prob * (mx - mn) + mn;
      }
      // constrain (emulator.py:727-748)
      float h = fminf(fmaxf(yp[0], p.hmin[n]), p.hmax[n]);
      const float r = p.b_raw[(s * p.N + n) * p.b_in];
      float qw = fmaxf(yp[1] + r - yp[2], 0.f) * (1.f - p.is_outfall[n]);
      bool f = false;
      if (p.if_flood) {
        f = pr > 0.5f;
        h = f ? p.hmax[n] : h;
      }
      if (p.epsilon > 0.f) qw *= ((p.hmax[n] - h) < p.epsilon) ? 1.f : 0.f;
      else if (p.epsilon == 0.f) {}
      else if (p.if_flood) qw *= f ? 1.f : 0.f;
      dst[0] = h; dst[1] = yp[1]; dst[2] = yp[2];
      if (p.if_flood) { dst[3] = pr; dst[4] = qw; } else dst[3] = qw;
    }
  }
}

// P3 pumped-storage depth (emulator.py:590-598) for the nodes `ps` (storage nodes that feed a pump): the depth is
// integrated from the predicted flows, de_0 = clip(h_in[last]), de_t = clip(de_{t-1} + (q_us - q_ds)_t / (area + 1e-6)),
// and replaces h BEFORE constrain, so the h-dependent parts of constrain (emulator.py:727-748) are redone here.
struct PumpStoreP {
  float* y; int ny;             // (Bc, To, N, ny) physical, already constrained by postproc_kernel
  const float* x; int n_in;     // (Bc, Ti, N, n_in) physical input states
  const float* b_raw; int b_in;
  const int* ps; int n_ps;      // node indices
  const float* area; const float* hmin; const float* hmax; const float* is_outfall;
  long long Bc; int Ti, To, N, if_flood; float epsilon;
};

__global__ void __launch_bounds__(128) pump_storage_kernel(PumpStoreP p) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= p.Bc * p.n_ps) return;
  const long long b = i / p.n_ps;
  const int n = p.ps[i % p.n_ps];
  const float lo = p.hmin[n], hi = p.hmax[n], ar = p.area[n] + 1e-6f;
  float de = 0.f;
  for (int t = 0; t < p.To; ++t) {
    float* yy = p.y + ((b * p.To + t) * p.N + n) * p.ny;
    const float v = t == 0 ? p.x[((b * p.Ti + p.Ti - 1) * p.N + n) * p.n_in] : de + (yy[1] - yy[2]) / ar;
    de = fminf(fmaxf(v, lo), hi);
    float h = de;                                               // clip(h, hmin, hmax) of constrain is the identity on de
    bool f = false;
    if (p.if_flood) { f = yy[3] > 0.5f; h = f ? hi : h; }
    yy[0] = h;
    if (p.epsilon > 0.f) {                                      // q_w gate depends on h: recompute
      const float r = p.b_raw[((b * p.To + t) * p.N + n) * p.b_in];
      float qw = fmaxf(yy[1] + r - yy[2], 0.f) * (1.f - p.is_outfall[n]);
      qw *= ((hi - h) < p.epsilon) ? 1.f : 0.f;
      yy[p.ny - 1] = qw;
    }
  }
}

// ---------------------------------------------------------------------------
// Feedback between autoregressive chunks, on the device.
//   MPC  (mpc.py:573-579):  x' = [h,q_us,q_ds,(q_w>0),r]      ex' = [ey, edge_action]
//   MBRL (mbrl.py:338-343): x' = [h,q_us,q_ds,(prob>0.5),r]   ex' = [ey, edge_action]
//   NORM (emulator.py:416-423): normalised y, flag = prob>0.5, b already normalised by the caller
// ---------------------------------------------------------------------------
struct FeedP {
  const float* y; int ny;       // (S,N,ny)
  const float* ey; int e_out;   // (S,M,e_out)
  const float* b; int b_in;     // (S,N,b_in) raw boundary of the chunk just predicted
  const float* a; int n_act; const int* act_of_edge;
  float* x; int n_in;           // (S,N,n_in)
  float* ex; int e_in;          // (S,M,e_in)
  long long S; int N, M, if_flood, rule;
};

__global__ void __launch_bounds__(256) feedback_kernel(FeedP p) {
  const long long tn = p.S * p.N, tm = p.S * p.M;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < tn + tm;
       g += (long long)gridDim.x * blockDim.x) {
    if (g < tn) {
      const float* ys = p.y + g * p.ny;
      float* xd = p.x + g * p.n_in;
      xd[0] = ys[0]; xd[1] = ys[1]; xd[2] = ys[2];
      int c = 3;
      if (p.if_flood) {
        float flag;
        if (p.rule == 1) flag = ys[p.ny - 1] > 0.f ? 1.f : 0.f;      // q_w > 0
        else flag = ys[3] > 0.5f ? 1.f : 0.f;                        // flood prob > 0.5
        xd[c++] = flag;
      }
      for (int k = 0; k < p.b_in && c < p.n_in; ++k, ++c) xd[c] = p.b[g * p.b_in + k];
    } else {
      const long long ge = g - tn;
      const int m = (int)(ge % p.M);
      const long long s = ge / p.M;
      const float* es = p.ey + ge * p.e_out;
      float* ed = p.ex + ge * p.e_in;
      for (int c = 0; c < p.e_out; ++c) ed[c] = es[c];
      float ae = 1.f;
      if (p.a) {
        const int k = p.act_of_edge[m];
        if (k >= 0) ae = p.a[s * p.n_act + k];
      }
      if (p.e_in > p.e_out) ed[p.e_out] = ae;
    }
  }
}

// gather/scatter of time slices: dst[b, t, :] = src[b, t0 + t, :] for row length L
__global__ void __launch_bounds__(256) slice_time_kernel(const float* __restrict__ src, float* __restrict__ dst,
                                                         long long B, int Tsrc, int t0, int Tdst, long long L,
                                                         int scatter) {
  const long long total = B * Tdst * L;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < total;
       g += (long long)gridDim.x * blockDim.x) {
    const long long l = g % L;
    const long long bt = g / L;
    const int t = (int)(bt % Tdst);
    const long long b = bt / Tdst;
    const long long big = (b * Tsrc + t0 + t) * L + l;
    if (scatter) dst[big] = src[g]; else dst[g] = src[big];
  }
}

// R5 astlingen objective (envs/scenario/astlingen.py:75-99), one CTA per candidate.
struct ObjP {
  const float* y; int ny;     // (B,T,N,ny): q_w = last channel, q_in = channel 1
  const float* q_in0;         // (B,N)
  const float* flood_w; const float* wwtp_w; const float* rough_w; const float* gamma;
  float* obj; long long B; int T, N;
};

__global__ void __launch_bounds__(128) objective_kernel(ObjP p) {
  __shared__ float red[4];
  for (long long b = blockIdx.x; b < p.B; b += gridDim.x) {
    float acc = 0.f;
    for (int g = threadIdx.x; g < p.T * p.N; g += blockDim.x) {
      const int t = g / p.N, n = g % p.N;
      const float* ys = p.y + ((b * p.T + t) * p.N + n) * p.ny;
      const float gm = p.gamma ? p.gamma[t] : 1.f;
      float v = ys[p.ny - 1] * p.flood_w[n] + ys[1] * p.wwtp_w[n];
      if (p.rough_w) {
        const float prev = (t == 0) ? p.q_in0[b * p.N + n] : (ys - (long long)p.N * p.ny)[1];
        v += fabsf(ys[1] - prev) * p.rough_w[n];
      }
      acc += v * gm;
    }
    acc = warp_sum(acc);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) p.obj[b] = red[0] + red[1] + red[2] + red[3];
  }
}

// ---- device-resident MPC evaluation (gnnuds_evaluate; mpc.py:210-259, 565-582) -------------------
// dst[i, t, :] = src[((b0 + i) % n_src), t0 + t, :]: one shared state / runoff broadcast over the candidates of a chunk
// (mpc.py:237-240 tiles them on the host)
__global__ void __launch_bounds__(256) bcast_time_kernel(const float* __restrict__ src, float* __restrict__ dst, long long nb,
                                                         long long b0, int n_src, int Tsrc, int t0, int Tdst, long long L) {
  const long long total = nb * Tdst * L;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (long long)gridDim.x * blockDim.x) {
    const long long l = g % L, it = g / L;
    const int t = (int)(it % Tdst);
    const long long i = it / Tdst;
    dst[g] = src[(((b0 + i) % n_src) * Tsrc + t0 + t) * L + l];
  }
}
// settings of chunk steps t0..t0+T-1 from the decision matrix (mpc.py:211-218: every setting held r_step steps, the last
// one repeated to the end of the horizon): dst[i, t, j] = dec[b0 + i, min((t0 + t) / r_step, n_step - 1), j]
__global__ void __launch_bounds__(256) dec_slice_kernel(const float* __restrict__ dec, float* __restrict__ dst, long long nb,
                                                        long long b0, int n_step, int r_step, int t0, int T, int n_act) {
  const long long total = nb * T * n_act;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (long long)gridDim.x * blockDim.x) {
    const int j = (int)(g % n_act);
    const long long it = g / n_act;
    const int t = (int)(it % T);
    const long long i = it / T;
    const int st = min((t0 + t) / r_step, n_step - 1);
    dst[g] = dec[((b0 + i) * n_step + st) * n_act + j];
  }
}
// prev_q[i, n] = q_in of the last observed step (state[-1, n, 1], astlingen.py:83), the same for every candidate
__global__ void __launch_bounds__(256) init_prevq_kernel(const float* __restrict__ x_last, float* __restrict__ prev_q, long long nb,
                                                         int N, int n_in) {
  const long long total = nb * N;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (long long)gridDim.x * blockDim.x)
    prev_q[g] = x_last[(g % N) * n_in + 1];
}
// astlingen objective accumulated chunk by chunk: thread `tid` of candidate i owns the terms g = t*N + n with
// g % 128 == tid over the WHOLE horizon, in increasing g - the same partial sums objective_kernel forms in one pass,
// so obj_final_kernel returns bitwise the same value.  prev_q carries q_in of the last step into the next chunk.
struct ObjAccP {
  const float* y; int ny;     // (nb,T,N,ny) predictions of this chunk
  float* prev_q;              // (nb,N) q_in of the step before this chunk (state[:, -1, :, 1] for the first)
  const float* flood_w; const float* wwtp_w; const float* rough_w; const float* gamma;
  float* part;                // (nb,128) running per-thread partial sums
  long long nb; int T, N, t0, first;
};
__global__ void __launch_bounds__(128) obj_acc_kernel(ObjAccP p) {
  for (long long i = blockIdx.x; i < p.nb; i += gridDim.x) {
    const int tid = threadIdx.x;
    float acc = p.first ? 0.f : p.part[i * 128 + tid];
    const long long g0 = (long long)p.t0 * p.N;
    const int start = (int)(((tid - g0) % 128 + 128) % 128);
    for (int g = start; g < p.T * p.N; g += 128) {
      const int t = g / p.N, n = g % p.N;
      const float* ys = p.y + ((i * p.T + t) * p.N + n) * p.ny;
      const float gm = p.gamma ? p.gamma[p.t0 + t] : 1.f;
      float v = ys[p.ny - 1] * p.flood_w[n] + ys[1] * p.wwtp_w[n];
      if (p.rough_w) {
        const float prev = (t == 0) ? p.prev_q[i * p.N + n] : (ys - (long long)p.N * p.ny)[1];
        v += fabsf(ys[1] - prev) * p.rough_w[n];
      }
      acc += v * gm;
    }
    p.part[i * 128 + tid] = acc;
    __syncthreads();
    for (int n = tid; n < p.N; n += 128) p.prev_q[i * p.N + n] = p.y[((i * p.T + p.T - 1) * p.N + n) * p.ny + 1];
    __syncthreads();
  }
}
__global__ void __launch_bounds__(128) obj_final_kernel(const float* __restrict__ part, float* __restrict__ obj, long long nb) {
  __shared__ float red[4];
  for (long long i = blockIdx.x; i < nb; i += gridDim.x) {
    float acc = warp_sum(part[i * 128 + threadIdx.x]);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) obj[i] = red[0] + red[1] + red[2] + red[3];
  }
}

// ---- event replay: teacher-forced windows built on the device (main.py:327-354, dataloader.py:93-95) -------------
// The reference expands an event of T steps into T windows of `seq` steps on the host (expand_seq: window idx holds the
// steps idx-seq .. idx-1, zero-padded on the left; ones for settings) - a seq-fold blow-up of the data before it reaches
// the model.  Here the raw series is uploaded once and every window element is gathered by index arithmetic:
//   x [w, s, n, :]  = [h, cuminflow - lateral, outflow, (cumflooding > 0), lateral] at step w - Ti + s      (0 when < 0)
//   b [w, s, n, 0]  = lateral at step w + To - seq + s  (seq = max(Ti, To); the reference's r = states[seq_out:, ..., -1:]
//                     keeps the FIRST To steps of a seq-step window, emulator.py:532);  b[..., 1] = h * is_outfall (tide)
//   a [w, s, :]     = settings at step w + s;            ex[w, s, m, :] = edge state at step w - Ti + s     (0 when < 0)
struct ReplayP {
  const float* st; const float* pf; const float* es; const float* se; const float* outfall;
  float* x; float* b; float* a; float* ex;
  long long w0, nw; int T, N, M, Ti, To, seq, n_in, b_in, e_in, n_act, if_flood;
};
__global__ void __launch_bounds__(256) replay_windows_kernel(ReplayP p) {
  const long long stride = (long long)gridDim.x * blockDim.x, g0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  {
    const long long total = p.nw * p.Ti * p.N * p.n_in;
    for (long long g = g0; g < total; g += stride) {
      const int c = (int)(g % p.n_in);
      long long q = g / p.n_in;
      const int n = (int)(q % p.N); q /= p.N;
      const int s = (int)(q % p.Ti);
      const long long t = p.w0 + q / p.Ti - p.Ti + s;
      float v = 0.f;
      if (t >= 0) {
        const float* r = p.st + (t * p.N + n) * 4;
        if (c == 0) v = r[0];
        else if (c == 1) v = r[1] - r[3];
        else if (c == 2) v = r[2];
        else if (c == p.n_in - 1) v = r[3];
        else v = p.pf[t * p.N + n] > 0.f ? 1.f : 0.f;
      }
      p.x[g] = v;
    }
  }
  {
    const long long total = p.nw * p.To * p.N * p.b_in;
    for (long long g = g0; g < total; g += stride) {
      const int c = (int)(g % p.b_in);
      long long q = g / p.b_in;
      const int n = (int)(q % p.N); q /= p.N;
      const int s = (int)(q % p.To);
      const long long t = p.w0 + q / p.To + p.To - p.seq + s;
      float v = 0.f;
      if (t >= 0) v = c == 0 ? p.st[(t * p.N + n) * 4 + 3] : p.st[(t * p.N + n) * 4] * p.outfall[n];
      p.b[g] = v;
    }
  }
  if (p.a) {
    const long long total = p.nw * p.To * p.n_act;
    for (long long g = g0; g < total; g += stride) {
      const int j = (int)(g % p.n_act);
      const long long q = g / p.n_act;
      const int s = (int)(q % p.To);
      p.a[g] = p.se[(p.w0 + q / p.To + s) * p.n_act + j];
    }
  }
  {
    const long long total = p.nw * p.Ti * p.M * p.e_in;
    for (long long g = g0; g < total; g += stride) {
      const long long L = (long long)p.M * p.e_in, l = g % L, q = g / L;
      const int s = (int)(q % p.Ti);
      const long long t = p.w0 + q / p.Ti - p.Ti + s;
      p.ex[g] = t >= 0 ? p.es[t * L + l] : 0.f;
    }
  }
}

// fp32 FFMA peak of the device (bench.py's roofline denominator of the fp32 parity path): 16 independent chains per thread
__global__ void fp32_peak_kernel(float* out, int iters) {
  float a[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) a[k] = (float)(threadIdx.x + k) * 1e-3f;
  const float b = 0.999f, c = 1e-3f;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int k = 0; k < 16; ++k) a[k] = fmaf(a[k], b, c);
  }
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < 16; ++k) s += a[k];
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

}  // namespace gu

// Reverse-mode kernels of the Emulator forward w.r.t. its INPUTS (control settings, states) - no weight gradients.
// Reference: surrogate/mpc.py:511-627 (mpc_problem_gr: tf.GradientTape through Emulator.predict_tf +
// objective_pred_tf; gradient_fn :596-607, pred_fit :512-548).  fp32 SIMT; every kernel is the hand-derived VJP of
// one forward kernel of kernels.cuh and reads the activations that forward left on the tape (gnnuds.cu: Tape).
// Sub-gradient conventions follow TensorFlow's for the same ops: relu'(0) = 0, clip_by_value passes the gradient
// on the closed interval, abs'(0) = 0, comparisons (flood flag, epsilon mask) carry no gradient.
#pragma once

namespace gu {

// derivative of an activation expressed through its OUTPUT value o = act(v)
__device__ __forceinline__ float dact_from_out(float o, int act) {
  switch (act) {
    case ACT_RELU: return o > 0.f ? 1.f : 0.f;
    case ACT_TANH: return 1.f - o * o;
    case ACT_SIGMOID: return o * (1.f - o);
    case ACT_HARD_SIGMOID: return (o > 0.f && o < 1.f) ? 0.2f : 0.f;
    default: return 1.f;
  }
}

// ---------------------------------------------------------------------------
// VJP of rows_gemm_f32 w.r.t. its input rows.  Forward (one source segment of width Kw at weight rows k0..):
//   Y[(b,t,n), :] = act( sum_taps X[(b, t + (Ts-Td) - shift_j, n), :] @ W_j + bias ),   t < Td, source time < 0 -> zero.
// Backward for the source rows (b, t', n), t' < Ts:
//   gX[(b,t',n), kk] (+)= sum_j sum_c  G[(b, t' - (Ts-Td) + shift_j, n), c] * act'(Y[..., c]) * W_j[kk, c]
// with W_j[kk, c] = W[(k0_j + kk) * Nout + c].  Dense layers are the one-tap case (shift 0, Ts == Td).
// 64 x 64 output tile per CTA, 256 threads x (4 x 4), the contraction over c in chunks of 16.
// ---------------------------------------------------------------------------
struct BwdTap { int shift; int k0; };
struct BwdGemmP {
  const float* G; int ldg;      // (Bc*Td*Nn, Nout) cotangent of the layer OUTPUT (post-activation)
  const float* Y; int ldy;      // the layer's output values (for act'), null when act == linear
  int act;
  const float* W; int Nout;     // forward weight (Ktot, Nout) row-major
  BwdTap tap[4]; int ntap;
  float* out; int ldo; int Kw;  // (Bc*Ts*Nn, Kw) cotangent of the source segment
  int accumulate;
  long long rows_out;           // Bc*Ts*Nn
  int Nn, Td, Ts;
  int pitch_g, off_g;           // cotangent / output-value rows of a (b,t) sample start at (b*Td+t)*pitch_g + off_g (joint node+edge buffers)
};

__global__ void __launch_bounds__(256) bwd_gemm_kernel(BwdGemmP p) {
  constexpr int BM = 64, BN = 64, BK = 16;
  __shared__ float As[BK][BM + 4];
  __shared__ float Ws[BK][BN + 4];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const long long row0 = (long long)blockIdx.x * BM;
  const int col0 = blockIdx.y * BN;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  // loader mapping: thread loads 4 consecutive c of one row (64 rows x 16 c = 256 threads x 4)
  const int lr = tid >> 2, lc = (tid & 3) * 4;
  const long long r_out = row0 + lr;
  const int dT = p.Ts - p.Td;
  for (int j = 0; j < p.ntap; ++j) {
    // source row of the cotangent for this tap
    long long r_src = -1;
    if (r_out < p.rows_out) {
      const long long bt = r_out / p.Nn;
      const int n = (int)(r_out % p.Nn);
      const int tp = (int)(bt % p.Ts);
      const long long b = bt / p.Ts;
      const int t = tp - dT + p.tap[j].shift;
      if (t >= 0 && t < p.Td) r_src = (b * p.Td + t) * p.pitch_g + p.off_g + n;
    }
    const int kk = col0 + lr;                      // W row handled by this loader thread
    for (int c0 = 0; c0 < p.Nout; c0 += BK) {
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
      if (r_src >= 0) {
        a = *reinterpret_cast<const float4*>(p.G + r_src * p.ldg + c0 + lc);
        if (p.Y) {
          const float4 y = *reinterpret_cast<const float4*>(p.Y + r_src * p.ldy + c0 + lc);
          a.x *= dact_from_out(y.x, p.act); a.y *= dact_from_out(y.y, p.act);
          a.z *= dact_from_out(y.z, p.act); a.w *= dact_from_out(y.w, p.act);
        }
      }
      float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
      if (kk < p.Kw) w = *reinterpret_cast<const float4*>(p.W + (size_t)(p.tap[j].k0 + kk) * p.Nout + c0 + lc);
      __syncthreads();
      As[lc + 0][lr] = a.x; As[lc + 1][lr] = a.y; As[lc + 2][lr] = a.z; As[lc + 3][lr] = a.w;
      Ws[lc + 0][lr] = w.x; Ws[lc + 1][lr] = w.y; Ws[lc + 2][lr] = w.z; Ws[lc + 3][lr] = w.w;
      __syncthreads();
#pragma unroll
      for (int c = 0; c < BK; ++c) {
        float av[4], wv[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) av[i] = As[c][ty * 4 + i];
#pragma unroll
        for (int i = 0; i < 4; ++i) wv[i] = Ws[c][tx * 4 + i];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int q = 0; q < 4; ++q) acc[i][q] = fmaf(av[i], wv[q], acc[i][q]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long r = row0 + ty * 4 + i;
    if (r >= p.rows_out) continue;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int k = col0 + tx * 4 + q;
      if (k >= p.Kw) continue;
      float* d = p.out + r * p.ldo + k;
      *d = p.accumulate ? *d + acc[i][q] : acc[i][q];
    }
  }
}

// Small row counts (the gradient MPC's single decision vector: 150 rows per chunk and side): one round of loads instead of
// Nout/16 dependent load -> compute rounds per tap.  A CTA owns 16 output rows x 32 source columns, stages the cotangent
// tile of every tap (already multiplied by act') and the matching weight rows in shared memory, and 128 threads (row x 4
// interleaved columns cg, cg+8, cg+16, cg+24) accumulate over taps and c in the order of bwd_gemm_kernel: bitwise the
// same results.  Dynamic smem: ntap * 48 * (Nout + 4) floats.
__global__ void __launch_bounds__(128) bwd_gemm_skinny_kernel(BwdGemmP p) {
  extern __shared__ __align__(16) float bws_sm[];
  const int LD = p.Nout + 4, nq = p.Nout >> 2;
  float* As = bws_sm;                               // [ntap][16][LD]
  float* Ws = bws_sm + (size_t)p.ntap * 16 * LD;    // [ntap][32][LD]
  const int tid = threadIdx.x;
  const long long row0 = (long long)blockIdx.x * 16;
  const int col0 = blockIdx.y * 32;
  const int dT = p.Ts - p.Td;
  for (int idx = tid; idx < p.ntap * 32 * nq; idx += 128) {
    const int j = idx / (32 * nq), rem = idx - j * 32 * nq, kk = rem / nq, c = (rem - kk * nq) * 4;
    float* dst = Ws + ((size_t)j * 32 + kk) * LD + c;
    if (col0 + kk < p.Kw) cp_async_f4(dst, p.W + (size_t)(p.tap[j].k0 + col0 + kk) * p.Nout + c);
    else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");       // programmatic dependent launch: only the weight rows above are independent of the previous kernel
#pragma unroll 4
  for (int idx = tid; idx < p.ntap * 16 * nq; idx += 128) {
    const int j = idx / (16 * nq), rem = idx - j * 16 * nq, lr = rem / nq, c = (rem - lr * nq) * 4;
    const long long r_out = row0 + lr;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r_out < p.rows_out) {
      const long long bt = r_out / p.Nn;
      const int n = (int)(r_out % p.Nn);
      const int tp = (int)(bt % p.Ts);
      const long long b = bt / p.Ts;
      const int t = tp - dT + p.tap[j].shift;
      if (t >= 0 && t < p.Td) {
        const long long r_src = (b * p.Td + t) * p.pitch_g + p.off_g + n;
        a = *reinterpret_cast<const float4*>(p.G + r_src * p.ldg + c);
        if (p.Y) {
          const float4 y = *reinterpret_cast<const float4*>(p.Y + r_src * p.ldy + c);
          a.x *= dact_from_out(y.x, p.act); a.y *= dact_from_out(y.y, p.act);
          a.z *= dact_from_out(y.z, p.act); a.w *= dact_from_out(y.w, p.act);
        }
      }
    }
    *reinterpret_cast<float4*>(As + ((size_t)j * 16 + lr) * LD + c) = a;
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int lr = tid >> 3, cg = tid & 7;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int j = 0; j < p.ntap; ++j) {
    const float* ar = As + ((size_t)j * 16 + lr) * LD;
    const float* wr = Ws + ((size_t)j * 32 + cg) * LD;
#pragma unroll 2
    for (int c = 0; c < p.Nout; c += 4) {
      const float4 a = *reinterpret_cast<const float4*>(ar + c);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 w = *reinterpret_cast<const float4*>(wr + (size_t)q * 8 * LD + c);
        acc[q] = fmaf(a.x, w.x, acc[q]); acc[q] = fmaf(a.y, w.y, acc[q]);
        acc[q] = fmaf(a.z, w.z, acc[q]); acc[q] = fmaf(a.w, w.w, acc[q]);
      }
    }
  }
  const long long r = row0 + lr;
  if (r < p.rows_out) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int k = col0 + cg + 8 * q;
      if (k >= p.Kw) continue;
      float* d = p.out + r * p.ldo + k;
      *d = p.accumulate ? *d + acc[q] : acc[q];
    }
  }
}

// ---------------------------------------------------------------------------
// VJP of nodeedge_kernel: out[s] = Mat @ in[s]  ->  gIn[s] = Mat^T @ gOut[s].  matT is the forward's transposed
// folded matrix (R_in, ldm): matT[j*ldm + i] = Mat[i][j].  One thread per (s, j, 4 features).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) nodeedge_bwd_kernel(const float* __restrict__ matT, int ldm, const float* __restrict__ gout,
                                                           float* __restrict__ gin, long long S, int R_in, int R_out, int F) {
  const int f4 = F >> 2;
  const long long total = S * R_in * f4;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(g % f4);
    const long long sj = g / f4;
    const int j = (int)(sj % R_in);
    const long long s = sj / R_in;
    const float* m = matT + (size_t)j * ldm;
    const float* go = gout + s * R_out * F + c * 4;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int i = 0; i < R_out; ++i) {
      const float w = m[i];
      const float4 v = *reinterpret_cast<const float4*>(go + (size_t)i * F);
      a.x = fmaf(w, v.x, a.x); a.y = fmaf(w, v.y, a.y); a.z = fmaf(w, v.z, a.z); a.w = fmaf(w, v.w, a.w);
    }
    *reinterpret_cast<float4*>(gin + sj * F + c * 4) = a;
  }
}

// ---------------------------------------------------------------------------
// VJP of agg_small_kernel (GCN / GAT aggregation of one (b,t) sample) w.r.t. h = x@W.
//   GCN: gh_j = sum_i val_ij gO_i
//   GAT: alpha_ij = softmax_j(lrelu(s_i + n_j)),  gO_i = g_i * act'(out_i)
//        dalpha_ij = gO_i . h_j ;  dz_ij = alpha_ij (dalpha_ij - sum_k alpha_ik dalpha_ik) * lrelu'(z_ij)
//        gh_j = sum_i alpha_ij gO_i + (sum_j' dz_jj') a_self + (sum_i dz_ij) a_neigh
// One CTA per sample; h, gO and the gh accumulator live in shared memory (3*R*E floats); column sums by shared atomics.
// ---------------------------------------------------------------------------
struct AggBwdP {
  const float* h;        // (S,R,E) forward h = x@W
  const float* out;      // (S,R,E) forward layer output (post-activation)
  const float* gout;     // (S,R,E) cotangent of the layer output
  float* gh;             // (S,R,E)
  const int* row_ptr; const int* col_idx; const float* val;
  const float* a_self; const float* a_neigh;
  long long S; int R, E, act, gat;
  // use_adj (emulator.py:268-271, 343-362): the per-sample attention mask of the forward (AggP), a constant of the VJP
  const int* act_entry; const unsigned char* keep; const float* a_dyn; int n_act;
};

__device__ __forceinline__ bool aggb_kept(const AggBwdP& p, long long s, int q, int i, int j) {
  if (!p.a_dyn || i == j) return true;
  const int k = p.act_entry[q];
  return k ? truncf(p.a_dyn[s * p.n_act + k - 1]) != 0.f : p.keep[q] != 0;
}

// VJP of nothing: packs the node / edge stream cotangents (S,N,E), (S,M,E) into the joint layout (S,N+M,E) of graph_base != 0
__global__ void __launch_bounds__(256) joint_pack_kernel(const float* __restrict__ gx, const float* __restrict__ ge, float* __restrict__ gj,
                                                         long long S, int N, int M, int E) {
  const int R = N + M, e4 = E >> 2;
  const long long total = S * R * e4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % e4);
    const long long sr = i / e4;
    const int r = (int)(sr % R);
    const long long s = sr / R;
    const float4 v = r < N ? *reinterpret_cast<const float4*>(gx + ((s * N + r) * E) + c * 4)
                           : *reinterpret_cast<const float4*>(ge + ((s * M + r - N) * E) + c * 4);
    *reinterpret_cast<float4*>(gj + sr * E + c * 4) = v;
  }
}

__global__ void __launch_bounds__(1024) agg_bwd_kernel(AggBwdP p) {      // 256 threads per sample, 1024 when there are few samples
  extern __shared__ __align__(16) float sm[];
  const int R = p.R, E = p.E;
  float* hs = sm;                       // R*E
  float* gs = hs + (size_t)R * E;       // R*E  gO
  float* dh = gs + (size_t)R * E;       // R*E  accumulator
  float* ss = dh + (size_t)R * E;       // R
  float* sn = ss + R;                   // R
  float* dss = sn + R;                  // R
  float* dsn = dss + R;                 // R
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
  for (long long s = blockIdx.x; s < p.S; s += gridDim.x) {
    const size_t base = (size_t)s * R * E;
    __syncthreads();
    for (int i = tid * 4; i < R * E; i += blockDim.x * 4) {
      *reinterpret_cast<float4*>(hs + i) = *reinterpret_cast<const float4*>(p.h + base + i);
      float4 g = *reinterpret_cast<const float4*>(p.gout + base + i);
      const float4 o = *reinterpret_cast<const float4*>(p.out + base + i);
      g.x *= dact_from_out(o.x, p.act); g.y *= dact_from_out(o.y, p.act);
      g.z *= dact_from_out(o.z, p.act); g.w *= dact_from_out(o.w, p.act);
      *reinterpret_cast<float4*>(gs + i) = g;
      *reinterpret_cast<float4*>(dh + i) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    for (int i = tid; i < R; i += blockDim.x) { dss[i] = 0.f; dsn[i] = 0.f; }
    __syncthreads();
    if (p.gat) {
      for (int i = warp; i < R; i += nwarp) {
        float d1 = 0.f, d2 = 0.f;
        for (int c = lane; c < E; c += 32) {
          const float hv = hs[(size_t)i * E + c];
          d1 = fmaf(hv, p.a_self[c], d1); d2 = fmaf(hv, p.a_neigh[c], d2);
        }
        d1 = warp_sum(d1); d2 = warp_sum(d2);
        if (lane == 0) { ss[i] = d1; sn[i] = d2; }
      }
      __syncthreads();
    }
    for (int i = warp; i < R; i += nwarp) {
      const int beg = p.row_ptr[i], end = p.row_ptr[i + 1];
      const float* gi = gs + (size_t)i * E;
      if (!p.gat) {
        for (int q = beg; q < end; ++q) {
          const int j = p.col_idx[q];
          const float w = p.val[q];
          for (int c = lane; c < E; c += 32) atomicAdd(dh + (size_t)j * E + c, w * gi[c]);
        }
        continue;
      }
      const float si = ss[i];
      float mx = -INFINITY;
      for (int q = beg + lane; q < end; q += 32) {
        float z = si + sn[p.col_idx[q]];
        z = z > 0.f ? z : 0.2f * z;
        if (aggb_kept(p, s, q, i, p.col_idx[q])) mx = fmaxf(mx, z);
      }
      mx = warp_max(mx);
      float den = 0.f;
      for (int q = beg + lane; q < end; q += 32) {
        float z = si + sn[p.col_idx[q]];
        z = z > 0.f ? z : 0.2f * z;
        if (aggb_kept(p, s, q, i, p.col_idx[q])) den += expf(z - mx);
      }
      den = warp_sum(den);
      const float inv = 1.f / den;
      // pass A: t_i = sum_q alpha_q dalpha_q
      float ti = 0.f;
      for (int q = beg; q < end; ++q) {
        const int j = p.col_idx[q];
        if (!aggb_kept(p, s, q, i, j)) continue;
        float z = si + sn[j];
        z = z > 0.f ? z : 0.2f * z;
        const float alpha = expf(z - mx) * inv;
        float d = 0.f;
        for (int c = lane; c < E; c += 32) d = fmaf(gi[c], hs[(size_t)j * E + c], d);
        d = warp_sum(d);
        ti = fmaf(alpha, d, ti);
      }
      // pass B: score cotangents and the alpha-weighted scatter of gO_i
      float dsi = 0.f;
      for (int q = beg; q < end; ++q) {
        const int j = p.col_idx[q];
        if (!aggb_kept(p, s, q, i, j)) continue;
        const float zr = si + sn[j];
        const float z = zr > 0.f ? zr : 0.2f * zr;
        const float alpha = expf(z - mx) * inv;
        float d = 0.f;
        for (int c = lane; c < E; c += 32) d = fmaf(gi[c], hs[(size_t)j * E + c], d);
        d = warp_sum(d);
        const float dz = alpha * (d - ti) * (zr > 0.f ? 1.f : 0.2f);
        dsi += dz;
        if (lane == 0) atomicAdd(dsn + j, dz);
        for (int c = lane; c < E; c += 32) atomicAdd(dh + (size_t)j * E + c, alpha * gi[c]);
      }
      if (lane == 0) dss[i] = dsi;
    }
    __syncthreads();
    for (int i = tid; i < R * E; i += blockDim.x) {
      float v = dh[i];
      if (p.gat) {
        const int r = i / E, c = i - r * E;
        v += dss[r] * p.a_self[c] + dsn[r] * p.a_neigh[c];
      }
      p.gh[base + i] = v;
    }
  }
}

// ---------------------------------------------------------------------------
// VJP of cumsum_res_act_kernel: x_t = act(cumsum_t(v) + res) -> gv_t = sum_{t' >= t} g_t' act'(x_t'), gres = gv_0.
// In place on g; gres (Bc*Nn, E) optional.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) cumsum_bwd_kernel(float* __restrict__ g, const float* __restrict__ xout, float* __restrict__ gres,
                                                         long long Bc, int Tn, int Nn, int E, int act) {
  const long long total = Bc * Nn * E;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % E);
    const long long bn = i / E;
    const int n = (int)(bn % Nn);
    const long long b = bn / Nn;
    float acc = 0.f;
    for (int t = Tn - 1; t >= 0; --t) {
      const size_t o = (size_t)((b * Tn + t) * Nn + n) * E + c;
      acc += g[o] * dact_from_out(xout[o], act);
      g[o] = acc;
    }
    if (gres) gres[i] = acc;
  }
}

// ---------------------------------------------------------------------------
// VJP of head_kernel (Nout <= 4 columns): gin[r, k] (+)= sum_o gout[r, col0+o] * act'(out[r, col0+o]) * W[k, o]
// ---------------------------------------------------------------------------
struct HeadBwdP {
  const float* gout; const float* out; int ldo; int col0; int Nout; int act;
  const float* W; int K;
  float* gin; int ld; int accumulate;
  long long rows;
};

__global__ void __launch_bounds__(256) head_bwd_kernel(HeadBwdP p) {
  const long long total = p.rows * p.K;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(i % p.K);
    const long long r = i / p.K;
    float v = 0.f;
    for (int o = 0; o < p.Nout; ++o) {
      const float ov = p.out[r * p.ldo + p.col0 + o];
      v = fmaf(p.gout[r * p.ldo + p.col0 + o] * dact_from_out(ov, p.act), p.W[k * p.Nout + o], v);
    }
    float* d = p.gin + r * p.ld + k;
    *d = p.accumulate ? *d + v : v;
  }
}

// ---------------------------------------------------------------------------
// VJP of embed_kernel.
//  mode 0: gin[r, c] = ( sum_e (g[r,e] act'(out[r,e]) + [t == T-1] gres[(b,n), e]) W[c, e] ) / (max - min)
//  mode 2 (control input, K == 1): ga[bt, k] += sum_e g[r,e] act'(out[r,e]) W[0,e] for the rows n with gather[n] == k
// One warp per row.
// ---------------------------------------------------------------------------
struct EmbedBwdP {
  const float* g; const float* out;     // (rows, Eo) cotangent / forward output (post-activation)
  const float* gres;                    // (Bc*Nn, Eo) or null
  const float* W;                       // (K, Eo)
  const float* nmax; const float* nmin; int cn;
  const int* gather; int n_act;
  float* gin;                           // mode 0: (rows, K); mode 2: (Bc*T, n_act), accumulated atomically
  long long rows; int T, Nn, K, Eo, act, mode;
};

__global__ void __launch_bounds__(256) embed_bwd_kernel(EmbedBwdP p) {
  const int lane = threadIdx.x & 31;
  const long long w0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nw = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long r = w0; r < p.rows; r += nw) {
    const int n = (int)(r % p.Nn);
    const long long bt = r / p.Nn;
    int k_act = -1;
    if (p.mode == 2) { k_act = p.gather[n]; if (k_act < 0) continue; }
    const bool last = p.gres && (int)(bt % p.T) == p.T - 1;
    const long long bn = (bt / p.T) * p.Nn + n;
    float acc[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) acc[c] = 0.f;
    for (int e = lane; e < p.Eo; e += 32) {
      float v = p.g[r * p.Eo + e] * dact_from_out(p.out[r * p.Eo + e], p.act);
      if (last) v += p.gres[bn * p.Eo + e];
#pragma unroll
      for (int c = 0; c < 8; ++c)
        if (c < p.K) acc[c] = fmaf(v, p.W[c * p.Eo + e], acc[c]);
    }
#pragma unroll
    for (int c = 0; c < 8; ++c)
      if (c < p.K) acc[c] = warp_sum(acc[c]);
    if (lane == 0) {
      if (p.mode == 2) {
        atomicAdd(p.gin + bt * p.n_act + k_act, acc[0]);
      } else {
        for (int c = 0; c < p.K; ++c) {
          float v = acc[c];
          if (p.nmax) v /= (p.nmax[n * p.cn + c] - p.nmin[n * p.cn + c]);
          p.gin[r * p.K + c] = v;
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------
// VJP of postproc_kernel (post_proc_tf + de-normalisation + depth clip + constrain_tf; emulator.py:604-641, 680-771)
// for pump-free, offset-free networks.  One CTA per (b,t) sample.  Cotangents gy (S,N,ny), gey (S,M,e_out) of the
// physical outputs -> cotangents of the head outputs g_out_n (S,N,ldn), g_out_e (S,M,lde) and the DIRECT part of the
// settings gradient ga (S,n_act), accumulated atomically (flow * setting products, emulator.py:704, :715).
// Forward values are recomputed from the taped head outputs.
// ---------------------------------------------------------------------------
struct PostBwdP {
  PostP f;                 // the forward's parameter block (out_n / out_e / a / norms / graph vectors; y, ey unused)
  const float* gy; const float* gey;
  float* g_out_n; float* g_out_e; float* ga;
  const float* y_final;    // (S,N,ny) the forward's physical node output (depth after the flood / pumped-storage rules)
  const int* ps; int n_ps; // pumped-storage nodes: their depth comes from the scan (pump_storage_bwd_kernel), not from the head
};

__global__ void __launch_bounds__(256) postproc_bwd_kernel(PostBwdP q) {
  extern __shared__ float sh[];     // gO[N] | gI[N] | gefl[M] | efl[M] | vpre[M] | pass[M] | hn[N]
  const PostP& p = q.f;
  float* gO = sh; float* gI = gO + p.N; float* gefl = gI + p.N; float* efl = gefl + p.M;
  float* vpre = efl + p.M;          // last edge channel after the offset / pump rules, before the setting product
  float* pass = vpre + p.M;         // 1 where that value still depends on the head output, 0 where a rule replaced it
  float* hn = pass + p.M;           // normalised node depth (tide applied): input of the rules' gates
  const bool need_h = p.pump || p.offset;
  for (long long s = blockIdx.x; s < p.S; s += gridDim.x) {
    __syncthreads();
    if (need_h) {
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
    // ---- recompute the physical edge flows (forward edge pass incl. the inlet-offset gate and the rated pump flow) ----
    for (int m = threadIdx.x; m < p.M; m += blockDim.x) {
      float v = p.out_e[(s * p.M + m) * p.lde + p.e_out - 1];
      float ps_ = 1.f;
      if (p.offset && p.offset[m] > 0.f) {                       // emulator.py:686-693
        const int fn = p.edge_from[m];
        const float mx = p.ymax[fn * p.cy], mn = p.ymin[fn * p.cy];
        const float inoff = (hn[fn] * (mx - mn) + mn) - p.hmin[fn];
        if (v > 0.f && !(inoff > p.offset[m])) { v = 0.f; ps_ = 0.f; }
      }
      if (p.act && p.pump && p.pump[m] != 0.f) {                 // emulator.py:698-701
        const float mx = p.emax[m * p.ce + p.e_out - 1];
        const float fl = p.pump[m] * (hn[p.edge_from[m]] > 0.01f ? 1.f : 0.f) * (mx > 1e-3f ? 1.f / mx : 0.f);
        if (fl != 0.f) { v = fl; ps_ = 0.f; }
      }
      vpre[m] = v; pass[m] = ps_;
      if (p.act) { const int k = p.act_of_edge[m]; if (k >= 0) v *= p.a[s * p.n_act + k]; }
      const float mx = p.emax[m * p.ce + p.e_out - 1], mn = p.emin[m * p.ce + p.e_out - 1];
      efl[m] = v * (mx - mn) + mn;
      gefl[m] = 0.f;
    }
    __syncthreads();
    // ---- node pass ----
    for (int n = threadIdx.x; n < p.N; n += blockDim.x) {
      const float* src = p.out_n + (s * p.N + n) * p.ldn;
      const float* g = q.gy + (s * p.N + n) * p.ny;
      float* go = q.g_out_n + (s * p.N + n) * p.ldn;
      for (int c = 0; c < p.ldn; ++c) go[c] = 0.f;
      float hN = src[0];
      if (p.tide) {
        const float braw = p.b_raw[(s * p.N + n) * p.b_in + p.b_in - 1];
        const float mx = p.bmax[n * p.cb + p.b_in - 1], mn = p.bmin[n * p.cb + p.b_in - 1];
        hN = hN * (1.f - p.is_outfall[n]) + (braw - mn) / (mx - mn);
      }
      float yn1, yn2, pre1 = 0.f, pre2 = 0.f;      // pre*: node-head flows before the setting product (non-fusion)
      float pass1 = 1.f, pass2 = 1.f;              // 0 where a rated pump flow replaced the head's value
      bool is_ps = false;
      for (int k = 0; k < q.n_ps; ++k) is_ps |= q.ps[k] == n;
      const float m1 = p.ymax[n * p.cy + 1], m2 = p.ymax[n * p.cy + 2];
      const float i1 = m1 > 1e-3f ? 1.f / m1 : 0.f, i2 = m2 > 1e-3f ? 1.f / m2 : 0.f;
      int ko = -1, ki = -1;
      if (p.edge_fusion) {
        float op = 0.f, on = 0.f, ip = 0.f, in_ = 0.f;
        for (int e = p.inc_ptr[n]; e < p.inc_ptr[n + 1]; ++e) {
          const float fl = efl[p.inc_edge[e]];
          const float pos = fmaxf(fl, 0.f), neg = fmaxf(-fl, 0.f);
          if (p.inc_sign[e] > 0) { op += pos; in_ += neg; } else { ip += pos; on += neg; }
        }
        yn2 = (op + on) * i2; yn1 = (ip + in_) * i1;
      } else {
        pre1 = src[1]; pre2 = src[2];
        if (p.act && p.pump_in) {                                // emulator.py:708-713: rated flows replace the head's
          const float on = hN > (p.tf_rules ? 0.f : 0.01f) ? 1.f : 0.f;
          const float fli = p.pump_in[n] * on / p.ymax[n * p.cy + 1], flo = p.pump_out[n] * on / p.ymax[n * p.cy + 2];
          if (fli != 0.f) { pre1 = fli; pass1 = 0.f; }
          if (flo != 0.f) { pre2 = flo; pass2 = 0.f; }
        }
        yn1 = pre1; yn2 = pre2;
        if (p.act) {
          ko = p.act_out_of_node[n]; ki = p.act_in_of_node[n];
          if (ko >= 0) yn2 *= p.a[s * p.n_act + ko];
          if (ki >= 0) yn1 *= p.a[s * p.n_act + ki];
        }
      }
      const float sc0 = p.ymax[n * p.cy] - p.ymin[n * p.cy];
      const float sc1 = p.ymax[n * p.cy + 1] - p.ymin[n * p.cy + 1], sc2 = p.ymax[n * p.cy + 2] - p.ymin[n * p.cy + 2];
      const float yp0 = hN * sc0 + p.ymin[n * p.cy];
      const float yp1 = yn1 * sc1 + p.ymin[n * p.cy + 1], yp2 = yn2 * sc2 + p.ymin[n * p.cy + 2];
      float pr = 0.f; bool fl = false;
      if (p.if_flood) {
        const float mx = p.ymax[n * p.cy + 3], mn = p.ymin[n * p.cy + 3];
        pr = src[p.n_out] * (mx - mn) + mn;
        fl = pr > 0.5f;
      }
      // constrain_tf backward
      const bool in_clip = yp0 >= p.hmin[n] && yp0 <= p.hmax[n];
      float h = fminf(fmaxf(yp0, p.hmin[n]), p.hmax[n]);
      if (fl) h = p.hmax[n];
      if (q.y_final) h = q.y_final[(s * p.N + n) * p.ny];          // pumped-storage nodes: the scan's depth
      const float r = p.b_raw[(s * p.N + n) * p.b_in];
      const float qraw = yp1 + r - yp2;
      float mask = 1.f - p.is_outfall[n];
      if (p.epsilon > 0.f) mask *= ((p.hmax[n] - h) < p.epsilon) ? 1.f : 0.f;
      else if (p.epsilon == 0.f) {}
      else if (p.if_flood) mask *= fl ? 1.f : 0.f;
      const float gqw = g[p.ny - 1] * mask * (qraw >= 0.f ? 1.f : 0.f);
      const float gyp0 = (fl || !in_clip || is_ps) ? 0.f : g[0];
      const float gyp1 = g[1] + gqw, gyp2 = g[2] - gqw;
      if (p.if_flood) go[p.n_out] = g[3] * (p.ymax[n * p.cy + 3] - p.ymin[n * p.cy + 3]);
      go[0] = gyp0 * sc0 * (p.tide ? (1.f - p.is_outfall[n]) : 1.f);
      const float gyn1 = gyp1 * sc1, gyn2 = gyp2 * sc2;
      if (p.edge_fusion) {
        gO[n] = gyn2 * i2; gI[n] = gyn1 * i1;
      } else {
        float a_o = 1.f, a_i = 1.f;
        if (ko >= 0) { a_o = p.a[s * p.n_act + ko]; atomicAdd(q.ga + s * p.n_act + ko, gyn2 * pre2); }
        if (ki >= 0) { a_i = p.a[s * p.n_act + ki]; atomicAdd(q.ga + s * p.n_act + ki, gyn1 * pre1); }
        go[1] = gyn1 * a_i * pass1; go[2] = gyn2 * a_o * pass2;
      }
    }
    __syncthreads();
    if (p.edge_fusion) {
      // mass-balance transpose: an edge's flow feeds the outflow of its from-node / inflow of its to-node when >= 0,
      // and the inflow of its from-node / outflow of its to-node (negated) when <= 0
      for (int n = threadIdx.x; n < p.N; n += blockDim.x) {
        for (int e = p.inc_ptr[n]; e < p.inc_ptr[n + 1]; ++e) {
          const int m = p.inc_edge[e];
          const float fl = efl[m];
          float g = 0.f;
          if (p.inc_sign[e] > 0) { if (fl >= 0.f) g += gO[n]; if (fl <= 0.f) g -= gI[n]; }
          else                   { if (fl >= 0.f) g += gI[n]; if (fl <= 0.f) g -= gO[n]; }
          atomicAdd(gefl + m, g);
        }
      }
      __syncthreads();
    }
    // ---- edge pass ----
    for (int m = threadIdx.x; m < p.M; m += blockDim.x) {
      const float* src = p.out_e + (s * p.M + m) * p.lde;
      const float* g = q.gey + (s * p.M + m) * p.e_out;
      float* go = q.g_out_e + (s * p.M + m) * p.lde;
      const int k = p.act ? p.act_of_edge[m] : -1;
      const float ae = k >= 0 ? p.a[s * p.n_act + k] : 1.f;
      for (int c = 0; c < p.e_out; ++c) {
        const float mx = p.emax[m * p.ce + c], mn = p.emin[m * p.ce + c];
        const bool lastc = c == p.e_out - 1;
        const float v = lastc ? vpre[m] * ae : src[c];
        const float phys = v * (mx - mn) + mn;
        float gp = g[c];
        if (c == 0 && !(phys >= 0.f && phys <= p.ehmax[m])) gp = 0.f;     // emulator.py:624 depth clip
        if (lastc) gp += gefl[m];
        const float gv = gp * (mx - mn);
        if (lastc) {
          go[c] = gv * ae * pass[m];
          if (k >= 0) atomicAdd(q.ga + s * p.n_act + k, gv * vpre[m]);
        } else {
          go[c] = gv;
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------
// VJP of pump_storage_kernel (emulator.py:626-637): the depth of a pumped-storage node is integrated from the predicted
// flows, de_0 = clip(h_in[last]), de_t = clip(de_{t-1} + (q_us - q_ds)_t / (area + 1e-6)), and replaces the head's depth.
// One thread per (scenario, pumped-storage node): the clip flags are recomputed forwards, then the depth cotangents run
// backwards through the scan into the flow cotangents gy[...,1] / gy[...,2] (in place) and the input state's depth
// (gx0 (Bc, n_ps), added to gx by ps_gx_add_kernel).  The node's own gy[...,0] is consumed (zeroed).
// ---------------------------------------------------------------------------
struct PumpStoreBwdP {
  const float* y; int ny;        // (Bc,To,N,ny) forward output
  const float* x; int n_in;      // (Bc,Ti,N,n_in)
  float* gy;                     // (Bc,To,N,ny) cotangents, modified in place
  float* gx0;                    // (Bc,n_ps)
  const int* ps; int n_ps;
  const float* area; const float* hmin; const float* hmax;
  long long Bc; int Ti, To, N, if_flood;
};

__global__ void __launch_bounds__(128) pump_storage_bwd_kernel(PumpStoreBwdP p) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= p.Bc * p.n_ps) return;
  const long long b = i / p.n_ps;
  const int n = p.ps[i % p.n_ps];
  const float lo = p.hmin[n], hi = p.hmax[n], ar = p.area[n] + 1e-6f;
  unsigned long long inclip[4] = {0ull, 0ull, 0ull, 0ull};       // To <= 256
  float de = 0.f;
  for (int t = 0; t < p.To; ++t) {
    const float* yy = p.y + ((b * p.To + t) * p.N + n) * p.ny;
    const float v = t == 0 ? p.x[((b * p.Ti + p.Ti - 1) * p.N + n) * p.n_in] : de + (yy[1] - yy[2]) / ar;
    if (v >= lo && v <= hi) inclip[t >> 6] |= 1ull << (t & 63);
    de = fminf(fmaxf(v, lo), hi);
  }
  float carry = 0.f;
  for (int t = p.To - 1; t >= 0; --t) {
    const size_t o = (size_t)((b * p.To + t) * p.N + n) * p.ny;
    const bool f = p.if_flood && p.y[o + 3] > 0.5f;
    const float g = (f ? 0.f : p.gy[o]) + carry;
    const float gv = (inclip[t >> 6] >> (t & 63)) & 1ull ? g : 0.f;
    p.gy[o] = 0.f;
    if (t > 0) { carry = gv; p.gy[o + 1] += gv / ar; p.gy[o + 2] -= gv / ar; }
    else p.gx0[i] = gv;
  }
}

__global__ void __launch_bounds__(128) ps_gx_add_kernel(float* __restrict__ gx, const float* __restrict__ gx0, const int* __restrict__ ps,
                                                        int n_ps, long long Bc, int Ti, int N, int n_in) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Bc * n_ps) return;
  const long long b = i / n_ps;
  const int n = ps[i % n_ps];
  gx[((b * Ti + Ti - 1) * N + n) * n_in] += gx0[i];
}

// ---------------------------------------------------------------------------
// VJP of objective_kernel (astlingen.objective_pred_tf) -> cotangent of y (B,T,N,ny); every channel written.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) objective_bwd_kernel(ObjP p, float* __restrict__ gy) {
  const long long total = p.B * p.T * p.N;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(g % p.N);
    const long long bt = g / p.N;
    const int t = (int)(bt % p.T);
    const long long b = bt / p.T;
    const float* ys = p.y + g * p.ny;
    float* d = gy + g * p.ny;
    for (int c = 0; c < p.ny; ++c) d[c] = 0.f;
    const float gm = p.gamma ? p.gamma[t] : 1.f;
    d[p.ny - 1] = gm * p.flood_w[n];
    float g1 = gm * p.wwtp_w[n];
    if (p.rough_w) {
      const float cur = ys[1];
      const float prev = (t == 0) ? p.q_in0[b * p.N + n] : (ys - (long long)p.N * p.ny)[1];
      const float dd = cur - prev;
      g1 += gm * p.rough_w[n] * (dd > 0.f ? 1.f : (dd < 0.f ? -1.f : 0.f));
      if (t + 1 < p.T) {
        const float nxt = (ys + (long long)p.N * p.ny)[1];
        const float dn = nxt - cur;
        const float gn = p.gamma ? p.gamma[t + 1] : 1.f;
        g1 -= gn * p.rough_w[n] * (dn > 0.f ? 1.f : (dn < 0.f ? -1.f : 0.f));
      }
    }
    d[1] = g1;
  }
}

// ---------------------------------------------------------------------------
// VJP of feedback_kernel (MPC / MBRL rules): the next chunk's state gradient flows into this chunk's output cotangents
//   gy[..., 0:3] += gx'[..., 0:3]      gey += gex'[..., :e_out]      ga[s, act_of_edge[m]] += gex'[s, m, e_out]
// (the flood flag and the runoff carry no gradient).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) feedback_bwd_kernel(const float* __restrict__ gx, int n_in, const float* __restrict__ gex, int e_in,
                                                           float* __restrict__ gy, int ny, float* __restrict__ gey, int e_out,
                                                           float* __restrict__ ga, int n_act, const int* __restrict__ act_of_edge,
                                                           long long S, int N, int M) {
  const long long tn = S * N, tm = S * M;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < tn + tm; g += (long long)gridDim.x * blockDim.x) {
    if (g < tn) {
      for (int c = 0; c < 3; ++c) gy[g * ny + c] += gx[g * n_in + c];
    } else {
      const long long ge = g - tn;
      const int m = (int)(ge % M);
      const long long s = ge / M;
      for (int c = 0; c < e_out; ++c) gey[ge * e_out + c] += gex[ge * e_in + c];
      if (ga && e_in > e_out) {
        const int k = act_of_edge[m];
        if (k >= 0) atomicAdd(ga + s * n_act + k, gex[ge * e_in + e_out]);
      }
    }
  }
}

// ---------------------------------------------------------------------------
// Back-propagation through time of gru_kernel / lstm_kernel (keras GRU reset_after / LSTM, emulator.py:159-163) w.r.t.
// the input projection mx = x@K + b and, through it, the layer input.  Same tiling as the forward: a CTA owns 32
// sequences x 8 column groups, R resident in shared memory; per step the gates are recomputed from the taped mx and
// h_{t-1} (one H x G mat-vec) and the state cotangent is pushed back through R^T (one G x H mat-vec).
//   GRU : z = s(mx_z + mh_z), r = s(mx_r + mh_r), hh = tanh(mx_h + r mh_h), h' = z h + (1 - z) hh,  mh = h R + b1
//   LSTM: z = zx + h R, [i f g o] = [s s tanh s](z), c' = f c + i g, h' = o tanh(c')
// gout (Bc,Td,Nn,H): cotangent of the last Td hidden states; gmx (Bc,Ts,Nn,G) written for every step.
// ---------------------------------------------------------------------------
struct RecBwdP {
  const float* mx;       // (Bc,Ts,Nn,G) taped input projection
  const float* hall;     // (Bc,Ts,Nn,H) taped hidden states
  const float* call;     // LSTM: taped cell states
  const float* R;        // (H,G)
  const float* b1;       // GRU: (G) recurrent bias
  const float* gout;     // (Bc,Td,Nn,H)
  float* gmx;            // (Bc,Ts,Nn,G)
  long long nseq; int Nn, Ts, Td, H;
};

template <bool LSTM>
__global__ void __launch_bounds__(256) rec_bwd_kernel(RecBwdP p) {
  extern __shared__ __align__(16) float sm[];
  const int H = p.H, G = (LSTM ? 4 : 3) * H;
  float* Rs = sm;                    // H*G
  float* hp = Rs + (size_t)H * G;    // 32*H  h_{t-1}
  float* mh = hp + 32 * H;           // 32*G  recurrent pre-activations, then their cotangents
  float* dh = mh + 32 * G;           // 32*H  state cotangent carried to step t-1
  float* dc = dh + 32 * H;           // 32*H  LSTM cell cotangent
  const int tid = threadIdx.x;
  for (int i = tid * 4; i < H * G; i += 256 * 4)
    *reinterpret_cast<float4*>(Rs + i) = *reinterpret_cast<const float4*>(p.R + i);
  for (int i = tid; i < 32 * H; i += 256) { dh[i] = 0.f; dc[i] = 0.f; }
  const int rr = tid >> 3, gg = tid & 7;
  const long long q = (long long)blockIdx.x * 32 + rr;
  const bool live = q < p.nseq;
  const long long b = live ? q / p.Nn : 0;
  const int n = live ? (int)(q % p.Nn) : 0;
  const int ncol4 = G / 32;
  __syncthreads();
  for (int t = p.Ts - 1; t >= 0; --t) {
    const size_t row = (size_t)((b * p.Ts + t) * p.Nn + n);
    const size_t rowp = (size_t)((b * p.Ts + t - 1) * p.Nn + n);
    for (int u = gg; u < H; u += 8) hp[rr * H + u] = (live && t > 0) ? p.hall[rowp * H + u] : 0.f;
    __syncthreads();
    for (int qq = 0; qq < ncol4; ++qq) {             // mh[rr][:] = hp[rr][:] @ R (+ b1)
      const int col = qq * 32 + gg * 4;
      float4 s = LSTM ? make_float4(0.f, 0.f, 0.f, 0.f) : *reinterpret_cast<const float4*>(p.b1 + col);
      for (int k = 0; k < H; ++k) {
        const float hv = hp[rr * H + k];
        const float4 rv = *reinterpret_cast<const float4*>(Rs + (size_t)k * G + col);
        s.x = fmaf(hv, rv.x, s.x); s.y = fmaf(hv, rv.y, s.y); s.z = fmaf(hv, rv.z, s.z); s.w = fmaf(hv, rv.w, s.w);
      }
      *reinterpret_cast<float4*>(mh + rr * G + col) = s;
    }
    __syncthreads();
    const int td = t - (p.Ts - p.Td);
    if (live) {
      const float* mxr = p.mx + row * G;
      float* gm = p.gmx + row * G;
      for (int u = gg; u < H; u += 8) {
        const float g = dh[rr * H + u] + (td >= 0 ? p.gout[((size_t)((b * p.Td + td) * p.Nn + n)) * H + u] : 0.f);
        if (!LSTM) {
          const float z = 1.f / (1.f + expf(-(mxr[u] + mh[rr * G + u])));
          const float r = 1.f / (1.f + expf(-(mxr[H + u] + mh[rr * G + H + u])));
          const float mhh = mh[rr * G + 2 * H + u];
          const float hh = tanhf(mxr[2 * H + u] + r * mhh);
          const float dz = g * (hp[rr * H + u] - hh), dhh = g * (1.f - z);
          const float dah = dhh * (1.f - hh * hh);
          const float daz = dz * z * (1.f - z), dar = dah * mhh * r * (1.f - r);
          gm[u] = daz; gm[H + u] = dar; gm[2 * H + u] = dah;
          mh[rr * G + u] = daz; mh[rr * G + H + u] = dar; mh[rr * G + 2 * H + u] = dah * r;
          dh[rr * H + u] = g * z;
        } else {
          const float ig = 1.f / (1.f + expf(-(mxr[u] + mh[rr * G + u])));
          const float fg = 1.f / (1.f + expf(-(mxr[H + u] + mh[rr * G + H + u])));
          const float cg = tanhf(mxr[2 * H + u] + mh[rr * G + 2 * H + u]);
          const float og = 1.f / (1.f + expf(-(mxr[3 * H + u] + mh[rr * G + 3 * H + u])));
          const float cprev = t > 0 ? p.call[rowp * H + u] : 0.f;
          const float tc = tanhf(p.call[row * H + u]);
          const float dcc = dc[rr * H + u] + g * og * (1.f - tc * tc);
          const float dzi = dcc * cg * ig * (1.f - ig), dzf = dcc * cprev * fg * (1.f - fg);
          const float dzg = dcc * ig * (1.f - cg * cg), dzo = g * tc * og * (1.f - og);
          gm[u] = dzi; gm[H + u] = dzf; gm[2 * H + u] = dzg; gm[3 * H + u] = dzo;
          mh[rr * G + u] = dzi; mh[rr * G + H + u] = dzf; mh[rr * G + 2 * H + u] = dzg; mh[rr * G + 3 * H + u] = dzo;
          dc[rr * H + u] = dcc * fg;
          dh[rr * H + u] = 0.f;
        }
      }
    }
    __syncthreads();
    // dh[rr][k] += sum_g dmh[rr][g] R[k][g]
    for (int k = gg; k < H; k += 8) {
      const float* rk = Rs + (size_t)k * G;
      const float* dm = mh + rr * G;
      float s = 0.f;
      for (int c = 0; c < G; c += 4) {
        const float4 rv = *reinterpret_cast<const float4*>(rk + c);
        const float4 dv = *reinterpret_cast<const float4*>(dm + c);
        s = fmaf(dv.x, rv.x, fmaf(dv.y, rv.y, fmaf(dv.z, rv.z, fmaf(dv.w, rv.w, s))));
      }
      dh[rr * H + k] += s;
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------
// NN-cat ends (csrc/nn_cat.cuh).  VJP of nn_unflatten_kernel + the head activation: the per-node cotangents of a head
// become the flattened row Gflat[r, n*nc + c] = g[(r*Nn+n), col0+c] * act'(out[...]) (zero in the padded columns).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) nn_flatten_bwd_kernel(const float* __restrict__ g, const float* __restrict__ out, int ldo, int col0,
                                                             int act, float* __restrict__ gflat, int ldi, long long rows, int Nn, int nc) {
  const long long total = rows * ldi;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / ldi;
    const int j = (int)(i - r * ldi);
    float v = 0.f;
    if (j < Nn * nc) {
      const int n = j / nc, c = j - n * nc;
      const long long o = (r * Nn + n) * ldo + col0 + c;
      v = g[o] * dact_from_out(out[o], act);
    }
    gflat[i] = v;
  }
}

// VJP of nn_prep_kernel.  mode 0: gin[r, n, c] = gflat[r, n*K + c] / (max - min);  mode 2 (control row, K == 1):
// ga[r, gather[n]] += gflat[r, n].
__global__ void __launch_bounds__(256) nn_prep_bwd_kernel(const float* __restrict__ gflat, int Kp, const float* __restrict__ nmax,
                                                          const float* __restrict__ nmin, int cn, const int* __restrict__ gather, int n_act,
                                                          float* __restrict__ gin, long long rows, int Nn, int K, int mode) {
  const int w = Nn * K;
  const long long total = rows * w;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / w;
    const int j = (int)(i - r * w);
    const int n = j / K, c = j - n * K;
    float v = gflat[r * Kp + j];
    if (mode == 2) {
      const int k = gather[n];
      if (k >= 0) atomicAdd(gin + r * n_act + k, v);
    } else {
      if (nmax) v /= (nmax[n * cn + c] - nmin[n * cn + c]);
      gin[i] = v;
    }
  }
}

// q0[b, n] = x[b, Ti-1, n, 1]: the last observed inflow, objective_pred_tf's q_in[:, 0] (astlingen.py:79)
__global__ void __launch_bounds__(256) pick_q0_kernel(const float* __restrict__ x, float* __restrict__ q0, long long B, int Ti, int N, int n_in) {
  const long long total = B * N;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (long long)gridDim.x * blockDim.x) {
    const long long b = g / N;
    const int n = (int)(g % N);
    q0[g] = x[((b * Ti + Ti - 1) * N + n) * n_in + 1];
  }
}

}  // namespace gu

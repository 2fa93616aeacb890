// NN-cat variant (conv falsy, emulator.py:180,197-211,231-233,284-286,322-337): the node / edge features of one
// time step are flattened into ONE row (N*C wide), so the whole network is a chain of row GEMMs over Bc*T rows
// (rows_gemm_f32 / the temporal kernels with Nn = 1).  Only the two ends need their own kernels:
//   nn_prep_kernel      raw (Bc,T,Nn,K) input -> min-max normalised (emulator.py:803-810), flattened row of width
//                       Nn*K zero-padded to a multiple of 16 (the K granule of rows_gemm_f32); mode 2 builds
//                       get_edge_action's row (emulator.py:385-398) from the settings;
//   nn_unflatten_kernel head output rows (Bc*T, Nn*nc [padded]) -> the per-node layout (Bc*T*Nn, ld) at column
//                       col0 that postproc_kernel consumes (emulator.py:325,331,337 reshape).
#pragma once
#include "kernels.cuh"

struct NnPrepP {
  const float* in;       // (rows, Nn, K) raw; mode 2: a (rows, n_act)
  const float* nmax;     // (Nn, cn) or null (already normalised)
  const float* nmin;
  int cn;
  const int* gather;     // mode 2: act_of_edge (Nn)
  int n_act;
  float* out;            // (rows, Kp)
  long long rows;
  int Nn, K, Kp, mode;
};

__global__ void __launch_bounds__(256) nn_prep_kernel(NnPrepP p) {
  const long long total = p.rows * p.Kp;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / p.Kp;
    const int j = (int)(i - r * p.Kp);
    float v = 0.f;
    if (j < p.Nn * p.K) {
      const int n = j / p.K, c = j - n * p.K;
      if (p.mode == 2) {
        const int k = p.gather[n];
        v = k >= 0 ? p.in[r * p.n_act + k] : 1.f;
      } else {
        v = p.in[r * (long long)(p.Nn * p.K) + j];
        if (p.nmax) {
          const float mx = p.nmax[n * p.cn + c], mn = p.nmin[n * p.cn + c];
          v = (v - mn) / (mx - mn);
        }
      }
    }
    p.out[i] = v;
  }
}

struct NnUnflatP {
  const float* in;       // (rows, ldi): column n*nc + c
  int ldi;
  float* out;            // (rows*Nn, ldo) at column col0
  int ldo, col0;
  long long rows;
  int Nn, nc;
};

__global__ void __launch_bounds__(256) nn_unflatten_kernel(NnUnflatP p) {
  const int w = p.Nn * p.nc;
  const long long total = p.rows * w;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / w;
    const int j = (int)(i - r * w);
    const int n = j / p.nc, c = j - n * p.nc;
    p.out[(r * p.Nn + n) * p.ldo + p.col0 + c] = p.in[r * p.ldi + j];
  }
}

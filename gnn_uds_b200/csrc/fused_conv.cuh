// Shared store helper of the fused tensor-core kernels (csrc/fused_tp.cuh).  The experimental conv + aggregation kernel
// that lived here (conv_agg_kernel, GNNUDS_FUSED=2) was superseded by sp_agg_kernel / sp_aggt_kernel (csrc/fused_agg.cuh,
// csrc/fused_aggt.cuh) and removed at the end of round 2.
#pragma once
#include <cuda.h>

#include "kernels.cuh"
#include "tc_gemm.cuh"
#include "umma.cuh"

namespace gu {

__device__ __forceinline__ void st_global_v8(void* p, const uint4& a, const uint4& b) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b.x),
               "r"(b.y), "r"(b.z), "r"(b.w)
               : "memory");
}

}  // namespace gu

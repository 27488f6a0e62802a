// Shared epilogue of the tcgen05 distance kernels (knn_gemm.cu, knn_gemm_pair.cu).
#pragma once
#include "knn_internal.cuh"
#include "ptx.cuh"
#include "topk.cuh"

namespace vqacx {

// Epilogue of one [128 x 256] accumulator tile, shared by the single-CTA and the CTA-pair kernels: this thread's
// TMEM lane (= query row) x 256 columns, v = |y~|^2 - 2 x~.y~, threshold filter + warp compaction (topk.cuh).
__device__ __forceinline__ void knn_epilogue_tile(RowFilter& st, uint32_t t_addr, const float* yn, uint32_t col_base) {
#pragma unroll 1
  for (int chunk = 0; chunk < kGemmBN / 32; ++chunk) {
    uint32_t r[32];
    tmem_ld_32x32(t_addr + chunk * 32, r);
    tmem_ld_wait();
    float v[32];
    float vmin = __int_as_float(0x7f800000);
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      const float4 y4 = *reinterpret_cast<const float4*>(yn + chunk * 32 + j);
      v[j + 0] = fmaf(-2.f, __uint_as_float(r[j + 0]), y4.x);
      v[j + 1] = fmaf(-2.f, __uint_as_float(r[j + 1]), y4.y);
      v[j + 2] = fmaf(-2.f, __uint_as_float(r[j + 2]), y4.z);
      v[j + 3] = fmaf(-2.f, __uint_as_float(r[j + 3]), y4.w);
      vmin = fminf(fminf(vmin, fminf(v[j], v[j + 1])), fminf(v[j + 2], v[j + 3]));
    }
    if (vmin < st.tau) {
      const uint32_t col0 = col_base + chunk * 32;
#pragma unroll
      for (int j = 0; j < 32; ++j) filter_push(st, v[j], col0 + j);
    }
    filter_compact_warp(st, 32);
  }
}

}  // namespace vqacx

// Shared epilogue of the tcgen05 distance kernels (knn_gemm.cu, knn_gemm_pair.cu).
#pragma once
#include "knn_internal.cuh"
#include "ptx.cuh"
#include "topk.cuh"

namespace vqacx {

// Per-thread staging row in shared memory for the values of one 32-column chunk: 36 floats (144 B) per thread, so the
// 8 x 16-byte stores of a quarter warp fall in distinct banks; 128 epilogue threads -> 18 KB.
constexpr int kEpiStagePitch = 36;
constexpr int kEpiStageBytes = 128 * kEpiStagePitch * 4;

// Epilogue of one [128 x 256] accumulator tile: this thread's TMEM lane (= query row) x 256 columns,
// v = |y~|^2 - 2 x~.y~, threshold filter + warp compaction (topk.cuh).
//
// A warp owns 32 rows, so "some lane has a hit in this chunk" is true for almost every chunk (a row accepts
// ~64 ln(N/64) of N columns). The hit path therefore must not cost a 32-step serial chain per chunk: the compare is
// folded into the FMA pass as a bitmask (independent instructions), the values are parked in this thread's shared-memory
// row only when the mask is non-zero, and the push loop runs popcount(mask) (typically 1) iterations.
template <bool kProf = false>
__device__ __forceinline__ void knn_epilogue_tile(RowFilter& st, uint32_t t_addr, const float* yn, uint32_t col_base,
                                                  float* vrow, long long* ld_cycles = nullptr,
                                                  long long* push_cycles = nullptr) {
#pragma unroll 1
  for (int chunk = 0; chunk < kGemmBN / 32; ++chunk) {
    uint32_t r[32];
    long long c0 = 0;
    if (kProf) c0 = clock64();
    tmem_ld_32x32(t_addr + chunk * 32, r);
    tmem_ld_wait();
    if (kProf) *ld_cycles += clock64() - c0;
    float v[32];
    uint32_t mask = 0;
    const float tau = st.tau;
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      const float4 y4 = *reinterpret_cast<const float4*>(yn + chunk * 32 + j);
      v[j + 0] = fmaf(-2.f, __uint_as_float(r[j + 0]), y4.x);
      v[j + 1] = fmaf(-2.f, __uint_as_float(r[j + 1]), y4.y);
      v[j + 2] = fmaf(-2.f, __uint_as_float(r[j + 2]), y4.z);
      v[j + 3] = fmaf(-2.f, __uint_as_float(r[j + 3]), y4.w);
      mask |= (v[j + 0] < tau ? 1u : 0u) << (j + 0);
      mask |= (v[j + 1] < tau ? 1u : 0u) << (j + 1);
      mask |= (v[j + 2] < tau ? 1u : 0u) << (j + 2);
      mask |= (v[j + 3] < tau ? 1u : 0u) << (j + 3);
    }
    if (kProf) c0 = clock64();
    if (mask != 0u) {
#pragma unroll
      for (int j = 0; j < 32; j += 4)
        *reinterpret_cast<float4*>(vrow + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
      const uint32_t col0 = col_base + chunk * 32;
      uint64_t* dst = st.buf + st.cnt;
      st.cnt += __popc(mask);
      while (mask != 0u) {
        const int j = __ffs(mask) - 1;
        mask &= mask - 1u;
        __stcg(reinterpret_cast<unsigned long long*>(dst++), make_key(vrow[j], col0 + j));
      }
    }
    filter_compact_warp(st, 32);
    if (kProf) *push_cycles += clock64() - c0;
  }
}

}  // namespace vqacx

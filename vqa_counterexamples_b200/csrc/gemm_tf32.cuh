// Internal (non-ABI) interface of the generic TF32 tcgen05 GEMM used by every dense product of stage 2
// (NeuralCX forward / backward). See gemm_tf32.cu for the kernel.
//
//   C[M,N] = epilogue( sum over segments s of  A_s[M,K_s] . B_s[N,K_s]^T )
//
// Operands are float32 in global memory and are fed to the tensor cores as TF32 (kind::tf32) straight from
// the fp32 master copies: there are no low-precision shadow weights to keep in sync. Each operand of a segment
// is described by how it sits in memory:
//   k-major  : row-major [rows = M or N][cols = K]          (nn.Linear weights, activations in forward)
//   mn-major : row-major [rows = K][cols = M or N]          (the "transposed" reads of the backward pass:
//                                                            dW = dY^T . X reduces over the row index)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace vqacx {

constexpr int kGemmMaxSeg = 3;

enum GemmAct : int { kActNone = 0, kActTanh = 1, kActRelu = 2 };

struct GemmOperand {
  const float* ptr;   // first element of the (sub-)matrix
  long long ld;       // elements between consecutive rows in memory (multiple of 4)
  int mn_major;       // 0: memory rows are M/N, cols are K;  1: memory rows are K, cols are M/N
};

struct GemmSegment {
  GemmOperand a, b;
  int k;              // reduction length of this segment
};

struct GemmEpilogue {
  const float* bias = nullptr;       // [N], added once (by k-split 0)
  const float* row_add = nullptr;    // += row_add[(m / row_add_div) * row_add_ld + n]   (per-question shared term)
  int row_add_div = 1;
  long long row_add_ld = 0;
  const float* row_add2 = nullptr;   // += row_add2[(m % row_add2_mod) * row_add2_ld + n] (per-rank bias table)
  int row_add2_mod = 1;
  long long row_add2_ld = 0;
  const float* row_add3 = nullptr;   // += row_add3[row_add3_idx[m] * row_add3_ld + n]  (a table row gathered per output row:
  const int* row_add3_idx = nullptr; //    the per-IMAGE part of layer 1, v_other . W1vo^T, looked up by the candidate's image)
  long long row_add3_ld = 0;
  int act = kActNone;
  // forward dropout (inverted): out *= keep(m, n) / (1 - p); keep is a counter-based hash of (seed, m, n)
  float drop_p = 0.f;
  unsigned long long drop_seed = 0;
  const unsigned long long* drop_seed_dev = nullptr;  // optional device counter added to drop_seed at run time
  const float* drop_mask = nullptr;  // optional explicit mask [M, mask_ld] of 0 / 1/(1-p) values (parity tests)
  long long drop_mask_ld = 0;
  // backward gate: out = acc * (gate[m, n] > 0 ? gate_scale : 0)  (relu + dropout derivative from the saved output)
  const float* gate = nullptr;
  long long gate_ld = 0;
  float gate_scale = 1.f;
  float alpha = 1.f;                 // out = alpha * (...)
  // optional second copy of the output in fp16 ([M, half_ld] __half): feeds kind::f16 consumers (mutan_fused.cu)
  void* c_half = nullptr;
  long long half_ld = 0;
  int half_tail = 0;                 // > 0: also write c_half[m, N] = 1 and c_half[m, N+1 .. N+half_tail) = 0 (a constant-one
                                     // input column, so that a consumer can carry its bias as one more k column)
  int half_pad_to = 0;               // > N: c_half[m, N .. half_pad_to) = 0 (zero padding up to the consumer's k extent)
  // rank-1 term: += rvec[m] * cvec[n] before the activation (layer 1: the pairwise-distance feature times its one
  // weight column, so that the 1-wide feature never occupies a k-block of the product)
  const float* rvec = nullptr;
  const float* cvec = nullptr;
};

struct GemmProblem {
  int M = 0, N = 0;
  int nseg = 0;
  // 1: every operand is IEEE fp16 (the `ptr` fields point at __half data, ld counted in halfs and a multiple of 8),
  // k-major only; products run as kind::f16 with fp32 accumulation: half the operand bytes and twice the tensor rate
  // of the TF32 path. C may be null when only the fp16 copy of the output (epi.c_half) is wanted.
  int f16 = 0;
  GemmSegment seg[kGemmMaxSeg];
  float* C = nullptr;
  long long ldc = 0;
  // Split along K (grid.z). 1 = none; > 1 = that many splits; 0 = choose (only ever > 1 when `scratch` is given).
  //   with scratch   : every split stores its raw fp32 partial tile to scratch[z][M][ceil32(N)] with plain vector stores
  //                    and a second kernel sums the partials IN ORDER (deterministic) and applies the whole epilogue
  //                    (bias, row terms, activation, dropout, gate) -> small-M products fill the chip without giving up
  //                    the fused epilogue
  //   without scratch: partial sums are atomically added into C, which the caller zeroed (act must be none)
  int split_k = 1;
  float* scratch = nullptr;
  size_t scratch_floats = 0;
  int accumulate = 0;   // 1: C += result (atomic add), without zeroing
  int bn = 0;           // N tile (multiple of 32, <= 256); 0 = choose
  GemmEpilogue epi;
};

// Enqueue on `s`. Returns cudaSuccess, or cudaErrorInvalidValue for unsupported shapes / alignment.
cudaError_t launch_gemm_tf32(const GemmProblem& p, cudaStream_t s);

// The dropout keep-decision shared by forward epilogues and anything that must reproduce it.
__host__ __device__ inline uint32_t drop_hash(unsigned long long seed, uint32_t m, uint32_t n) {
  unsigned long long x = seed ^ (static_cast<unsigned long long>(m) << 32 | n);
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL;
  x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL;
  x ^= x >> 33;
  return static_cast<uint32_t>(x >> 8) & 0xffffffu;  // 24 uniform bits
}
__host__ __device__ inline bool drop_keep(unsigned long long seed, uint32_t m, uint32_t n, float p) {
  return static_cast<float>(drop_hash(seed, m, n)) * (1.f / 16777216.f) >= p;
}

}  // namespace vqacx

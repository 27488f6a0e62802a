// C-ABI entry points for stage 2 (NeuralCX scoring). Host-side orchestration only.
#include "../../include/vqacx.h"
#include "gemm_tf32.cuh"
#include "knn_internal.cuh"

using namespace vqacx;

#define CU_TRY(expr)                                      \
  do {                                                    \
    cudaError_t e__ = (expr);                             \
    if (e__ != cudaSuccess) return static_cast<int>(e__); \
  } while (0)

extern "C" {

int vqacx_gemm_tf32(int m, int n, int k, const float* a, int64_t lda, int a_mn_major, const float* b, int64_t ldb,
                    int b_mn_major, float* c, int64_t ldc, const float* bias, int act, int split_k, int bn,
                    int accumulate, void* stream) {
  if (m < 0 || n < 0 || k <= 0 || act < 0 || act > 2) return VQACX_EINVAL;
  if (m == 0 || n == 0) return VQACX_OK;
  if (!a || !b || !c) return VQACX_EINVAL;
  if ((lda % 4) || (ldb % 4) || (reinterpret_cast<uintptr_t>(a) & 15) || (reinterpret_cast<uintptr_t>(b) & 15))
    return VQACX_EALIGN;
  GemmProblem p;
  p.M = m; p.N = n; p.nseg = 1;
  p.seg[0].a = {a, lda, a_mn_major};
  p.seg[0].b = {b, ldb, b_mn_major};
  p.seg[0].k = k;
  p.C = c; p.ldc = ldc; p.split_k = split_k; p.bn = bn; p.accumulate = accumulate;
  p.epi.bias = bias; p.epi.act = act;
  const cudaError_t e = launch_gemm_tf32(p, static_cast<cudaStream_t>(stream));
  if (e == cudaErrorInvalidValue) return VQACX_EINVAL;
  CU_TRY(e);
  return VQACX_OK;
}

}  // extern "C"

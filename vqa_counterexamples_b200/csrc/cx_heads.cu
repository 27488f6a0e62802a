// Heads of the other CX models that share CXModelBase.vqa_forward (reference vqa/models/cx.py):
//   BlackBox         cx.py:114-136   score = -softmax(a_knn)[answer]
//   SemanticBaseline cx.py:159-205   score = softmax_over_knn( lam (S[answer,:].p - p[answer]) - (1-lam) log(p[answer]+1e-8) )
// plus the one-time cosine-similarity table S of the answer embedding (cx.py:172-175, sklearn cosine_similarity).
// The softmax probabilities p [R, n_ans] are the ones the shared front end (abi_ncx.cu vqa_front) leaves in the
// workspace; both heads are one pass over them: HBM/L2-bound, 4 * n_ans bytes per candidate.
#include "knn_internal.cuh"
#include "ncx_internal.cuh"

namespace vqacx {

namespace {

__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// One CTA per question, one warp per candidate (knn <= 32 warps).
__global__ void __launch_bounds__(1024) cx_head_kernel(const float* __restrict__ p, long long ld_p, int n_ans,
                                                       const long long* __restrict__ answer_aid, int knn, int head,
                                                       const float* __restrict__ sim, long long ld_sim, float lam,
                                                       float* __restrict__ scores) {
  __shared__ double raw[32];
  const int b = blockIdx.x, w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long aid = answer_aid[b];
  if (w < knn) {
    const float* row = p + (static_cast<long long>(b) * knn + w) * ld_p;
    const float pa = row[aid];
    if (head == kHeadBlackBox) {
      // highest VQA probability of the original answer = worst counterexample: flip the sign (cx.py:131-134)
      if (lane == 0) scores[static_cast<long long>(b) * knn + w] = -pa;
    } else {
      // float64 like the reference's numpy loop: emb_pairs[aid, :].dot(nb) - nb[aid]   (cx.py:190-193)
      const float* srow = sim + aid * ld_sim;
      double acc = 0.0;
      for (int i = lane; i < n_ans; i += 32) acc = fma(static_cast<double>(__ldg(srow + i)), static_cast<double>(row[i]), acc);
      acc = warp_sum_f64(acc);
      const double weighted = acc - static_cast<double>(pa);
      const double logp = log(static_cast<double>(pa) + 1e-8);
      if (lane == 0) raw[w] = static_cast<double>(lam) * weighted - (1.0 - static_cast<double>(lam)) * logp;
    }
  }
  if (head == kHeadBlackBox) return;
  __syncthreads();
  if (w == 0) {
    // softmax over the knn raw scores WITHOUT max subtraction, as the reference's helper does (cx.py:177-180)
    const double e = lane < knn ? exp(raw[lane]) : 0.0;
    const double tot = warp_sum_f64(e);
    if (lane < knn) scores[static_cast<long long>(b) * knn + lane] = static_cast<float>(e / tot);
  }
}

// S[i, j] = <x_i, x_j> / (max(|x_i|, tiny) max(|x_j|, tiny)) with float64 accumulation; a zero row gives S = 0 like
// sklearn's normalize (which leaves zero rows untouched). 32 x 32 output tile per CTA, 32-deep K slabs in shared memory.
__global__ void __launch_bounds__(1024) cx_cosine_kernel(const float* __restrict__ x, int n, int d, long long ld,
                                                         float* __restrict__ out, long long ld_out) {
  __shared__ float sa[32][33], sb[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int i0 = blockIdx.y * 32, j0 = blockIdx.x * 32;
  double dot = 0.0, na = 0.0, nb = 0.0;
  for (int k0 = 0; k0 < d; k0 += 32) {
    const int k = k0 + tx;
    sa[ty][tx] = (i0 + ty < n && k < d) ? x[static_cast<long long>(i0 + ty) * ld + k] : 0.f;
    sb[ty][tx] = (j0 + ty < n && k < d) ? x[static_cast<long long>(j0 + ty) * ld + k] : 0.f;
    __syncthreads();
#pragma unroll 8
    for (int kk = 0; kk < 32; ++kk) {
      const double a = sa[ty][kk], b = sb[tx][kk];
      dot = fma(a, b, dot);
      na = fma(a, a, na);
      nb = fma(b, b, nb);
    }
    __syncthreads();
  }
  const int i = i0 + ty, j = j0 + tx;
  if (i < n && j < n) {
    const double den = (na > 0.0 ? sqrt(na) : 1.0) * (nb > 0.0 ? sqrt(nb) : 1.0);
    out[static_cast<long long>(i) * ld_out + j] = static_cast<float>(dot / den);
  }
}

}  // namespace

cudaError_t launch_cx_head(const float* p, long long ld_p, int n_ans, const long long* answer_aid, int batch, int knn,
                           int head, const float* sim, long long ld_sim, float lam, float* scores, cudaStream_t s) {
  if (batch <= 0) return cudaSuccess;
  if (knn < 1 || knn > 32) return cudaErrorInvalidValue;
  count_launch();
  cx_head_kernel<<<batch, 32 * knn, 0, s>>>(p, ld_p, n_ans, answer_aid, knn, head, sim, ld_sim, lam, scores);
  return cudaGetLastError();
}

cudaError_t launch_cx_cosine(const float* x, int n, int d, long long ld, float* out, long long ld_out, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  count_launch();
  const dim3 grid((n + 31) / 32, (n + 31) / 32);
  cx_cosine_kernel<<<grid, 1024, 0, s>>>(x, n, d, ld, out, ld_out);
  return cudaGetLastError();
}

}  // namespace vqacx

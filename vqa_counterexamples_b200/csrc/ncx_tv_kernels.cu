// Elementwise kernels of the TRAINABLE-VQA variant of the NeuralCX step (cx.py:73-74, options/cx/
// neuralcx_trainable_vqa.yaml): the MUTAN head in train mode (fusion.py:84-121 with dropout_v / dropout_q, noatt.py:24-29
// with the classifier dropout) and its backward. The products are gemm_tf32 calls (abi_ncx.cu); these kernels are the
// glue between them. Rows are in the library's order: [batch * knn candidate rows | batch original-image rows].
#include "gemm_tf32.cuh"
#include "knn_internal.cuh"
#include "ncx_internal.cuh"

namespace vqacx {

namespace {

constexpr int kThreads = 256;
inline unsigned grid_for(long long items) {
  const long long b = (items + kThreads - 1) / kThreads;
  return static_cast<unsigned>(b < 1 ? 1 : (b > 148 * 16 ? 148 * 16 : b));
}

// source row of a per-question tensor for library row `row`
__device__ __forceinline__ int question_of(int row, int n_cand_rows, int knn) {
  return row < n_cand_rows ? row / knn : row - n_cand_rows;
}

__device__ __forceinline__ float keep_scale(const float* mask, long long mask_ld, unsigned long long seed, int row, int c,
                                            float p) {
  if (mask) return __ldg(mask + static_cast<long long>(row) * mask_ld + c);
  if (p <= 0.f) return 1.f;
  return drop_keep(seed, static_cast<uint32_t>(row), static_cast<uint32_t>(c), p) ? 1.f / (1.f - p) : 0.f;
}

// dst[row, c] = src[src_row(row), c] * keep(row, c) / (1 - p); columns [cols, cols_pad) are zeroed
__global__ void __launch_bounds__(kThreads) tv_dropout_rows_kernel(const float* __restrict__ src, long long ld_src,
                                                                   int per_question, int n_cand_rows, int knn,
                                                                   float* __restrict__ dst, long long ld_dst, int rows,
                                                                   int cols, int cols_pad, float p,
                                                                   unsigned long long seed, const float* __restrict__ mask,
                                                                   float* __restrict__ mask_out) {
  const long long total = static_cast<long long>(rows) * cols_pad;
  for (long long i = blockIdx.x * static_cast<long long>(kThreads) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * kThreads) {
    const int row = static_cast<int>(i / cols_pad), c = static_cast<int>(i % cols_pad);
    float v = 0.f, k = 0.f;
    if (c < cols) {
      const int sr = per_question ? question_of(row, n_cand_rows, knn) : row;
      k = keep_scale(mask, cols, seed, row, c, p);
      v = __ldg(src + static_cast<long long>(sr) * ld_src + c) * k;
    }
    dst[static_cast<long long>(row) * ld_dst + c] = v;
    if (mask_out) mask_out[static_cast<long long>(row) * ld_dst + c] = k;
  }
}

// z[row, c] = sum_r hv[row, r * dim_mm + c] * hq[row, r * dim_mm + c]  (fusion.py:103-116 with a per-row question side)
__global__ void __launch_bounds__(kThreads) tv_mutan_rows_kernel(const float* __restrict__ hv, const float* __restrict__ hq,
                                                                 long long ld_h, int rows, int n_cand_rows, int R,
                                                                 int dim_mm, int dim_mmp, float* __restrict__ z,
                                                                 long long ld_z, float* __restrict__ z_orig,
                                                                 long long ld_zo) {
  const long long total = static_cast<long long>(rows) * dim_mmp;
  for (long long i = blockIdx.x * static_cast<long long>(kThreads) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * kThreads) {
    const int row = static_cast<int>(i / dim_mmp), c = static_cast<int>(i % dim_mmp);
    float acc = 0.f;
    if (c < dim_mm) {
      const float* a = hv + static_cast<long long>(row) * ld_h + c;
      const float* b = hq + static_cast<long long>(row) * ld_h + c;
      for (int r = 0; r < R; ++r) acc = fmaf(a[r * dim_mm], b[r * dim_mm], acc);
    }
    if (row < n_cand_rows) z[static_cast<long long>(row) * ld_z + c] = acc;
    else z_orig[static_cast<long long>(row - n_cand_rows) * ld_zo + c] = acc;
  }
}

// in place: hv <- dz * hq, hq <- dz * hv  (d hv_r = dz . hq_r, d hq_r = dz . hv_r)
__global__ void __launch_bounds__(kThreads) tv_mutan_bwd_kernel(float* __restrict__ hv, float* __restrict__ hq,
                                                                long long ld_h, const float* __restrict__ dz,
                                                                long long ld_dz, int rows, int R, int dim_mm) {
  const long long total = static_cast<long long>(rows) * R * dim_mm;
  for (long long i = blockIdx.x * static_cast<long long>(kThreads) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * kThreads) {
    const int row = static_cast<int>(i / (R * dim_mm)), rc = static_cast<int>(i % (R * dim_mm)), c = rc % dim_mm;
    const float g = dz[static_cast<long long>(row) * ld_dz + c];
    const long long o = static_cast<long long>(row) * ld_h + rc;
    const float a = hv[o], b = hq[o];
    hv[o] = g * b;
    hq[o] = g * a;
  }
}

// in place on dp [rows, ld]: dA = P * (dP - sum_n dP P)   (backward of F.softmax, cx.py:281); one warp per row
__global__ void __launch_bounds__(kThreads) tv_softmax_bwd_kernel(const float* __restrict__ p, float* __restrict__ dp,
                                                                  long long ld, int rows, int n) {
  const int row = blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float* pr = p + static_cast<long long>(row) * ld;
  float* dr = dp + static_cast<long long>(row) * ld;
  float dot = 0.f;
  for (int c = lane; c < n; c += 32) dot = fmaf(pr[c], dr[c], dot);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
  for (int c = lane; c < n; c += 32) dr[c] = pr[c] * (dr[c] - dot);
}

// d[row, c] *= 1 - x[row, c]^2   (backward of tanh from its output)
__global__ void __launch_bounds__(kThreads) tv_tanh_bwd_kernel(float* __restrict__ d, long long ld_d,
                                                               const float* __restrict__ x, long long ld_x, int rows,
                                                               int cols) {
  const long long total = static_cast<long long>(rows) * cols;
  for (long long i = blockIdx.x * static_cast<long long>(kThreads) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * kThreads) {
    const int row = static_cast<int>(i / cols), c = static_cast<int>(i % cols);
    const float xv = x[static_cast<long long>(row) * ld_x + c];
    d[static_cast<long long>(row) * ld_d + c] *= 1.f - xv * xv;
  }
}

// acc[row, c] += add[row, c] * scale[row, c]
__global__ void __launch_bounds__(kThreads) tv_add_scaled_kernel(float* __restrict__ acc, long long ld_acc,
                                                                 const float* __restrict__ add, const float* __restrict__ scale,
                                                                 long long ld_add, int rows, int cols) {
  const long long total = static_cast<long long>(rows) * cols;
  for (long long i = blockIdx.x * static_cast<long long>(kThreads) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * kThreads) {
    const int row = static_cast<int>(i / cols), c = static_cast<int>(i % cols);
    const long long o = static_cast<long long>(row) * ld_add + c;
    acc[static_cast<long long>(row) * ld_acc + c] += add[o] * scale[o];
  }
}

// dq[b, c] += sum over the knn + 1 rows of question b of dxq0[row, c] * keep(row, c) / (1 - p)
__global__ void __launch_bounds__(kThreads) tv_dq_reduce_kernel(const float* __restrict__ dxq0, long long ld_d, int batch,
                                                                int knn, int cols, float p, unsigned long long seed,
                                                                const float* __restrict__ mask, float* __restrict__ dq,
                                                                long long ld_dq) {
  const long long total = static_cast<long long>(batch) * cols;
  const int n_cand_rows = batch * knn;
  for (long long i = blockIdx.x * static_cast<long long>(kThreads) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * kThreads) {
    const int b = static_cast<int>(i / cols), c = static_cast<int>(i % cols);
    float acc = 0.f;
    for (int j = 0; j <= knn; ++j) {
      const int row = j < knn ? b * knn + j : n_cand_rows + b;
      acc = fmaf(dxq0[static_cast<long long>(row) * ld_d + c], keep_scale(mask, cols, seed, row, c, p), acc);
    }
    dq[static_cast<long long>(b) * ld_dq + c] += acc;
  }
}

}  // namespace

#define TV_LAUNCH(kernel, items, ...)                                  \
  do {                                                                 \
    if ((items) <= 0) return cudaSuccess;                              \
    count_launch();                                                    \
    kernel<<<grid_for(items), kThreads, 0, s>>>(__VA_ARGS__);          \
    return cudaGetLastError();                                         \
  } while (0)

cudaError_t launch_tv_dropout_rows(const float* src, long long ld_src, int per_question, int n_cand_rows, int knn,
                                   float* dst, long long ld_dst, int rows, int cols, int cols_pad, float p,
                                   unsigned long long seed, const float* mask, float* mask_out, cudaStream_t s) {
  TV_LAUNCH(tv_dropout_rows_kernel, static_cast<long long>(rows) * cols_pad, src, ld_src, per_question, n_cand_rows, knn,
            dst, ld_dst, rows, cols, cols_pad, p, seed, mask, mask_out);
}
cudaError_t launch_tv_mutan_rows(const float* hv, const float* hq, long long ld_h, int rows, int n_cand_rows, int R,
                                 int dim_mm, int dim_mmp, float* z, long long ld_z, float* z_orig, long long ld_zo,
                                 cudaStream_t s) {
  TV_LAUNCH(tv_mutan_rows_kernel, static_cast<long long>(rows) * dim_mmp, hv, hq, ld_h, rows, n_cand_rows, R, dim_mm,
            dim_mmp, z, ld_z, z_orig, ld_zo);
}
cudaError_t launch_tv_mutan_bwd(float* hv, float* hq, long long ld_h, const float* dz, long long ld_dz, int rows, int R,
                                int dim_mm, cudaStream_t s) {
  TV_LAUNCH(tv_mutan_bwd_kernel, static_cast<long long>(rows) * R * dim_mm, hv, hq, ld_h, dz, ld_dz, rows, R, dim_mm);
}
cudaError_t launch_tv_softmax_bwd(const float* p, float* dp, long long ld, int rows, int n, cudaStream_t s) {
  if (rows <= 0) return cudaSuccess;
  count_launch();
  tv_softmax_bwd_kernel<<<(rows + kThreads / 32 - 1) / (kThreads / 32), kThreads, 0, s>>>(p, dp, ld, rows, n);
  return cudaGetLastError();
}
cudaError_t launch_tv_tanh_bwd(float* d, long long ld_d, const float* x, long long ld_x, int rows, int cols,
                               cudaStream_t s) {
  TV_LAUNCH(tv_tanh_bwd_kernel, static_cast<long long>(rows) * cols, d, ld_d, x, ld_x, rows, cols);
}
cudaError_t launch_tv_add_scaled(float* acc, long long ld_acc, const float* add, const float* scale, long long ld_add,
                                 int rows, int cols, cudaStream_t s) {
  TV_LAUNCH(tv_add_scaled_kernel, static_cast<long long>(rows) * cols, acc, ld_acc, add, scale, ld_add, rows, cols);
}
cudaError_t launch_tv_dq_reduce(const float* dxq0, long long ld_d, int batch, int knn, int cols, float p,
                                unsigned long long seed, const float* mask, float* dq, long long ld_dq, cudaStream_t s) {
  TV_LAUNCH(tv_dq_reduce_kernel, static_cast<long long>(batch) * cols, dxq0, ld_d, batch, knn, cols, p, seed, mask, dq,
            ld_dq);
}

}  // namespace vqacx

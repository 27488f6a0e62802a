// Stage-2 glue kernels (everything of NeuralCX that is not a dense product). All are HBM- or latency-bound:
// coalesced float4 traffic, one pass over their data, no host synchronisation.
#include "knn_internal.cuh"
#include "ncx_internal.cuh"
#include "ptx.cuh"

namespace vqacx {

namespace {
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
// block-wide sum / max for blockDim.x = 256 (8 warps); `red` is 8 floats of shared memory
__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) t += red[i];
  return t;
}
__device__ __forceinline__ float block_max(float v, float* red) {
  v = warp_max(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = red[0];
#pragma unroll
  for (int i = 1; i < 8; ++i) t = fmaxf(t, red[i]);
  return t;
}
}  // namespace

// ------------------------------------------------------------------------------------------------
// feature construction: gather of the 25 image rows of a question + v_mult + v_dist (cx.py:289-303) and the
// per-question rows (v_orig, q_emb, E[answer]; cx.py:280). One block per (question, image slot).
// Replaces getDataFromBatch's host gather + H2D copy (counterexamples.py:540-541) when `features` is the
// device-resident feature matrix.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ncx_features_kernel(const FeatArgs a) {
  pdl_enter();   // programmatic dependent launch: see ptx.cuh
  __shared__ float red[8];
  const int b = blockIdx.x / (a.knn + 1), slot = blockIdx.x % (a.knn + 1);
  const int tid = threadIdx.x;
  const float4* vo = reinterpret_cast<const float4*>(a.features + static_cast<long long>(a.img_idx[b * (a.knn + 1)]) * a.ld_features);
  const int nv = a.dim_v / 4;
  if (slot == 0) {
    float* row = a.xs + static_cast<long long>(b) * a.ld_xs;
    float4* d = reinterpret_cast<float4*>(row);
    // the original image's features also go to row (batch*knn + b) of the candidate matrix, so that the MUTAN visual
    // products run once over [24 B candidate rows | B original rows] instead of twice
    float* row2 = a.xc + (static_cast<long long>(a.batch) * a.knn + b) * a.ld_xc;
    float4* d2 = reinterpret_cast<float4*>(row2);
    for (int i = tid; i < nv; i += 256) {
      const float4 t = __ldg(vo + i);
      d[i] = t;
      d2[i] = t;
    }
    for (int i = a.dim_v + tid; i < a.dim_vp; i += 256) {
      row[i] = 0.f;
      row2[i] = 0.f;
    }
    const float4* q = reinterpret_cast<const float4*>(a.q_emb + static_cast<long long>(b) * a.ld_q);
    float4* dq = reinterpret_cast<float4*>(row + a.col_q);
    for (int i = tid; i < a.dim_q / 4; i += 256) dq[i] = __ldg(q + i);
    for (int i = a.dim_q + tid; i < a.dim_qp; i += 256) row[a.col_q + i] = 0.f;
    if (a.emb) {   // null for the heads without an answer embedding (BlackBox, SemanticBaseline, LinearContext)
      const float4* e = reinterpret_cast<const float4*>(a.emb + a.answer_aid[b] * static_cast<long long>(a.dim_a));
      float4* de = reinterpret_cast<float4*>(row + a.col_agt);
      for (int i = tid; i < a.dim_a / 4; i += 256) de[i] = __ldg(e + i);
    }
    return;
  }
  const long long r = static_cast<long long>(b) * a.knn + (slot - 1);
  const float4* vk = reinterpret_cast<const float4*>(a.features + static_cast<long long>(a.img_idx[b * (a.knn + 1) + slot]) * a.ld_features);
  float* row = a.xc + r * a.ld_xc;
  float4* d_other = reinterpret_cast<float4*>(row);
  float4* d_mult = reinterpret_cast<float4*>(row + a.col_vmult);
  float acc = 0.f;
  for (int i = tid; i < nv; i += 256) {
    const float4 o = __ldg(vo + i), k = __ldg(vk + i);
    d_other[i] = k;
    d_mult[i] = a.no_vmult ? make_float4(0.f, 0.f, 0.f, 0.f) : make_float4(o.x * k.x, o.y * k.y, o.z * k.z, o.w * k.w);
    // F.pairwise_distance: || x1 - x2 + eps ||_2 (cx.py:300)
    const float dx = o.x - k.x + a.eps, dy = o.y - k.y + a.eps, dz = o.z - k.z + a.eps, dw = o.w - k.w + a.eps;
    acc += dx * dx + dy * dy + dz * dz + dw * dw;
  }
  for (int i = a.dim_v + tid; i < a.dim_vp; i += 256) {
    row[i] = 0.f;
    row[a.col_vmult + i] = 0.f;
  }
  const float tot = block_sum(acc, red);
  if (tid < a.dist_pad) row[a.col_dist + tid] = (tid == 0 && !a.no_vdist) ? sqrtf(tot) : 0.f;
}

cudaError_t launch_ncx_features(const FeatArgs& a, cudaStream_t s) {
  if (a.batch <= 0) return cudaSuccess;
  count_launch();
  (void)launch_pdl(ncx_features_kernel, dim3(a.batch * (a.knn + 1)), dim3(256), 0, s, a);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// MUTAN rank reduction (fusion.py:99-113): z = sum_r (W_hq^r x_q + b) * (W_hv^r x_v + b)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) ncx_mutan_reduce_kernel(const float* __restrict__ hv, long long ld_hv,
                                                               const float* __restrict__ hq, long long ld_hq,
                                                               int rows_per_q, int n_cand_rows, int R, int dim_mm,
                                                               int dim_mmp, float* __restrict__ z, long long ld_z,
                                                               float* __restrict__ z_orig, long long ld_zo) {
  // rows [0, n_cand_rows): candidate rows (question = row / rows_per_q) -> z; rows beyond: one original image per
  // question -> z_orig
  const long long row = blockIdx.x;
  const bool cand = row < n_cand_rows;
  const long long qi = cand ? row / rows_per_q : row - n_cand_rows;
  const float* v = hv + row * ld_hv;
  const float* q = hq + qi * ld_hq;
  float* out = cand ? z + row * ld_z : z_orig + qi * ld_zo;
  for (int j = threadIdx.x; j < dim_mmp; j += blockDim.x) {
    float acc = 0.f;
    if (j < dim_mm)
      for (int r = 0; r < R; ++r) acc = fmaf(__ldcs(v + r * dim_mm + j), __ldg(q + r * dim_mm + j), acc);
    out[j] = acc;
  }
}
cudaError_t launch_ncx_mutan_reduce(const float* hv, long long ld_hv, const float* hq, long long ld_hq, int rows,
                                    int rows_per_q, int n_cand_rows, int R, int dim_mm, int dim_mmp, float* z,
                                    long long ld_z, float* z_orig, long long ld_zo, cudaStream_t s) {
  if (rows <= 0) return cudaSuccess;
  count_launch();
  ncx_mutan_reduce_kernel<<<rows, 128, 0, s>>>(hv, ld_hv, hq, ld_hq, rows_per_q, n_cand_rows, R, dim_mm, dim_mmp, z,
                                               ld_z, z_orig, ld_zo);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// softmax over the answer logits (cx.py:281), in place, one block per candidate row
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ncx_softmax_kernel(float* __restrict__ x, long long ld, int n, int n_pad) {
  pdl_enter();   // programmatic dependent launch: see ptx.cuh
  __shared__ float red[8];
  float* row = x + static_cast<long long>(blockIdx.x) * ld;
  float mx = -INFINITY;
  for (int i = threadIdx.x; i < n; i += 256) mx = fmaxf(mx, row[i]);
  mx = block_max(mx, red);
  float sum = 0.f;
  for (int i = threadIdx.x; i < n; i += 256) {
    const float e = expf(row[i] - mx);
    row[i] = e;
    sum += e;
  }
  sum = block_sum(sum, red);
  const float inv = 1.f / sum;
  for (int i = threadIdx.x; i < n_pad; i += 256) row[i] = i < n ? row[i] * inv : 0.f;
}
// Rows of up to 2048 logits (the 2000 answers): one WARP per row, the whole row in registers (16 float4 per lane, all
// loads issued before the first use), shuffle reductions only -- one read and one write of the row, no block barriers.
__global__ void __launch_bounds__(256) ncx_softmax_warp_kernel(float* __restrict__ x, long long ld, int rows, int n,
                                                               int n_pad) {
  pdl_enter();   // programmatic dependent launch: see ptx.cuh
  const int lane = threadIdx.x & 31;
  const long long r = static_cast<long long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (r >= rows) return;
  float4* row = reinterpret_cast<float4*>(x + r * ld);
  const int n4 = n_pad / 4;                  // n_pad is a multiple of 32
  float4 v[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const int i = j * 32 + lane;
    v[j] = i < n4 ? row[i] : make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
  }
  float mx = -INFINITY;
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const int c = (j * 32 + lane) * 4;
    if (c + 0 >= n) v[j].x = -INFINITY;
    if (c + 1 >= n) v[j].y = -INFINITY;
    if (c + 2 >= n) v[j].z = -INFINITY;
    if (c + 3 >= n) v[j].w = -INFINITY;
    mx = fmaxf(mx, fmaxf(fmaxf(v[j].x, v[j].y), fmaxf(v[j].z, v[j].w)));
  }
  mx = warp_max(mx);
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    v[j].x = expf(v[j].x - mx); v[j].y = expf(v[j].y - mx); v[j].z = expf(v[j].z - mx); v[j].w = expf(v[j].w - mx);
    sum += (v[j].x + v[j].y) + (v[j].z + v[j].w);
  }
  sum = warp_sum(sum);
  const float inv = 1.f / sum;
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const int i = j * 32 + lane;
    if (i < n4) row[i] = make_float4(v[j].x * inv, v[j].y * inv, v[j].z * inv, v[j].w * inv);   // exp(-inf) = 0 pads
  }
}

cudaError_t launch_ncx_softmax(float* x, long long ld, int rows, int n, int n_pad, cudaStream_t s) {
  if (rows <= 0) return cudaSuccess;
  count_launch();
  if (n_pad <= 2048 && (n_pad % 4) == 0 && (ld % 4) == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0)
    (void)launch_pdl(ncx_softmax_warp_kernel, dim3((rows + 7) / 8), dim3(256), 0, s, x, ld, rows, n, n_pad);
  else
    (void)launch_pdl(ncx_softmax_kernel, dim3(rows), dim3(256), 0, s, x, ld, n, n_pad);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// transposed column slices of linear_1.weight (the 24 one-hot rank columns act as a per-rank bias table)
// ------------------------------------------------------------------------------------------------
__global__ void ncx_slice_t_kernel(const float* __restrict__ w, long long ld_w, int rows, int col0, int cnt,
                                   float* __restrict__ out, long long ld_out) {
  pdl_enter();   // programmatic dependent launch: see ptx.cuh
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= rows) return;
  for (int i = 0; i < cnt; ++i) out[i * ld_out + n] = w[n * ld_w + col0 + i];
}
__global__ void ncx_unslice_t_kernel(const float* __restrict__ in, long long ld_in, int rows, int col0, int cnt,
                                     float* __restrict__ w, long long ld_w) {
  pdl_enter();   // programmatic dependent launch: see ptx.cuh
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= rows) return;
  for (int i = 0; i < cnt; ++i) w[n * ld_w + col0 + i] = in[i * ld_in + n];
}
cudaError_t launch_ncx_slice_t(const float* w, long long ld_w, int rows, int col0, int cnt, float* out, long long ld_out,
                               cudaStream_t s) {
  count_launch();
  (void)launch_pdl(ncx_slice_t_kernel, dim3((rows + 127) / 128), dim3(128), 0, s, w, ld_w, rows, col0, cnt, out, ld_out);
  return cudaGetLastError();
}
cudaError_t launch_ncx_unslice_t(const float* in, long long ld_in, int rows, int col0, int cnt, float* w, long long ld_w,
                                 cudaStream_t s) {
  count_launch();
  (void)launch_pdl(ncx_unslice_t_kernel, dim3((rows + 127) / 128), dim3(128), 0, s, in, ld_in, rows, col0, cnt, w, ld_w);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// out layer (cx.py:327): one warp per candidate row
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ncx_score_kernel(const float* __restrict__ h, long long ld_h, int rows, int n,
                                                        const float* __restrict__ w, const float* __restrict__ b,
                                                        float* __restrict__ scores) {
  pdl_enter();   // programmatic dependent launch: see ptx.cuh
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float* hr = h + static_cast<long long>(row) * ld_h;
  float acc = 0.f;
  for (int i = lane; i < n; i += 32) acc = fmaf(hr[i], __ldg(w + i), acc);
  acc = warp_sum(acc);
  if (lane == 0) scores[row] = acc + __ldg(b);
}
cudaError_t launch_ncx_score(const float* h, long long ld_h, int rows, int n, const float* w, const float* b,
                             float* scores, cudaStream_t s) {
  if (rows <= 0) return cudaSuccess;
  count_launch();
  (void)launch_pdl(ncx_score_kernel, dim3((rows + 7) / 8), dim3(256), 0, s, h, ld_h, rows, n, w, b, scores);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// K3: fused 24-way softmax cross-entropy + gradient + Recall@1/5. One warp per question, lane = candidate.
// Replaces nn.CrossEntropyLoss(size_average=False) (counterexamples.py:310,334), its backward, and recallAtK
// (counterexamples.py:501-506) including that function's device->host round trip per step.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ncx_ce_recall_kernel(const float* __restrict__ scores,
                                                            const long long* __restrict__ gt, long long batch, int knn,
                                                            float* __restrict__ loss_sum, float* __restrict__ dscores,
                                                            float grad_scale, int* __restrict__ hits) {
  pdl_enter();   // programmatic dependent launch: see ptx.cuh
  __shared__ float s_loss[8];
  __shared__ int s_h1[8], s_h5[8];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long b = static_cast<long long>(blockIdx.x) * 8 + w;
  float loss = 0.f;
  int h1 = 0, h5 = 0;
  if (b < batch) {
    const long long g64 = gt[b];
    const bool g_ok = g64 >= 0 && g64 < knn;   // a label outside [0, knn) poisons the loss (NaN) and never counts as a hit
    const int g = g_ok ? static_cast<int>(g64) : 0;
    const float sc = lane < knn ? scores[b * knn + lane] : -INFINITY;
    const float mx = warp_max(sc);
    const float e = lane < knn ? expf(sc - mx) : 0.f;
    const float sum = warp_sum(e);
    const float sg = __shfl_sync(0xffffffffu, sc, g);
    loss = g_ok ? logf(sum) + mx - sg : __int_as_float(0x7fc00000);
    if (dscores && lane < knn)
      dscores[b * knn + lane] = g_ok ? grad_scale * (e / sum - (lane == g ? 1.f : 0.f)) : 0.f;
    const bool ahead = lane < knn && (sc > sg || (sc == sg && lane < g));
    const int rank = __popc(__ballot_sync(0xffffffffu, ahead));
    // NaN scores make every comparison false (rank 0): a diverged model must not report Recall 1.0
    const bool countable = g_ok && isfinite(sg) && !__any_sync(0xffffffffu, lane < knn && sc != sc);
    h1 = countable && rank < 1;
    h5 = countable && rank < 5;
  }
  if (lane == 0) { s_loss[w] = loss; s_h1[w] = h1; s_h5[w] = h5; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float l = 0.f; int a = 0, c = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) { l += s_loss[i]; a += s_h1[i]; c += s_h5[i]; }
    if (loss_sum) atomicAdd(loss_sum, l);
    if (hits) { atomicAdd(hits, a); atomicAdd(hits + 1, c); }
  }
}
cudaError_t launch_ncx_ce_recall(const float* scores, const long long* gt, long long batch, int knn, float* loss_sum,
                                 float* dscores, float grad_scale, int* hits, cudaStream_t s) {
  if (batch <= 0) return cudaSuccess;
  count_launch();
  (void)launch_pdl(ncx_ce_recall_kernel, dim3(static_cast<unsigned>((batch + 7) / 8)), dim3(256), 0, s, scores, gt, batch, knn, loss_sum, dscores,
                                                                            grad_scale, hits);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// backward glue
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ncx_dout_kernel(const float* __restrict__ dscores, const float* __restrict__ w_out,
                                                       const float* __restrict__ h, long long ld_h, int rows, int n,
                                                       float gate_scale, float* __restrict__ dpre, long long ld_d) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<long long>(rows) * n) return;
  const int r = static_cast<int>(i / n), c = static_cast<int>(i % n);
  const float hv = h[static_cast<long long>(r) * ld_h + c];
  dpre[static_cast<long long>(r) * ld_d + c] = hv > 0.f ? dscores[r] * __ldg(w_out + c) * gate_scale : 0.f;
}
cudaError_t launch_ncx_dout(const float* dscores, const float* w_out, const float* h, long long ld_h, int rows, int n,
                            float gate_scale, float* dpre, long long ld_d, cudaStream_t s) {
  if (rows <= 0) return cudaSuccess;
  count_launch();
  const long long tot = static_cast<long long>(rows) * n;
  ncx_dout_kernel<<<static_cast<unsigned>((tot + 255) / 256), 256, 0, s>>>(dscores, w_out, h, ld_h, rows, n, gate_scale, dpre, ld_d);
  return cudaGetLastError();
}

// out[n] += sum_r x[r, n] * w[r]: blocks of 32 columns x 128 rows, 8 row lanes, one atomicAdd per (block, column);
// `out` must be zeroed by the caller (vqacx_ncx_backward clears the small gradient tensors wholesale).
__global__ void __launch_bounds__(256) ncx_colsum_kernel(const float* __restrict__ x, long long ld, int rows, int n_cols,
                                                         const float* __restrict__ w, float* __restrict__ out) {
  pdl_enter();   // programmatic dependent launch: see ptx.cuh
  __shared__ float part[8][33];
  const int c = blockIdx.x * 32 + (threadIdx.x & 31), g = threadIdx.x >> 5;
  const int r0 = blockIdx.y * 128, r1 = min(rows, r0 + 128);
  float acc = 0.f;
  if (c < n_cols) {
#pragma unroll 4
    for (int r = r0 + g; r < r1; r += 8) acc = fmaf(x[static_cast<long long>(r) * ld + c], w ? __ldg(w + r) : 1.f, acc);
  }
  part[g][threadIdx.x & 31] = acc;
  __syncthreads();
  if (g == 0 && c < n_cols) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += part[i][threadIdx.x];
    atomicAdd(out + c, t);
  }
}
cudaError_t launch_ncx_colsum(const float* x, long long ld, int rows, int n_cols, const float* w, float* out,
                              cudaStream_t s) {
  if (n_cols <= 0 || rows <= 0) return cudaSuccess;
  count_launch();
  (void)launch_pdl(ncx_colsum_kernel, dim3(dim3((n_cols + 31) / 32, (rows + 127) / 128)), dim3(256), 0, s, x, ld, rows, n_cols, w, out);
  return cudaGetLastError();
}

// Backward of the `out` Linear and of the last hidden layer's relu/dropout in one pass over h_last [rows, n]
// (cx.py:329-331 reversed):  dpre[r, c] = h[r, c] > 0 ? dscores[r] * w_out[c] * gate_scale : 0
//   g_wout[c] += sum_r dscores[r] h[r, c]     g_bout += sum_r dscores[r]     g_blast[c] += sum_r dpre[r, c]
// (the three sums are atomically added into the zeroed gradient buffer; replaces colsum x3 + dout).
__global__ void __launch_bounds__(256) ncx_out_backward_kernel(const float* __restrict__ dscores,
                                                               const float* __restrict__ w_out,
                                                               const float* __restrict__ h, long long ld_h, int rows,
                                                               int n, float gate_scale, float* __restrict__ dpre,
                                                               long long ld_d, float* __restrict__ g_wout,
                                                               float* __restrict__ g_bout, float* __restrict__ g_blast) {
  pdl_enter();   // programmatic dependent launch: see ptx.cuh
  __shared__ float pw[8][33], pb[8][33], ps[8];
  const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  const int r0 = blockIdx.y * 128, r1 = min(rows, r0 + 128);
  const float wo = c < n ? __ldg(w_out + c) * gate_scale : 0.f;
  float sw = 0.f, sb = 0.f, sd = 0.f;
  if (c < n) {
#pragma unroll 4
    for (int r = r0 + g; r < r1; r += 8) {
      const float ds = __ldg(dscores + r);
      const float hv = h[static_cast<long long>(r) * ld_h + c];
      const float dp = hv > 0.f ? ds * wo : 0.f;
      dpre[static_cast<long long>(r) * ld_d + c] = dp;
      sw = fmaf(ds, hv, sw);
      sb += dp;
      sd += ds;
    }
  }
  pw[g][lane] = sw;
  pb[g][lane] = sb;
  if (lane == 0) ps[g] = sd;
  __syncthreads();
  if (g == 0 && c < n) {
    float tw = 0.f, tb = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { tw += pw[i][lane]; tb += pb[i][lane]; }
    atomicAdd(g_wout + c, tw);
    atomicAdd(g_blast + c, tb);
    if (blockIdx.x == 0 && lane == 0) {
      float td = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) td += ps[i];
      atomicAdd(g_bout, td);
    }
  }
}
cudaError_t launch_ncx_out_backward(const float* dscores, const float* w_out, const float* h, long long ld_h, int rows,
                                    int n, float gate_scale, float* dpre, long long ld_d, float* g_wout, float* g_bout,
                                    float* g_blast, cudaStream_t s) {
  if (rows <= 0 || n <= 0) return cudaSuccess;
  count_launch();
  (void)launch_pdl(ncx_out_backward_kernel, dim3(dim3((n + 31) / 32, (rows + 127) / 128)), dim3(256), 0, s, dscores, w_out, h, ld_h, rows, n,
                                                                               gate_scale, dpre, ld_d, g_wout, g_bout,
                                                                               g_blast);
  return cudaGetLastError();
}

// One pass over dpre1 [batch*knn, n]:  ds[b, :] = sum_i d[b*knn+i, :] (per-question term of layer 1),
// drank_t[i, :] += sum_b d[b*knn+i, :] (one-hot rank columns), g_b1[:] += sum of all rows (nullable).
// Blocks of 32 columns x 4 questions; drank_t and g_b1 must be zeroed by the caller.
__global__ void __launch_bounds__(256) ncx_group_reduce_kernel(const float* __restrict__ d, long long ld, int batch, int knn,
                                                               int n, float* __restrict__ ds, long long ld_ds,
                                                               float* __restrict__ drank_t, long long ld_dr,
                                                               float* __restrict__ g_b1) {
  pdl_enter();   // programmatic dependent launch: see ptx.cuh
  __shared__ float part[8][33];
  const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  const int q0 = blockIdx.y * 4, q1 = min(batch, q0 + 4);
  float rk[4] = {0.f, 0.f, 0.f, 0.f};
  float bsum = 0.f;
  for (int q = q0; q < q1; ++q) {
    float acc = 0.f;
    if (c < n) {
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const int i = g + 8 * t;
        if (i < knn) {
          const float v = d[(static_cast<long long>(q) * knn + i) * ld + c];
          rk[t] += v;
          acc += v;
        }
      }
    }
    part[g][lane] = acc;
    __syncthreads();
    if (g == 0 && c < n) {
      float t = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) t += part[i][lane];
      ds[static_cast<long long>(q) * ld_ds + c] = t;
      bsum += t;
    }
    __syncthreads();
  }
  if (c < n) {
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int i = g + 8 * t;
      if (i < knn) atomicAdd(drank_t + static_cast<long long>(i) * ld_dr + c, rk[t]);
    }
    if (g == 0 && g_b1) atomicAdd(g_b1 + c, bsum);
  }
}
cudaError_t launch_ncx_group_reduce(const float* d, long long ld, int batch, int knn, int n, float* ds, long long ld_ds,
                                    float* drank_t, long long ld_dr, float* g_b1, cudaStream_t s) {
  if (batch <= 0 || n <= 0) return cudaSuccess;
  count_launch();
  (void)launch_pdl(ncx_group_reduce_kernel, dim3(dim3((n + 31) / 32, (batch + 3) / 4)), dim3(256), 0, s, d, ld, batch, knn, n, ds, ld_ds, drank_t,
                                                                             ld_dr, g_b1);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) ncx_scatter_add_rows_kernel(const float* __restrict__ src, long long ld_src,
                                                                   const long long* __restrict__ idx, int n,
                                                                   float* __restrict__ dst, long long ld_dst) {
  pdl_enter();   // programmatic dependent launch: see ptx.cuh
  const int b = blockIdx.x;
  float* drow = dst + idx[b] * ld_dst;
  const float* srow = src + static_cast<long long>(b) * ld_src;
  for (int c = threadIdx.x; c < n; c += 256) atomicAdd(drow + c, srow[c]);
}
cudaError_t launch_ncx_scatter_add_rows(const float* src, long long ld_src, const long long* idx, int batch, int n,
                                        float* dst, long long ld_dst, cudaStream_t s) {
  if (batch <= 0) return cudaSuccess;
  count_launch();
  (void)launch_pdl(ncx_scatter_add_rows_kernel, dim3(batch), dim3(256), 0, s, src, ld_src, idx, n, dst, ld_dst);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// Adam (counterexamples.py:275-276 torch.optim.Adam(lr); :339 optimizer.step()) over the flat trainable buffer.
// HBM-bound: reads p, g, m, v and writes p, m, v = 28 B per parameter.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                   float* __restrict__ v, long long n4, long long n, float lr, float b1,
                                                   float b2, float eps, float bc1, float bc2, float gs,
                                                   const int* __restrict__ step_dev) {
  pdl_enter();   // programmatic dependent launch: see ptx.cuh
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  if (step_dev) {  // same double-precision bias corrections the host computes for vqacx_adam_step
    const double st = static_cast<double>(__ldg(step_dev));
    bc1 = static_cast<float>(1.0 - pow(static_cast<double>(b1), st));
    bc2 = static_cast<float>(1.0 - pow(static_cast<double>(b2), st));
  }
  const float step_size = lr / bc1;
  const float inv_sqrt_bc2 = rsqrtf(bc2);
  for (long long j = i; j < n4; j += stride) {
    float4 pp = reinterpret_cast<float4*>(p)[j];
    const float4 gg = reinterpret_cast<const float4*>(g)[j];
    float4 mm = reinterpret_cast<float4*>(m)[j], vv = reinterpret_cast<float4*>(v)[j];
    float* pa = &pp.x; const float* ga = &gg.x; float* ma = &mm.x; float* va = &vv.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float gr = ga[k] * gs;
      ma[k] = b1 * ma[k] + (1.f - b1) * gr;
      va[k] = b2 * va[k] + (1.f - b2) * gr * gr;
      pa[k] -= step_size * ma[k] / (sqrtf(va[k]) * inv_sqrt_bc2 + eps);
    }
    reinterpret_cast<float4*>(p)[j] = pp;
    reinterpret_cast<float4*>(m)[j] = mm;
    reinterpret_cast<float4*>(v)[j] = vv;
  }
  for (long long j = n4 * 4 + i; j < n; j += stride) {
    const float gr = g[j] * gs;
    m[j] = b1 * m[j] + (1.f - b1) * gr;
    v[j] = b2 * v[j] + (1.f - b2) * gr * gr;
    p[j] -= step_size * m[j] / (sqrtf(v[j]) * inv_sqrt_bc2 + eps);
  }
}
cudaError_t launch_adam(float* p, const float* g, float* m, float* v, long long n, float lr, float b1, float b2,
                        float eps, float bc1, float bc2, float grad_scale, const int* step_dev, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  count_launch();
  const long long n4 = n / 4;
  long long blocks = (n4 + 255) / 256;
  if (blocks > 148LL * 8) blocks = 148LL * 8;
  if (blocks < 1) blocks = 1;
  (void)launch_pdl(adam_kernel, dim3(static_cast<unsigned>(blocks)), dim3(256), 0, s, p, g, m, v, n4, n, lr, b1, b2, eps, bc1, bc2, grad_scale, step_dev);
  return cudaGetLastError();
}

__global__ void step_tick_kernel(int* step_dev, unsigned long long* seed_dev, unsigned long long seed_inc) {
  pdl_enter();   // programmatic dependent launch: see ptx.cuh
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    if (step_dev) step_dev[0] += 1;
    if (seed_dev) seed_dev[0] += seed_inc;
  }
}
cudaError_t launch_step_tick(int* step_dev, unsigned long long* seed_dev, unsigned long long seed_inc, cudaStream_t s) {
  count_launch();
  (void)launch_pdl(step_tick_kernel, dim3(1), dim3(32), 0, s, step_dev, seed_dev, seed_inc);
  return cudaGetLastError();
}

}  // namespace vqacx

// ------------------------------------------------------------------------------------------------
// Example builder join (vqacx_valset_builder.ipynb cell 13 / vqacx_trainset_builder.ipynb cell 18): for a question q1
// on image i1 whose complementary question q2 sits on image i2, keep the example iff i2 is one of i1's 24 nearest
// neighbours; `comp.knn_index` = knns[i1].index(i2). Consumes the stage-1 output directly: knn_idx [n_img, k1] with
// the image itself among its own k1 = 25 neighbours (knn.py -k 25). The 24-list is that row minus the image itself
// (minus the last entry if exact duplicates displaced it, as NearestNeighbors.kneighbors(None) does). Produces the
// img_idx [n_q, k1] / comp_idx [n_q] tensors vqacx_ncx_forward and vqacx_ce_recall take. One thread per question.
// ------------------------------------------------------------------------------------------------
namespace vqacx {
__global__ void __launch_bounds__(256) cx_build_examples_kernel(const long long* __restrict__ knn_idx, long long n_img,
                                                                int k1, const int* __restrict__ q_img,
                                                                const int* __restrict__ q_comp, long long n_q,
                                                                int* __restrict__ img_idx_out,
                                                                int* __restrict__ comp_idx_out) {
  const long long q = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (q >= n_q) return;
  const int i1 = q_img[q];
  const int q2 = q_comp[q];
  int comp = -1;
  int* out = img_idx_out + q * k1;
  if (i1 < 0 || i1 >= n_img) {
    for (int j = 0; j < k1; ++j) out[j] = -1;
    comp_idx_out[q] = -1;
    return;
  }
  // q2 must itself be a question with a complement (the notebook's "q2 in intersection")
  const int i2 = (q2 >= 0 && q2 < n_q && q_comp[q2] >= 0) ? q_img[q2] : -1;
  const long long* row = knn_idx + static_cast<long long>(i1) * k1;
  out[0] = i1;
  int w = 0;                       // neighbours written so far
  bool self_seen = false;
  for (int j = 0; j < k1 && w < k1 - 1; ++j) {
    const int nb = static_cast<int>(row[j]);
    if (!self_seen && nb == i1) { self_seen = true; continue; }
    out[1 + w] = nb;
    if (comp < 0 && nb == i2 && i2 >= 0) comp = w;
    ++w;
  }
  comp_idx_out[q] = comp;
}

cudaError_t launch_cx_build_examples(const long long* knn_idx, long long n_img, int k1, const int* q_img, const int* q_comp,
                                     long long n_q, int* img_idx_out, int* comp_idx_out, cudaStream_t s) {
  if (n_q <= 0) return cudaSuccess;
  count_launch();
  cx_build_examples_kernel<<<static_cast<unsigned>((n_q + 255) / 256), 256, 0, s>>>(knn_idx, n_img, k1, q_img, q_comp, n_q,
                                                                                  img_idx_out, comp_idx_out);
  return cudaGetLastError();
}
}  // namespace vqacx

// Internal (non-ABI) declarations of the stage-2 (NeuralCX) element-wise / reduction kernels.
// The dense products go through gemm_tf32.cuh; everything here is HBM- or latency-bound glue around them.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace vqacx {

// Per-candidate and per-question input rows of layer 1 (cx.py:289-320), written once per batch:
//   xs[b] = [ v_orig | q_emb | (z_orig: filled later) | E[answer] ]      xc[r = b*24+i] = [ v_other | v_orig*v_other |
//   (z_other: filled later) | ||v_orig - v_other + eps||_2 ]  with zeroed padding between the column groups.
struct FeatArgs {
  const float* features; long long ld_features;
  const int* img_idx;            // [batch, knn+1]
  const float* q_emb; long long ld_q;
  const float* emb;              // [n_ans, dim_a]
  const long long* answer_aid;   // [batch]
  float* xs; long long ld_xs;
  float* xc; long long ld_xc;
  int batch, knn, dim_v, dim_vp, dim_q, dim_qp, dim_a;
  int col_q, col_agt;            // in xs
  int col_vmult, col_dist, dist_pad;  // in xc
  float eps;
  int no_vmult, no_vdist;        // lesions: write zeros instead of v_orig * v_other / the pairwise distance
};
cudaError_t launch_ncx_features(const FeatArgs& a, cudaStream_t s);

// ---- fp16 fast path of the evaluation forward (ncx_fast_kernels.cu) ----------------------------------------------------
struct Feat16Args {
  const float* features; long long ld_features;
  const __half* features16; long long ld_f16;   // optional fp16 copy of `features` (per-image tables buffer): the 24
                                                // neighbour rows of a question are gathered from it at half the bytes
  const int* img_idx;            // [batch, knn+1]
  const float* q_emb; long long ld_q;
  const float* emb;              // [n_ans, dim_a]
  const long long* answer_aid;   // [batch]
  float* xs; long long ld_xs;    // per-question fp32 rows (same layout as FeatArgs::xs)
  __half* x16; long long ld_x16; // [batch*knn, 2*dim_vp16] = [v_other | v_orig * v_other]
  float* dist;                   // [batch*knn]
  int* cand_img;                 // [batch*knn] image row of every candidate (nullable)
  int with_vother;               // 1: x16 row = [v_other | v_mult]; 0: x16 row = [v_mult] only (v_other's part of layer 1
                                 //    comes from a per-image table)
  int batch, knn, dim_v, dim_vp, dim_vp16, dim_q, dim_qp, dim_a;
  int col_q, col_agt;            // in xs
  float eps;
  int no_vmult, no_vdist;
};
cudaError_t launch_ncx_features16(const Feat16Args& a, cudaStream_t s);
// z (fp16 [batch*knn, ld_z16], zero-padded to zp16 columns) and z_orig (fp32 rows, zero-padded to zo_pad) from the
// per-image table tab [n_img, ld_tab] (fp16, R*dim_mm used columns, [R][dim_mm] order) and hq [batch, ld_hq] (fp32,
// same order)
cudaError_t launch_ncx_mutan_apply(const void* tab, long long ld_tab, const int* img_idx, int batch, int knn,
                                   const float* hq, long long ld_hq, int R, int dim_mm, void* z16, long long ld_z16,
                                   int zp16, float* z_orig, long long ld_zo, int zo_pad, cudaStream_t s,
                                   float* z32 = nullptr, long long ld_z32 = 0, int z32_pad = 0);
// (z32 non-null: the candidates' z goes there in fp32, rows of ld_z32 floats zero-padded to z32_pad, instead of z16)
cudaError_t launch_ncx_softmax16(void* x, long long ld, int rows, int n, int n_pad, cudaStream_t s);
cudaError_t launch_ncx_to_half(const float* src, long long ld_src, int rows, int width, void* dst, long long ld_dst,
                               int dst_width, cudaStream_t s);

// z[row, j] = sum_r hv[row, r*dim_mm + j] * hq[row / rows_per_q, r*dim_mm + j]   (fusion.py:99-113), pad zeroed
cudaError_t launch_ncx_mutan_reduce(const float* hv, long long ld_hv, const float* hq, long long ld_hq, int rows,
                                    int rows_per_q, int n_cand_rows, int R, int dim_mm, int dim_mmp, float* z,
                                    long long ld_z, float* z_orig, long long ld_zo, cudaStream_t s);

// in-place row softmax over the first n columns of x[rows, ld]; columns [n, n_pad) are zeroed (cx.py:281)
cudaError_t launch_ncx_softmax(float* x, long long ld, int rows, int n, int n_pad, cudaStream_t s);

// out[i, n] = w[n * ld_w + col0 + i]  for i < cnt, n < rows  (transposed slice: the one-hot rank columns of linear_1)
cudaError_t launch_ncx_slice_t(const float* w, long long ld_w, int rows, int col0, int cnt, float* out, long long ld_out,
                               cudaStream_t s);
// inverse: w[n * ld_w + col0 + i] = in[i, n]
cudaError_t launch_ncx_unslice_t(const float* in, long long ld_in, int rows, int col0, int cnt, float* w, long long ld_w,
                                 cudaStream_t s);

// scores[r] = h[r, :n] . w + b[0]   (the `out` Linear, cx.py:327)
cudaError_t launch_ncx_score(const float* h, long long ld_h, int rows, int n, const float* w, const float* b,
                             float* scores, cudaStream_t s);

// Fused 24-way softmax cross-entropy (sum-reduced, counterexamples.py:310), its gradient, and Recall@1 / Recall@5
// (counterexamples.py:501-506 as a rank test). loss_sum[0] += sum_b CE_b;  hits[0] += #(rank < 1), hits[1] += #(rank < 5);
// dscores (nullable) = grad_scale * (softmax - onehot). Ties: a candidate with an equal score and a lower index is ahead.
cudaError_t launch_ncx_ce_recall(const float* scores, const long long* gt, long long batch, int knn, float* loss_sum,
                                 float* dscores, float grad_scale, int* hits, cudaStream_t s);

// d_pre[r, n] = dscores[r] * w_out[n] * (h[r, n] > 0 ? gate_scale : 0)
cudaError_t launch_ncx_dout(const float* dscores, const float* w_out, const float* h, long long ld_h, int rows, int n,
                            float gate_scale, float* dpre, long long ld_d, cudaStream_t s);

// Fused MUTAN rank products: z = sum_r (W_hv^r x_v + b_hv^r) * hq_r without materialising hv (mutan_fused.cu).
// cudaErrorNotSupported = shape outside the resident-tile design (caller falls back to GEMM + mutan_reduce).
cudaError_t launch_ncx_mutan_fused(const float* xv, long long ld_xv, const void* xv16, long long ld_xv16, int rows,
                                   int dim_hv, const float* whv, const void* whv16, long long ld_whv16,
                                   long long ld_whv, const float* bhv, const float* hq, long long ld_hq, int R, int dim_mm, int dim_mmp, int rows_per_q,
                                   int n_cand_rows, float* z, long long ld_z, float* z_orig, long long ld_zo,
                                   cudaStream_t s);

// dpre, g_wout, g_bout, g_blast of the `out` Linear + last hidden layer in one pass (sums atomically added: zero first)
cudaError_t launch_ncx_out_backward(const float* dscores, const float* w_out, const float* h, long long ld_h, int rows,
                                    int n, float gate_scale, float* dpre, long long ld_d, float* g_wout, float* g_bout,
                                    float* g_blast, cudaStream_t s);

// out[n] += sum_r x[r, n] * (w ? w[r] : 1)   for n < n_cols (atomic per 128-row block: zero `out` first)
cudaError_t launch_ncx_colsum(const float* x, long long ld, int rows, int n_cols, const float* w, float* out,
                              cudaStream_t s);

// ds[b, n] = sum_i d[b*knn + i, n];   drank_t[i, n] += sum_b d[b*knn + i, n];   g_b1[n] += sum of all rows (nullable)
cudaError_t launch_ncx_group_reduce(const float* d, long long ld, int batch, int knn, int n, float* ds, long long ld_ds,
                                    float* drank_t, long long ld_dr, float* g_b1, cudaStream_t s);

// dst[idx[b], :n] += src[b, :n]  (embedding-row gradient, atomics because answers repeat inside a batch)
cudaError_t launch_ncx_scatter_add_rows(const float* src, long long ld_src, const long long* idx, int batch, int n,
                                        float* dst, long long ld_dst, cudaStream_t s);

// torch.optim.Adam (no weight decay, no amsgrad) over a flat buffer; bias corrections passed in by the host
cudaError_t launch_adam(float* p, const float* g, float* m, float* v, long long n, float lr, float b1, float b2,
                        float eps, float bc1, float bc2, float grad_scale, const int* step_dev, cudaStream_t s);
// step_dev non-null: bc1 / bc2 are computed on the device from step_dev[0] (the launch can sit in a CUDA graph)

// step_dev[0] += 1; seed_dev[0] += seed_inc (both nullable): end-of-step bookkeeping of a graph-captured train step
cudaError_t launch_step_tick(int* step_dev, unsigned long long* seed_dev, unsigned long long seed_inc, cudaStream_t s);

// Example builder join: kNN output + complementary question pairs -> img_idx [n_q, k1], comp_idx [n_q] (-1 = dropped)
cudaError_t launch_cx_build_examples(const long long* knn_idx, long long n_img, int k1, const int* q_img, const int* q_comp,
                                     long long n_q, int* img_idx_out, int* comp_idx_out, cudaStream_t s);

// Heads of the other CX models on the softmax probabilities p [batch*knn, ld_p] (cx_heads.cu)
constexpr int kHeadBlackBox = 0, kHeadSemantic = 1;
cudaError_t launch_cx_head(const float* p, long long ld_p, int n_ans, const long long* answer_aid, int batch, int knn,
                           int head, const float* sim, long long ld_sim, float lam, float* scores, cudaStream_t s);
// out[n, n] = cosine similarity of the rows of x [n, d] (float64 accumulation)
cudaError_t launch_cx_cosine(const float* x, int n, int d, long long ld, float* out, long long ld_out, cudaStream_t s);

// ---- trainable-VQA variant (ncx_tv_kernels.cu): rows in library order [batch*knn candidate rows | batch original rows]
cudaError_t launch_tv_dropout_rows(const float* src, long long ld_src, int per_question, int n_cand_rows, int knn,
                                   float* dst, long long ld_dst, int rows, int cols, int cols_pad, float p,
                                   unsigned long long seed, const float* mask, float* mask_out, cudaStream_t s);
cudaError_t launch_tv_mutan_rows(const float* hv, const float* hq, long long ld_h, int rows, int n_cand_rows, int R,
                                 int dim_mm, int dim_mmp, float* z, long long ld_z, float* z_orig, long long ld_zo,
                                 cudaStream_t s);
cudaError_t launch_tv_mutan_bwd(float* hv, float* hq, long long ld_h, const float* dz, long long ld_dz, int rows, int R,
                                int dim_mm, cudaStream_t s);
cudaError_t launch_tv_softmax_bwd(const float* p, float* dp, long long ld, int rows, int n, cudaStream_t s);
cudaError_t launch_tv_tanh_bwd(float* d, long long ld_d, const float* x, long long ld_x, int rows, int cols,
                               cudaStream_t s);
cudaError_t launch_tv_add_scaled(float* acc, long long ld_acc, const float* add, const float* scale, long long ld_add,
                                 int rows, int cols, cudaStream_t s);
cudaError_t launch_tv_dq_reduce(const float* dxq0, long long ld_d, int batch, int knn, int cols, float p,
                                unsigned long long seed, const float* mask, float* dq, long long ld_dq, cudaStream_t s);

}  // namespace vqacx

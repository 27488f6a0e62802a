// Question encoder in TRAIN mode and its backward (back-propagation through time): BayesianUniSkip over BayesianGRU with
// its six SequentialDropout masks (skipthoughts.py:265-300, gru.py:46-84,132-155, dropout.py:44-66) -- what
// `trainable_vqa: true` (cx.py:73-74) adds to SURVEY §8 row f1: the reference back-propagates the NeuralCX loss through
// q_emb into seq2vec. Not a benchmarked path: one stream, every product a gemm_tf32 launch.
//
//   masks  m_x[g] [B, E], m_h[g] [B, H] (g = r, i, n): 0 or 1/(1-p), one draw per sequence, the same at every step
//   Xm[g]  [T*B, E] = emb[wid] * m_x[g]                       (step-major rows t*B + b)
//   GX[g]  [T*B, H] = Xm[g] . W_ig^T + b_ig                    three GEMMs for all steps
//   step t: GH[g] = (h * m_h[g]) . W_hg^T ; r, i = sigmoid(GX + GH) ; n = tanh(GX_n + r * GH_n) ; h' = (1-i) n + i h
//           rows with t >= length keep their state (so the last state IS "select last", skipthoughts.py:113-122)
//   kept for the backward: every h, r, i, n, GH_n and the masked copies Xm / Hm.
// Backward, t = T-1 .. 0 from dq = dLoss/dq_emb:  di = dh (h - n), dn = dh (1 - i), dh_direct = dh i ;
//   dn_pre = dn (1 - n^2), dr = dn_pre GH_n, dGH_n = dn_pre r ; di_pre = di i (1-i), dr_pre = dr r (1-r) ;
//   dh_prev = dh_direct + sum_g m_h[g] * (dG_g . W_hg) ; afterwards, over all steps at once:
//   dW_hg = dG_g^T . Hm[g],  dW_ig = dGx_g^T . Xm[g],  db_ig = colsum(dGx_g),  d emb[wid] += sum_g m_x[g] * (dGx_g . W_ig).
#include "../../include/vqacx.h"
#include "gemm_tf32.cuh"
#include "knn_internal.cuh"
#include "ncx_internal.cuh"

namespace vqacx {

namespace {

inline long long r32ll(long long x) { return (x + 31) / 32 * 32; }

struct GruTrainWs {
  float *mx, *mh;             // [3][B][E], [3][B][H]
  float *Xm, *GX, *GH;        // [3][TB][Ep], [3][TB][H], [3][B][H]
  float *Hall, *Hm;           // [T+1][B][H], [3][TB][H]
  float *R, *I, *N, *HN;      // [TB][H]
  float *DG;                  // [4][TB][H]: dr_pre, di_pre, dn_pre, dGH_n
  float *dh, *tmp, *dX;       // [2][B][H], [3][B][H], [3][TB][Ep]
  float* scratch;
  int* lengths;
  size_t scratch_floats, bytes;
};

GruTrainWs carve(void* base, long long B, int T, int E, int H) {
  GruTrainWs w{};
  size_t off = 0;
  auto take = [&](size_t n_floats) {
    off = (off + 255) & ~static_cast<size_t>(255);
    float* p = base ? reinterpret_cast<float*>(static_cast<uint8_t*>(base) + off) : nullptr;
    off += n_floats * 4;
    return p;
  };
  const size_t TB = static_cast<size_t>(T) * B, Ep = static_cast<size_t>(r32ll(E));
  w.mx = take(3 * B * E);
  w.mh = take(3 * B * static_cast<size_t>(H));
  w.Xm = take(3 * TB * Ep);
  w.GX = take(3 * TB * H);
  w.GH = take(3 * B * static_cast<size_t>(H));
  w.Hall = take((static_cast<size_t>(T) + 1) * B * H);
  w.Hm = take(3 * TB * H);
  w.R = take(TB * H); w.I = take(TB * H); w.N = take(TB * H); w.HN = take(TB * H);
  w.DG = take(4 * TB * H);
  w.dh = take(2 * B * static_cast<size_t>(H));
  w.tmp = take(3 * B * static_cast<size_t>(H));
  w.dX = take(3 * TB * Ep);
  w.lengths = reinterpret_cast<int*>(take(static_cast<size_t>(B)));
  w.scratch_floats = static_cast<size_t>(2) * 148 * 128 * 256;
  w.scratch = take(w.scratch_floats);
  w.bytes = off + 256;
  return w;
}

constexpr int kThreads = 256;
inline unsigned blocks_for(long long items) { return static_cast<unsigned>((items + kThreads - 1) / kThreads); }

// masks: explicit copy, or keep(seed + site, b, c) / (1 - p)
__global__ void __launch_bounds__(kThreads) gru_mask_kernel(const float* __restrict__ src, float* __restrict__ dst, int B,
                                                            int dim, float p, unsigned long long seed) {
  const long long i = static_cast<long long>(blockIdx.x) * kThreads + threadIdx.x;
  if (i >= 3LL * B * dim) return;
  if (src) { dst[i] = src[i]; return; }
  const int g = static_cast<int>(i / (static_cast<long long>(B) * dim));
  const long long r = i % (static_cast<long long>(B) * dim);
  const int b = static_cast<int>(r / dim), c = static_cast<int>(r % dim);
  dst[i] = p <= 0.f ? 1.f : (drop_keep(seed + static_cast<unsigned long long>(g), static_cast<uint32_t>(b),
                                       static_cast<uint32_t>(c), p) ? 1.f / (1.f - p) : 0.f);
}

__global__ void __launch_bounds__(kThreads) gru_train_gather_kernel(const float* __restrict__ emb, long long n_rows, int E,
                                                                    int Ep, const long long* __restrict__ wids, int B,
                                                                    int T, const float* __restrict__ mx,
                                                                    float* __restrict__ Xm, int* __restrict__ lengths) {
  const int t = blockIdx.x / B, b = blockIdx.x % B;
  long long w = wids[static_cast<long long>(b) * T + t];
  if (w < 0 || w >= n_rows) w = 0;
  const float* src = emb + w * E;
  const long long TB = static_cast<long long>(T) * B, row = static_cast<long long>(t) * B + b;
  for (int c = threadIdx.x; c < Ep; c += blockDim.x) {
    const float x = c < E ? __ldg(src + c) : 0.f;
#pragma unroll
    for (int g = 0; g < 3; ++g)
      Xm[(g * TB + row) * Ep + c] = c < E ? x * mx[(static_cast<long long>(g) * B + b) * E + c] : 0.f;
  }
  if (t == 0 && threadIdx.x == 0) {
    int n = 0;
    for (int j = 0; j < T; ++j) n += wids[static_cast<long long>(b) * T + j] != 0 ? 1 : 0;
    lengths[b] = n;
  }
}

__global__ void __launch_bounds__(kThreads) gru_train_gate_kernel(const float* __restrict__ GX, const float* __restrict__ GH,
                                                                  const int* __restrict__ lengths, int B, int H, int T, int t,
                                                                  const float* __restrict__ h_prev, float* __restrict__ h_new,
                                                                  const float* __restrict__ mh, float* __restrict__ Hm,
                                                                  float* __restrict__ R, float* __restrict__ I,
                                                                  float* __restrict__ N, float* __restrict__ HN) {
  const long long idx = static_cast<long long>(blockIdx.x) * kThreads + threadIdx.x;
  const long long BH = static_cast<long long>(B) * H;
  if (idx >= BH) return;
  const int b = static_cast<int>(idx / H), c = static_cast<int>(idx % H);
  const long long TB = static_cast<long long>(T) * B, row = static_cast<long long>(t) * B + b;
  const int len = lengths[b];
  const bool active = (len == 0) || (t < len);
  const float hp = h_prev[idx];
  float r = 0.f, i = 0.f, n = 0.f, hn = 0.f, h = hp;
  if (active) {
    float hr = 0.f, hi = 0.f;
    if (t > 0) { hr = GH[idx]; hi = GH[BH + idx]; hn = GH[2 * BH + idx]; }
    r = 1.f / (1.f + expf(-(GX[(0 * TB + row) * H + c] + hr)));
    i = 1.f / (1.f + expf(-(GX[(1 * TB + row) * H + c] + hi)));
    n = tanhf(GX[(2 * TB + row) * H + c] + r * hn);
    h = (1.f - i) * n + i * hp;
  }
  h_new[idx] = h;
  R[row * H + c] = r; I[row * H + c] = i; N[row * H + c] = n; HN[row * H + c] = hn;
  if (t + 1 < T) {
#pragma unroll
    for (int g = 0; g < 3; ++g) Hm[(g * TB + row + B) * H + c] = h * mh[(static_cast<long long>(g) * B + b) * H + c];
  }
}

__global__ void __launch_bounds__(kThreads) gru_train_gate_bwd_kernel(const float* __restrict__ dh, const float* __restrict__ R,
                                                                      const float* __restrict__ I, const float* __restrict__ N,
                                                                      const float* __restrict__ HN,
                                                                      const float* __restrict__ h_prev,
                                                                      const int* __restrict__ lengths, int B, int H, int T,
                                                                      int t, float* __restrict__ DG, float* __restrict__ dh_out) {
  const long long idx = static_cast<long long>(blockIdx.x) * kThreads + threadIdx.x;
  if (idx >= static_cast<long long>(B) * H) return;
  const int b = static_cast<int>(idx / H), c = static_cast<int>(idx % H);
  const long long TBH = static_cast<long long>(T) * B * H, o = (static_cast<long long>(t) * B + b) * H + c;
  const int len = lengths[b];
  const bool active = (len == 0) || (t < len);
  const float g = dh[idx];
  float dr_pre = 0.f, di_pre = 0.f, dn_pre = 0.f, dhn = 0.f, direct = g;
  if (active) {
    const float r = R[o], i = I[o], n = N[o], hn = HN[o], hp = h_prev[idx];
    const float di = g * (hp - n), dn = g * (1.f - i);
    direct = g * i;
    dn_pre = dn * (1.f - n * n);
    dhn = dn_pre * r;
    di_pre = di * i * (1.f - i);
    dr_pre = dn_pre * hn * r * (1.f - r);
  }
  DG[o] = dr_pre; DG[TBH + o] = di_pre; DG[2 * TBH + o] = dn_pre; DG[3 * TBH + o] = dhn;
  dh_out[idx] = direct;
}

// dh[b, c] += sum_g tmp[g][b, c] * m_h[g][b, c]
__global__ void __launch_bounds__(kThreads) gru_train_dh_kernel(float* __restrict__ dh, const float* __restrict__ tmp,
                                                                const float* __restrict__ mh, long long BH) {
  const long long idx = static_cast<long long>(blockIdx.x) * kThreads + threadIdx.x;
  if (idx >= BH) return;
  dh[idx] += tmp[idx] * mh[idx] + tmp[BH + idx] * mh[BH + idx] + tmp[2 * BH + idx] * mh[2 * BH + idx];
}

__global__ void __launch_bounds__(kThreads) gru_copy_rows_kernel(const float* __restrict__ src, long long ld_src,
                                                                 float* __restrict__ dst, long long ld_dst, int rows, int cols) {
  const long long idx = static_cast<long long>(blockIdx.x) * kThreads + threadIdx.x;
  if (idx >= static_cast<long long>(rows) * cols) return;
  const int r = static_cast<int>(idx / cols), c = static_cast<int>(idx % cols);
  dst[r * ld_dst + c] = src[r * ld_src + c];
}

// d_emb[wid(b, t), c] += sum_g dX[g][t*B + b, c] * m_x[g][b, c]   (padding row 0 and out-of-vocabulary ids get nothing)
__global__ void __launch_bounds__(kThreads) gru_train_emb_scatter_kernel(const float* __restrict__ dX, int Ep,
                                                                         const float* __restrict__ mx, long long n_rows, int E,
                                                                         const long long* __restrict__ wids, int B, int T,
                                                                         float* __restrict__ d_emb) {
  const int t = blockIdx.x / B, b = blockIdx.x % B;
  const long long w = wids[static_cast<long long>(b) * T + t];
  if (w <= 0 || w >= n_rows) return;
  const long long TB = static_cast<long long>(T) * B, row = static_cast<long long>(t) * B + b;
  for (int c = threadIdx.x; c < E; c += blockDim.x) {
    float acc = 0.f;
#pragma unroll
    for (int g = 0; g < 3; ++g) acc += dX[(g * TB + row) * Ep + c] * mx[(static_cast<long long>(g) * B + b) * E + c];
    atomicAdd(d_emb + w * E + c, acc);
  }
}

cudaError_t gemm(const GruTrainWs& w, int M, int N, int K, const float* a, long long lda, int a_mn, const float* b,
                 long long ldb, int b_mn, float* c, long long ldc, const float* bias, cudaStream_t s) {
  GemmProblem g;
  g.M = M; g.N = N; g.nseg = 1;
  g.seg[0].a = {a, lda, a_mn};
  g.seg[0].b = {b, ldb, b_mn};
  g.seg[0].k = K;
  g.C = c; g.ldc = ldc;
  g.split_k = 0; g.bn = 0; g.scratch = w.scratch; g.scratch_floats = w.scratch_floats;
  g.epi.bias = bias;
  return launch_gemm_tf32(g, s);
}

int status(cudaError_t e) {
  if (e == cudaSuccess) return VQACX_OK;
  return e == cudaErrorInvalidValue ? VQACX_EINVAL : static_cast<int>(e);
}

bool shapes_ok(int64_t batch, int T, int E, int H) {
  return batch >= 0 && T > 0 && E > 0 && H > 0 && (E % 4) == 0 && (H % 4) == 0 &&
         batch * static_cast<int64_t>(T) < (1LL << 24);
}

}  // namespace
}  // namespace vqacx

using namespace vqacx;

#define GT_TRY(expr)                                   \
  do {                                                 \
    const cudaError_t e__ = (expr);                    \
    if (e__ != cudaSuccess) return status(e__);        \
  } while (0)
#define GT_LAUNCH(kernel, grid, ...)                   \
  do {                                                 \
    count_launch();                                    \
    kernel<<<(grid), kThreads, 0, s>>>(__VA_ARGS__);   \
    GT_TRY(cudaGetLastError());                        \
  } while (0)

extern "C" {

size_t vqacx_gru_train_workspace_bytes(int64_t batch, int T, int E, int H) {
  if (!shapes_ok(batch, T, E, H)) return 0;
  return carve(nullptr, batch, T, E, H).bytes;
}

int vqacx_gru_forward_train(const float* emb, int64_t n_emb_rows, int E, const int64_t* wids, int64_t batch, int T,
                            const float* w_i, const float* b_i, const float* w_h, int H, float p_drop, uint64_t seed,
                            const float* masks_x, const float* masks_h, float* out, void* ws, size_t ws_bytes,
                            void* stream) {
  if (!shapes_ok(batch, T, E, H) || n_emb_rows <= 0 || !(p_drop >= 0.f && p_drop < 1.f)) return VQACX_EINVAL;
  if (batch == 0) return VQACX_OK;
  if (!emb || !wids || !w_i || !b_i || !w_h || !out || !ws) return VQACX_EINVAL;
  if ((reinterpret_cast<uintptr_t>(w_i) & 15) || (reinterpret_cast<uintptr_t>(w_h) & 15) ||
      (reinterpret_cast<uintptr_t>(ws) & 255))
    return VQACX_EALIGN;
  const GruTrainWs w = carve(ws, batch, T, E, H);
  if (w.bytes > ws_bytes) return VQACX_EWORKSPACE;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int B = static_cast<int>(batch), Ep = static_cast<int>(r32ll(E));
  const long long TB = static_cast<long long>(T) * B, BH = static_cast<long long>(B) * H;
  GT_LAUNCH(gru_mask_kernel, blocks_for(3LL * B * E), masks_x, w.mx, B, E, p_drop, seed + 1);
  GT_LAUNCH(gru_mask_kernel, blocks_for(3 * BH), masks_h, w.mh, B, H, p_drop, seed + 4);
  GT_LAUNCH(gru_train_gather_kernel, static_cast<unsigned>(TB), emb, n_emb_rows, E, Ep,
            reinterpret_cast<const long long*>(wids), B, T, w.mx, w.Xm, w.lengths);
  for (int g = 0; g < 3; ++g)
    GT_TRY(gemm(w, static_cast<int>(TB), H, E, w.Xm + g * TB * Ep, Ep, 0, w_i + static_cast<long long>(g) * H * E, E, 0,
                w.GX + g * TB * H, H, b_i + g * H, s));
  GT_TRY(cudaMemsetAsync(w.Hall, 0, static_cast<size_t>(BH) * 4, s));            // h_0 = 0 (gru.py:73-74)
  for (int t = 0; t < T; ++t) {
    if (t > 0)
      for (int g = 0; g < 3; ++g)
        GT_TRY(gemm(w, B, H, H, w.Hm + (g * TB + static_cast<long long>(t) * B) * H, H, 0,
                    w_h + static_cast<long long>(g) * H * H, H, 0, w.GH + g * BH, H, nullptr, s));
    GT_LAUNCH(gru_train_gate_kernel, blocks_for(BH), w.GX, w.GH, w.lengths, B, H, T, t, w.Hall + t * BH,
              w.Hall + (t + 1) * BH, w.mh, w.Hm, w.R, w.I, w.N, w.HN);
  }
  GT_TRY(cudaMemcpyAsync(out, w.Hall + static_cast<long long>(T) * BH, static_cast<size_t>(BH) * 4, cudaMemcpyDeviceToDevice, s));
  return VQACX_OK;
}

int vqacx_gru_backward(int64_t n_emb_rows, int E, const int64_t* wids, int64_t batch, int T, const float* w_i,
                       const float* w_h, int H, const float* dq, int64_t ld_dq, void* ws, size_t ws_bytes, float* d_emb,
                       float* d_w_i, float* d_b_i, float* d_w_h, void* stream) {
  if (!shapes_ok(batch, T, E, H) || n_emb_rows <= 0 || ld_dq < H) return VQACX_EINVAL;
  if (!wids || !w_i || !w_h || !ws || !d_emb || !d_w_i || !d_b_i || !d_w_h || (batch > 0 && !dq)) return VQACX_EINVAL;
  if ((reinterpret_cast<uintptr_t>(w_i) & 15) || (reinterpret_cast<uintptr_t>(w_h) & 15) ||
      (reinterpret_cast<uintptr_t>(d_w_i) & 15) || (reinterpret_cast<uintptr_t>(d_w_h) & 15) ||
      (reinterpret_cast<uintptr_t>(ws) & 255))
    return VQACX_EALIGN;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  GT_TRY(cudaMemsetAsync(d_emb, 0, static_cast<size_t>(n_emb_rows) * E * 4, s));
  GT_TRY(cudaMemsetAsync(d_b_i, 0, static_cast<size_t>(3) * H * 4, s));
  if (batch == 0) {
    GT_TRY(cudaMemsetAsync(d_w_i, 0, static_cast<size_t>(3) * H * E * 4, s));
    GT_TRY(cudaMemsetAsync(d_w_h, 0, static_cast<size_t>(3) * H * H * 4, s));
    return VQACX_OK;
  }
  const GruTrainWs w = carve(ws, batch, T, E, H);
  if (w.bytes > ws_bytes) return VQACX_EWORKSPACE;
  const int B = static_cast<int>(batch), Ep = static_cast<int>(r32ll(E));
  const long long TB = static_cast<long long>(T) * B, BH = static_cast<long long>(B) * H;
  float* dh_cur = w.dh;
  float* dh_nxt = w.dh + BH;
  GT_LAUNCH(gru_copy_rows_kernel, blocks_for(BH), dq, ld_dq, dh_cur, H, B, H);
  const int gate_of_h[3] = {0, 1, 3};   // recurrent side: dr_pre, di_pre, dGH_n
  for (int t = T - 1; t >= 0; --t) {
    GT_LAUNCH(gru_train_gate_bwd_kernel, blocks_for(BH), dh_cur, w.R, w.I, w.N, w.HN, w.Hall + t * BH, w.lengths, B, H, T, t,
              w.DG, dh_nxt);
    if (t > 0) {
      for (int g = 0; g < 3; ++g)
        GT_TRY(gemm(w, B, H, H, w.DG + (gate_of_h[g] * TB + static_cast<long long>(t) * B) * H, H, 0,
                    w_h + static_cast<long long>(g) * H * H, H, 1, w.tmp + g * BH, H, nullptr, s));
      GT_LAUNCH(gru_train_dh_kernel, blocks_for(BH), dh_nxt, w.tmp, w.mh, BH);
    }
    float* sw = dh_cur; dh_cur = dh_nxt; dh_nxt = sw;
  }
  for (int g = 0; g < 3; ++g) {
    // recurrent weights: the t = 0 rows of Hm are h_0 * m = 0 (never written: clear them), so all T*B rows take part
    GT_TRY(cudaMemsetAsync(w.Hm + g * TB * H, 0, static_cast<size_t>(BH) * 4, s));
    GT_TRY(gemm(w, H, H, static_cast<int>(TB), w.DG + gate_of_h[g] * TB * H, H, 1, w.Hm + g * TB * H, H, 1,
                d_w_h + static_cast<long long>(g) * H * H, H, nullptr, s));
    GT_TRY(gemm(w, H, E, static_cast<int>(TB), w.DG + g * TB * H, H, 1, w.Xm + g * TB * Ep, Ep, 1,
                d_w_i + static_cast<long long>(g) * H * E, E, nullptr, s));
    GT_TRY(launch_ncx_colsum(w.DG + g * TB * H, H, static_cast<int>(TB), H, nullptr, d_b_i + g * H, s));
    GT_TRY(gemm(w, static_cast<int>(TB), E, H, w.DG + g * TB * H, H, 0, w_i + static_cast<long long>(g) * H * E, E, 1,
                w.dX + g * TB * Ep, Ep, nullptr, s));
  }
  GT_LAUNCH(gru_train_emb_scatter_kernel, static_cast<unsigned>(TB), w.dX, Ep, w.mx, n_emb_rows, E,
            reinterpret_cast<const long long*>(wids), B, T, d_emb);
  return VQACX_OK;
}

}  // extern "C"

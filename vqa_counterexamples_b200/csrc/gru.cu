// Question encoder on the device: BayesianUniSkip in eval mode (skipthoughts.py:265-300 over BayesianGRU, gru.py:46-84,
// 132-155) -- SURVEY §8 rows a7 / f1. The encoder is frozen and in eval mode on the NeuralCX path (cx.py:76-80), so
// every SequentialDropout is the identity and q_emb is a pure function of the word ids.
//
//   X  [T*B, E]  = embedding rows, step-major (row t*B + b), one gather kernel (also derives the lengths)
//   GX [T*B, 3H] = X . [W_ir; W_ii; W_in]^T + [b_ir; b_ii; b_in]      ONE tcgen05 GEMM for all steps
//   per step t:  GH [B, 3H] = h . [W_hr; W_hi; W_hn]^T                  (gemm_tf32, split-K: M = B is small)
//                r = sigmoid(GX_r + GH_r), i = sigmoid(GX_i + GH_i), n = tanh(GX_n + r * GH_n), h = (1-i) n + i h
//                applied only to rows with t < length (so the final h IS "select last", skipthoughts.py:113-122)
#include "../../include/vqacx.h"
#include "gemm_tf32.cuh"
#include "knn_internal.cuh"

namespace vqacx {

namespace {
inline long long r32ll(long long x) { return (x + 31) / 32 * 32; }
}

__global__ void __launch_bounds__(256) gru_gather_kernel(const float* __restrict__ emb, long long n_rows, int E, int Ep,
                                                         const long long* __restrict__ wids, int B, int T,
                                                         float* __restrict__ X, int* __restrict__ lengths) {
  const int t = blockIdx.x / B, b = blockIdx.x % B;
  long long w = wids[static_cast<long long>(b) * T + t];
  if (w < 0 || w >= n_rows) w = 0;   // out-of-vocabulary ids read the padding row
  const float* src = emb + w * E;
  float* dst = X + (static_cast<long long>(t) * B + b) * Ep;
  for (int c = threadIdx.x; c < Ep; c += blockDim.x) dst[c] = c < E ? __ldg(src + c) : 0.f;
  if (t == 0 && threadIdx.x == 0) {
    int n = 0;
    for (int j = 0; j < T; ++j) n += wids[static_cast<long long>(b) * T + j] != 0 ? 1 : 0;   // T - #zeros (:131-134)
    lengths[b] = n;
  }
}

__global__ void __launch_bounds__(256) gru_gate_kernel(const float* __restrict__ gx, const float* __restrict__ gh,
                                                       const int* __restrict__ lengths, int B, int H, int t,
                                                       float* __restrict__ h) {
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= static_cast<long long>(B) * H) return;
  const int b = static_cast<int>(idx / H), c = static_cast<int>(idx % H);
  const int len = lengths[b];
  if (len != 0 && t >= len) return;            // finished rows keep their state; an all-padding row runs every step
  const float* x = gx + (static_cast<long long>(t) * B + b) * 3 * H;
  float hr = 0.f, hi = 0.f, hn = 0.f;
  if (gh) {
    const float* g = gh + static_cast<long long>(b) * 3 * H;
    hr = g[c]; hi = g[H + c]; hn = g[2 * H + c];
  }
  const float r = 1.f / (1.f + expf(-(x[c] + hr)));
  const float i = 1.f / (1.f + expf(-(x[H + c] + hi)));
  const float n = tanhf(x[2 * H + c] + r * hn);
  const float hp = h[idx];
  h[idx] = (1.f - i) * n + i * hp;
}

struct GruWs {
  float *X, *GX, *GH, *scratch;
  int* lengths;
  size_t scratch_floats, bytes;
};

static GruWs carve_gru(void* base, long long B, int T, int E, int H) {
  GruWs w{};
  size_t off = 0;
  auto take = [&](size_t bytes) {
    off = (off + 255) & ~static_cast<size_t>(255);
    void* p = base ? static_cast<uint8_t*>(base) + off : nullptr;
    off += bytes;
    return p;
  };
  w.X = static_cast<float*>(take(static_cast<size_t>(T) * B * r32ll(E) * 4));
  w.GX = static_cast<float*>(take(static_cast<size_t>(T) * B * 3 * H * 4));
  w.GH = static_cast<float*>(take(static_cast<size_t>(B) * 3 * H * 4));
  w.lengths = static_cast<int*>(take(static_cast<size_t>(B) * 4));
  w.scratch_floats = static_cast<size_t>(2) * 148 * 128 * 256;
  w.scratch = static_cast<float*>(take(w.scratch_floats * 4));
  w.bytes = off + 256;
  return w;
}

}  // namespace vqacx

using namespace vqacx;

extern "C" {

size_t vqacx_gru_workspace_bytes(int64_t batch, int T, int E, int H) {
  if (batch < 0 || T <= 0 || E <= 0 || H <= 0) return 0;
  return carve_gru(nullptr, batch, T, E, H).bytes;
}

int vqacx_gru_encode(const float* emb, int64_t n_emb_rows, int E, const int64_t* wids, int64_t batch, int T,
                     const float* w_i, const float* b_i, const float* w_h, int H, int max_steps, float* out, void* ws,
                     size_t ws_bytes, void* stream) {
  if (batch < 0 || T <= 0 || E <= 0 || H <= 0 || (E % 4) || (H % 4) || n_emb_rows <= 0 || max_steps < 0 || max_steps > T)
    return VQACX_EINVAL;
  if (batch == 0) return VQACX_OK;
  if (batch * static_cast<int64_t>(T) >= (1LL << 31) - 256) return VQACX_EINVAL;
  if (!emb || !wids || !w_i || !b_i || !w_h || !out || !ws) return VQACX_EINVAL;
  if ((reinterpret_cast<uintptr_t>(w_i) & 15) || (reinterpret_cast<uintptr_t>(w_h) & 15) ||
      (reinterpret_cast<uintptr_t>(out) & 15) || (reinterpret_cast<uintptr_t>(ws) & 255))
    return VQACX_EALIGN;
  GruWs w = carve_gru(ws, batch, T, E, H);
  if (w.bytes > ws_bytes) return VQACX_EWORKSPACE;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int B = static_cast<int>(batch), steps = max_steps ? max_steps : T;
  const int Ep = static_cast<int>(r32ll(E));
  cudaError_t e = cudaMemsetAsync(out, 0, static_cast<size_t>(B) * H * sizeof(float), s);   // h_0 = 0 (gru.py:73-74)
  if (e != cudaSuccess) return static_cast<int>(e);
  count_launch();
  gru_gather_kernel<<<B * T, 256, 0, s>>>(emb, n_emb_rows, E, Ep, reinterpret_cast<const long long*>(wids), B, T, w.X,
                                          w.lengths);
  if ((e = cudaGetLastError()) != cudaSuccess) return static_cast<int>(e);
  GemmProblem g;
  g.M = steps * B; g.N = 3 * H; g.nseg = 1;
  g.seg[0].a = {w.X, Ep, 0};
  g.seg[0].b = {w_i, E, 0};
  g.seg[0].k = E;
  g.C = w.GX; g.ldc = 3LL * H;
  g.split_k = 0; g.bn = 0; g.scratch = w.scratch; g.scratch_floats = w.scratch_floats;
  g.epi.bias = b_i;
  e = launch_gemm_tf32(g, s);
  if (e != cudaSuccess) return e == cudaErrorInvalidValue ? VQACX_EINVAL : static_cast<int>(e);
  const long long tot = static_cast<long long>(B) * H;
  const unsigned blocks = static_cast<unsigned>((tot + 255) / 256);
  for (int t = 0; t < steps; ++t) {
    if (t > 0) {   // h_0 = 0 -> GH = 0 at the first step
      GemmProblem r;
      r.M = B; r.N = 3 * H; r.nseg = 1;
      r.seg[0].a = {out, H, 0};
      r.seg[0].b = {w_h, H, 0};
      r.seg[0].k = H;
      r.C = w.GH; r.ldc = 3LL * H;
      r.split_k = 0; r.bn = 0; r.scratch = w.scratch; r.scratch_floats = w.scratch_floats;
      e = launch_gemm_tf32(r, s);
      if (e != cudaSuccess) return e == cudaErrorInvalidValue ? VQACX_EINVAL : static_cast<int>(e);
    }
    count_launch();
    gru_gate_kernel<<<blocks, 256, 0, s>>>(w.GX, t > 0 ? w.GH : nullptr, w.lengths, B, H, t, out);
    if ((e = cudaGetLastError()) != cudaSuccess) return static_cast<int>(e);
  }
  return VQACX_OK;
}

}  // extern "C"

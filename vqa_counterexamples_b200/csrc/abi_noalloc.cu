// C-ABI entry point that allocates nothing: the deterministic split-K GEMM with its partial-tile scratch in a
// file-scope device array instead of a caller buffer. For callers that cannot provide scratch memory (CUDA-graph
// capture set-ups, serving loops and harnesses that forbid cudaMalloc). No counterpart in the reference.
#include "../../include/vqacx.h"
#include "gemm_tf32.cuh"
#include <atomic>
#include <mutex>

using namespace vqacx;

namespace {

constexpr size_t kStaticFloats = VQACX_STATIC_SCRATCH_BYTES / sizeof(float);
constexpr int kMaxDevices = 64;

// Held by every GPU that loads the library, until the process exits. In HOST code this name is only a handle for
// cudaGetSymbolAddress -- it is not a device address and must never be passed to a kernel.
static __device__ __align__(256) float g_splitk_scratch[kStaticFloats];

std::mutex g_mu;
std::atomic<float*> g_scratch_ptr[kMaxDevices];

// Device address of the array on the CURRENT device, resolved once per device.
int static_scratch(float** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return static_cast<int>(e);
  if (dev < 0 || dev >= kMaxDevices) return VQACX_EINVAL;
  float* p = g_scratch_ptr[dev].load(std::memory_order_acquire);
  if (!p) {
    std::lock_guard<std::mutex> lock(g_mu);
    p = g_scratch_ptr[dev].load(std::memory_order_relaxed);
    if (!p) {
      void* addr = nullptr;
      e = cudaGetSymbolAddress(&addr, g_splitk_scratch);
      if (e != cudaSuccess) return static_cast<int>(e);
      if (!addr || (reinterpret_cast<uintptr_t>(addr) & 255)) return VQACX_EALIGN;
      p = static_cast<float*>(addr);
      g_scratch_ptr[dev].store(p, std::memory_order_release);
    }
  }
  *out = p;
  return VQACX_OK;
}

}  // namespace

extern "C" int vqacx_gemm_tf32_splitk_static(int m, int n, int k, const float* a, int64_t lda, int a_mn_major,
                                             const float* b, int64_t ldb, int b_mn_major, float* c, int64_t ldc,
                                             const float* bias, int act, int split_k, int bn, void* stream) {
  if (m < 0 || n < 0 || k <= 0 || act < 0 || act > 2 || split_k < 0) return VQACX_EINVAL;
  if (m == 0 || n == 0) return VQACX_OK;
  if (!a || !b || !c) return VQACX_EINVAL;
  if ((lda % 4) || (ldb % 4) || (reinterpret_cast<uintptr_t>(a) & 15) || (reinterpret_cast<uintptr_t>(b) & 15))
    return VQACX_EALIGN;
  // one raw partial tile [m, ceil32(n)] per K-slice; with split_k = 0 the cost model may pick any split that fits,
  // so at least two slices have to (otherwise the call would not be a split-K product at all)
  const size_t slice = static_cast<size_t>(m) * (static_cast<size_t>((n + 31) / 32) * 32);
  const size_t slices = split_k > 1 ? static_cast<size_t>(split_k) : 2;
  if (slice > kStaticFloats / slices) return VQACX_ESTATIC;
  float* scratch = nullptr;
  const int rc = static_scratch(&scratch);
  if (rc != VQACX_OK) return rc;
  GemmProblem p;
  p.M = m; p.N = n; p.nseg = 1;
  p.seg[0].a = {a, lda, a_mn_major};
  p.seg[0].b = {b, ldb, b_mn_major};
  p.seg[0].k = k;
  p.C = c; p.ldc = ldc; p.split_k = split_k; p.bn = bn;
  p.scratch = scratch; p.scratch_floats = kStaticFloats;
  p.epi.bias = bias; p.epi.act = act;
  const cudaError_t e = launch_gemm_tf32(p, static_cast<cudaStream_t>(stream));
  if (e == cudaErrorInvalidValue) return VQACX_EINVAL;
  return e == cudaSuccess ? VQACX_OK : static_cast<int>(e);
}

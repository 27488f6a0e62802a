// Pooled-feature producer, the step before stage 1 (extract.py:121-124):
//     output_noatt = output_att.sum(3).sum(2).div(nb_regions)          [b, C, H, W] -> [b, C]      (C = 2048, H*W = 196)
// fused with what stage 1's `fit` needs from every row (vqacx_knn_prep): the fp16 copy of the pooled row, the squared
// norm of the ROUNDED row and the global max norm -- so that the 1.6 MB activation map of an image is read exactly once
// and the kNN index is ready when the last image has been pooled. HBM-bound: C*HW*4 bytes read per image.
// One CTA per image; one warp per channel (HW contiguous floats), 8 channels in flight per CTA.
#include "../../include/vqacx.h"
#include "knn_internal.cuh"
#include <cuda_fp16.h>

namespace vqacx {

__global__ void __launch_bounds__(256) pool_features_kernel(const float* __restrict__ att, int C, int HW,
                                                            float* __restrict__ out, __half* __restrict__ out_lp,
                                                            float* __restrict__ norm_sq,
                                                            unsigned int* __restrict__ max_norm_sq_bits,
                                                            int* __restrict__ bad_count) {
  __shared__ double s_acc[8];
  __shared__ int s_bad[8];
  const long long img = blockIdx.x;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const float* base = att + img * static_cast<long long>(C) * HW;
  const float inv = 1.f / static_cast<float>(HW);
  const bool vec = (HW % 4) == 0 && (reinterpret_cast<uintptr_t>(att) & 15) == 0;
  double acc = 0.0;   // squared norm of the fp16-rounded pooled row (lane 0 of each warp)
  int bad = 0;
  for (int c = w; c < C; c += 8) {
    const float* src = base + static_cast<long long>(c) * HW;
    float s = 0.f;
    if (vec) {
      const float4* s4 = reinterpret_cast<const float4*>(src);
      for (int i = lane; i < HW / 4; i += 32) {
        const float4 v = __ldcs(s4 + i);
        s += (v.x + v.y) + (v.z + v.w);
      }
    } else {
      for (int i = lane; i < HW; i += 32) s += __ldcs(src + i);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) {
      const float m = s * inv;                                     // .div(nb_regions)
      out[img * C + c] = m;
      if (out_lp) {
        const __half h = __float2half_rn(m);
        out_lp[img * C + c] = h;
        const float r = __half2float(h);
        bad |= !isfinite(r);
        acc += static_cast<double>(r) * r;
      }
    }
  }
  if (!out_lp) return;
  if (lane == 0) { s_acc[w] = acc; s_bad[w] = bad; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    int b = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) { t += s_acc[i]; b |= s_bad[i]; }
    const float ns = static_cast<float>(t);
    if (norm_sq) norm_sq[img] = ns;
    if (max_norm_sq_bits) atomicMax(max_norm_sq_bits, __float_as_uint(ns));
    if (b && bad_count) atomicAdd(bad_count, 1);
  }
}

}  // namespace vqacx

using namespace vqacx;

extern "C" int vqacx_pool_features(const float* att, int64_t n, int C, int HW, float* out_f32, void* out_lp,
                                   float* norm_sq, uint32_t* max_norm_sq_bits, int32_t* bad_count, void* stream) {
  if (n < 0 || C <= 0 || HW <= 0) return VQACX_EINVAL;
  if (n == 0) return VQACX_OK;
  if (!att || !out_f32) return VQACX_EINVAL;
  if (out_lp && (!norm_sq || !max_norm_sq_bits || !bad_count)) return VQACX_EINVAL;
  if (n >= (1LL << 31)) return VQACX_EINVAL;
  count_launch();
  pool_features_kernel<<<static_cast<unsigned>(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      att, C, HW, out_f32, static_cast<__half*>(out_lp), norm_sq, max_norm_sq_bits, bad_count);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? VQACX_OK : static_cast<int>(e);
}

// Glue kernels of the fp16 fast path of NeuralCX evaluation (vqacx_ncx_fast_forward, abi_ncx.cu): the operands of the
// big products are written ONCE, in fp16, by the kernel that produces them; nothing here is a dense product.
//   ncx_features16   gather of the 25 image rows of a question -> X16[r] = [v_other | v_orig * v_other] (fp16), the
//                    pairwise distance as a separate fp32 vector (it enters layer 1 as a rank-1 epilogue term), and the
//                    per-question fp32 row [v_orig | q_emb | . | E[answer]]                         (cx.py:280, 289-303)
//   ncx_mutan_apply  z = sum_r hv_r[image] * hq_r[question] from the per-IMAGE table of the frozen head's visual rank
//                    projections (fusion.py:99-113: hv depends on the image only, hq on the question only)
//   ncx_softmax16    row softmax over fp16 logits, in place                                                   (cx.py:281)
//   ncx_to_half      fp32 -> fp16 copy of a column block with zero padding (weight twins)
#include "knn_internal.cuh"
#include "ncx_internal.cuh"
#include "ptx.cuh"
#include <cuda_fp16.h>

namespace vqacx {

namespace {
__device__ __forceinline__ float wsum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float wmax(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ uint2 pack4(float a, float b, float c, float d) {
  const __half2 lo = __floats2half2_rn(a, b), hi = __floats2half2_rn(c, d);
  uint2 r;
  r.x = *reinterpret_cast<const uint32_t*>(&lo);
  r.y = *reinterpret_cast<const uint32_t*>(&hi);
  return r;
}
}  // namespace

// One block per question, one WARP per image slot (slots w, w + 8, ...): the original image's row sits in the lane's
// registers (dim_v <= 2048: 16 float4 per lane) and every slot is 16 independent 512-byte warp loads, the products, the
// fp16 stores and ONE warp reduction for the distance -- no block barrier, ~16 loads in flight per lane. (A first
// version with one 256-thread block per (question, slot), two dependent loads per thread and a block reduction, ran
// at 3.9 TB/s of DRAM traffic.) Rows wider than 2048 take the per-slot-block kernel below.
constexpr int kFeatMaxV4 = 16;   // float4 per lane
constexpr int kFeatQWarps = 12;   // 24 neighbour slots = two per warp
__global__ void __launch_bounds__(32 * kFeatQWarps, 2) ncx_features16_q_kernel(const Feat16Args a) {
  // v_orig lives in shared memory (it was 64 registers per lane: 174 registers, one block of 8 warps per SM and 64 KB of
  // gathers in flight; now 12 warps per block gather at once and a second block fits beside it)
  __shared__ float4 s_o[kFeatMaxV4 * 32];
  const int b = blockIdx.x, lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int nv = a.dim_v / 4;
  const float4* vo = reinterpret_cast<const float4*>(a.features + static_cast<long long>(a.img_idx[b * (a.knn + 1)]) * a.ld_features);
  for (int i = threadIdx.x; i < kFeatMaxV4 * 32; i += blockDim.x) s_o[i] = i < nv ? __ldg(vo + i) : make_float4(0.f, 0.f, 0.f, 0.f);
  __syncthreads();
  const float4* o = s_o + lane;     // o[32 * t] = this lane's t-th piece
  if (w == 0) {   // per-question fp32 row: [v_orig | q_emb | (z_orig: mutan_apply) | E[answer]]
    float* row = a.xs + static_cast<long long>(b) * a.ld_xs;
    float4* d = reinterpret_cast<float4*>(row);
#pragma unroll
    for (int t = 0; t < kFeatMaxV4; ++t) {
      const int i = lane + 32 * t;
      if (i < nv) d[i] = o[32 * t];
    }
    for (int i = a.dim_v + lane; i < a.dim_vp; i += 32) row[i] = 0.f;
  } else if (w == 1) {
    float* row = a.xs + static_cast<long long>(b) * a.ld_xs;
    const float4* q = reinterpret_cast<const float4*>(a.q_emb + static_cast<long long>(b) * a.ld_q);
    float4* dq = reinterpret_cast<float4*>(row + a.col_q);
    for (int i = lane; i < a.dim_q / 4; i += 32) dq[i] = __ldg(q + i);
    for (int i = a.dim_q + lane; i < a.dim_qp; i += 32) row[a.col_q + i] = 0.f;
  } else if (w == 2) {
    float* row = a.xs + static_cast<long long>(b) * a.ld_xs;
    const float4* e = reinterpret_cast<const float4*>(a.emb + a.answer_aid[b] * static_cast<long long>(a.dim_a));
    float4* de = reinterpret_cast<float4*>(row + a.col_agt);
    for (int i = lane; i < a.dim_a / 4; i += 32) de[i] = __ldg(e + i);
  }
  const int mult0 = a.with_vother ? a.dim_vp16 : 0;
  for (int slot = 1 + w; slot <= a.knn; slot += kFeatQWarps) {
    const long long r = static_cast<long long>(b) * a.knn + (slot - 1);
    const int img = a.img_idx[b * (a.knn + 1) + slot];
    if (a.cand_img && lane == 0) a.cand_img[r] = img;
    if (a.features16) {
      // neighbour row from the fp16 copy: 16-byte pieces of 8 halfs (half the gathered bytes; v_other's fp16 value is the
      // same rounding the fp32 path applies, v_mult and the distance use it against the exact fp32 v_orig)
      const uint4* vk16 = reinterpret_cast<const uint4*>(a.features16 + static_cast<long long>(img) * a.ld_f16);
      __half* row = a.x16 + r * a.ld_x16;
      uint4* d_other = reinterpret_cast<uint4*>(row);
      uint4* d_mult = reinterpret_cast<uint4*>(row + mult0);
      const int n8 = a.dim_v / 8;
      uint4 k8[kFeatMaxV4 / 2];
#pragma unroll
      for (int t = 0; t < kFeatMaxV4 / 2; ++t) {
        const int i = lane + 32 * t;
        k8[t] = i < n8 ? __ldg(vk16 + i) : make_uint4(0u, 0u, 0u, 0u);
      }
      float acc = 0.f;
#pragma unroll
      for (int t = 0; t < kFeatMaxV4 / 2; ++t) {
        const int i = lane + 32 * t;
        if (i < n8) {
          const float4 o0 = s_o[2 * i], o1 = s_o[2 * i + 1];
          const __half2* h = reinterpret_cast<const __half2*>(&k8[t]);
          const float2 a0 = __half22float2(h[0]), a1 = __half22float2(h[1]), a2 = __half22float2(h[2]), a3 = __half22float2(h[3]);
          if (a.with_vother) d_other[i] = k8[t];
          uint4 m;
          if (a.no_vmult) {
            m = make_uint4(0u, 0u, 0u, 0u);
          } else {
            const uint2 p0 = pack4(o0.x * a0.x, o0.y * a0.y, o0.z * a1.x, o0.w * a1.y);
            const uint2 p1 = pack4(o1.x * a2.x, o1.y * a2.y, o1.z * a3.x, o1.w * a3.y);
            m = make_uint4(p0.x, p0.y, p1.x, p1.y);
          }
          d_mult[i] = m;
          const float e0 = o0.x - a0.x + a.eps, e1 = o0.y - a0.y + a.eps, e2 = o0.z - a1.x + a.eps, e3 = o0.w - a1.y + a.eps;
          const float e4 = o1.x - a2.x + a.eps, e5 = o1.y - a2.y + a.eps, e6 = o1.z - a3.x + a.eps, e7 = o1.w - a3.y + a.eps;
          acc += e0 * e0 + e1 * e1 + e2 * e2 + e3 * e3 + e4 * e4 + e5 * e5 + e6 * e6 + e7 * e7;
        }
      }
      for (int i = a.dim_v + lane; i < a.dim_vp16; i += 32) {
        if (a.with_vother) row[i] = __float2half_rn(0.f);
        row[mult0 + i] = __float2half_rn(0.f);
      }
      acc = wsum(acc);
      if (lane == 0) a.dist[r] = a.no_vdist ? 0.f : sqrtf(acc);
      continue;
    }
    const float4* vk = reinterpret_cast<const float4*>(a.features + static_cast<long long>(img) * a.ld_features);
    __half* row = a.x16 + r * a.ld_x16;
    uint2* d_other = reinterpret_cast<uint2*>(row);
    uint2* d_mult = reinterpret_cast<uint2*>(row + mult0);
    float4 k[kFeatMaxV4];
#pragma unroll
    for (int t = 0; t < kFeatMaxV4; ++t) {
      const int i = lane + 32 * t;
      k[t] = i < nv ? __ldg(vk + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float acc = 0.f;
#pragma unroll
    for (int t = 0; t < kFeatMaxV4; ++t) {
      const int i = lane + 32 * t;
      if (i < nv) {
        const float4 ov = o[32 * t];
        if (a.with_vother) d_other[i] = pack4(k[t].x, k[t].y, k[t].z, k[t].w);
        d_mult[i] = a.no_vmult ? make_uint2(0u, 0u)
                               : pack4(ov.x * k[t].x, ov.y * k[t].y, ov.z * k[t].z, ov.w * k[t].w);
        const float dx = ov.x - k[t].x + a.eps, dy = ov.y - k[t].y + a.eps, dz = ov.z - k[t].z + a.eps,
                    dw = ov.w - k[t].w + a.eps;
        acc += dx * dx + dy * dy + dz * dz + dw * dw;
      }
    }
    for (int i = a.dim_v + lane; i < a.dim_vp16; i += 32) {
      if (a.with_vother) row[i] = __float2half_rn(0.f);
      row[mult0 + i] = __float2half_rn(0.f);
    }
    acc = wsum(acc);
    if (lane == 0) a.dist[r] = a.no_vdist ? 0.f : sqrtf(acc);
  }
}

__global__ void __launch_bounds__(256) ncx_features16_kernel(const Feat16Args a) {
  __shared__ float red[8];
  const int b = blockIdx.x / (a.knn + 1), slot = blockIdx.x % (a.knn + 1);
  const int tid = threadIdx.x;
  const float4* vo = reinterpret_cast<const float4*>(a.features + static_cast<long long>(a.img_idx[b * (a.knn + 1)]) * a.ld_features);
  const int nv = a.dim_v / 4;
  if (slot == 0) {
    float* row = a.xs + static_cast<long long>(b) * a.ld_xs;
    float4* d = reinterpret_cast<float4*>(row);
    for (int i = tid; i < nv; i += 256) d[i] = __ldg(vo + i);
    for (int i = a.dim_v + tid; i < a.dim_vp; i += 256) row[i] = 0.f;
    const float4* q = reinterpret_cast<const float4*>(a.q_emb + static_cast<long long>(b) * a.ld_q);
    float4* dq = reinterpret_cast<float4*>(row + a.col_q);
    for (int i = tid; i < a.dim_q / 4; i += 256) dq[i] = __ldg(q + i);
    for (int i = a.dim_q + tid; i < a.dim_qp; i += 256) row[a.col_q + i] = 0.f;
    const float4* e = reinterpret_cast<const float4*>(a.emb + a.answer_aid[b] * static_cast<long long>(a.dim_a));
    float4* de = reinterpret_cast<float4*>(row + a.col_agt);
    for (int i = tid; i < a.dim_a / 4; i += 256) de[i] = __ldg(e + i);
    return;
  }
  const long long r = static_cast<long long>(b) * a.knn + (slot - 1);
  const int img = a.img_idx[b * (a.knn + 1) + slot];
  if (a.cand_img && tid == 0) a.cand_img[r] = img;
  const float4* vk = reinterpret_cast<const float4*>(a.features + static_cast<long long>(img) * a.ld_features);
  __half* row = a.x16 + r * a.ld_x16;
  uint2* d_other = reinterpret_cast<uint2*>(row);
  const int mult0 = a.with_vother ? a.dim_vp16 : 0;
  uint2* d_mult = reinterpret_cast<uint2*>(row + mult0);
  float acc = 0.f;
  for (int i = tid; i < nv; i += 256) {
    const float4 o = __ldg(vo + i), k = __ldg(vk + i);
    if (a.with_vother) d_other[i] = pack4(k.x, k.y, k.z, k.w);
    d_mult[i] = a.no_vmult ? make_uint2(0u, 0u) : pack4(o.x * k.x, o.y * k.y, o.z * k.z, o.w * k.w);
    // F.pairwise_distance: || x1 - x2 + eps ||_2 (cx.py:300), in fp32 from the fp32 rows
    const float dx = o.x - k.x + a.eps, dy = o.y - k.y + a.eps, dz = o.z - k.z + a.eps, dw = o.w - k.w + a.eps;
    acc += dx * dx + dy * dy + dz * dz + dw * dw;
  }
  for (int i = a.dim_v + tid; i < a.dim_vp16; i += 256) {
    if (a.with_vother) row[i] = __float2half_rn(0.f);
    row[mult0 + i] = __float2half_rn(0.f);
  }
  // block sum
  acc = wsum(acc);
  if ((tid & 31) == 0) red[tid >> 5] = acc;
  __syncthreads();
  if (tid == 0) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i];
    a.dist[r] = a.no_vdist ? 0.f : sqrtf(t);
  }
}
cudaError_t launch_ncx_features16(const Feat16Args& a, cudaStream_t s) {
  if (a.batch <= 0) return cudaSuccess;
  count_launch();
  if (a.dim_v <= kFeatMaxV4 * 32 * 4 && a.batch >= 512) {   // enough questions to fill the chip with one block each
    ncx_features16_q_kernel<<<a.batch, 32 * kFeatQWarps, 0, s>>>(a);
    return cudaGetLastError();
  }
  ncx_features16_kernel<<<a.batch * (a.knn + 1), 256, 0, s>>>(a);
  return cudaGetLastError();
}

// z = sum_r hv_r[image] * hq_r[question] from the per-image table. One block per question, one HALF-WARP per image slot
// (13 warps for the 25 slots). Table row and hq row are both [R][dim_mm] (the order of the stacked W_hv / W_hq). A lane
// owns 8 consecutive output columns and walks r: for a given r the 16 lanes of a row read 16 consecutive 16-byte pieces
// (two whole 128-byte lines per load instruction) and the sum over r stays in the lane's registers. The question's hq
// row is copied to shared memory as it is (float4 copies). 45 column groups over 16 lanes = 3 rounds at 94 % lane
// utilisation; at 40 registers three blocks share an SM, so one block's staging / barrier overlaps the others' gathers.
// History (batch of 4096, 0.74 GB gathered):
//   4-byte loads, [R][dim_mm] layout ........................................................ 348 us (0.97 GB of DRAM reads)
//   GROUP-MAJOR table [dim_mm/8][R][8]: a lane reads the R pieces of its group contiguously .. 283 us -- per-lane
//     contiguity is the wrong goal: every load instruction touched 32 different 128-byte lines (lane stride 160 B), and
//     the kernel was bound by L1 wavefronts (~900 cycles per row), not by HBM
//   this layout, 16-byte loads coalesced across lanes ......................................... see DESIGN.md
// (one block per question walking its 25 rows serially took 107 us for a batch of 64: pure load latency.)
__global__ void __launch_bounds__(512, 3) ncx_mutan_apply_kernel(
    const __half* __restrict__ tab, long long ld_tab, const int* __restrict__ img_idx, int knn,
    const float* __restrict__ hq, long long ld_hq, int R, int dim_mm, __half* __restrict__ z16, long long ld_z16,
    int zp16, float* __restrict__ z_orig, long long ld_zo, int zo_pad, float* __restrict__ z32, long long ld_z32,
    int z32_pad) {
  pdl_enter();   // programmatic dependent launch: see ptx.cuh
  extern __shared__ float s_hq[];          // [R][dim_mm] floats: the question's hq row
  const int b = blockIdx.x, lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int G8 = dim_mm / 8, n4 = R * dim_mm / 4;
  const int slot = w * 2 + (lane >> 4), l16 = lane & 15;
  const bool live = slot <= knn;
  const uint4* row = reinterpret_cast<const uint4*>(tab + static_cast<long long>(img_idx[b * (knn + 1) + (live ? slot : 0)]) * ld_tab);
  {
    const float4* src = reinterpret_cast<const float4*>(hq + static_cast<long long>(b) * ld_hq);
    float4* dst = reinterpret_cast<float4*>(s_hq);
    for (int i = threadIdx.x; i < n4; i += blockDim.x) dst[i] = __ldg(src + i);
  }
  __syncthreads();
  if (!live) return;
  const long long r_out = static_cast<long long>(b) * knn + slot - 1;
  for (int g = l16; g < G8; g += 16) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    const uint4* src = row + g;
    const float* q = s_hq + g * 8;
    // (R is small: unrolled by 10 the loads of all ranks are in flight together)
#pragma unroll 10
    for (int r = 0; r < R; ++r) {
      const uint4 u = __ldg(src + static_cast<long long>(r) * G8);
      const __half2* h = reinterpret_cast<const __half2*>(&u);
      const float4 q0 = *reinterpret_cast<const float4*>(q + r * dim_mm), q1 = *reinterpret_cast<const float4*>(q + r * dim_mm + 4);
      const float2 a = __half22float2(h[0]), c = __half22float2(h[1]), e = __half22float2(h[2]), f = __half22float2(h[3]);
      acc[0] = fmaf(a.x, q0.x, acc[0]); acc[1] = fmaf(a.y, q0.y, acc[1]);
      acc[2] = fmaf(c.x, q0.z, acc[2]); acc[3] = fmaf(c.y, q0.w, acc[3]);
      acc[4] = fmaf(e.x, q1.x, acc[4]); acc[5] = fmaf(e.y, q1.y, acc[5]);
      acc[6] = fmaf(f.x, q1.z, acc[6]); acc[7] = fmaf(f.y, q1.w, acc[7]);
    }
    if (slot == 0 || z32) {
      float* dst = slot == 0 ? z_orig + static_cast<long long>(b) * ld_zo + g * 8 : z32 + r_out * ld_z32 + g * 8;
      *reinterpret_cast<float4*>(dst) = make_float4(acc[0], acc[1], acc[2], acc[3]);
      *reinterpret_cast<float4*>(dst + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
    } else {
      uint4 o;
      __half2* oh = reinterpret_cast<__half2*>(&o);
      oh[0] = __floats2half2_rn(acc[0], acc[1]); oh[1] = __floats2half2_rn(acc[2], acc[3]);
      oh[2] = __floats2half2_rn(acc[4], acc[5]); oh[3] = __floats2half2_rn(acc[6], acc[7]);
      *reinterpret_cast<uint4*>(z16 + r_out * ld_z16 + g * 8) = o;
    }
  }
  if (slot == 0) {
    for (int j = dim_mm + l16; j < zo_pad; j += 16) z_orig[static_cast<long long>(b) * ld_zo + j] = 0.f;
  } else if (z32) {
    for (int j = dim_mm + l16; j < z32_pad; j += 16) z32[r_out * ld_z32 + j] = 0.f;
  } else {
    for (int j = dim_mm + l16; j < zp16; j += 16) z16[r_out * ld_z16 + j] = __float2half_rn(0.f);
  }
}
cudaError_t launch_ncx_mutan_apply(const void* tab, long long ld_tab, const int* img_idx, int batch, int knn,
                                   const float* hq, long long ld_hq, int R, int dim_mm, void* z16, long long ld_z16,
                                   int zp16, float* z_orig, long long ld_zo, int zo_pad, cudaStream_t s, float* z32,
                                   long long ld_z32, int z32_pad) {
  if (batch <= 0) return cudaSuccess;
  if ((dim_mm % 8) || R < 1) return cudaErrorNotSupported;
  const size_t smem = static_cast<size_t>(R) * dim_mm * sizeof(float);
  const int warps = (knn + 1 + 1) / 2;
  if (smem > 64 * 1024 || warps > 16 || (ld_tab % 8) || (ld_z16 % 8) || (ld_z32 % 4) || (ld_zo % 4) || (ld_hq % 4) ||
      (reinterpret_cast<uintptr_t>(tab) & 15) || (reinterpret_cast<uintptr_t>(hq) & 15) ||
      (z16 && (reinterpret_cast<uintptr_t>(z16) & 15)) || (z32 && (reinterpret_cast<uintptr_t>(z32) & 15)) ||
      (reinterpret_cast<uintptr_t>(z_orig) & 15))
    return cudaErrorNotSupported;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(ncx_mutan_apply_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  count_launch();
  (void)launch_pdl(ncx_mutan_apply_kernel, dim3(batch), dim3(32 * warps), smem, s, static_cast<const __half*>(tab), ld_tab,
                   img_idx, knn, hq, ld_hq, R, dim_mm, static_cast<__half*>(z16), ld_z16, zp16, z_orig, ld_zo, zo_pad, z32,
                   ld_z32, z32_pad);
  return cudaGetLastError();
}

// Row softmax over fp16 logits, in place (p stays fp16: the largest probability of a row is >= 1 / n, a normal fp16
// number, and what underflows to a subnormal is < 6e-5 of it). One warp per row of up to 2048 logits, the row held in
// registers (8 x 16 bytes per lane); columns [n, n_pad) become 0.
__global__ void __launch_bounds__(256) ncx_softmax16_kernel(__half* __restrict__ x, long long ld, int rows, int n,
                                                            int n_pad) {
  const int lane = threadIdx.x & 31;
  const long long r = static_cast<long long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (r >= rows) return;
  uint4* row = reinterpret_cast<uint4*>(x + r * ld);
  const int n8 = n_pad / 8;                  // n_pad is a multiple of 64
  float v[8][8];
  float mx = -INFINITY;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int i = j * 32 + lane;
    uint4 u = make_uint4(0u, 0u, 0u, 0u);
    if (i < n8) u = row[i];
    const __half2* h = reinterpret_cast<const __half2*>(&u);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = __half22float2(h[e]);
      const int c = i * 8 + 2 * e;
      v[j][2 * e] = (i < n8 && c < n) ? f.x : -INFINITY;
      v[j][2 * e + 1] = (i < n8 && c + 1 < n) ? f.y : -INFINITY;
      mx = fmaxf(mx, fmaxf(v[j][2 * e], v[j][2 * e + 1]));
    }
  }
  mx = wmax(mx);
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      v[j][e] = __expf(v[j][e] - mx);
      sum += v[j][e];
    }
  }
  sum = wsum(sum);
  const float inv = 1.f / sum;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int i = j * 32 + lane;
    if (i < n8) {
      uint4 u;
      __half2* h = reinterpret_cast<__half2*>(&u);
#pragma unroll
      for (int e = 0; e < 4; ++e) h[e] = __floats2half2_rn(v[j][2 * e] * inv, v[j][2 * e + 1] * inv);
      row[i] = u;
    }
  }
}
cudaError_t launch_ncx_softmax16(void* x, long long ld, int rows, int n, int n_pad, cudaStream_t s) {
  if (rows <= 0) return cudaSuccess;
  if (n_pad > 2048 || (n_pad % 8) || (ld % 8) || (reinterpret_cast<uintptr_t>(x) & 15)) return cudaErrorNotSupported;
  count_launch();
  ncx_softmax16_kernel<<<(rows + 7) / 8, 256, 0, s>>>(static_cast<__half*>(x), ld, rows, n, n_pad);
  return cudaGetLastError();
}

// dst[r, c] = half(src[r, c]) for c < width, 0 for width <= c < dst_width
__global__ void ncx_to_half_kernel(const float* __restrict__ src, long long ld_src, int rows, int width,
                                   __half* __restrict__ dst, long long ld_dst, int dst_width) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<long long>(rows) * dst_width) return;
  const int r = static_cast<int>(i / dst_width), c = static_cast<int>(i % dst_width);
  dst[r * ld_dst + c] = __float2half_rn(c < width ? __ldg(src + r * ld_src + c) : 0.f);
}
cudaError_t launch_ncx_to_half(const float* src, long long ld_src, int rows, int width, void* dst, long long ld_dst,
                               int dst_width, cudaStream_t s) {
  if (rows <= 0 || dst_width <= 0) return cudaSuccess;
  count_launch();
  const long long tot = static_cast<long long>(rows) * dst_width;
  ncx_to_half_kernel<<<static_cast<unsigned>((tot + 255) / 256), 256, 0, s>>>(src, ld_src, rows, width,
                                                                              static_cast<__half*>(dst), ld_dst, dst_width);
  return cudaGetLastError();
}

}  // namespace vqacx

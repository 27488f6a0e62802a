// Fused MUTAN rank products (fusion.py:99-113 on the NeuralCX path, reached from cx.py:87 `_fusion`):
//
//     z[row, c] = sum_r ( W_hv^r x_v[row] + b_hv^r )[c] * hq_r[question(row)][c]          r < R (= 10), c < dim_mm (= 360)
//
// The unfused form materialises hv = [rows, R*dim_mm] (1.47 GB at an eval batch of 4096 questions: written by the GEMM,
// read back by the reduction). Here one CTA owns a 128-row tile of x_v, keeps it RESIDENT in shared memory
// (K = dim_hv <= 384 fp32 -> 12 k-blocks x 16 KB = 192 KB), streams the stacked weights W_hv through a small TMA ring
// and, for each 64-column slice of z, runs the R products back to back on the tensor cores (tcgen05.mma kind::tf32,
// M 128 x N 64, two TMEM accumulators ping-pong). The epilogue warps multiply each finished accumulator by the
// question-side factor hq_r (+ bias) and keep the running sum over r in registers; only z leaves the SM.
//
// Warp roles as in gemm_tf32.cu: warp 0 TMA producer, warp 1 MMA issuer, warp 2 TMEM allocator, warps 4-7 epilogue
// (thread = row).
#include "knn_internal.cuh"
#include "ncx_internal.cuh"
#include "ptx.cuh"

namespace vqacx {

namespace {
constexpr int BM = 128;
constexpr int BN = 64;              // z columns per unit
constexpr int kMaxKB = 12;          // resident A: up to 384 fp32 of K
constexpr int kABlock = BM * 128;   // 16 KB
constexpr int kBStage = BN * 128;   // 8 KB
constexpr int kMaxStages = 12;
constexpr int kAuxBytes = 1024;
constexpr int kSmemMax = 227 * 1024;

struct MArgs {
  CUtensorMap ta, tb;
  int rows, n_cand_rows, rows_per_q;
  int kb;                 // k-blocks of dim_hv (32 fp32 or 64 fp16 each: one 128 B swizzle row)
  int stages;             // depth of the W_hv ring
  int R, dim_mm, dim_mmp;
  int c_tiles, c_groups;
  const float* bhv;
  const float* hq;
  long long ld_hq;
  float* z;
  long long ld_z;
  float* z_orig;
  long long ld_zo;
};

struct Aux {
  uint64_t a_full;
  uint64_t full[kMaxStages];
  uint64_t empty[kMaxStages];
  uint64_t tfull[2];
  uint64_t tempty[2];
  uint32_t tmem_base;
};
static_assert(sizeof(Aux) <= kAuxBytes, "aux too small");
}  // namespace

// kHalf: operands are fp16 copies (x_v = tanh(.) in (-1, 1) and the frozen W_hv: same 10 mantissa bits as TF32, half the
// bytes, twice the tensor rate); otherwise fp32 read as TF32.
template <bool kHalf>
__global__ void __launch_bounds__(384, 1) ncx_mutan_fused_kernel(const __grid_constant__ MArgs args) {
  constexpr int BKE = kHalf ? 64 : 32;   // elements per k-block
  constexpr bool kBiasInK = kHalf;       // fp16 twins carry the bias as one more k column (x_v16[:, dim_hv] = 1)
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* smem_b = smem + args.kb * kABlock;
  Aux* aux = reinterpret_cast<Aux*>(smem_b + args.stages * kBStage);

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * BM;
  // this CTA's slice of the z column tiles
  const int ct0 = static_cast<int>(static_cast<long long>(args.c_tiles) * blockIdx.y / args.c_groups);
  const int ct1 = static_cast<int>(static_cast<long long>(args.c_tiles) * (blockIdx.y + 1) / args.c_groups);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&args.ta);
    tma_prefetch_desc(&args.tb);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(&aux->a_full, 1);
    for (int s = 0; s < args.stages; ++s) {
      mbar_init(&aux->full[s], 1);
      mbar_init(&aux->empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&aux->tfull[a], 1);
      mbar_init(&aux->tempty[a], 8);  // one arrive per epilogue warp
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(&aux->tmem_base, 2 * BN);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = aux->tmem_base;

  if (warp == 0) {
    // ============================== TMA producer ==============================
    if (elect_one()) {
      mbar_arrive_expect_tx(&aux->a_full, args.kb * kABlock);
      for (int kb = 0; kb < args.kb; ++kb) tma_load_2d(smem + kb * kABlock, &args.ta, &aux->a_full, kb * BKE, m0);
    }
    __syncwarp();
    int stage = 0;
    uint32_t phase = 0;
    for (int ct = ct0; ct < ct1; ++ct) {
      for (int r = 0; r < args.R; ++r) {
        const int w_row0 = r * args.dim_mm + ct * BN;   // first row of W_hv (stacked [R*dim_mm, dim_hv]) of this unit
        for (int kb = 0; kb < args.kb; ++kb) {
          mbar_wait(&aux->empty[stage], phase ^ 1);
          if (elect_one()) {
            mbar_arrive_expect_tx(&aux->full[stage], kBStage);
            tma_load_2d(smem_b + stage * kBStage, &args.tb, &aux->full[stage], kb * BKE, w_row0);
          }
          __syncwarp();
          if (++stage == args.stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ============================== MMA issuer ==============================
    constexpr uint32_t idesc = kHalf ? make_idesc(kFmtF16, kFmtF16, BM, BN, 0, 0) : make_idesc(kFmtTF32, kFmtTF32, BM, BN, 0, 0);
    mbar_wait(&aux->a_full, 0);
    tc_fence_after();
    int stage = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
    const uint32_t a_base = smem_u32(smem);
    for (int ct = ct0; ct < ct1; ++ct) {
      for (int r = 0; r < args.R; ++r) {
        mbar_wait(&aux->tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < args.kb; ++kb) {
          mbar_wait(&aux->full[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = a_base + kb * kABlock;
          const uint32_t b_addr = smem_u32(smem_b + stage * kBStage);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {   // 32 bytes of K per instruction: 8 tf32 or 16 fp16
              const uint64_t da = make_sdesc_sw128(a_addr + k * 32, 16, 1024);
              const uint64_t db = make_sdesc_sw128(b_addr + k * 32, 16, 1024);
              if (kHalf) mma_f16_ss(d_tmem, da, db, idesc, (kb > 0 || k > 0) ? 1u : 0u);
              else mma_tf32_ss(d_tmem, da, db, idesc, (kb > 0 || k > 0) ? 1u : 0u);
            }
            tc_commit(&aux->empty[stage]);
          }
          __syncwarp();
          if (++stage == args.stages) { stage = 0; phase ^= 1; }
        }
        if (elect_one()) tc_commit(&aux->tfull[acc]);
        __syncwarp();
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else if (warp >= 4) {
    // ============================== epilogue: acc * hq (+ bias), summed over r in registers ==============================
    // 8 warps: warps 4-7 own columns [0, 32) of every unit, warps 8-11 columns [32, 64) (a warp reaches TMEM lanes
    // 32*(warp % 4) .. +31). One unit costs a warp: wait -> tcgen05.ld -> 32 FMAs; the question-side factors of the
    // NEXT unit are fetched before the wait, so no global-memory latency sits between two accumulators.
    const int e = warp & 3, h = (warp - 4) >> 2;
    const int m = m0 + e * 32 + lane;
    const bool row_ok = m < args.rows;
    const bool cand = m < args.n_cand_rows;
    const long long qi = !row_ok ? 0 : (cand ? m / args.rows_per_q : m - args.n_cand_rows);
    const float* hq_row = args.hq + qi * args.ld_hq + h * 32;
    const float* b_row = args.bhv + h * 32;
    float* out_row = (cand ? args.z + static_cast<long long>(m) * args.ld_z : args.z_orig + qi * args.ld_zo) + h * 32;
    const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(e * 32) << 16) + h * 32;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int ct = ct0; ct < ct1; ++ct) {
      const int c0 = ct * BN;                              // first z column of the unit; this warp: c0 + h*32 ..
      const int n_ok = args.dim_mm - (c0 + h * 32);        // valid columns of this warp's slice (may be <= 0)
      const bool full_tile = n_ok >= 32;                   // warp-uniform
      float zacc[32], hq_next[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) zacc[j] = 0.f;
      auto load_hq = [&](int r, float (&dst)[32]) {
        const float* src = hq_row + r * args.dim_mm + c0;
        if (full_tile) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 q4 = __ldg(reinterpret_cast<const float4*>(src) + j);
            dst[4 * j] = q4.x; dst[4 * j + 1] = q4.y; dst[4 * j + 2] = q4.z; dst[4 * j + 3] = q4.w;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) dst[j] = j < n_ok ? __ldg(src + j) : 0.f;
        }
      };
      load_hq(0, hq_next);
      for (int r = 0; r < args.R; ++r) {
        float hq_cur[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) hq_cur[j] = hq_next[j];
        if (r + 1 < args.R) load_hq(r + 1, hq_next);       // in flight across the wait below
        if (!kBiasInK) {
          // fp32 / TF32 operands: the bias is added here:  (acc + b) * hq = acc * hq + b * hq
          const float* bsrc = b_row + r * args.dim_mm + c0;
#pragma unroll
          for (int j = 0; j < 32; ++j) zacc[j] = fmaf(j < n_ok ? __ldg(bsrc + j) : 0.f, hq_cur[j], zacc[j]);
        }
        mbar_wait(&aux->tfull[acc], acc_phase);
        tc_fence_after();
        uint32_t v[32];
        tmem_ld_32x32(t_lane + acc * BN, v);
        tmem_ld_wait_regs(v);
#pragma unroll
        for (int j = 0; j < 32; ++j) zacc[j] = fmaf(__uint_as_float(v[j]), hq_cur[j], zacc[j]);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&aux->tempty[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
      if (row_ok) {
        const int n_out = args.dim_mmp - (c0 + h * 32);    // columns up to the padded width are written (zeros beyond dim_mm)
        if (n_out >= 32) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            reinterpret_cast<float4*>(out_row + c0)[j] = make_float4(zacc[4 * j], zacc[4 * j + 1], zacc[4 * j + 2], zacc[4 * j + 3]);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < n_out) out_row[c0 + j] = zacc[j];
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 2 * BN);
}

// Returns cudaErrorNotSupported when the shape does not fit the resident-A design (the caller then runs the unfused
// GEMM + ncx_mutan_reduce_kernel path). xv16 / whv16 (both non-null): fp16 operands with the bias folded into K:
// k16 = roundup8(dim_hv + 1) columns, x_v16[:, dim_hv] = 1, whv16[:, dim_hv] = b_hv, zeros after (ld_whv16 >= k16).
cudaError_t launch_ncx_mutan_fused(const float* xv, long long ld_xv, const void* xv16, long long ld_xv16, int rows,
                                   int dim_hv, const float* whv, const void* whv16, long long ld_whv16,
                                   long long ld_whv, const float* bhv,
                                   const float* hq, long long ld_hq, int R, int dim_mm, int dim_mmp, int rows_per_q,
                                   int n_cand_rows, float* z, long long ld_z, float* z_orig, long long ld_zo,
                                   cudaStream_t s) {
  if (rows <= 0) return cudaSuccess;
  const int k16 = (dim_hv + 1 + 7) / 8 * 8;
  const bool half = xv16 && whv16 && (ld_xv16 % 8) == 0 && (ld_whv16 % 8) == 0 && ld_xv16 >= k16 && ld_whv16 >= k16 &&
                    !(reinterpret_cast<uintptr_t>(xv16) & 15) && !(reinterpret_cast<uintptr_t>(whv16) & 15);
  const int k_eff = half ? k16 : dim_hv;
  const int bke = half ? 64 : 32;
  const int kb = (k_eff + bke - 1) / bke;
  if (kb > kMaxKB || (dim_hv % 4) || (dim_mm % 4) || (ld_xv % 4) || (ld_whv % 4) || (ld_hq % 4) ||
      (reinterpret_cast<uintptr_t>(xv) & 15) || (reinterpret_cast<uintptr_t>(whv) & 15) ||
      (reinterpret_cast<uintptr_t>(bhv) & 15) || (reinterpret_cast<uintptr_t>(hq) & 15) ||
      (reinterpret_cast<uintptr_t>(z) & 15) || (reinterpret_cast<uintptr_t>(z_orig) & 15) || (ld_z % 4) || (ld_zo % 4))
    return cudaErrorNotSupported;
  MArgs a{};
  const int eb = half ? 2 : 4;
  if (make_tmap_2d(&a.ta, half ? xv16 : static_cast<const void*>(xv), eb, static_cast<uint64_t>(rows),
                   static_cast<uint64_t>(k_eff), static_cast<uint64_t>(half ? ld_xv16 : ld_xv) * eb, BM) != 0)
    return cudaErrorNotSupported;
  if (make_tmap_2d(&a.tb, half ? whv16 : static_cast<const void*>(whv), eb, static_cast<uint64_t>(R) * dim_mm,
                   static_cast<uint64_t>(k_eff), static_cast<uint64_t>(half ? ld_whv16 : ld_whv) * eb, BN) != 0)
    return cudaErrorNotSupported;
  a.rows = rows; a.n_cand_rows = n_cand_rows; a.rows_per_q = rows_per_q;
  a.kb = kb; a.R = R; a.dim_mm = dim_mm; a.dim_mmp = dim_mmp;
  a.stages = (kSmemMax - kAuxBytes - 1024 - kb * kABlock) / kBStage;
  if (a.stages > kMaxStages) a.stages = kMaxStages;
  if (a.stages < 2) return cudaErrorNotSupported;
  a.c_tiles = (dim_mmp + BN - 1) / BN;
  const int m_tiles = (rows + BM - 1) / BM;
  // few row tiles (a training batch): split the column tiles over CTAs so the chip is busy; many row tiles: one CTA
  // sweeps all column tiles and the x_v tile is fetched once
  int groups = 1;
  if (m_tiles < 148) groups = (148 + m_tiles - 1) / m_tiles;
  if (groups > a.c_tiles) groups = a.c_tiles;
  a.c_groups = groups;
  a.bhv = bhv; a.hq = hq; a.ld_hq = ld_hq;
  a.z = z; a.ld_z = ld_z; a.z_orig = z_orig; a.ld_zo = ld_zo;
  const int smem = kb * kABlock + a.stages * kBStage + kAuxBytes + 1024;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(ncx_mutan_fused_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemMax);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(ncx_mutan_fused_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemMax);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  count_launch();
  if (half) ncx_mutan_fused_kernel<true><<<dim3(m_tiles, groups), 384, smem, s>>>(a);
  else ncx_mutan_fused_kernel<false><<<dim3(m_tiles, groups), 384, smem, s>>>(a);
  return cudaGetLastError();
}

}  // namespace vqacx

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
constexpr int BK = 32;              // fp32 per k-block (128 B swizzle row)
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
__global__ void __launch_bounds__(256, 1) ncx_mutan_fused_kernel(const __grid_constant__ MArgs args) {
  constexpr int BKE = kHalf ? 64 : 32;   // elements per k-block
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
      mbar_init(&aux->tempty[a], 4);  // one arrive per epilogue warp
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
    // ============================== epilogue: (acc + b_hv) * hq, summed over r in registers ==============================
    const int e = warp - 4;
    const int m = m0 + e * 32 + lane;
    const bool row_ok = m < args.rows;
    const bool cand = m < args.n_cand_rows;
    const long long qi = !row_ok ? 0 : (cand ? m / args.rows_per_q : m - args.n_cand_rows);
    const float* hq_row = args.hq + qi * args.ld_hq;
    float* out_row = cand ? args.z + static_cast<long long>(m) * args.ld_z : args.z_orig + qi * args.ld_zo;
    const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(e * 32) << 16);
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int ct = ct0; ct < ct1; ++ct) {
      const int c0 = ct * BN;
      const bool full_tile = c0 + BN <= args.dim_mm;   // warp-uniform
      float zacc[BN];
#pragma unroll
      for (int j = 0; j < BN; ++j) zacc[j] = 0.f;
      for (int r = 0; r < args.R; ++r) {
        mbar_wait(&aux->tfull[acc], acc_phase);
        tc_fence_after();
        const int col0 = r * args.dim_mm + c0;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t v[32];
          tmem_ld_32x32(t_lane + acc * BN + half * 32, v);
          float hb[32], hqv[32];
          if (full_tile) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 b4 = __ldg(reinterpret_cast<const float4*>(args.bhv + col0 + half * 32) + j);
              hb[4 * j] = b4.x; hb[4 * j + 1] = b4.y; hb[4 * j + 2] = b4.z; hb[4 * j + 3] = b4.w;
              const float4 q4 = __ldg(reinterpret_cast<const float4*>(hq_row + col0 + half * 32) + j);
              hqv[4 * j] = q4.x; hqv[4 * j + 1] = q4.y; hqv[4 * j + 2] = q4.z; hqv[4 * j + 3] = q4.w;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const bool ok = c0 + half * 32 + j < args.dim_mm;
              hb[j] = ok ? __ldg(args.bhv + col0 + half * 32 + j) : 0.f;
              hqv[j] = ok ? __ldg(hq_row + col0 + half * 32 + j) : 0.f;
            }
          }
          tmem_ld_wait_regs(v);
#pragma unroll
          for (int j = 0; j < 32; ++j) zacc[half * 32 + j] = fmaf(__uint_as_float(v[j]) + hb[j], hqv[j], zacc[half * 32 + j]);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&aux->tempty[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
      if (row_ok) {
        if (c0 + BN <= args.dim_mmp) {
#pragma unroll
          for (int j = 0; j < BN / 4; ++j)
            reinterpret_cast<float4*>(out_row + c0)[j] = make_float4(zacc[4 * j], zacc[4 * j + 1], zacc[4 * j + 2], zacc[4 * j + 3]);
        } else {
#pragma unroll
          for (int j = 0; j < BN; ++j)
            if (c0 + j < args.dim_mmp) out_row[c0 + j] = zacc[j];
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 2 * BN);
}

// Returns cudaErrorNotSupported when the shape does not fit the resident-A design (the caller then runs the unfused
// GEMM + ncx_mutan_reduce_kernel path). xv16 / whv16 (both non-null, dim_hv % 8 == 0): fp16 operands.
cudaError_t launch_ncx_mutan_fused(const float* xv, long long ld_xv, const void* xv16, long long ld_xv16, int rows,
                                   int dim_hv, const float* whv, const void* whv16, long long ld_whv, const float* bhv,
                                   const float* hq, long long ld_hq, int R, int dim_mm, int dim_mmp, int rows_per_q,
                                   int n_cand_rows, float* z, long long ld_z, float* z_orig, long long ld_zo,
                                   cudaStream_t s) {
  if (rows <= 0) return cudaSuccess;
  const bool half = xv16 && whv16 && (dim_hv % 8) == 0 && (ld_xv16 % 8) == 0 && (ld_whv % 8) == 0 &&
                    !(reinterpret_cast<uintptr_t>(xv16) & 15) && !(reinterpret_cast<uintptr_t>(whv16) & 15);
  const int bke = half ? 64 : 32;
  const int kb = (dim_hv + bke - 1) / bke;
  if (kb > kMaxKB || (dim_hv % 4) || (dim_mm % 4) || (ld_xv % 4) || (ld_whv % 4) || (ld_hq % 4) ||
      (reinterpret_cast<uintptr_t>(xv) & 15) || (reinterpret_cast<uintptr_t>(whv) & 15) ||
      (reinterpret_cast<uintptr_t>(bhv) & 15) || (reinterpret_cast<uintptr_t>(hq) & 15) ||
      (reinterpret_cast<uintptr_t>(z) & 15) || (reinterpret_cast<uintptr_t>(z_orig) & 15) || (ld_z % 4) || (ld_zo % 4))
    return cudaErrorNotSupported;
  MArgs a{};
  const int eb = half ? 2 : 4;
  if (make_tmap_2d(&a.ta, half ? xv16 : static_cast<const void*>(xv), eb, static_cast<uint64_t>(rows),
                   static_cast<uint64_t>(dim_hv), static_cast<uint64_t>(half ? ld_xv16 : ld_xv) * eb, BM) != 0)
    return cudaErrorNotSupported;
  if (make_tmap_2d(&a.tb, half ? whv16 : static_cast<const void*>(whv), eb, static_cast<uint64_t>(R) * dim_mm,
                   static_cast<uint64_t>(dim_hv), static_cast<uint64_t>(ld_whv) * eb, BN) != 0)
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
  if (half) ncx_mutan_fused_kernel<true><<<dim3(m_tiles, groups), 256, smem, s>>>(a);
  else ncx_mutan_fused_kernel<false><<<dim3(m_tiles, groups), 256, smem, s>>>(a);
  return cudaGetLastError();
}

}  // namespace vqacx

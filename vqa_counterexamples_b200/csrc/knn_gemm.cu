// K1: kNN candidate generation on the 5th-gen tensor cores.
//
// Replaces the inner arithmetic of `nbrs.kneighbors(features[i:i+b])` (knn.py:49 -> sklearn ArgKmin:
// chunked  -2 X.Y^T  GEMM + |y|^2, per-row heap) with one persistent, warp-specialised sm_100a kernel:
//
//   warp 0      TMA producer   : query tile [128 x 64] + database tile [256 x 64] fp16, 128 B swizzle,
//                                4-stage mbarrier ring (cp.async.bulk.tensor -> shared)
//   warp 1      MMA issuer     : tcgen05.mma kind::f16, M=128 N=256 K=16, fp32 accumulators in TMEM,
//                                two 256-column accumulators (all 512 TMEM columns) ping-pong
//   warp 2      TMEM allocator
//   warps 4-7   epilogue       : tcgen05.ld 32 lanes x 32 columns; thread == query row;
//                                v = |y~|^2 - 2 x~.y~ ; threshold filter + warp bitonic compaction
//                                (topk.cuh). The Q x N distance matrix never exists in memory.
//
// Work decomposition: host-built list of (query tile, database-tile range, scratch slot) per CTA, so all
// CTAs sweep the database in the same direction at the same pace (database tiles are shared through L2).
#include "knn_internal.cuh"
#include "ptx.cuh"
#include "topk.cuh"
#include "knn_epilogue.cuh"

namespace vqacx {

namespace {
constexpr int kStages = 4;
constexpr int kABytes = kGemmBM * kGemmBK * 2;  // 16 KB
constexpr int kBBytes = kGemmBN * kGemmBK * 2;  // 32 KB
constexpr int kStageBytes = kABytes + kBBytes;
constexpr int kNumAcc = 2;
constexpr int kTmemCols = kNumAcc * kGemmBN;  // 512
constexpr int kUmmaK = 16;
constexpr int kAuxBytes = 4096;
constexpr int kSmemBytes = kStages * kStageBytes + kAuxBytes + kEpiStageBytes + 1024;  // +1024 for manual alignment
constexpr uint32_t kIdesc = make_idesc(kFmtF16, kFmtF16, kGemmBM, kGemmBN);

struct Aux {
  uint64_t full[kStages];
  uint64_t empty[kStages];
  uint64_t tfull[kNumAcc];
  uint64_t tempty[kNumAcc];
  uint32_t tmem_base;
  uint32_t pad[3];
  float ynorm[kNumAcc][kGemmBN];
};
static_assert(sizeof(Aux) <= kAuxBytes, "aux region too small");
}  // namespace

__global__ void __launch_bounds__(256, 1)
knn_gemm_topk_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_db,
                     const KnnWorkItem* __restrict__ items, const int* __restrict__ cta_off,
                     const float* __restrict__ db_norm_sq, int n_db, int kblocks, uint64_t* __restrict__ cand,
                     int* __restrict__ cnt) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  Aux* aux = reinterpret_cast<Aux*>(smem + kStages * kStageBytes);
  float* epi_stage = reinterpret_cast<float*>(smem + kStages * kStageBytes + kAuxBytes);

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_q);
    tma_prefetch_desc(&tm_db);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&aux->full[s], 1);
      mbar_init(&aux->empty[s], 1);
    }
    for (int a = 0; a < kNumAcc; ++a) {
      mbar_init(&aux->tfull[a], 1);
      mbar_init(&aux->tempty[a], 4);  // one arrive per epilogue warp
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(&aux->tmem_base, kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = aux->tmem_base;

  const int it_begin = cta_off[blockIdx.x], it_end = cta_off[blockIdx.x + 1];

  // Producer / MMA loops are warp-uniform (all lanes run the control flow, one elected lane issues): inside a divergent
  // `if (lane == 0)` region the compiler wraps every UTMALDG / UTCMMA in a waterfall loop (measured ~600 issue cycles per
  // k-block in gemm_tf32.cu before the same change).
  if (warp == 0) {
    {
      // ============================== TMA producer ==============================
      int stage = 0;
      uint32_t phase = 0;
      for (int it = it_begin; it < it_end; ++it) {
        const KnnWorkItem item = items[it];
        for (int t = item.db_tile_begin; t < item.db_tile_end; ++t) {
          for (int kb = 0; kb < kblocks; ++kb) {
            mbar_wait(&aux->empty[stage], phase ^ 1);
            uint8_t* sa = smem + stage * kStageBytes;
            if (elect_one()) {
              mbar_arrive_expect_tx(&aux->full[stage], kStageBytes);
              tma_load_2d(sa, &tm_q, &aux->full[stage], kb * kGemmBK, item.q_tile * kGemmBM);
              tma_load_2d(sa + kABytes, &tm_db, &aux->full[stage], kb * kGemmBK, t * kGemmBN);
            }
            __syncwarp();
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    {
      // ============================== MMA issuer ==============================
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int it = it_begin; it < it_end; ++it) {
        const KnnWorkItem item = items[it];
        for (int t = item.db_tile_begin; t < item.db_tile_end; ++t) {
          mbar_wait(&aux->tempty[acc], acc_phase ^ 1);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + acc * kGemmBN;
          for (int kb = 0; kb < kblocks; ++kb) {
            mbar_wait(&aux->full[stage], phase);
            tc_fence_after();
            const uint32_t a_addr = smem_u32(smem + stage * kStageBytes);
            const uint32_t b_addr = a_addr + kABytes;
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < kGemmBK / kUmmaK; ++k) {
                const uint64_t da = make_sdesc_sw128(a_addr + k * kUmmaK * 2, 16, 1024);
                const uint64_t db = make_sdesc_sw128(b_addr + k * kUmmaK * 2, 16, 1024);
                mma_f16_ss(d_tmem, da, db, kIdesc, (kb | k) != 0 ? 1u : 0u);
              }
              tc_commit(&aux->empty[stage]);  // frees this smem stage once the MMAs have read it
            }
            __syncwarp();
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
          if (elect_one()) tc_commit(&aux->tfull[acc]);  // accumulator complete -> epilogue
          __syncwarp();
          acc ^= 1;
          if (acc == 0) acc_phase ^= 1;
        }
      }
    }
  } else if (warp >= 4) {
    // ============================== epilogue: fused top-k' ==============================
    const int e = warp - 4;             // TMEM lane group of this warp (warp % 4)
    const int row_in_tile = e * 32 + lane;
    const int et = threadIdx.x - 128;   // 0..127 among epilogue threads
    float* vrow = epi_stage + et * kEpiStagePitch;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int it = it_begin; it < it_end; ++it) {
      const KnnWorkItem item = items[it];
      RowFilter st;
      filter_reset(st, cand + (static_cast<size_t>(item.slot) * kGemmBM + row_in_tile) * kTopkCap);
      for (int t = item.db_tile_begin; t < item.db_tile_end; ++t) {
        float* yn = aux->ynorm[acc];
        for (int i = et; i < kGemmBN; i += 128) {
          const int col = t * kGemmBN + i;
          yn[i] = col < n_db ? __ldg(db_norm_sq + col) : __int_as_float(0x7f800000);
        }
        named_bar_sync(1, 128);
        mbar_wait(&aux->tfull[acc], acc_phase);
        tc_fence_after();
        const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(e * 32) << 16) + acc * kGemmBN;
        knn_epilogue_tile<false>(st, t_addr, yn, static_cast<uint32_t>(t * kGemmBN), vrow);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&aux->tempty[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
      cnt[static_cast<size_t>(item.slot) * kGemmBM + row_in_tile] = st.cnt;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
}

cudaError_t launch_knn_gemm_topk(const CUtensorMap& tm_q, const CUtensorMap& tm_db, const KnnWorkItem* items,
                                 const int* cta_off, int n_ctas, const float* db_norm_sq, int n_db, int kblocks,
                                 uint64_t* cand, int* cnt, cudaStream_t s) {
  if (n_ctas <= 0) return cudaSuccess;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(knn_gemm_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kSmemBytes);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  count_launch();
  knn_gemm_topk_kernel<<<n_ctas, 256, kSmemBytes, s>>>(tm_q, tm_db, items, cta_off, db_norm_sq, n_db, kblocks,
                                                       cand, cnt);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// tensor-map construction through the runtime's driver entry point (no link-time libcuda dependency)
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int make_tmap_2d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t rows, uint64_t cols,
                 uint64_t row_pitch_bytes, uint32_t box_rows, bool swizzle_atom32) {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess || !p)
      return -1;
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (row_pitch_bytes & 15) || box_rows == 0 || box_rows > 256) return -2;
  const cuuint64_t gdim[2] = {cols, rows};
  const cuuint64_t gstride[1] = {row_pitch_bytes};
  const cuuint32_t box[2] = {static_cast<cuuint32_t>(128 / elem_bytes), box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUtensorMapDataType dt = elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  const CUresult r = fn(out, dt, 2, const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        swizzle_atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -3 - static_cast<int>(r);
}

int make_tmap_3d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t planes, uint64_t rows, uint64_t cols,
                 uint64_t row_pitch_bytes, uint64_t plane_pitch_bytes, uint32_t box_planes, uint32_t box_rows) {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess || !p)
      return -1;
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (row_pitch_bytes & 15) || (plane_pitch_bytes & 15) || box_rows == 0 ||
      box_rows > 256 || box_planes == 0 || box_planes > 256)
    return -2;
  const cuuint64_t gdim[3] = {cols, rows, planes};
  const cuuint64_t gstride[2] = {row_pitch_bytes, plane_pitch_bytes};
  const cuuint32_t box[3] = {static_cast<cuuint32_t>(128 / elem_bytes), box_rows, box_planes};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUtensorMapDataType dt = elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  const CUresult r = fn(out, dt, 3, const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -3 - static_cast<int>(r);
}

}  // namespace vqacx

// Internal (non-ABI) declarations shared by the kNN translation units.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace vqacx {

// indices into the int32 stats block returned by vqacx_knn_search
enum : int {
  kStatFlaggedRows = 0,      // rows whose low-precision exclusion bound failed -> redone by the exact path
  kStatUnresolvedRows = 1,   // rows still ambiguous after the exact path (ties within fp32 eps among > k'-k rows)
  kStatBoundViolations = 2,  // candidates whose exact distance fell outside the modelled error interval
  kStatUsedTensorPath = 3,   // 1 if the tcgen05 kernel produced the candidates, 0 if the exact fp32 kernel did
  kStatRerankEvaluated = 4,  // sum over every 16th query of the candidates whose exact distance was evaluated (pruning)
  kStatMirrorKeys = 5,       // symmetric self-join: sum over every 16th query of its column-side (mirror) appends
  kStatCount = 8
};

// One unit of work for the persistent tcgen05 kernel: a 128-query tile against a range of 256-row DB tiles.
struct KnnWorkItem {
  int q_tile;
  int db_tile_begin;
  int db_tile_end;
  int slot;
  int flags;   // kItemMirror: symmetric self-join block, the tile also feeds the COLUMN side (see KnnSymArgs)
};
// one mirror append counter per 32-byte sector: in a lock-step sweep every CTA pair appends to the same 256 columns at
// the same time, and counters packed 8 to a sector serialise in the L2 atomic unit
constexpr int kMirrorCntStride = 8;
constexpr int kItemMirror = 1;   // informational: launches are homogeneous (KnnSymArgs::mirror_launch)

// Symmetric self-join (queries = database, pair kernel only). d(i, j) = d(j, i), so an off-diagonal block of the
// distance matrix is computed ONCE and consumed from both sides: the row side is the usual per-thread filter, the
// column side ("mirror") tests u = |x_i|^2 - 2 x_i.y_j against the column's published threshold and appends
// (u, row i) to query j's mirror buffer.
//   gtau [n]        ordered-float (topk.cuh float_to_ordered) threshold per query: atomicMin'ed by the query's row-side
//                   owner whenever its tau tightens; every value ever stored is a valid exclusion bound (some buffer of
//                   the query holds >= kTopkKeep keys below it), so stale reads only cost extra appends
//   mbuf [n, mcap]  mirror keys, mcnt [n * kMirrorCntStride] append counters (filled by knn_mirror_scatter_kernel from the
//                   mirror warps' append logs) (a count above mcap = overflow -> the finalize kernel flags
//                   the row for the exact fallback)
// All null = plain search.
struct KnnSymArgs {
  uint32_t* gtau = nullptr;
  uint64_t* mbuf = nullptr;
  int* mcnt = nullptr;
  int mcap = 0;
  uint4* log = nullptr;    // append logs of the kernel's mirror warps: [n_regions, log_cap] entries {key lo, key hi, column}
  int* log_len = nullptr;  // [n_regions] entries written by each mirror warp (above log_cap = overflow)
  int log_cap = 0;
  int n_regions = 0;       // = CTAs x 8 mirror warps
  int mirror_warps = 8;    // 8 (two per TMEM lane quarter) or 4 (VQACX_OPT_KNN_MIRROR_WARPS, A/B runs)
  int debug = 0;           // VQACX_KNN_DEBUG bits 16 / 32 (timing experiments)
  int mirror_launch = 0;   // 1: every item of this launch is a mirrored block (the kernel's mirror warps take part)
};

// Process-wide tuning options (include/vqacx.h VQACX_OPT_*): read ONCE from the environment when the library is first
// used, changed afterwards only through vqacx_set_option. Options >= kOptFirstDebug exist in -DVQACX_DEBUG builds only
// (measurement variants and timing experiments, some of which invalidate results): in the default build they are
// compiled out and opt() returns their neutral value.
enum : int {
  kOptKnnLockstep = 1, kOptKnnSymGroups = 2, kOptKnnSymMirrorCap = 3, kOptKnnSymLogCap = 4, kOptKnnNoSymmetric = 5,
  kOptKnnNoQsort = 6, kOptNcxSerial = 7, kOptNcxNoMutanFuse = 8, kOptKnnSingleExchange = 9, kOptKnnShardSymGroups = 10, kOptKnnSymLockstep = 11, kOptNcxNoPdl = 12, kOptGemmNoPair = 13, kOptKnnMirrorWarps = 14,
  kOptFirstDebug = 16, kOptKnnSingleCta = 16, kOptKnnOverlap = 17, kOptKnnWaveTrace = 18, kOptKnnDebug = 19,
  kOptGemmDebug = 20, kOptCount = 21
};
int opt(int id);

// Optional device buffer [n_ctas * 8] of cycle counters written by the pair kernel's warp roles (tools/ only):
// [0] producer wait-empty [1] producer total [2] MMA wait-full [3] MMA wait-tempty [4] MMA total
// [5] epilogue wait-tfull [6] epilogue total. Null (default) = no instrumentation.
unsigned long long* knn_debug_profile_buffer();

// bumped by every launcher (vqacx_launch_count)
void count_launch(int n = 1);

constexpr int kGemmBM = 128;  // queries per tile (UMMA M, TMEM lanes)
constexpr int kGemmBN = 256;  // database rows per tile (UMMA N, TMEM columns per accumulator)
constexpr int kGemmBK = 64;   // fp16 elements per k-block = one 128 B swizzle row

cudaError_t launch_knn_prep(const float* x, long long n, int d, void* x_lp, float* norm_sq,
                            unsigned int* max_norm_sq_bits, int* bad_count, cudaStream_t s);
cudaError_t launch_knn_exact(const float* q, const int* q_rows, const int* n_q_ptr, int n_q, const float* db,
                             int n_db, int d, int nsplit, uint64_t* cand, int* cnt, cudaStream_t s);
cudaError_t launch_knn_finalize(const uint64_t* cand, const int* cnt, const int* tile_slot0, const int* tile_nslots,
                                int nslots_uniform, int rows_per_tile, const int* n_q_ptr, int n_q, float* cand_val,
                                int* cand_idx, cudaStream_t s);
// Symmetric self-join: also merges each row's mirror buffer; rows whose mirror buffer overflowed are appended to
// flag_rows (exact fallback).
cudaError_t launch_knn_finalize_sym(const uint64_t* cand, const int* cnt, const int* tile_slot0, const int* tile_nslots,
                                    int rows_per_tile, int n_q, const KnnSymArgs& sym, float* cand_val, int* cand_idx,
                                    int* flag_count, int* flag_rows, int* stats, cudaStream_t s);
// gtau = ordered(+inf), mcnt = 0, log_len = 0
cudaError_t launch_knn_sym_reset(const KnnSymArgs& sym, int n, cudaStream_t s);
// Files the mirror warps' log entries under their queries (mbuf / mcnt); *overflow_flag |= 1 if a log overflowed.
cudaError_t launch_knn_mirror_scatter(const KnnSymArgs& sym, int* overflow_flag, cudaStream_t s);
cudaError_t launch_knn_rerank(const float* q, const int* q_rows, const int* n_q_ptr, int n_q, const float* db,
                              int n_db, int d, const float* cand_val, const int* cand_idx, const float* q_norm_sq_lp,
                              const unsigned int* db_max_norm_sq_bits, float eps_d, float d_abs, float eps_sq_abs,
                              float eps_sq_rel, int k, long long db_row0, double* out_dist, long long* out_idx,
                              int* flag_count, int* flag_rows, int* stats, const float* ext_bound, cudaStream_t s,
                              int out_ld = 0);
void knn_support_prefer_max_shared_carveout();
cudaError_t launch_knn_finalize_range(const uint64_t* cand, const int* cnt, const int* tile_slot0, const int* tile_nslots,
                                      int nslots_uniform, int rows_per_tile, const int* n_q_ptr, int row0, int row1,
                                      float* cand_val, int* cand_idx, cudaStream_t s);
cudaError_t launch_knn_rerank_range(const float* q, const int* q_rows, const int* n_q_ptr, int row0, int row1,
                                    const float* db, int n_db, int d, const float* cand_val, const int* cand_idx,
                                    const float* q_norm_sq_lp, const unsigned int* db_max_norm_sq_bits, float eps_d,
                                    float d_abs, float eps_sq_abs, float eps_sq_rel, int k, long long db_row0,
                                    double* out_dist, long long* out_idx, int* flag_count, int* flag_rows, int* stats,
                                    const float* ext_bound, cudaStream_t s, bool cand_by_qrow = false,
                                    const uint32_t* excl_thr = nullptr, int out_ld = 0);
// per-query thresholds <-> plain floats (threshold exchange of the block-partitioned symmetric self-join)
cudaError_t launch_knn_thr_export(const uint32_t* gtau, int n, float* thr, cudaStream_t s);
cudaError_t launch_knn_thr_import(const float* thr, int n, uint32_t* gtau, cudaStream_t s);
cudaError_t launch_knn_flag_to_float(const int* flag, float* dst, int n_fill, cudaStream_t s);
// candidate exchange of the block-partitioned self-join (knn_support.cu): packed per-query rows of kPackInts int32
constexpr int kPackInts = 130;   // 2 * kTopkKeep + 2
cudaError_t launch_knn_cand_pack(const float* cand_val, const int* cand_idx, const uint32_t* gtau, int n, int rows_alloc,
                                 const int* overflow_flag, const int* mcnt, int mcap, int* pack, cudaStream_t s);
cudaError_t launch_knn_cand_merge(const int* recv, int G, int per, int n_slice, float* cand_val, int* cand_idx,
                                  uint32_t* excl_ord, cudaStream_t s);
// perm [n_q] = the query rows ordered so that queries with overlapping candidate lists are adjacent (re-rank L2 reuse);
// key [n_q], hist [n_db + 1] scratch
cudaError_t launch_knn_locality_order(const int* cand_idx, int n_q, int n_db, int* key, int* hist, int* perm,
                                      cudaStream_t s);
cudaError_t launch_knn_local_hi(const float* q, int n_q, int d, const float* cand_val, const int* cand_idx,
                                const float* q_norm_sq_lp, const unsigned int* db_max_norm_sq_bits, float eps_d,
                                float d_abs, float eps_sq_abs, float eps_sq_rel, int k, float* hi_out, cudaStream_t s);
cudaError_t launch_knn_kth_bound(const float* hi_all, int G, long long n_q, int k, float* bound, cudaStream_t s);
cudaError_t launch_knn_merge(const double* dist, const long long* idx, int G, long long n_q, int k,
                             double* out_dist, long long* out_idx, cudaStream_t s, int dist_ld = 0,
                             int* fail_count = nullptr, int* fail_rows = nullptr);

// tcgen05 distance GEMM + fused top-k'. Tensor maps are built by the caller (abi.cu).
cudaError_t launch_knn_gemm_topk(const CUtensorMap& tm_q, const CUtensorMap& tm_db, const KnnWorkItem* items,
                                 const int* cta_off, int n_ctas, const float* db_norm_sq, int n_db, int kblocks,
                                 uint64_t* cand, int* cnt, cudaStream_t s);

// CTA-pair (cta_group::2) version: items address 256-query tiles, one item list per pair; tm_db box = 128 rows.
cudaError_t launch_knn_gemm_topk_pair(const CUtensorMap& tm_q, const CUtensorMap& tm_db, const KnnWorkItem* items,
                                      const int* pair_off, int n_pairs, const float* db_norm_sq, int n_db,
                                      int kblocks, uint64_t* cand, int* cnt, cudaStream_t s);
// Same, restricted to items [item_skip, item_skip + item_max) of every pair's list (one launch per GEMM wave).
cudaError_t launch_knn_gemm_topk_pair_range(const CUtensorMap& tm_q, const CUtensorMap& tm_db, const KnnWorkItem* items,
                                            const int* pair_off, int n_pairs, const float* db_norm_sq, int n_db,
                                            int kblocks, uint64_t* cand, int* cnt, int item_skip, int item_max,
                                            cudaStream_t s);
// Lock-step throttle of the pair kernel (knn_gemm_pair.cu producer loop): kLockstepMaxPairs ints of workspace that the
// launches of the calling host thread use as their progress array (reset before every launch); null = off.
constexpr int kLockstepMaxPairs = 96;
constexpr int kLockstepWindowDefault = 8;   // database tiles a pair may run ahead of the slowest (VQACX_KNN_LOCKSTEP)
void knn_pair_set_lockstep_buffer(int* buf);
// Symmetric self-join launch (tm_q and tm_db describe the same matrix; db_norm_sq doubles as the query norms).
cudaError_t launch_knn_gemm_topk_pair_sym(const CUtensorMap& tm_q, const CUtensorMap& tm_db, const KnnWorkItem* items,
                                          const int* pair_off, int n_pairs, const float* db_norm_sq, int n_db,
                                          int kblocks, uint64_t* cand, int* cnt, const KnnSymArgs& sym, cudaStream_t s);

// Build a 2-D row-major tensor map with 128 B swizzle (box = box_rows x 128 bytes of the inner dim).
// elem_bytes is 2 (fp16) or 4 (fp32/tf32). Returns 0 on success.
// swizzle_atom32 = true selects the "128 B swizzle, 32 B atom" pattern that MN-major TF32 UMMA operands need.
int make_tmap_2d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t rows, uint64_t cols,
                 uint64_t row_pitch_bytes, uint32_t box_rows, bool swizzle_atom32 = false);
// Same for a tensor of rank 3 ([planes][rows][cols], cols innermost); box = box_planes x box_rows x 128 bytes.
int make_tmap_3d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t planes, uint64_t rows, uint64_t cols,
                 uint64_t row_pitch_bytes, uint64_t plane_pitch_bytes, uint32_t box_planes, uint32_t box_rows);

// Kernel launch with the programmatic-stream-serialization attribute (ptx.cuh pdl_*): used for the chain of small
// dependent launches of a NeuralCX step, where the gap between a kernel's last CTA and the next kernel's first
// instruction is a visible part of the step. VQACX_OPT_NCX_NO_PDL = 1 launches the same kernels without it (A/B).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = opt(kOptNcxNoPdl) ? 0 : 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

}  // namespace vqacx

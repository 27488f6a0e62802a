/* vqacx.h — C ABI of the B200-native VQA-CX hot path (libvqacx.so).
 *
 * The reference (gabegrand/VQA-Counterexamples) is pure Python and has no FFI of its own; the calls
 * below are what its two hot-path stages bind to when they are moved onto the GPU:
 *
 *   stage 1  knn.py:41-56            NearestNeighbors(n_neighbors=k).fit(X); kneighbors(X[i:i+b])
 *   stage 2  vqa/models/cx.py:261-333 NeuralModel.forward  (+ CXModelBase.vqa_forward cx.py:64-104,
 *            MutanFusion.forward fusion.py:78-121, AbstractNoAtt._classif noatt.py:24-29)
 *            counterexamples.py:310,334 CrossEntropyLoss(size_average=False), :501-506 recallAtK,
 *            :275-276,336-339 Adam
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host;
 *   - the caller owns all memory including workspaces (sizes from the *_bytes functions);
 *   - `stream` is a cudaStream_t passed as void*; work is enqueued on it. Functions documented as
 *     "synchronises" wait for that stream before returning, all others are asynchronous;
 *   - return value: 0 = ok, < 0 = VQACX_E* below, > 0 = a cudaError_t value;
 *   - no function throws, exits, allocates device memory or keeps global mutable state other than
 *     cached function attributes.
 */
#ifndef VQACX_H_
#define VQACX_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VQACX_OK 0
#define VQACX_EINVAL (-1)      /* bad shape / null pointer / unsupported size                       */
#define VQACX_EALIGN (-2)      /* pointer or pitch not 16-byte aligned                              */
#define VQACX_EWORKSPACE (-3)  /* workspace too small                                               */
#define VQACX_ETMA (-4)        /* cuTensorMapEncodeTiled unavailable or rejected the descriptor     */
#define VQACX_ERANGE (-5)      /* input not representable in the low-precision copy (inf/nan)       */

#define VQACX_KNN_MAX_K 64     /* k' kept per row by the candidate generator; k must be <= this     */
#define VQACX_KNN_FORCE_EXACT 1u /* flags bit: skip the tensor-core path, use the exact fp32 kernel */

#define VQACX_KNN_NPHASES 4    /* phase_ms_host: [0] tcgen05/exact candidate kernel, [1] finalize,
                                * [2] fp64 re-rank, [3] exact fallback rounds (all in milliseconds)  */
#define VQACX_KNN_NSTATS 8
/* stats_host[0] rows redone by the exact path, [1] rows still tie-ambiguous afterwards,
 * [2] error-model violations seen, [3] 1 if the tcgen05 kernel generated the candidates            */

const char* vqacx_version(void);
const char* vqacx_error_string(int code);
/* Number of kernels this library has launched in this process so far (monotonic; bench.py's gpu_launches). */
int64_t vqacx_launch_count(void);

/* ------------------------------------------------------------------------------------------------
 * Stage 1: exact k-nearest-neighbour search (replaces sklearn NearestNeighbors as used by knn.py)
 * ---------------------------------------------------------------------------------------------- */

/* Low-precision side arrays of a row-major [n, d] float32 matrix: fp16 copy (round to nearest),
 * squared norms of the rounded rows, and max squared norm (float bits in a uint32, caller zeroes it
 * before the first call; several calls may accumulate into it, e.g. one per shard).
 * Replaces: the `fit` step knn.py:41-42 (sklearn stores X; we also stage what the GEMM needs).
 * Requires d % 4 == 0. bad_count (int32, caller zeroes) counts rows with non-finite fp16 values. */
int vqacx_knn_prep(const float* x, int64_t n, int d, void* x_lp, float* norm_sq, uint32_t* max_norm_sq_bits,
                   int32_t* bad_count, void* stream);

/* Workspace bytes for one vqacx_knn_search call with these sizes (valid for any k <= VQACX_KNN_MAX_K). */
size_t vqacx_knn_workspace_bytes(int64_t n_q, int64_t n_db, int d);

/* Exact top-k of every query row against the database rows, ascending by (distance, index).
 * Replaces: `nbrs.kneighbors(features[i:i+b])` knn.py:49-51 (any b, including the whole matrix).
 *   db_f32 [n_db,d], db_lp fp16 [n_db,d], db_norm_sq [n_db], db_max_norm_sq_bits [1]  (from vqacx_knn_prep)
 *   q_f32 [n_q,d], q_lp, q_norm_sq  likewise (pass the db arrays again for a self-join)
 *   db_row0       added to every returned index (row offset of this shard in the global database)
 *   out_dist      float64 [n_q,k]  Euclidean distances (sqrt), as sklearn returns
 *   out_idx       int64   [n_q,k]  global row indices, -1 / +inf where fewer than k rows exist
 *   stats_host    optional HOST int32[VQACX_KNN_NSTATS]
 *   phase_ms_host optional HOST float[VQACX_KNN_NPHASES]; when non-null CUDA events are recorded on
 *                 `stream` around each phase and their durations returned (measurement only)
 * Synchronises `stream` (it must read the count of rows that need the exact fallback). */
int vqacx_knn_search(const float* db_f32, const void* db_lp, const float* db_norm_sq,
                     const uint32_t* db_max_norm_sq_bits, int64_t n_db, int64_t db_row0, const float* q_f32,
                     const void* q_lp, const float* q_norm_sq, int64_t n_q, int d, int k, double* out_dist,
                     int64_t* out_idx, void* workspace, size_t workspace_bytes, uint32_t flags,
                     int32_t* stats_host, float* phase_ms_host, void* stream);

/* Merge G per-shard results laid out [G, n_q, k] (what all_gather produces) into the global top-k.
 * New (the reference is single-process); G * k <= 256. Asynchronous. */
int vqacx_topk_merge(const double* dist, const int64_t* idx, int G, int64_t n_q, int k, double* out_dist,
                     int64_t* out_idx, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Stage 2: NeuralCX candidate scoring (replaces vqa/models/cx.py NeuralModel.forward and its autograd)
 * ---------------------------------------------------------------------------------------------- */

/* C[m,n] (=|+=) act(A . B^T + bias) on the tensor cores in TF32 with fp32 accumulation: the one dense product
 * behind every nn.Linear / torch.bmm of the path (cx.py:87-88, 281-282, 322-327; fusion.py:86-111).
 *   a_mn_major = 0: A is row-major [m, k] (lda elements between rows); 1: A is row-major [k, m]
 *   b_mn_major = 0: B is row-major [n, k] (an nn.Linear weight);        1: B is row-major [k, n]
 *   act: 0 none, 1 tanh, 2 relu.  split_k > 1 or accumulate = 1: results are atomically added into C
 *   (split_k > 1 needs a zeroed C and act = 0).  bn: N tile (0 = choose; multiple of 32 <= 256).
 * lda, ldb multiples of 4; a, b 16-byte aligned. Asynchronous. */
int vqacx_gemm_tf32(int m, int n, int k, const float* a, int64_t lda, int a_mn_major, const float* b, int64_t ldb,
                    int b_mn_major, float* c, int64_t ldc, const float* bias, int act, int split_k, int bn,
                    int accumulate, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VQACX_H_ */

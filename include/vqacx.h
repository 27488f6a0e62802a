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
 *   - the caller owns all memory including workspaces (sizes from the *_bytes functions); the one exception is
 *     vqacx_gemm_tf32_splitk_static, whose scratch is a fixed device array inside the library;
 *   - `stream` is a cudaStream_t passed as void*; work is enqueued on it. Functions documented as
 *     "synchronises" wait for that stream before returning, all others are asynchronous;
 *   - return value: 0 = ok, < 0 = VQACX_E* below, > 0 = a cudaError_t value;
 *   - no function throws, exits or allocates device memory.
 *
 * State the library owns (everything else is the caller's):
 *   - cached kernel attributes (dynamic shared memory opt-in), the launch counter, and the option table below
 *     (environment read ONCE at first use; afterwards only vqacx_set_option changes it);
 *   - per host thread and device, created lazily: three non-blocking SIDE STREAMS + 32 events that the NeuralCX
 *     forward / backward fork independent products onto and join back into the caller's stream (under CUDA-graph
 *     capture they become parallel branches of the caller's graph), and ONE COPY STREAM + 17 events used by
 *     vqacx_knn_selfjoin_host to overlap the chunked upload with the search. All work forked there is joined back
 *     into `stream` before the call returns to it, so the caller's stream order is the only order it has to reason
 *     about; the streams are never destroyed (process lifetime). VQACX_OPT_NCX_SERIAL = 1 keeps NeuralCX on the
 *     caller's stream alone;
 *   - the progress array of the kNN lock-step throttle lives in the CALLER's workspace; the pointer handed to the
 *     next launch is remembered per host thread between carving the workspace and launching (same call).
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
#define VQACX_ESTATIC (-6)     /* problem needs more than the static scratch of a *_static entry    */

#define VQACX_KNN_MAX_K 64     /* k' kept per row by the candidate generator; k must be <= this     */
#define VQACX_KNN_FORCE_EXACT 1u /* flags bit: skip the tensor-core path, use the exact fp32 kernel */
#define VQACX_KNN_NO_SYMMETRIC 2u /* flags bit: a self-join (queries == database pointers) still runs the plain
                                     sweep instead of the symmetric one (A/B runs; results are identical) */

#define VQACX_KNN_NPHASES 4    /* phase_ms_host: [0] tcgen05/exact candidate kernel, [1] finalize,
                                * [2] fp64 re-rank, [3] exact fallback rounds (all in milliseconds)  */
#define VQACX_KNN_NSTATS 8
/* stats_host[0] rows redone by the exact path, [1] rows still tie-ambiguous afterwards,
 * [2] error-model violations seen, [3] 1 if the tcgen05 kernel generated the candidates,
 * [4] candidates whose exact distance was evaluated, summed over every 16th query (re-rank pruning)  */

const char* vqacx_version(void);   /* contains "debug build" iff compiled with -DVQACX_DEBUG */

/* Process-wide tuning options. The defaults are the product configuration; the options exist for tests (shrinking a
 * buffer to exercise an overflow path) and A/B measurements, and none of them changes a result. Initial values come
 * from the environment variable named beside each, read once when the library is first used.
 * vqacx_set_option returns the previous value (>= 0) or VQACX_EINVAL; not synchronised with concurrent calls. */
#define VQACX_OPT_KNN_LOCKSTEP 1        /* VQACX_KNN_LOCKSTEP: lock-step window of the pair kernel in database tiles, 0 = off (8) */
#define VQACX_OPT_KNN_SYM_GROUPS 2      /* VQACX_KNN_SYM_GROUPS: row groups of the symmetric self-join, odd, 3..15 (7)      */
#define VQACX_OPT_KNN_SYM_MIRROR_CAP 3  /* VQACX_KNN_SYM_MIRROR_CAP: mirror keys kept per query (512)                       */
#define VQACX_OPT_KNN_SYM_LOG_CAP 4     /* VQACX_KNN_SYM_LOG_CAP: entries per mirror-warp append log, 0 = sized from n (0)  */
#define VQACX_OPT_KNN_NO_SYMMETRIC 5    /* VQACX_KNN_NO_SYMMETRIC: 1 = self-joins run the plain sweep (0)                   */
#define VQACX_OPT_KNN_NO_QSORT 6        /* VQACX_KNN_NO_QSORT: 1 = re-rank in list order instead of locality order (0)      */
#define VQACX_OPT_NCX_SERIAL 7          /* VQACX_NCX_SERIAL: 1 = no side streams in the NeuralCX step (0)                   */
#define VQACX_OPT_NCX_NO_MUTAN_FUSE 8   /* VQACX_NO_MUTAN_FUSE: 1 = unfused MUTAN product + reduce (0)                      */
#define VQACX_OPT_KNN_SINGLE_EXCHANGE 9 /* VQACX_KNN_SINGLE_EXCHANGE: host protocol hint read by knn.py (0)                 */
#define VQACX_OPT_KNN_SHARD_SYM_GROUPS 10 /* VQACX_KNN_SHARD_SYM_GROUPS: row groups of the block-partitioned self-join over
                                          * several ranks, 0 = auto (15 for world > 1) (0)                              */
#define VQACX_OPT_KNN_SYM_LOCKSTEP 11   /* VQACX_KNN_SYM_LOCKSTEP: lock-step window of the MIRRORED launch of the symmetric
                                          * self-join, whose tiles take unequal time (24)                              */
#define VQACX_OPT_NCX_NO_PDL 12         /* VQACX_NCX_NO_PDL: 1 = the NeuralCX kernels are launched without programmatic
                                          * dependent launch (0)                                                      */
#define VQACX_OPT_GEMM_NO_PAIR 13       /* VQACX_GEMM_NO_PAIR: 1 = large-M fp16 products run one CTA per tile instead of
                                          * CTA pairs sharing the weight tile (0)                                      */
#define VQACX_OPT_KNN_MIRROR_WARPS 14   /* VQACX_KNN_MIRROR_WARPS: column-side warps per CTA of the mirrored launch, 8 or 4 (8) */
/* Options 16..20 (single-CTA kernel, wave overlap, wave trace, kernel timing experiments -- some of which skip work and
 * INVALIDATE results) exist only in a -DVQACX_DEBUG build; the default build rejects them with VQACX_EINVAL and carries
 * neither their branches nor the clock64() role counters in its kernels. */
int vqacx_get_option(int option);
int vqacx_set_option(int option, int value);
const char* vqacx_error_string(int code);
/* Number of kernels this library has launched in this process so far (monotonic; bench.py's gpu_launches). */
int64_t vqacx_launch_count(void);

/* Measurement aid (tools/ only, -DVQACX_DEBUG builds; ignored otherwise): device buffer of 8 uint64 cycle counters per
 * CTA written by the tcgen05 kernels' warp roles (see csrc/knn_internal.cuh); NULL (the default) = off. */
void vqacx_debug_set_knn_profile(void* device_buf);

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

/* Row-sharded search (one process per GPU, database rows split over the ranks) runs the search in two halves around
 * one small exchange, so that the float64 re-rank gathers ~k/G rows per query instead of ~k:
 *   phase 1: candidates + finalize; hi_io [n_q, k] (float, out) = upper bounds on the TRUE distances of this shard's
 *            k best candidates (+inf padding); out_dist / out_idx may be NULL
 *   (caller: all_gather hi over the ranks, vqacx_knn_kth_bound -> bound [n_q])
 *   phase 2: re-rank of the candidates phase 1 left in `workspace`; hi_io [n_q] (in) = the agreed upper bound on the
 *            global k-th distance. Entries of out_dist / out_idx that cannot belong to the global top k are +inf / -1.
 * Same arguments as vqacx_knn_search otherwise; both phases must use the same workspace, queries and database. */
int vqacx_knn_search_phase(const float* db_f32, const void* db_lp, const float* db_norm_sq,
                           const uint32_t* db_max_norm_sq_bits, int64_t n_db, int64_t db_row0, const float* q_f32,
                           const void* q_lp, const float* q_norm_sq, int64_t n_q, int d, int k, double* out_dist,
                           int64_t* out_idx, void* workspace, size_t workspace_bytes, uint32_t flags,
                           int32_t* stats_host, float* phase_ms_host, void* stream, int phase, float* hi_io);

/* The knn.py workload in one call with a HOST feature matrix (knn.py:35-56: load, fit(features), kneighbors over the
 * same rows): x_host [n, d] float32 (pinned memory gives DMA speed) is uploaded in n_chunks pieces on an internal copy
 * stream; each piece is prepped as it lands (vqacx_knn_prep) and the candidate GEMM runs, shell by shell, every
 * (query rows, database rows) block whose rows are already resident, so all but the last shell of the search overlaps
 * the upload. Results as vqacx_knn_search (self-join: db = queries, db_row0 = 0) in DEVICE memory; x_dev [n, d],
 * lp_dev [n, d] fp16, norm_sq_dev [n], max_norm_sq_bits_dev [1], bad_count_dev [1] receive the fitted index (valid for
 * later vqacx_knn_search calls). d % 64 == 0. Returns VQACX_ERANGE if a row is not representable in fp16.
 * Synchronises `stream` before returning. n_chunks = 1 is upload-then-search. */
size_t vqacx_knn_selfjoin_workspace_bytes(int64_t n, int d, int n_chunks);
int vqacx_knn_selfjoin_host(const float* x_host, int64_t n, int d, int k, int n_chunks, float* x_dev, void* lp_dev,
                            float* norm_sq_dev, uint32_t* max_norm_sq_bits_dev, int32_t* bad_count_dev,
                            double* out_dist, int64_t* out_idx, void* workspace, size_t workspace_bytes,
                            int32_t* stats_host, void* stream);

/* Self-join over `world` ranks, one process per GPU, EVERY rank holding the whole matrix (the queries of a self-join
 * are replicated anyway): instead of a row shard each rank takes a contiguous share of the blocks of the SYMMETRIC
 * sweep -- every unordered pair of row groups is computed once, by one rank, and feeds both its row side and its column
 * side -- so the tensor work per rank is (groups + 1) / (2 groups) / world of the matrix instead of 1 / world. Each rank
 * ends with, per query, the exact top-k over the database blocks it covered; the lists are exchanged and merged like
 * row-shard results (vqacx_topk_merge). Three phases around two small exchanges the caller performs:
 *   phase 1: this rank's share of the diagonal blocks; thr_io [n] (float, out) = per-query candidate thresholds
 *            (+inf where this rank saw nothing)                       -> caller: all_reduce(MIN) of thr_io over the ranks
 *   phase 2: thr_io (in) = the reduced thresholds; this rank's share of the off-diagonal blocks; hi_io [(n + 1), k] (float,
 *            out) = upper bounds on the true distances of its k best candidates per query; row n, column 0 = 1 if one
 *            of its append logs overflowed                            -> caller: all_gather, vqacx_knn_kth_bound -> bound
 *   phase 3: hi_io (in) = bound [n]; float64 re-rank pruned by it -> out_dist [n, k + 1] / out_idx [n, k] (+inf / -1
 *            where a candidate cannot belong to the global top k; column k of out_dist = this rank's exclusion bound, see
 *            vqacx_topk_merge_checked, which completes the proof after the exchange). Synchronises `stream`. If ANY
 *            rank reported an overflow the caller must discard the result and fall back to vqacx_knn_search_phase on
 *            row shards.
 * x_f32 / x_lp / norm_sq / max_norm_sq_bits: the whole matrix and its vqacx_knn_prep side arrays (all n rows). All
 * phases use the same workspace. _supported: 1 if this (n, d, world) can take this path (d % 64 == 0, enough tiles). */
int vqacx_knn_selfjoin_shard_supported(int64_t n, int d, int world);
size_t vqacx_knn_selfjoin_shard_workspace_bytes(int64_t n, int d, int world);
int vqacx_knn_selfjoin_shard(const float* x_f32, const void* x_lp, const float* norm_sq,
                             const uint32_t* max_norm_sq_bits, int64_t n, int d, int k, int rank, int world, int phase,
                             float* thr_io, float* hi_io, double* out_dist, int64_t* out_idx, void* workspace,
                             size_t workspace_bytes, int32_t* stats_host, float* phase_ms_host, void* stream);

/* CANDIDATE exchange (the protocol knn.py uses): after phase 2 (called with hi_io = NULL) a rank holds, per query, the
 * <= 64 best candidates among the blocks it covered. They travel to the rank that OWNS the query, which merges the G
 * lists and runs the ordinary re-rank + exactness proof on its n / world queries -- instead of every rank re-ranking
 * all n queries and exchanging bounds and results (phase 3 + vqacx_topk_merge_checked).
 *   _pack : pack_out [rows_alloc >= n, VQACX_KNN_PACK_INTS] int32 = per query 64 approximate values (float bits), 64 rows
 *           (-1 = none), the exclusion bound (float bits; -inf if this rank dropped keys), pad. Rows [n, rows_alloc) are
 *           exchange padding. The caller performs ONE all_to_all over equal slices of `per` rows.
 *   _owner: pack_recv [world, per, VQACX_KNN_PACK_INTS] = everybody's rows of this rank's query slice [q0, q0 + n_slice);
 *           out_dist / out_idx [n_slice, k] = the final neighbours of those queries. Synchronises `stream`. */
#define VQACX_KNN_PACK_INTS 130
int vqacx_knn_selfjoin_shard_pack(int64_t n, int d, int world, int64_t rows_alloc, int32_t* pack_out, void* workspace,
                                  size_t workspace_bytes, void* stream);
int vqacx_knn_selfjoin_shard_owner(const float* x_f32, const float* norm_sq, const uint32_t* max_norm_sq_bits, int64_t n,
                                   int d, int k, int world, const int32_t* pack_recv, int64_t per, int64_t q0,
                                   int64_t n_slice, double* out_dist, int64_t* out_idx, void* workspace,
                                   size_t workspace_bytes, int32_t* stats_host, float* phase_ms_host, void* stream);

/* bound[q] = k-th smallest of the G*k phase-1 bounds hi_all [G][n_q][k] of query q (G*k <= 256). */
int vqacx_knn_kth_bound(const float* hi_all, int G, int64_t n_q, int k, float* bound, void* stream);

/* Merge G per-shard results laid out [G, n_q, k] (what all_gather / all_to_all produce) into the global top-k by
 * (distance, index); an index reported by several shards is kept once. New (the reference is single-process);
 * G * k <= 256. Asynchronous. */
int vqacx_topk_merge(const double* dist, const int64_t* idx, int G, int64_t n_q, int k, double* out_dist,
                     int64_t* out_idx, void* stream);

/* Merge + completion of the exactness proof, for the lists vqacx_knn_selfjoin_shard produces: dist is [G, n_q, k + 1],
 * column k of a row = that rank's lower bound on the true distance of every (query, row) pair it EXCLUDED (+inf: none);
 * idx is [G, n_q, k]. After the merge the k-th merged distance must lie below every rank's bound, otherwise an excluded
 * pair could belong to the top k: such rows are appended to fail_rows [n_q] (count in fail_count [1], caller zeroes it)
 * and the caller recomputes them with the exact search (flags bit VQACX_KNN_FORCE_EXACT). Asynchronous. */
int vqacx_topk_merge_checked(const double* dist, const int64_t* idx, int G, int64_t n_q, int k, double* out_dist,
                             int64_t* out_idx, int32_t* fail_count, int32_t* fail_rows, void* stream);

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

/* Same product with DETERMINISTIC split-K: each of `split_k` K-slices (0 = chosen by the cost model) stores its raw
 * partial tile to `scratch` (>= split_k * m * ceil32(n) floats) and a second kernel sums the slices in order and
 * applies bias / activation. This is how the small-M products of a batch-64 NeuralCX step fill all 148 SMs. */
int vqacx_gemm_tf32_splitk(int m, int n, int k, const float* a, int64_t lda, int a_mn_major, const float* b,
                           int64_t ldb, int b_mn_major, float* c, int64_t ldc, const float* bias, int act,
                           int split_k, int bn, float* scratch, size_t scratch_bytes, void* stream);

/* The same deterministic split-K product for callers that cannot provide scratch memory (set-ups under CUDA-graph
 * capture, serving loops and harnesses that forbid cudaMalloc): the partial tiles go to a file-scope device array of
 * VQACX_STATIC_SCRATCH_BYTES that every GPU loading the library holds until the process exits. NO COUNTERPART IN THE
 * REFERENCE. Calls no cudaMalloc / cudaFree / allocating routine.
 *   - limit: max(split_k, 2) * m * ceil32(n) * 4 bytes <= VQACX_STATIC_SCRATCH_BYTES, else VQACX_ESTATIC and nothing
 *     is launched (split_k = 0: the cost model chooses among the splits that fit);
 *   - every call on a device shares the one array: calls that overlap on that GPU (different streams, or graph
 *     branches) RACE -- keep them in one stream order. Nothing in it has to start at zero: each K-slice overwrites
 *     its whole partial tile before the second kernel reads it, so the previous call's contents are harmless;
 *   - results are bit-identical to vqacx_gemm_tf32_splitk given a scratch of VQACX_STATIC_SCRATCH_BYTES.
 * Other arguments, alignment rules and return codes as above. Asynchronous. */
#define VQACX_STATIC_SCRATCH_BYTES (8u << 20)
int vqacx_gemm_tf32_splitk_static(int m, int n, int k, const float* a, int64_t lda, int a_mn_major, const float* b,
                                  int64_t ldb, int b_mn_major, float* c, int64_t ldc, const float* bias, int act,
                                  int split_k, int bn, void* stream);


/* Model dimensions. Reference sources: vqa_model.opt['fusion'] (cx.py:230-232), len(vocab_answers), dim_a = 2400
 * (cx.py:234-235), dim_h / n_layers (cx.py:237-238), knn_size = 24 (cx.py:226). Every dim_* must be a multiple of 4. */
typedef struct {
  int32_t dim_v, dim_q, dim_hv, dim_hq, dim_mm, R, n_ans, dim_a, dim_h, n_layers, knn;
} VqacxNcxDims;

/* Frozen VQA head (MutanNoAtt), device pointers in the reference's own layouts (nn.Linear weight = [out, in]
 * row-major): fusion.linear_v / linear_q (fusion.py:62-66), the R list_linear_hv / list_linear_hq stacked along
 * the output dimension ([R*dim_mm, dim_hv] / [R*dim_mm, dim_hq]; fusion.py:70-76), linear_classif (noatt.py:19). */
typedef struct {
  const float *wv, *bv, *wq, *bq, *whv, *bhv, *whq, *bhq, *wc, *bc;
  const void* whv_f16;   /* optional: [R*dim_mm, k16] IEEE fp16, k16 = roundup8(dim_hv + 1): columns [0, dim_hv) = whv,
                          * column dim_hv = bhv (the bias rides as one more k column against a constant-one input),
                          * zeros after. With it the fused MUTAN kernel runs kind::f16 (x_v = tanh(.) and the frozen
                          * weights keep the 10 mantissa bits TF32 keeps, at half the bytes and twice the tensor
                          * rate); NULL = TF32 from the fp32 weights.                                               */
} VqacxNcxFrozen;

/* Placement of the trainable tensors of NeuralModel (cx.py:239-259) inside ONE flat float32 buffer (the same
 * layout is used for gradients and both Adam moments, so the optimiser and the data-parallel all-reduce are one
 * pass each). linear_1.weight is stored [dim_h, ld_w1] with its 14 089 reference columns regrouped into 16-byte
 * aligned blocks; col_* are the first packed column of each reference column group (cx.py:309-320 order:
 * v_orig, v_other, v_mult, v_dist, v_rank, q_emb, z_orig, z_other, a_emb_gt, a_emb_other). Padding is zero. */
typedef struct {
  int64_t total;
  int64_t off_emb, off_w1, off_b1, off_w2, off_b2, off_w3, off_b3, off_wout, off_bout;
  int64_t ld_w1, ld_h;
  int64_t col_vother, col_vmult, col_zother, col_dist;        /* candidate block  [0, kc)           */
  int64_t col_vorig, col_q, col_zorig, col_agt;               /* per-question block [kc, kc + ks)   */
  int64_t col_rank, col_aother;
  int64_t kc, ks;
} VqacxNcxLayout;

/* One batch of questions. `features` is the (device-resident) [n_rows, dim_v] matrix; img_idx[b, 0] is the row of
 * the original image and img_idx[b, 1..knn] its neighbours (what getDataFromBatch gathers on the host,
 * counterexamples.py:527-541). q_emb is the question encoder output (cx.py:71), the boundary of this path. */
typedef struct {
  int64_t batch;
  const float* features;
  int64_t ld_features;
  const int32_t* img_idx;
  const float* q_emb;
  int64_t ld_q;
  const int64_t* answer_aid;
  float drop_p;               /* nn.Dropout(p) of cx.py:259; used only when training != 0                       */
  uint64_t drop_seed;         /* counter-based mask: keep(seed + layer, row, col)                               */
  const float* drop_mask;     /* optional explicit masks [n_layers][batch*knn][dim_h] of 0 or 1/(1-p) (tests)   */
  const uint64_t* drop_seed_dev; /* optional DEVICE counter added to drop_seed when the kernels run: lets a captured
                                  * CUDA graph of the step draw a fresh mask on every replay (vqacx_step_tick)    */
  uint32_t lesion;            /* VQACX_LESION_* bits = model_spec flags that are false (cx.py:265-307); 0 = none  */
  const float* v_rank;        /* VQACX_LESION_V_RANK: [batch*knn, 32] rows standing in for the one-hot rank vector
                               * (cx.py:306 draws torch.rand([B, 24]) per neighbour); columns >= knn must be zero */
  const void* hv_table;       /* optional: per-image table of the frozen head's visual rank projections for the rows of
                               * `features` (vqacx_ncx_fast_build_tables). With it vqacx_ncx_forward / _backward skip the
                               * per-candidate linear_v and rank products (they are functions of the image alone) and
                               * only multiply table rows by the questions' hq. NULL = evaluate them per candidate.   */
} VqacxNcxBatch;

/* Lesioned feature groups (options/cx/neuralcx_lesion_*.yaml). v_mult / v_dist are replaced by zeros (cx.py:297,301),
 * v_rank by the caller's random rows. The v_emb lesion replaces the image features themselves (cx.py:265-266) and is
 * applied by the caller to `features`; q_emb / z_emb alone change nothing in the reference (cx.py:270-277 still runs
 * vqa_forward when either is true). */
#define VQACX_LESION_V_MULT 1u
#define VQACX_LESION_V_DIST 2u
#define VQACX_LESION_V_RANK 4u

int vqacx_ncx_layout(const VqacxNcxDims* dims, VqacxNcxLayout* out);
size_t vqacx_ncx_workspace_bytes(const VqacxNcxDims* dims, int64_t batch, int training);

/* scores[batch, knn] = NeuralModel.forward (cx.py:261-333, all model_spec flags true) including the frozen
 * vqa_forward (cx.py:64-104). training != 0 applies dropout and keeps the activations backward needs in `ws`.
 * Asynchronous; `ws` must stay untouched until the matching backward has run. */
int vqacx_ncx_forward(const VqacxNcxDims* dims, const VqacxNcxFrozen* frozen, const float* params,
                      const VqacxNcxBatch* batch, void* ws, size_t ws_bytes, int training, float* scores,
                      void* stream);

/* fp16 fast path of the EVALUATION forward (eval_model, counterexamples.py:450-490): same scores as vqacx_ncx_forward with
 * training = 0 within the fp16 / TF32 tolerance (both keep 10 mantissa bits per operand; fp32 accumulation), for a batch
 * whose images are rows of a RESIDENT feature matrix (batch->features + batch->img_idx).
 *   tables  per-IMAGE activations of the frozen head's visual side, hv_r = W_hv^r tanh(W_v v + b_v) + b_hv^r for every row
 *           of the feature matrix ([n_img, R * dim_mm] fp16): a function of the image alone (fusion.py:86-101), evaluated
 *           once per feature matrix instead of once per (question, candidate); behind it an fp16 copy of the feature
 *           matrix itself, from which a batch gathers its neighbour rows. Rebuild when the features or the frozen head
 *           change.
 *   twins   fp16 copies of the weights the big products read (linear_1's candidate columns, M^T = W1a . E^T, the
 *           classifier, the hidden layers). Rebuild whenever `params` (or the frozen head) change.
 * The *_bytes functions size the caller-owned buffers; the workspace of vqacx_ncx_fast_workspace_bytes(dims, batch) also
 * serves the two builders (any batch >= 0). n_ans <= 2048. The random-rank lesion is not supported here (VQACX_EINVAL):
 * use vqacx_ncx_forward. All asynchronous. */
size_t vqacx_ncx_fast_tables_bytes(const VqacxNcxDims* dims, int64_t n_img);
size_t vqacx_ncx_fast_twins_bytes(const VqacxNcxDims* dims);
size_t vqacx_ncx_fast_workspace_bytes(const VqacxNcxDims* dims, int64_t batch);
int vqacx_ncx_fast_build_tables(const VqacxNcxDims* dims, const VqacxNcxFrozen* frozen, const float* features,
                                int64_t ld_features, int64_t n_img, void* tables, size_t tables_bytes, void* ws,
                                size_t ws_bytes, void* stream);
int vqacx_ncx_fast_build_twins(const VqacxNcxDims* dims, const VqacxNcxFrozen* frozen, const float* params, void* twins,
                               size_t twins_bytes, void* ws, size_t ws_bytes, void* stream);
/* p1 (optional, may be NULL): [n_img, ld_h] fp32 table P1[img] = features[img] . W1[:, v_other]^T
 * (vqacx_ncx_fast_build_p1; rebuild when `params` or the features change). With it, linear_1's v_other columns are a
 * table lookup by the candidate's image instead of 2048 k-columns of the per-candidate product. */
size_t vqacx_ncx_fast_p1_bytes(const VqacxNcxDims* dims, int64_t n_img);
int vqacx_ncx_fast_build_p1(const VqacxNcxDims* dims, const float* params, const float* features, int64_t ld_features,
                            int64_t n_img, void* p1, size_t p1_bytes, void* ws, size_t ws_bytes, void* stream);
int vqacx_ncx_fast_forward(const VqacxNcxDims* dims, const VqacxNcxFrozen* frozen, const float* params,
                           const VqacxNcxBatch* batch, const void* tables, int64_t n_img, const void* twins,
                           const void* p1, void* ws, size_t ws_bytes, float* scores, void* stream);

/* c [m, n] (fp32, nullable) and / or c_half [m, n] (fp16, nullable) = act(a . b^T + bias) with IEEE fp16 operands
 * a [m, k], b [n, k] (row-major, lda / ldb in halfs and multiples of 8) as kind::f16 tensor-core products with fp32
 * accumulation: the product behind the fast path above. act: 0 none, 1 tanh, 2 relu. Asynchronous. */
int vqacx_gemm_f16(int m, int n, int k, const void* a, int64_t lda, const void* b, int64_t ldb, float* c, int64_t ldc,
                   void* c_half, int64_t ld_half, const float* bias, int act, void* stream);

/* Fused 24-way softmax cross-entropy (sum over the batch, counterexamples.py:310), its gradient and Recall@1/5
 * (counterexamples.py:501-506). loss_sum[0] and hits[0..1] are ACCUMULATED (caller zeroes them);
 * dscores (nullable) = grad_scale * dLoss/dscores, grad_scale = 1/len(batch) reproduces counterexamples.py:334. */
int vqacx_ce_recall(const float* scores, const int64_t* comp_idx, int64_t batch, int knn, float* loss_sum,
                    float* dscores, float grad_scale, int32_t* hits, void* stream);

/* Gradients of every trainable tensor into `grads` (flat, VqacxNcxLayout), overwriting it. Replaces loss.backward()
 * (counterexamples.py:337) for the trainable_vqa = false case (the frozen head gets no gradient, cx.py:76-80). */
int vqacx_ncx_backward(const VqacxNcxDims* dims, const float* params, const VqacxNcxBatch* batch, void* ws,
                       size_t ws_bytes, const float* dscores, float* grads, void* stream);

/* ---- trainable_vqa = true (cx.py:73-74, options/cx/neuralcx_trainable_vqa.yaml) --------------------------------------
 * The MUTAN head runs in TRAIN mode (the reference leaves vqa_model in train() when it is trainable): dropout_v on the
 * image features and dropout_q on the question vector of EVERY (question, image) row (cx.py:83-84 duplicates q_emb
 * before fusion.py:86,94 drop it), classifier dropout (noatt.py:27), and its weights receive gradients through
 * z_orig, z_knns, softmax(a_knns) and q_emb (cx.py:309-320). The head's ten tensors live in ONE flat float32 buffer
 * (VqacxNcxHeadLayout; same layout for gradients and Adam moments), each in the reference's own nn.Linear layout, the R
 * list_linear_hv / list_linear_hq stacked along the output dimension. Rows of the masks and of the activations are in
 * the library's order: [batch*knn candidate rows (b*knn + j) | batch original-image rows].
 * The question encoder is a separate pair of calls: q_emb is an input here (vqacx_gru_forward_train produces it), dq_emb
 * the gradient handed back for vqacx_gru_backward. */
typedef struct {
  int64_t total;
  int64_t off_wv, off_bv, off_wq, off_bq, off_whv, off_bhv, off_whq, off_bhq, off_wc, off_bc;
} VqacxNcxHeadLayout;

typedef struct {
  float p_v, p_q, p_c;      /* fusion dropout_v, dropout_q; classif dropout                                          */
  uint64_t seed;            /* counter-based masks keep(seed + site, row, col); sites 101 (v), 102 (q), 103 (classif) */
  const float *mask_v, *mask_q, *mask_c; /* optional explicit masks of 0 or 1/(1-p) (tests): dense [batch*(knn+1), dim_v],
                                          * [batch*(knn+1), dim_q], [batch*knn, dim_mm]                                */
} VqacxNcxHeadTrain;

int vqacx_ncx_head_layout(const VqacxNcxDims* dims, VqacxNcxHeadLayout* out);
size_t vqacx_ncx_tv_workspace_bytes(const VqacxNcxDims* dims, int64_t batch);
/* Forward of a training step with the trainable head: scores as vqacx_ncx_forward(training = 1), activations kept in
 * `ws` (vqacx_ncx_workspace_bytes(.., training = 1)) and `tv_ws`. batch->hv_table must be NULL. Asynchronous. */
int vqacx_ncx_forward_tv(const VqacxNcxDims* dims, const float* head, const float* params, const VqacxNcxBatch* batch,
                         const VqacxNcxHeadTrain* ht, void* ws, size_t ws_bytes, void* tv_ws, size_t tv_ws_bytes,
                         float* scores, void* stream);
/* Backward: `grads` as vqacx_ncx_backward, `head_grads` (flat, VqacxNcxHeadLayout, overwritten) and dq_emb
 * [batch, ld_dq] (nullable) = dLoss/dq_emb for the question encoder. Replaces loss.backward() with trainable_vqa = true
 * (counterexamples.py:337). */
int vqacx_ncx_backward_tv(const VqacxNcxDims* dims, const float* head, const float* params, const VqacxNcxBatch* batch,
                          const VqacxNcxHeadTrain* ht, void* ws, size_t ws_bytes, void* tv_ws, size_t tv_ws_bytes,
                          const float* dscores, float* grads, float* head_grads, float* dq_emb, int64_t ld_dq,
                          void* stream);

/* Test / inspection hook: copies the post-activation hidden layer `layer` (1-based; relu and dropout applied,
 * cx.py:322-329) of the last forward that used `ws` into out [batch*knn, dim_h] (dense). Lets a checker evaluate the
 * reference's backward with the same relu decisions the TF32 forward took. */
int vqacx_ncx_hidden(const VqacxNcxDims* dims, int64_t batch, int training, const void* ws, size_t ws_bytes, int layer,
                     float* out, void* stream);

/* torch.optim.Adam step (counterexamples.py:275-276,339) over a flat buffer; step counts from 1. */
int vqacx_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                    float beta1, float beta2, float eps, int32_t step, float grad_scale, void* stream);

/* Adam with the step count read from DEVICE memory (step_dev[0] >= 1) so that the launch can live inside a captured
 * CUDA graph; same arithmetic as vqacx_adam_step. */
int vqacx_adam_step_dev(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                        float beta1, float beta2, float eps, const int32_t* step_dev, float grad_scale, void* stream);

/* End-of-step bookkeeping on the device: step_dev[0] += 1 (nullable), drop_seed_dev[0] += seed_inc (nullable). */
int vqacx_step_tick(int32_t* step_dev, uint64_t* drop_seed_dev, uint64_t seed_inc, void* stream);

/* Example builder join between stage 1 and stage 2 (vqacx_valset_builder.ipynb cell 13, vqacx_trainset_builder.ipynb
 * cell 18): knn_idx [n_img, k1] = stage-1 output (each image among its own k1 = 25 neighbours); q_img [n_q] = image row
 * of every question; q_comp [n_q] = index of its complementary question, -1 if it has none in the set. Out:
 * img_idx_out [n_q, k1] = [image, its k1-1 neighbours] (what getDataFromBatch gathers, counterexamples.py:527-541) and
 * comp_idx_out [n_q] = knns[i1].index(i2) (`comp.knn_index`), -1 where the notebook drops the example. */
int vqacx_cx_build_examples(const int64_t* knn_idx, int64_t n_img, int k1, const int32_t* q_img, const int32_t* q_comp,
                            int64_t n_q, int32_t* img_idx_out, int32_t* comp_idx_out, void* stream);

/* ------------------------------------------------------------------------------------------------
 * The other CX heads that share CXModelBase.vqa_forward (vqa/models/cx.py:64-104); same dims / frozen / batch structs
 * and workspace as vqacx_ncx_forward (dims->dim_h, n_layers only size the workspace; answer_aid is read by the heads).
 * ---------------------------------------------------------------------------------------------- */

/* vqa_forward on the gathered rows: z_knns [batch*knn, dim_mm], z_orig [batch, dim_mm], p_knns [batch*knn, n_ans] =
 * softmax of the neighbours' answer logits (each nullable). LinearContext (cx.py:139-156) feeds z_knns to its Linear. */
int vqacx_cx_vqa_forward(const VqacxNcxDims* dims, const VqacxNcxFrozen* frozen, const VqacxNcxBatch* batch, void* ws,
                         size_t ws_bytes, float* z_knns, float* z_orig, float* p_knns, void* stream);

#define VQACX_HEAD_BLACKBOX 0 /* BlackBox.forward, cx.py:120-136: scores = -softmax(a_knns)[answer] */
#define VQACX_HEAD_SEMANTIC 1 /* SemanticBaseline.forward, cx.py:182-205 (emb_pairs [n_ans, n_ans], lam) */
/* scores [batch, knn] of one of the parameter-free heads. */
int vqacx_cx_head_scores(const VqacxNcxDims* dims, const VqacxNcxFrozen* frozen, const VqacxNcxBatch* batch, void* ws,
                         size_t ws_bytes, int head, const float* emb_pairs, float lam, float* scores, void* stream);

/* LinearContext.forward (cx.py:149-156): scores [batch, knn] = z_knns.view(batch, knn*dim_mm) . weight^T + bias with
 * weight [knn, knn*dim_mm] (nn.Linear layout); z_flat [batch, knn*dim_mm] receives the flattened z_knns, which the
 * backward needs. Backward of the Linear only (the VQA head is frozen): g_weight = dscores^T . z_flat, g_bias =
 * column sums of dscores [batch, knn] (knn % 4 == 0). */
int vqacx_cx_linear_context_forward(const VqacxNcxDims* dims, const VqacxNcxFrozen* frozen, const VqacxNcxBatch* batch,
                                    void* ws, size_t ws_bytes, const float* weight, const float* bias, float* z_flat,
                                    float* scores, void* stream);
int vqacx_cx_linear_context_backward(const VqacxNcxDims* dims, int64_t batch, void* ws, size_t ws_bytes,
                                     const float* z_flat, const float* dscores, float* g_weight, float* g_bias,
                                     void* stream);

/* out [n, n] = cosine similarity between the rows of x [n, d] (SemanticBaseline.set_answer_embedding, cx.py:172-175:
 * sklearn.metrics.pairwise.cosine_similarity); float64 accumulation, zero rows give 0. */
int vqacx_cosine_similarity(const float* x, int n, int d, float* out, void* stream);

/* Question encoder seq2vec = BayesianUniSkip in eval mode (skipthoughts.py:265-300; BayesianGRU gru.py:46-84,132-155):
 * q_emb[b] = GRU state after the last word of question b. emb [n_emb_rows, E] = embedding.weight (row 0 = padding),
 * wids [batch, T] right-padded with 0, w_i [3H, E] = [weight_ir; weight_ii; weight_in], b_i [3H] their biases,
 * w_h [3H, H] = [weight_hr; weight_hi; weight_hn] (no bias, gru.py:10). max_steps = max(lengths) if the caller
 * knows it (0 = T: finished rows simply stop updating). out [batch, H]. */
size_t vqacx_gru_workspace_bytes(int64_t batch, int T, int E, int H);
int vqacx_gru_encode(const float* emb, int64_t n_emb_rows, int E, const int64_t* wids, int64_t batch, int T,
                     const float* w_i, const float* b_i, const float* w_h, int H, int max_steps, float* out, void* ws,
                     size_t ws_bytes, void* stream);

/* The same encoder in TRAIN mode with its backward (back-propagation through time), for trainable_vqa = true (cx.py:73-74:
 * the reference back-propagates the NeuralCX loss through q_emb into seq2vec): BayesianGRUCell.forward with its six
 * SequentialDropout masks (gru.py:71-84; one Bernoulli(1-p)/(1-p) draw per gate input and per SEQUENCE, the same at
 * every step, dropout.py:44-66). masks_x [3][batch][E], masks_h [3][batch][H] (gate order r, i, n; values 0 or 1/(1-p))
 * are optional explicit masks (tests); NULL = counter-based keep(seed + site, b, c). w_i / b_i / w_h as for
 * vqacx_gru_encode. The forward keeps every state, gate and masked input in `ws` (vqacx_gru_train_workspace_bytes) for the
 * backward, which overwrites d_emb [n_emb_rows, E] (row 0, the padding index, stays zero), d_w_i [3H, E], d_b_i [3H],
 * d_w_h [3H, H] given dq [batch, ld_dq] = dLoss/dq_emb (what vqacx_ncx_backward_tv returns). Asynchronous. */
size_t vqacx_gru_train_workspace_bytes(int64_t batch, int T, int E, int H);
int vqacx_gru_forward_train(const float* emb, int64_t n_emb_rows, int E, const int64_t* wids, int64_t batch, int T,
                            const float* w_i, const float* b_i, const float* w_h, int H, float p_drop, uint64_t seed,
                            const float* masks_x, const float* masks_h, float* out, void* ws, size_t ws_bytes,
                            void* stream);
int vqacx_gru_backward(int64_t n_emb_rows, int E, const int64_t* wids, int64_t batch, int T, const float* w_i,
                       const float* w_h, int H, const float* dq, int64_t ld_dq, void* ws, size_t ws_bytes, float* d_emb,
                       float* d_w_i, float* d_b_i, float* d_w_h, void* stream);

/* Pooled-feature producer, the step before stage 1 (extract.py:121-124): att [n, C, HW] fp32 activation maps ->
 * out_f32 [n, C] = mean over the HW regions (`output_att.sum(3).sum(2).div(nb_regions)`). If out_lp is non-NULL the
 * kernel also emits everything vqacx_knn_prep would compute from out_f32 (fp16 copy [n, C], norm_sq [n], max norm
 * bits (caller zeroes), non-representable count (caller zeroes)) in the same pass over the maps. */
int vqacx_pool_features(const float* att, int64_t n, int C, int HW, float* out_f32, void* out_lp, float* norm_sq,
                        uint32_t* max_norm_sq_bits, int32_t* bad_count, void* stream);

/* Adds n to the launch counter: a replayed CUDA graph launches its kernels without passing through the library. */
void vqacx_launch_count_add(int64_t n);

#ifdef __cplusplus
}
#endif
#endif /* VQACX_H_ */

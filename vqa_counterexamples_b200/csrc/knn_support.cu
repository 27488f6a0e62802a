// kNN support kernels: low-precision prep, exact fp32 distance path (first-correct path and the
// fallback for rows whose error-bound check fails), candidate finalisation, fp64 re-rank with the
// exactness check, and the cross-shard top-k merge.
//
// Reference behaviour being reproduced (knn.py:41-56 -> sklearn NearestNeighbors(n_neighbors=k),
// algorithm 'brute', metric euclidean): for every query row the k smallest Euclidean distances to
// the fitted rows, ascending, as (distances f64, indices i64). sklearn evaluates
// |x|^2 - 2 x.y + |y|^2 in float64; we evaluate sum((x-y)^2) in float64 for the surviving
// candidates, which is at least as accurate, and order ties by (distance, index).
#include "knn_internal.cuh"
#include "topk.cuh"
#include <cuda_fp16.h>

namespace vqacx {

// ---------------------------------------------------------------------------------------------
// prep: fp32 rows -> fp16 rows (round-to-nearest) + squared norm of the ROUNDED row (so that
// |y~|^2 - 2 x~.y~ + |x~|^2 is exactly |x~ - y~|^2 up to accumulation error) + global max norm.
// One warp per row; HBM-bound: reads 4*d, writes 2*d + 4 bytes per row.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) knn_prep_kernel(const float* __restrict__ x, long long n, int d,
                                                       __half* __restrict__ x_lp, float* __restrict__ norm_sq,
                                                       unsigned int* __restrict__ max_norm_sq_bits,
                                                       int* __restrict__ bad_count) {
  const int lane = threadIdx.x & 31;
  const long long warp = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  for (long long row = warp; row < n; row += nwarps) {
    const float4* src = reinterpret_cast<const float4*>(x + row * d);
    uint2* dst = reinterpret_cast<uint2*>(x_lp + row * d);
    double acc = 0.0;
    int bad = 0;
    for (int i = lane; i < d / 4; i += 32) {
      float4 v = __ldcs(src + i);
      __half2 h01 = __floats2half2_rn(v.x, v.y);
      __half2 h23 = __floats2half2_rn(v.z, v.w);
      float2 f01 = __half22float2(h01), f23 = __half22float2(h23);
      bad |= !(isfinite(f01.x) && isfinite(f01.y) && isfinite(f23.x) && isfinite(f23.y));
      acc += static_cast<double>(f01.x) * f01.x + static_cast<double>(f01.y) * f01.y +
             static_cast<double>(f23.x) * f23.x + static_cast<double>(f23.y) * f23.y;
      uint2 out;
      out.x = *reinterpret_cast<uint32_t*>(&h01);
      out.y = *reinterpret_cast<uint32_t*>(&h23);
      dst[i] = out;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    bad = __any_sync(0xffffffffu, bad);
    if (lane == 0) {
      // round UP so the stored norm never under-states |y~|^2 by more than 1 ulp either way
      const float ns = static_cast<float>(acc);
      norm_sq[row] = ns;
      atomicMax(max_norm_sq_bits, __float_as_uint(ns));
      if (bad) atomicAdd(bad_count, 1);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// exact fp32 path: d2 = sum((x-y)^2) with fp32 FMA, 64x64 tile per CTA, fused streaming top-k'.
// grid = (q tiles of 64, db splits). slot = blockIdx.x * gridDim.y + blockIdx.y.
// ---------------------------------------------------------------------------------------------
constexpr int SBM = 64, SBN = 64, SBK = 16, SPITCH = 68;

__global__ void __launch_bounds__(256) knn_exact_kernel(const float* __restrict__ q,
                                                        const int* __restrict__ q_rows, const int* n_q_ptr,
                                                        int n_q_static, const float* __restrict__ db, int n_db,
                                                        int d, int tiles_per_split, uint64_t* __restrict__ cand,
                                                        int* __restrict__ cnt) {
  __shared__ __align__(16) float As[SBK][SPITCH];
  __shared__ __align__(16) float Bs[SBK][SPITCH];
  __shared__ float Ds[SBM][SBN + 1];

  const int n_q = n_q_ptr ? min(*n_q_ptr, n_q_static) : n_q_static;
  const int q0 = blockIdx.x * SBM;
  if (q0 >= n_q) return;
  const int slot = blockIdx.x * gridDim.y + blockIdx.y;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int lm = tid >> 2, lk = (tid & 3) * 4;  // loader coordinates

  const int qi = q0 + lm;
  long long a_row = -1;
  if (qi < n_q) a_row = q_rows ? q_rows[qi] : qi;

  RowFilter st;
  if (tid < SBM) filter_reset(st, cand + (static_cast<size_t>(slot) * SBM + tid) * kTopkCap);

  const int n_tiles = (n_db + SBN - 1) / SBN;
  const int t_begin = blockIdx.y * tiles_per_split;
  const int t_end = min(n_tiles, t_begin + tiles_per_split);

  for (int t = t_begin; t < t_end; ++t) {
    const int col0 = t * SBN;
    const long long b_row = (col0 + lm < n_db) ? (col0 + lm) : -1;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    float4 ra = make_float4(0.f, 0.f, 0.f, 0.f), rb = ra;
    if (a_row >= 0 && lk < d) ra = *reinterpret_cast<const float4*>(q + a_row * d + lk);
    if (b_row >= 0 && lk < d) rb = *reinterpret_cast<const float4*>(db + b_row * d + lk);

    for (int k0 = 0; k0 < d; k0 += SBK) {
      __syncthreads();
      As[lk + 0][lm] = ra.x; As[lk + 1][lm] = ra.y; As[lk + 2][lm] = ra.z; As[lk + 3][lm] = ra.w;
      Bs[lk + 0][lm] = rb.x; Bs[lk + 1][lm] = rb.y; Bs[lk + 2][lm] = rb.z; Bs[lk + 3][lm] = rb.w;
      __syncthreads();
      const int kn = k0 + SBK + lk;
      ra = make_float4(0.f, 0.f, 0.f, 0.f); rb = ra;
      if (kn < d) {
        if (a_row >= 0) ra = *reinterpret_cast<const float4*>(q + a_row * d + kn);
        if (b_row >= 0) rb = *reinterpret_cast<const float4*>(db + b_row * d + kn);
      }
#pragma unroll
      for (int k = 0; k < SBK; ++k) {
        const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
        const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
        const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float df = av[i] - bv[j];
            acc[i][j] = fmaf(df, df, acc[i][j]);
          }
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) Ds[ty * 4 + i][tx * 4 + j] = acc[i][j];
    __syncthreads();
    if (tid < SBM) {  // warps 0 and 1: one thread per query row
#pragma unroll 1
      for (int half = 0; half < 2; ++half) {
        for (int c = half * 32; c < half * 32 + 32; ++c) {
          const int col = col0 + c;
          if (col < n_db) filter_push(st, Ds[tid][c], static_cast<uint32_t>(col));
        }
        filter_compact_warp(st, 32);
      }
    }
    // Ds is rewritten only after the next tile's k-loop, which contains block-wide barriers.
  }
  if (tid < SBM) cnt[static_cast<size_t>(slot) * SBM + tid] = st.cnt;
}

// ---------------------------------------------------------------------------------------------
// finalize: merge the per-slot candidate buffers of one query row into its KP smallest keys.
// One warp per row. tile_slot0 / tile_nslots may be null -> uniform `nslots_uniform` slots per tile.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) knn_finalize_kernel(const uint64_t* __restrict__ cand,
                                                           const int* __restrict__ cnt,
                                                           const int* __restrict__ tile_slot0,
                                                           const int* __restrict__ tile_nslots,
                                                           int nslots_uniform, int rows_per_tile,
                                                           const int* n_q_ptr, int n_q_static,
                                                           float* __restrict__ cand_val,
                                                           int* __restrict__ cand_idx, int row0) {
  const int n_q = n_q_ptr ? min(*n_q_ptr, n_q_static) : n_q_static;
  const int lane = threadIdx.x & 31;
  const int row = row0 + ((blockIdx.x * blockDim.x + threadIdx.x) >> 5);   // rows [row0, n_q) of this launch
  if (row >= n_q) return;
  const int t = row / rows_per_tile, r = row % rows_per_tile;
  const int s0 = tile_slot0 ? tile_slot0[t] : t * nslots_uniform;
  const int ns = tile_nslots ? tile_nslots[t] : nslots_uniform;
  uint64_t k[4] = {kKeyPad, kKeyPad, kKeyPad, kKeyPad};
  for (int s = 0; s < ns; ++s) {
    const size_t base = static_cast<size_t>(s0 + s) * rows_per_tile + r;
    const int c = cnt[base];
    const uint64_t* b = cand + base * kTopkCap;
    for (int off = 0; off < c; off += 64) {
      k[2] = (off + lane < c) ? __ldcg(reinterpret_cast<const unsigned long long*>(b + off + lane)) : kKeyPad;
      k[3] = (off + 32 + lane < c) ? __ldcg(reinterpret_cast<const unsigned long long*>(b + off + 32 + lane))
                                   : kKeyPad;
      warp_sort128(k);
    }
  }
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const bool pad = (k[j] == kKeyPad);
    cand_val[static_cast<size_t>(row) * kTopkKeep + j * 32 + lane] = pad ? __int_as_float(0x7f800000) : key_value(k[j]);
    cand_idx[static_cast<size_t>(row) * kTopkKeep + j * 32 + lane] = pad ? -1 : static_cast<int>(key_col(k[j]));
  }
}

// Symmetric self-join finalize: a row's candidates are spread over several row-side slots plus its mirror buffer
// (~10x the keys of the plain search), almost all of them above the final 64-th key. Keys are therefore pre-filtered
// against the running 64-th best and only the survivors are staged (shared memory, 128 keys per warp) and merged 64
// at a time, so the number of 128-key sorts stays close to (#keys that ever beat the running bound) / 64.
__global__ void __launch_bounds__(256) knn_finalize_sym_kernel(const uint64_t* __restrict__ cand,
                                                               const int* __restrict__ cnt,
                                                               const int* __restrict__ tile_slot0,
                                                               const int* __restrict__ tile_nslots, int rows_per_tile,
                                                               int n_q, const KnnSymArgs sym,
                                                               float* __restrict__ cand_val, int* __restrict__ cand_idx,
                                                               int* __restrict__ flag_count, int* __restrict__ flag_rows,
                                                               int* __restrict__ stats) {
  __shared__ uint64_t s_stage[8][128];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= n_q) return;
  uint64_t* stage = s_stage[w];
  const int t = row / rows_per_tile, r = row % rows_per_tile;
  const int s0 = tile_slot0[t], ns = tile_nslots[t];
  uint64_t k[4] = {kKeyPad, kKeyPad, kKeyPad, kKeyPad};
  uint64_t thr = kKeyPad;   // the 64-th best key so far
  int staged = 0;
  const unsigned lt_mask = (1u << lane) - 1u;
  const int subcap = sym.mcap;
  int mc_total = 0;
  bool overflow = false;
  for (int s = 0; s <= ns; ++s) {
    const uint64_t* b;
    int c;
    if (s < ns) {
      const size_t base = static_cast<size_t>(s0 + s) * rows_per_tile + r;
      c = cnt[base];
      b = cand + base * kTopkCap;
    } else {
      const int raw = sym.mcnt[static_cast<size_t>(row) * kMirrorCntStride];
      mc_total += raw;
      overflow |= raw > subcap;
      c = min(raw, subcap);
      b = sym.mbuf + static_cast<size_t>(row) * sym.mcap;
    }
    for (int off = 0; off < c; off += 32) {
      const uint64_t key = (off + lane < c) ? __ldcg(reinterpret_cast<const unsigned long long*>(b + off + lane)) : kKeyPad;
      const bool pred = key < thr;
      const unsigned bal = __ballot_sync(0xffffffffu, pred);
      if (pred) stage[staged + __popc(bal & lt_mask)] = key;
      staged += __popc(bal);
      if (staged >= 64) {
        __syncwarp();
        k[2] = stage[lane];
        k[3] = stage[32 + lane];
        const int rem = staged - 64;
        const uint64_t carry = lane < rem ? stage[64 + lane] : kKeyPad;
        __syncwarp();
        if (lane < rem) stage[lane] = carry;
        staged = rem;
        warp_sort128(k);
        thr = shfl_u64(k[1], 31);
      }
    }
  }
  if (stats && lane == 0 && (row & 15) == 0) atomicAdd(&stats[kStatMirrorKeys], mc_total);   // sampled: 1 query in 16
  if (staged > 0) {
    __syncwarp();
    k[2] = lane < staged ? stage[lane] : kKeyPad;
    k[3] = 32 + lane < staged ? stage[32 + lane] : kKeyPad;
    warp_sort128(k);
  }
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const bool pad = (k[j] == kKeyPad);
    cand_val[static_cast<size_t>(row) * kTopkKeep + j * 32 + lane] = pad ? __int_as_float(0x7f800000) : key_value(k[j]);
    cand_idx[static_cast<size_t>(row) * kTopkKeep + j * 32 + lane] = pad ? -1 : static_cast<int>(key_col(k[j]));
  }
  if (lane == 0 && overflow) {   // mirror overflow: keys were dropped -> exact fallback for this row
    const int pos = atomicAdd(flag_count, 1);
    flag_rows[pos] = row;
  }
}

// ---------------------------------------------------------------------------------------------
// re-rank: exact float64 distances for the KP candidates of each query, sort by (d2, idx),
// write the k best, and prove (or flag) that no non-candidate can belong to the top k.
// One warp per query. HBM/L2 gather-bound: KP * 4*d bytes per query.
// ---------------------------------------------------------------------------------------------
// float -> double widening on the integer pipe. cvt.f64.f32 issues at 16 per clock per SM, and the re-rank needs 1.5 of
// them per candidate element, which made conversions the kernel's bound; this is five integer operations: re-bias
// the exponent (+896) and shift the 23-bit mantissa to the top of the 52-bit field — exact for every normal float.
// Zeros (~45 % of post-ReLU features) and subnormals (|x| < 2^-126) map to 0.0 by a select: a subnormal coordinate
// moves a squared distance by < 2^-250, far below one ulp of any float64 sum this kernel can produce (no branch and no
// cvt must remain: the compiler evaluates a rare-path cvt speculatively for every element).
__device__ __forceinline__ double widen_exact(float x) {
  const uint32_t b = __float_as_uint(x);
  const uint32_t hi = (b & 0x80000000u) | (((b & 0x7fffffffu) >> 3) + 0x38000000u);
  const uint32_t lo = b << 29;
  const bool tiny = (b & 0x7f800000u) == 0u;
  // (the low word needs no select: it is zero for a zero input, and for a subnormal input it only adds < 2^-1045)
  return __hiloint2double(static_cast<int>(tiny ? 0u : hi), static_cast<int>(lo));
}
// The re-rank's inner loop was bound by the integer pipe (12 ALU operations per candidate element when every value is
// widened by widen_exact). It now splits the widening between the two pipes BY COLUMN: components .x/.z of every
// float4 go through cvt.f64.f32 (16 per clock per SM, its own pipe), .y/.w through the integer sequence. Both
// operands of a difference always take the same route, so equal inputs still subtract to exactly zero (the routes
// differ only for subnormal inputs, by < 2^-126).
struct Wide4 {
  double x, y, z, w;
};
__device__ __forceinline__ Wide4 widen4(const float4 v) {
  return {static_cast<double>(v.x), widen_exact(v.y), static_cast<double>(v.z), widen_exact(v.w)};
}
__device__ __forceinline__ double sqdiff_acc(const Wide4& a, const float4 v, double acc) {
  const Wide4 b = widen4(v);
  double t;
  t = a.x - b.x; acc = fma(t, t, acc);
  t = a.y - b.y; acc = fma(t, t, acc);
  t = a.z - b.z; acc = fma(t, t, acc);
  t = a.w - b.w; acc = fma(t, t, acc);
  return acc;
}

struct PairLess {
  __device__ __forceinline__ bool operator()(double a, int ia, double b, int ib) const {
    return (a < b) || (a == b && ia < ib);
  }
};

__global__ void __launch_bounds__(256, 3) knn_rerank_kernel(
    const float* __restrict__ q, const int* __restrict__ q_rows, const int* n_q_ptr, int n_q_static,
    const float* __restrict__ db, int n_db, int d, const float* __restrict__ cand_val,
    const int* __restrict__ cand_idx, const float* __restrict__ q_norm_sq_lp,
    const unsigned int* __restrict__ db_max_norm_sq_bits, float eps_d, float d_abs, float eps_sq_abs, float eps_sq_rel, int k,
    long long db_row0, double* __restrict__ out_dist, long long* __restrict__ out_idx, int* __restrict__ flag_count,
    int* __restrict__ flag_rows, int* __restrict__ stats, const float* __restrict__ ext_bound, int row0,
    int cand_by_qrow, const uint32_t* __restrict__ excl_thr, int out_ld) {
  __shared__ double s_d2[8][kTopkKeep];
  __shared__ int s_ix[8][kTopkKeep];
  const int n_q = n_q_ptr ? min(*n_q_ptr, n_q_static) : n_q_static;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int lq = row0 + blockIdx.x * 8 + w;   // rows [row0, n_q) of this launch
  if (lq >= n_q) return;
  const long long qrow = q_rows ? q_rows[lq] : lq;
  const float4* xr = reinterpret_cast<const float4*>(q + qrow * d);
  const int nv = d / 4;

  double xn2 = 0.0;
  for (int i = lane; i < nv; i += 32) {
    const float4 v = xr[i];
    xn2 += static_cast<double>(v.x) * v.x + static_cast<double>(v.y) * v.y + static_cast<double>(v.z) * v.z +
           static_cast<double>(v.w) * v.w;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) xn2 += __shfl_xor_sync(0xffffffffu, xn2, o);

  // candidate lists are stored per list position (fallback rows) or per query row (locality-ordered full search)
  const size_t crow = cand_by_qrow ? static_cast<size_t>(qrow) : static_cast<size_t>(lq);
  const int* ci = cand_idx + crow * kTopkKeep;
  const float* cv = cand_val + crow * kTopkKeep;

  // --- error model of the approximate values: every candidate's true distance lies in [lo_c, hi_c] ---------
  const float ymax = db_max_norm_sq_bits ? sqrtf(__uint_as_float(*db_max_norm_sq_bits)) : 0.f;
  const double xn = sqrt(xn2);
  const double xlp2 = q_norm_sq_lp ? static_cast<double>(q_norm_sq_lp[qrow]) : 0.0;
  const double abs_term = static_cast<double>(eps_sq_abs) * xn * ymax;
  const double d_term = static_cast<double>(eps_d) * (xn + ymax) + d_abs;
  double lo_c[2], hi_c[2];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const double ap2 = static_cast<double>(cv[j * 32 + lane]) + xlp2;
    lo_c[j] = sqrt(fmax(ap2 * (1.0 - eps_sq_rel) - abs_term, 0.0)) - d_term;
    hi_c[j] = sqrt(fmax(ap2 * (1.0 + eps_sq_rel) + abs_term, 0.0)) + d_term;
  }
  // Pruning: the candidates arrive sorted by approximate value, so the first k of them all have a true distance
  // <= hi_{k-1}, i.e. the true k-th distance is <= hi_{k-1}; a candidate with lo_c > hi_{k-1} cannot be among the k
  // nearest and its 4*d-byte row is never gathered (on the COCO-shaped set this halves the gather, the cost of this
  // kernel). lo is monotone in the approximate value, so the survivors are a prefix.
  int n_valid = __popc(__ballot_sync(0xffffffffu, ci[lane] >= 0)) + __popc(__ballot_sync(0xffffffffu, ci[32 + lane] >= 0));
  int n_eval = n_valid;
  // ext_bound (row-sharded search): an upper bound on the GLOBAL k-th distance, agreed across the shards before this
  // kernel runs; it is usually far below this shard's own k-th distance, so most local candidates are never gathered
  const double hi_ext = ext_bound ? static_cast<double>(ext_bound[qrow]) : __longlong_as_double(0x7ff0000000000000LL);
  double hi_k = hi_ext;
  if (k <= n_valid && k <= kTopkKeep) {
    const int kj = (k - 1) >> 5, kl = (k - 1) & 31;
    const double h0 = __shfl_sync(0xffffffffu, hi_c[0], kl), h1 = __shfl_sync(0xffffffffu, hi_c[1], kl);
    hi_k = fmin(hi_k, kj ? h1 : h0);
  }
  if (hi_k < __longlong_as_double(0x7ff0000000000000LL)) {
    n_eval = __popc(__ballot_sync(0xffffffffu, ci[lane] >= 0 && lo_c[0] <= hi_k)) +
             __popc(__ballot_sync(0xffffffffu, ci[32 + lane] >= 0 && lo_c[1] <= hi_k));
    if (!ext_bound) n_eval = max(n_eval, k);
  }
  // four candidates per iteration: memory-level parallelism, and the query row is widened once per four rows
  const double kInf = __longlong_as_double(0x7ff0000000000000LL);
  for (int c = 0; c < kTopkKeep; c += 4) {
    // Early stop on the EXACT k-th distance: once k candidates have been evaluated, the k-th smallest exact distance so
    // far bounds the true k-th distance from above -- usually well below the a-priori hi_{k-1} -- and every remaining
    // candidate whose lower bound exceeds it is skipped like the ones pruned above (the survivors stay a prefix).
    if (c >= k && c < n_eval && k <= kTopkKeep) {
      __syncwarp();
      const double v0 = lane < c ? s_d2[w][lane] : kInf;
      const double v1 = 32 + lane < c ? s_d2[w][32 + lane] : kInf;
      int r0 = 0, r1 = 0;
      for (int j = 0; j < c; ++j) {      // rank of this lane's two values among the c evaluated ones (ties by slot)
        const double t = s_d2[w][j];
        r0 += (t < v0 || (t == v0 && j < lane)) ? 1 : 0;
        r1 += (t < v1 || (t == v1 && j < 32 + lane)) ? 1 : 0;
      }
      double d2k = (lane < c && r0 == k - 1) ? v0 : ((32 + lane < c && r1 == k - 1) ? v1 : -1.0);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) d2k = fmax(d2k, __shfl_xor_sync(0xffffffffu, d2k, o));
      if (d2k >= 0.0 && d2k < kInf) {
        const double dk_now = sqrt(d2k);
        const int n_new = __popc(__ballot_sync(0xffffffffu, ci[lane] >= 0 && lo_c[0] <= dk_now)) +
                          __popc(__ballot_sync(0xffffffffu, ci[32 + lane] >= 0 && lo_c[1] <= dk_now));
        n_eval = min(n_eval, max(n_new, c));
      }
    }
    int ix[4];
    bool any = false;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      ix[j] = (c + j < n_eval) ? ci[c + j] : -1;
      any |= ix[j] >= 0;
    }
    double a[4] = {0.0, 0.0, 0.0, 0.0};
    if (any) {
      int first = ix[0] >= 0 ? ix[0] : (ix[1] >= 0 ? ix[1] : (ix[2] >= 0 ? ix[2] : ix[3]));
      const float4* y[4];
#pragma unroll
      for (int j = 0; j < 4; ++j)   // a skipped slot re-reads a live row (its result is discarded): no divergent loads
        y[j] = reinterpret_cast<const float4*>(db + static_cast<long long>(ix[j] < 0 ? first : ix[j]) * d);
#pragma unroll 2
      for (int i = lane; i < nv; i += 32) {
        const float4 x4 = xr[i];
        const float4 u0 = __ldg(y[0] + i), u1 = __ldg(y[1] + i), u2 = __ldg(y[2] + i), u3 = __ldg(y[3] + i);
        const Wide4 xw = widen4(x4);
        a[0] = sqdiff_acc(xw, u0, a[0]);
        a[1] = sqdiff_acc(xw, u1, a[1]);
        a[2] = sqdiff_acc(xw, u2, a[2]);
        a[3] = sqdiff_acc(xw, u3, a[3]);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int j = 0; j < 4; ++j) a[j] += __shfl_xor_sync(0xffffffffu, a[j], o);
      }
    }
    if (lane < 4) {
      const int j = lane;
      const int idx = j == 0 ? ix[0] : (j == 1 ? ix[1] : (j == 2 ? ix[2] : ix[3]));
      const double val = j == 0 ? a[0] : (j == 1 ? a[1] : (j == 2 ? a[2] : a[3]));
      s_d2[w][c + j] = idx >= 0 ? val : kInf;
      s_ix[w][c + j] = idx >= 0 ? idx : 0x7fffffff;
    }
  }
  __syncwarp();
  if (stats && lane == 0 && (lq & 15) == 0) atomicAdd(&stats[kStatRerankEvaluated], n_eval);  // sampled: 1 query in 16

  // --- validate the error model on the evaluated candidates, then the exclusion bound -------------------
  int violated = 0;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int c = j * 32 + lane;
    if (c < n_eval && ci[c] >= 0) {
      // the model claims  lo_c <= true_d <= hi_c
      const double td = sqrt(s_d2[w][c]);
      if (td < lo_c[j] * (1.0 - 1e-9) - 1e-30 || td > hi_c[j] * (1.0 + 1e-9) + 1e-30) violated = 1;
    }
  }
  violated = __any_sync(0xffffffffu, violated);
  // What the candidate generator excluded: everything above the KP-th key it kept, and -- block-partitioned symmetric
  // self-join (vqacx_knn_selfjoin_shard) -- everything above the query's published threshold excl_thr, which was
  // warmed on ANOTHER rank: the >= KP keys that justify it live there, so this rank's own list may hold fewer than
  // KP keys although pairs were excluded. The exclusion bound is the smaller of the two.
  float a_last = cv[kTopkKeep - 1];
  bool have_all = (ci[kTopkKeep - 1] < 0);  // fewer than KP keys and no external threshold: nothing was excluded
  if (excl_thr) {
    const float thr = ordered_to_float(excl_thr[qrow]);
    if (have_all) {
      a_last = thr;
      have_all = !(thr < __int_as_float(0x7f800000));
    } else {
      a_last = fminf(a_last, thr);
    }
  }

  // --- bitonic sort of 64 (d2, idx) pairs in shared memory -----------------------------------
  PairLess less;
  for (int size = 2; size <= kTopkKeep; size <<= 1) {
    for (int dd = size >> 1; dd > 0; dd >>= 1) {
      __syncwarp();
      const int e = ((lane & ~(dd - 1)) << 1) | (lane & (dd - 1));  // lower index of the pair
      const int p = e | dd;
      const bool asc = ((e & size) == 0);
      const double da = s_d2[w][e], dbv = s_d2[w][p];
      const int ia = s_ix[w][e], ib = s_ix[w][p];
      const bool swap = asc ? less(dbv, ib, da, ia) : less(da, ia, dbv, ib);
      if (swap) {
        s_d2[w][e] = dbv; s_ix[w][e] = ib;
        s_d2[w][p] = da;  s_ix[w][p] = ia;
      }
    }
  }
  __syncwarp();

  for (int e = lane; e < k; e += 32) {
    const bool valid = e < kTopkKeep && s_ix[w][e] != 0x7fffffff;
    out_dist[qrow * out_ld + e] = valid ? sqrt(s_d2[w][e]) : __longlong_as_double(0x7ff0000000000000LL);
    out_idx[qrow * k + e] = valid ? (static_cast<long long>(s_ix[w][e]) + db_row0) : -1LL;
  }
  if (lane == 0) {
    bool ok = true;
    double lb = __longlong_as_double(0x7ff0000000000000LL);
    if (!have_all) {
      const int kk = min(k, kTopkKeep);
      const double dk = fmin(sqrt(s_d2[w][kk - 1]), hi_ext);   // sharded: the global k-th distance is <= hi_ext
      const double ap2 = static_cast<double>(a_last) + xlp2;
      lb = sqrt(fmax(ap2 * (1.0 - eps_sq_rel) - abs_term, 0.0)) - d_term;
      ok = (k <= kTopkKeep) && (lb > dk);
    }
    if (out_ld > k) {
      // column k of the output row = lower bound on the true distance of every pair this search EXCLUDED for the row
      // (+inf: nothing was excluded). Block-partitioned self-join: the proof "lb > global k-th distance" is completed
      // after the cross-rank merge, against the exact merged distance (vqacx_topk_merge_checked), not here against the
      // looser pre-exchange bound hi_ext -- so a failing comparison here is not a flag.
      out_dist[qrow * out_ld + k] = excl_thr ? lb : __longlong_as_double(0x7ff0000000000000LL);
      if (excl_thr && k <= kTopkKeep) ok = true;
    }
    if (violated) atomicAdd(&stats[kStatBoundViolations], 1);
    if (!ok || violated) {
      const int pos = atomicAdd(flag_count, 1);
      flag_rows[pos] = static_cast<int>(qrow);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// locality order for the re-rank: queries whose candidate lists overlap should run close in time, so that a gathered
// 8 KB row is served from L2 to the next query that wants it instead of from HBM again (every database row is gathered
// ~k' times per self-join). Key = smallest database index among a query's 32 best candidates: near-duplicates and
// members of one visual cluster share it. Counting sort of the queries by key (histogram, single-block exclusive scan,
// scatter); the order inside a bucket is arbitrary and does not affect any result (outputs are written per query row).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) knn_order_key_kernel(const int* __restrict__ cand_idx, int n_q, int n_db,
                                                            int* __restrict__ key, int* __restrict__ hist) {
  const int lane = threadIdx.x & 31;
  const int q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (q >= n_q) return;
  const int c = cand_idx[static_cast<size_t>(q) * kTopkKeep + lane];
  int m = __reduce_min_sync(0xffffffffu, c < 0 ? 0x7fffffff : c);
  m = min(m, n_db);   // a query without candidates sorts last
  if (lane == 0) {
    key[q] = m;
    atomicAdd(hist + m, 1);
  }
}

__global__ void __launch_bounds__(1024) knn_order_scan_kernel(int* __restrict__ hist, int n) {
  __shared__ int part[1024];
  const int t = threadIdx.x;
  const int per = (n + 1023) / 1024;
  const int b = min(n, t * per), e = min(n, b + per);
  int sum = 0;
  for (int i = b; i < e; ++i) sum += hist[i];
  part[t] = sum;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1) {   // inclusive Hillis-Steele scan of the 1024 partial sums
    const int v = t >= o ? part[t - o] : 0;
    __syncthreads();
    part[t] += v;
    __syncthreads();
  }
  int run = part[t] - sum;               // exclusive prefix of this thread's segment
  for (int i = b; i < e; ++i) {
    const int h = hist[i];
    hist[i] = run;
    run += h;
  }
}

__global__ void __launch_bounds__(256) knn_order_scatter_kernel(const int* __restrict__ key, int n_q,
                                                                int* __restrict__ cursor, int* __restrict__ perm) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n_q) return;
  perm[atomicAdd(cursor + key[q], 1)] = q;
}

cudaError_t launch_knn_locality_order(const int* cand_idx, int n_q, int n_db, int* key, int* hist, int* perm,
                                      cudaStream_t s) {
  if (n_q <= 0) return cudaSuccess;
  cudaError_t e = cudaMemsetAsync(hist, 0, static_cast<size_t>(n_db + 1) * sizeof(int), s);
  if (e != cudaSuccess) return e;
  count_launch(3);
  knn_order_key_kernel<<<(n_q + 7) / 8, 256, 0, s>>>(cand_idx, n_q, n_db, key, hist);
  knn_order_scan_kernel<<<1, 1024, 0, s>>>(hist, n_db + 1);
  knn_order_scatter_kernel<<<(n_q + 255) / 256, 256, 0, s>>>(key, n_q, hist, perm);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// cross-shard merge: G sorted (dist, idx) lists of length k per query -> best k by (dist, idx).
// One warp per query, G*k <= 256. Inputs laid out [G][n_q][k] (what an all_gather produces).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) knn_merge_kernel(const double* __restrict__ dist,
                                                        const long long* __restrict__ idx, int G, long long n_q,
                                                        int k, double* __restrict__ out_dist,
                                                        long long* __restrict__ out_idx, int dist_ld,
                                                        int* __restrict__ fail_count, int* __restrict__ fail_rows) {
  __shared__ double s_d[4][256];
  __shared__ long long s_i[4][256];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const long long row = static_cast<long long>(blockIdx.x) * 4 + w;
  if (row >= n_q) return;
  const int total = G * k;
  int n = 32;
  while (n < total) n <<= 1;
  for (int e = lane; e < n; e += 32) {
    if (e < total) {
      const int g = e / k, j = e % k;
      const size_t src = (static_cast<size_t>(g) * n_q + row) * k + j;
      const long long ix = idx[src];
      s_d[w][e] = ix >= 0 ? dist[(static_cast<size_t>(g) * n_q + row) * dist_ld + j] : __longlong_as_double(0x7ff0000000000000LL);
      s_i[w][e] = ix >= 0 ? ix : 0x7fffffffffffffffLL;
    } else {
      s_d[w][e] = __longlong_as_double(0x7ff0000000000000LL);
      s_i[w][e] = 0x7fffffffffffffffLL;
    }
  }
  for (int size = 2; size <= n; size <<= 1) {
    for (int dd = size >> 1; dd > 0; dd >>= 1) {
      __syncwarp();
      for (int t = lane; t < n / 2; t += 32) {
        const int e = ((t & ~(dd - 1)) << 1) | (t & (dd - 1));
        const int p = e | dd;
        const bool asc = ((e & size) == 0);
        const double da = s_d[w][e], dbv = s_d[w][p];
        const long long ia = s_i[w][e], ib = s_i[w][p];
        const bool b_lt_a = (dbv < da) || (dbv == da && ib < ia);
        const bool a_lt_b = (da < dbv) || (da == dbv && ia < ib);
        if (asc ? b_lt_a : a_lt_b) {
          s_d[w][e] = dbv; s_i[w][e] = ib;
          s_d[w][p] = da;  s_i[w][p] = ia;
        }
      }
    }
  }
  __syncwarp();
  // Equal (distance, index) pairs are adjacent after the sort. They only occur when a rank of the block-partitioned
  // self-join redid a flagged row with the exact search over ALL rows (its list then overlaps the other ranks'): keep
  // the first of each run. Lane l scans elements [l * per, (l + 1) * per); exclusive warp scan of the kept counts.
  const int per = n / 32;
  const int e0 = lane * per;
  int kept = 0;
  for (int j = 0; j < per; ++j) {
    const int e = e0 + j;
    const bool dup = e > 0 && s_i[w][e] == s_i[w][e - 1] && s_i[w][e] != 0x7fffffffffffffffLL;
    kept += dup ? 0 : 1;
  }
  int pos = kept;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, pos, o);
    if (lane >= o) pos += t;
  }
  pos -= kept;
  const bool any_dup = __reduce_add_sync(0xffffffffu, kept) != n;
  if (!any_dup) {
    for (int e = lane; e < k; e += 32) {
      const bool valid = s_i[w][e] != 0x7fffffffffffffffLL;
      out_dist[row * k + e] = s_d[w][e];
      out_idx[row * k + e] = valid ? s_i[w][e] : -1LL;
    }
  } else {
    for (int j = 0; j < per; ++j) {
      const int e = e0 + j;
      const bool dup = e > 0 && s_i[w][e] == s_i[w][e - 1] && s_i[w][e] != 0x7fffffffffffffffLL;
      if (!dup) {
        if (pos < k) {
          const bool valid = s_i[w][e] != 0x7fffffffffffffffLL;
          out_dist[row * k + pos] = s_d[w][e];
          out_idx[row * k + pos] = valid ? s_i[w][e] : -1LL;
        }
        ++pos;
      }
    }
  }
  // Completion of the exactness proof for lists that carry their exclusion bound in column k (dist_ld > k): every
  // shard excluded only pairs whose true distance exceeds its lb; the merged k-th distance must lie below all of them.
  if (dist_ld > k && fail_count) {
    __syncwarp();
    double lb_min = __longlong_as_double(0x7ff0000000000000LL);
    for (int g = lane; g < G; g += 32) lb_min = fmin(lb_min, dist[(static_cast<size_t>(g) * n_q + row) * dist_ld + k]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) lb_min = fmin(lb_min, __shfl_xor_sync(0xffffffffu, lb_min, o));
    if (lane == 0) {
      const double dk = out_dist[row * k + (k - 1)];
      // (fewer than k rows exist: dk = +inf and every lb is +inf as well -> nothing was excluded)
      const bool ok = (lb_min > dk) || (lb_min == __longlong_as_double(0x7ff0000000000000LL));
      if (!ok) fail_rows[atomicAdd(fail_count, 1)] = static_cast<int>(row);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// host-side launchers (called from abi.cu)
// ---------------------------------------------------------------------------------------------
cudaError_t launch_knn_prep(const float* x, long long n, int d, void* x_lp, float* norm_sq,
                            unsigned int* max_norm_sq_bits, int* bad_count, cudaStream_t s) {
  if (n == 0) return cudaSuccess;
  long long blocks = (n + 7) / 8;
  if (blocks > 148LL * 16) blocks = 148LL * 16;
  count_launch();
  knn_prep_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(x, n, d, static_cast<__half*>(x_lp), norm_sq,
                                                                max_norm_sq_bits, bad_count);
  return cudaGetLastError();
}

cudaError_t launch_knn_exact(const float* q, const int* q_rows, const int* n_q_ptr, int n_q, const float* db,
                             int n_db, int d, int nsplit, uint64_t* cand, int* cnt, cudaStream_t s) {
  if (n_q == 0 || n_db == 0) return cudaSuccess;
  const int n_tiles = (n_db + SBN - 1) / SBN;
  const int tps = (n_tiles + nsplit - 1) / nsplit;
  dim3 grid((n_q + SBM - 1) / SBM, nsplit);
  count_launch();
  knn_exact_kernel<<<grid, 256, 0, s>>>(q, q_rows, n_q_ptr, n_q, db, n_db, d, tps, cand, cnt);
  return cudaGetLastError();
}

// Opt-in wave overlap (abi_knn.cu): finalize / re-rank blocks run beside the candidate GEMM's CTAs. An SM holds one
// shared-memory carveout at a time, so they must ask for the same (maximum) one as the GEMM to be co-resident.
void knn_support_prefer_max_shared_carveout() {
  cudaFuncSetAttribute(knn_finalize_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  cudaFuncSetAttribute(knn_rerank_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  cudaGetLastError();
}

cudaError_t launch_knn_finalize(const uint64_t* cand, const int* cnt, const int* tile_slot0, const int* tile_nslots,
                                int nslots_uniform, int rows_per_tile, const int* n_q_ptr, int n_q, float* cand_val,
                                int* cand_idx, cudaStream_t s) {
  return launch_knn_finalize_range(cand, cnt, tile_slot0, tile_nslots, nslots_uniform, rows_per_tile, n_q_ptr, 0, n_q,
                                   cand_val, cand_idx, s);
}

cudaError_t launch_knn_finalize_sym(const uint64_t* cand, const int* cnt, const int* tile_slot0, const int* tile_nslots,
                                    int rows_per_tile, int n_q, const KnnSymArgs& sym, float* cand_val, int* cand_idx,
                                    int* flag_count, int* flag_rows, int* stats, cudaStream_t s) {
  if (n_q <= 0) return cudaSuccess;
  count_launch();
  knn_finalize_sym_kernel<<<(n_q + 7) / 8, 256, 0, s>>>(cand, cnt, tile_slot0, tile_nslots, rows_per_tile, n_q, sym,
                                                        cand_val, cand_idx, flag_count, flag_rows, stats);
  return cudaGetLastError();
}

// rows [row0, row1) only: lets the search re-rank the query tiles of one GEMM wave while the next wave runs
cudaError_t launch_knn_finalize_range(const uint64_t* cand, const int* cnt, const int* tile_slot0, const int* tile_nslots,
                                      int nslots_uniform, int rows_per_tile, const int* n_q_ptr, int row0, int row1,
                                      float* cand_val, int* cand_idx, cudaStream_t s) {
  if (row1 <= row0) return cudaSuccess;
  count_launch();
  knn_finalize_kernel<<<(row1 - row0 + 7) / 8, 256, 0, s>>>(cand, cnt, tile_slot0, tile_nslots, nslots_uniform,
                                                            rows_per_tile, n_q_ptr, row1, cand_val, cand_idx, row0);
  return cudaGetLastError();
}

cudaError_t launch_knn_rerank(const float* q, const int* q_rows, const int* n_q_ptr, int n_q, const float* db,
                              int n_db, int d, const float* cand_val, const int* cand_idx, const float* q_norm_sq_lp,
                              const unsigned int* db_max_norm_sq_bits, float eps_d, float d_abs, float eps_sq_abs,
                              float eps_sq_rel, int k, long long db_row0, double* out_dist, long long* out_idx,
                              int* flag_count, int* flag_rows, int* stats, const float* ext_bound, cudaStream_t s,
                              int out_ld) {
  return launch_knn_rerank_range(q, q_rows, n_q_ptr, 0, n_q, db, n_db, d, cand_val, cand_idx, q_norm_sq_lp,
                                 db_max_norm_sq_bits, eps_d, d_abs, eps_sq_abs, eps_sq_rel, k, db_row0, out_dist, out_idx,
                                 flag_count, flag_rows, stats, ext_bound, s, false, nullptr, out_ld);
}

cudaError_t launch_knn_rerank_range(const float* q, const int* q_rows, const int* n_q_ptr, int row0, int row1,
                                    const float* db, int n_db, int d, const float* cand_val, const int* cand_idx,
                                    const float* q_norm_sq_lp, const unsigned int* db_max_norm_sq_bits, float eps_d,
                                    float d_abs, float eps_sq_abs, float eps_sq_rel, int k, long long db_row0,
                                    double* out_dist, long long* out_idx, int* flag_count, int* flag_rows, int* stats,
                                    const float* ext_bound, cudaStream_t s, bool cand_by_qrow,
                                    const uint32_t* excl_thr, int out_ld) {
  if (row1 <= row0) return cudaSuccess;
  count_launch();
  knn_rerank_kernel<<<(row1 - row0 + 7) / 8, 256, 0, s>>>(q, q_rows, n_q_ptr, row1, db, n_db, d, cand_val, cand_idx,
                                                          q_norm_sq_lp, db_max_norm_sq_bits, eps_d, d_abs, eps_sq_abs,
                                                          eps_sq_rel, k, db_row0, out_dist, out_idx, flag_count,
                                                          flag_rows, stats, ext_bound, row0, cand_by_qrow ? 1 : 0,
                                                          excl_thr, out_ld > k ? out_ld : k);
  return cudaGetLastError();
}

// Threshold exchange of the block-partitioned symmetric self-join: the per-query thresholds travel between ranks as
// plain floats (an all_reduce(MIN) over float32 is what every backend offers); +inf = no threshold yet.
__global__ void knn_thr_export_kernel(const uint32_t* __restrict__ gtau, int n, float* __restrict__ thr) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) thr[i] = ordered_to_float(gtau[i]);
}
__global__ void knn_thr_import_kernel(const float* __restrict__ thr, int n, uint32_t* __restrict__ gtau) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) gtau[i] = float_to_ordered(thr[i]);
}
cudaError_t launch_knn_thr_export(const uint32_t* gtau, int n, float* thr, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  count_launch();
  knn_thr_export_kernel<<<(n + 255) / 256, 256, 0, s>>>(gtau, n, thr);
  return cudaGetLastError();
}
cudaError_t launch_knn_thr_import(const float* thr, int n, uint32_t* gtau, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  count_launch();
  knn_thr_import_kernel<<<(n + 255) / 256, 256, 0, s>>>(thr, n, gtau);
  return cudaGetLastError();
}
// ---- candidate exchange of the block-partitioned self-join ---------------------------------------------------------------
// After its share of the blocks a rank holds, per query, the <= KP best candidates among the database blocks it covered.
// Instead of re-ranking them where they are (every rank walking all n queries) the lists travel to the rank that OWNS
// the query (all_to_all), which merges the G lists to the global KP best and runs the ordinary single-list re-rank and
// exactness proof on its n / G queries. One packed row per query: kPackInts int32 =
//   [0, KP) approximate values (float bits) | [KP, 2 KP) database rows (-1 = none) | [2 KP] exclusion bound (float bits):
//   every pair this rank excluded for the query has an approximate value >= it (+inf: nothing excluded) | [2 KP + 1] pad.
__global__ void __launch_bounds__(256) knn_cand_pack_kernel(const float* __restrict__ cand_val,
                                                            const int* __restrict__ cand_idx,
                                                            const uint32_t* __restrict__ gtau, int n, int rows_alloc,
                                                            const int* __restrict__ overflow_flag,
                                                            const int* __restrict__ mcnt, int mcap,
                                                            int* __restrict__ pack) {
  const int lane = threadIdx.x & 31;
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= rows_alloc) return;
  int* out = pack + static_cast<size_t>(row) * kPackInts;
  if (row >= n) {   // exchange padding
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      out[j * 32 + lane] = 0x7f800000;
      out[kTopkKeep + j * 32 + lane] = -1;
    }
    if (lane < 2) out[2 * kTopkKeep + lane] = 0x7f800000;
    return;
  }
  float last = 0.f;
  int last_idx = 0;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const float v = cand_val[static_cast<size_t>(row) * kTopkKeep + j * 32 + lane];
    const int ix = cand_idx[static_cast<size_t>(row) * kTopkKeep + j * 32 + lane];
    out[j * 32 + lane] = __float_as_int(v);
    out[kTopkKeep + j * 32 + lane] = ix;
    if (j == 1) { last = v; last_idx = ix; }
  }
  last = __shfl_sync(0xffffffffu, last, 31);
  last_idx = __shfl_sync(0xffffffffu, last_idx, 31);
  if (lane == 0) {
    const float thr = ordered_to_float(gtau[row]);
    float excl = last_idx >= 0 ? fminf(last, thr) : thr;      // a full list also excluded everything above its last key
    // keys were dropped (an append log, or this query's mirror buffer, overflowed): the list proves nothing -> -inf makes
    // the owner's proof fail for the query, which is then recomputed exactly
    if (*overflow_flag != 0 || mcnt[static_cast<size_t>(row) * kMirrorCntStride] > mcap) excl = __int_as_float(0xff800000);
    out[2 * kTopkKeep] = __float_as_int(excl);
    out[2 * kTopkKeep + 1] = 0;
  }
}
cudaError_t launch_knn_cand_pack(const float* cand_val, const int* cand_idx, const uint32_t* gtau, int n, int rows_alloc,
                                 const int* overflow_flag, const int* mcnt, int mcap, int* pack, cudaStream_t s) {
  if (rows_alloc <= 0) return cudaSuccess;
  count_launch();
  knn_cand_pack_kernel<<<(rows_alloc + 7) / 8, 256, 0, s>>>(cand_val, cand_idx, gtau, n, rows_alloc, overflow_flag, mcnt,
                                                            mcap, pack);
  return cudaGetLastError();
}

// Owner side: G packed rows per query (recv [G, per, kPackInts]) -> the KP smallest keys by (value, row) and the smallest
// exclusion bound. One warp per query.
__global__ void __launch_bounds__(256) knn_cand_merge_kernel(const int* __restrict__ recv, int G, int per, int n_slice,
                                                             float* __restrict__ cand_val, int* __restrict__ cand_idx,
                                                             uint32_t* __restrict__ excl_ord) {
  const int lane = threadIdx.x & 31;
  const int j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (j >= n_slice) return;
  uint64_t k[4] = {kKeyPad, kKeyPad, kKeyPad, kKeyPad};
  float excl = __int_as_float(0x7f800000);
  for (int g = 0; g < G; ++g) {
    const int* in = recv + (static_cast<size_t>(g) * per + j) * kPackInts;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int ix = in[kTopkKeep + h * 32 + lane];
      const float v = __int_as_float(in[h * 32 + lane]);
      k[2 + h] = ix >= 0 ? make_key(v, static_cast<uint32_t>(ix)) : kKeyPad;
    }
    warp_sort128(k);
    excl = fminf(excl, __int_as_float(in[2 * kTopkKeep]));
  }
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const bool pad = (k[h] == kKeyPad);
    cand_val[static_cast<size_t>(j) * kTopkKeep + h * 32 + lane] = pad ? __int_as_float(0x7f800000) : key_value(k[h]);
    cand_idx[static_cast<size_t>(j) * kTopkKeep + h * 32 + lane] = pad ? -1 : static_cast<int>(key_col(k[h]));
  }
  if (lane == 0) excl_ord[j] = float_to_ordered(excl);
}
cudaError_t launch_knn_cand_merge(const int* recv, int G, int per, int n_slice, float* cand_val, int* cand_idx,
                                  uint32_t* excl_ord, cudaStream_t s) {
  if (n_slice <= 0) return cudaSuccess;
  count_launch();
  knn_cand_merge_kernel<<<(n_slice + 7) / 8, 256, 0, s>>>(recv, G, per, n_slice, cand_val, cand_idx, excl_ord);
  return cudaGetLastError();
}

// flag [1] int -> dst[0] = flag != 0 ? 1.f : 0.f, dst[1 .. n_fill) = 0 (the overflow flag rides in the bound exchange)
__global__ void knn_flag_to_float_kernel(const int* __restrict__ flag, float* __restrict__ dst, int n_fill) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_fill) dst[i] = (i == 0 && *flag != 0) ? 1.f : 0.f;
}
cudaError_t launch_knn_flag_to_float(const int* flag, float* dst, int n_fill, cudaStream_t s) {
  count_launch();
  knn_flag_to_float_kernel<<<(n_fill + 255) / 256, 256, 0, s>>>(flag, dst, n_fill);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Row-sharded search, phase 1 -> 2 glue.
// knn_local_hi: for each query, upper bounds hi_c on the TRUE distances of this shard's k best approximate candidates
// (same error model as the re-rank; +inf where the shard has fewer than k rows). One warp per query.
// knn_kth_bound: the k-th smallest of the G*k bounds gathered from all shards = an upper bound on the global k-th
// distance (k candidates somewhere are provably that close).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) knn_local_hi_kernel(const float* __restrict__ q, int n_q, int d,
                                                           const float* __restrict__ cand_val,
                                                           const int* __restrict__ cand_idx,
                                                           const float* __restrict__ q_norm_sq_lp,
                                                           const unsigned int* __restrict__ db_max_norm_sq_bits,
                                                           float eps_d, float d_abs, float eps_sq_abs, float eps_sq_rel,
                                                           int k, float* __restrict__ hi_out) {
  const int lane = threadIdx.x & 31;
  const int lq = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (lq >= n_q) return;
  const float4* xr = reinterpret_cast<const float4*>(q + static_cast<long long>(lq) * d);
  double xn2 = 0.0;
  for (int i = lane; i < d / 4; i += 32) {
    const float4 v = xr[i];
    xn2 += static_cast<double>(v.x) * v.x + static_cast<double>(v.y) * v.y + static_cast<double>(v.z) * v.z +
           static_cast<double>(v.w) * v.w;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) xn2 += __shfl_xor_sync(0xffffffffu, xn2, o);
  const float ymax = db_max_norm_sq_bits ? sqrtf(__uint_as_float(*db_max_norm_sq_bits)) : 0.f;
  const double xn = sqrt(xn2);
  const double xlp2 = q_norm_sq_lp ? static_cast<double>(q_norm_sq_lp[lq]) : 0.0;
  const double abs_term = static_cast<double>(eps_sq_abs) * xn * ymax;
  const double d_term = static_cast<double>(eps_d) * (xn + ymax) + d_abs;
  for (int c = lane; c < k; c += 32) {
    float hi = __int_as_float(0x7f800000);
    if (c < kTopkKeep && cand_idx[static_cast<size_t>(lq) * kTopkKeep + c] >= 0) {
      const double ap2 = static_cast<double>(cand_val[static_cast<size_t>(lq) * kTopkKeep + c]) + xlp2;
      const double h = sqrt(fmax(ap2 * (1.0 + eps_sq_rel) + abs_term, 0.0)) + d_term;
      hi = __double2float_ru(h);   // round up: it must stay an upper bound
    }
    hi_out[static_cast<size_t>(lq) * k + c] = hi;
  }
}
cudaError_t launch_knn_local_hi(const float* q, int n_q, int d, const float* cand_val, const int* cand_idx,
                                const float* q_norm_sq_lp, const unsigned int* db_max_norm_sq_bits, float eps_d,
                                float d_abs, float eps_sq_abs, float eps_sq_rel, int k, float* hi_out, cudaStream_t s) {
  if (n_q == 0) return cudaSuccess;
  count_launch();
  knn_local_hi_kernel<<<(n_q + 7) / 8, 256, 0, s>>>(q, n_q, d, cand_val, cand_idx, q_norm_sq_lp, db_max_norm_sq_bits,
                                                    eps_d, d_abs, eps_sq_abs, eps_sq_rel, k, hi_out);
  return cudaGetLastError();
}

// hi_all [G][n_q][k] (non-negative floats, +inf padding) -> bound[q] = k-th smallest of the G*k values (G*k <= 256).
// One warp per query; bisection on the float bit pattern (order-preserving for non-negative floats).
__global__ void __launch_bounds__(256) knn_kth_bound_kernel(const float* __restrict__ hi_all, int G, long long n_q, int k,
                                                            float* __restrict__ bound) {
  const int lane = threadIdx.x & 31;
  const long long qi = static_cast<long long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (qi >= n_q) return;
  const int tot = G * k;
  uint32_t v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int e = j * 32 + lane;
    float f = __int_as_float(0x7f800000);
    if (e < tot) f = hi_all[(static_cast<long long>(e / k) * n_q + qi) * k + (e % k)];
    v[j] = __float_as_uint(fmaxf(f, 0.f));
  }
  // smallest T with #{v <= T} >= k
  uint32_t lo = 0u, hi = 0x7f800000u;   // #{v <= hi} = 256 >= k
  while (lo < hi) {
    const uint32_t mid = lo + ((hi - lo) >> 1);
    int c = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) c += (v[j] <= mid) ? 1 : 0;
    c = __reduce_add_sync(0xffffffffu, c);
    if (c >= k) hi = mid;
    else lo = mid + 1u;
  }
  if (lane == 0) bound[qi] = __uint_as_float(lo);
}
cudaError_t launch_knn_kth_bound(const float* hi_all, int G, long long n_q, int k, float* bound, cudaStream_t s) {
  if (n_q == 0) return cudaSuccess;
  count_launch();
  knn_kth_bound_kernel<<<static_cast<unsigned>((n_q + 7) / 8), 256, 0, s>>>(hi_all, G, n_q, k, bound);
  return cudaGetLastError();
}

cudaError_t launch_knn_merge(const double* dist, const long long* idx, int G, long long n_q, int k,
                             double* out_dist, long long* out_idx, cudaStream_t s, int dist_ld, int* fail_count,
                             int* fail_rows) {
  if (n_q == 0) return cudaSuccess;
  count_launch();
  knn_merge_kernel<<<static_cast<unsigned>((n_q + 3) / 4), 128, 0, s>>>(dist, idx, G, n_q, k, out_dist, out_idx,
                                                                      dist_ld > k ? dist_ld : k, fail_count, fail_rows);
  return cudaGetLastError();
}

}  // namespace vqacx

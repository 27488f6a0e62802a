// Streaming top-k' selection shared by the tcgen05 distance kernel and the exact fp32 kernel.
//
// Replaces sklearn's per-row max-heap (`heap_push` rejecting val >= heap[0], then
// `simultaneous_sort`; reached from knn.py:49 `nbrs.kneighbors`) with a design that suits one
// thread per query row consuming accumulator columns: a register threshold `tau` rejects almost every
// candidate with one compare; survivors are appended to a per-row buffer of CAP 64-bit keys
// (ordered-float distance << 32 | column) in L2-resident global scratch; when a row's buffer is
// nearly full the whole warp sorts it (bitonic, 4 keys per lane) and keeps the KP smallest, which
// tightens tau. With CAP = 2*KP the seen-count grows geometrically between compactions, so a row
// pays O(log(N/KP)) sorts per sweep. Keys are unique (column ids differ) so ties order by index.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace vqacx {

constexpr int kTopkCap = 256;      // per-row candidate buffer entries
constexpr int kTopkKeep = 64;      // k' handed to the re-rank; a compaction keeps between kTopkKeep and kTopkKeepMax
constexpr int kTopkKeepMax = 96;
constexpr uint64_t kKeyPad = ~0ull;

__device__ __forceinline__ uint32_t float_to_ordered(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ordered_to_float(uint32_t o) {
  uint32_t u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
  return __uint_as_float(u);
}
__device__ __forceinline__ uint64_t make_key(float v, uint32_t col) {
  return (static_cast<uint64_t>(float_to_ordered(v)) << 32) | col;
}
__device__ __forceinline__ float key_value(uint64_t k) { return ordered_to_float(static_cast<uint32_t>(k >> 32)); }
__device__ __forceinline__ uint32_t key_col(uint64_t k) { return static_cast<uint32_t>(k); }

__device__ __forceinline__ uint64_t shfl_xor_u64(uint64_t v, int mask) {
  uint32_t lo = __shfl_xor_sync(0xffffffffu, static_cast<uint32_t>(v), mask);
  uint32_t hi = __shfl_xor_sync(0xffffffffu, static_cast<uint32_t>(v >> 32), mask);
  return (static_cast<uint64_t>(hi) << 32) | lo;
}
__device__ __forceinline__ uint64_t shfl_u64(uint64_t v, int src) {
  uint32_t lo = __shfl_sync(0xffffffffu, static_cast<uint32_t>(v), src);
  uint32_t hi = __shfl_sync(0xffffffffu, static_cast<uint32_t>(v >> 32), src);
  return (static_cast<uint64_t>(hi) << 32) | lo;
}

// Ascending bitonic sort of 128 keys held by one warp: element e = j*32 + lane lives in k[j].
__device__ __forceinline__ void warp_sort128(uint64_t (&k)[4]) {
  const uint32_t lane = threadIdx.x & 31;
#pragma unroll
  for (int size = 2; size <= 128; size <<= 1) {
#pragma unroll
    for (int d = size >> 1; d > 0; d >>= 1) {
      if (d >= 32) {
        const int dj = d >> 5;  // partner register distance
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if ((j & dj) == 0) {
            const int e = j * 32;  // lane bits do not reach `size` >= 64 ... but size may be 64 or 128
            const bool asc = (((e | lane) & size) == 0);
            uint64_t a = k[j], b = k[j | dj];
            uint64_t lo = a < b ? a : b, hi = a < b ? b : a;
            k[j] = asc ? lo : hi;
            k[j | dj] = asc ? hi : lo;
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t e = j * 32 + lane;
          const bool asc = ((e & size) == 0);
          const bool lower = ((lane & d) == 0);
          uint64_t other = shfl_xor_u64(k[j], d);
          uint64_t mn = k[j] < other ? k[j] : other, mx = k[j] < other ? other : k[j];
          k[j] = (lower == asc) ? mn : mx;
        }
      }
    }
  }
}

// Per-thread (= per query row) filter state.
struct RowFilter {
  float tau;      // accept v < tau
  int cnt;        // entries currently in buf
  uint64_t* buf;  // kTopkCap keys in global scratch
};

__device__ __forceinline__ void filter_reset(RowFilter& st, uint64_t* buf) {
  st.tau = __int_as_float(0x7f800000);  // +inf
  st.cnt = 0;
  st.buf = buf;
}

__device__ __forceinline__ void filter_push(RowFilter& st, float v, uint32_t col) {
  if (v < st.tau) {
    __stcg(reinterpret_cast<unsigned long long*>(st.buf + st.cnt), make_key(v, col));
    st.cnt++;
  }
}

// Warp-collective: compact every lane's row whose buffer cannot take `room` more appends.
// Must be called by all 32 lanes (converged). After it, every lane has cnt <= kTopkCap - room.
//
// A compaction does NOT sort. It needs (a) a threshold T with  kTopkKeep <= #{keys < T} <= kTopkKeepMax  and (b) those
// keys moved to the front of the buffer in any order (the finalize kernel sorts once at the end). T is found by
// bisection on the 32-bit ordered distance between the buffer's min and max: each step is 8 compares + one warp
// reduction, and because any count in [64, 96] is acceptable it converges in a handful of steps (~10x fewer
// instructions than the bitonic sort this replaces, which made the epilogue the critical path at the start of every
// database sweep, when tau = +inf and all 32 rows of a warp overflow together). tau becomes T: every kept key is < T
// and at least kTopkKeep of them are, so a later candidate with v >= T cannot be among the kTopkKeep smallest.
// Rows that need compaction are processed one after another with the next row's keys already in flight.
__device__ __forceinline__ void filter_load_keys(const uint64_t* b, int c, uint32_t lane, uint64_t (&k)[kTopkCap / 32]) {
#pragma unroll
  for (int j = 0; j < kTopkCap / 32; ++j) {
    const int e = j * 32 + static_cast<int>(lane);
    k[j] = (e < c) ? __ldcg(reinterpret_cast<const unsigned long long*>(b + e)) : kKeyPad;
  }
}

__device__ __forceinline__ void filter_compact_warp(RowFilter& st, int room) {
  constexpr int kRegs = kTopkCap / 32;
  const uint32_t lane = threadIdx.x & 31;
  unsigned need = __ballot_sync(0xffffffffu, st.cnt > kTopkCap - room);
  if (need == 0u) return;
  __syncwarp();  // order every owner's appends before the other lanes' reads
  int src = __ffs(need) - 1;
  uint64_t* b = reinterpret_cast<uint64_t*>(shfl_u64(reinterpret_cast<uint64_t>(st.buf), src));
  int c = __shfl_sync(0xffffffffu, st.cnt, src);
  uint64_t k[kRegs];
  filter_load_keys(b, c, lane, k);
  while (need) {
    need &= need - 1;
    // prefetch the next row that needs compaction
    const int src_n = need ? __ffs(need) - 1 : src;
    uint64_t* b_n = reinterpret_cast<uint64_t*>(shfl_u64(reinterpret_cast<uint64_t>(st.buf), src_n));
    const int c_n = __shfl_sync(0xffffffffu, st.cnt, src_n);
    uint64_t kn[kRegs];
    if (need) filter_load_keys(b_n, c_n, lane, kn);

    uint32_t h[kRegs];
    uint32_t mn = 0xffffffffu, mx = 0u;
#pragma unroll
    for (int j = 0; j < kRegs; ++j) {
      h[j] = static_cast<uint32_t>(k[j] >> 32);           // pad keys read 0xffffffff and never count as < T
      const bool valid = (j * 32 + static_cast<int>(lane)) < c;
      mn = min(mn, h[j]);
      mx = valid ? max(mx, h[j]) : mx;
    }
    uint32_t lo = __reduce_min_sync(0xffffffffu, mn);      // #{< lo} = 0 < kTopkKeep
    uint32_t hi = __reduce_max_sync(0xffffffffu, mx) + 1u; // #{< hi} = c > kTopkKeepMax
    uint32_t T = hi;
    int tie_quota = -1;                                    // >= 0: keep keys < lo plus this many keys == lo
    if (c > kTopkKeepMax) {
      for (;;) {
        if (hi - lo <= 1u) {  // more than kTopkKeepMax - kTopkKeep keys tie exactly at `lo` (duplicate rows)
          int below = 0;
#pragma unroll
          for (int j = 0; j < kRegs; ++j) below += (h[j] < lo) ? 1 : 0;
          below = __reduce_add_sync(0xffffffffu, below);
          tie_quota = kTopkKeep - below;
          T = lo + 1u;
          break;
        }
        const uint32_t mid = lo + ((hi - lo) >> 1);
        int n_lt = 0;
#pragma unroll
        for (int j = 0; j < kRegs; ++j) n_lt += (h[j] < mid) ? 1 : 0;
        n_lt = __reduce_add_sync(0xffffffffu, n_lt);
        if (n_lt < kTopkKeep) lo = mid;
        else if (n_lt > kTopkKeepMax) hi = mid;
        else { T = mid; break; }
      }
    }
    // move the kept keys to the front (stable in slot order)
    int base = 0;
#pragma unroll
    for (int j = 0; j < kRegs; ++j) {
      bool keep = h[j] < (tie_quota >= 0 ? lo : T);
      if (tie_quota >= 0) {
        const bool tie = (h[j] == lo);
        const unsigned tb = __ballot_sync(0xffffffffu, tie);
        const int rank = __popc(tb & ((1u << lane) - 1u));
        if (tie && rank < tie_quota) keep = true;
        tie_quota = max(0, tie_quota - __popc(tb));
      }
      const unsigned kb = __ballot_sync(0xffffffffu, keep);
      if (keep) __stcg(reinterpret_cast<unsigned long long*>(b + base + __popc(kb & ((1u << lane) - 1u))), k[j]);
      base += __popc(kb);
    }
    if (lane == static_cast<uint32_t>(src)) {
      st.cnt = base;
      if (c > kTopkKeepMax) st.tau = ordered_to_float(T);
    }
    __syncwarp();
    src = src_n; b = b_n; c = c_n;
#pragma unroll
    for (int j = 0; j < kRegs; ++j) k[j] = kn[j];
  }
}

}  // namespace vqacx

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

constexpr int kTopkCap = 128;  // per-row candidate buffer entries
constexpr int kTopkKeep = 64;  // k' kept by a compaction (and handed to the re-rank)
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
__device__ __forceinline__ void filter_compact_warp(RowFilter& st, int room) {
  const uint32_t lane = threadIdx.x & 31;
  unsigned need = __ballot_sync(0xffffffffu, st.cnt > kTopkCap - room);
  while (need) {
    const int src = __ffs(need) - 1;
    need &= need - 1;
    __syncwarp();  // order the owner's appends before the other lanes' reads
    uint64_t* b = reinterpret_cast<uint64_t*>(shfl_u64(reinterpret_cast<uint64_t>(st.buf), src));
    const int c = __shfl_sync(0xffffffffu, st.cnt, src);
    uint64_t k[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int e = j * 32 + lane;
      k[j] = (e < c) ? __ldcg(reinterpret_cast<const unsigned long long*>(b + e)) : kKeyPad;
    }
    warp_sort128(k);
    __stcg(reinterpret_cast<unsigned long long*>(b + lane), k[0]);
    __stcg(reinterpret_cast<unsigned long long*>(b + 32 + lane), k[1]);
    const uint32_t kth = __shfl_sync(0xffffffffu, static_cast<uint32_t>(k[1] >> 32), 31);
    if (lane == src) {
      st.cnt = c < kTopkKeep ? c : kTopkKeep;
      if (c >= kTopkKeep) st.tau = ordered_to_float(kth);
    }
    __syncwarp();
  }
}

}  // namespace vqacx

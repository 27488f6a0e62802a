// Shared epilogue of the tcgen05 distance kernels (knn_gemm.cu, knn_gemm_pair.cu).
#pragma once
#include "knn_internal.cuh"
#include "ptx.cuh"
#include "topk.cuh"

namespace vqacx {

// Per-thread staging row in shared memory for the values of one 32-column chunk: 36 floats (144 B) per thread, so the
// 8 x 16-byte stores of a quarter warp fall in distinct banks; 128 epilogue threads -> 18 KB.
constexpr int kEpiStagePitch = 36;
constexpr int kEpiStageBytes = 128 * kEpiStagePitch * 4;

// Epilogue of one [128 x 256] accumulator tile: this thread's TMEM lane (= query row) x 256 columns,
// v = |y~|^2 - 2 x~.y~, threshold filter + warp compaction (topk.cuh).
//
// A warp owns 32 rows, so "some lane has a hit in this chunk" is true for almost every chunk (a row accepts
// ~64 ln(N/64) of N columns). The hit path therefore must not cost a 32-step serial chain per chunk: the compare is
// folded into the FMA pass as a bitmask (independent instructions), the values are parked in this thread's shared-memory
// row only when the mask is non-zero, and the push loop runs popcount(mask) (typically 1) iterations.
template <bool kProf = false>
__device__ __forceinline__ void knn_epilogue_tile(RowFilter& st, uint32_t t_addr, const float* yn, uint32_t col_base,
                                                  float* vrow, long long* ld_cycles = nullptr,
                                                  long long* push_cycles = nullptr) {
#pragma unroll 1
  for (int chunk = 0; chunk < kGemmBN / 32; ++chunk) {
    uint32_t r[32];
    long long c0 = 0;
    if (kProf) c0 = clock64();
    tmem_ld_32x32(t_addr + chunk * 32, r);
    tmem_ld_wait();
    if (kProf) *ld_cycles += clock64() - c0;
    float v[32];
    uint32_t mask = 0;
    const float tau = st.tau;
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      const float4 y4 = *reinterpret_cast<const float4*>(yn + chunk * 32 + j);
      v[j + 0] = fmaf(-2.f, __uint_as_float(r[j + 0]), y4.x);
      v[j + 1] = fmaf(-2.f, __uint_as_float(r[j + 1]), y4.y);
      v[j + 2] = fmaf(-2.f, __uint_as_float(r[j + 2]), y4.z);
      v[j + 3] = fmaf(-2.f, __uint_as_float(r[j + 3]), y4.w);
      mask |= (v[j + 0] < tau ? 1u : 0u) << (j + 0);
      mask |= (v[j + 1] < tau ? 1u : 0u) << (j + 1);
      mask |= (v[j + 2] < tau ? 1u : 0u) << (j + 2);
      mask |= (v[j + 3] < tau ? 1u : 0u) << (j + 3);
    }
    if (kProf) c0 = clock64();
    if (mask != 0u) {
#pragma unroll
      for (int j = 0; j < 32; j += 4)
        *reinterpret_cast<float4*>(vrow + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
      const uint32_t col0 = col_base + chunk * 32;
      uint64_t* dst = st.buf + st.cnt;
      st.cnt += __popc(mask);
      while (mask != 0u) {
        const int j = __ffs(mask) - 1;
        mask &= mask - 1u;
        __stcg(reinterpret_cast<unsigned long long*>(dst++), make_key(vrow[j], col0 + j));
      }
    }
    filter_compact_warp(st, 32);
    if (kProf) *push_cycles += clock64() - c0;
  }
}

// Symmetric self-join, COLUMN side of one [128 x 256] accumulator tile (run by the pair kernel's mirror warps, beside
// the row-side warps that read the same TMEM lanes): the thread (database row i of the block = TMEM lane, squared fp16
// norm xn, global row id q_row) tests u = xn - 2 x~.y~ against the threshold ct[j] of the column's query (shared
// memory, refreshed per tile from the published per-query thresholds). u is bit-identical to the value the plain kernel
// computes for (query j, row i): the dot product is the same accumulator and the norm term is the same array entry.
// Hits go to the warp's private APPEND LOG (16-byte entries {key, column}; slot = warp-aggregated prefix over the
// lanes that hit, the running length lives in a register): coalesced fire-and-forget stores, no atomics and no
// round trip inside the GEMM. knn_mirror_scatter_kernel files the entries under their queries afterwards.
// History (all measured at 123 287 x 2048, mirrored launch; 17.5 ms when the column side only compares):
//   both sides in the same warps, atomicAdd slot reservation per hit ............ 32-42 ms (epilogue-bound)
//   separate mirror warps, per-thread pending lists flushed per tile ............ 26-29 ms
//   + reservation pipelined one tile ahead, accumulator released before it ...... 26-29 ms (no change)
//   + 8 / 16 replicas of every counter (same-address contention?) ............... 29 ms    (no change)
//   appends dropped (timing experiment) ......................................... 20.5 ms
// i.e. the returning atomics themselves, not their latency or contention, were the cost.
//
// r[j] for a run-time j without dynamic register indexing (which would spill r[] to local memory): 31 selects.
__device__ __forceinline__ uint32_t select32(const uint32_t (&r)[32], int j) {
  uint32_t a[16], b[8], c[4], d[2];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = (j & 16) ? r[i + 16] : r[i];
#pragma unroll
  for (int i = 0; i < 8; ++i) b[i] = (j & 8) ? a[i + 8] : a[i];
#pragma unroll
  for (int i = 0; i < 4; ++i) c[i] = (j & 4) ? b[i + 4] : b[i];
#pragma unroll
  for (int i = 0; i < 2; ++i) d[i] = (j & 2) ? c[i + 2] : c[i];
  return (j & 1) ? d[1] : d[0];
}

// One mirror warp's append log. Hits are first parked in a 64-entry ring in shared memory and leave it 32 at a time:
// one warp-wide store of 512 contiguous bytes (whole 128-byte lines) per 32 hits instead of one store instruction of
// 2-5 scattered 16-byte lanes per loop iteration. Measured (123 287 x 2048, mirrored launch, role counters): with the
// per-iteration stores the MMA issuer waited 23.6 Mcycles for operands against 11.1 Mcycles with the stores compiled
// out -- the partial-sector stores slowed the TMA operand stream of the whole SM, not the mirror warps themselves.
// `len` may exceed `cap`: entries past it are dropped and the launcher's scatter pass reports the overflow, which makes
// the host redo the search with the plain sweep.
constexpr int kMirrorRing = 64;                             // entries per mirror warp
constexpr int kMirrorWarps = 8;                              // two per TMEM lane quarter: even / odd 32-column chunks
constexpr int kMirrorStageBytes = kMirrorWarps * kMirrorRing * 16;
struct MirrorLog {
  uint4* log;      // this warp's region of the global log
  uint4* ring;     // kMirrorRing entries of shared memory
  int len;         // entries already written to `log` (warp-uniform)
  int cap;
  int head, fill;  // ring state (warp-uniform); fill < 32 between calls
};

__device__ __forceinline__ void mirror_log_flush32(MirrorLog& m, uint32_t lane, int debug = 0) {
  __syncwarp();
  const uint4 e = m.ring[(m.head + lane) & (kMirrorRing - 1)];
  if (m.len + static_cast<int>(lane) < m.cap) {
#ifdef VQACX_DEBUG
    // timing experiments: 512 = no global store (results invalid), 128 = L2 evict_first hint, 256 = st.global.cs
    uint4* dst = m.log + m.len + lane;
    if (debug & 512) {
    } else if (debug & 128) {
      const uint64_t pol = l2_policy_evict_first();
      asm volatile("st.global.L2::cache_hint.v4.b32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(dst), "r"(e.x), "r"(e.y),
                   "r"(e.z), "r"(e.w), "l"(pol) : "memory");
    } else if (debug & 256) {
      __stcs(dst, e);
    } else {
      __stcg(dst, e);
    }
#else
    (void)debug;
    __stcg(m.log + m.len + lane, e);
#endif
  }
  m.len += 32;
  m.head = (m.head + 32) & (kMirrorRing - 1);
  m.fill -= 32;
  __syncwarp();
}

// end of the launch: the < 32 entries still parked
__device__ __forceinline__ void mirror_log_drain(MirrorLog& m, uint32_t lane) {
  __syncwarp();
  if (static_cast<int>(lane) < m.fill && m.len + static_cast<int>(lane) < m.cap)
    __stcg(m.log + m.len + lane, m.ring[(m.head + lane) & (kMirrorRing - 1)]);
  m.len += m.fill;
  m.fill = 0;
  __syncwarp();
}

// Per-lane queue of hits waiting for a warp-wide append. A lane meets ~1.1 hits per tile, so with three slots the warp
// synchronises (ballots, slot arithmetic, ring writes) once every few TILES instead of once or twice per 32-column
// chunk -- the per-chunk hit loop, ~550 dependent cycles per chunk, was what made a tile's column-side scan take 11.4
// Mcycles per launch against 5.4 for the compares alone (role counters).
constexpr int kMirrorQueueDepth = 4;
struct MirrorQueue {
  float v0, v1, v2, v3;
  uint32_t c0, c1, c2, c3;
  int n;
};
__device__ __forceinline__ void mirror_queue_reset(MirrorQueue& q) {
  q.v0 = q.v1 = q.v2 = q.v3 = 0.f;
  q.c0 = q.c1 = q.c2 = q.c3 = 0u;
  q.n = 0;
}
__device__ __forceinline__ void mirror_queue_put(MirrorQueue& q, float v, uint32_t c) {   // caller checked q.n < depth
  if (q.n == 0) { q.v0 = v; q.c0 = c; }
  else if (q.n == 1) { q.v1 = v; q.c1 = c; }
  else if (q.n == 2) { q.v2 = v; q.c2 = c; }
  else { q.v3 = v; q.c3 = c; }
  ++q.n;
}

// warp-wide: every lane's queued hits go to the ring (and on to the log, 32 at a time); queues are empty afterwards
__device__ __forceinline__ void mirror_queue_flush(MirrorQueue& q, MirrorLog& m, uint32_t q_row, uint32_t lane, int debug) {
  const unsigned lt_mask = (1u << lane) - 1u;
#pragma unroll
  for (int i = 0; i < kMirrorQueueDepth; ++i) {
    const bool has = q.n > i;
    const unsigned any = __ballot_sync(0xffffffffu, has);
    if (any == 0u) break;
    if (has) {
      const float v = i == 0 ? q.v0 : (i == 1 ? q.v1 : (i == 2 ? q.v2 : q.v3));
      const uint32_t c = i == 0 ? q.c0 : (i == 1 ? q.c1 : (i == 2 ? q.c2 : q.c3));
      const uint64_t key = make_key(v, q_row);
      const int slot = (m.head + m.fill + __popc(any & lt_mask)) & (kMirrorRing - 1);
      if (!(debug & 16)) m.ring[slot] = make_uint4(static_cast<uint32_t>(key), static_cast<uint32_t>(key >> 32), c, 0u);
    }
    m.fill += __popc(any);
    if (m.fill >= 32) mirror_log_flush32(m, lane, debug);   // fill < 32 before, <= 63 after
  }
  q.n = 0;
}

__device__ __forceinline__ void knn_mirror_scan_tile(uint32_t t_addr, const float* ct, uint32_t col_base, float xn,
                                                     uint32_t q_row, MirrorLog& m, MirrorQueue& q, int debug_arg,
                                                     int chunk0 = 0, int chunk_step = 1) {
#ifdef VQACX_DEBUG
  const int debug = debug_arg;
#else
  constexpr int debug = 0;
  (void)debug_arg;
#endif
  const uint32_t lane = threadIdx.x & 31;
  const unsigned lt_mask = (1u << lane) - 1u;
#pragma unroll 1
  for (int chunk = chunk0; chunk < kGemmBN / 32; chunk += chunk_step) {
    uint32_t r[32];
    tmem_ld_32x32(t_addr + chunk * 32, r);
    tmem_ld_wait();
    // compare pass; alongside the hit mask, each of EIGHT interleaved chains (column index mod 8) remembers the value of
    // its last hit, so that a lane's last hit -- and its second-to-last, unless both fall in the same chain -- need no
    // run-time index into r[] afterwards
    uint32_t cmask = 0;
    float hv[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      const float4 c4 = *reinterpret_cast<const float4*>(ct + chunk * 32 + j);
      const float u0 = fmaf(-2.f, __uint_as_float(r[j + 0]), xn), u1 = fmaf(-2.f, __uint_as_float(r[j + 1]), xn);
      const float u2 = fmaf(-2.f, __uint_as_float(r[j + 2]), xn), u3 = fmaf(-2.f, __uint_as_float(r[j + 3]), xn);
      const bool h0 = u0 < c4.x, h1 = u1 < c4.y, h2 = u2 < c4.z, h3 = u3 < c4.w;
      cmask |= (h0 ? 1u : 0u) << (j + 0);
      cmask |= (h1 ? 1u : 0u) << (j + 1);
      cmask |= (h2 ? 1u : 0u) << (j + 2);
      cmask |= (h3 ? 1u : 0u) << (j + 3);
      hv[(j + 0) & 7] = h0 ? u0 : hv[(j + 0) & 7];
      hv[(j + 1) & 7] = h1 ? u1 : hv[(j + 1) & 7];
      hv[(j + 2) & 7] = h2 ? u2 : hv[(j + 2) & 7];
      hv[(j + 3) & 7] = h3 ? u3 : hv[(j + 3) & 7];
    }
    auto chain = [&](int j) -> float {   // hv[j & 7] without dynamic register indexing
      const float a = (j & 1) ? hv[1] : hv[0], b = (j & 1) ? hv[3] : hv[2];
      const float c = (j & 1) ? hv[5] : hv[4], d = (j & 1) ? hv[7] : hv[6];
      const float ab = (j & 2) ? b : a, cd = (j & 2) ? d : c;
      return (j & 4) ? cd : ab;
    };
    const uint32_t col0 = col_base + chunk * 32;
    if (debug & 32) cmask = 0u;   // timing experiment: compare pass only (results invalid)
    // the lane's last (two) hits go to its own queue: no warp-wide step in the common case
    bool over = false;
    if ((debug & 64) && q.n >= kMirrorQueueDepth) q.n = 0;       // timing experiment: queues never overflow (results invalid)
    if (cmask != 0u) {
      if (q.n < kMirrorQueueDepth) {
        const int j1 = 31 - __clz(cmask);
        cmask ^= 1u << j1;
        mirror_queue_put(q, chain(j1), col0 + j1);
        if (cmask != 0u && q.n < kMirrorQueueDepth) {
          const int j2 = 31 - __clz(cmask);
          if (((j1 ^ j2) & 7) != 0) {       // another chain: its last hit IS column j2
            cmask ^= 1u << j2;
            mirror_queue_put(q, chain(j2), col0 + j2);
          }
        }
      }
      over = cmask != 0u;      // more hits in this chunk than the fast path takes, or the queue was full
      if (debug & 64) over = false;
    }
    if (!__any_sync(0xffffffffu, over)) continue;
    // rare path (a lane's queue is full or it has several hits in this chunk): empty all queues, then the remaining
    // hits of this chunk one warp-wide iteration at a time
    mirror_queue_flush(q, m, q_row, lane, debug);
    for (;;) {
      const bool has = cmask != 0u;
      const unsigned any = __ballot_sync(0xffffffffu, has);
      if (any == 0u) break;
      if (has) {
        const int j = __ffs(cmask) - 1;
        cmask &= cmask - 1u;
        const uint64_t key = make_key(fmaf(-2.f, __uint_as_float(select32(r, j)), xn), q_row);
        const int slot = (m.head + m.fill + __popc(any & lt_mask)) & (kMirrorRing - 1);
        if (!(debug & 16))
          m.ring[slot] = make_uint4(static_cast<uint32_t>(key), static_cast<uint32_t>(key >> 32), col0 + j, 0u);
      }
      m.fill += __popc(any);
      if (m.fill >= 32) mirror_log_flush32(m, lane, debug);   // fill < 32 before the iteration, <= 63 after it
    }
  }
}

}  // namespace vqacx

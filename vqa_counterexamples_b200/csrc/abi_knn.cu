// C-ABI entry points for stage 1 (kNN). Host-side orchestration only; kernels live in knn_*.cu.
#include "../../include/vqacx.h"
#include "knn_internal.cuh"
#include "topk.cuh"
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

using namespace vqacx;

namespace vqacx {
static std::atomic<long long> g_launches{0};
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
static std::atomic<unsigned long long*> g_knn_prof{nullptr};
unsigned long long* knn_debug_profile_buffer() {
#ifdef VQACX_DEBUG
  return g_knn_prof.load(std::memory_order_relaxed);
#else
  return nullptr;
#endif
}

// ---- process-wide options: environment read once, then vqacx_set_option only ------------------------------------------
static std::atomic<int> g_opt[kOptCount];
static std::once_flag g_opt_once;
static int env_int(const char* name, int dflt) {
  const char* e = std::getenv(name);
  return (e && e[0]) ? std::atoi(e) : dflt;
}
static void load_options() {
  for (auto& o : g_opt) o.store(0, std::memory_order_relaxed);
  g_opt[kOptKnnLockstep] = env_int("VQACX_KNN_LOCKSTEP", kLockstepWindowDefault);
  g_opt[kOptKnnSymGroups] = env_int("VQACX_KNN_SYM_GROUPS", 7);
  g_opt[kOptKnnSymMirrorCap] = env_int("VQACX_KNN_SYM_MIRROR_CAP", 512);
  g_opt[kOptKnnSymLogCap] = env_int("VQACX_KNN_SYM_LOG_CAP", 0);
  g_opt[kOptKnnNoSymmetric] = env_int("VQACX_KNN_NO_SYMMETRIC", 0);
  g_opt[kOptKnnNoQsort] = env_int("VQACX_KNN_NO_QSORT", 0);
  g_opt[kOptNcxSerial] = env_int("VQACX_NCX_SERIAL", 0);
  g_opt[kOptNcxNoMutanFuse] = env_int("VQACX_NO_MUTAN_FUSE", 0);
  g_opt[kOptKnnSingleExchange] = env_int("VQACX_KNN_SINGLE_EXCHANGE", 0);
  g_opt[kOptKnnShardSymGroups] = env_int("VQACX_KNN_SHARD_SYM_GROUPS", 0);
  g_opt[kOptKnnSymLockstep] = env_int("VQACX_KNN_SYM_LOCKSTEP", 24);
  g_opt[kOptNcxNoPdl] = env_int("VQACX_NCX_NO_PDL", 0);
  g_opt[kOptGemmNoPair] = env_int("VQACX_GEMM_NO_PAIR", 0);
  g_opt[kOptKnnMirrorWarps] = env_int("VQACX_KNN_MIRROR_WARPS", 8);
#ifdef VQACX_DEBUG
  g_opt[kOptKnnSingleCta] = env_int("VQACX_KNN_SINGLE_CTA", 0);
  g_opt[kOptKnnOverlap] = env_int("VQACX_KNN_OVERLAP", 0);
  g_opt[kOptKnnWaveTrace] = env_int("VQACX_KNN_WAVE_TRACE", 0);
  g_opt[kOptKnnDebug] = env_int("VQACX_KNN_DEBUG", 0);
  g_opt[kOptGemmDebug] = env_int("VQACX_GEMM_DEBUG", 0);
#endif
}
int opt(int id) {
  std::call_once(g_opt_once, load_options);
  if (id <= 0 || id >= kOptCount) return 0;
#ifndef VQACX_DEBUG
  if (id >= kOptFirstDebug) return 0;
#endif
  return g_opt[id].load(std::memory_order_relaxed);
}
}  // namespace vqacx

namespace {

// Optional per-phase timing: events recorded on the caller's stream, read after the final synchronise.
struct PhaseTimer {
  cudaEvent_t ev[8];
  int n = 0;
  bool on;
  cudaStream_t s;
  PhaseTimer(bool enable, cudaStream_t st) : on(enable), s(st) {
    if (on)
      for (auto& e : ev)
        if (cudaEventCreate(&e) != cudaSuccess) { on = false; break; }
  }
  void mark() { mark_on(s); }
  void mark_on(cudaStream_t st) {
    if (on && n < 8) cudaEventRecord(ev[n++], st);
  }
  float ms(int a, int b) {
    float t = 0.f;
    if (on && b < n && cudaEventElapsedTime(&t, ev[a], ev[b]) != cudaSuccess) t = 0.f;
    return t;
  }
  ~PhaseTimer() {
    if (on)
      for (auto& e : ev) cudaEventDestroy(e);
  }
  // Deferred read-out: the phases of a sharded search that need no host synchronisation of their own hand their events
  // to a per-thread pending list; the phase that synchronises anyway (the last one) reads them all. host_dst[i] must
  // stay valid until then (the Python wrappers keep the arrays in the shard object).
  void defer(float* host_dst, int i, int a, int b);
};

struct PendingSpan { cudaEvent_t a, b; float* dst; };
thread_local std::vector<PendingSpan> t_pending_spans;
thread_local std::vector<cudaEvent_t> t_pending_events;

void PhaseTimer::defer(float* host_dst, int i, int a, int b) {
  if (!on || b >= n || !host_dst) return;
  t_pending_spans.push_back({ev[a], ev[b], host_dst + i});
}
void release_timer_events(PhaseTimer& pt) {   // ownership of the events moves to the pending list
  if (!pt.on) return;
  for (auto& e : pt.ev) t_pending_events.push_back(e);
  pt.on = false;
}
void drop_pending_timers() {
  for (cudaEvent_t e : t_pending_events) cudaEventDestroy(e);
  t_pending_events.clear();
  t_pending_spans.clear();
}
void resolve_pending_timers() {   // call after a stream synchronisation
  for (const PendingSpan& p : t_pending_spans) {
    float t = 0.f;
    if (cudaEventElapsedTime(&t, p.a, p.b) == cudaSuccess) *p.dst = t;
    else cudaGetLastError();
  }
  drop_pending_timers();
}

constexpr int kFbRowsMax = 4096;  // rows the exact fallback handles per round
constexpr int kFbSplit = 8;       // database splits of the fallback grid
constexpr int kMaxSplit = 16;     // max database splits of one query tile in the tensor path

int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
        n <= 0) {
      cudaGetLastError();
      n = 148;  // B200; only reached without a device (size queries in CPU-only tests)
    }
  }
  return n;
}

struct Plan {
  std::vector<KnnWorkItem> items;
  std::vector<int> cta_off;      // [n_ctas + 1]
  std::vector<int> tile_slot0;   // [n_q_tiles]
  std::vector<int> tile_nslots;  // [n_q_tiles]
  int n_ctas = 0, n_q_tiles = 0, n_db_tiles = 0, n_slots = 0;
};

// Remainder query tiles (those past the last full wave of G) are split s ways along the database so the
// last wave costs ~ceil(r*s/G)/s of a full sweep; pick the s minimising that (mild penalty per split).
int pick_split(int r, int G, int D) {
  int s_best = 1;
  if (r > 0) {
    double best = 1e30;
    const int s_max = std::min(kMaxSplit, std::max(1, D));
    for (int s = 1; s <= s_max; ++s) {
      const double cost = std::ceil(static_cast<double>(r) * s / G) / s + 0.01 * s;
      if (cost < best - 1e-12) { best = cost; s_best = s; }
    }
  }
  return s_best;
}

// Full waves: query tile t -> CTA t % G, whole database; then the split remainder, round-robin.
void make_plan(int64_t n_q, int64_t n_db, int G, int tile_rows, Plan& p) {
  p.n_q_tiles = static_cast<int>((n_q + tile_rows - 1) / tile_rows);
  p.n_db_tiles = static_cast<int>((n_db + kGemmBN - 1) / kGemmBN);
  const int T = p.n_q_tiles, D = p.n_db_tiles;
  const int full = (T / G) * G, r = T - full;
  const int s_best = pick_split(r, G, D);
  p.n_ctas = std::min(G, full > 0 ? G : r * s_best);
  std::vector<std::vector<KnnWorkItem>> per(p.n_ctas);
  p.tile_slot0.assign(T, 0);
  p.tile_nslots.assign(T, 0);
  int slot = 0;
  for (int t = 0; t < full; ++t) {
    p.tile_slot0[t] = slot;
    p.tile_nslots[t] = 1;
    per[t % G].push_back({t, 0, D, slot++});
  }
  int rr = 0;
  for (int t = full; t < T; ++t) {
    p.tile_slot0[t] = slot;
    p.tile_nslots[t] = s_best;
    for (int s = 0; s < s_best; ++s) {
      const int b = static_cast<int>(static_cast<int64_t>(D) * s / s_best);
      const int e = static_cast<int>(static_cast<int64_t>(D) * (s + 1) / s_best);
      per[rr % p.n_ctas].push_back({t, b, e, slot++});
      ++rr;
    }
  }
  p.n_slots = slot;
  p.cta_off.assign(p.n_ctas + 1, 0);
  p.items.clear();
  for (int c = 0; c < p.n_ctas; ++c) {
    p.cta_off[c] = static_cast<int>(p.items.size());
    p.items.insert(p.items.end(), per[c].begin(), per[c].end());
  }
  p.cta_off[p.n_ctas] = static_cast<int>(p.items.size());
}

int max_slots(int64_t n_q, int64_t n_db, int G, int tile_rows) {
  const int T = static_cast<int>((n_q + tile_rows - 1) / tile_rows);
  const int D = static_cast<int>((n_db + kGemmBN - 1) / kGemmBN);
  const int full = (T / G) * G, r = T - full;
  return full + r * pick_split(r, G, D) + 1;
}

// The CTA-pair kernel (knn_gemm_pair.cu) is the product path; the single-CTA kernel (knn_gemm.cu) serves a device with
// an odd SM count, and A/B measurements in -DVQACX_DEBUG builds (VQACX_OPT_KNN_SINGLE_CTA).
bool use_pair_kernel() { return opt(kOptKnnSingleCta) == 0 && sm_count() % 2 == 0; }

// Wave overlap (phase 0, pair kernel; OPT-IN with VQACX_KNN_OVERLAP=1): the tensor-bound candidate GEMM runs one wave
// of query tiles per launch on an internal high-priority stream, and the gather-bound finalize + re-rank of wave w run
// on the caller's stream while the GEMM is busy with wave w+1 (its CTAs leave ~12 KB of shared memory and half the
// registers of every SM free: one re-rank block per SM fits beside them).
// Measured on B200 (DESIGN.md "kNN: overlap experiment", VQACX_KNN_WAVE_TRACE=1): the co-resident re-rank blocks do run,
// but at ~1/12 of their stand-alone rate (8 warps per SM queueing behind the GEMM's TMA operand stream on the
// L2->SM path, which the GEMM already uses ~70 % of), the GEMM waves slow by ~12 %, and the step time is unchanged
// (60.7 vs 61.1 ms sustained). The serial three-launch schedule therefore stays the default.
constexpr int kMaxWaveGroups = 16;
struct OverlapCtx {
  cudaStream_t hi = nullptr;
  cudaEvent_t ev_in = nullptr;
  cudaEvent_t ev[kMaxWaveGroups + 1] = {};
  bool ok = false;
};

OverlapCtx* overlap_ctx() {
  static OverlapCtx ctx[16];
  static bool tried[16] = {};
  const bool disabled = opt(kOptKnnOverlap) == 0;   // -DVQACX_DEBUG builds only (a documented negative result)
  int dev = 0;
  if (disabled || cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) return nullptr;
  OverlapCtx& c = ctx[dev];
  if (!tried[dev]) {
    tried[dev] = true;
    int least = 0, greatest = 0;
    bool ok = cudaDeviceGetStreamPriorityRange(&least, &greatest) == cudaSuccess &&
              cudaStreamCreateWithPriority(&c.hi, cudaStreamNonBlocking, greatest) == cudaSuccess &&
              cudaEventCreateWithFlags(&c.ev_in, cudaEventDisableTiming) == cudaSuccess;
    for (int i = 0; ok && i <= kMaxWaveGroups; ++i)
      ok = cudaEventCreateWithFlags(&c.ev[i], cudaEventDisableTiming) == cudaSuccess;
    if (!ok) cudaGetLastError();
    if (ok) knn_support_prefer_max_shared_carveout();
    c.ok = ok;
  }
  return c.ok ? &c : nullptr;
}

struct Carver {
  uint8_t* base;
  size_t off = 0;
  explicit Carver(void* b) : base(static_cast<uint8_t*>(b)) {}
  template <class T>
  T* take(size_t n) {
    off = (off + 255) & ~static_cast<size_t>(255);
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += n * sizeof(T);
    return p;
  }
};

struct Workspace {
  KnnWorkItem* items;
  int *cta_off, *tile_slot0, *tile_nslots;
  uint64_t* cand;
  int* cnt;
  float* cand_val;
  int* cand_idx;
  int* counters;  // [0] flag_count, [1] unresolved count, [8..8+kStatCount) stats; [32..32+kLockstepMaxPairs) pair-kernel progress
  int *flag_rows, *flag_rows2;
  uint64_t* fb_cand;
  int* fb_cnt;
  float* fb_val;
  int* fb_idx;
  int *order_key, *order_hist, *order_perm;   // re-rank locality order (launch_knn_locality_order)
  KnnSymArgs sym;                              // symmetric self-join scratch (null unless sym_eligible)
  size_t bytes;
};

// ---- symmetric self-join (queries = database) -------------------------------------------------------------------------
// d(i, j) = d(j, i): the 256-row tiles are cut into an ODD number of groups; a group's diagonal block is computed in
// full (launch 0: it also warms every query's threshold up on a 1/groups sample of the database), and each unordered
// pair of groups {a, b} is computed ONCE (launch 1), by the group that sees the other within the next (groups-1)/2
// groups cyclically, with the column side fed through the mirror buffers (knn_internal.cuh KnnSymArgs). Tensor work:
// (groups + 1) / (2 groups) of the plain sweep (0.571 at 7 groups).
int sym_mirror_cap() {   // keys per query; VQACX_OPT_KNN_SYM_MIRROR_CAP shrinks it to exercise the overflow fallback in tests
  return std::max(8, std::min(opt(kOptKnnSymMirrorCap), 4096));
}
int sym_groups() {
  int g = opt(kOptKnnSymGroups);
  if (g < 3) g = 3;
  if (g > 15) g = 15;   // one upload event per group in the pipelined host self-join
  return g | 1;
}
bool sym_eligible(int64_t n_q, int64_t n_db) {
  if (opt(kOptKnnNoSymmetric)) return false;
  if (n_q != n_db || sm_count() % 2 != 0) return false;
  const int64_t T = (n_q + 2 * kGemmBM - 1) / (2 * kGemmBM);
  return T >= 4 * sym_groups() && T >= sm_count();   // enough tiles for the halved sweep to matter
}

struct SymPlan {
  std::vector<KnnWorkItem> items;   // launch 0 (diagonal blocks), then launch 1 (mirrored blocks); pair-major inside
  std::vector<int> pair_off;        // 2 x (G + 1), relative to the launch's first item
  std::vector<int> tile_slot0, tile_nslots;
  int item0[3] = {0, 0, 0};
  int pairs[2] = {0, 0};
  int T = 0, slots_per_tile = 0;
};

void make_sym_plan(int64_t n, int G, int groups, SymPlan& p) {
  const int T = static_cast<int>((n + 2 * kGemmBM - 1) / (2 * kGemmBM));
  const int half = (groups - 1) / 2;
  p.T = T;
  p.slots_per_tile = 1 + half;
  std::vector<int> gb(groups + 1);
  for (int g = 0; g <= groups; ++g) gb[g] = static_cast<int>(static_cast<int64_t>(T) * g / groups);
  p.tile_slot0.assign(T, 0);
  p.tile_nslots.assign(T, 1 + half);
  for (int t = 0; t < T; ++t) p.tile_slot0[t] = t * (1 + half);
  p.items.clear();
  p.pair_off.assign(2 * (G + 1), 0);
  for (int L = 0; L < 2; ++L) {
    // items of one database group are adjacent and dealt round-robin, so concurrent pairs sweep the same database
    // tiles at about the same time (they are shared through L2, as in the plain lock-step sweep)
    std::vector<KnnWorkItem> list;
    if (L == 0) {
      for (int g = 0; g < groups; ++g)
        for (int t = gb[g]; t < gb[g + 1]; ++t) list.push_back({t, gb[g], gb[g + 1], p.tile_slot0[t], 0});
    } else {
      for (int b = 0; b < groups; ++b)
        for (int h = 1; h <= half; ++h) {
          const int a = (b - h + groups) % groups;
          for (int t = gb[a]; t < gb[a + 1]; ++t) list.push_back({t, gb[b], gb[b + 1], p.tile_slot0[t] + h, kItemMirror});
        }
    }
    std::vector<std::vector<KnnWorkItem>> per(G);
    for (size_t i = 0; i < list.size(); ++i) per[i % G].push_back(list[i]);
    p.item0[L] = static_cast<int>(p.items.size());
    int off = 0, used = 0;
    for (int g = 0; g < G; ++g) {
      p.pair_off[static_cast<size_t>(L) * (G + 1) + g] = off;
      off += static_cast<int>(per[g].size());
      if (!per[g].empty()) used = g + 1;
      p.items.insert(p.items.end(), per[g].begin(), per[g].end());
    }
    p.pair_off[static_cast<size_t>(L) * (G + 1) + G] = off;
    p.pairs[L] = used;
  }
  p.item0[2] = static_cast<int>(p.items.size());
}

void carve_sym(Carver& c, int64_t n, KnnSymArgs& sym, int mcap_override = 0);

// ---- block-partitioned symmetric self-join over `world` ranks (vqacx_knn_selfjoin_shard) ------------------------------
// The single-GPU symmetric plan (make_sym_plan) cut across ranks: the launch-0 list (diagonal blocks) and the launch-1
// list (one block per unordered pair of groups, row side + mirror side) are each split into `world` contiguous chunks;
// rank r runs chunk r. Contiguous, so that a rank's concurrent CTA pairs sweep the same few database groups (shared
// through L2 under the lock-step throttle). Every (query, database row) pair is still covered exactly once over all
// ranks; a rank's candidate lists therefore cover, per query, a subset of the database blocks, and the per-rank exact
// top-k lists are merged across ranks like row shards are. More groups than on one GPU (15 by default): the items must
// be short enough for world x 74 pairs to balance, and the diagonal share (1 / groups of the matrix) is all a rank can
// do before the first threshold exchange.
int shard_sym_groups(int world) {
  int g = opt(kOptKnnShardSymGroups);
  if (g <= 0) g = world > 1 ? 15 : opt(kOptKnnSymGroups);
  if (g < 3) g = 3;
  if (g > 15) g = 15;
  return g | 1;
}

bool shard_sym_eligible(int64_t n, int world) {
  if (opt(kOptKnnNoSymmetric) || sm_count() % 2 != 0 || world < 1) return false;
  const int64_t T = (n + 2 * kGemmBM - 1) / (2 * kGemmBM);
  return T >= 4 * shard_sym_groups(world) && T >= sm_count();
}

struct ShardSymPlan {
  std::vector<KnnWorkItem> items;   // this rank's diagonal items, then its mirrored items; pair-major inside
  std::vector<int> pair_off;        // 2 x (G + 1)
  std::vector<int> tile_slot0, tile_nslots;
  int item0[3] = {0, 0, 0};
  int pairs[2] = {0, 0};
  int T = 0, slots_per_tile = 0;
};

void make_shard_sym_plan(int64_t n, int G, int groups, int rank, int world, ShardSymPlan& p) {
  const int T = static_cast<int>((n + 2 * kGemmBM - 1) / (2 * kGemmBM));
  const int half = (groups - 1) / 2;
  p.T = T;
  p.slots_per_tile = 1 + half;
  std::vector<int> gb(groups + 1);
  for (int g = 0; g <= groups; ++g) gb[g] = static_cast<int>(static_cast<int64_t>(T) * g / groups);
  p.tile_slot0.assign(T, 0);
  p.tile_nslots.assign(T, 1 + half);
  for (int t = 0; t < T; ++t) p.tile_slot0[t] = t * (1 + half);
  p.items.clear();
  p.pair_off.assign(2 * (G + 1), 0);
  for (int L = 0; L < 2; ++L) {
    std::vector<KnnWorkItem> list;
    if (L == 0) {
      for (int g = 0; g < groups; ++g)
        for (int t = gb[g]; t < gb[g + 1]; ++t) list.push_back({t, gb[g], gb[g + 1], p.tile_slot0[t], 0});
    } else {
      for (int b = 0; b < groups; ++b)
        for (int h = 1; h <= half; ++h) {
          const int a = (b - h + groups) % groups;
          for (int t = gb[a]; t < gb[a + 1]; ++t) list.push_back({t, gb[b], gb[b + 1], p.tile_slot0[t] + h, kItemMirror});
        }
    }
    const size_t lo = list.size() * static_cast<size_t>(rank) / world, hi = list.size() * static_cast<size_t>(rank + 1) / world;
    std::vector<std::vector<KnnWorkItem>> per(G);
    for (size_t i = lo; i < hi; ++i) per[(i - lo) % G].push_back(list[i]);
    p.item0[L] = static_cast<int>(p.items.size());
    int off = 0, used = 0;
    for (int g = 0; g < G; ++g) {
      p.pair_off[static_cast<size_t>(L) * (G + 1) + g] = off;
      off += static_cast<int>(per[g].size());
      if (!per[g].empty()) used = g + 1;
      p.items.insert(p.items.end(), per[g].begin(), per[g].end());
    }
    p.pair_off[static_cast<size_t>(L) * (G + 1) + G] = off;
    p.pairs[L] = used;
  }
  p.item0[2] = static_cast<int>(p.items.size());
}

// Re-rank in locality order when there are enough queries for reuse to matter. VQACX_OPT_KNN_NO_QSORT: list order (A/B).
bool use_locality_order(int64_t n_q) { return n_q >= 4096 && opt(kOptKnnNoQsort) == 0; }

int exact_split_for(int64_t n_q, int64_t n_db) {
  const int qt = static_cast<int>((n_q + 63) / 64);
  const int dt = static_cast<int>((n_db + 63) / 64);
  int s = (2 * sm_count() + qt - 1) / std::max(qt, 1);
  s = std::max(1, std::min(std::min(s, kMaxSplit), dt));
  return s;
}

// Symmetric self-join fed by a chunked upload (vqacx_knn_selfjoin_host): chunk = group. When group L has landed, launch
// 2L runs its diagonal block and launch 2L+1 every mirrored block between L and an earlier group (row side chosen by the
// same circulant rule as make_sym_plan, so every tile keeps 1 + (groups-1)/2 row-side slots and every query gets about
// half of its off-diagonal blocks through the mirror side).
struct SymChunkPlan {
  std::vector<KnnWorkItem> items;
  std::vector<int> pair_off;   // n_launch x (G + 1)
  std::vector<int> item0, pairs, gb;
  std::vector<int> tile_slot0, tile_nslots;
  int T = 0, groups = 0, n_launch = 0;
};

void make_sym_chunk_plan(int64_t n, int G, int groups, SymChunkPlan& p) {
  const int T = static_cast<int>((n + 2 * kGemmBM - 1) / (2 * kGemmBM));
  const int half = (groups - 1) / 2;
  p.T = T;
  p.groups = groups;
  p.n_launch = 2 * groups;
  p.gb.resize(groups + 1);
  for (int g = 0; g <= groups; ++g) p.gb[g] = static_cast<int>(static_cast<int64_t>(T) * g / groups);
  const std::vector<int>& gb = p.gb;
  p.tile_slot0.assign(T, 0);
  p.tile_nslots.assign(T, 1 + half);
  for (int t = 0; t < T; ++t) p.tile_slot0[t] = t * (1 + half);
  p.items.clear();
  p.pair_off.assign(static_cast<size_t>(p.n_launch) * (G + 1), 0);
  p.item0.assign(p.n_launch + 1, 0);
  p.pairs.assign(p.n_launch, 0);
  for (int L = 0; L < groups; ++L) {
    for (int m = 0; m < 2; ++m) {
      std::vector<KnnWorkItem> list;
      if (m == 0) {
        for (int t = gb[L]; t < gb[L + 1]; ++t) list.push_back({t, gb[L], gb[L + 1], p.tile_slot0[t], 0});
      } else {
        for (int pass = 0; pass < 2; ++pass)     // blocks whose COLUMN group is L first (they share database tiles)
          for (int a = 0; a < L; ++a) {
            const int h_fwd = L - a;                       // a sees L h_fwd groups ahead
            const bool rows_a = h_fwd <= half;
            if ((pass == 0) != rows_a) continue;
            const int rows = rows_a ? a : L, cols = rows_a ? L : a;
            const int h = rows_a ? h_fwd : groups - h_fwd;
            for (int t = gb[rows]; t < gb[rows + 1]; ++t)
              list.push_back({t, gb[cols], gb[cols + 1], p.tile_slot0[t] + h, kItemMirror});
          }
      }
      const int launch = 2 * L + m;
      std::vector<std::vector<KnnWorkItem>> per(G);
      for (size_t i = 0; i < list.size(); ++i) per[i % G].push_back(list[i]);
      p.item0[launch] = static_cast<int>(p.items.size());
      int off = 0, used = 0;
      for (int g = 0; g < G; ++g) {
        p.pair_off[static_cast<size_t>(launch) * (G + 1) + g] = off;
        off += static_cast<int>(per[g].size());
        if (!per[g].empty()) used = g + 1;
        p.items.insert(p.items.end(), per[g].begin(), per[g].end());
      }
      p.pair_off[static_cast<size_t>(launch) * (G + 1) + G] = off;
      p.pairs[launch] = used;
    }
  }
  p.item0[p.n_launch] = static_cast<int>(p.items.size());
}

void carve_sym(Carver& c, int64_t n, KnnSymArgs& sym, int mcap_override) {
  sym.gtau = c.take<uint32_t>(n);
  sym.mcnt = c.take<int>(static_cast<size_t>(n) * kMirrorCntStride);
  sym.mcap = mcap_override > 0 ? mcap_override : sym_mirror_cap();
  // append logs of the GEMM's mirror warps: as many entries in total as the mirror buffers hold, split evenly
  sym.n_regions = sm_count() * 8;   // one append log per mirror warp (knn_epilogue.cuh kMirrorWarps)
  sym.mirror_warps = opt(kOptKnnMirrorWarps) == 4 ? 4 : 8;
  // (VQACX_OPT_KNN_SYM_LOG_CAP = entries per log: shrinks them to exercise the log-overflow fallback in tests)
  const int lc = opt(kOptKnnSymLogCap);
  sym.log_cap = lc > 0 ? lc
                   : static_cast<int>((static_cast<size_t>(n) * std::min(std::max(sym.mcap, 512), 1024) + sym.n_regions - 1) /
                                      sym.n_regions);
  sym.log_len = c.take<int>(sym.n_regions);
  sym.log = c.take<uint4>(static_cast<size_t>(sym.n_regions) * sym.log_cap);
  sym.mbuf = c.take<uint64_t>(static_cast<size_t>(n) * sym.mcap);
}

Workspace carve(void* base, int64_t n_q, int64_t n_db) {
  Workspace w{};
  Carver c(base);
  const int T = static_cast<int>((n_q + kGemmBM - 1) / kGemmBM);  // >= tiles of either kernel
  // candidate rows: slots x tile rows, for whichever tensor kernel runs (pair: 256-row tiles over SMs/2 pairs)
  const size_t rows_single = static_cast<size_t>(max_slots(n_q, n_db, sm_count(), kGemmBM)) * kGemmBM;
  const size_t rows_pair = static_cast<size_t>(max_slots(n_q, n_db, std::max(1, sm_count() / 2), 2 * kGemmBM)) * 2 * kGemmBM;
  const int ms = std::max(max_slots(n_q, n_db, sm_count(), kGemmBM), max_slots(n_q, n_db, std::max(1, sm_count() / 2), 2 * kGemmBM));
  // exact primary path: 64-row tiles x uniform split
  const size_t exact_rows = static_cast<size_t>((n_q + 63) / 64) * 64 * exact_split_for(n_q, n_db);
  const bool sym = sym_eligible(n_q, n_db);
  const int T256 = static_cast<int>((n_q + 2 * kGemmBM - 1) / (2 * kGemmBM));
  const size_t sym_slots = sym ? static_cast<size_t>(T256) * (1 + (sym_groups() - 1) / 2) : 0;
  const size_t rows = std::max(std::max(std::max(rows_single, rows_pair), exact_rows), sym_slots * 2 * kGemmBM);
  w.items = c.take<KnnWorkItem>(std::max(static_cast<size_t>(ms), sym_slots + 1));
  w.cta_off = c.take<int>(2 * (sm_count() + 1));
  w.tile_slot0 = c.take<int>(T + 1);
  w.tile_nslots = c.take<int>(T + 1);
  w.cand = c.take<uint64_t>(rows * kTopkCap);
  w.cnt = c.take<int>(rows);
  w.cand_val = c.take<float>(static_cast<size_t>(n_q) * kTopkKeep);
  w.cand_idx = c.take<int>(static_cast<size_t>(n_q) * kTopkKeep);
  w.counters = c.take<int>(32 + kLockstepMaxPairs);   // [32..): lock-step progress array of the pair kernel
  w.flag_rows = c.take<int>(2 * n_q + 8);   // re-rank flags + mirror-overflow flags of the symmetric finalize
  w.flag_rows2 = c.take<int>(kFbRowsMax);
  w.fb_cand = c.take<uint64_t>(static_cast<size_t>(kFbRowsMax) * kFbSplit * kTopkCap);
  w.fb_cnt = c.take<int>(static_cast<size_t>(kFbRowsMax) * kFbSplit);
  w.fb_val = c.take<float>(static_cast<size_t>(kFbRowsMax) * kTopkKeep);
  w.fb_idx = c.take<int>(static_cast<size_t>(kFbRowsMax) * kTopkKeep);
  w.order_key = c.take<int>(n_q);
  w.order_hist = c.take<int>(n_db + 2);
  w.order_perm = c.take<int>(n_q);
  if (sym) {
    carve_sym(c, n_q, w.sym);
  }
  w.bytes = c.off + 256;
  return w;
}

// ---- pipelined self-join (vqacx_knn_selfjoin_host) -----------------------------------------------------------------
// The rows are cut into C chunks of whole 256-row tiles. Chunk c's host->device copy + prep run on an internal copy
// stream; as soon as it has landed, launch c of the candidate GEMM processes every (query tile, database chunk) block
// that became computable: (tiles of chunks <= c) x (chunk c) and (tiles of chunk c) x (chunks < c) -- the distance
// matrix is filled in L-shaped shells, so all of it except the last shell overlaps the upload. Each block is one work
// item with its own candidate slot (slot = tile * C + chunk: uniform C slots per tile for the finalize kernel).
struct ChunkPlan {
  int C = 1, T = 0, tiles_per_chunk = 0, G = 1;
  std::vector<KnnWorkItem> items;   // launch-major, then pair-major
  std::vector<int> pair_off;        // C x (G + 1), offsets relative to the launch's first item
  std::vector<int> launch_item0;    // [C + 1]
  std::vector<int> launch_pairs;    // [C] pairs with work
};

int selfjoin_chunks(int64_t n, int n_chunks) {
  const int T = static_cast<int>((n + 2 * kGemmBM - 1) / (2 * kGemmBM));
  return std::max(1, std::min(std::min(n_chunks, 16), T));
}

void make_chunk_plan(int64_t n, int n_chunks, int G, ChunkPlan& p) {
  p.T = static_cast<int>((n + 2 * kGemmBM - 1) / (2 * kGemmBM));   // 256-row tiles: query tile = database tile
  p.C = selfjoin_chunks(n, n_chunks);
  p.G = G;
  p.tiles_per_chunk = (p.T + p.C - 1) / p.C;
  p.C = (p.T + p.tiles_per_chunk - 1) / p.tiles_per_chunk;        // no empty trailing chunk
  const int C = p.C, T = p.T, tpc = p.tiles_per_chunk;
  p.items.clear();
  p.pair_off.assign(static_cast<size_t>(C) * (G + 1), 0);
  p.launch_item0.assign(C + 1, 0);
  p.launch_pairs.assign(C, 0);
  for (int L = 0; L < C; ++L) {
    // blocks of shell L, longest first, each to the least-loaded pair (lengths differ only in the last chunk)
    std::vector<KnnWorkItem> blocks;
    for (int t = 0; t < T; ++t) {
      const int ct = t / tpc;
      if (ct > L) break;
      for (int c = (ct == L ? 0 : L); c <= L; ++c) {
        const int b = c * tpc, e = std::min(T, (c + 1) * tpc);
        blocks.push_back({t, b, e, t * C + c});
      }
    }
    std::stable_sort(blocks.begin(), blocks.end(), [](const KnnWorkItem& a, const KnnWorkItem& b) {
      return (a.db_tile_end - a.db_tile_begin) > (b.db_tile_end - b.db_tile_begin);
    });
    std::vector<std::vector<KnnWorkItem>> per(G);
    std::vector<long long> load(G, 0);
    for (const KnnWorkItem& it : blocks) {
      int g = 0;
      for (int i = 1; i < G; ++i)
        if (load[i] < load[g]) g = i;
      per[g].push_back(it);
      load[g] += it.db_tile_end - it.db_tile_begin;
    }
    p.launch_item0[L] = static_cast<int>(p.items.size());
    int used = 0, off = 0;
    for (int g = 0; g < G; ++g) {
      p.pair_off[static_cast<size_t>(L) * (G + 1) + g] = off;
      off += static_cast<int>(per[g].size());
      if (!per[g].empty()) used = g + 1;
      p.items.insert(p.items.end(), per[g].begin(), per[g].end());
    }
    p.pair_off[static_cast<size_t>(L) * (G + 1) + G] = off;
    p.launch_pairs[L] = used;
  }
  p.launch_item0[C] = static_cast<int>(p.items.size());
}

struct SelfjoinWs {
  Workspace w;
  int* pair_off;
  size_t bytes;
};

SelfjoinWs carve_selfjoin(void* base, int64_t n, int n_chunks) {
  SelfjoinWs sw{};
  Workspace& w = sw.w;
  Carver c(base);
  const int T = static_cast<int>((n + 2 * kGemmBM - 1) / (2 * kGemmBM));
  const int C = selfjoin_chunks(n, n_chunks);
  const bool sym = sym_eligible(n, n);
  const size_t sym_slots = sym ? static_cast<size_t>(T) * (1 + (sym_groups() - 1) / 2) : 0;
  const size_t slots = std::max(static_cast<size_t>(T) * C, sym_slots);
  const size_t rows = slots * 2 * kGemmBM;
  w.items = c.take<KnnWorkItem>(slots + 1);
  sw.pair_off = c.take<int>(static_cast<size_t>(std::max(C, 2 * sym_groups())) * (sm_count() / 2 + 1) + 1);
  w.cta_off = sw.pair_off;
  w.tile_slot0 = w.tile_nslots = nullptr;
  if (sym) {
    w.tile_slot0 = c.take<int>(T + 1);
    w.tile_nslots = c.take<int>(T + 1);
  }
  w.cand = c.take<uint64_t>(rows * kTopkCap);
  w.cnt = c.take<int>(rows);
  w.cand_val = c.take<float>(static_cast<size_t>(n) * kTopkKeep);
  w.cand_idx = c.take<int>(static_cast<size_t>(n) * kTopkKeep);
  w.counters = c.take<int>(32 + kLockstepMaxPairs);   // [32..): lock-step progress array of the pair kernel
  w.flag_rows = c.take<int>(2 * n + 8);
  w.flag_rows2 = c.take<int>(kFbRowsMax);
  w.fb_cand = c.take<uint64_t>(static_cast<size_t>(kFbRowsMax) * kFbSplit * kTopkCap);
  w.fb_cnt = c.take<int>(static_cast<size_t>(kFbRowsMax) * kFbSplit);
  w.fb_val = c.take<float>(static_cast<size_t>(kFbRowsMax) * kTopkKeep);
  w.fb_idx = c.take<int>(static_cast<size_t>(kFbRowsMax) * kTopkKeep);
  w.order_key = c.take<int>(n);
  w.order_hist = c.take<int>(n + 2);
  w.order_perm = c.take<int>(n);
  if (sym) carve_sym(c, n, w.sym);
  sw.bytes = w.bytes = c.off + 256;
  return sw;
}

// Workspace of vqacx_knn_selfjoin_shard: like carve() for a self-join, with the slot count of the shard plan's groups.
Workspace carve_shard_sym(void* base, int64_t n, int world) {
  Workspace w{};
  Carver c(base);
  const int T = static_cast<int>((n + 2 * kGemmBM - 1) / (2 * kGemmBM));
  const int half = (shard_sym_groups(world) - 1) / 2;
  const size_t slots = static_cast<size_t>(T) * (1 + half);
  const size_t rows = slots * 2 * kGemmBM;
  w.items = c.take<KnnWorkItem>(slots + 1);
  w.cta_off = c.take<int>(2 * (sm_count() / 2 + 1) + 1);
  w.tile_slot0 = c.take<int>(T + 1);
  w.tile_nslots = c.take<int>(T + 1);
  w.cand = c.take<uint64_t>(rows * kTopkCap);
  w.cnt = c.take<int>(rows);
  w.cand_val = c.take<float>(static_cast<size_t>(n) * kTopkKeep);
  w.cand_idx = c.take<int>(static_cast<size_t>(n) * kTopkKeep);
  w.counters = c.take<int>(32 + kLockstepMaxPairs);
  w.flag_rows = c.take<int>(2 * n + 8);
  w.flag_rows2 = c.take<int>(kFbRowsMax);
  w.fb_cand = c.take<uint64_t>(static_cast<size_t>(kFbRowsMax) * kFbSplit * kTopkCap);
  w.fb_cnt = c.take<int>(static_cast<size_t>(kFbRowsMax) * kFbSplit);
  w.fb_val = c.take<float>(static_cast<size_t>(kFbRowsMax) * kTopkKeep);
  w.fb_idx = c.take<int>(static_cast<size_t>(kFbRowsMax) * kTopkKeep);
  w.order_key = c.take<int>(n);
  w.order_hist = c.take<int>(n + 2);
  w.order_perm = c.take<int>(n);
  // Mirror buffers: a rank that does not own a query's row-side blocks never sees that query's threshold tighten
  // after the exchange, so its column side keeps appending at the warm-up level (the k'-th best of a 1 / groups sample:
  // ~ n k' groups / n = a fixed FRACTION of all covered rows, several hundred keys per query at world = 2). Four times
  // the single-GPU capacity unless the option was lowered explicitly (tests of the overflow path).
  carve_sym(c, n, w.sym, opt(kOptKnnSymMirrorCap) < 512 ? 0 : 4 * sym_mirror_cap());
  w.bytes = c.off + 256;
  return w;
}

struct CopyCtx {
  cudaStream_t copy = nullptr;
  cudaEvent_t ev_in = nullptr, ev[16] = {};
  bool ok = false;
};

CopyCtx* copy_ctx() {
  static CopyCtx ctx[16];
  static bool tried[16] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) return nullptr;
  CopyCtx& c = ctx[dev];
  if (!tried[dev]) {
    tried[dev] = true;
    bool ok = cudaStreamCreateWithFlags(&c.copy, cudaStreamNonBlocking) == cudaSuccess &&
              cudaEventCreateWithFlags(&c.ev_in, cudaEventDisableTiming) == cudaSuccess;
    for (int i = 0; ok && i < 16; ++i) ok = cudaEventCreateWithFlags(&c.ev[i], cudaEventDisableTiming) == cudaSuccess;
    if (!ok) cudaGetLastError();
    c.ok = ok;
  }
  return c.ok ? &c : nullptr;
}

#define CU_TRY(expr)                         \
  do {                                       \
    cudaError_t e__ = (expr);                \
    if (e__ != cudaSuccess) return static_cast<int>(e__); \
  } while (0)

}  // namespace

extern "C" {

void vqacx_debug_set_knn_profile(void* device_buf) {
  vqacx::g_knn_prof.store(static_cast<unsigned long long*>(device_buf), std::memory_order_relaxed);
}

const char* vqacx_version(void) {
#ifdef VQACX_DEBUG
  return "vqacx-b200 0.2 (sm_100a, debug build: measurement options compiled in)";
#else
  return "vqacx-b200 0.2 (sm_100a)";
#endif
}

int vqacx_get_option(int option) {
  if (option <= 0 || option >= kOptCount) return VQACX_EINVAL;
  return opt(option);
}

int vqacx_set_option(int option, int value) {
  if (option <= 0 || option >= kOptCount) return VQACX_EINVAL;
#ifndef VQACX_DEBUG
  if (option >= kOptFirstDebug) return VQACX_EINVAL;   // measurement options do not exist in this build
#endif
  const int old = opt(option);
  g_opt[option].store(value, std::memory_order_relaxed);
  return old;
}

int64_t vqacx_launch_count(void) { return vqacx::g_launches.load(std::memory_order_relaxed); }

const char* vqacx_error_string(int code) {
  switch (code) {
    case VQACX_OK: return "ok";
    case VQACX_EINVAL: return "invalid argument (shape, null pointer or unsupported size)";
    case VQACX_EALIGN: return "pointer or pitch not 16-byte aligned";
    case VQACX_EWORKSPACE: return "workspace too small";
    case VQACX_ETMA: return "TMA descriptor creation failed";
    case VQACX_ERANGE: return "input not representable in fp16 (non-finite after rounding)";
    case VQACX_ESTATIC: return "problem exceeds the static scratch (VQACX_STATIC_SCRATCH_BYTES) of a *_static entry point";
    default: break;
  }
  if (code > 0) return cudaGetErrorString(static_cast<cudaError_t>(code));
  return "unknown vqacx error";
}

int vqacx_knn_prep(const float* x, int64_t n, int d, void* x_lp, float* norm_sq, uint32_t* max_norm_sq_bits,
                   int32_t* bad_count, void* stream) {
  if (n < 0 || d <= 0 || (d % 4) != 0) return VQACX_EINVAL;
  if (n == 0) return VQACX_OK;
  if (!x || !x_lp || !norm_sq || !max_norm_sq_bits || !bad_count) return VQACX_EINVAL;
  if ((reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(x_lp) & 15)) return VQACX_EALIGN;
  CU_TRY(launch_knn_prep(x, n, d, x_lp, norm_sq, max_norm_sq_bits, bad_count, static_cast<cudaStream_t>(stream)));
  return VQACX_OK;
}

size_t vqacx_knn_workspace_bytes(int64_t n_q, int64_t n_db, int d) {
  (void)d;
  if (n_q < 0 || n_db < 0) return 0;
  return carve(nullptr, n_q, n_db).bytes;
}

}  // extern "C"

// Exact fp32 path for the rows the proof flagged (w.flag_rows[0..flagged)), kFbRowsMax at a time; its own flags count
// "unresolved" rows (more than k' - k exact ties).
static int resolve_flagged(const Workspace& w, const float* q_f32, const float* db_f32, int64_t n_db, int d, int k,
                           int64_t db_row0, double* out_dist, int64_t* out_idx, int flagged, int* unresolved,
                           cudaStream_t s, int out_ld = 0) {
  const float eps_sq_rel = (d + 8) * 1.1920929e-7f;
  for (int off = 0; off < flagged; off += kFbRowsMax) {
    const int n = std::min(kFbRowsMax, flagged - off);
    // grid = (query tiles of 64) x (database splits): a handful of flagged rows must still fill the chip (6 rows on 8
    // CTAs swept the 1 GB matrix in 34 ms), so split further while the candidate scratch (kFbRowsMax / 64 * kFbSplit
    // tile slots) allows
    const int q_tiles = (n + 63) / 64;
    const int slots_max = kFbRowsMax / 64 * kFbSplit;
    int split = std::max(kFbSplit, (2 * sm_count() + q_tiles - 1) / q_tiles);
    split = std::max(1, std::min(std::min(split, slots_max / q_tiles), static_cast<int>((n_db + 63) / 64)));
    CU_TRY(cudaMemsetAsync(&w.counters[2], 0, sizeof(int), s));
    CU_TRY(launch_knn_exact(q_f32, w.flag_rows + off, nullptr, n, db_f32, static_cast<int>(n_db), d, split, w.fb_cand,
                            w.fb_cnt, s));
    CU_TRY(launch_knn_finalize(w.fb_cand, w.fb_cnt, nullptr, nullptr, split, 64, nullptr, n, w.fb_val, w.fb_idx, s));
    CU_TRY(launch_knn_rerank(q_f32, w.flag_rows + off, nullptr, n, db_f32, static_cast<int>(n_db), d, w.fb_val, w.fb_idx,
                             nullptr, nullptr, 0.f, 0.f, 0.f, eps_sq_rel, k, db_row0, out_dist,
                             reinterpret_cast<long long*>(out_idx), &w.counters[2], w.flag_rows2, &w.counters[8], nullptr,
                             s, out_ld));
    int u = 0;
    CU_TRY(cudaMemcpyAsync(&u, &w.counters[2], sizeof(int), cudaMemcpyDeviceToHost, s));
    CU_TRY(cudaStreamSynchronize(s));
    *unresolved += u;
  }
  return VQACX_OK;
}

// phase 0: the whole search. Row-sharded search (N > 1 GPUs) runs it in two halves around one small exchange:
// phase 1 = candidates + finalize + per-query upper bounds hi_io[n_q, k] of this shard's k best candidates;
// phase 2 = re-rank of the candidates phase 1 left in the workspace, pruned by hi_io[n_q] = agreed bound on the
// global k-th distance (vqacx_knn_kth_bound of the gathered phase-1 bounds).
static int knn_search_impl(const float* db_f32, const void* db_lp, const float* db_norm_sq,
                           const uint32_t* db_max_norm_sq_bits, int64_t n_db, int64_t db_row0, const float* q_f32,
                           const void* q_lp, const float* q_norm_sq, int64_t n_q, int d, int k, double* out_dist,
                           int64_t* out_idx, void* workspace, size_t workspace_bytes, uint32_t flags,
                           int32_t* stats_host, float* phase_ms_host, void* stream, int phase, float* hi_io) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (phase != 0 && !hi_io) return VQACX_EINVAL;
  if (phase_ms_host) std::memset(phase_ms_host, 0, sizeof(float) * VQACX_KNN_NPHASES);
  if (stats_host) std::memset(stats_host, 0, sizeof(int32_t) * VQACX_KNN_NSTATS);
  if (n_q < 0 || n_db <= 0 || d <= 0 || (d % 4) != 0 || k <= 0 || k > VQACX_KNN_MAX_K) return VQACX_EINVAL;
  if (n_db >= (1LL << 31) - 512 || n_q >= (1LL << 31) - 512) return VQACX_EINVAL;
  if (n_q == 0) return VQACX_OK;
  if (!db_f32 || !q_f32 || !workspace || (phase != 1 && (!out_dist || !out_idx))) return VQACX_EINVAL;
  if ((reinterpret_cast<uintptr_t>(db_f32) & 15) || (reinterpret_cast<uintptr_t>(q_f32) & 15) ||
      (reinterpret_cast<uintptr_t>(workspace) & 255))
    return VQACX_EALIGN;
  Workspace w = carve(workspace, n_q, n_db);
  if (w.bytes > workspace_bytes) return VQACX_EWORKSPACE;
  knn_pair_set_lockstep_buffer(w.counters + 32);

  const bool tensor_ok = !(flags & VQACX_KNN_FORCE_EXACT) && db_lp && q_lp && db_norm_sq && q_norm_sq &&
                         db_max_norm_sq_bits && (d % kGemmBK) == 0;
  CU_TRY(cudaMemsetAsync(w.counters, 0, 32 * sizeof(int), s));
  PhaseTimer pt(phase_ms_host != nullptr, s);

  const float* ext_bound = phase == 2 ? hi_io : nullptr;
  bool overlapped = false;
  cudaEvent_t tr_t0 = nullptr, tr_g[kMaxWaveGroups + 1], tr_r[kMaxWaveGroups + 1];
  int tr_n = 0;
  if (tensor_ok) {
    if ((reinterpret_cast<uintptr_t>(db_lp) & 15) || (reinterpret_cast<uintptr_t>(q_lp) & 15)) return VQACX_EALIGN;
    // error model (DESIGN.md "kNN exactness"): fp16 rounding moves each point by <= 2^-11 |x| (+ subnormal
    // absolute term), so |x~ - y~| is within eps_d (|x| + |y|max) of |x - y|; fp32 tensor accumulation and the
    // fp32 norm / epilogue roundings perturb the squared value by <= eps_sq_abs |x| |y|max.
    const float eps_d = 1.01f * 0.00048828125f;
    const float d_abs = 2.f * std::sqrt(static_cast<float>(d)) * 2.98023224e-8f;
    const float eps_sq_abs = 2.f * (d / 8 + 8) * 1.1920929e-7f;
    const float eps_sq_rel = 4.f * 1.1920929e-7f;
    const bool symmetric = phase == 0 && !(flags & VQACX_KNN_NO_SYMMETRIC) && use_pair_kernel() && q_lp == db_lp &&
                           q_f32 == db_f32 && q_norm_sq == db_norm_sq && w.sym.gtau != nullptr;
    if (symmetric) {
      const int G = sm_count() / 2;
      SymPlan p;
      make_sym_plan(n_q, G, sym_groups(), p);
      CUtensorMap tm_q, tm_db;
      if (make_tmap_2d(&tm_q, q_lp, 2, static_cast<uint64_t>(n_q), static_cast<uint64_t>(d),
                       static_cast<uint64_t>(d) * 2, kGemmBM) != 0 ||
          make_tmap_2d(&tm_db, db_lp, 2, static_cast<uint64_t>(n_db), static_cast<uint64_t>(d),
                       static_cast<uint64_t>(d) * 2, kGemmBN / 2) != 0)
        return VQACX_ETMA;
      CU_TRY(cudaMemcpyAsync(w.items, p.items.data(), p.items.size() * sizeof(KnnWorkItem), cudaMemcpyHostToDevice, s));
      CU_TRY(cudaMemcpyAsync(w.cta_off, p.pair_off.data(), p.pair_off.size() * sizeof(int), cudaMemcpyHostToDevice, s));
      CU_TRY(cudaMemcpyAsync(w.tile_slot0, p.tile_slot0.data(), p.tile_slot0.size() * sizeof(int),
                             cudaMemcpyHostToDevice, s));
      CU_TRY(cudaMemcpyAsync(w.tile_nslots, p.tile_nslots.data(), p.tile_nslots.size() * sizeof(int),
                             cudaMemcpyHostToDevice, s));
      pt.mark();
      CU_TRY(launch_knn_sym_reset(w.sym, static_cast<int>(n_q), s));
      const bool trace = opt(kOptKnnWaveTrace) != 0;   // debug builds: per-launch times on stderr
      cudaEvent_t tr[3] = {};
      if (trace) {
        for (auto& e : tr) cudaEventCreate(&e);
        cudaEventRecord(tr[0], s);
      }
      for (int L = 0; L < 2; ++L) {
        KnnSymArgs sa = w.sym;
        sa.mirror_launch = L;   // launch 0: diagonal blocks (row side only); launch 1: mirrored blocks
        CU_TRY(launch_knn_gemm_topk_pair_sym(tm_q, tm_db, w.items + p.item0[L], w.cta_off + L * (G + 1), p.pairs[L],
                                             db_norm_sq, static_cast<int>(n_db), d / kGemmBK, w.cand, w.cnt, sa, s));
        if (trace) cudaEventRecord(tr[L + 1], s);
      }
      if (trace) {
        cudaEventSynchronize(tr[2]);
        float a = 0.f, b = 0.f;
        cudaEventElapsedTime(&a, tr[0], tr[1]);
        cudaEventElapsedTime(&b, tr[1], tr[2]);
        std::fprintf(stderr, "[vqacx symmetric] diagonal launch %.2f ms (%d items), mirrored launch %.2f ms (%d items)\n", a,
                     p.item0[1] - p.item0[0], b, p.item0[2] - p.item0[1]);
        for (auto& e : tr) cudaEventDestroy(e);
      }
      CU_TRY(launch_knn_mirror_scatter(w.sym, &w.counters[3], s));
      pt.mark();
      CU_TRY(launch_knn_finalize_sym(w.cand, w.cnt, w.tile_slot0, w.tile_nslots, 2 * kGemmBM, static_cast<int>(n_q),
                                     w.sym, w.cand_val, w.cand_idx, &w.counters[0], w.flag_rows, &w.counters[8], s));
      pt.mark();
    } else if (phase != 2) {
      const bool pair = use_pair_kernel();
      const int tile_rows = pair ? 2 * kGemmBM : kGemmBM;
      Plan p;
      make_plan(n_q, n_db, pair ? sm_count() / 2 : sm_count(), tile_rows, p);
      CUtensorMap tm_q, tm_db;
      if (make_tmap_2d(&tm_q, q_lp, 2, static_cast<uint64_t>(n_q), static_cast<uint64_t>(d),
                       static_cast<uint64_t>(d) * 2, kGemmBM) != 0)
        return VQACX_ETMA;
      // pair kernel: each CTA stages half (128 rows) of a 256-row database tile
      if (make_tmap_2d(&tm_db, db_lp, 2, static_cast<uint64_t>(n_db), static_cast<uint64_t>(d),
                       static_cast<uint64_t>(d) * 2, pair ? kGemmBN / 2 : kGemmBN) != 0)
        return VQACX_ETMA;
      CU_TRY(cudaMemcpyAsync(w.items, p.items.data(), p.items.size() * sizeof(KnnWorkItem), cudaMemcpyHostToDevice, s));
      CU_TRY(cudaMemcpyAsync(w.cta_off, p.cta_off.data(), p.cta_off.size() * sizeof(int), cudaMemcpyHostToDevice, s));
      CU_TRY(cudaMemcpyAsync(w.tile_slot0, p.tile_slot0.data(), p.tile_slot0.size() * sizeof(int),
                             cudaMemcpyHostToDevice, s));
      CU_TRY(cudaMemcpyAsync(w.tile_nslots, p.tile_nslots.data(), p.tile_nslots.size() * sizeof(int),
                             cudaMemcpyHostToDevice, s));
      // the plan vectors are pageable: the copies above are staged before returning, so `p` may die here
      pt.mark();
      const int G = pair ? sm_count() / 2 : sm_count();
      const int n_waves = p.n_q_tiles / G;   // full waves: tile t is item t / G of CTA pair t % G
      OverlapCtx* oc = (pair && phase == 0 && n_waves >= 1 && p.n_q_tiles > G) ? overlap_ctx() : nullptr;
      if (oc) {
        const int wpg = (n_waves + kMaxWaveGroups - 1) / kMaxWaveGroups;   // waves per launch
        const int64_t wave_rows = static_cast<int64_t>(G) * tile_rows;
        CU_TRY(cudaEventRecord(oc->ev_in, s));
        CU_TRY(cudaStreamWaitEvent(oc->hi, oc->ev_in, 0));
        const bool trace = opt(kOptKnnWaveTrace) != 0;   // debug builds: per-wave completion times on stderr
        if (trace) {
          cudaEventCreate(&tr_t0);
          cudaEventRecord(tr_t0, s);
        }
        int g = 0;
        for (int w0 = 0; w0 <= n_waves; w0 += wpg, ++g) {
          const bool rest = w0 >= n_waves;     // last launch: the split remainder tiles (items n_waves.. of each pair)
          if (rest) w0 = n_waves;
          const int w1 = rest ? n_waves : std::min(n_waves, w0 + wpg);
          const int row0 = static_cast<int>(std::min<int64_t>(n_q, w0 * wave_rows));
          const int row1 = rest ? static_cast<int>(n_q) : static_cast<int>(std::min<int64_t>(n_q, w1 * wave_rows));
          if (row1 > row0) {
            CU_TRY(launch_knn_gemm_topk_pair_range(tm_q, tm_db, w.items, w.cta_off, p.n_ctas, db_norm_sq,
                                                   static_cast<int>(n_db), d / kGemmBK, w.cand, w.cnt, w0,
                                                   rest ? 0x7fffffff : w1 - w0, oc->hi));
            CU_TRY(cudaEventRecord(oc->ev[g], oc->hi));
            if (trace) {
              cudaEventCreate(&tr_g[tr_n]);
              cudaEventRecord(tr_g[tr_n], oc->hi);
            }
            CU_TRY(cudaStreamWaitEvent(s, oc->ev[g], 0));
            if (rest) { pt.mark_on(oc->hi); pt.mark_on(oc->hi); }   // GEMM end; finalize is inside the overlap
            CU_TRY(launch_knn_finalize_range(w.cand, w.cnt, w.tile_slot0, w.tile_nslots, 0, tile_rows, nullptr, row0,
                                             row1, w.cand_val, w.cand_idx, s));
            CU_TRY(launch_knn_rerank_range(q_f32, nullptr, nullptr, row0, row1, db_f32, static_cast<int>(n_db), d,
                                           w.cand_val, w.cand_idx, q_norm_sq, db_max_norm_sq_bits, eps_d, d_abs,
                                           eps_sq_abs, eps_sq_rel, k, db_row0, out_dist,
                                           reinterpret_cast<long long*>(out_idx), &w.counters[0], w.flag_rows,
                                           &w.counters[8], ext_bound, s));
            if (trace) {
              cudaEventCreate(&tr_r[tr_n]);
              cudaEventRecord(tr_r[tr_n], s);
              ++tr_n;
            }
          } else if (rest) {
            pt.mark_on(oc->hi); pt.mark_on(oc->hi);
          }
          if (rest) break;
        }
        overlapped = true;
      } else if (pair) {
        CU_TRY(launch_knn_gemm_topk_pair(tm_q, tm_db, w.items, w.cta_off, p.n_ctas, db_norm_sq, static_cast<int>(n_db),
                                         d / kGemmBK, w.cand, w.cnt, s));
      } else {
        CU_TRY(launch_knn_gemm_topk(tm_q, tm_db, w.items, w.cta_off, p.n_ctas, db_norm_sq, static_cast<int>(n_db),
                                    d / kGemmBK, w.cand, w.cnt, s));
      }
      if (!overlapped) {
        pt.mark();
        CU_TRY(launch_knn_finalize(w.cand, w.cnt, w.tile_slot0, w.tile_nslots, 0, tile_rows, nullptr,
                                   static_cast<int>(n_q), w.cand_val, w.cand_idx, s));
        pt.mark();
      }
    } else {
      pt.mark(); pt.mark(); pt.mark();
    }
    if (overlapped) {
      // finalize + re-rank were issued per wave above
    } else if (phase == 1) {
      CU_TRY(launch_knn_local_hi(q_f32, static_cast<int>(n_q), d, w.cand_val, w.cand_idx, q_norm_sq, db_max_norm_sq_bits,
                                 eps_d, d_abs, eps_sq_abs, eps_sq_rel, k, hi_io, s));
    } else {
      const bool ordered = use_locality_order(n_q);
      if (ordered)
        CU_TRY(launch_knn_locality_order(w.cand_idx, static_cast<int>(n_q), static_cast<int>(n_db), w.order_key,
                                         w.order_hist, w.order_perm, s));
      CU_TRY(launch_knn_rerank_range(q_f32, ordered ? w.order_perm : nullptr, nullptr, 0, static_cast<int>(n_q), db_f32,
                                     static_cast<int>(n_db), d, w.cand_val, w.cand_idx, q_norm_sq, db_max_norm_sq_bits,
                                     eps_d, d_abs, eps_sq_abs, eps_sq_rel, k, db_row0, out_dist,
                                     reinterpret_cast<long long*>(out_idx), &w.counters[0], w.flag_rows, &w.counters[8],
                                     ext_bound, s, ordered));
    }
  } else {
    const int split = exact_split_for(n_q, n_db);
    const float eps_sq_rel = (d + 8) * 1.1920929e-7f;
    if (phase != 2) {
      pt.mark();
      CU_TRY(launch_knn_exact(q_f32, nullptr, nullptr, static_cast<int>(n_q), db_f32, static_cast<int>(n_db), d, split,
                              w.cand, w.cnt, s));
      pt.mark();
      CU_TRY(launch_knn_finalize(w.cand, w.cnt, nullptr, nullptr, split, 64, nullptr, static_cast<int>(n_q),
                                 w.cand_val, w.cand_idx, s));
      pt.mark();
    } else {
      pt.mark(); pt.mark(); pt.mark();
    }
    if (phase == 1) {
      CU_TRY(launch_knn_local_hi(q_f32, static_cast<int>(n_q), d, w.cand_val, w.cand_idx, nullptr, nullptr, 0.f, 0.f,
                                 0.f, eps_sq_rel, k, hi_io, s));
    } else {
      CU_TRY(launch_knn_rerank(q_f32, nullptr, nullptr, static_cast<int>(n_q), db_f32, static_cast<int>(n_db), d,
                               w.cand_val, w.cand_idx, nullptr, nullptr, 0.f, 0.f, 0.f, eps_sq_rel, k, db_row0, out_dist,
                               reinterpret_cast<long long*>(out_idx), &w.counters[1], w.flag_rows, &w.counters[8],
                               ext_bound, s));
    }
  }

  pt.mark();  // end of re-rank
  if (phase == 1) {
    drop_pending_timers();
    pt.defer(phase_ms_host, 0, 0, 1);   // read by phase 2, which synchronises anyway (no host sync in phase 1)
    pt.defer(phase_ms_host, 1, 1, 2);
    release_timer_events(pt);
    if (stats_host) stats_host[kStatUsedTensorPath] = tensor_ok ? 1 : 0;
    return VQACX_OK;
  }
  int host_counters[32];
  CU_TRY(cudaMemcpyAsync(host_counters, w.counters, sizeof(host_counters), cudaMemcpyDeviceToHost, s));
  CU_TRY(cudaStreamSynchronize(s));
  if (phase == 2) resolve_pending_timers();
  if (tr_t0) {
    for (int i = 0; i < tr_n; ++i) {
      float a = 0.f, b = 0.f;
      cudaEventElapsedTime(&a, tr_t0, tr_g[i]);
      cudaEventElapsedTime(&b, tr_t0, tr_r[i]);
      std::fprintf(stderr, "[vqacx wave %d] gemm done %.2f ms, re-rank done %.2f ms\n", i, a, b);
      cudaEventDestroy(tr_g[i]);
      cudaEventDestroy(tr_r[i]);
    }
    cudaEventDestroy(tr_t0);
  }
  if (host_counters[3] != 0 && !(flags & VQACX_KNN_NO_SYMMETRIC))   // a mirror log overflowed: keys were dropped
    return knn_search_impl(db_f32, db_lp, db_norm_sq, db_max_norm_sq_bits, n_db, db_row0, q_f32, q_lp, q_norm_sq, n_q, d, k,
                           out_dist, out_idx, workspace, workspace_bytes, flags | VQACX_KNN_NO_SYMMETRIC, stats_host,
                           phase_ms_host, stream, phase, hi_io);
  int flagged = tensor_ok ? host_counters[0] : 0;
  int unresolved = tensor_ok ? 0 : host_counters[1];
  int violations = host_counters[8 + kStatBoundViolations];

  if (flagged > 0) {
    const int st = resolve_flagged(w, q_f32, db_f32, n_db, d, k, db_row0, out_dist, out_idx, flagged, &unresolved, s);
    if (st != VQACX_OK) return st;
  }
  if (phase_ms_host) {
    pt.mark();
    CU_TRY(cudaStreamSynchronize(s));
    phase_ms_host[0] = pt.ms(0, 1);
    phase_ms_host[1] = pt.ms(1, 2);
    phase_ms_host[2] = pt.ms(2, 3);
    phase_ms_host[3] = pt.ms(3, 4);
  }
  if (stats_host) {
    stats_host[kStatFlaggedRows] = flagged;
    stats_host[kStatUnresolvedRows] = unresolved;
    stats_host[kStatBoundViolations] = violations;
    stats_host[kStatUsedTensorPath] = tensor_ok ? 1 : 0;
    stats_host[kStatRerankEvaluated] = host_counters[8 + kStatRerankEvaluated];
    stats_host[kStatMirrorKeys] = host_counters[8 + kStatMirrorKeys];
  }
  return VQACX_OK;
}

extern "C" {

size_t vqacx_knn_selfjoin_workspace_bytes(int64_t n, int d, int n_chunks) {
  (void)d;
  if (n < 0) return 0;
  return carve_selfjoin(nullptr, n, n_chunks).bytes;
}

int vqacx_knn_selfjoin_host(const float* x_host, int64_t n, int d, int k, int n_chunks, float* x_dev, void* lp_dev,
                            float* norm_sq_dev, uint32_t* max_norm_sq_bits_dev, int32_t* bad_count_dev,
                            double* out_dist, int64_t* out_idx, void* workspace, size_t workspace_bytes,
                            int32_t* stats_host, void* stream) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (stats_host) std::memset(stats_host, 0, sizeof(int32_t) * VQACX_KNN_NSTATS);
  if (n < 0 || d <= 0 || (d % kGemmBK) != 0 || k <= 0 || k > VQACX_KNN_MAX_K || n_chunks < 1) return VQACX_EINVAL;
  if (n >= (1LL << 31) - 512) return VQACX_EINVAL;
  if (n == 0) return VQACX_OK;
  if (!x_host || !x_dev || !lp_dev || !norm_sq_dev || !max_norm_sq_bits_dev || !bad_count_dev || !out_dist || !out_idx ||
      !workspace)
    return VQACX_EINVAL;
  if ((reinterpret_cast<uintptr_t>(x_dev) & 15) || (reinterpret_cast<uintptr_t>(lp_dev) & 15) ||
      (reinterpret_cast<uintptr_t>(workspace) & 255))
    return VQACX_EALIGN;
  if (!use_pair_kernel()) return VQACX_EINVAL;   // the shell schedule is written for the CTA-pair kernel's 256-row tiles
  SelfjoinWs sw = carve_selfjoin(workspace, n, n_chunks);
  if (sw.bytes > workspace_bytes) return VQACX_EWORKSPACE;
  const Workspace& w = sw.w;
  knn_pair_set_lockstep_buffer(w.counters + 32);
  CopyCtx* cc = copy_ctx();
  if (!cc) return static_cast<int>(cudaErrorUnknown);

  const int G = sm_count() / 2;
  CUtensorMap tm_q, tm_db;
  if (make_tmap_2d(&tm_q, lp_dev, 2, static_cast<uint64_t>(n), static_cast<uint64_t>(d), static_cast<uint64_t>(d) * 2,
                   kGemmBM) != 0 ||
      make_tmap_2d(&tm_db, lp_dev, 2, static_cast<uint64_t>(n), static_cast<uint64_t>(d), static_cast<uint64_t>(d) * 2,
                   kGemmBN / 2) != 0)
    return VQACX_ETMA;
  CU_TRY(cudaMemsetAsync(w.counters, 0, 32 * sizeof(int), s));
  CU_TRY(cudaMemsetAsync(max_norm_sq_bits_dev, 0, sizeof(uint32_t), s));
  CU_TRY(cudaMemsetAsync(bad_count_dev, 0, sizeof(int32_t), s));
  const bool trace = opt(kOptKnnWaveTrace) != 0;   // debug builds: shell / finalize / re-rank end times on stderr
  cudaEvent_t tr[24] = {};
  int tr_n = 0, n_shells = 0;
  auto tmark = [&]() {
    if (trace && tr_n < 24 && cudaEventCreate(&tr[tr_n]) == cudaSuccess) cudaEventRecord(tr[tr_n++], s);
  };
  const bool symmetric = w.sym.gtau != nullptr;   // carve_selfjoin: sym_eligible(n, n)
  if (symmetric) {
    // Symmetric sweep (see make_sym_plan): upload chunk = group; n_chunks is not used. As soon as group L has landed,
    // its diagonal block runs, then every mirrored block between L and the groups before it.
    SymChunkPlan p;
    make_sym_chunk_plan(n, G, sym_groups(), p);
    n_shells = p.groups;
    CU_TRY(cudaMemcpyAsync(w.items, p.items.data(), p.items.size() * sizeof(KnnWorkItem), cudaMemcpyHostToDevice, s));
    CU_TRY(cudaMemcpyAsync(sw.pair_off, p.pair_off.data(), p.pair_off.size() * sizeof(int), cudaMemcpyHostToDevice, s));
    CU_TRY(cudaMemcpyAsync(w.tile_slot0, p.tile_slot0.data(), p.tile_slot0.size() * sizeof(int), cudaMemcpyHostToDevice, s));
    CU_TRY(cudaMemcpyAsync(w.tile_nslots, p.tile_nslots.data(), p.tile_nslots.size() * sizeof(int),
                           cudaMemcpyHostToDevice, s));
    CU_TRY(launch_knn_sym_reset(w.sym, static_cast<int>(n), s));
    CU_TRY(cudaEventRecord(cc->ev_in, s));
    CU_TRY(cudaStreamWaitEvent(cc->copy, cc->ev_in, 0));
    tmark();
    for (int L = 0; L < p.groups; ++L) {
      const int64_t r0 = static_cast<int64_t>(p.gb[L]) * 2 * kGemmBM;
      const int64_t rows = std::min<int64_t>(n, static_cast<int64_t>(p.gb[L + 1]) * 2 * kGemmBM) - r0;
      CU_TRY(cudaMemcpyAsync(x_dev + r0 * d, x_host + r0 * d, static_cast<size_t>(rows) * d * sizeof(float),
                             cudaMemcpyHostToDevice, cc->copy));
      CU_TRY(launch_knn_prep(x_dev + r0 * d, rows, d, static_cast<uint8_t*>(lp_dev) + static_cast<size_t>(r0) * d * 2,
                             norm_sq_dev + r0, max_norm_sq_bits_dev, bad_count_dev, cc->copy));
      CU_TRY(cudaEventRecord(cc->ev[L], cc->copy));
      CU_TRY(cudaStreamWaitEvent(s, cc->ev[L], 0));
      for (int m = 0; m < 2; ++m) {
        const int launch = 2 * L + m;
        KnnSymArgs sa = w.sym;
        sa.mirror_launch = m;
        CU_TRY(launch_knn_gemm_topk_pair_sym(tm_q, tm_db, w.items + p.item0[launch],
                                             sw.pair_off + static_cast<size_t>(launch) * (G + 1), p.pairs[launch],
                                             norm_sq_dev, static_cast<int>(n), d / kGemmBK, w.cand, w.cnt, sa, s));
      }
      tmark();
    }
    CU_TRY(launch_knn_mirror_scatter(w.sym, &w.counters[3], s));
    CU_TRY(launch_knn_finalize_sym(w.cand, w.cnt, w.tile_slot0, w.tile_nslots, 2 * kGemmBM, static_cast<int>(n), w.sym,
                                   w.cand_val, w.cand_idx, &w.counters[0], w.flag_rows, &w.counters[8], s));
    tmark();
  } else {
  ChunkPlan p;
  make_chunk_plan(n, n_chunks, G, p);
  const int C = p.C;
  n_shells = C;
  const int64_t chunk_rows = static_cast<int64_t>(p.tiles_per_chunk) * 2 * kGemmBM;
  CU_TRY(cudaMemcpyAsync(w.items, p.items.data(), p.items.size() * sizeof(KnnWorkItem), cudaMemcpyHostToDevice, s));
  CU_TRY(cudaMemcpyAsync(sw.pair_off, p.pair_off.data(), p.pair_off.size() * sizeof(int), cudaMemcpyHostToDevice, s));
  CU_TRY(cudaEventRecord(cc->ev_in, s));
  CU_TRY(cudaStreamWaitEvent(cc->copy, cc->ev_in, 0));
  tmark();
  // Issue order copy 0, shell 0, copy 1, shell 1, ...: with pinned memory every call is asynchronous and the copy
  // stream runs ahead on its own; with pageable memory cudaMemcpyAsync blocks the host while it stages a chunk, and
  // the shell queued just before keeps the GPU busy meanwhile.
  for (int L = 0; L < C; ++L) {
    const int64_t r0 = L * chunk_rows, rows = std::min<int64_t>(n, r0 + chunk_rows) - r0;
    CU_TRY(cudaMemcpyAsync(x_dev + r0 * d, x_host + r0 * d, static_cast<size_t>(rows) * d * sizeof(float),
                           cudaMemcpyHostToDevice, cc->copy));
    CU_TRY(launch_knn_prep(x_dev + r0 * d, rows, d, static_cast<uint8_t*>(lp_dev) + static_cast<size_t>(r0) * d * 2,
                           norm_sq_dev + r0, max_norm_sq_bits_dev, bad_count_dev, cc->copy));
    CU_TRY(cudaEventRecord(cc->ev[L], cc->copy));
    CU_TRY(cudaStreamWaitEvent(s, cc->ev[L], 0));
    CU_TRY(launch_knn_gemm_topk_pair(tm_q, tm_db, w.items + p.launch_item0[L], sw.pair_off + static_cast<size_t>(L) * (G + 1),
                                     p.launch_pairs[L], norm_sq_dev, static_cast<int>(n), d / kGemmBK, w.cand, w.cnt, s));
    tmark();
  }
  CU_TRY(launch_knn_finalize(w.cand, w.cnt, nullptr, nullptr, C, 2 * kGemmBM, nullptr, static_cast<int>(n), w.cand_val,
                             w.cand_idx, s));
  tmark();
  }
  // same error model as vqacx_knn_search (DESIGN.md "kNN exactness")
  const float eps_d = 1.01f * 0.00048828125f;
  const float d_abs = 2.f * std::sqrt(static_cast<float>(d)) * 2.98023224e-8f;
  const float eps_sq_abs = 2.f * (d / 8 + 8) * 1.1920929e-7f;
  const float eps_sq_rel = 4.f * 1.1920929e-7f;
  const bool ordered = use_locality_order(n);
  if (ordered)
    CU_TRY(launch_knn_locality_order(w.cand_idx, static_cast<int>(n), static_cast<int>(n), w.order_key, w.order_hist,
                                     w.order_perm, s));
  CU_TRY(launch_knn_rerank_range(x_dev, ordered ? w.order_perm : nullptr, nullptr, 0, static_cast<int>(n), x_dev,
                                 static_cast<int>(n), d, w.cand_val, w.cand_idx, norm_sq_dev, max_norm_sq_bits_dev, eps_d,
                                 d_abs, eps_sq_abs, eps_sq_rel, k, 0, out_dist, reinterpret_cast<long long*>(out_idx),
                                 &w.counters[0], w.flag_rows, &w.counters[8], nullptr, s, ordered));
  int host_counters[32];
  int bad = 0;
  CU_TRY(cudaMemcpyAsync(host_counters, w.counters, sizeof(host_counters), cudaMemcpyDeviceToHost, s));
  tmark();
  CU_TRY(cudaMemcpyAsync(&bad, bad_count_dev, sizeof(int), cudaMemcpyDeviceToHost, s));
  CU_TRY(cudaStreamSynchronize(s));
  for (int i = 1; i < tr_n; ++i) {
    float t = 0.f;
    cudaEventElapsedTime(&t, tr[0], tr[i]);
    std::fprintf(stderr, "[vqacx selfjoin] %s %d done at %.2f ms\n",
                 i <= n_shells ? "shell" : (i == n_shells + 1 ? "finalize" : "re-rank"), i <= n_shells ? i - 1 : 0, t);
  }
  for (int i = 0; i < tr_n; ++i) cudaEventDestroy(tr[i]);
  if (bad != 0) return VQACX_ERANGE;
  if (host_counters[3] != 0)   // a mirror log overflowed (keys dropped): redo the search with the plain sweep
    return knn_search_impl(x_dev, lp_dev, norm_sq_dev, max_norm_sq_bits_dev, n, 0, x_dev, lp_dev, norm_sq_dev, n, d, k,
                           out_dist, out_idx, workspace, workspace_bytes, VQACX_KNN_NO_SYMMETRIC, stats_host, nullptr,
                           stream, 0, nullptr);
  const int flagged = host_counters[0];
  int unresolved = 0;
  if (flagged > 0) {
    const int st = resolve_flagged(w, x_dev, x_dev, n, d, k, 0, out_dist, out_idx, flagged, &unresolved, s);
    if (st != VQACX_OK) return st;
  }
  if (stats_host) {
    stats_host[kStatFlaggedRows] = flagged;
    stats_host[kStatUnresolvedRows] = unresolved;
    stats_host[kStatBoundViolations] = host_counters[8 + kStatBoundViolations];
    stats_host[kStatUsedTensorPath] = 1;
    stats_host[kStatRerankEvaluated] = host_counters[8 + kStatRerankEvaluated];
    stats_host[kStatMirrorKeys] = host_counters[8 + kStatMirrorKeys];
  }
  return VQACX_OK;
}

int vqacx_knn_selfjoin_shard_supported(int64_t n, int d, int world) {
  return (n > 0 && d > 0 && (d % kGemmBK) == 0 && n < (1LL << 31) - 512 && use_pair_kernel() &&
          shard_sym_eligible(n, world)) ? 1 : 0;
}

size_t vqacx_knn_selfjoin_shard_workspace_bytes(int64_t n, int d, int world) {
  (void)d;
  if (n < 0 || world < 1) return 0;
  return carve_shard_sym(nullptr, n, world).bytes;
}

int vqacx_knn_selfjoin_shard(const float* x_f32, const void* x_lp, const float* norm_sq,
                             const uint32_t* max_norm_sq_bits, int64_t n, int d, int k, int rank, int world, int phase,
                             float* thr_io, float* hi_io, double* out_dist, int64_t* out_idx, void* workspace,
                             size_t workspace_bytes, int32_t* stats_host, float* phase_ms_host, void* stream) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (phase < 1 || phase > 3 || world < 1 || rank < 0 || rank >= world) return VQACX_EINVAL;
  if (!vqacx_knn_selfjoin_shard_supported(n, d, world) || k <= 0 || k > VQACX_KNN_MAX_K) return VQACX_EINVAL;
  if (!x_f32 || !x_lp || !norm_sq || !max_norm_sq_bits || !workspace || !thr_io) return VQACX_EINVAL;
  if (phase == 3 && (!hi_io || !out_dist || !out_idx)) return VQACX_EINVAL;
  if ((reinterpret_cast<uintptr_t>(x_f32) & 15) || (reinterpret_cast<uintptr_t>(x_lp) & 15) ||
      (reinterpret_cast<uintptr_t>(workspace) & 255))
    return VQACX_EALIGN;
  Workspace w = carve_shard_sym(workspace, n, world);
  if (w.bytes > workspace_bytes) return VQACX_EWORKSPACE;
  knn_pair_set_lockstep_buffer(w.counters + 32);
  const int G = sm_count() / 2;
  const int n_i = static_cast<int>(n);
  PhaseTimer pt(phase_ms_host != nullptr, s);
  if (phase_ms_host) std::memset(phase_ms_host, 0, sizeof(float) * VQACX_KNN_NPHASES);
  // same error model as vqacx_knn_search (DESIGN.md "kNN exactness")
  const float eps_d = 1.01f * 0.00048828125f;
  const float d_abs = 2.f * std::sqrt(static_cast<float>(d)) * 2.98023224e-8f;
  const float eps_sq_abs = 2.f * (d / 8 + 8) * 1.1920929e-7f;
  const float eps_sq_rel = 4.f * 1.1920929e-7f;

  if (phase == 1 || phase == 2) {
    ShardSymPlan p;
    make_shard_sym_plan(n, G, shard_sym_groups(world), rank, world, p);
    CUtensorMap tm_q, tm_db;
    if (make_tmap_2d(&tm_q, x_lp, 2, static_cast<uint64_t>(n), static_cast<uint64_t>(d), static_cast<uint64_t>(d) * 2,
                     kGemmBM) != 0 ||
        make_tmap_2d(&tm_db, x_lp, 2, static_cast<uint64_t>(n), static_cast<uint64_t>(d), static_cast<uint64_t>(d) * 2,
                     kGemmBN / 2) != 0)
      return VQACX_ETMA;
    const int L = phase - 1;
    if (phase == 1) {
      if (stats_host) std::memset(stats_host, 0, sizeof(int32_t) * VQACX_KNN_NSTATS);
      CU_TRY(cudaMemsetAsync(w.counters, 0, 32 * sizeof(int), s));
      // slots this rank never writes must read as empty in the finalize
      CU_TRY(cudaMemsetAsync(w.cnt, 0, static_cast<size_t>(p.T) * p.slots_per_tile * 2 * kGemmBM * sizeof(int), s));
      CU_TRY(cudaMemcpyAsync(w.items, p.items.data(), p.items.size() * sizeof(KnnWorkItem), cudaMemcpyHostToDevice, s));
      CU_TRY(cudaMemcpyAsync(w.cta_off, p.pair_off.data(), p.pair_off.size() * sizeof(int), cudaMemcpyHostToDevice, s));
      CU_TRY(cudaMemcpyAsync(w.tile_slot0, p.tile_slot0.data(), p.tile_slot0.size() * sizeof(int),
                             cudaMemcpyHostToDevice, s));
      CU_TRY(cudaMemcpyAsync(w.tile_nslots, p.tile_nslots.data(), p.tile_nslots.size() * sizeof(int),
                             cudaMemcpyHostToDevice, s));
      CU_TRY(launch_knn_sym_reset(w.sym, n_i, s));
    } else {
      CU_TRY(launch_knn_thr_import(thr_io, n_i, w.sym.gtau, s));
    }
    pt.mark();
    KnnSymArgs sa = w.sym;
    sa.mirror_launch = L;
    CU_TRY(launch_knn_gemm_topk_pair_sym(tm_q, tm_db, w.items + p.item0[L], w.cta_off + L * (G + 1), p.pairs[L], norm_sq,
                                         n_i, d / kGemmBK, w.cand, w.cnt, sa, s));
    if (phase == 1) {
      CU_TRY(launch_knn_thr_export(w.sym.gtau, n_i, thr_io, s));
      pt.mark();
      drop_pending_timers();             // a new search: nothing of an abandoned one may be written later
      pt.defer(phase_ms_host, 0, 0, 1);   // read by phase 3, which synchronises anyway
      release_timer_events(pt);
      return VQACX_OK;
    }
    CU_TRY(launch_knn_mirror_scatter(w.sym, &w.counters[3], s));
    pt.mark();
    CU_TRY(launch_knn_finalize_sym(w.cand, w.cnt, w.tile_slot0, w.tile_nslots, 2 * kGemmBM, n_i, w.sym, w.cand_val,
                                   w.cand_idx, &w.counters[0], w.flag_rows, &w.counters[8], s));
    if (hi_io) {   // result-exchange protocol (phase 3 follows); the candidate-exchange protocol packs the lists instead
      CU_TRY(launch_knn_local_hi(x_f32, n_i, d, w.cand_val, w.cand_idx, norm_sq, max_norm_sq_bits, eps_d, d_abs,
                                 eps_sq_abs, eps_sq_rel, k, hi_io, s));
      // row n of the exchanged bounds: [0] = 1 if one of this rank's mirror logs overflowed (keys were dropped)
      CU_TRY(launch_knn_flag_to_float(&w.counters[3], hi_io + static_cast<size_t>(n) * k, k, s));
    }
    pt.mark();
    pt.defer(phase_ms_host, 0, 0, 1);
    pt.defer(phase_ms_host, 1, 1, 2);
    release_timer_events(pt);
    return VQACX_OK;
  }

  // phase 3: float64 re-rank of this rank's candidates, pruned by the agreed bound on the global k-th distance; the
  // exclusion bound of a row is its threshold as it stands after this rank's launches (w.sym.gtau)
  pt.mark();
  const bool ordered = use_locality_order(n);
  if (ordered)
    CU_TRY(launch_knn_locality_order(w.cand_idx, n_i, n_i, w.order_key, w.order_hist, w.order_perm, s));
  CU_TRY(launch_knn_rerank_range(x_f32, ordered ? w.order_perm : nullptr, nullptr, 0, n_i, x_f32, n_i, d, w.cand_val,
                                 w.cand_idx, norm_sq, max_norm_sq_bits, eps_d, d_abs, eps_sq_abs, eps_sq_rel, k, 0, out_dist,
                                 reinterpret_cast<long long*>(out_idx), &w.counters[0], w.flag_rows, &w.counters[8], hi_io, s,
                                 ordered, w.sym.gtau, k + 1));
  pt.mark();
  int host_counters[32];
  CU_TRY(cudaMemcpyAsync(host_counters, w.counters, sizeof(host_counters), cudaMemcpyDeviceToHost, s));
  CU_TRY(cudaStreamSynchronize(s));
  resolve_pending_timers();
  const int flagged = host_counters[0];
  int unresolved = 0;
  if (flagged > 0) {
    // the exact search covers ALL rows, so a redone row's list overlaps the other ranks' lists: vqacx_topk_merge
    // drops the duplicates
    const int st = resolve_flagged(w, x_f32, x_f32, n, d, k, 0, out_dist, out_idx, flagged, &unresolved, s, k + 1);
    if (st != VQACX_OK) return st;
  }
  if (phase_ms_host) {
    pt.mark();
    CU_TRY(cudaStreamSynchronize(s));
    phase_ms_host[2] = pt.ms(0, 1);
    phase_ms_host[3] = pt.ms(1, 2);
  }
  if (stats_host) {
    stats_host[kStatFlaggedRows] = flagged;
    stats_host[kStatUnresolvedRows] = unresolved;
    stats_host[kStatBoundViolations] = host_counters[8 + kStatBoundViolations];
    stats_host[kStatUsedTensorPath] = 1;
    stats_host[kStatRerankEvaluated] = host_counters[8 + kStatRerankEvaluated];
    stats_host[kStatMirrorKeys] = host_counters[8 + kStatMirrorKeys];
    stats_host[6] = host_counters[3];   // 1: a mirror log of THIS rank overflowed (every rank sees it in the bound exchange)
  }
  return VQACX_OK;
}

int vqacx_knn_selfjoin_shard_pack(int64_t n, int d, int world, int64_t rows_alloc, int32_t* pack_out, void* workspace,
                                  size_t workspace_bytes, void* stream) {
  if (!vqacx_knn_selfjoin_shard_supported(n, d, world) || rows_alloc < n || !pack_out || !workspace) return VQACX_EINVAL;
  if (reinterpret_cast<uintptr_t>(workspace) & 255) return VQACX_EALIGN;
  Workspace w = carve_shard_sym(workspace, n, world);
  if (w.bytes > workspace_bytes) return VQACX_EWORKSPACE;
  CU_TRY(launch_knn_cand_pack(w.cand_val, w.cand_idx, w.sym.gtau, static_cast<int>(n), static_cast<int>(rows_alloc),
                              &w.counters[3], w.sym.mcnt, w.sym.mcap, pack_out, static_cast<cudaStream_t>(stream)));
  return VQACX_OK;
}

int vqacx_knn_selfjoin_shard_owner(const float* x_f32, const float* norm_sq, const uint32_t* max_norm_sq_bits, int64_t n,
                                   int d, int k, int world, const int32_t* pack_recv, int64_t per, int64_t q0,
                                   int64_t n_slice, double* out_dist, int64_t* out_idx, void* workspace,
                                   size_t workspace_bytes, int32_t* stats_host, float* phase_ms_host, void* stream) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (stats_host) std::memset(stats_host, 0, sizeof(int32_t) * VQACX_KNN_NSTATS);
  if (!vqacx_knn_selfjoin_shard_supported(n, d, world) || k <= 0 || k > VQACX_KNN_MAX_K) return VQACX_EINVAL;
  if (q0 < 0 || n_slice < 0 || n_slice > per || q0 + n_slice > n) return VQACX_EINVAL;
  if (!x_f32 || !norm_sq || !max_norm_sq_bits || !pack_recv || !workspace) return VQACX_EINVAL;
  if (reinterpret_cast<uintptr_t>(workspace) & 255) return VQACX_EALIGN;
  Workspace w = carve_shard_sym(workspace, n, world);
  if (w.bytes > workspace_bytes) return VQACX_EWORKSPACE;
  PhaseTimer pt(phase_ms_host != nullptr, s);
  if (phase_ms_host) std::memset(phase_ms_host, 0, sizeof(float) * VQACX_KNN_NPHASES);
  if (n_slice > 0 && (!out_dist || !out_idx)) return VQACX_EINVAL;
  const float eps_d = 1.01f * 0.00048828125f;
  const float d_abs = 2.f * std::sqrt(static_cast<float>(d)) * 2.98023224e-8f;
  const float eps_sq_abs = 2.f * (d / 8 + 8) * 1.1920929e-7f;
  const float eps_sq_rel = 4.f * 1.1920929e-7f;
  const int ns = static_cast<int>(n_slice), n_i = static_cast<int>(n);
  // counters [0..2] and the stats block restart for the owner's re-rank ([3] = this rank's log-overflow flag stays)
  CU_TRY(cudaMemsetAsync(w.counters, 0, 3 * sizeof(int), s));
  CU_TRY(cudaMemsetAsync(w.counters + 8, 0, 8 * sizeof(int), s));
  pt.mark();
  // everything below works on LOCAL query rows j = 0 .. n_slice (query row q0 + j): the query-side pointers are offset
  const float* q_f32 = x_f32 + q0 * d;
  CU_TRY(launch_knn_cand_merge(pack_recv, world, static_cast<int>(per), ns, w.cand_val, w.cand_idx, w.sym.gtau, s));
  pt.mark();
  const bool ordered = use_locality_order(n_slice);
  if (ordered) CU_TRY(launch_knn_locality_order(w.cand_idx, ns, n_i, w.order_key, w.order_hist, w.order_perm, s));
  CU_TRY(launch_knn_rerank_range(q_f32, ordered ? w.order_perm : nullptr, nullptr, 0, ns, x_f32, n_i, d, w.cand_val,
                                 w.cand_idx, norm_sq + q0, max_norm_sq_bits, eps_d, d_abs, eps_sq_abs, eps_sq_rel, k, 0,
                                 out_dist, reinterpret_cast<long long*>(out_idx), &w.counters[0], w.flag_rows,
                                 &w.counters[8], nullptr, s, ordered, w.sym.gtau));
  pt.mark();
  int host_counters[32];
  CU_TRY(cudaMemcpyAsync(host_counters, w.counters, sizeof(host_counters), cudaMemcpyDeviceToHost, s));
  CU_TRY(cudaStreamSynchronize(s));
  resolve_pending_timers();
  const int flagged = host_counters[0];
  int unresolved = 0;
  if (flagged > 0) {
    const int st = resolve_flagged(w, q_f32, x_f32, n, d, k, 0, out_dist, out_idx, flagged, &unresolved, s);
    if (st != VQACX_OK) return st;
  }
  if (phase_ms_host) {
    pt.mark();
    CU_TRY(cudaStreamSynchronize(s));
    phase_ms_host[1] = pt.ms(0, 1);
    phase_ms_host[2] = pt.ms(1, 2);
    phase_ms_host[3] = pt.ms(2, 3);
  }
  if (stats_host) {
    stats_host[kStatFlaggedRows] = flagged;
    stats_host[kStatUnresolvedRows] = unresolved;
    stats_host[kStatBoundViolations] = host_counters[8 + kStatBoundViolations];
    stats_host[kStatUsedTensorPath] = 1;
    stats_host[kStatRerankEvaluated] = host_counters[8 + kStatRerankEvaluated];
    stats_host[6] = host_counters[3];
  }
  return VQACX_OK;
}

int vqacx_knn_search(const float* db_f32, const void* db_lp, const float* db_norm_sq,
                     const uint32_t* db_max_norm_sq_bits, int64_t n_db, int64_t db_row0, const float* q_f32,
                     const void* q_lp, const float* q_norm_sq, int64_t n_q, int d, int k, double* out_dist,
                     int64_t* out_idx, void* workspace, size_t workspace_bytes, uint32_t flags,
                     int32_t* stats_host, float* phase_ms_host, void* stream) {
  return knn_search_impl(db_f32, db_lp, db_norm_sq, db_max_norm_sq_bits, n_db, db_row0, q_f32, q_lp, q_norm_sq, n_q, d, k,
                         out_dist, out_idx, workspace, workspace_bytes, flags, stats_host, phase_ms_host, stream, 0,
                         nullptr);
}

int vqacx_knn_search_phase(const float* db_f32, const void* db_lp, const float* db_norm_sq,
                           const uint32_t* db_max_norm_sq_bits, int64_t n_db, int64_t db_row0, const float* q_f32,
                           const void* q_lp, const float* q_norm_sq, int64_t n_q, int d, int k, double* out_dist,
                           int64_t* out_idx, void* workspace, size_t workspace_bytes, uint32_t flags,
                           int32_t* stats_host, float* phase_ms_host, void* stream, int phase, float* hi_io) {
  if (phase != 1 && phase != 2) return VQACX_EINVAL;
  return knn_search_impl(db_f32, db_lp, db_norm_sq, db_max_norm_sq_bits, n_db, db_row0, q_f32, q_lp, q_norm_sq, n_q, d, k,
                         out_dist, out_idx, workspace, workspace_bytes, flags, stats_host, phase_ms_host, stream, phase,
                         hi_io);
}

int vqacx_knn_kth_bound(const float* hi_all, int G, int64_t n_q, int k, float* bound, void* stream) {
  if (G <= 0 || k <= 0 || G * k > 256 || n_q < 0) return VQACX_EINVAL;
  if (n_q == 0) return VQACX_OK;
  if (!hi_all || !bound) return VQACX_EINVAL;
  CU_TRY(launch_knn_kth_bound(hi_all, G, n_q, k, bound, static_cast<cudaStream_t>(stream)));
  return VQACX_OK;
}

int vqacx_topk_merge_checked(const double* dist, const int64_t* idx, int G, int64_t n_q, int k, double* out_dist,
                             int64_t* out_idx, int32_t* fail_count, int32_t* fail_rows, void* stream) {
  if (G <= 0 || k <= 0 || G * k > 256 || n_q < 0 || n_q >= (1LL << 31)) return VQACX_EINVAL;
  if (n_q == 0) return VQACX_OK;
  if (!dist || !idx || !out_dist || !out_idx || !fail_count || !fail_rows) return VQACX_EINVAL;
  CU_TRY(launch_knn_merge(dist, reinterpret_cast<const long long*>(idx), G, n_q, k, out_dist,
                          reinterpret_cast<long long*>(out_idx), static_cast<cudaStream_t>(stream), k + 1, fail_count,
                          fail_rows));
  return VQACX_OK;
}

int vqacx_topk_merge(const double* dist, const int64_t* idx, int G, int64_t n_q, int k, double* out_dist,
                     int64_t* out_idx, void* stream) {
  if (G <= 0 || k <= 0 || G * k > 256 || n_q < 0) return VQACX_EINVAL;
  if (n_q == 0) return VQACX_OK;
  if (!dist || !idx || !out_dist || !out_idx) return VQACX_EINVAL;
  CU_TRY(launch_knn_merge(dist, reinterpret_cast<const long long*>(idx), G, n_q, k, out_dist,
                          reinterpret_cast<long long*>(out_idx), static_cast<cudaStream_t>(stream)));
  return VQACX_OK;
}

}  // extern "C"

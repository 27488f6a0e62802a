// K1 (CTA-pair version): kNN candidate generation with tcgen05.mma.cta_group::2.
//
// Same job as knn_gemm.cu (the inner arithmetic of `nbrs.kneighbors`, knn.py:49 -> sklearn ArgKmin), but two CTAs of
// one TPC form a pair (cluster of 2) that works on a [256 query x 256 database] tile:
//
//   - each CTA stages ITS 128 query rows (A, 16 KB per k-block) and ITS HALF of the database tile (B, 128 rows,
//     16 KB per k-block): 32 KB per CTA per k-block instead of 48 KB -> the L2 -> SM operand stream, which bounds the
//     single-CTA kernel (ncu: ~10 KB/cycle chip-wide at 71 % tensor activity), drops by a third, and the ring gets
//     6 stages instead of 4 in the same shared memory;
//   - the leader CTA's MMA warp issues ONE tcgen05.mma.cta_group::2 (M 256 = 128 per CTA, N 256, K 16) per k-step;
//     the hardware reads both CTAs' A rows and both halves of B; each CTA's TMEM receives its own 128 x 256 fp32
//     accumulator (two of them ping-pong, 512 columns);
//   - both CTAs run the same fused top-k' epilogue on their own 128 rows (knn_epilogue.cuh).
//
// Barrier protocol (sm = shared memory of one stage, acc = one TMEM accumulator):
//   full[s]    lives in the LEADER, count 1: the leader's producer arrives with expect_tx(bytes of BOTH CTAs);
//              both producers' TMA loads (.cta_group::2) complete_tx on it
//   empty[s]   one per CTA, count 1: the leader's tcgen05.commit multicasts the arrive to both CTAs
//   tfull[a]   one per CTA, count 1: multicast commit after the last k-block of a tile
//   tempty[a]  lives in the LEADER, count 8: the 4 epilogue warps of each CTA arrive (the peer's remotely)
#include "knn_epilogue.cuh"
#include "knn_internal.cuh"
#include "ptx.cuh"
#include "topk.cuh"
#include <algorithm>
#include <cstdlib>

namespace vqacx {

namespace {
constexpr int kStages = 6;
constexpr int kRowsPerCta = 128;                       // query rows and database rows staged per CTA
constexpr int kABytes = kRowsPerCta * kGemmBK * 2;     // 16 KB
constexpr int kBBytes = kRowsPerCta * kGemmBK * 2;     // 16 KB (half of the 256-row database tile)
constexpr int kStageBytes = kABytes + kBBytes;         // 32 KB per CTA
constexpr int kNumAcc = 2;
constexpr int kTmemCols = kNumAcc * kGemmBN;           // 512
constexpr int kUmmaK = 16;
constexpr int kAuxBytes = 8192;
constexpr int kSmemBytes = kStages * kStageBytes + kAuxBytes + kEpiStageBytes + kMirrorStageBytes + 1024;
constexpr int kThreads = 128 + 128 + 32 * kMirrorWarps;   // warps 0-2 roles, 3 idle, 4-7 row-side epilogue, 8-15 mirror (column-side) epilogue
static_assert(kSmemBytes <= 232448, "shared memory budget");
constexpr uint32_t kIdescPair = make_idesc(kFmtF16, kFmtF16, 2 * kRowsPerCta, kGemmBN);

struct Aux {
  uint64_t full[kStages];
  uint64_t empty[kStages];
  uint64_t tfull[kNumAcc];
  uint64_t tempty[kNumAcc];
  uint32_t tmem_base;
  uint32_t pad[3];
  float ynorm[kNumAcc][kGemmBN];
  float ctau[kNumAcc][kGemmBN];   // symmetric self-join: per-column thresholds of the tile
};
static_assert(sizeof(Aux) <= kAuxBytes, "aux region too small");
}  // namespace

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
knn_gemm_topk_pair_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_db,
                          const KnnWorkItem* __restrict__ items, const int* __restrict__ pair_off,
                          const float* __restrict__ db_norm_sq, int n_db, int kblocks, uint64_t* __restrict__ cand,
                          int* __restrict__ cnt, unsigned long long* __restrict__ prof_arg, int debug_arg, int item_skip,
                          int item_max, const KnnSymArgs sym, int* progress, int ls_window) {
#ifdef VQACX_DEBUG
  // measurement build: clock64() role counters (tools/time_knn_phases.py) and the VQACX_OPT_KNN_DEBUG timing experiments
  // (bits 1 / 8 / 16 / 32 skip work and INVALIDATE results; 2 / 4 change the L2 policy of the database loads)
  unsigned long long* const prof = prof_arg;
  const int debug_flags = debug_arg;
#else
  constexpr unsigned long long* prof = nullptr;   // default build: none of this exists in the binary
  constexpr int debug_flags = 0;
  (void)prof_arg; (void)debug_arg;
#endif
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  Aux* aux = reinterpret_cast<Aux*>(smem + kStages * kStageBytes);
  float* epi_stage = reinterpret_cast<float*>(smem + kStages * kStageBytes + kAuxBytes);
  const bool mirror_on = sym.gtau != nullptr && sym.mirror_launch != 0;

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();  // 0 = leader
  const int pair = blockIdx.x >> 1;

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
      // 4 epilogue warps x 2 CTAs (+ the 4 x 2 mirror warps in a mirrored launch); only the leader's copy is used
      mbar_init(&aux->tempty[a], mirror_on ? 8 + 2 * sym.mirror_warps : 8);
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc_pair(&aux->tmem_base, kTmemCols);
    tmem_relinquish_pair();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = aux->tmem_base;

  // this launch's slice of the pair's item list (the whole list, or one GEMM wave of it)
  const int it_begin = min(pair_off[pair] + item_skip, pair_off[pair + 1]);
  const int it_end = (item_max < pair_off[pair + 1] - it_begin) ? it_begin + item_max : pair_off[pair + 1];

  // Producer / MMA loops are warp-uniform (all lanes run the control flow, one elected lane issues): inside a divergent
  // `if (lane == 0)` region the compiler wraps every UTMALDG / UTCMMA in a waterfall loop (measured ~600 issue cycles per
  // k-block in gemm_tf32.cu before the same change).
  if (warp == 0) {
    {
      // ============================== TMA producer (both CTAs) ==============================
      int stage = 0;
      uint32_t phase = 0;
      long long w_empty = 0, w_ls = 0, w_issue = 0;
      const long long t_start = clock64();
      const uint64_t pol_keep = l2_policy_evict_last(), pol_stream = l2_policy_evict_first();
      // Lock-step throttle: the pairs of a launch walk item lists of the same shape (same database tiles at the same
      // list position), and a database tile is only shared through L2 if they reach it within a short window. Left
      // alone they drift apart by hundreds of tiles over a launch (ncu: the mirrored launch read 93 GB from DRAM for a
      // 0.5 GB matrix). Every pair publishes its position (item << 16 | tile) and a producer does not start a tile
      // more than ls_window tiles ahead of the slowest pair. A hint only: waits are bounded, and a CTA that ever
      // times out stops throttling (no forward-progress dependence between CTAs).
      bool ls_live = progress != nullptr;
      int ls_pm[3] = {0, 0, 0};    // positions sampled mid-tile (kLockstepMaxPairs = 96 = 3 per lane); all start at 0
      const int n_pairs_ls = static_cast<int>(gridDim.x >> 1);
      for (int it = it_begin; it < it_end; ++it) {
        const KnnWorkItem item = items[it];
        const int q_row0 = item.q_tile * (2 * kRowsPerCta) + static_cast<int>(rank) * kRowsPerCta;
        for (int t = item.db_tile_begin; t < item.db_tile_end; ++t) {
          const int db_row0 = t * kGemmBN + static_cast<int>(rank) * kRowsPerCta;
          const long long c_ls0 = prof ? clock64() : 0;
          if (ls_live) {
            const int my_it = it - it_begin, my_t = t - item.db_tile_begin;
            if (rank == 0 && lane == 0)
              *reinterpret_cast<volatile int*>(progress + pair) = (my_it << 16) | my_t;
            const int item_len = item.db_tile_end - item.db_tile_begin;
            // first look at the minimum sampled in the middle of the previous tile (positions only grow, so a stale
            // minimum is conservative): in the common case the producer does not sit out an L2 round trip here while
            // its ring drains
            int m = __reduce_min_sync(0xffffffffu, min(ls_pm[0], min(ls_pm[1], ls_pm[2])));
            for (int polls = 0;; ++polls) {
              const int ahead = (my_it - (m >> 16)) * item_len + (my_t - (m & 0xffff));
              if (ahead <= ls_window) break;
              if (polls >= 4096) { ls_live = false; break; }
              if (polls) __nanosleep(64);
              m = 0x7fffffff;
              for (int j = lane; j < n_pairs_ls; j += 32) m = min(m, *reinterpret_cast<volatile int*>(progress + j));
              m = __reduce_min_sync(0xffffffffu, m);
            }
          }
          if (prof) w_ls += clock64() - c_ls0;
          for (int kb = 0; kb < kblocks; ++kb) {
            if (ls_live && kb == (kblocks >> 1)) {   // loads only: their values are first used at the next tile
#pragma unroll
              for (int j = 0; j < 3; ++j)
                ls_pm[j] = (lane + 32 * j < n_pairs_ls) ? *reinterpret_cast<volatile int*>(progress + lane + 32 * j)
                                                        : 0x7fffffff;
            }
            if (prof) {
              const long long c0 = clock64();
              mbar_wait(&aux->empty[stage], phase ^ 1);
              w_empty += clock64() - c0;
            } else {
              mbar_wait(&aux->empty[stage], phase ^ 1);
            }
            uint8_t* sa = smem + stage * kStageBytes;
            const uint32_t full_leader = mapa_u32(smem_u32(&aux->full[stage]), 0);
            const long long c_is0 = prof ? clock64() : 0;
            if (elect_one()) {
              if (rank == 0) mbar_arrive_expect_tx(&aux->full[stage], 2 * kStageBytes);
              // the query tile is re-read for every database tile (keep it in L2); the database tile is read by all
              // pairs within a short window and then not again for a whole sweep (let it go first)
              tma_load_2d_pair_hint(sa, &tm_q, full_leader, kb * kGemmBK, q_row0, pol_keep);
              if (debug_flags & 2) tma_load_2d_pair(sa + kABytes, &tm_db, full_leader, kb * kGemmBK, db_row0);
              else if (debug_flags & 4) tma_load_2d_pair_hint(sa + kABytes, &tm_db, full_leader, kb * kGemmBK, db_row0, pol_keep);
              else tma_load_2d_pair_hint(sa + kABytes, &tm_db, full_leader, kb * kGemmBK, db_row0, pol_stream);
            }
            __syncwarp();
            if (prof) w_issue += clock64() - c_is0;
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
        }
      }
      if (progress != nullptr && rank == 0 && lane == 0) *reinterpret_cast<volatile int*>(progress + pair) = 0x7fffffff;
      if (prof && lane == 0) {
        prof[blockIdx.x * 8 + 0] = static_cast<unsigned long long>(w_empty);
        prof[blockIdx.x * 8 + 1] = static_cast<unsigned long long>(clock64() - t_start);
        // second bank (gridDim.x * 8 onwards): [0] lock-step wait, [1] TMA issue, [2..4] mirror warp 8
        prof[(gridDim.x + blockIdx.x) * 8 + 0] = static_cast<unsigned long long>(w_ls);
        prof[(gridDim.x + blockIdx.x) * 8 + 1] = static_cast<unsigned long long>(w_issue);
      }
    }
  } else if (warp == 1) {
    if (rank == 0) {
      // ============================== MMA issuer (leader CTA only) ==============================
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      long long w_full = 0, w_tempty = 0;
      const long long t_start = clock64();
      for (int it = it_begin; it < it_end; ++it) {
        const KnnWorkItem item = items[it];
        for (int t = item.db_tile_begin; t < item.db_tile_end; ++t) {
          if (prof) {
            const long long c0 = clock64();
            mbar_wait(&aux->tempty[acc], acc_phase ^ 1);
            w_tempty += clock64() - c0;
          } else {
            mbar_wait(&aux->tempty[acc], acc_phase ^ 1);
          }
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + acc * kGemmBN;
          for (int kb = 0; kb < kblocks; ++kb) {
            if (prof) {
              const long long c0 = clock64();
              mbar_wait(&aux->full[stage], phase);
              w_full += clock64() - c0;
            } else {
              mbar_wait(&aux->full[stage], phase);
            }
            tc_fence_after();
            const uint32_t a_addr = smem_u32(smem + stage * kStageBytes);
            const uint32_t b_addr = a_addr + kABytes;
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < kGemmBK / kUmmaK; ++k) {
                const uint64_t da = make_sdesc_sw128(a_addr + k * kUmmaK * 2, 16, 1024);
                const uint64_t db = make_sdesc_sw128(b_addr + k * kUmmaK * 2, 16, 1024);
                mma_f16_ss_pair(d_tmem, da, db, kIdescPair, (kb | k) != 0 ? 1u : 0u);
              }
              tc_commit_pair(&aux->empty[stage], 0x3);  // frees this stage in BOTH CTAs once the MMAs have read it
            }
            __syncwarp();
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
          if (elect_one()) tc_commit_pair(&aux->tfull[acc], 0x3);  // accumulator complete -> both epilogues
          __syncwarp();
          acc ^= 1;
          if (acc == 0) acc_phase ^= 1;
        }
      }
      if (prof && lane == 0) {
        prof[blockIdx.x * 8 + 2] = static_cast<unsigned long long>(w_full);
        prof[blockIdx.x * 8 + 3] = static_cast<unsigned long long>(w_tempty);
        prof[blockIdx.x * 8 + 4] = static_cast<unsigned long long>(clock64() - t_start);
      }
    }
  } else if (warp >= 8) {
    // ============================== mirror warps: column side of a symmetric block (both CTAs) ==============
    if (mirror_on && warp - 8 < sym.mirror_warps) {
      // Eight mirror warps: two per TMEM lane quarter (= warp % 4, as for the row-side warps), taking the even and the
      // odd 32-column chunks of a tile. The column-side scan is latency-bound per warp (a hub row -- a row that is close
      // to many queries -- makes its lane loop over several hits per chunk), and its per-tile time, not its instruction
      // count, is what the pairs' lock-step couples: two warps halve it.
      const int mw = warp - 8;                     // 0..7
      const int e = mw & 3, par = mw >> 2, n_mt = 32 * sym.mirror_warps;   // (4-warp mode: every warp takes all chunks)
      const int row_in_cta = e * 32 + lane;
      const int mt = threadIdx.x - 256;            // 0..255; the first 128 stage the tile's column thresholds
      int acc = 0;
      uint32_t acc_phase = 0;
      // this warp's append log (continues where the previous launch of the same search left it)
      const int region = static_cast<int>(blockIdx.x) * kMirrorWarps + mw;
      MirrorLog mlog;
      mlog.log = sym.log + static_cast<size_t>(region) * sym.log_cap;
      mlog.ring = reinterpret_cast<uint4*>(smem + kStages * kStageBytes + kAuxBytes + kEpiStageBytes) + mw * kMirrorRing;
      mlog.len = sym.log_len[region];
      mlog.cap = sym.log_cap;
      mlog.head = 0;
      mlog.fill = 0;
      MirrorQueue mq;
      mirror_queue_reset(mq);
      long long mw_tfull = 0, mw_scan = 0;
      const long long mt_start = clock64();
      for (int it = it_begin; it < it_end; ++it) {
        const KnnWorkItem item = items[it];
        const int q_row = item.q_tile * (2 * kRowsPerCta) + static_cast<int>(rank) * kRowsPerCta + row_in_cta;
        // a padding row (past n) never appends: +inf norm -> u = +inf
        const float xn = q_row < n_db ? __ldg(db_norm_sq + q_row) : __int_as_float(0x7f800000);
        // column thresholds of a tile: one per mirror thread, fetched one tile ahead (an L2 round trip otherwise sits in
        // front of every tile); -inf for columns past n
        static_assert(32 * kMirrorWarps == kGemmBN, "one threshold per mirror thread");
        auto load_ct = [&](int t, int half) -> float {
          const int col = t * kGemmBN + mt + half * 128;
          return col < n_db ? ordered_to_float(__ldcg(sym.gtau + col)) : __int_as_float(0xff800000);
        };
        const bool two = n_mt < kGemmBN;      // 4-warp mode: two thresholds per thread
        float ct0 = load_ct(item.db_tile_begin, 0), ct1 = two ? load_ct(item.db_tile_begin, 1) : 0.f;
        for (int t = item.db_tile_begin; t < item.db_tile_end; ++t) {
          float* ct = aux->ctau[acc];
          ct[mt] = ct0;
          if (two) ct[mt + 128] = ct1;
          if (t + 1 < item.db_tile_end) {
            ct0 = load_ct(t + 1, 0);
            if (two) ct1 = load_ct(t + 1, 1);
          }
          named_bar_sync(2, n_mt);
          const long long mc0 = prof ? clock64() : 0;
          mbar_wait(&aux->tfull[acc], acc_phase);
          tc_fence_after();
          const long long mc1 = prof ? clock64() : 0;
          const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(e * 32) << 16) + acc * kGemmBN;
          if (!(debug_flags & 8))
            knn_mirror_scan_tile(t_addr, ct, static_cast<uint32_t>(t * kGemmBN), xn, static_cast<uint32_t>(q_row), mlog, mq,
                                 debug_flags, two ? 0 : par, two ? 1 : 2);
          if (prof) { mw_tfull += mc1 - mc0; mw_scan += clock64() - mc1; }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster_relaxed(mapa_u32(smem_u32(&aux->tempty[acc]), 0));
          acc ^= 1;
          if (acc == 0) acc_phase ^= 1;
        }
        // queued hits carry (value, column); the row id is this item's: empty the queues before it changes
        mirror_queue_flush(mq, mlog, static_cast<uint32_t>(q_row), static_cast<uint32_t>(lane), debug_flags);
      }
      mirror_log_drain(mlog, static_cast<uint32_t>(lane));
      if (lane == 0) sym.log_len[region] = mlog.len;
      if (prof && warp == 8 && lane == 0) {
        prof[(gridDim.x + blockIdx.x) * 8 + 2] = static_cast<unsigned long long>(mw_tfull);
        prof[(gridDim.x + blockIdx.x) * 8 + 3] = static_cast<unsigned long long>(mw_scan);
        prof[(gridDim.x + blockIdx.x) * 8 + 4] = static_cast<unsigned long long>(clock64() - mt_start);
      }
    }
  } else if (warp >= 4) {
    // ============================== epilogue: fused top-k' (both CTAs) ==============================
    const int e = warp - 4;
    const int row_in_cta = e * 32 + lane;
    const int et = threadIdx.x - 128;
    float* vrow = epi_stage + et * kEpiStagePitch;
    int acc = 0;
    uint32_t acc_phase = 0;
    long long w_tfull = 0, c_ld = 0, c_push = 0;
    const long long t_start = clock64();
    for (int it = it_begin; it < it_end; ++it) {
      const KnnWorkItem item = items[it];
      const size_t slot_row = static_cast<size_t>(item.slot) * (2 * kRowsPerCta) + rank * kRowsPerCta + row_in_cta;
      RowFilter st;
      filter_reset(st, cand + slot_row * kTopkCap);
      // symmetric self-join: start from the query's published threshold and keep publishing what this item tightens
      const int q_row = item.q_tile * (2 * kRowsPerCta) + static_cast<int>(rank) * kRowsPerCta + row_in_cta;
      const bool sym_row = sym.gtau != nullptr && q_row < n_db;
      uint32_t published = 0xff800000u;   // ordered(+inf)
      if (sym_row) {
        published = __ldcg(sym.gtau + q_row);
        st.tau = ordered_to_float(published);
      }
      for (int t = item.db_tile_begin; t < item.db_tile_end; ++t) {
        float* yn = aux->ynorm[acc];
        for (int i = et; i < kGemmBN; i += 128) {
          const int col = t * kGemmBN + i;
          yn[i] = col < n_db ? __ldg(db_norm_sq + col) : __int_as_float(0x7f800000);
        }
        named_bar_sync(1, 128);
        if (prof) {
          const long long c0 = clock64();
          mbar_wait(&aux->tfull[acc], acc_phase);
          w_tfull += clock64() - c0;
        } else {
          mbar_wait(&aux->tfull[acc], acc_phase);
        }
        tc_fence_after();
        const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(e * 32) << 16) + acc * kGemmBN;
        if (debug_flags & 1) { /* timing experiment: no epilogue work (results invalid) */ }
        else if (prof) knn_epilogue_tile<true>(st, t_addr, yn, static_cast<uint32_t>(t * kGemmBN), vrow, &c_ld, &c_push);
        else knn_epilogue_tile<false>(st, t_addr, yn, static_cast<uint32_t>(t * kGemmBN), vrow);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster_relaxed(mapa_u32(smem_u32(&aux->tempty[acc]), 0));
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
        if (sym_row) {
          const uint32_t o = float_to_ordered(st.tau);
          if (o < published) {
            atomicMin(sym.gtau + q_row, o);
            published = o;
          }
        }
      }
      if (sym.gtau != nullptr) {
        // The threshold a row publishes decides how many column-side hits every OTHER block produces for it. Between
        // compactions up to kTopkCap - 32 keys pass tau; a last compaction (any row above kTopkKeepMax keys) brings
        // that back to 64..96 keys before the value is published for the next launch.
        filter_compact_warp(st, kTopkCap - kTopkKeepMax);
        if (sym_row) {
          const uint32_t o = float_to_ordered(st.tau);
          if (o < published) atomicMin(sym.gtau + q_row, o);
        }
      }
      cnt[slot_row] = st.cnt;
    }
    if (prof && threadIdx.x == 128) {
      prof[blockIdx.x * 8 + 5] = static_cast<unsigned long long>(w_tfull);
      prof[blockIdx.x * 8 + 6] = static_cast<unsigned long long>(clock64() - t_start);
      prof[blockIdx.x * 8 + 7] = static_cast<unsigned long long>(c_ld);
      prof[blockIdx.x * 8 + 0] = rank == 0 ? prof[blockIdx.x * 8 + 0] : static_cast<unsigned long long>(c_push);
    }
  }

  // both CTAs must be done with each other's shared memory / TMEM before either exits or frees
  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) tmem_dealloc_pair(tmem_base, kTmemCols);
}

static cudaError_t launch_pair_impl(const CUtensorMap& tm_q, const CUtensorMap& tm_db, const KnnWorkItem* items,
                                    const int* pair_off, int n_pairs, const float* db_norm_sq, int n_db, int kblocks,
                                    uint64_t* cand, int* cnt, int item_skip, int item_max, const KnnSymArgs& sym,
                                    cudaStream_t s);

cudaError_t launch_knn_gemm_topk_pair(const CUtensorMap& tm_q, const CUtensorMap& tm_db, const KnnWorkItem* items,
                                      const int* pair_off, int n_pairs, const float* db_norm_sq, int n_db,
                                      int kblocks, uint64_t* cand, int* cnt, cudaStream_t s) {
  return launch_knn_gemm_topk_pair_range(tm_q, tm_db, items, pair_off, n_pairs, db_norm_sq, n_db, kblocks, cand, cnt, 0,
                                         0x7fffffff, s);
}

cudaError_t launch_knn_gemm_topk_pair_range(const CUtensorMap& tm_q, const CUtensorMap& tm_db, const KnnWorkItem* items,
                                            const int* pair_off, int n_pairs, const float* db_norm_sq, int n_db,
                                            int kblocks, uint64_t* cand, int* cnt, int item_skip, int item_max,
                                            cudaStream_t s) {
  return launch_pair_impl(tm_q, tm_db, items, pair_off, n_pairs, db_norm_sq, n_db, kblocks, cand, cnt, item_skip,
                          item_max, KnnSymArgs{}, s);
}

// Lock-step progress array of the searches this host thread issues: kLockstepMaxPairs ints of the caller's workspace
// (knn_pair_set_lockstep_buffer, called by the ABI entry points after carving); null = no throttle.
static thread_local int* t_lockstep_buf = nullptr;
void knn_pair_set_lockstep_buffer(int* buf) { t_lockstep_buf = buf; }

static cudaError_t launch_pair_impl(const CUtensorMap& tm_q, const CUtensorMap& tm_db, const KnnWorkItem* items,
                                    const int* pair_off, int n_pairs, const float* db_norm_sq, int n_db, int kblocks,
                                    uint64_t* cand, int* cnt, int item_skip, int item_max, const KnnSymArgs& sym,
                                    cudaStream_t s) {
  if (n_pairs <= 0) return cudaSuccess;
  unsigned long long* prof = knn_debug_profile_buffer();
  const int debug_flags = opt(kOptKnnDebug);   // 0 unless this is a -DVQACX_DEBUG build
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(knn_gemm_topk_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kSmemBytes);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  count_launch();
  KnnSymArgs sym_dbg = sym;
  sym_dbg.debug = debug_flags;
  // lock-step throttle (see the producer loop): window in database tiles, 0 = off (VQACX_OPT_KNN_LOCKSTEP)
  // The mirrored launch of the symmetric self-join has its own, wider window (VQACX_OPT_KNN_SYM_LOCKSTEP): its tiles
  // take unequal time (the column-side hit loop runs ~35 times per warp and tile, in bursts), and with 8 tiles of slack
  // every pair spent ~38 % of the launch waiting for whichever pair was slowest at the moment (producer role counter:
  // 13.8 of 36.2 Mcycles). Interleaved A/B, candidate phase of the 123 287-row self-join (ten rounds, one box): window
  // 8: 31-43 ms, 16: 28-38, 24: 27.2-30.9, 32: 26.7-27.8, 48: 26.1-26.6 (the default), 64: 26.3-35.4 (bistable: an item
  // is 68-69 tiles long), 96: 26.8-27.6, 0 (off): 28.4-29.5. The plain sweep keeps 8 (config 4: 358 ms vs 382 at 64).
  // With EIGHT mirror warps (two per TMEM lane quarter) a tile's column side takes half as long and the pairs drift
  // less: medians of ten rounds 8: 30.2, 12: 28.2, 16: 25.4, 24: 24.8 (the default now), 32: 25.6, 48: 26.7 ms.
  const int ls_window = std::max(0, opt(sym.gtau != nullptr && sym.mirror_launch != 0 ? kOptKnnSymLockstep : kOptKnnLockstep));
  int* progress = nullptr;
  if (ls_window > 0 && t_lockstep_buf && n_pairs > 1 && n_pairs <= kLockstepMaxPairs &&
      (n_db + kGemmBN - 1) / kGemmBN < 65536) {
    progress = t_lockstep_buf;
    cudaError_t e = cudaMemsetAsync(progress, 0, sizeof(int) * kLockstepMaxPairs, s);
    if (e != cudaSuccess) return e;
  }
  knn_gemm_topk_pair_kernel<<<2 * n_pairs, kThreads, kSmemBytes, s>>>(tm_q, tm_db, items, pair_off, db_norm_sq, n_db,
                                                                 kblocks, cand, cnt, prof, debug_flags, item_skip, item_max,
                                                                 sym_dbg, progress, ls_window);
  return cudaGetLastError();
}

cudaError_t launch_knn_gemm_topk_pair_sym(const CUtensorMap& tm_q, const CUtensorMap& tm_db, const KnnWorkItem* items,
                                          const int* pair_off, int n_pairs, const float* db_norm_sq, int n_db,
                                          int kblocks, uint64_t* cand, int* cnt, const KnnSymArgs& sym, cudaStream_t s) {
  return launch_pair_impl(tm_q, tm_db, items, pair_off, n_pairs, db_norm_sq, n_db, kblocks, cand, cnt, 0, 0x7fffffff,
                          sym, s);
}

__global__ void knn_sym_reset_kernel(uint32_t* gtau, int* mcnt, int n, int* log_len, int n_regions) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    gtau[i] = 0xff800000u;   // float_to_ordered(+inf)
    mcnt[static_cast<size_t>(i) * kMirrorCntStride] = 0;
  }
  if (i < n_regions) log_len[i] = 0;
}

// One block per mirror-warp log: every entry reserves a slot in its query's mirror buffer. This is where the
// returning atomics live now: a bandwidth-shaped pass with thousands of them in flight per SM.
__global__ void __launch_bounds__(1024) knn_mirror_scatter_kernel(const KnnSymArgs sym, int* __restrict__ overflow_flag) {
  const int region = blockIdx.x;
  int n = sym.log_len[region];
  if (n > sym.log_cap) {
    if (threadIdx.x == 0) atomicOr(overflow_flag, 1);
    n = sym.log_cap;
  }
  const uint4* log = sym.log + static_cast<size_t>(region) * sym.log_cap;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const uint4 e = __ldcs(log + i);
    const int pos = atomicAdd(sym.mcnt + static_cast<size_t>(e.z) * kMirrorCntStride, 1);
    if (pos < sym.mcap)
      __stcg(reinterpret_cast<unsigned long long*>(sym.mbuf + static_cast<size_t>(e.z) * sym.mcap + pos),
             (static_cast<unsigned long long>(e.y) << 32) | e.x);
  }
}

cudaError_t launch_knn_sym_reset(const KnnSymArgs& sym, int n, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  count_launch();
  knn_sym_reset_kernel<<<(std::max(n, sym.n_regions) + 255) / 256, 256, 0, s>>>(sym.gtau, sym.mcnt, n, sym.log_len,
                                                                              sym.n_regions);
  return cudaGetLastError();
}

cudaError_t launch_knn_mirror_scatter(const KnnSymArgs& sym, int* overflow_flag, cudaStream_t s) {
  if (sym.n_regions <= 0) return cudaSuccess;
  count_launch();
  knn_mirror_scatter_kernel<<<sym.n_regions, 1024, 0, s>>>(sym, overflow_flag);
  return cudaGetLastError();
}

}  // namespace vqacx

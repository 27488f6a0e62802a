// Generic TF32 tensor-core GEMM for stage 2 (NeuralCX), sm_100a only.
//
// Replaces every dense product the reference issues through nn.Linear / torch.bmm on the NeuralCX path
// (cx.py:87-88 MUTAN + classifier, cx.py:281-282 softmax.E, cx.py:322-327 MLP) and the autograd products of
// their backward (counterexamples.py:337 loss.backward()).
//
// One 128 x bn output tile per CTA, optional split along K (grid.z). Warp roles as in knn_gemm.cu:
//   warp 0  TMA producer (fp32 tiles, 128 B swizzle; k-major operands: one box per tile, mn-major operands:
//           one 32-wide box per 32 columns, which lands exactly in the UMMA "MN-major SW128" canonical layout)
//   warp 1  tcgen05.mma kind::tf32, M = 128, N = bn, K = 8 per instruction, fp32 accumulator in TMEM
//   warp 2  TMEM allocator
//   warps 4-7 epilogue: tcgen05.ld, bias / per-question and per-rank row terms / activation / dropout / gate,
//           plain or atomic store
// Several (A, B) segments may accumulate into the same tile (e.g. layer 1: X_cand.W1c^T + p.M).
#include "gemm_tf32.cuh"
#include "knn_internal.cuh"
#include "ptx.cuh"
#include <algorithm>
#include <cuda_fp16.h>
#include <cmath>
#include <cstdlib>

// Timing experiments and clock64() role counters exist in -DVQACX_DEBUG builds only (the default build's kernel carries
// neither the branches nor the counters).
#ifdef VQACX_DEBUG
#define GEMM_DBG(a) ((a).debug)
#define GEMM_PROF(a) ((a).prof)
#else
#define GEMM_DBG(a) 0
#define GEMM_PROF(a) (static_cast<unsigned long long*>(nullptr))
#endif

namespace vqacx {

namespace {
constexpr int BM = 128;
constexpr int BK = 32;             // fp32 elements = one 128 B swizzle row
constexpr int kABytes = BM * 128;  // 16 KB
constexpr int kMaxStages = 8;
constexpr int kAuxBytes = 1024;
constexpr int kEpiWarps = 8;                            // two epilogue warps per TMEM lane quarter (even / odd chunks)
constexpr int kSmemBudget = 232448 - 2048 - kEpiWarps * 32 * 36 * 4;   // operand ring: what aux + staging + alignment leave
constexpr int kStgPitch = 36;                            // floats per staging row: 16-byte aligned, conflict-free both ways
constexpr int kStgBytes = kEpiWarps * 32 * kStgPitch * 4;  // one 32-row staging block per epilogue warp (36 KB)

struct KArgs {
  CUtensorMap ta[kGemmMaxSeg];
  CUtensorMap tb[kGemmMaxSeg];
  int kb_end[kGemmMaxSeg];  // exclusive prefix of k-blocks per segment
  int a_mn[kGemmMaxSeg];
  int b_mn[kGemmMaxSeg];
  int nseg;
  int f16;       // operands are fp16: a k-block is 64 elements (still one 128 B swizzle row), kind::f16 MMAs of K = 16
  int M, N, bn, stages, tmem_cols;
  int ntn, ntm, split, total_tiles;   // persistent tile loop: tile -> (n tile fastest, m tile, k split)
  float* C;
  long long ldc;
  int atomic;  // 1: atomicAdd into C
  int epi_vec; // 1: every pointer / pitch the epilogue touches allows 16-byte accesses (host-checked)
  int pair;    // 1: two CTAs of a TPC (cluster of 2) share every B tile: tcgen05.mma.cta_group::2 with M = 256 (128 rows of A
               // per CTA), each CTA stages its own A tile and HALF of the B rows -- the weight tile, which every CTA of a
               // large-M product re-reads per m-tile and which dominated the L2 -> SM operand stream (56 KB per k-block and
               // CTA for the N = 300 layer-1 product, 40 KB of it B), crosses L2 -> SM once per PAIR. fp16 k-major operands,
               // no split-K. ntm then counts PAIRS of m-tiles.
  int dual;    // 1: N spans exactly two n-tiles and one CTA computes BOTH from one staged A tile (accumulator 0 = columns
               // [0, bn), accumulator 1 = [bn, 2 bn)): A crosses L2 -> SM (and HBM) once instead of twice. Used when there
               // are enough m-tiles to fill the chip; the epilogue then no longer overlaps the next tile's mainloop.
  float* partial;             // non-null: store the raw accumulator tile to partial[z][m][n] (no epilogue, no C)
  long long partial_ld;       // floats per partial row (multiple of 32)
  long long partial_stride;   // floats per split
  GemmEpilogue epi;
  int debug;                 // tools only: timing experiments (bit0 no stores, bit1 no tmem ld, bit2 no addend loads)
  unsigned long long* prof;  // tools only: 8 cycle counters per CTA (see tools/time_gemm.py); null = off
};

struct Aux {
  uint64_t full[kMaxStages];
  uint64_t empty[kMaxStages];
  uint64_t tfull[2];
  uint64_t tempty[2];
  uint32_t tmem_base;
};
static_assert(sizeof(Aux) <= kAuxBytes, "aux too small");

// 32 consecutive floats of a row starting at column c0 (row has n valid columns): vector loads when the whole chunk is
// in range and 16-byte aligned, otherwise clamped scalar loads (no branches per element -> the loads are batched).
__device__ __forceinline__ void load_row32(const float* __restrict__ row, int c0, int n, bool full, float (&out)[32]) {
  const float* p = row + c0;
  if (full && (reinterpret_cast<uintptr_t>(p) & 15) == 0) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float4 t = __ldg(reinterpret_cast<const float4*>(p) + j);
      out[4 * j] = t.x; out[4 * j + 1] = t.y; out[4 * j + 2] = t.z; out[4 * j + 3] = t.w;
    }
  } else {
    const int last = n - 1 - c0;  // >= 0 because c0 < n
#pragma unroll
    for (int j = 0; j < 32; ++j) out[j] = __ldg(p + (j < last ? j : last));
  }
}

__device__ __forceinline__ float apply_act(float x, int act) {
  if (act == kActTanh) return tanhf(x);
  if (act == kActRelu) return fmaxf(x, 0.f);
  return x;
}
}  // namespace

// kPair: the CTA-pair variant (KArgs::pair) -- a separate instantiation, because a kernel that contains cluster
// instructions must be launched as a cluster, and because the single-CTA variant should not carry the extra state.
template <bool kPair>
__device__ __forceinline__ void gemm_tf32_body(const KArgs& args) {
  extern __shared__ uint8_t smem_raw[];
  const long long t_entry = clock64();
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int bn = args.bn;
  const int nsub = args.dual ? 2 : 1;
  constexpr bool pair = kPair;
  uint32_t rank = 0u;                                    // 0 = leader (issues the MMAs of the pair)
  if constexpr (kPair) rank = cluster_ctarank();
  const int bh = pair ? bn / 2 : bn;                     // B rows of one n sub-tile staged by THIS CTA
  const int stage_bytes = kABytes + nsub * bh * 128;
  Aux* aux = reinterpret_cast<Aux*>(smem + args.stages * stage_bytes);
  const int tile_first = pair ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
  const int tile_step = pair ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  // Persistent tile loop: CTA c works on tiles c, c + gridDim.x, ... (n tile fastest, so the CTAs running at the same
  // time share A row tiles through L2). Two TMEM accumulators ping-pong: the epilogue of tile t overlaps the TMA
  // loads and MMAs of tile t + 1, which is what the K <= 360 products (classifier, hidden layers) were missing.
  const int kb_total = args.kb_end[args.nseg - 1];
  const uint32_t acc_stride = static_cast<uint32_t>(args.tmem_cols >> 1);
  auto decode = [&](int tile, int& m0, int& n0, int& kb0, int& kb1, int& z) {
    const int ni = tile % args.ntn, r = tile / args.ntn;
    const int mi = r % args.ntm;
    z = r / args.ntm;
    m0 = pair ? (mi * 2 + static_cast<int>(rank)) * BM : mi * BM;
    n0 = ni * bn;
    kb0 = static_cast<int>(static_cast<long long>(kb_total) * z / args.split);
    kb1 = static_cast<int>(static_cast<long long>(kb_total) * (z + 1) / args.split);
  };

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < args.nseg; ++s) {
      tma_prefetch_desc(&args.ta[s]);
      tma_prefetch_desc(&args.tb[s]);
    }
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < args.stages; ++s) {
      mbar_init(&aux->full[s], 1);
      mbar_init(&aux->empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&aux->tfull[a], 1);
      mbar_init(&aux->tempty[a], pair ? 2 * kEpiWarps : kEpiWarps);   // one arrive per epilogue warp (of both CTAs)
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    if constexpr (kPair) {
      tmem_alloc_pair(&aux->tmem_base, args.tmem_cols);
      tmem_relinquish_pair();
    } else {
      tmem_alloc(&aux->tmem_base, args.tmem_cols);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  if constexpr (kPair) cluster_sync_all();
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = aux->tmem_base;
  // Programmatic dependent launch: everything above (barriers, TMEM, descriptor prefetch) may run while the previous
  // kernel of the stream is still finishing; the next kernel may be scheduled from here on (AFTER this CTA owns its
  // TMEM columns, so a younger grid can never take them from an older one that it is waiting for).
  pdl_trigger();
  pdl_wait();

  // Producer and MMA loops run WARP-UNIFORM (all 32 lanes execute the control flow; one elected lane issues the
  // TMA / tcgen05 instructions): inside a divergent `if (lane == 0)` region the compiler cannot keep addresses and
  // descriptors in uniform registers and wraps every UTMALDG / UTCMMA in a waterfall loop, which measured ~600 cycles
  // of issue overhead per k-block (more than the MMAs of a narrow tile).
  if (warp == 0) {
    int stage = 0;
    uint32_t phase = 0;
    const long long t0 = clock64();
    long long w_empty = 0;
    const int bk_elems = args.f16 ? 2 * BK : BK;   // elements per 128-byte k-block
    for (int tile = tile_first; tile < args.total_tiles; tile += tile_step) {
    int m0, n0, kb0, kb1, z, seg = 0;
    decode(tile, m0, n0, kb0, kb1, z);
    for (int kb = kb0; kb < kb1; ++kb) {
      while (kb >= args.kb_end[seg]) ++seg;
      const int kloc = kb - (seg ? args.kb_end[seg - 1] : 0);
      if (GEMM_PROF(args)) {
        const long long c0 = clock64();
        mbar_wait(&aux->empty[stage], phase ^ 1);
        w_empty += clock64() - c0;
      } else {
        mbar_wait(&aux->empty[stage], phase ^ 1);
      }
      uint8_t* sa = smem + stage * stage_bytes;
      uint8_t* sb = sa + kABytes;
      const int a_mn = args.a_mn[seg], b_mn = args.b_mn[seg];
      if constexpr (kPair) {
        // both CTAs' loads complete on the LEADER's full barrier, which expects the bytes of both
        const uint32_t full_leader = mapa_u32(smem_u32(&aux->full[stage]), 0);
        if (elect_one()) {
          if (rank == 0) mbar_arrive_expect_tx(&aux->full[stage], 2 * stage_bytes);
          tma_load_2d_pair(sa, &args.ta[seg], full_leader, kloc * bk_elems, m0);
          for (int h = 0; h < nsub; ++h)
            tma_load_2d_pair(sb + h * bh * 128, &args.tb[seg], full_leader, kloc * bk_elems,
                             n0 + h * bn + static_cast<int>(rank) * bh);
        }
      } else if (elect_one()) {
        mbar_arrive_expect_tx(&aux->full[stage], stage_bytes);
        if (!a_mn) {
          tma_load_2d(sa, &args.ta[seg], &aux->full[stage], kloc * bk_elems, m0);
        } else {
          for (int j = 0; j < BM / 32; ++j)
            tma_load_2d(sa + j * 4096, &args.ta[seg], &aux->full[stage], m0 + 32 * j, kloc * BK);
        }
        for (int h = 0; h < nsub; ++h) {
          uint8_t* sbh = sb + h * bn * 128;
          const int nh = n0 + h * bn;
          if (!b_mn) {
            tma_load_2d(sbh, &args.tb[seg], &aux->full[stage], kloc * bk_elems, nh);
          } else {
            for (int j = 0; j < bn / 32; ++j)
              tma_load_2d(sbh + j * 4096, &args.tb[seg], &aux->full[stage], nh + 32 * j, kloc * BK);
          }
        }
      }
      __syncwarp();
      if (++stage == args.stages) { stage = 0; phase ^= 1; }
    }
    }
    if (GEMM_PROF(args) && lane == 0) {
      unsigned long long* pr = GEMM_PROF(args) + 8 * blockIdx.x;
      pr[0] = static_cast<unsigned long long>(w_empty);
      pr[1] = static_cast<unsigned long long>(clock64() - t0);
      pr[7] = static_cast<unsigned long long>(t0 - t_entry);
    }
  } else if (warp == 1 && rank == 0) {
    int stage = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
    const long long t0 = clock64();
    long long w_full = 0;
    for (int tile = tile_first; tile < args.total_tiles; tile += tile_step) {
    int m0, n0, kb0, kb1, z, seg = 0;
    decode(tile, m0, n0, kb0, kb1, z);
    mbar_wait(&aux->tempty[acc], acc_phase ^ 1);   // the epilogue has drained this accumulator (first use: free)
    if (args.dual) mbar_wait(&aux->tempty[1], acc_phase ^ 1);
    tc_fence_after();
    const uint32_t d_tmem = tmem_base + acc * acc_stride;
    for (int kb = kb0; kb < kb1; ++kb) {
      while (kb >= args.kb_end[seg]) ++seg;
      const int a_mn = args.a_mn[seg], b_mn = args.b_mn[seg];
      const uint32_t idesc = args.f16 ? make_idesc(kFmtF16, kFmtF16, BM, bn, 0, 0)
                                      : make_idesc(kFmtTF32, kFmtTF32, BM, bn, a_mn, b_mn);
      if (GEMM_PROF(args)) {
        const long long c0 = clock64();
        mbar_wait(&aux->full[stage], phase);
        w_full += clock64() - c0;
      } else {
        mbar_wait(&aux->full[stage], phase);
      }
      tc_fence_after();
      const uint32_t a_addr = smem_u32(smem + stage * stage_bytes);
      const uint32_t b_addr = a_addr + kABytes;
      if constexpr (kPair) {
        // one instruction covers the pair: A rows [0,128) from this CTA and [128,256) from the peer, the two halves of
        // the B tile from the same offsets of both CTAs' stages; every CTA's TMEM receives its own 128 x bn accumulator
        const uint32_t idesc2 = make_idesc(kFmtF16, kFmtF16, 2 * BM, bn, 0, 0);
        if (elect_one()) {
          for (int h = 0; h < nsub; ++h) {
#pragma unroll
            for (int k = 0; k < BK / 8; ++k) {
              const uint64_t da = make_sdesc_sw128(a_addr + k * 32, 16, 1024);
              const uint64_t db = make_sdesc_sw128(b_addr + h * bh * 128 + k * 32, 16, 1024);
              mma_f16_ss_pair(d_tmem + h * acc_stride, da, db, idesc2, (kb > kb0 || k > 0) ? 1u : 0u);
            }
          }
          tc_commit_pair(&aux->empty[stage], 0x3);   // frees this stage in BOTH CTAs once the MMAs have read it
        }
      } else if (elect_one()) {
#pragma unroll
        for (int k = 0; k < BK / 8; ++k) {
          // k-major: 8 tf32 = 32 B further along the swizzled row; mn-major: the next 8 k-rows (two 4-row groups)
          const uint64_t da = a_mn ? make_sdesc_sw128_atom32(a_addr + k * 1024, 4096, 512)
                                   : make_sdesc_sw128(a_addr + k * 32, 16, 1024);
          const uint64_t db = b_mn ? make_sdesc_sw128_atom32(b_addr + k * 1024, 4096, 512)
                                   : make_sdesc_sw128(b_addr + k * 32, 16, 1024);
          // (fp16: 16 halfs = the same 32 bytes further along the swizzled row per K step)
          if (args.f16) mma_f16_ss(d_tmem, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          else mma_tf32_ss(d_tmem, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
        }
        if (args.dual) {   // second n-tile: same A stage, the B rows staged behind the first half, accumulator 1
#pragma unroll
          for (int k = 0; k < BK / 8; ++k) {
            const uint32_t b2 = b_addr + bn * 128;
            const uint64_t da = a_mn ? make_sdesc_sw128_atom32(a_addr + k * 1024, 4096, 512)
                                     : make_sdesc_sw128(a_addr + k * 32, 16, 1024);
            const uint64_t db = b_mn ? make_sdesc_sw128_atom32(b2 + k * 1024, 4096, 512)
                                     : make_sdesc_sw128(b2 + k * 32, 16, 1024);
            if (args.f16) mma_f16_ss(d_tmem + acc_stride, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            else mma_tf32_ss(d_tmem + acc_stride, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
        }
        tc_commit(&aux->empty[stage]);
      }
      __syncwarp();
      if (++stage == args.stages) { stage = 0; phase ^= 1; }
    }
    if (elect_one()) {
      if constexpr (kPair) {
        tc_commit_pair(&aux->tfull[acc], 0x3);
        if (args.dual) tc_commit_pair(&aux->tfull[1], 0x3);
      } else {
        tc_commit(&aux->tfull[acc]);
        if (args.dual) tc_commit(&aux->tfull[1]);
      }
    }
    __syncwarp();
    if (args.dual) {
      acc_phase ^= 1;            // both accumulators are used by every tile
    } else {
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
    }
    if (GEMM_PROF(args) && lane == 0) {
      unsigned long long* pr = GEMM_PROF(args) + 8 * blockIdx.x;
      pr[2] = static_cast<unsigned long long>(w_full);
      pr[3] = static_cast<unsigned long long>(clock64() - t0);
    }
  } else if (warp >= 4) {
    // ============================== epilogue ==============================
    // Two phases per 32-column chunk. (A) thread = accumulator row (that is how tcgen05.ld hands the data out): the raw
    // 32 values go to this warp's staging rows in shared memory. (B) lanes re-read them TRANSPOSED -- 8 lanes cover the
    // 32 columns of one row, 4 rows per pass -- so that every global access of the epilogue (bias / row terms / gate /
    // mask loads, fp32 and fp16 stores) touches whole 64-128 byte row segments. With thread = row in phase B every
    // warp-wide store scattered 16-byte pieces over 32 rows: measured on the classifier product of an eval batch
    // (98 304 x 2000 x 384, fp16 out) 440 us, of which ~300 us were those stores and ~110 us the per-thread addend loads.
    // A warp may only read the TMEM lanes of its quarter (warp % 4); the two warps of a quarter take the even and the
    // odd 32-column chunks. One epilogue warp per scheduler could not hide its own latencies (measured: the epilogue,
    // not the mainloop, bound every product with K <= 400).
    const int e = warp & 3, par = (warp - 4) >> 2;
    const GemmEpilogue& ep = args.epi;
    float* stg = reinterpret_cast<float*>(smem + args.stages * stage_bytes + kAuxBytes) + (warp - 4) * (32 * kStgPitch);
    const int lr = lane >> 3, lc = (lane & 7) * 4;
    int acc = 0;
    uint32_t acc_phase = 0;
    long long t_wait = 0, t_work = 0;
    const float drop_scale = ep.drop_p > 0.f ? 1.f / (1.f - ep.drop_p) : 1.f;
    const unsigned long long drop_seed = ep.drop_seed + ((ep.drop_seed_dev && ep.drop_p > 0.f) ? __ldg(ep.drop_seed_dev) : 0ull);
    auto load4 = [](const float* p, int nvalid) -> float4 {   // p[0..4) with the first nvalid elements in range
      if (nvalid >= 4 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) return __ldg(reinterpret_cast<const float4*>(p));
      float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
      if (nvalid > 0) t.x = __ldg(p);
      if (nvalid > 1) t.y = __ldg(p + 1);
      if (nvalid > 2) t.z = __ldg(p + 2);
      if (nvalid > 3) t.w = __ldg(p + 3);
      return t;
    };
    for (int tile = tile_first; tile < args.total_tiles; tile += tile_step) {
    for (int half = 0; half < nsub; ++half) {
    int m0, n0, kb0, kb1, z;
    decode(tile, m0, n0, kb0, kb1, z);
    if (args.dual) { n0 = half * bn; acc = half; }
    const int mw = m0 + e * 32;              // first row of this warp
    const bool first_split = z == 0;
    const long long t_e0 = clock64();
    if (kb1 > kb0) {
      mbar_wait(&aux->tfull[acc], acc_phase);
      tc_fence_after();
    }
    const long long t_e1 = clock64();
    const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(e * 32) << 16) + acc * acc_stride;
    const bool has_bias = ep.bias && first_split && !args.partial;
    const bool has_r1 = ep.rvec && first_split && !args.partial;
    const bool has_ra = ep.row_add && first_split && !args.partial;
    const bool has_ra2 = ep.row_add2 && first_split && !args.partial;
    const bool has_ra3 = ep.row_add3 && first_split && !args.partial;
    const float rv_lane = (has_r1 && mw + lane < args.M) ? __ldg(ep.rvec + mw + lane) : 0.f;
    // Per-tile, per-lane row bookkeeping of phase B (rows rr_i = 4 i + lr, i < 8): validity, the offsets of the
    // per-question / per-rank addend rows (one integer division per row and TILE instead of per chunk) and the rank-1
    // row factors. Everything the inner loop needs then sits in registers.
    uint32_t rmask = 0;
    int ra_off[8], ra2_off[8], ra3_off[8];
    float rvs[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int m = mw + 4 * i + lr;
      if (m < args.M) rmask |= 1u << i;
      const int mc = m < args.M ? m : 0;
      ra_off[i] = has_ra ? (mc / ep.row_add_div) * static_cast<int>(ep.row_add_ld) : 0;
      ra2_off[i] = has_ra2 ? (mc % ep.row_add2_mod) * static_cast<int>(ep.row_add2_ld) : 0;
      ra3_off[i] = has_ra3 ? __ldg(ep.row_add3_idx + mc) * static_cast<int>(ep.row_add3_ld) : 0;
      rvs[i] = __shfl_sync(0xffffffffu, rv_lane, 4 * i + lr);
    }
    const bool extras = ep.drop_mask || ep.drop_p > 0.f || ep.gate || ep.alpha != 1.f;
#pragma unroll 1
    for (int chunk = par; chunk < bn / 32; chunk += kEpiWarps / 4) {
      const int c0 = n0 + chunk * 32;
      if (c0 >= args.N) break;  // warp-uniform
      // ---- phase A: accumulator rows -> staging ----
      {
        uint32_t r[32];
        if (kb1 > kb0 && !(GEMM_DBG(args) & 2)) {
          tmem_ld_32x32(t_addr + chunk * 32, r);
          tmem_ld_wait_regs(r);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) r[j] = 0u;
        }
        float4* srow = reinterpret_cast<float4*>(stg + lane * kStgPitch);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          srow[j] = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]), __uint_as_float(r[4 * j + 2]),
                                __uint_as_float(r[4 * j + 3]));
      }
      __syncwarp();
      // ---- phase B: lane = 4 consecutive columns of one row ----
      const int c = c0 + lc;
      const int nv = args.N - c;             // valid columns among this lane's four (<= 0: none)
      if (args.partial) {
        // deterministic split-K, stage 1: raw partial sums (ld is a multiple of 32: always a full aligned float4)
        float* pp = args.partial + z * args.partial_stride + static_cast<long long>(mw + lr) * args.partial_ld + c;
#pragma unroll
        for (int i = 0; i < 8; ++i)
          if ((rmask >> i) & 1u)
            __stcg(reinterpret_cast<float4*>(pp + static_cast<long long>(4 * i) * args.partial_ld),
                   *reinterpret_cast<const float4*>(stg + (4 * i + lr) * kStgPitch + lc));
      } else if (args.epi_vec && nv >= 4 && !(GEMM_DBG(args) & 16)) {
        // fast path: everything 16-byte aligned and in range -> unrolled, branch-light, loads batched by the compiler
        float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f), cv4 = b4;
        if (has_bias) b4 = __ldg(reinterpret_cast<const float4*>(ep.bias + c));
        if (has_r1) cv4 = __ldg(reinterpret_cast<const float4*>(ep.cvec + c));
        float4 v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          v[i] = *reinterpret_cast<const float4*>(stg + (4 * i + lr) * kStgPitch + lc);
          v[i].x += b4.x; v[i].y += b4.y; v[i].z += b4.z; v[i].w += b4.w;
        }
        if (has_ra) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 t = __ldg(reinterpret_cast<const float4*>(ep.row_add + ra_off[i] + c));
            v[i].x += t.x; v[i].y += t.y; v[i].z += t.z; v[i].w += t.w;
          }
        }
        if (has_ra2) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 t = __ldg(reinterpret_cast<const float4*>(ep.row_add2 + ra2_off[i] + c));
            v[i].x += t.x; v[i].y += t.y; v[i].z += t.z; v[i].w += t.w;
          }
        }
        if (has_ra3) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 t = __ldg(reinterpret_cast<const float4*>(ep.row_add3 + ra3_off[i] + c));
            v[i].x += t.x; v[i].y += t.y; v[i].z += t.z; v[i].w += t.w;
          }
        }
        if (has_r1) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            v[i].x = fmaf(rvs[i], cv4.x, v[i].x); v[i].y = fmaf(rvs[i], cv4.y, v[i].y);
            v[i].z = fmaf(rvs[i], cv4.z, v[i].z); v[i].w = fmaf(rvs[i], cv4.w, v[i].w);
          }
        }
        if (ep.act == kActRelu) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            v[i].x = fmaxf(v[i].x, 0.f); v[i].y = fmaxf(v[i].y, 0.f); v[i].z = fmaxf(v[i].z, 0.f); v[i].w = fmaxf(v[i].w, 0.f);
          }
        } else if (ep.act == kActTanh) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            v[i].x = tanhf(v[i].x); v[i].y = tanhf(v[i].y); v[i].z = tanhf(v[i].z); v[i].w = tanhf(v[i].w);
          }
        }
        if (extras) {
          const long long mb = mw + lr;
          if (ep.drop_mask) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              if (!((rmask >> i) & 1u)) continue;
              const float4 t = __ldg(reinterpret_cast<const float4*>(ep.drop_mask + (mb + 4 * i) * ep.drop_mask_ld + c));
              v[i].x *= t.x; v[i].y *= t.y; v[i].z *= t.z; v[i].w *= t.w;
            }
          } else if (ep.drop_p > 0.f) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const uint32_t mu = static_cast<uint32_t>(mb + 4 * i), cu = static_cast<uint32_t>(c);
              v[i].x = drop_keep(drop_seed, mu, cu, ep.drop_p) ? v[i].x * drop_scale : 0.f;
              v[i].y = drop_keep(drop_seed, mu, cu + 1, ep.drop_p) ? v[i].y * drop_scale : 0.f;
              v[i].z = drop_keep(drop_seed, mu, cu + 2, ep.drop_p) ? v[i].z * drop_scale : 0.f;
              v[i].w = drop_keep(drop_seed, mu, cu + 3, ep.drop_p) ? v[i].w * drop_scale : 0.f;
            }
          }
          if (ep.gate) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              if (!((rmask >> i) & 1u)) continue;
              const float4 t = __ldg(reinterpret_cast<const float4*>(ep.gate + (mb + 4 * i) * ep.gate_ld + c));
              v[i].x = t.x > 0.f ? v[i].x * ep.gate_scale : 0.f; v[i].y = t.y > 0.f ? v[i].y * ep.gate_scale : 0.f;
              v[i].z = t.z > 0.f ? v[i].z * ep.gate_scale : 0.f; v[i].w = t.w > 0.f ? v[i].w * ep.gate_scale : 0.f;
            }
          }
          if (ep.alpha != 1.f) {
#pragma unroll
            for (int i = 0; i < 8; ++i) { v[i].x *= ep.alpha; v[i].y *= ep.alpha; v[i].z *= ep.alpha; v[i].w *= ep.alpha; }
          }
        }
        if (ep.c_half && !(GEMM_DBG(args) & 8)) {
          __half* hp_ = reinterpret_cast<__half*>(ep.c_half) + static_cast<long long>(mw + lr) * ep.half_ld + c;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            if (!((rmask >> i) & 1u)) continue;
            const __half2 h0 = __floats2half2_rn(v[i].x, v[i].y), h1 = __floats2half2_rn(v[i].z, v[i].w);
            uint2 pk;
            pk.x = *reinterpret_cast<const uint32_t*>(&h0); pk.y = *reinterpret_cast<const uint32_t*>(&h1);
            *reinterpret_cast<uint2*>(hp_ + static_cast<long long>(4 * i) * ep.half_ld) = pk;
          }
        }
        if (args.C && !(GEMM_DBG(args) & 1)) {
          float* cp = args.C + static_cast<long long>(mw + lr) * args.ldc + c;
          if (!args.atomic) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
              if ((rmask >> i) & 1u) *reinterpret_cast<float4*>(cp + static_cast<long long>(4 * i) * args.ldc) = v[i];
          } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              if (!((rmask >> i) & 1u)) continue;
              float* q = cp + static_cast<long long>(4 * i) * args.ldc;
              atomicAdd(q, v[i].x); atomicAdd(q + 1, v[i].y); atomicAdd(q + 2, v[i].z); atomicAdd(q + 3, v[i].w);
            }
          }
        }
      } else if (nv > 0 && !(GEMM_DBG(args) & 16)) {
        // careful path: ragged last columns or unaligned operands (element-wise guards)
        float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f), cv4 = b4;
        if (has_bias) b4 = load4(ep.bias + c, nv);
        if (has_r1) cv4 = load4(ep.cvec + c, nv);
#pragma unroll
        for (int i = 0; i < 8; ++i) {   // (unrolled: the per-row register arrays must keep static indices)
          if (!((rmask >> i) & 1u)) continue;
          const int rr = 4 * i + lr;
          const int m = mw + rr;
          float4 v = *reinterpret_cast<const float4*>(stg + rr * kStgPitch + lc);
          v.x += b4.x; v.y += b4.y; v.z += b4.z; v.w += b4.w;
          if (has_ra) {
            const float4 t = load4(ep.row_add + ra_off[i] + c, nv);
            v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w;
          }
          if (has_ra2) {
            const float4 t = load4(ep.row_add2 + ra2_off[i] + c, nv);
            v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w;
          }
          if (has_ra3) {
            const float4 t = load4(ep.row_add3 + ra3_off[i] + c, nv);
            v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w;
          }
          if (has_r1) { v.x = fmaf(rvs[i], cv4.x, v.x); v.y = fmaf(rvs[i], cv4.y, v.y); v.z = fmaf(rvs[i], cv4.z, v.z); v.w = fmaf(rvs[i], cv4.w, v.w); }
          if (ep.act == kActTanh) { v.x = tanhf(v.x); v.y = tanhf(v.y); v.z = tanhf(v.z); v.w = tanhf(v.w); }
          else if (ep.act == kActRelu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
          if (ep.drop_mask) {
            const float4 t = load4(ep.drop_mask + static_cast<long long>(m) * ep.drop_mask_ld + c, nv);
            v.x *= t.x; v.y *= t.y; v.z *= t.z; v.w *= t.w;
          } else if (ep.drop_p > 0.f) {
            const uint32_t mu = static_cast<uint32_t>(m), cu = static_cast<uint32_t>(c);
            v.x = drop_keep(drop_seed, mu, cu, ep.drop_p) ? v.x * drop_scale : 0.f;
            v.y = drop_keep(drop_seed, mu, cu + 1, ep.drop_p) ? v.y * drop_scale : 0.f;
            v.z = drop_keep(drop_seed, mu, cu + 2, ep.drop_p) ? v.z * drop_scale : 0.f;
            v.w = drop_keep(drop_seed, mu, cu + 3, ep.drop_p) ? v.w * drop_scale : 0.f;
          }
          if (ep.gate) {
            const float4 t = load4(ep.gate + static_cast<long long>(m) * ep.gate_ld + c, nv);
            v.x = t.x > 0.f ? v.x * ep.gate_scale : 0.f; v.y = t.y > 0.f ? v.y * ep.gate_scale : 0.f;
            v.z = t.z > 0.f ? v.z * ep.gate_scale : 0.f; v.w = t.w > 0.f ? v.w * ep.gate_scale : 0.f;
          }
          if (ep.alpha != 1.f) { v.x *= ep.alpha; v.y *= ep.alpha; v.z *= ep.alpha; v.w *= ep.alpha; }
          if (ep.c_half) {
            __half* hp_ = reinterpret_cast<__half*>(ep.c_half) + static_cast<long long>(m) * ep.half_ld + c;
            hp_[0] = __float2half_rn(v.x);
            if (nv > 1) hp_[1] = __float2half_rn(v.y);
            if (nv > 2) hp_[2] = __float2half_rn(v.z);
            if (nv > 3) hp_[3] = __float2half_rn(v.w);
          }
          if (args.C) {
            float* cp = args.C + static_cast<long long>(m) * args.ldc + c;
            if (args.atomic) {
              atomicAdd(cp, v.x);
              if (nv > 1) atomicAdd(cp + 1, v.y);
              if (nv > 2) atomicAdd(cp + 2, v.z);
              if (nv > 3) atomicAdd(cp + 3, v.w);
            } else {
              cp[0] = v.x;
              if (nv > 1) cp[1] = v.y;
              if (nv > 2) cp[2] = v.z;
              if (nv > 3) cp[3] = v.w;
            }
          }
        }
      }
      __syncwarp();   // the staging rows are rewritten by the next chunk
    }
    {
      // constant-one column / zero padding of the fp16 copy (thread = row; once per row, by the tile that holds column N-1)
      const int m = mw + lane;
      if (par == 0 && ep.c_half && !args.partial && m < args.M && n0 < args.N && n0 + bn >= args.N) {
        __half* hrow = reinterpret_cast<__half*>(ep.c_half) + static_cast<long long>(m) * ep.half_ld;
        for (int j = 0; j < ep.half_tail; ++j) hrow[args.N + j] = __float2half_rn(j == 0 ? 1.f : 0.f);
        for (int j = args.N; j < ep.half_pad_to; ++j) hrow[j] = __float2half_rn(0.f);
      }
    }
    // this accumulator may be overwritten by the MMAs of the tile after next
    tc_fence_before();
    __syncwarp();
    if (lane == 0) {
      if constexpr (kPair) mbar_arrive_cluster_relaxed(mapa_u32(smem_u32(&aux->tempty[acc]), 0));   // the leader's barrier
      else mbar_arrive(&aux->tempty[acc]);
    }
    if (args.dual) {
      if (half == 1) acc_phase ^= 1;
    } else {
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
    t_wait += t_e1 - t_e0;
    t_work += clock64() - t_e1;
    }
    }
    if (GEMM_PROF(args) && threadIdx.x == 128) {
      unsigned long long* pr = GEMM_PROF(args) + 8 * blockIdx.x;
      pr[4] = static_cast<unsigned long long>(t_wait);
      pr[5] = static_cast<unsigned long long>(t_work);
      pr[6] = static_cast<unsigned long long>(clock64() - t_entry);
    }
  }

  tc_fence_before();
  if constexpr (kPair) {   // both CTAs must be done with each other's shared memory / TMEM before either exits or frees
    cluster_sync_all();
    if (warp == 2) tmem_dealloc_pair(tmem_base, args.tmem_cols);
  } else {
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, args.tmem_cols);
  }
}

__global__ void __launch_bounds__(128 + 32 * kEpiWarps, 1) gemm_tf32_kernel(const __grid_constant__ KArgs args) {
  gemm_tf32_body<false>(args);
}
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128 + 32 * kEpiWarps, 1)
gemm_pair_kernel(const __grid_constant__ KArgs args) {
  gemm_tf32_body<true>(args);
}

// Deterministic split-K, stage 2: C = epilogue( sum_z partial[z] ), z in ascending order. One thread per 4 consecutive
// columns of a row (coalesced 16 B loads of every partial; they are L2-resident, written microseconds earlier).
struct SplitEpiArgs {
  const float* partial;
  long long partial_ld, partial_stride;
  int split, M, N;
  float* C;
  long long ldc;
  int accumulate;
  GemmEpilogue epi;
};

__global__ void __launch_bounds__(256) gemm_splitk_epilogue_kernel(const SplitEpiArgs a) {
  pdl_enter();   // programmatic dependent launch: see ptx.cuh
  const int n4 = (a.N + 3) / 4;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<long long>(a.M) * n4) return;
  const int m = static_cast<int>(i / n4), c0 = static_cast<int>(i % n4) * 4;
  const GemmEpilogue& ep = a.epi;
  const float* src = a.partial + static_cast<long long>(m) * a.partial_ld + c0;
  float4 acc = __ldcg(reinterpret_cast<const float4*>(src));
  for (int z = 1; z < a.split; ++z) {
    const float4 t = __ldcg(reinterpret_cast<const float4*>(src + z * a.partial_stride));
    acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
  }
  float v[4] = {acc.x, acc.y, acc.z, acc.w};
  const float drop_scale = ep.drop_p > 0.f ? 1.f / (1.f - ep.drop_p) : 1.f;
  const unsigned long long drop_seed = ep.drop_seed + ((ep.drop_seed_dev && ep.drop_p > 0.f) ? __ldg(ep.drop_seed_dev) : 0ull);
  float* crow = a.C ? a.C + static_cast<long long>(m) * a.ldc : nullptr;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int c = c0 + j;
    if (c >= a.N) break;
    float x = v[j];
    if (ep.bias) x += __ldg(ep.bias + c);
    if (ep.row_add) x += __ldg(ep.row_add + static_cast<long long>(m / ep.row_add_div) * ep.row_add_ld + c);
    if (ep.row_add2) x += __ldg(ep.row_add2 + static_cast<long long>(m % ep.row_add2_mod) * ep.row_add2_ld + c);
    if (ep.row_add3) x += __ldg(ep.row_add3 + static_cast<long long>(__ldg(ep.row_add3_idx + m)) * ep.row_add3_ld + c);
    if (ep.rvec) x = fmaf(__ldg(ep.rvec + m), __ldg(ep.cvec + c), x);
    x = apply_act(x, ep.act);
    if (ep.drop_mask) x *= __ldg(ep.drop_mask + static_cast<long long>(m) * ep.drop_mask_ld + c);
    else if (ep.drop_p > 0.f)
      x = drop_keep(drop_seed, static_cast<uint32_t>(m), static_cast<uint32_t>(c), ep.drop_p) ? x * drop_scale : 0.f;
    if (ep.gate) x = __ldg(ep.gate + static_cast<long long>(m) * ep.gate_ld + c) > 0.f ? x * ep.gate_scale : 0.f;
    x *= ep.alpha;
    if (ep.c_half) {
      __half* hrow = reinterpret_cast<__half*>(ep.c_half) + static_cast<long long>(m) * ep.half_ld;
      hrow[c] = __float2half_rn(x);
      if (c == a.N - 1) {
        for (int t = 0; t < ep.half_tail; ++t) hrow[a.N + t] = __float2half_rn(t == 0 ? 1.f : 0.f);
        for (int t = a.N; t < ep.half_pad_to; ++t) hrow[t] = __float2half_rn(0.f);
      }
    }
    if (!a.C) continue;
    if (a.accumulate) crow[c] += x;
    else crow[c] = x;
  }
}

// Balanced N tile: the fewest tiles of at most 256 columns, each rounded up to 32 (N = 300 -> 2 x 160, not 256 + 44).
static int balanced_bn(int N) {
  const int nt = (N + 255) / 256;
  const int bn = ((N + nt - 1) / nt + 31) / 32 * 32;
  return bn < 32 ? 32 : (bn > 256 ? 256 : bn);
}

cudaError_t launch_gemm_tf32(const GemmProblem& p, cudaStream_t s) {
  if (p.M <= 0 || p.N <= 0) return cudaSuccess;
  if (p.nseg < 1 || p.nseg > kGemmMaxSeg || (!p.C && !p.epi.c_half)) return cudaErrorInvalidValue;
  if (!p.C && (p.accumulate || (p.split_k > 1 && !p.scratch))) return cudaErrorInvalidValue;
  if ((p.epi.rvec != nullptr) != (p.epi.cvec != nullptr)) return cudaErrorInvalidValue;
  if ((p.epi.row_add3 != nullptr) != (p.epi.row_add3_idx != nullptr)) return cudaErrorInvalidValue;
  KArgs a{};
  const int mt = (p.M + BM - 1) / BM;
  const int bk_el = p.f16 ? 2 * BK : BK;        // elements per 128-byte k-block
  const int el_bytes = p.f16 ? 2 : 4;
  int kb_all = 0;
  for (int i = 0; i < p.nseg; ++i) kb_all += (p.seg[i].k + bk_el - 1) / bk_el;
  const bool has_scratch = p.scratch != nullptr && p.scratch_floats > 0;
  const long long n_pad32 = (p.N + 31) / 32 * 32;
  int split = p.split_k < 0 ? 1 : p.split_k;
  int bn = p.bn;
  if (split == 0 || bn == 0) {
    // Tile / split choice by a small cost model (cycles). These products are bound by the L2 -> shared-memory operand
    // stream (fp32 operands: 16 KB + 128 B x bn per k-block), ~5400 B/clk chip-wide and <= ~70 B/clk for one SM, so the
    // model trades total traffic (wide tiles) against the number of CTAs pulling it (split-K with the stage-2 kernel).
    double best = 1e30;
    int best_bn = 32, best_split = 1;
    const int bal = balanced_bn(p.N);
    const int bn_cands[5] = {32, 64, 128, bal, 256};
    const int sp_cands[12] = {1, 2, 3, 4, 6, 8, 10, 12, 16, 20, 24, 32};
    for (int bi = 0; bi < 5; ++bi) {
      const int bc = bn != 0 ? bn : bn_cands[bi];
      if (bn != 0 && bi > 0) break;
      if (bn == 0 && bc > n_pad32 && bc != 32 && bc != bal) continue;
      const int ntn = (p.N + bc - 1) / bc;
      for (int si = 0; si < 12; ++si) {
        const int sc = split != 0 ? split : sp_cands[si];
        if (split != 0 && si > 0) break;
        if (sc > 1) {
          if (split == 0 && sc * 2 > kb_all) continue;
          if (!has_scratch && (p.epi.act != kActNone)) continue;
          if (has_scratch && static_cast<size_t>(sc) * p.M * n_pad32 > p.scratch_floats) continue;
          if (!has_scratch && split == 0) continue;  // never choose atomics on our own
        }
        const long long ctas = static_cast<long long>(mt) * ntn * sc;
        const double conc = static_cast<double>(ctas < 148 ? ctas : 148);
        const double bpc = std::min(70.0, 5400.0 / conc);
        // MMA time of a k-block: 4 instructions of N = bc (TF32 K = 8, fp16 K = 16: the same issue count per 128 B)
        const double per_kb = std::max(2.0 * bc, (16384.0 + 128.0 * bc) / bpc);
        const double kbc = std::ceil(static_cast<double>(kb_all) / sc);
        const double epi = sc > 1 ? 800.0 + 300.0 * (bc / 32) : 1500.0 + 600.0 * (bc / 32);
        const double waves = std::ceil(ctas / 148.0);
        double cost = waves * (1500.0 + kbc * per_kb + epi) + 4000.0;
        if (sc > 1) cost += 7000.0 + static_cast<double>(p.M) * p.N * sc * 4.0 / 5400.0;
        if (cost < best) { best = cost; best_bn = bc; best_split = sc; }
      }
    }
    if (best >= 1e30) return cudaErrorInvalidValue;
    bn = best_bn;
    split = best_split;
  }
  if (bn % 32 != 0 || bn < 32 || bn > 256) return cudaErrorInvalidValue;
  // CTA-pair mode (KArgs::pair): large-M fp16 products whose weight tile every CTA would re-read per m-tile
  // (short-K products with many n-tiles are bound by their epilogue, not by operands: the classifier of an eval batch,
  // K = 384, N = 2000, measured 225 us single / 237 us paired -- pairs only where the operand stream is the bound)
  const int ntn_plain = (p.N + bn - 1) / bn;
  bool pair = p.f16 && split == 1 && !p.accumulate && mt >= 2 * 148 && (kb_all >= 16 || ntn_plain == 2) &&
              !opt(kOptGemmNoPair);
  for (int i = 0; i < p.nseg; ++i) pair = pair && !p.seg[i].a.mn_major && !p.seg[i].b.mn_major;
  const int b_box = pair ? bn / 2 : bn;
  int kb = 0;
  for (int i = 0; i < p.nseg; ++i) {
    const GemmSegment& g = p.seg[i];
    const int ld_mult = p.f16 ? 8 : 4;          // 16-byte rows
    if (g.k <= 0 || !g.a.ptr || !g.b.ptr || (g.a.ld % ld_mult) || (g.b.ld % ld_mult) ||
        (reinterpret_cast<uintptr_t>(g.a.ptr) & 15) || (reinterpret_cast<uintptr_t>(g.b.ptr) & 15))
      return cudaErrorInvalidValue;
    if (p.f16 && (g.a.mn_major || g.b.mn_major)) return cudaErrorInvalidValue;   // fp16 operands: k-major only
    int rc;
    if (!g.a.mn_major) rc = make_tmap_2d(&a.ta[i], g.a.ptr, el_bytes, p.M, g.k, g.a.ld * el_bytes, BM);
    else rc = make_tmap_2d(&a.ta[i], g.a.ptr, 4, g.k, p.M, g.a.ld * 4, 32, true);
    if (rc != 0) return cudaErrorInvalidValue;
    if (!g.b.mn_major) rc = make_tmap_2d(&a.tb[i], g.b.ptr, el_bytes, p.N, g.k, g.b.ld * el_bytes, b_box);
    else rc = make_tmap_2d(&a.tb[i], g.b.ptr, 4, g.k, p.N, g.b.ld * 4, 32, true);
    if (rc != 0) return cudaErrorInvalidValue;
    kb += (g.k + bk_el - 1) / bk_el;
    a.kb_end[i] = kb;
    a.a_mn[i] = g.a.mn_major ? 1 : 0;
    a.b_mn[i] = g.b.mn_major ? 1 : 0;
  }
  if (split > kb) split = kb;
  const bool partial_mode = split > 1 && has_scratch && !p.accumulate &&
                            static_cast<size_t>(split) * p.M * n_pad32 <= p.scratch_floats;
  if (split > 1 && !partial_mode && p.epi.act != kActNone) return cudaErrorInvalidValue;
  a.nseg = p.nseg;
  a.f16 = p.f16 ? 1 : 0;
  a.M = p.M; a.N = p.N; a.bn = bn;
  const int stage_bytes = kABytes + bn * 128;
  a.ntn = (p.N + bn - 1) / bn;
  // both n-tiles of a two-tile-wide product from one staged A tile, when the m-tiles alone fill the chip
  // (not in pair mode: there the two n-tiles run one after the other on ping-pong accumulators, so the epilogue of one
  // overlaps the mainloop of the next -- in dual mode both accumulators belong to one tile and nothing overlaps -- and
  // the A tile's second trip costs little once B is shared. Eval batch of 4096: 1.300 -> 1.205 ms.)
  a.dual = (!pair && a.ntn == 2 && static_cast<long long>(mt) * split >= 2 * 148 &&
            kSmemBudget / (kABytes + 2 * bn * 128) >= 3) ? 1 : 0;
  a.pair = pair ? 1 : 0;
  const int stage_bytes_k = kABytes + (a.dual ? 2 : 1) * b_box * 128;
  a.stages = kSmemBudget / stage_bytes_k;
  if (a.stages > kMaxStages) a.stages = kMaxStages;
  a.tmem_cols = 64;
  while (a.tmem_cols < 2 * bn) a.tmem_cols <<= 1;   // two accumulators of bn columns
  if (a.dual) a.ntn = 1;     // a tile of the loop is the PAIR of n-tiles
  a.ntm = pair ? (mt + 1) / 2 : mt;
  a.split = split;
  a.total_tiles = a.ntn * a.ntm * split;
  a.C = p.C; a.ldc = p.ldc;
  a.atomic = (!partial_mode && (split > 1 || p.accumulate)) ? 1 : 0;
  a.epi = p.epi;
  {
    auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    const GemmEpilogue& e = p.epi;
    a.epi_vec = (!p.C || (al16(p.C) && p.ldc % 4 == 0)) && (!e.bias || al16(e.bias)) && (!e.cvec || al16(e.cvec)) &&
                (!e.row_add || (al16(e.row_add) && e.row_add_ld % 4 == 0)) &&
                (!e.row_add2 || (al16(e.row_add2) && e.row_add2_ld % 4 == 0)) &&
                (!e.row_add3 || (al16(e.row_add3) && e.row_add3_ld % 4 == 0)) &&
                (!e.drop_mask || (al16(e.drop_mask) && e.drop_mask_ld % 4 == 0)) &&
                (!e.gate || (al16(e.gate) && e.gate_ld % 4 == 0)) &&
                (!e.c_half || ((reinterpret_cast<uintptr_t>(e.c_half) & 7) == 0 && e.half_ld % 4 == 0));
  }
  if (partial_mode) {
    a.partial = p.scratch;
    a.partial_ld = n_pad32;
    a.partial_stride = static_cast<long long>(p.M) * n_pad32;
  }
  a.prof = knn_debug_profile_buffer();
  a.debug = opt(kOptGemmDebug);   // 0 unless this is a -DVQACX_DEBUG build
  const int smem = a.stages * stage_bytes_k + kAuxBytes + kStgBytes + 1024;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tf32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kSmemBudget + kAuxBytes + kStgBytes + 1024);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  dim3 grid(a.total_tiles < 148 ? a.total_tiles : 148);   // persistent: one CTA per SM, tiles strided over them
  count_launch();
  cudaError_t err;
  if (pair) {   // 74 CTA pairs (clusters of 2: the two SMs of a TPC), pair-tiles strided over them
    static bool configured_pair = false;
    if (!configured_pair) {
      cudaError_t e = cudaFuncSetAttribute(gemm_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           kSmemBudget + kAuxBytes + kStgBytes + 1024);
      if (e != cudaSuccess) return e;
      configured_pair = true;
    }
    err = launch_pdl(gemm_pair_kernel, dim3(2 * (a.total_tiles < 74 ? a.total_tiles : 74)), dim3(128 + 32 * kEpiWarps),
                     static_cast<size_t>(smem), s, a);
  } else {
    err = launch_pdl(gemm_tf32_kernel, grid, dim3(128 + 32 * kEpiWarps), static_cast<size_t>(smem), s, a);
  }
  if (err == cudaSuccess) err = cudaGetLastError();
  if (err != cudaSuccess || !partial_mode) return err;
  SplitEpiArgs e{};
  e.partial = a.partial; e.partial_ld = a.partial_ld; e.partial_stride = a.partial_stride;
  e.split = split; e.M = p.M; e.N = p.N; e.C = p.C; e.ldc = p.ldc; e.accumulate = 0; e.epi = p.epi;
  const long long items = static_cast<long long>(p.M) * ((p.N + 3) / 4);
  count_launch();
  (void)launch_pdl(gemm_splitk_epilogue_kernel, dim3(static_cast<unsigned>((items + 255) / 256)), dim3(256), 0, s, e);
  return cudaGetLastError();
}

}  // namespace vqacx

// Grouped expert GEMMs on the 5th-generation tensor cores (tcgen05 + TMEM + TMA), sm_100a only.
//
// Replaces the per-expert F.linear calls of the reference (mlp.py:50,52 -> linear.py:75,152),
// with SiLU-and-mul (activation.py:13-14) fused into the gate_up epilogue and the routing-weight
// scale (moe.py:149) fused into the down epilogue.
//
// Orientation ("swap-AB"): the *weights* are the M=128 operand and the *token rows* are the N
// operand, D[feature, token] = sum_k W[e][feature, k] * X[row(token), k].  An expert with n_e
// rows costs ceil(n_e / BN) MMAs of N = BN (16..128) per weight tile, so the decode regime
// (n_e = 1..8) streams every weight byte through shared memory exactly once while the tensor
// pipe stays far from its limit, and the prefill regime uses the same kernel with BN = 128.
//
// CTA = 8 warps, warp-specialised, persistent over a device-side work list (tile x feature-tile):
//   warp 0  TMA producer: weights [128 x 64] (+ the matching up_proj tile when GATED) and the
//           activation tile [BN x 64] per k-block into a kStages-deep 128B-swizzled smem ring
//           (a `tile::gather4` variant that read the un-permuted tokens directly was measured
//           2.4-4x slower at 256-row tiles and only tied at decode: profiles/r1_gather4.md)
//   warp 1  MMA issuer (one elected lane): 4 x tcgen05.mma (K=16) per k-block per accumulator
//   warp 2  TMEM allocator / deallocator
//   warps 4-7 epilogue: tcgen05.ld 32 lanes x 16 columns, fused math, coalesced 2-byte stores
//           (lane = output feature, so a warp writes 64 contiguous bytes per token)
// Pipelines: smem full/empty mbarriers (TMA <-> MMA) and 2 TMEM accumulator stages
// (MMA <-> epilogue) so the epilogue of tile i overlaps the MMAs of tile i+1.
#include "umma_common.cuh"
#include "umma_gemm.cuh"

namespace glm4moe {

constexpr int kBK = 64;    // 64 two-byte elements = 128 B swizzle atom

// Epilogue modes
constexpr int kModeDown = 0;    // y = acc * routing weight              (mlp.py:52, moe.py:149)
constexpr int kModeGated = 1;   // act = silu(acc_gate) * acc_up         (mlp.py:50-51)
constexpr int kModePlain = 2;   // out = acc (router logits, moe.py:32), features bounded by n_feat_valid
// Down with TWO 128-feature weight tiles per CTA tile (256 features x BN tokens): the token tile is
// fetched once for both, so the shared-memory fill per MMA cycle drops from 96 to 64 B/clk at BN = 256
// (the plain down tile is bound by that fill rate at prefill: profiles/r1d.md).  The two accumulators
// fill TMEM like the gated tile's (one accumulator stage).
constexpr int kModeDown2 = 3;

template <int BN, int MODE, int BK = kBK>
struct UmmaCfg {
  static constexpr bool kGated = MODE == kModeGated || MODE == kModeDown2;   // two weight tiles per stage
  static constexpr int kFeatPerTile = MODE == kModeDown2 ? 2 * kBM : kBM;
  static constexpr int kABytes = kBM * BK * 2;                  // 16 KB per weight tile (8 KB at BK = 32)
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kStageBytes = (kGated ? 2 : 1) * kABytes + kBBytes;
  // BK = 32 (64-byte swizzle atoms): half-size stages, ring sized to the whole 227 KB.  The 256-token
  // two-weight-tile stage is 64 KB at BK = 64, i.e. 3 stages = 2 in flight while one is consumed
  // (1.4 us of look-ahead; ncu: the MMA warp waits for data 27 % of the time); 7 x 32 KB keeps 6 in
  // flight (2.1 us).
  static constexpr int kSmemBudget = BK == 32 ? 232448 - 1536 : 212 * 1024;
  static constexpr int kStages = kSmemBudget / kStageBytes > 12 ? 12 : kSmemBudget / kStageBytes;
  static constexpr int kAccCols = (kGated ? 2 : 1) * BN;        // TMEM columns per accumulator stage
  static constexpr int kAccStages = 2 * kAccCols <= 512 ? 2 : 1;   // BN=256 gated fills TMEM: 1 stage
  static constexpr int kTmemColsRaw = kAccStages * kAccCols;
  static constexpr int kTmemCols = kTmemColsRaw <= 32 ? 32 : kTmemColsRaw <= 64 ? 64
                                   : kTmemColsRaw <= 128 ? 128 : kTmemColsRaw <= 256 ? 256 : 512;
  // wide tiles: two warps per TMEM lane quadrant; the gated 256-row tile has a single accumulator
  // stage (TMEM is full), so its epilogue is on the MMA's critical path: four warps per quadrant
  static constexpr int kEpiWarps = (kGated && BN == 256) ? 16 : BN >= 64 ? 8 : 4;
  static constexpr int kThreads = 128 + 32 * kEpiWarps;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 512 /*barriers*/;
  static_assert(kTmemColsRaw <= 512, "accumulators do not fit TMEM");
  static_assert(kBBytes % 1024 == 0, "operand tiles must stay 1024-byte aligned");
  static_assert(kStages >= 3, "pipeline too shallow");
};


// ------------------------------------------------------------------------------------------
// the kernel
// ------------------------------------------------------------------------------------------
// Activation tensor maps with fewer box rows than BN over the same buffer (ascending `rows`, 0 = unused).
// The producer picks the smallest box that covers a tile's rows: the ragged last token tile of an
// expert (256 +- 16 rows per expert at Air T = 4096: half of the experts own a 1..48-row second tile)
// then brings its rows across L2 -> SM, not a whole BN-row box.
struct TailMaps {
  CUtensorMap m[3];
  int rows[3];
  int pad;
};

constexpr int kSharedWorkBase = 1 << 30;   // direct mode: work id of the shared expert's feature tile f
constexpr int kSchedDepth = 4;   // work-index queue between the producer and the consumer warps

template <typename T, int BN, int MODE, int BK = kBK>
__global__ void __launch_bounds__(UmmaCfg<BN, MODE, BK>::kThreads, 1) umma_grouped_kernel(
    const __grid_constant__ CUtensorMap tm_w,   // weights  [rows_w_total, K], box {64, 128}
    const __grid_constant__ CUtensorMap tm_x,   // activation rows [rows_x_total, K], box {64, BN}
    const __grid_constant__ TailMaps tail,      // same rows, smaller boxes
    const int4* __restrict__ tiles, const int32_t* __restrict__ num_tiles_ptr, int dense_rows,
    T* __restrict__ out, int ld_out,            // out[row, feature]
    const float* __restrict__ row_weight,       // kModeDown: per-row scale for rows < n_routed_rows
    int n_routed_rows, int n_feat_tiles, int n_feat_valid, int n_kblocks,
    int w_rows_per_expert,                      // gated: 2I, down: H, plain: 0
    int up_row_offset,                          // gated: I
    int32_t* __restrict__ sched,                // [0] next work index - gridDim, [1] CTAs finished
    const DirectArgs direct) {                  // decode "direct" mode (tokens > 0), see umma_gemm.cuh
  using Cfg = UmmaCfg<BN, MODE, BK>;
  constexpr bool GATED = Cfg::kGated;
  constexpr int AS = Cfg::kAccStages;
  auto kdesc = [](uint32_t saddr) { return BK == 32 ? make_kmajor_desc_sw64(saddr) : make_kmajor_desc(saddr); };
  constexpr int EW = Cfg::kEpiWarps;
  extern __shared__ unsigned char smem_dyn[];
  unsigned char* smem = reinterpret_cast<unsigned char*>(
      (reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + Cfg::kStages;
  uint64_t* tmem_full = bars + 2 * Cfg::kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint64_t* sched_full = tmem_empty + 2;
  uint64_t* sched_empty = sched_full + kSchedDepth;
  int32_t* sched_w = reinterpret_cast<int32_t*>(sched_empty + kSchedDepth);   // [depth][2]: work item, n_tiles
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(sched_w + 2 * kSchedDepth);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tm_w);
    prefetch_tmap(&tm_x);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], EW);   // one arrive per epilogue warp
    }
    for (int s = 0; s < kSchedDepth; ++s) {
      mbar_init(&sched_full[s], 1);
      mbar_init(&sched_empty[s], 1 + EW);   // MMA warp + epilogue warps
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<Cfg::kTmemCols>(tmem_ptr_smem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  // consumer side of the work queue: n-th work item of this CTA (-1 = done) and the tile count
  auto next_work = [&](int n, int& n_tiles) -> int {
    const int q = n % kSchedDepth;
    mbar_wait(&sched_full[q], (n / kSchedDepth) & 1u);
    const int w = sched_w[2 * q];
    n_tiles = sched_w[2 * q + 1];
    __syncwarp();
    if (lane == 0) mbar_arrive(&sched_empty[q]);
    return w;
  };

  if (warp == 0) {
    // ===================== TMA producer + dynamic tile scheduler =====================
    if (elect_one()) {
      const int n_tiles = direct.tokens > 0 ? 0 : tiles ? *num_tiles_ptr : (dense_rows + BN - 1) / BN;
      const int total_work = n_tiles * n_feat_tiles;
      int n_pushed = 0;
      bool first = true;
      // dynamic schedule: the first item is static (blockIdx.x); later ones come from a global counter so
      // CTAs always take the lowest unprocessed index: balanced, and neighbours in time share
      // activation / weight tiles in L2.
      auto fetch_push = [&]() -> int {
        int w = first ? static_cast<int>(blockIdx.x) : atomicAdd(&sched[0], 1) + static_cast<int>(gridDim.x);
        first = false;
        if (w >= total_work) w = -1;
        const int q = n_pushed % kSchedDepth;
        mbar_wait(&sched_empty[q], ((n_pushed / kSchedDepth) & 1u) ^ 1u);
        sched_w[2 * q] = w;
        sched_w[2 * q + 1] = n_tiles;
        mbar_arrive(&sched_full[q]);
        ++n_pushed;
        return w;
      };
      WorkRef wr{0, 0, 0, 0};
      const bool is_direct = direct.tokens > 0;
      const CUtensorMap* mx = &tm_x;
      uint32_t x_bytes = Cfg::kBBytes;
      auto pick_x = [&]() {   // after every change of `wr` (not in direct mode: the box is the token count)
        const int need = max(16, (wr.nrows + 15) & ~15);
        mx = &tm_x;
        x_bytes = Cfg::kBBytes;
        if (tail.rows[2] >= need) { mx = &tail.m[2]; x_bytes = tail.rows[2] * BK * 2; }
        if (tail.rows[1] >= need) { mx = &tail.m[1]; x_bytes = tail.rows[1] * BK * 2; }
        if (tail.rows[0] >= need) { mx = &tail.m[0]; x_bytes = tail.rows[0] * BK * 2; }
      };
      auto issue = [&](int kb, int stage, bool do_w, bool do_x) {
        unsigned char* sa = smem + stage * Cfg::kStageBytes;
        if (do_w) {
          const int w_row = wr.expert * w_rows_per_expert + wr.f * Cfg::kFeatPerTile;
          mbar_expect_tx(&full_bar[stage], (GATED ? 2 : 1) * Cfg::kABytes + x_bytes);
          tma_load_2d(&tm_w, &full_bar[stage], sa, kb * BK, w_row, kEvictFirst);
          if (GATED)
            tma_load_2d(&tm_w, &full_bar[stage], sa + Cfg::kABytes, kb * BK, w_row + up_row_offset, kEvictFirst);
        }
        if (do_x)
          tma_load_2d(mx, &full_bar[stage], sa + (GATED ? 2 : 1) * Cfg::kABytes, kb * BK,
                      is_direct ? 0 : wr.row_start, kEvictLast);
      };
      if (is_direct) {
        // Decode "direct" mode: the token tile is the layer input itself (all T <= BN tokens are the
        // MMA columns of every expert), so nothing here depends on the routing except WHICH experts
        // run.  The shared expert's feature tiles therefore stream before the dependency wait,
        // overlapping the router; the routed items come from the global counter afterwards.
        int stage = 0;
        uint32_t phase = 0;
        auto stream_item = [&]() {
          for (int kb = 0; kb < n_kblocks; ++kb) {
            mbar_wait(&empty_bar[stage], phase ^ 1u);
            issue(kb, stage, true, true);
            if (++stage == Cfg::kStages) { stage = 0; phase ^= 1u; }
          }
        };
        auto push = [&](int w, int nt) {
          const int q = n_pushed % kSchedDepth;
          mbar_wait(&sched_empty[q], ((n_pushed / kSchedDepth) & 1u) ^ 1u);
          sched_w[2 * q] = w;
          sched_w[2 * q + 1] = nt;
          mbar_arrive(&sched_full[q]);
          ++n_pushed;
        };
        const int n_sh_items = direct.sh_hi > direct.sh_lo ? n_feat_tiles : 0;
        for (int f = blockIdx.x; f < n_sh_items; f += gridDim.x) {
          wr = WorkRef{direct.shared_expert, 0, 0, f};
          push(kSharedWorkBase + f, 0);
          stream_item();
        }
        pdl_wait();
        pdl_launch();
        const int n_routed_tiles = *num_tiles_ptr - (n_sh_items ? 1 : 0);
        const int total_routed = n_routed_tiles * n_feat_tiles;
        for (;;) {
          const int w = atomicAdd(&sched[0], 1);
          if (w >= total_routed) {
            push(-1, 0);
            break;
          }
          push(w, n_routed_tiles);
          wr = get_work<BN>(tiles, w, n_feat_tiles, n_routed_tiles, 0);
          stream_item();
        }
      } else {
      int w = fetch_push();
      if (w >= 0) {
        wr = get_work<BN>(tiles, w, n_feat_tiles, n_tiles, dense_rows);
        pick_x();
      }
      // (1) Before the dependency wait: weight tiles of the first item for the first pass over
      //     the ring.  Weights and the work list are at least two launches old (see common.cuh),
      //     so this overlaps the predecessor's tail with our pipeline fill.
      int n_pre = 0;
      if (w >= 0)
        for (; n_pre < Cfg::kStages && n_pre < n_kblocks; ++n_pre) issue(n_pre, n_pre, true, false);
      pdl_wait();
      pdl_launch();
      // (2) the activation tiles of those stages (written by the direct predecessor)
      for (int i = 0; i < n_pre; ++i) issue(i, i, false, true);
      int kb = n_pre;
      int stage = n_pre == Cfg::kStages ? 0 : n_pre;
      uint32_t phase = n_pre == Cfg::kStages ? 1u : 0u;
      // (3) steady state
      while (w >= 0) {
        if (kb == n_kblocks) {
          w = fetch_push();
          if (w >= 0) {
            wr = get_work<BN>(tiles, w, n_feat_tiles, n_tiles, dense_rows);
            pick_x();
          }
          kb = 0;
          continue;
        }
        mbar_wait(&empty_bar[stage], phase ^ 1u);
        issue(kb, stage, true, true);
        ++kb;
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1u; }
      }
      }
      // last CTA out re-arms the counters for the next launch
      __threadfence();
      if (atomicAdd(&sched[1], 1) == static_cast<int>(gridDim.x) - 1) {
        sched[0] = 0;
        sched[1] = 0;
        __threadfence();
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc_base = make_idesc(sizeof(T) == 2 && !std::is_same<T, __half>::value, 0);
    int stage = 0;
    uint32_t phase = 0;
    for (int it = 0;; ++it) {
      int n_tiles;
      const int w = next_work(it, n_tiles);
      if (w < 0) break;
      // MMA N follows the rows this tile really holds (multiple of 16): a ragged tail tile of an
      // expert costs tensor time in proportion to its rows, not to BN.  Direct mode: all T tokens.
      int n_eff;
      if (direct.tokens > 0) {
        n_eff = min(BN, (direct.tokens + 15) & ~15);
      } else {
        const WorkRef wr = get_work<BN>(tiles, w, n_feat_tiles, n_tiles, dense_rows);
        n_eff = min(BN, (wr.nrows + 15) & ~15);
      }
      const uint32_t idesc = idesc_base | (static_cast<uint32_t>(n_eff >> 3) << 17);
      const int as = AS == 2 ? (it & 1) : 0;
      const uint32_t aphase = AS == 2 ? ((it >> 1) & 1u) : (it & 1u);
      mbar_wait(&tmem_empty[as], aphase ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * Cfg::kAccCols;
      for (int kb = 0; kb < n_kblocks; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(smem + stage * Cfg::kStageBytes);
          const uint64_t a_desc = kdesc(sa);
          const uint64_t u_desc = kdesc(sa + Cfg::kABytes);
          const uint64_t b_desc = kdesc(sa + (GATED ? 2 : 1) * Cfg::kABytes);
#pragma unroll
          for (int ks = 0; ks < BK / 16; ++ks) {
            const uint32_t acc = (kb > 0 || ks > 0) ? 1u : 0u;
            umma_f16(d_tmem, a_desc + 2 * ks, b_desc + 2 * ks, idesc, acc);
            if (GATED) umma_f16(d_tmem + BN, u_desc + 2 * ks, b_desc + 2 * ks, idesc, acc);
          }
          umma_commit(&empty_bar[stage]);                          // smem slot free when MMAs finish
          if (kb == n_kblocks - 1) umma_commit(&tmem_full[as]);    // accumulator ready
        }
        __syncwarp();
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    // EW = 4: warp q owns TMEM lanes [32q, 32q+32) and all token columns.
    // EW = 8: two warps share a lane quadrant and take alternate 16-column chunks.
    const int ew = (warp - 4) & 3;
    const int chunk0 = (warp - 4) >> 2;           // 0, or 1 for the second warp of a quadrant
    constexpr int kChunkStep = 16 * (EW / 4);
    const uint32_t lane_base = static_cast<uint32_t>(ew * 32) << 16;
    bool dep_waited = false;
    for (int it = 0;; ++it) {
      int n_tiles;
      const int w = next_work(it, n_tiles);
      if (w < 0) break;
      WorkRef wr;
      int tok_lane = -1;     // direct mode: token (= MMA column) held by row (row_start + lane) of this item
      if (MODE == kModeGated && direct.tokens > 0) {
        if (w >= kSharedWorkBase) {
          wr = WorkRef{direct.shared_expert, direct.shared_row0 + direct.sh_lo, direct.sh_hi - direct.sh_lo,
                       w - kSharedWorkBase};
          if (lane < wr.nrows) tok_lane = direct.sh_lo + lane;
        } else {
          if (!dep_waited) {   // the work list and row_token are the direct predecessor's output
            pdl_wait();
            dep_waited = true;
          }
          wr = get_work<BN>(tiles, w, n_feat_tiles, n_tiles, 0);
          if (lane < wr.nrows) tok_lane = __ldcg(direct.row_token + wr.row_start + lane);
        }
      } else {
        wr = get_work<BN>(tiles, w, n_feat_tiles, n_tiles, dense_rows);
      }
      const int as = AS == 2 ? (it & 1) : 0;
      const uint32_t aphase = AS == 2 ? ((it >> 1) & 1u) : (it & 1u);
      mbar_wait(&tmem_full[as], aphase);
      tc_fence_after();
      const uint32_t t_acc = tmem_base + lane_base + as * Cfg::kAccCols;
      const int feat = wr.f * Cfg::kFeatPerTile + ew * 32 + lane;
      // output features past n_feat_valid (the ragged last feature tile of a dense MLP whose width is not a
      // multiple of 128, or the router's E) are computed on whatever rows the box covers and not stored
      const bool feat_ok = feat < n_feat_valid;
      T* out_col = out + static_cast<size_t>(wr.row_start) * ld_out + feat;
#pragma unroll 1
      const int n_cols = (MODE == kModeGated && direct.tokens > 0) ? direct.tokens : wr.nrows;
      for (int c0 = 16 * chunk0; c0 < BN; c0 += kChunkStep) {
        if (c0 >= n_cols) break;                // warp-uniform: remaining token columns unused
        uint32_t g[16];
        tmem_ld16(t_acc + c0, g);
        if (MODE == kModeGated) {
          uint32_t u[16];
          tmem_ld16(t_acc + BN + c0, u);
          tmem_ld_wait();
          if (direct.tokens > 0) {
            // column = token; its output row inside this expert is the lane that holds the token
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const uint32_t m = __ballot_sync(0xffffffffu, tok_lane == c0 + j);
              if (m && feat_ok) {
                const float v = silu_fast(__uint_as_float(g[j])) * __uint_as_float(u[j]);
                out_col[static_cast<size_t>(__ffs(m) - 1) * ld_out] = Dt<T>::from_f(v);
              }
            }
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              if (c0 + j < wr.nrows && feat_ok) {
                const float v = silu_fast(__uint_as_float(g[j])) * __uint_as_float(u[j]);
                out_col[static_cast<size_t>(c0 + j) * ld_out] = Dt<T>::from_f(v);
              }
            }
          }
        } else if (MODE == kModeDown || MODE == kModeDown2) {
          // lane j (< 16) fetches the routing weight of token column c0 + j once; broadcast by shuffle
          float rw_lane = 1.0f;
          {
            const int row = wr.row_start + c0 + (lane & 15);
            if (c0 + (lane & 15) < wr.nrows && row < n_routed_rows) rw_lane = __ldg(row_weight + row);
          }
          uint32_t g2[16];
          if (MODE == kModeDown2) tmem_ld16(t_acc + BN + c0, g2);   // features [128, 256) of the tile
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float rw = __shfl_sync(0xffffffffu, rw_lane, j);
            if (c0 + j < wr.nrows) {
              if (feat_ok) out_col[static_cast<size_t>(c0 + j) * ld_out] = Dt<T>::from_f(__uint_as_float(g[j]) * rw);
              if (MODE == kModeDown2 && feat + kBM < n_feat_valid)
                out_col[static_cast<size_t>(c0 + j) * ld_out + kBM] = Dt<T>::from_f(__uint_as_float(g2[j]) * rw);
            }
          }
        } else {
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            if (c0 + j < wr.nrows && feat_ok)
              out_col[static_cast<size_t>(c0 + j) * ld_out] = Dt<T>::from_f(__uint_as_float(g[j]));
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[as]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
}

// ------------------------------------------------------------------------------------------
// host side: tensor maps + launch
// ------------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess)
    return nullptr;
  fn = reinterpret_cast<PFN_encodeTiled>(p);
  return fn;
}

int umma_make_tensor_map(void* map_out, const void* base, int dtype, uint64_t rows, uint64_t cols,
                         uint32_t box_rows) {
  return umma_make_tensor_map_bk(map_out, base, dtype, rows, cols, box_rows, kBK);
}

int umma_make_tensor_map_bk(void* map_out, const void* base, int dtype, uint64_t rows, uint64_t cols,
                            uint32_t box_rows, int bk) {
  PFN_encodeTiled enc = get_encode_fn();
  if (!enc) return GLM4MOE_ERR_DEVICE;
  if (dtype != GLM4MOE_BF16 && dtype != GLM4MOE_F16) return GLM4MOE_ERR_UNSUPPORTED;
  if (cols % 8 != 0 || box_rows == 0 || box_rows > 256 || rows == 0 || (bk != 64 && bk != 32)) return GLM4MOE_ERR_ARG;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {cols * 2};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(bk), box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(static_cast<CUtensorMap*>(map_out),
                   dtype == GLM4MOE_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16,
                   2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   bk == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? GLM4MOE_OK : GLM4MOE_ERR_ARG;
}

template <typename T, int BN, int MODE, int BK = kBK>
static int launch_one(const UmmaLaunch& a, cudaStream_t st) {
  using Cfg = UmmaCfg<BN, MODE, BK>;
  auto kern = umma_grouped_kernel<T, BN, MODE, BK>;
  static unsigned long long attr_done = 0;   // per instantiation, one bit per device
  GLM4MOE_CUDA_CHECK(ensure_dynamic_smem(kern, Cfg::kSmemBytes, &attr_done));
  TailMaps tail{};
  if (BK == kBK && a.direct.tokens == 0) {
    int n = 0;
    for (int i = 0; i < 3; ++i)
      if (a.tm_x_tail[i] && a.tail_rows[i] > 0 && a.tail_rows[i] < BN && a.tail_rows[i] % 16 == 0 &&
          (n == 0 || a.tail_rows[i] > tail.rows[n - 1])) {
        tail.m[n] = *static_cast<const CUtensorMap*>(a.tm_x_tail[i]);
        tail.rows[n++] = a.tail_rows[i];
      }
  }
  GLM4MOE_CUDA_CHECK(launch_pdl(
      kern, dim3(a.grid), dim3(Cfg::kThreads), Cfg::kSmemBytes, st, *static_cast<const CUtensorMap*>(a.tm_w),
      *static_cast<const CUtensorMap*>(a.tm_x), tail, reinterpret_cast<const int4*>(a.tiles), a.num_tiles,
      a.dense_rows,
      static_cast<T*>(a.out), a.ld_out, a.row_weight, a.n_routed_rows, a.n_feat_tiles, a.n_feat_valid, a.n_kblocks,
      a.w_rows_per_expert, a.up_row_offset, a.sched, a.direct));
  return GLM4MOE_OK;
}

template <typename T, int MODE>
static int launch_bn(const UmmaLaunch& a, cudaStream_t st) {
  switch (a.tile_rows) {
    case 16: return launch_one<T, 16, MODE>(a, st);
    case 32: return launch_one<T, 32, MODE>(a, st);
    case 64: return launch_one<T, 64, MODE>(a, st);
    case 128: return launch_one<T, 128, MODE>(a, st);
    case 256: return launch_one<T, 256, MODE>(a, st);
    default: return GLM4MOE_ERR_ARG;
  }
}

template <typename T>
static int launch_mode(const UmmaLaunch& a, cudaStream_t st) {
  if (a.bk == 32) {   // half-size stages: the 256-token two-weight-tile configurations only
    if (a.tile_rows != 256) return GLM4MOE_ERR_ARG;
    if (a.mode == kModeDown2) return launch_one<T, 256, kModeDown2, 32>(a, st);
    if (a.mode == kModeGated) return launch_one<T, 256, kModeGated, 32>(a, st);
    return GLM4MOE_ERR_ARG;
  }
  if (a.mode == kModeDown2) return a.tile_rows == 256 ? launch_one<T, 256, kModeDown2>(a, st) : GLM4MOE_ERR_ARG;
  switch (a.mode) {
    case kModeDown: return launch_bn<T, kModeDown>(a, st);
    case kModeGated: return launch_bn<T, kModeGated>(a, st);
    case kModePlain: return launch_bn<T, kModePlain>(a, st);
    default: return GLM4MOE_ERR_ARG;
  }
}

int umma_launch(const UmmaLaunch& a, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (a.dtype == GLM4MOE_BF16) return launch_mode<__nv_bfloat16>(a, st);
  if (a.dtype == GLM4MOE_F16) return launch_mode<__half>(a, st);
  return GLM4MOE_ERR_UNSUPPORTED;
}

bool umma_shape_ok(int H, int I, int dtype) {
  return (dtype == GLM4MOE_BF16 || dtype == GLM4MOE_F16) && H % kBM == 0 && I % kBM == 0;
}

}  // namespace glm4moe

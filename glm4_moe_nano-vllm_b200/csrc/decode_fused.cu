// Decode-regime expert MLP (T <= 32): gate_up + SiLU*mul + down in ONE persistent kernel.
//
// Replaces, for decode-size batches, every Glm4MoeMLP.forward of the reference loop
// (moe.py:147-149 -> mlp.py:49-54) and of the shared expert (moe.py:165).  The regime is weight
// streaming: a hit expert's 34.6 MB (bf16) cross HBM once and the tensor pipe idles, so the only
// thing that matters is that all 148 SMs keep bulk copies in flight from the first microsecond
// to the last.  Two back-to-back GEMM launches lose ~15 us of a 70 us layer at T = 1 to the tail
// of the first and the ramp of the second (profiles/r1_decode.md); here both projections are work
// items of one device-wide queue:
//
//   S   gate_up of the shared expert, feature tile f  -> CTA f, BEFORE the programmatic dependency
//       wait: the token tile is the layer input itself ("direct" mode, see umma_gemm.cuh) and the
//       shared expert is always on, so these items overlap the router
//   G   gate_up of routed tile t, feature tile f      (tile-major; one tile per hit expert)
//   D   down of tile t (shared first), feature tile f (tile-major)
//
// A down item needs the activation rows of ALL gate_up feature tiles of its expert: every gate_up
// epilogue warp release-increments ready[tile]; the down item's producer streams its WEIGHT atoms
// first, then acquire-spins on the counter and only then loads the activation atoms.  Items are
// claimed in queue order by co-resident CTAs (grid <= SM count, 1 CTA/SM), so every gate_up item
// is already running when a down item waits on it: no deadlock.
//
// Geometry is expressed in bytes so the bf16 and (later) e4m3 tiers share it: an "atom" is a
// [rows x 128 B] K-major box with the 128-byte swizzle = 64 two-byte or 128 one-byte elements of
// K.  A pipeline stage holds two weight atoms + two token atoms (36 / 40 KB):
//   gate_up: gate atom, up atom (same K range), one token atom
//   down   : weight atoms of two consecutive K ranges and the two matching activation atoms
// Warp roles as umma_gemm.cu: warp 0 TMA producer + scheduler, warp 1 MMA issuer, warp 2 TMEM
// allocator, warps 4-7 epilogue (tcgen05.ld, fused math, stores with lane = output feature).
#include <cuda_fp8.h>

#include "route_device.cuh"

#include "umma_common.cuh"
#include "umma_gemm.cuh"

namespace glm4moe {

template <int BN, bool FP8>
struct FusedCfg {
  static constexpr int kAtomA = kBM * 128;          // 16 KB
  static constexpr int kAtomB = BN * 128;
  static constexpr int kStageBytes = 2 * kAtomA + 2 * kAtomB;
  static constexpr int kStages = (232448 - 2048) / kStageBytes;      // 6 (BN=16) / 5 (BN=32)
  static constexpr int kAccCols = 2 * BN;           // gate | up  (down uses the first half)
  // bf16: two accumulator slots (the epilogue of item i overlaps the MMAs of item i + 1).
  // e4m3: every stage's raw partial product lands in its own slot (ring of 8) and the epilogue
  // warps promote it with the fp32 block scales.
  static constexpr int kAccSlots = FP8 ? 8 : 2;
  static constexpr int kAtomK = FP8 ? 128 : 64;     // K elements per 128-byte atom
  static constexpr int kTmemColsRaw = kAccSlots * kAccCols;
  static constexpr int kTmemCols = kTmemColsRaw <= 32 ? 32 : kTmemColsRaw <= 64 ? 64 : kTmemColsRaw <= 128 ? 128
                                   : kTmemColsRaw <= 256 ? 256 : 512;
  // e4m3: 8 epilogue warps, two per TMEM lane quadrant, each owning half of the token columns (the
  // per-stage promotion is the busiest role there: ncu showed 4 warps ~90 % busy and HBM at 62 %)
  static constexpr int kEpiWarps = FP8 ? 8 : 4;
  static constexpr int kColsPerWarp = BN / (kEpiWarps / 4);
  static constexpr int kThreads = 128 + 32 * kEpiWarps;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 1024 /*barriers, queue, amax*/;
  static_assert(kStages >= 4, "pipeline too shallow");
};

struct FusedItem {
  int kind;        // 0 gate_up, 1 down
  int expert;      // index into the local weight stack
  int row_start;   // first row of the tile in the row space
  int nrows;
  int f;           // feature tile
  int flag;        // index into ready[]: 0 shared, 1 + routed tile
  int is_shared;
};

constexpr int kFusedQueue = 4;

// expert-parallel fast path: tell every rank (incl. this one) that this rank's rows of call e are complete
__device__ __forceinline__ void ep_publish_ready(const EpFastSignal& ep, int e) {
  __threadfence_system();
  for (int r = 0; r < ep.world; ++r)
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(ep.peer_ready[r]), "r"(e * 4 + kEpModeFast) : "memory");
}

template <typename T, int BN, bool FP8>
__global__ void __launch_bounds__(FusedCfg<BN, FP8>::kThreads, 1) decode_fused_kernel(
    const __grid_constant__ CUtensorMap tm_wgu,   // gate_up stack [n_stack * 2I, H], box {64, 128}
    const __grid_constant__ CUtensorMap tm_wd,    // down stack    [n_stack * H, I],  box {64, 128}
    const __grid_constant__ CUtensorMap tm_x,     // the layer input [T, H],          box {64, BN}
    const __grid_constant__ CUtensorMap tm_act,   // activation rows [R, I],          box {64, BN}
    const int4* tiles, const int32_t* num_tiles_ptr,       // (may be written by this kernel: RouteMerged)
    const int32_t* row_token, const float* row_weight, T* __restrict__ act, T* __restrict__ y, int H, int I,
    int tokens, int n_routed_rows, int shared_expert, int sh_lo, int sh_hi,
    int32_t* __restrict__ sched,                  // [0] queue head, [1] CTAs finished
    int32_t* __restrict__ ready,                  // [1 + tile capacity] gate_up completion counters
    const int32_t* slot_row, T* __restrict__ out, int k,
    int32_t* __restrict__ fdone,                  // [H/128] down items finished per feature tile
    const unsigned char* __restrict__ pf0, size_t pf0_bytes,    // shared expert's gate_up / down weights:
    const unsigned char* __restrict__ pf1, size_t pf1_bytes,    // pulled into L2 while the router runs
    const unsigned char* __restrict__ pf2, size_t pf2_bytes,    // speculative: the first local experts' weights
    const unsigned char* __restrict__ pf3, size_t pf3_bytes,    // (probably hit at T >= 24), same L2 prefetch
    const Fp8Args q,                              // e4m3 tier: scales and the quantised activation buffer
    const EpFastSignal ep,                        // expert-parallel fast path: parity row buffers + ready flags
    const RouteMerged rt) {                       // T <= 2: router front-end inside this kernel
  using Cfg = FusedCfg<BN, FP8>;
  constexpr int AK = Cfg::kAtomK;
  constexpr int NS = Cfg::kAccSlots;
  constexpr int EW = Cfg::kEpiWarps;
  constexpr int NC = Cfg::kColsPerWarp;     // token columns one epilogue warp owns
  constexpr int ET = 32 * EW;               // epilogue threads (named barrier 1)
  extern __shared__ unsigned char smem_dyn[];
  unsigned char* smem = reinterpret_cast<unsigned char*>(
      (reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + Cfg::kStages;
  uint64_t* tmem_full = bars + 2 * Cfg::kStages;
  uint64_t* tmem_empty = tmem_full + NS;
  uint64_t* q_full = tmem_empty + NS;
  uint64_t* q_empty = q_full + kFusedQueue;
  int32_t* q_w = reinterpret_cast<int32_t*>(q_empty + kFusedQueue);   // [depth][2]: work id, routed tiles
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(q_w + 2 * kFusedQueue);
  int32_t* s_last = reinterpret_cast<int32_t*>(tmem_ptr_smem + 1);   // epilogue-warp broadcast slot
  float* s_amax = reinterpret_cast<float*>(s_last + 2);               // [4 warps][32 columns] (e4m3 act quantiser)
  // e4m3: scales of the quantised input tokens, [H/128][BN], staged once per CTA (every gate_up item
  // of every expert multiplies by the same ones)
  float* s_sx = reinterpret_cast<float*>(smem + Cfg::kStages * Cfg::kStageBytes + 1024);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int FG = I / kBM, FD = H / kBM;           // feature tiles of the two projections
  const int n_sh = sh_hi - sh_lo;
  const int has_shared = n_sh > 0 ? 1 : 0;
  const int shared_row0 = n_routed_rows + sh_lo;   // T*k + sh_lo

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tm_wgu);
    prefetch_tmap(&tm_wd);
    prefetch_tmap(&tm_x);
    prefetch_tmap(&tm_act);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < NS; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], EW);
    }
    for (int s = 0; s < kFusedQueue; ++s) {
      mbar_init(&q_full[s], 1);
      mbar_init(&q_empty[s], 1 + EW);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<Cfg::kTmemCols>(tmem_ptr_smem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  // work id -> item.  ids: kSharedBase + f = phase S; [0, G) phase G; [G, G + D) phase D.
  constexpr int kSharedBase = 1 << 30;
  constexpr int kZeroBase = 1 << 29;     // no local work at all: feature tile (w - kZeroBase) of `out` is zeroed
  auto decode = [&](int w, int n_routed) -> FusedItem {
    FusedItem it;
    if (w >= kSharedBase) {
      it = FusedItem{0, shared_expert, shared_row0, n_sh, w - kSharedBase, 0, 1};
      return it;
    }
    const int G = n_routed * FG;
    if (w < G) {
      const int t = w / FG;
      const int4 tl = __ldcg(tiles + t);
      it = FusedItem{0, tl.x, tl.y, tl.z, w - t * FG, 1 + t, 0};
      return it;
    }
    const int d = w - G;
    const int td = d / FD;
    const int f = d - td * FD;
    if (has_shared && td == 0) {
      it = FusedItem{1, shared_expert, shared_row0, n_sh, f, 0, 1};
    } else {
      const int t = td - has_shared;
      const int4 tl = __ldcg(tiles + t);
      it = FusedItem{1, tl.x, tl.y, tl.z, f, 1 + t, 0};
    }
    return it;
  };
  auto next_work = [&](int n, int& n_routed) -> int {
    const int q = n % kFusedQueue;
    mbar_wait(&q_full[q], (n / kFusedQueue) & 1u);
    const int w = q_w[2 * q];
    n_routed = q_w[2 * q + 1];
    __syncwarp();
    if (lane == 0) mbar_arrive(&q_empty[q]);
    return w;
  };
  const int atoms_gu = H / AK;    // K atoms (128 B = 64 two-byte or 128 one-byte elements) of gate_up
  const int atoms_d = I / AK;     // K atoms of down

  if (warp == 0) {
    // ===================== TMA producer + scheduler =====================
    if (elect_one()) {
      int n_pushed = 0, stage = 0;
      uint32_t phase = 0;
      auto push = [&](int w, int nr) {
        const int q = n_pushed % kFusedQueue;
        mbar_wait(&q_empty[q], ((n_pushed / kFusedQueue) & 1u) ^ 1u);
        q_w[2 * q] = w;
        q_w[2 * q + 1] = nr;
        mbar_arrive(&q_full[q]);
        ++n_pushed;
      };
      auto advance = [&]() {
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1u; }
      };
      // gate_up item: stage s <- {gate atom, up atom, token atom} of K atom kb
      auto stream_gate_up = [&](const FusedItem& it) {
        const int w_row = it.expert * 2 * I + it.f * kBM;
        for (int kb = 0; kb < atoms_gu; ++kb) {
          unsigned char* sa = smem + stage * Cfg::kStageBytes;
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          mbar_expect_tx(&full_bar[stage], 2 * Cfg::kAtomA + Cfg::kAtomB);
          tma_load_2d(&tm_wgu, &full_bar[stage], sa, kb * AK, w_row, kEvictFirst);
          tma_load_2d(&tm_wgu, &full_bar[stage], sa + Cfg::kAtomA, kb * AK, w_row + I, kEvictFirst);
          tma_load_2d(&tm_x, &full_bar[stage], sa + 2 * Cfg::kAtomA, kb * AK, 0, kEvictLast);
          advance();
        }
      };
      // down item: stage s <- weight atoms {2j, 2j+1} then, once the expert's activation rows are
      // complete, the matching activation atoms
      auto stream_down = [&](const FusedItem& it) {
        const int w_row = it.expert * H + it.f * kBM;
        const int n_st = (atoms_d + 1) / 2;
        auto load_w = [&](int j, int s) {
          unsigned char* sa = smem + s * Cfg::kStageBytes;
          const int na = min(2, atoms_d - 2 * j);
          mbar_expect_tx(&full_bar[s], na * (Cfg::kAtomA + Cfg::kAtomB));
          for (int a = 0; a < na; ++a)
            tma_load_2d(&tm_wd, &full_bar[s], sa + a * Cfg::kAtomA, (2 * j + a) * AK, w_row, kEvictFirst);
        };
        auto load_b = [&](int j, int s) {
          unsigned char* sa = smem + s * Cfg::kStageBytes + 2 * Cfg::kAtomA;
          const int na = min(2, atoms_d - 2 * j);
          for (int a = 0; a < na; ++a)
            tma_load_2d(&tm_act, &full_bar[s], sa + a * Cfg::kAtomB, (2 * j + a) * AK, it.row_start, kEvictLast);
        };
        const int target = EW * FG;
        const int32_t* flag = ready + it.flag;
        int v;
        asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
        if (v >= target) {
          // common case in the down phase: the expert's activation rows are already complete, so every
          // stage is issued whole (weights + activations) as soon as its slot frees up
          asm volatile("fence.proxy.async.global;" ::: "memory");
          for (int j = 0; j < n_st; ++j) {
            mbar_wait(&empty_bar[stage], phase ^ 1u);
            load_w(j, stage);
            load_b(j, stage);
            advance();
          }
          return;
        }
        // not yet: stream the WEIGHT atoms of the first ring pass, then wait, then the activation atoms
        const int n_pre = min(n_st, Cfg::kStages);
        int s = stage;
        uint32_t ph = phase;
        for (int j = 0; j < n_pre; ++j) {
          mbar_wait(&empty_bar[s], ph ^ 1u);
          load_w(j, s);
          if (++s == Cfg::kStages) { s = 0; ph ^= 1u; }
        }
        // acquire: all gate_up feature tiles of this expert have stored their activation columns
        do {
          asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
        } while (v < target);
        asm volatile("fence.proxy.async.global;" ::: "memory");   // generic-proxy stores -> async-proxy (TMA) reads
        for (int j = 0; j < n_pre; ++j) {
          load_b(j, stage);
          advance();
        }
        for (int j = n_pre; j < n_st; ++j) {
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          load_w(j, stage);
          load_b(j, stage);
          advance();
        }
      };
      // ---- before the dependency wait: the shared expert is always on.  Every CTA pulls its slice of
      // the shared expert's 3*H*I weights into L2 (HBM would otherwise idle while the router runs;
      // they fit the 126 MB L2 many times over), then phase S streams gate_up from there. ----
      {
        constexpr size_t kChunk = 32768;
        auto prefetch_slice = [&](const unsigned char* base, size_t bytes) {
          const size_t n_chunks = (bytes + kChunk - 1) / kChunk;
          for (size_t c = blockIdx.x; c < n_chunks; c += gridDim.x) {
            const size_t off = c * kChunk;
            const uint32_t sz = static_cast<uint32_t>(min(kChunk, bytes - off));
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(base + off), "r"(sz) : "memory");
          }
        };
        if (has_shared) {
          prefetch_slice(pf0, pf0_bytes);
          prefetch_slice(pf1, pf1_bytes);
        }
        // Speculation: with >= ~24 tokens each expert is hit with probability >= 0.8 and the work list visits
        // the local experts in ascending order, so the first ones are the first routed items after the wait.
        // Their weights are pulled towards L2 in the router's shadow too (HBM is idle there); a miss costs
        // bandwidth nobody was using.
        if (pf2_bytes) prefetch_slice(pf2, pf2_bytes);
        if (pf3_bytes) prefetch_slice(pf3, pf3_bytes);
      }
      // merged front-end: this kernel is the first of the layer, so x itself is the predecessor's output
      if (rt.enabled) {
        pdl_wait();
        pdl_launch();
      }
      // (e4m3: the token tile is the router kernel's quantised copy of x, so phase S follows the wait)
      if (has_shared && !FP8)
        for (int f = blockIdx.x; f < FG; f += gridDim.x) {
          push(kSharedBase + f, 0);
          stream_gate_up(FusedItem{0, shared_expert, shared_row0, n_sh, f, 0, 1});
        }
      if (rt.enabled) {
        // the routing of this call: released by the router CTA's epilogue warps (below)
        int v;
        const long long t0 = clock64();
        do {
          asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(rt.sync + 1) : "memory");
          if (v == 0 && clock64() - t0 > (1ll << 32)) __trap();     // never in a correct run
        } while (v == 0);
      } else {
        pdl_wait();
        pdl_launch();
      }
      if (has_shared && FP8)
        for (int f = blockIdx.x; f < FG; f += gridDim.x) {
          push(kSharedBase + f, 0);
          stream_gate_up(FusedItem{0, shared_expert, shared_row0, n_sh, f, 0, 1});
        }
      const int n_routed = *reinterpret_cast<const volatile int32_t*>(num_tiles_ptr) - has_shared;
      const int total = n_routed * FG + (n_routed + has_shared) * FD;
      if (n_routed + has_shared == 0) {   // expert-parallel rank without a hit expert: the partial output is zero
        for (int f = blockIdx.x; f < FD; f += gridDim.x) push(kZeroBase + f, 0);
        // ... and it has no row to write: "complete" right away (the epoch is final after the wait above)
        if (ep.enabled && blockIdx.x == 0)
          ep_publish_ready(ep, *reinterpret_cast<const volatile int32_t*>(ep.ctl) + 1);
      }
      for (;;) {
        const int w = atomicAdd(&sched[0], 1);
        if (w >= total) {
          push(-1, 0);
          break;
        }
        push(w, n_routed);
        const FusedItem it = decode(w, n_routed);
        if (it.kind == 0) stream_gate_up(it); else stream_down(it);
      }
      // last CTA out re-arms the queue head and the completion counters for the next launch; every
      // producer is past its last acquire by then, so every counter has reached its final value
      __threadfence();
      if (atomicAdd(&sched[1], 1) == static_cast<int>(gridDim.x) - 1) {
        sched[0] = 0;
        sched[1] = 0;
        for (int i = 0; i <= n_routed; ++i) ready[i] = 0;
        if (rt.enabled) rt.sync[1] = 0;      // every producer is past its acquire of the routing flag
        __threadfence();
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc_base = make_idesc(sizeof(T) == 2 && !std::is_same<T, __half>::value, 0);
    int stage = 0;
    uint32_t phase = 0;
    uint32_t pc = 0;        // e4m3: stages issued so far (TMEM slot ring position)
    bool dep_waited = false;
    for (int qn = 0, n = 0;; ++qn) {
      int n_routed;
      const int w = next_work(qn, n_routed);
      if (w < 0) break;
      if (w >= kZeroBase && w < kSharedBase) continue;
      if (w < kSharedBase && !dep_waited) {   // decode() reads the work list: the direct predecessor's output
        pdl_wait();
        dep_waited = true;
      }
      const FusedItem it = decode(w, n_routed);
      const int n_eff = min(BN, ((it.kind == 0 ? tokens : it.nrows) + 15) & ~15);
      const int n_st = it.kind == 0 ? atoms_gu : (atoms_d + 1) / 2;
      if constexpr (FP8) {
        // one fresh TMEM slot per stage: [gate | up] partials of a K atom, or the partials of the
        // two K atoms of a down stage; the epilogue warps apply the block scales
        const uint32_t idesc = make_idesc_f8(0) | (static_cast<uint32_t>(n_eff >> 3) << 17);
        for (int j = 0; j < n_st; ++j, ++pc) {
          const int ps = pc % NS;
          mbar_wait(&tmem_empty[ps], ((pc / NS) & 1u) ^ 1u);
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t sa = smem_u32(smem + stage * Cfg::kStageBytes);
            const uint64_t a0 = make_kmajor_desc(sa);
            const uint64_t a1 = make_kmajor_desc(sa + Cfg::kAtomA);
            const uint64_t b0 = make_kmajor_desc(sa + 2 * Cfg::kAtomA);
            const uint64_t b1 = make_kmajor_desc(sa + 2 * Cfg::kAtomA + Cfg::kAtomB);
            const uint32_t d_tmem = tmem_base + ps * Cfg::kAccCols;
            const bool second = it.kind == 0 || atoms_d - 2 * j >= 2;
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
              umma_f8(d_tmem, a0 + 2 * ks, b0 + 2 * ks, idesc, ks > 0 ? 1u : 0u);
              if (second) umma_f8(d_tmem + BN, a1 + 2 * ks, (it.kind == 0 ? b0 : b1) + 2 * ks, idesc, ks > 0 ? 1u : 0u);
            }
            umma_commit(&empty_bar[stage]);
            umma_commit(&tmem_full[ps]);
          }
          __syncwarp();
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1u; }
        }
      } else {
      const uint32_t idesc = idesc_base | (static_cast<uint32_t>(n_eff >> 3) << 17);
      const int as = n & 1;
      mbar_wait(&tmem_empty[as], ((n >> 1) & 1u) ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * Cfg::kAccCols;
      for (int j = 0; j < n_st; ++j) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(smem + stage * Cfg::kStageBytes);
          const uint64_t a0 = make_kmajor_desc(sa);
          const uint64_t a1 = make_kmajor_desc(sa + Cfg::kAtomA);
          const uint64_t b0 = make_kmajor_desc(sa + 2 * Cfg::kAtomA);
          const uint64_t b1 = make_kmajor_desc(sa + 2 * Cfg::kAtomA + Cfg::kAtomB);
          if (it.kind == 0) {
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
              const uint32_t acc = (j > 0 || ks > 0) ? 1u : 0u;
              umma_f16(d_tmem, a0 + 2 * ks, b0 + 2 * ks, idesc, acc);
              umma_f16(d_tmem + BN, a1 + 2 * ks, b0 + 2 * ks, idesc, acc);
            }
          } else {
            const int na = min(2, atoms_d - 2 * j);
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) umma_f16(d_tmem, a0 + 2 * ks, b0 + 2 * ks, idesc, (j > 0 || ks > 0) ? 1u : 0u);
            if (na == 2) {
#pragma unroll
              for (int ks = 0; ks < 4; ++ks) umma_f16(d_tmem, a1 + 2 * ks, b1 + 2 * ks, idesc, 1u);
            }
          }
          umma_commit(&empty_bar[stage]);
          if (j == n_st - 1) umma_commit(&tmem_full[as]);
        }
        __syncwarp();
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1u; }
      }
      }
      ++n;
    }
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    const int ewi = warp - 4;                 // epilogue warp index
    const int ew = ewi & 3;                   // TMEM lane quadrant
    const int col0 = (ewi >> 2) * NC;         // first token column this warp owns
    const uint32_t lane_base = static_cast<uint32_t>(ew * 32) << 16;
    auto epi_sync = [&]() { asm volatile("bar.sync 1, %0;" ::"n"(ET) : "memory"); };
    uint32_t pc = 0;        // e4m3: stages consumed so far (TMEM slot ring position)
    bool dep_waited = false;
    // The LAST down item of feature tile f sums, for every token, its k routed rows in ascending
    // expert order + the shared row over these 128 features: same order and single rounding as
    // combine_kernel (bit-identical, no atomics on data).
    auto combine_f = [&](int f) {
      __threadfence();
      const size_t col = static_cast<size_t>(f) * kBM + 4 * lane;
      for (int t = ewi; t < tokens; t += EW) {
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        auto add_row = [&](int r) {
          const uint2 raw = __ldcg(reinterpret_cast<const uint2*>(y + static_cast<size_t>(r) * H + col));
          const T* h4 = reinterpret_cast<const T*>(&raw);
#pragma unroll
          for (int i = 0; i < 4; ++i) acc[i] += Dt<T>::to_f(h4[i]);
        };
        for (int sl = 0; sl < k; ++sl) {
          const int r = __ldcg(slot_row + static_cast<size_t>(t) * k + sl);
          if (r >= 0) add_row(r);
        }
        if (t >= sh_lo && t < sh_hi) add_row(n_routed_rows + t);
        __align__(8) T o[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) o[i] = Dt<T>::from_f(acc[i]);
        *reinterpret_cast<uint2*>(out + static_cast<size_t>(t) * H + col) = *reinterpret_cast<uint2*>(o);
      }
    };
    // ---- merged router front-end (T <= 2): logits of expert blockIdx.x (+ gridDim.x, ...) by this CTA's 4
    // epilogue warps in route_small_kernel's summation order (its 256 threads = 2 "virtual" threads per
    // thread here: vectors v, v + 256, ... accumulated in order, warp sums over the same lanes, the 8 warp
    // partials added in order), then ticket -> selection -> permutation -> flag. ----
    if constexpr (!FP8) {
      if (rt.enabled) {
        pdl_wait();
        constexpr int V = Dt<T>::kVec;
        const int E = rt.E, nvec = H / V;
        const T* xin = static_cast<const T*>(rt.x);
        const T* gw = static_cast<const T*>(rt.gate_w);
        T* lg = static_cast<T*>(rt.logits);
        const int tid = ewi * 32 + lane;                       // 0..127 (4 epilogue warps in this tier)
        const int n_rt = min(static_cast<int>(gridDim.x), E);  // CTAs that own at least one expert
        for (int e = blockIdx.x; e < E; e += gridDim.x) {
          float acc[2][2] = {{0.f, 0.f}, {0.f, 0.f}};         // [virtual thread][token]
#pragma unroll
          for (int vt = 0; vt < 2; ++vt) {
            for (int c = tid + 128 * vt; c < nvec; c += 256) {
              float wv[V], xv[2][V];
              unpack16<T>(ld_cached16(reinterpret_cast<const uint4*>(gw + static_cast<size_t>(e) * H) + c), wv);
#pragma unroll
              for (int i = 0; i < 2; ++i)
                unpack16<T>(ld_cached16(reinterpret_cast<const uint4*>(xin + static_cast<size_t>(min(i, tokens - 1)) * H) + c), xv[i]);
#pragma unroll
              for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int qv = 0; qv < V; ++qv) acc[vt][i] = fmaf(xv[i][qv], wv[qv], acc[vt][i]);
            }
          }
#pragma unroll
          for (int vt = 0; vt < 2; ++vt)
#pragma unroll
            for (int i = 0; i < 2; ++i) {
              const float s = warp_sum(acc[vt][i]);
              if (lane == 0) s_amax[(ewi + 4 * vt) * 2 + i] = s;       // virtual warp ewi + 4 vt
            }
          epi_sync();
          if (tid < tokens) {
            float s = 0.f;
#pragma unroll
            for (int wdx = 0; wdx < 8; ++wdx) s += s_amax[wdx * 2 + tid];
            lg[static_cast<size_t>(tid) * E + e] = Dt<T>::from_f(s);
            __threadfence();
          }
          epi_sync();
        }
        if (tid == 0) {
          int last = 0;
          if (static_cast<int>(blockIdx.x) < n_rt) {
            __threadfence();
            last = atomicAdd(rt.sync, 1) == n_rt - 1;
            if (last) rt.sync[0] = 0;                         // re-armed for the next launch
          }
          s_last[0] = last;
        }
        epi_sync();
        if (s_last[0]) {
          __threadfence();
          if (ewi < tokens)
            select_token<T>(lg + static_cast<size_t>(ewi) * E, rt.bias, E, rt.k, rt.scale, lane,
                            rt.idx + static_cast<size_t>(ewi) * rt.k, rt.w + static_cast<size_t>(ewi) * rt.k);
          __threadfence();
          epi_sync();
          if (ewi == 0) {
            permute_tiny_warp(rt.idx, rt.w, tokens, rt.k, E, rt.e_lo, rt.e_hi, BN, sh_lo, sh_hi, rt.expert_offsets,
                              rt.row_token, rt.row_weight, rt.slot_row, reinterpret_cast<int4*>(rt.tiles),
                              rt.tile_capacity, rt.num_tiles, lane);
            __threadfence();
            __syncwarp();
            if (lane == 0) asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(rt.sync + 1), "r"(1) : "memory");
          }
        }
        epi_sync();     // s_last / s_amax are reused by the main loop
      }
    }
    int pend_f = -1, pend_done = 0, pend_target = 0;   // down item whose feature-tile count is in flight
    int pend_slot = 0;                                  // broadcast slot (alternates, so one barrier suffices)
    T* y_rows = y;      // expert-parallel fast path: the buffer of this call's parity (resolved at the first down item)
    int ep_e = 0;
    // publish(): thread 0 turns the count that came back for the PREVIOUS down item into a flag in
    // s_last[slot]; after the caller's next epi_sync every epilogue thread reads it (resolve()).
    auto publish = [&]() {
      if (pend_f >= 0 && ewi == 0 && lane == 0) {
        const int last = pend_done == pend_target;
        if (last) {
          fdone[pend_f] = 0;                            // re-armed for the next launch
          if (ep.enabled) {
            // feature tile complete on this rank; the thread that completes the last one publishes the
            // ready flags (its acq_rel chain has seen every other CTA's row stores)
            int c;
            asm volatile("atom.add.acq_rel.gpu.global.s32 %0, [%1], 1;" : "=r"(c) : "l"(fdone + FD + 1) : "memory");
            if (c == FD - 1) {
              fdone[FD + 1] = 0;
              ep_publish_ready(ep, ep_e);
            }
          }
        }
        s_last[pend_slot] = last;
      }
    };
    auto resolve = [&]() {
      if (pend_f < 0) return;                           // uniform over the epilogue warps
      const int last = reinterpret_cast<volatile int32_t*>(s_last)[pend_slot];
      if (last && out != nullptr) combine_f(pend_f);   // out == nullptr: rows only (peer-memory combine follows)
      pend_f = -1;
      pend_slot ^= 1;
    };
    for (int qn = 0, n = 0;; ++qn) {
      int n_routed;
      const int w = next_work(qn, n_routed);
      if (w < 0) break;
      if (w >= kZeroBase && w < kSharedBase) {
        const int col = (w - kZeroBase) * kBM + 4 * lane;
        if (out != nullptr)
          for (int t = ewi; t < tokens; t += EW) *reinterpret_cast<uint2*>(out + static_cast<size_t>(t) * H + col) = make_uint2(0u, 0u);
        continue;
      }
      if ((FP8 || w < kSharedBase) && !dep_waited) {   // e4m3: the token scales are the router kernel's output
        pdl_wait();
        dep_waited = true;
        if constexpr (FP8) {
          for (int i = ewi * 32 + lane; i < atoms_gu * BN; i += ET) {
            const int a = i / BN, c = i - a * BN;
            s_sx[i] = c < tokens ? __ldcg(q.sx_t + static_cast<size_t>(a) * q.scale_ld + c) : 0.f;
          }
          epi_sync();
        }
      }
      const FusedItem it = decode(w, n_routed);
      const int as = n & 1;
      const uint32_t t_acc = tmem_base + lane_base + as * Cfg::kAccCols;
      const int feat = it.f * kBM + ew * 32 + lane;
      if (it.kind == 0) {
        // column c of the accumulator is token c; its row inside this expert is the lane holding it
        int tok_lane = -1;
        if (lane < it.nrows) tok_lane = it.is_shared ? sh_lo + lane : __ldcg(row_token + it.row_start + lane);
        T* out_col = act + static_cast<size_t>(it.row_start) * I + feat;
        if constexpr (FP8) {
          // promote every K atom's raw partial with s_w[feature block][atom] * s_x[token][atom];
          // this warp owns token columns [col0, col0 + NC)
          // ... and skips the 8-column chunks none of this expert's tokens fall into (at T = 32 an expert
          // holds ~2 of the 32 columns)
          constexpr int LD = 8;
          const uint32_t need = __reduce_or_sync(0xffffffffu, tok_lane >= 0 ? 1u << tok_lane : 0u);
          float acc_g[NC], acc_u[NC];
#pragma unroll
          for (int c = 0; c < NC; ++c) acc_g[c] = acc_u[c] = 0.f;
          const float* swg = q.s_gate_up + (static_cast<size_t>(it.expert) * (2 * I / 128) + it.f) * atoms_gu;
          const float* swu = swg + static_cast<size_t>(I / 128) * atoms_gu;
          // weight scales of the whole item in two registers per lane (atoms <= 64), fetched before the
          // first partial lands; per stage they come back by shuffle (no global latency per stage)
          const float wg_lo = lane < atoms_gu ? __ldg(swg + lane) : 0.f;
          const float wg_hi = 32 + lane < atoms_gu ? __ldg(swg + 32 + lane) : 0.f;
          const float wu_lo = lane < atoms_gu ? __ldg(swu + lane) : 0.f;
          const float wu_hi = 32 + lane < atoms_gu ? __ldg(swu + 32 + lane) : 0.f;
          for (int j = 0; j < atoms_gu; ++j, ++pc) {
            const int ps = pc % NS;
            const float sg = __shfl_sync(0xffffffffu, j < 32 ? wg_lo : wg_hi, j & 31);
            const float su = __shfl_sync(0xffffffffu, j < 32 ? wu_lo : wu_hi, j & 31);
            const float sx_l = s_sx[j * BN + (lane & (BN - 1))];      // lane c holds column c's scale
            mbar_wait(&tmem_full[ps], (pc / NS) & 1u);
            tc_fence_after();
            const uint32_t t_p = tmem_base + lane_base + ps * Cfg::kAccCols + col0;
#pragma unroll
            for (int cc = 0; cc < NC; cc += LD) {
              if ((need >> (col0 + cc)) & 0xffu) {
                uint32_t g[LD], u[LD];
                tmem_ldn<LD>(t_p + cc, g);
                tmem_ldn<LD>(t_p + BN + cc, u);
                tmem_ld_wait();
#pragma unroll
                for (int jj = 0; jj < LD; ++jj) {
                  const float sx = __shfl_sync(0xffffffffu, sx_l, col0 + cc + jj);
                  acc_g[cc + jj] = fmaf(__uint_as_float(g[jj]), sx * sg, acc_g[cc + jj]);
                  acc_u[cc + jj] = fmaf(__uint_as_float(u[jj]), sx * su, acc_u[cc + jj]);
                }
              }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty[ps]);
          }
          // activation value in the model dtype, then the 1x128 quantiser over this tile's 128 features
          // (scale = amax / 448, q = rn_e4m3(v / scale): the formula of quant_rows_fp8_kernel)
#pragma unroll
          for (int c = 0; c < NC; ++c) {
            if ((need >> (col0 + c)) & 1u) {
              acc_g[c] = Dt<T>::round(silu_fast(acc_g[c]) * acc_u[c]);
              float m = fabsf(acc_g[c]);
#pragma unroll
              for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
              if (lane == 0) s_amax[ew * 32 + col0 + c] = m;
            }
          }
          epi_sync();
#pragma unroll
          for (int c = 0; c < NC; ++c) {
            const int col = col0 + c;
            if ((need >> col) & 1u) {
              const uint32_t m = __ballot_sync(0xffffffffu, tok_lane == col);
              if (m) {
                const int row = it.row_start + __ffs(m) - 1;
                const float amax = fmaxf(fmaxf(s_amax[col], s_amax[32 + col]), fmaxf(s_amax[64 + col], s_amax[96 + col]));
                const float scale = act_scale(amax, q.pow2);
                act[static_cast<size_t>(row) * I + feat] = Dt<T>::from_f(acc_g[c]);
                q.actq[static_cast<size_t>(row) * I + feat] = static_cast<unsigned char>(
                    __nv_cvt_float_to_fp8(__fdiv_rn(acc_g[c], scale), __NV_SATFINITE, __NV_E4M3));
                if (ew == 0 && lane == 0) q.sact_t[static_cast<size_t>(it.f) * q.scale_ld + row] = scale;
              }
            }
          }
          epi_sync();   // s_amax is reused by the next item
          if (lane == 0) {
            // release this warp's activation columns to the down items of the expert (TMA = async proxy)
            asm volatile("fence.proxy.async.global;" ::: "memory");
            asm volatile("red.release.gpu.global.add.s32 [%0], 1;" ::"l"(ready + it.flag) : "memory");
          }
        } else {
        mbar_wait(&tmem_full[as], (n >> 1) & 1u);
        tc_fence_after();
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 16) {
          if (c0 >= tokens) break;
          uint32_t g[16], u[16];
          tmem_ld16(t_acc + c0, g);
          tmem_ld16(t_acc + BN + c0, u);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const uint32_t m = __ballot_sync(0xffffffffu, tok_lane == c0 + j);
            if (m) {
              const float v = silu_fast(__uint_as_float(g[j])) * __uint_as_float(u[j]);
              out_col[static_cast<size_t>(__ffs(m) - 1) * I] = Dt<T>::from_f(v);
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&tmem_empty[as]);
          // release the activation columns to the down items of this expert (TMA = async proxy); the
          // other lanes' stores are ordered before by the __syncwarp above (release is cumulative)
          asm volatile("fence.proxy.async.global;" ::: "memory");
          asm volatile("red.release.gpu.global.add.s32 [%0], 1;" ::"l"(ready + it.flag) : "memory");
        }
        }
      } else {
        if (ep.enabled && ep_e == 0) {   // after the dependency wait above: the previous combine has retired
          ep_e = *reinterpret_cast<const volatile int32_t*>(ep.ctl) + 1;
          y_rows = static_cast<T*>(ep.y_alt[ep_e & 1]);
        }
        T* out_col = y_rows + static_cast<size_t>(it.row_start) * H + feat;
        if constexpr (FP8) {
          // The activation scales were written by other CTAs' gate_up epilogues.  This CTA's producer
          // acquired ready[flag] before it loaded the activation atoms whose MMAs these warps wait on,
          // and the scales are read from L2 (ld.cg) after that wait.  No spin here: the counter may
          // already have been re-armed by the last producer to finish.
          int flag_seen = 0;
          if (lane == 0)
            asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(flag_seen) : "l"(ready + it.flag) : "memory");
          flag_seen = __shfl_sync(0xffffffffu, flag_seen, 0);
          constexpr int LD = NC >= 16 ? 16 : 8;
          float acc[NC];
#pragma unroll
          for (int c = 0; c < NC; ++c) acc[c] = 0.f;
          const float* sw = q.s_down + (static_cast<size_t>(it.expert) * (H / 128) + it.f) * atoms_d;
          const bool col_ok = lane < it.nrows;
          const float* sa = q.sact_t + it.row_start + lane;
          const int n_st = (atoms_d + 1) / 2;
          auto load_sa = [&](int atom) -> float {
            return col_ok && atom < atoms_d ? __ldcg(sa + static_cast<size_t>(atom) * q.scale_ld) : 0.f;
          };
          float a0 = 0.f, a1 = 0.f, nx0 = 0.f, nx1 = 0.f;
          const float wd_lo = lane < atoms_d ? __ldg(sw + lane) : 0.f;
          const float wd_hi = 32 + lane < atoms_d ? __ldg(sw + 32 + lane) : 0.f;
          float rw_lane = 1.0f;
          if (col_ok && it.row_start + lane < n_routed_rows) rw_lane = __ldcg(row_weight + it.row_start + lane);
          // In the down phase the expert's rows are normally complete long before this item starts: peek
          // at the counter and, if so, fetch the first pair of activation scales before the first wait.
          const bool early = flag_seen >= EW * FG;
          if (early) { a0 = load_sa(0); a1 = load_sa(1); }
          for (int j = 0; j < n_st; ++j, ++pc) {
            const int ps = pc % NS;
            const bool two = atoms_d - 2 * j >= 2;
            const float w0 = __shfl_sync(0xffffffffu, 2 * j < 32 ? wd_lo : wd_hi, (2 * j) & 31);
            const float w1 = __shfl_sync(0xffffffffu, 2 * j + 1 < 32 ? wd_lo : wd_hi, (2 * j + 1) & 31);
            mbar_wait(&tmem_full[ps], (pc / NS) & 1u);
            tc_fence_after();
            // activation scales only after the first partial has landed (= the rows are complete);
            // the next stage's pair is fetched one stage ahead
            if (j == 0) { if (!early) { a0 = load_sa(0); a1 = load_sa(1); } } else { a0 = nx0; a1 = nx1; }
            nx0 = load_sa(2 * j + 2);
            nx1 = load_sa(2 * j + 3);
            const uint32_t t_p = tmem_base + lane_base + ps * Cfg::kAccCols + col0;
#pragma unroll
            for (int cc = 0; cc < NC; cc += LD) {
              if (col0 + cc < it.nrows) {
                uint32_t p0[LD], p1[LD];
                tmem_ldn<LD>(t_p + cc, p0);
                if (two) tmem_ldn<LD>(t_p + BN + cc, p1);
                tmem_ld_wait();
#pragma unroll
                for (int jj = 0; jj < LD; ++jj) {
                  const float s0 = __shfl_sync(0xffffffffu, a0, col0 + cc + jj);
                  const float s1 = __shfl_sync(0xffffffffu, a1, col0 + cc + jj);
                  acc[cc + jj] = fmaf(__uint_as_float(p0[jj]), s0 * w0, acc[cc + jj]);
                  if (two) acc[cc + jj] = fmaf(__uint_as_float(p1[jj]), s1 * w1, acc[cc + jj]);
                }
              }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty[ps]);
          }
#pragma unroll
          for (int c = 0; c < NC; ++c) {
            const float rw = __shfl_sync(0xffffffffu, rw_lane, col0 + c);
            if (col0 + c < it.nrows) out_col[static_cast<size_t>(col0 + c) * H] = Dt<T>::from_f(acc[c] * rw);
          }
        } else {
        mbar_wait(&tmem_full[as], (n >> 1) & 1u);
        tc_fence_after();
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 16) {
          if (c0 >= it.nrows) break;
          uint32_t g[16];
          tmem_ld16(t_acc + c0, g);
          float rw_lane = 1.0f;
          {
            const int row = it.row_start + c0 + (lane & 15);
            if (c0 + (lane & 15) < it.nrows && row < n_routed_rows) rw_lane = __ldcg(row_weight + row);
          }
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float rw = __shfl_sync(0xffffffffu, rw_lane, j);
            if (c0 + j < it.nrows)
              out_col[static_cast<size_t>(c0 + j) * H] = Dt<T>::from_f(__uint_as_float(g[j]) * rw);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty[as]);
        }
        // ---- combine (moe.py:151,171), fused: count this item on its feature tile.  The counter's
        // round trip to L2 is not waited for here: the result is looked at after the NEXT down item
        // (finish_pending), so the epilogue never stalls on it.
        publish();
        epi_sync();                // all epilogue threads' y stores happen-before the release below
        resolve();
        if (ewi == 0 && lane == 0) {
          asm volatile("atom.add.acq_rel.gpu.global.s32 %0, [%1], 1;" : "=r"(pend_done) : "l"(fdone + it.f) : "memory");
          pend_done += 1;
        }
        pend_f = it.f;
        pend_target = n_routed + has_shared;
      }
      ++n;
    }
    publish();
    epi_sync();
    resolve();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
}

template <typename T, int BN, bool FP8>
static int launch_fused_one(const DecodeFusedLaunch& a, cudaStream_t st) {
  using Cfg = FusedCfg<BN, FP8>;
  auto kern = decode_fused_kernel<T, BN, FP8>;
  static unsigned long long attr_done = 0;   // per instantiation, one bit per device
  GLM4MOE_CUDA_CHECK(ensure_dynamic_smem(kern, 232448, &attr_done));
  if (FP8 && (a.H / 128 > 64 || a.I / 128 > 64 || Cfg::kSmemBytes + a.H / 128 * BN * 4 > 232448))
    return GLM4MOE_ERR_UNSUPPORTED;
  const size_t smem_bytes = Cfg::kSmemBytes + (FP8 ? a.H / 128 * BN * 4 : 0);
  GLM4MOE_CUDA_CHECK(launch_pdl(
      kern, dim3(a.grid), dim3(Cfg::kThreads), smem_bytes, st, *static_cast<const CUtensorMap*>(a.tm_wgu),
      *static_cast<const CUtensorMap*>(a.tm_wd), *static_cast<const CUtensorMap*>(a.tm_x),
      *static_cast<const CUtensorMap*>(a.tm_act), reinterpret_cast<const int4*>(a.tiles), a.num_tiles, a.row_token,
      a.row_weight, static_cast<T*>(a.act), static_cast<T*>(a.y), a.H, a.I, a.tokens, a.n_routed_rows,
      a.shared_expert, a.sh_lo, a.sh_hi, a.sched, a.ready, a.slot_row, static_cast<T*>(a.out), a.k, a.fdone,
      static_cast<const unsigned char*>(a.pf0), a.pf0_bytes, static_cast<const unsigned char*>(a.pf1), a.pf1_bytes,
      static_cast<const unsigned char*>(a.pf2), a.pf2_bytes, static_cast<const unsigned char*>(a.pf3), a.pf3_bytes,
      a.q, a.ep, a.rt));
  return GLM4MOE_OK;
}

template <typename T>
static int launch_fused_t(const DecodeFusedLaunch& a, cudaStream_t st) {
  if (a.fp8)
    return a.tile_rows == 16 ? launch_fused_one<T, 16, true>(a, st) : launch_fused_one<T, 32, true>(a, st);
  return a.tile_rows == 16 ? launch_fused_one<T, 16, false>(a, st) : launch_fused_one<T, 32, false>(a, st);
}

int decode_fused_launch(const DecodeFusedLaunch& a, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (a.tokens <= 0 || a.tokens > a.tile_rows || (a.tile_rows != 16 && a.tile_rows != 32)) return GLM4MOE_ERR_ARG;
  if (a.fp8 && (!a.q.s_gate_up || !a.q.s_down || !a.q.sx_t || !a.q.sact_t || !a.q.actq)) return GLM4MOE_ERR_ARG;
  if (a.rt.enabled && (a.fp8 || a.ep.enabled || a.tokens > 2 || a.rt.k * a.tokens > 32 || !a.rt.sync)) return GLM4MOE_ERR_ARG;
  if (a.dtype == GLM4MOE_BF16) return launch_fused_t<__nv_bfloat16>(a, st);
  if (a.dtype == GLM4MOE_F16) return launch_fused_t<__half>(a, st);
  return GLM4MOE_ERR_UNSUPPORTED;
}

}  // namespace glm4moe

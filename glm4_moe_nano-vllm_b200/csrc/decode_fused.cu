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
#include "umma_common.cuh"
#include "umma_gemm.cuh"

namespace glm4moe {

template <int BN>
struct FusedCfg {
  static constexpr int kAtomA = kBM * 128;          // 16 KB
  static constexpr int kAtomB = BN * 128;
  static constexpr int kStageBytes = 2 * kAtomA + 2 * kAtomB;
  static constexpr int kStages = (232448 - 2048) / kStageBytes;      // 6 (BN=16) / 5 (BN=32)
  static constexpr int kAccCols = 2 * BN;           // gate | up  (down uses the first half)
  static constexpr int kTmemCols = 2 * kAccCols <= 32 ? 32 : 2 * kAccCols <= 64 ? 64 : 128;
  static constexpr int kThreads = 256;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 512;
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

template <typename T, int BN>
__global__ void __launch_bounds__(FusedCfg<BN>::kThreads, 1) decode_fused_kernel(
    const __grid_constant__ CUtensorMap tm_wgu,   // gate_up stack [n_stack * 2I, H], box {64, 128}
    const __grid_constant__ CUtensorMap tm_wd,    // down stack    [n_stack * H, I],  box {64, 128}
    const __grid_constant__ CUtensorMap tm_x,     // the layer input [T, H],          box {64, BN}
    const __grid_constant__ CUtensorMap tm_act,   // activation rows [R, I],          box {64, BN}
    const int4* __restrict__ tiles, const int32_t* __restrict__ num_tiles_ptr,
    const int32_t* row_token, const float* row_weight, T* __restrict__ act, T* __restrict__ y, int H, int I,
    int tokens, int n_routed_rows, int shared_expert, int sh_lo, int sh_hi,
    int32_t* __restrict__ sched,                  // [0] queue head, [1] CTAs finished
    int32_t* __restrict__ ready,                  // [1 + tile capacity] gate_up completion counters
    const int32_t* slot_row, T* __restrict__ out, int k,
    int32_t* __restrict__ fdone,                  // [H/128] down items finished per feature tile
    const unsigned char* __restrict__ pf0, size_t pf0_bytes,    // shared expert's gate_up / down weights:
    const unsigned char* __restrict__ pf1, size_t pf1_bytes) {  // pulled into L2 while the router runs
  using Cfg = FusedCfg<BN>;
  extern __shared__ unsigned char smem_dyn[];
  unsigned char* smem = reinterpret_cast<unsigned char*>(
      (reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + Cfg::kStages;
  uint64_t* tmem_full = bars + 2 * Cfg::kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint64_t* q_full = tmem_empty + 2;
  uint64_t* q_empty = q_full + kFusedQueue;
  int32_t* q_w = reinterpret_cast<int32_t*>(q_empty + kFusedQueue);   // [depth][2]: work id, routed tiles
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(q_w + 2 * kFusedQueue);
  int32_t* s_last = reinterpret_cast<int32_t*>(tmem_ptr_smem + 1);   // epilogue-warp broadcast slot

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
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], 4);
    }
    for (int s = 0; s < kFusedQueue; ++s) {
      mbar_init(&q_full[s], 1);
      mbar_init(&q_empty[s], 1 + 4);
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
      const int4 tl = tiles[t];
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
      const int4 tl = tiles[t];
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
  const int atoms_gu = H / 64;    // K atoms (128 B = 64 two-byte elements) of gate_up
  const int atoms_d = I / 64;     // K atoms of down

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
          tma_load_2d(&tm_wgu, &full_bar[stage], sa, kb * 64, w_row, kEvictFirst);
          tma_load_2d(&tm_wgu, &full_bar[stage], sa + Cfg::kAtomA, kb * 64, w_row + I, kEvictFirst);
          tma_load_2d(&tm_x, &full_bar[stage], sa + 2 * Cfg::kAtomA, kb * 64, 0, kEvictLast);
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
            tma_load_2d(&tm_wd, &full_bar[s], sa + a * Cfg::kAtomA, (2 * j + a) * 64, w_row, kEvictFirst);
        };
        auto load_b = [&](int j, int s) {
          unsigned char* sa = smem + s * Cfg::kStageBytes + 2 * Cfg::kAtomA;
          const int na = min(2, atoms_d - 2 * j);
          for (int a = 0; a < na; ++a)
            tma_load_2d(&tm_act, &full_bar[s], sa + a * Cfg::kAtomB, (2 * j + a) * 64, it.row_start, kEvictLast);
        };
        const int n_pre = min(n_st, Cfg::kStages);
        int s = stage;
        uint32_t ph = phase;
        for (int j = 0; j < n_pre; ++j) {
          mbar_wait(&empty_bar[s], ph ^ 1u);
          load_w(j, s);
          if (++s == Cfg::kStages) { s = 0; ph ^= 1u; }
        }
        // acquire: all gate_up feature tiles of this expert have stored their activation columns
        const int target = 4 * FG;
        const int32_t* flag = ready + it.flag;
        int v;
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
      if (has_shared) {
        constexpr size_t kChunk = 32768;
        auto prefetch_slice = [&](const unsigned char* base, size_t bytes) {
          const size_t n_chunks = (bytes + kChunk - 1) / kChunk;
          for (size_t c = blockIdx.x; c < n_chunks; c += gridDim.x) {
            const size_t off = c * kChunk;
            const uint32_t sz = static_cast<uint32_t>(min(kChunk, bytes - off));
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(base + off), "r"(sz) : "memory");
          }
        };
        prefetch_slice(pf0, pf0_bytes);
        prefetch_slice(pf1, pf1_bytes);
      }
      if (has_shared)
        for (int f = blockIdx.x; f < FG; f += gridDim.x) {
          push(kSharedBase + f, 0);
          stream_gate_up(FusedItem{0, shared_expert, shared_row0, n_sh, f, 0, 1});
        }
      pdl_wait();
      pdl_launch();
      const int n_routed = *num_tiles_ptr - has_shared;
      const int total = n_routed * FG + (n_routed + has_shared) * FD;
      if (n_routed + has_shared == 0)   // expert-parallel rank without a hit expert: the partial output is zero
        for (int f = blockIdx.x; f < FD; f += gridDim.x) push(kZeroBase + f, 0);
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
        __threadfence();
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc_base = make_idesc(sizeof(T) == 2 && !std::is_same<T, __half>::value, 0);
    int stage = 0;
    uint32_t phase = 0;
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
      const uint32_t idesc = idesc_base | (static_cast<uint32_t>(n_eff >> 3) << 17);
      const int as = n & 1;
      mbar_wait(&tmem_empty[as], ((n >> 1) & 1u) ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * Cfg::kAccCols;
      const int n_st = it.kind == 0 ? atoms_gu : (atoms_d + 1) / 2;
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
      ++n;
    }
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    const int ew = warp - 4;
    const uint32_t lane_base = static_cast<uint32_t>(ew * 32) << 16;
    bool dep_waited = false;
    for (int qn = 0, n = 0;; ++qn) {
      int n_routed;
      const int w = next_work(qn, n_routed);
      if (w < 0) break;
      if (w >= kZeroBase && w < kSharedBase) {
        const int col = (w - kZeroBase) * kBM + 4 * lane;
        for (int t = ew; t < tokens; t += 4) *reinterpret_cast<uint2*>(out + static_cast<size_t>(t) * H + col) = make_uint2(0u, 0u);
        continue;
      }
      if (w < kSharedBase && !dep_waited) {
        pdl_wait();
        dep_waited = true;
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
          // release the activation columns to the down items of this expert (TMA = async proxy)
          asm volatile("fence.proxy.async.global;" ::: "memory");
          __threadfence();
          atomicAdd(ready + it.flag, 1);
        }
      } else {
        T* out_col = y + static_cast<size_t>(it.row_start) * H + feat;
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
        // ---- combine (moe.py:151,171), fused: the LAST down item of feature tile f sums, for every
        // token, its k routed rows in ascending expert order + the shared row over these 128
        // features.  Same order and single rounding as combine_kernel: bit-identical, no atomics on data.
        __threadfence();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (ew == 0 && lane == 0) {
          const int done = atomicAdd(fdone + it.f, 1) + 1;
          const int last = done == n_routed + has_shared;
          if (last) fdone[it.f] = 0;          // re-armed for the next launch
          *s_last = last;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        const int last = *reinterpret_cast<volatile int32_t*>(s_last);
        asm volatile("bar.sync 1, 128;" ::: "memory");   // slot may be rewritten by the next item
        if (last) {
          __threadfence();
          const size_t col = static_cast<size_t>(it.f) * kBM + 4 * lane;
          for (int t = ew; t < tokens; t += 4) {
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            auto add_row = [&](int r) {
              const uint2 raw = __ldcg(reinterpret_cast<const uint2*>(y + static_cast<size_t>(r) * H + col));
              const T* h4 = reinterpret_cast<const T*>(&raw);
#pragma unroll
              for (int q = 0; q < 4; ++q) acc[q] += Dt<T>::to_f(h4[q]);
            };
            for (int sl = 0; sl < k; ++sl) {
              const int r = __ldcg(slot_row + static_cast<size_t>(t) * k + sl);
              if (r >= 0) add_row(r);
            }
            if (t >= sh_lo && t < sh_hi) add_row(n_routed_rows + t);
            __align__(8) T o[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) o[q] = Dt<T>::from_f(acc[q]);
            *reinterpret_cast<uint2*>(out + static_cast<size_t>(t) * H + col) = *reinterpret_cast<uint2*>(o);
          }
        }
      }
      ++n;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
}

template <typename T, int BN>
static int launch_fused_one(const DecodeFusedLaunch& a, cudaStream_t st) {
  using Cfg = FusedCfg<BN>;
  auto kern = decode_fused_kernel<T, BN>;
  static bool attr_set = false;
  if (!attr_set) {
    GLM4MOE_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    attr_set = true;
  }
  GLM4MOE_CUDA_CHECK(launch_pdl(
      kern, dim3(a.grid), dim3(Cfg::kThreads), Cfg::kSmemBytes, st, *static_cast<const CUtensorMap*>(a.tm_wgu),
      *static_cast<const CUtensorMap*>(a.tm_wd), *static_cast<const CUtensorMap*>(a.tm_x),
      *static_cast<const CUtensorMap*>(a.tm_act), reinterpret_cast<const int4*>(a.tiles), a.num_tiles, a.row_token,
      a.row_weight, static_cast<T*>(a.act), static_cast<T*>(a.y), a.H, a.I, a.tokens, a.n_routed_rows,
      a.shared_expert, a.sh_lo, a.sh_hi, a.sched, a.ready, a.slot_row, static_cast<T*>(a.out), a.k, a.fdone,
      static_cast<const unsigned char*>(a.pf0), a.pf0_bytes, static_cast<const unsigned char*>(a.pf1), a.pf1_bytes));
  return GLM4MOE_OK;
}

int decode_fused_launch(const DecodeFusedLaunch& a, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (a.tokens <= 0 || a.tokens > a.tile_rows || (a.tile_rows != 16 && a.tile_rows != 32)) return GLM4MOE_ERR_ARG;
  if (a.dtype == GLM4MOE_BF16)
    return a.tile_rows == 16 ? launch_fused_one<__nv_bfloat16, 16>(a, st) : launch_fused_one<__nv_bfloat16, 32>(a, st);
  if (a.dtype == GLM4MOE_F16)
    return a.tile_rows == 16 ? launch_fused_one<__half, 16>(a, st) : launch_fused_one<__half, 32>(a, st);
  return GLM4MOE_ERR_UNSUPPORTED;
}

}  // namespace glm4moe

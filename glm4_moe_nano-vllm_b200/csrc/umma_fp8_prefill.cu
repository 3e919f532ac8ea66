// FP8 (e4m3) grouped expert GEMMs for the prefill regime (hundreds of rows per expert), sm_100a.
//
// Same arithmetic as umma_fp8.cu (128x128 fp32 weight block scales, per-token 1x128 activation
// scales, every k-block's raw tcgen05 `kind::f8f6f4` product promoted in fp32), different geometry:
// the TOKEN rows are the M=128 operand and the WEIGHT rows the N=256 operand,
//     D[token, feature] = sum_k Xq[row(token), k] * Wq[e][feature, k],
// so after tcgen05.ld a thread holds ONE token row.  Its activation scale is then a per-thread
// scalar and the weight scale a warp-uniform scalar: the promotion is exactly one FFMA per element
// (the swap-AB kernel needs a shuffle per token column and is limited to 64-token tiles, which
// re-streams every weight slab 4+ times at prefill).  A tile is 128 tokens x 256 weight rows:
//   gate_up: N = [128 gate rows | 128 up rows] of feature tile f; gate[c] and up[c] of a token
//            sit in the same thread, so SiLU*mul, the rounding to the model dtype and the e4m3
//            re-quantisation for the down projection (one 1x128 group per token per tile) are
//            fused into the epilogue: the bf16 activation rows never exist in HBM.
//   down:    N = 256 output features; epilogue multiplies by the routing weight (moe.py:149).
// Replaces mlp.py:50-52 for the quant tier (BASELINE.json configs[2], prefill 4k tokens).
//
// CTA = 20 warps: warp 0 TMA producer + dynamic tile scheduler, warp 1 MMA issuer, warp 2 TMEM
// allocator, warps 4-19 epilogue (4 per TMEM lane quadrant, 64 accumulator columns each).
// Smem ring: 4 stages x (16 KB tokens + 32 KB weights).  TMEM: 2 x 256 columns of raw partials.
#include <cuda_fp8.h>

#include "umma_common.cuh"
#include "umma_gemm.cuh"

namespace glm4moe {

namespace {

constexpr int kPK = 128;                         // e4m3 elements (= bytes) per k-block = one scale block
constexpr int kPM = 128;                         // token rows per tile (MMA M)
constexpr int kPN = 256;                         // weight rows per tile (MMA N)
constexpr int kPStages = 4;
constexpr int kPABytes = kPM * kPK;              // 16 KB
constexpr int kPBBytes = kPN * kPK;              // 32 KB
constexpr int kPStageBytes = kPABytes + kPBBytes;
constexpr int kPEpiWarps = 16;
constexpr int kPEpiThreads = 32 * kPEpiWarps;
constexpr int kPThreads = 128 + kPEpiThreads;    // 640
constexpr int kPAmaxBytes = 2 * 4 * kPM * 4;     // [tile parity][column group][row] fp32
constexpr int kPSmemBytes = kPStages * kPStageBytes + 1024 /*align*/ + 512 /*barriers*/ + kPAmaxBytes;
constexpr int kPSchedDepth = 4;

__device__ __forceinline__ void epi_bar_sync() {   // the 16 epilogue warps only
  asm volatile("bar.sync 1, %0;" ::"n"(kPEpiThreads) : "memory");
}

__device__ __forceinline__ uint32_t pack_e4m3x4(float a, float b, float c, float d) {
  return static_cast<uint32_t>(__nv_cvt_float_to_fp8(a, __NV_SATFINITE, __NV_E4M3)) |
         static_cast<uint32_t>(__nv_cvt_float_to_fp8(b, __NV_SATFINITE, __NV_E4M3)) << 8 |
         static_cast<uint32_t>(__nv_cvt_float_to_fp8(c, __NV_SATFINITE, __NV_E4M3)) << 16 |
         static_cast<uint32_t>(__nv_cvt_float_to_fp8(d, __NV_SATFINITE, __NV_E4M3)) << 24;
}

template <typename T>
__device__ __forceinline__ uint32_t pack2(float a, float b);
template <>
__device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
template <>
__device__ __forceinline__ uint32_t pack2<__half>(float a, float b) {
  __half2 v = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

}  // namespace

template <typename T, bool GATED>
__global__ void __launch_bounds__(kPThreads, 1) umma_fp8_prefill_kernel(
    const __grid_constant__ CUtensorMap tm_w,   // e4m3 weight stack [rows_w_total, K] bytes, box {128, 128}
    const __grid_constant__ CUtensorMap tm_x,   // e4m3 activation rows [rows, K], box {128, 128}
    const float* __restrict__ w_scale,          // [n_stack, w_rows_per_expert/128, K/128]
    const float* __restrict__ x_scale_t,        // [K/128, scale_ld]   (k-block major, row minor)
    int scale_ld,
    const int4* __restrict__ tiles, const int32_t* __restrict__ num_tiles_ptr,
    T* __restrict__ out, int ld_out,            // down: y[row, H]
    uint8_t* __restrict__ out_q,                // gate_up: e4m3 activation rows [row, I]
    float* __restrict__ out_scale_t,            // gate_up: their scales [I/128, scale_ld]
    const float* __restrict__ row_weight, int n_routed_rows,
    int n_feat_tiles,                           // gate_up: I/128, down: H/256
    int n_kblocks, int w_rows_per_expert, int up_row_offset, int32_t* __restrict__ sched,
    int pow2, uint8_t* __restrict__ out_sf) {   // gate_up: power-of-two scales (+ packed UE8M0 copy, nullable)
  extern __shared__ unsigned char smem_dyn[];
  unsigned char* smem = reinterpret_cast<unsigned char*>(
      (reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kPStages * kPStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kPStages;
  uint64_t* tmem_full = bars + 2 * kPStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint64_t* sched_full = tmem_empty + 2;
  uint64_t* sched_empty = sched_full + kPSchedDepth;
  int32_t* sched_w = reinterpret_cast<int32_t*>(sched_empty + kPSchedDepth);
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(sched_w + kPSchedDepth);
  float* amax_x = reinterpret_cast<float*>(smem + kPStages * kPStageBytes + 512);   // [2][4][kPM]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tm_w);
    prefetch_tmap(&tm_x);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kPStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], kPEpiWarps);
    }
    for (int s = 0; s < kPSchedDepth; ++s) {
      mbar_init(&sched_full[s], 1);
      mbar_init(&sched_empty[s], 1 + kPEpiWarps);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_ptr_smem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  const int n_tiles = *num_tiles_ptr;
  const int total_work = n_tiles * n_feat_tiles;

  auto next_work = [&](int n) -> int {
    const int q = n % kPSchedDepth;
    mbar_wait(&sched_full[q], (n / kPSchedDepth) & 1u);
    const int w = sched_w[q];
    __syncwarp();
    if (lane == 0) mbar_arrive(&sched_empty[q]);
    return w;
  };

  if (warp == 0) {
    // ===================== TMA producer + dynamic tile scheduler =====================
    if (elect_one()) {
      int n_pushed = 0;
      auto push_work = [&](int w) {
        const int q = n_pushed % kPSchedDepth;
        mbar_wait(&sched_empty[q], ((n_pushed / kPSchedDepth) & 1u) ^ 1u);
        sched_w[q] = w;
        mbar_arrive(&sched_full[q]);
        ++n_pushed;
      };
      WorkRef wr{0, 0, 0, 0};
      auto issue = [&](int kb, int stage, bool do_w, bool do_x) {
        unsigned char* sa = smem + stage * kPStageBytes;
        if (do_w) {
          // gate_up: weight rows [f*128, +128) and the matching up_proj rows; down: rows [f*256, +256)
          const int w_row = wr.expert * w_rows_per_expert + wr.f * (GATED ? 128 : 256);
          mbar_expect_tx(&full_bar[stage], kPStageBytes);
          tma_load_2d(&tm_w, &full_bar[stage], sa + kPABytes, kb * kPK, w_row, kEvictFirst);
          tma_load_2d(&tm_w, &full_bar[stage], sa + kPABytes + kPBBytes / 2, kb * kPK,
                      w_row + (GATED ? up_row_offset : 128), kEvictFirst);
        }
        if (do_x) tma_load_2d(&tm_x, &full_bar[stage], sa, kb * kPK, wr.row_start, kEvictLast);
      };
      int w = static_cast<int>(blockIdx.x) < total_work ? static_cast<int>(blockIdx.x) : -1;
      push_work(w);
      if (w >= 0) wr = get_work<kPM>(tiles, w, n_feat_tiles, n_tiles, 0);
      // weight slabs of the first ring pass before the dependency wait (weights and the tile list
      // are at least two launches old, see common.cuh); the token tiles after it
      int n_pre = 0;
      if (w >= 0)
        for (; n_pre < kPStages && n_pre < n_kblocks; ++n_pre) issue(n_pre, n_pre, true, false);
      pdl_wait();
      pdl_launch();
      for (int i = 0; i < n_pre; ++i) issue(i, i, false, true);
      int kb = n_pre;
      int stage = n_pre == kPStages ? 0 : n_pre;
      uint32_t phase = n_pre == kPStages ? 1u : 0u;
      while (w >= 0) {
        if (kb == n_kblocks) {
          kb = 0;
          const int nw = atomicAdd(&sched[0], 1) + static_cast<int>(gridDim.x);
          w = nw < total_work ? nw : -1;
          push_work(w);
          if (w >= 0) wr = get_work<kPM>(tiles, w, n_feat_tiles, n_tiles, 0);
          continue;
        }
        mbar_wait(&empty_bar[stage], phase ^ 1u);
        issue(kb, stage, true, true);
        ++kb;
        if (++stage == kPStages) { stage = 0; phase ^= 1u; }
      }
      __threadfence();
      if (atomicAdd(&sched[1], 1) == static_cast<int>(gridDim.x) - 1) {
        sched[0] = 0;
        sched[1] = 0;
        __threadfence();
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer: one fresh 128 x 256 partial per k-block =====================
    constexpr uint32_t idesc = make_idesc_f8(kPN);
    int stage = 0;
    uint32_t phase = 0;
    uint32_t pcount = 0;
    for (int it = 0;; ++it) {
      const int w = next_work(it);
      if (w < 0) break;
      for (int kb = 0; kb < n_kblocks; ++kb, ++pcount) {
        const int ps = pcount & 1u;
        mbar_wait(&tmem_empty[ps], ((pcount >> 1) & 1u) ^ 1u);
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(smem + stage * kPStageBytes);
          const uint64_t a_desc = make_kmajor_desc(sa);              // tokens  [128 x 128 B]
          const uint64_t b_desc = make_kmajor_desc(sa + kPABytes);   // weights [256 x 128 B]
          const uint32_t d_tmem = tmem_base + ps * kPN;
#pragma unroll
          for (int ks = 0; ks < kPK / 32; ++ks)
            umma_f8(d_tmem, a_desc + 2 * ks, b_desc + 2 * ks, idesc, ks > 0 ? 1u : 0u);
          umma_commit(&empty_bar[stage]);
          umma_commit(&tmem_full[ps]);
        }
        __syncwarp();
        if (++stage == kPStages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue: promote every k-block, then the fused output math =====================
    // warp -> TMEM lane quadrant q (hardware: warp id % 4) and column group cg:
    //   gate_up: gate columns [32cg, 32cg+32) and up columns [128+32cg, +32)  (32 outputs, 64 accumulators)
    //   down:    columns [64cg, 64cg+64)                                      (64 outputs)
    const int q = warp & 3;
    const int cg = (warp - 4) >> 2;
    const uint32_t lane_base = static_cast<uint32_t>(q * 32) << 16;
    const int r = q * 32 + lane;                    // token row inside the tile held by this thread
    const int wblocks = w_rows_per_expert / 128;    // scale block rows per expert
    uint32_t pcount = 0;
    // the activation rows and their scales are written by the direct predecessor
    pdl_wait();
    for (int it = 0;; ++it) {
      const int w = next_work(it);
      if (w < 0) break;
      const WorkRef tile = get_work<kPM>(tiles, w, n_feat_tiles, n_tiles, 0);
      const bool row_ok = r < tile.nrows;
      const int row = tile.row_start + r;
      float acc[64];
#pragma unroll
      for (int j = 0; j < 64; ++j) acc[j] = 0.f;
      // weight scale rows: gate_up -> block rows f (gate) and I/128 + f (up); down -> block row 2f + cg/2
      const float* sw0 = w_scale + (static_cast<size_t>(tile.expert) * wblocks +
                                    (GATED ? tile.f : 2 * tile.f + (cg >> 1))) * n_kblocks;
      const float* sw1 = GATED ? sw0 + static_cast<size_t>(up_row_offset / 128) * n_kblocks : sw0;
      const float* sx = x_scale_t + row;
      const int col0 = GATED ? 32 * cg : 64 * cg;
      const int col1 = GATED ? 128 + 32 * cg : 64 * cg + 32;
      // scales are software-pipelined one k-block ahead: their L2 latency (~700 cycles under load)
      // must not sit between two promotions, which gate the MMA's reuse of the partial buffers
      float sx_n = row_ok ? __ldg(sx) : 0.f;
      float sw0_n = __ldg(sw0);
      float sw1_n = GATED ? __ldg(sw1) : 0.f;
      for (int kb = 0; kb < n_kblocks; ++kb, ++pcount) {
        const int ps = pcount & 1u;
        const float s0 = sx_n * sw0_n;
        const float s1 = GATED ? sx_n * sw1_n : s0;
        if (kb + 1 < n_kblocks) {
          sx_n = row_ok ? __ldg(sx + static_cast<size_t>(kb + 1) * scale_ld) : 0.f;
          sw0_n = __ldg(sw0 + kb + 1);
          if (GATED) sw1_n = __ldg(sw1 + kb + 1);
        }
        mbar_wait(&tmem_full[ps], (pcount >> 1) & 1u);
        tc_fence_after();
        const uint32_t t_acc = tmem_base + lane_base + ps * kPN;
        uint32_t a[16], b[16];
        tmem_ld16(t_acc + col0, a);
        tmem_ld16(t_acc + col1, b);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          acc[j] = fmaf(__uint_as_float(a[j]), s0, acc[j]);
          acc[32 + j] = fmaf(__uint_as_float(b[j]), s1, acc[32 + j]);
        }
        tmem_ld16(t_acc + col0 + 16, a);
        tmem_ld16(t_acc + col1 + 16, b);
        tmem_ld_wait();
        // both halves are in registers: hand the partial back before the last FMAs
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty[ps]);
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          acc[16 + j] = fmaf(__uint_as_float(a[j]), s0, acc[16 + j]);
          acc[48 + j] = fmaf(__uint_as_float(b[j]), s1, acc[48 + j]);
        }
      }
      if (GATED) {
        // acc[0..31] = gate, acc[32..63] = up of output features f*128 + 32cg + j.
        // SiLU*mul (activation.py:13-14), one rounding to the model dtype (the reference stores the
        // intermediate in it), then the 1x128 e4m3 quantiser of the down projection's input:
        // the group = this tile's 128 features of the token = the 4 column groups of the quadrant.
        float v[32];
        float amax = 0.f;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          v[j] = Dt<T>::round(silu_fast(acc[j]) * acc[32 + j]);
          amax = fmaxf(amax, fabsf(v[j]));
        }
        float* ax = amax_x + (it & 1) * 4 * kPM;
        ax[cg * kPM + r] = amax;
        epi_bar_sync();
        amax = fmaxf(fmaxf(ax[r], ax[kPM + r]), fmaxf(ax[2 * kPM + r], ax[3 * kPM + r]));
        uint32_t ebyte = 0;
        const float sc = pow2 ? pow2_scale(amax, &ebyte) : __fdiv_rn(fmaxf(amax, 1e-10f), 448.0f);
        if (row_ok) {
          uint32_t p[8];
#pragma unroll
          for (int j = 0; j < 8; ++j)
            p[j] = pack_e4m3x4(__fdiv_rn(v[4 * j], sc), __fdiv_rn(v[4 * j + 1], sc), __fdiv_rn(v[4 * j + 2], sc),
                               __fdiv_rn(v[4 * j + 3], sc));
          uint4* dst = reinterpret_cast<uint4*>(out_q + static_cast<size_t>(row) * ld_out + tile.f * 128 + 32 * cg);
          dst[0] = make_uint4(p[0], p[1], p[2], p[3]);
          dst[1] = make_uint4(p[4], p[5], p[6], p[7]);
          if (cg == 0) {
            out_scale_t[static_cast<size_t>(tile.f) * scale_ld + row] = sc;
            if (out_sf) out_sf[(static_cast<size_t>(tile.f >> 2) * scale_ld + row) * 4 + (tile.f & 3)] = static_cast<uint8_t>(ebyte);
          }
        }
      } else {
        // y = acc * routing weight (mlp.py:52, moe.py:149); shared-expert rows carry weight 1
        if (row_ok) {
          const float rw = row < n_routed_rows ? __ldg(row_weight + row) : 1.0f;
          uint4* dst = reinterpret_cast<uint4*>(out + static_cast<size_t>(row) * ld_out + tile.f * 256 + 64 * cg);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            dst[j] = make_uint4(pack2<T>(acc[8 * j] * rw, acc[8 * j + 1] * rw),
                                pack2<T>(acc[8 * j + 2] * rw, acc[8 * j + 3] * rw),
                                pack2<T>(acc[8 * j + 4] * rw, acc[8 * j + 5] * rw),
                                pack2<T>(acc[8 * j + 6] * rw, acc[8 * j + 7] * rw));
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<512>(tmem_base);
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
template <typename T, bool GATED>
static int launch_prefill_one(const UmmaFp8PrefillLaunch& a, cudaStream_t st) {
  auto kern = umma_fp8_prefill_kernel<T, GATED>;
  static unsigned long long attr_done = 0;   // per instantiation, one bit per device
  GLM4MOE_CUDA_CHECK(ensure_dynamic_smem(kern, kPSmemBytes, &attr_done));
  GLM4MOE_CUDA_CHECK(launch_pdl(kern, dim3(a.grid), dim3(kPThreads), kPSmemBytes, st,
                                *static_cast<const CUtensorMap*>(a.tm_w), *static_cast<const CUtensorMap*>(a.tm_x),
                                a.w_scale, a.x_scale_t, a.scale_ld, reinterpret_cast<const int4*>(a.tiles),
                                a.num_tiles, static_cast<T*>(a.out), a.ld_out, static_cast<uint8_t*>(a.out_q),
                                a.out_scale_t, a.row_weight, a.n_routed_rows, a.n_feat_tiles, a.n_kblocks,
                                a.w_rows_per_expert, a.up_row_offset, a.sched, a.pow2,
                                static_cast<uint8_t*>(a.out_sf)));
  return GLM4MOE_OK;
}

bool umma_fp8_prefill_shape_ok(int H, int I) { return H % 256 == 0 && I % 128 == 0; }

int umma_fp8_prefill_launch(const UmmaFp8PrefillLaunch& a, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (a.gated ? (!a.out_q || !a.out_scale_t) : !a.out) return GLM4MOE_ERR_ARG;
  if (a.dtype == GLM4MOE_BF16)
    return a.gated ? launch_prefill_one<__nv_bfloat16, true>(a, st) : launch_prefill_one<__nv_bfloat16, false>(a, st);
  if (a.dtype == GLM4MOE_F16)
    return a.gated ? launch_prefill_one<__half, true>(a, st) : launch_prefill_one<__half, false>(a, st);
  return GLM4MOE_ERR_UNSUPPORTED;
}

}  // namespace glm4moe

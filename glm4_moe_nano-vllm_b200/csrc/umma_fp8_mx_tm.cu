// FP8 (e4m3) gate_up GEMM of the prefill regime with HARDWARE block scaling, token-major tiles, sm_100a.
//
// Same scheme as umma_fp8_mx.cu (tcgen05.mma.kind::mxf8f6f4.block_scale, UE8M0 scales staged to TMEM by
// tcgen05.cp, one TMEM accumulator per tile, read once), other orientation: tokens = MMA M (128), weight rows
// = MMA N (256: gate 128 + up 128 of one feature tile) as ONE instruction per K = 32 slice.  A thread of the
// epilogue owns one token row, so SiLU*mul, the per-token amax, the power-of-two re-quantisation and two
// 16-byte stores are all thread-local (the feature-major epilogue of umma_fp8_mx.cu needs a warp reduction per
// token and byte stores), and each operand byte is read from shared memory once per k-block (the
// two-slab form reads the token tile twice).  Measured (profiles/r2*.md, Air): gate_up 0.498 / 0.810 ms at
// T = 4096 / 8192 against 0.520 / 0.995 ms for the feature-major kernel, although a ragged 14-row expert tail
// costs a full 128-row tile here; the down projection (no re-quantiser, K = 1408) is faster feature-major
// (0.291 vs 0.325 ms), so api.cu uses this kernel for gate_up and umma_fp8_mx.cu for down.
//
// Scale operands: SFA = the tile's 128 token rows (4 TMEM columns), SFB = the 256 weight rows (8 columns, one
// value per 128-row slab), per 4 consecutive k-blocks one uint32 per row; see umma_fp8_mx.cu.
#include <cuda_fp8.h>

#include "umma_common.cuh"
#include "umma_gemm.cuh"

namespace glm4moe {

namespace {

constexpr int kXK = 128;                         // e4m3 elements (= bytes) per k-block = one scale block
constexpr int kXM = 128;                         // token rows per tile (MMA M)
constexpr int kXN = 256;                         // weight rows per tile (MMA N)
constexpr int kXStages = 4;
constexpr int kXABytes = kXM * kXK;              // 16 KB
constexpr int kXBBytes = kXN * kXK;              // 32 KB
constexpr int kXStageBytes = kXABytes + kXBBytes;
constexpr int kXEpiWarps = 16;
constexpr int kXEpiThreads = 32 * kXEpiWarps;
constexpr int kXThreads = 128 + kXEpiThreads;    // 640
constexpr int kXSfSlotBytes = (kXM + kXN) * 4;   // 128 SFA words + 256 SFB words
constexpr int kXAmaxBytes = 2 * 4 * kXM * 4;     // [tile parity][column group][row] fp32
constexpr int kXBarBytes = 512;
constexpr int kXSmemBytes = kXStages * kXStageBytes + 1024 /*align*/ + kXBarBytes + 2 * kXSfSlotBytes + kXAmaxBytes;
constexpr int kXSchedDepth = 4;
constexpr uint32_t kXSfCol0 = 256;               // first scale-factor column in TMEM
constexpr uint32_t kXSfColStride = 16;           // per slot: SFA at +0 (4 columns), SFB at +4 (8 columns)

__device__ __forceinline__ void xepi_bar_sync() {   // the 16 epilogue warps only
  asm volatile("bar.sync 1, %0;" ::"n"(kXEpiThreads) : "memory");
}

__device__ __forceinline__ uint32_t xpack_e4m3x4(float a, float b, float c, float d) {
  return static_cast<uint32_t>(__nv_cvt_float_to_fp8(a, __NV_SATFINITE, __NV_E4M3)) |
         static_cast<uint32_t>(__nv_cvt_float_to_fp8(b, __NV_SATFINITE, __NV_E4M3)) << 8 |
         static_cast<uint32_t>(__nv_cvt_float_to_fp8(c, __NV_SATFINITE, __NV_E4M3)) << 16 |
         static_cast<uint32_t>(__nv_cvt_float_to_fp8(d, __NV_SATFINITE, __NV_E4M3)) << 24;
}

template <typename T>
__device__ __forceinline__ uint32_t xpack2(float a, float b);
template <>
__device__ __forceinline__ uint32_t xpack2<__nv_bfloat16>(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
template <>
__device__ __forceinline__ uint32_t xpack2<__half>(float a, float b) {
  __half2 v = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// D[tmem] (+)= (A * SFA) x (B * SFB): e4m3 operands, UE8M0 scale per row per K = 32, fp32 accumulate
__device__ __forceinline__ void umma_f8_block_scaled(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                                     uint32_t sfa_tmem, uint32_t sfb_tmem, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf8f6f4.block_scale [%0], %1, %2, %3, [%5], [%6], p;\n\t"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate), "r"(sfa_tmem), "r"(sfb_tmem)
      : "memory");
}
// instruction descriptor, kind::mxf8f6f4.block_scale: A/B e4m3 (format 0) K-major, UE8M0 scales,
// M = 128, N = n, scale-factor byte ids of this K = 32 slice's k-block
__device__ __forceinline__ uint32_t make_idesc_mx(int n, uint32_t a_sf_id, uint32_t b_sf_id) {
  return (b_sf_id << 4) | (static_cast<uint32_t>(n >> 3) << 17) | (1u << 23) | (static_cast<uint32_t>(128 >> 4) << 24) |
         (a_sf_id << 29);
}
// shared memory -> tensor memory: 32 rows x 128 bits, replicated to the 4 lane quadrants
__device__ __forceinline__ void utccp_32x128b_x4(uint32_t tmem_dst, uint64_t smem_desc) {
  asm volatile("tcgen05.cp.cta_group::1.32x128b.warpx4 [%0], %1;" ::"r"(tmem_dst), "l"(smem_desc) : "memory");
}
// K-major, no swizzle, 8-row groups 128 bytes apart (one 16-byte atom along K)
__device__ __forceinline__ uint64_t make_sf_desc(uint32_t saddr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr >> 4) & 0x3fffu);
  d |= static_cast<uint64_t>(128u >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  return d;
}

}  // namespace

template <typename T, bool GATED>
__global__ void __launch_bounds__(kXThreads, 1) umma_fp8_mx_tm_kernel(
    const __grid_constant__ CUtensorMap tm_w,   // e4m3 weight stack [rows_w_total, K] bytes, box {128, 128}
    const __grid_constant__ CUtensorMap tm_x,   // e4m3 activation rows [rows, K], box {128, 128}
    const uint32_t* __restrict__ w_sf,          // [n_stack, w_rows_per_expert/128, n_kgroups] packed UE8M0
    const uint32_t* __restrict__ x_sf,          // [n_kgroups, scale_ld] packed UE8M0 (k-group major, row minor)
    int scale_ld,
    const int4* __restrict__ tiles, const int32_t* __restrict__ num_tiles_ptr,
    T* __restrict__ out, int ld_out,            // down: y[row, H]
    uint8_t* __restrict__ out_q,                // gate_up: e4m3 activation rows [row, I]
    float* __restrict__ out_scale_t,            // gate_up: their fp32 scales [I/128, scale_ld]
    uint8_t* __restrict__ out_sf,               // gate_up: the same as packed UE8M0 bytes [ceil(I/512), scale_ld] x 4
    const float* __restrict__ row_weight, int n_routed_rows,
    int n_feat_tiles,                           // gate_up: I/128, down: H/256
    int n_kblocks, int w_rows_per_expert, int up_row_offset, int32_t* __restrict__ sched) {
  extern __shared__ unsigned char smem_dyn[];
  unsigned char* smem = reinterpret_cast<unsigned char*>(
      (reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kXStages * kXStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kXStages;
  uint64_t* tmem_full = bars + 2 * kXStages;           // [1]
  uint64_t* tmem_empty = tmem_full + 1;                // [1]
  uint64_t* sf_full = tmem_empty + 1;                  // [2]
  uint64_t* sf_empty = sf_full + 2;                    // [2]
  uint64_t* sched_full = sf_empty + 2;
  uint64_t* sched_empty = sched_full + kXSchedDepth;
  int32_t* sched_w = reinterpret_cast<int32_t*>(sched_empty + kXSchedDepth);
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(sched_w + kXSchedDepth);
  uint32_t* sf_smem = reinterpret_cast<uint32_t*>(smem + kXStages * kXStageBytes + kXBarBytes);   // [2][128 + 256]
  float* amax_x = reinterpret_cast<float*>(smem + kXStages * kXStageBytes + kXBarBytes + 2 * kXSfSlotBytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int n_kgroups = (n_kblocks + 3) >> 2;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tm_w);
    prefetch_tmap(&tm_x);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kXStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&tmem_full[0], 1);
    mbar_init(&tmem_empty[0], kXEpiWarps);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&sf_full[s], 1);
      mbar_init(&sf_empty[s], 1);
    }
    for (int s = 0; s < kXSchedDepth; ++s) {
      mbar_init(&sched_full[s], 1);
      mbar_init(&sched_empty[s], 2 + kXEpiWarps);   // MMA warp, scale-factor warp, epilogue warps
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
    const int q = n % kXSchedDepth;
    mbar_wait(&sched_full[q], (n / kXSchedDepth) & 1u);
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
        const int q = n_pushed % kXSchedDepth;
        mbar_wait(&sched_empty[q], ((n_pushed / kXSchedDepth) & 1u) ^ 1u);
        sched_w[q] = w;
        mbar_arrive(&sched_full[q]);
        ++n_pushed;
      };
      WorkRef wr{0, 0, 0, 0};
      auto issue = [&](int kb, int stage, bool do_w, bool do_x) {
        unsigned char* sa = smem + stage * kXStageBytes;
        if (do_w) {
          // gate_up: weight rows [f*128, +128) and the matching up_proj rows; down: rows [f*256, +256)
          const int w_row = wr.expert * w_rows_per_expert + wr.f * (GATED ? 128 : 256);
          mbar_expect_tx(&full_bar[stage], kXStageBytes);
          tma_load_2d(&tm_w, &full_bar[stage], sa + kXABytes, kb * kXK, w_row, kEvictFirst);
          tma_load_2d(&tm_w, &full_bar[stage], sa + kXABytes + kXBBytes / 2, kb * kXK,
                      w_row + (GATED ? up_row_offset : 128), kEvictFirst);
        }
        if (do_x) tma_load_2d(&tm_x, &full_bar[stage], sa, kb * kXK, wr.row_start, kEvictLast);
      };
      int w = static_cast<int>(blockIdx.x) < total_work ? static_cast<int>(blockIdx.x) : -1;
      push_work(w);
      if (w >= 0) wr = get_work<kXM>(tiles, w, n_feat_tiles, n_tiles, 0);
      // weight slabs of the first ring pass before the dependency wait (weights and the tile list
      // are at least two launches old, see common.cuh); the token tiles after it
      int n_pre = 0;
      if (w >= 0)
        for (; n_pre < kXStages && n_pre < n_kblocks; ++n_pre) issue(n_pre, n_pre, true, false);
      pdl_wait();
      pdl_launch();
      for (int i = 0; i < n_pre; ++i) issue(i, i, false, true);
      int kb = n_pre;
      int stage = n_pre == kXStages ? 0 : n_pre;
      uint32_t phase = n_pre == kXStages ? 1u : 0u;
      while (w >= 0) {
        if (kb == n_kblocks) {
          kb = 0;
          const int nw = atomicAdd(&sched[0], 1) + static_cast<int>(gridDim.x);
          w = nw < total_work ? nw : -1;
          push_work(w);
          if (w >= 0) wr = get_work<kXM>(tiles, w, n_feat_tiles, n_tiles, 0);
          continue;
        }
        mbar_wait(&empty_bar[stage], phase ^ 1u);
        issue(kb, stage, true, true);
        ++kb;
        if (++stage == kXStages) { stage = 0; phase ^= 1u; }
      }
      __threadfence();
      if (atomicAdd(&sched[1], 1) == static_cast<int>(gridDim.x) - 1) {
        sched[0] = 0;
        sched[1] = 0;
        __threadfence();
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer: scale factors to TMEM, then 4 block-scaled MMAs per k-block =====================
    int stage = 0;
    uint32_t phase = 0;
    uint32_t gcount = 0;      // k-groups consumed so far (scale-factor slot ring)
    for (int it = 0;; ++it) {
      const int w = next_work(it);
      if (w < 0) break;
      mbar_wait(&tmem_empty[0], (it & 1u) ^ 1u);
      tc_fence_after();
      for (int kb = 0; kb < n_kblocks; ++kb) {
        const int slot = gcount & 1u;
        const uint32_t sfa_t = tmem_base + kXSfCol0 + slot * kXSfColStride;
        const uint32_t sfb_t = sfa_t + 4;
        if ((kb & 3) == 0) {
          mbar_wait(&sf_full[slot], (gcount >> 1) & 1u);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t ss = smem_u32(sf_smem + slot * (kXM + kXN));
            utccp_32x128b_x4(sfa_t, make_sf_desc(ss));                 // 128 token rows   -> 4 columns
            utccp_32x128b_x4(sfb_t, make_sf_desc(ss + 512));           // weight rows 0..127   -> 4 columns
            utccp_32x128b_x4(sfb_t + 4, make_sf_desc(ss + 1024));      // weight rows 128..255 -> 4 columns
          }
          __syncwarp();
        }
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(smem + stage * kXStageBytes);
          const uint64_t a_desc = make_kmajor_desc(sa);              // tokens  [128 x 128 B]
          const uint64_t b_desc = make_kmajor_desc(sa + kXABytes);   // weights [256 x 128 B]
          const uint32_t idesc = make_idesc_mx(kXN, kb & 3, kb & 3);
#pragma unroll
          for (int ks = 0; ks < kXK / 32; ++ks)
            umma_f8_block_scaled(tmem_base, a_desc + 2 * ks, b_desc + 2 * ks, idesc, sfa_t, sfb_t,
                                 (kb > 0 || ks > 0) ? 1u : 0u);
          umma_commit(&empty_bar[stage]);
          if ((kb & 3) == 3 || kb == n_kblocks - 1) umma_commit(&sf_empty[slot]);   // group's scale factors consumed
          if (kb == n_kblocks - 1) umma_commit(&tmem_full[0]);
        }
        __syncwarp();
        if ((kb & 3) == 3 || kb == n_kblocks - 1) ++gcount;
        if (++stage == kXStages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 3) {
    // ===================== scale-factor loader =====================
    // per k-group: 128 token words (one coalesced 512-byte read per 32-row block) and the two weight
    // block words of this tile, written in the order tcgen05.cp.32x128b.warpx4 reads them
    pdl_wait();      // x_sf is the direct predecessor's output
    // The words of group g + 1 are fetched into registers BEFORE group g is published, so the global-load
    // latency of a group hides behind the previous group's slot wait (with the loads issued only after the
    // publish, the MMA warp waited on sf_full 4-26 % of its time, profiles/r2k).
    const int wblocks = w_rows_per_expert / 128;
    struct SfWords { uint32_t a[4], b0, b1; };
    auto fetch = [&](const WorkRef& tile, int g) {
      const uint32_t* wb0 = w_sf + (static_cast<size_t>(tile.expert) * wblocks + (GATED ? tile.f : 2 * tile.f)) * n_kgroups;
      const uint32_t* wb1 = GATED ? wb0 + static_cast<size_t>(up_row_offset / 128) * n_kgroups : wb0 + n_kgroups;
      const uint32_t* xs = x_sf + static_cast<size_t>(g) * scale_ld;
      SfWords r;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int row = min(tile.row_start + i * 32 + lane, scale_ld - 1);   // rows past the tile: any in-bounds word
        r.a[i] = __ldcg(xs + row);
      }
      r.b0 = __ldg(wb0 + g);
      r.b1 = __ldg(wb1 + g);
      return r;
    };
    uint32_t gcount = 0;
    int it = 0;
    int w = next_work(it);
    if (w >= 0) {
      WorkRef tile = get_work<kXM>(tiles, w, n_feat_tiles, n_tiles, 0);
      int g = 0;
      SfWords cur = fetch(tile, g);
      for (;;) {
        bool more = true;
        WorkRef ntile = tile;
        int ng = g + 1;
        if (ng == n_kgroups) {
          w = next_work(++it);
          more = w >= 0;
          ng = 0;
          if (more) ntile = get_work<kXM>(tiles, w, n_feat_tiles, n_tiles, 0);
        }
        SfWords nxt = cur;
        if (more) nxt = fetch(ntile, ng);
        const int slot = gcount & 1u;
        mbar_wait(&sf_empty[slot], ((gcount >> 1) & 1u) ^ 1u);
        uint32_t* dst = sf_smem + slot * (kXM + kXN);
        *reinterpret_cast<uint4*>(dst + lane * 4) = make_uint4(cur.a[0], cur.a[1], cur.a[2], cur.a[3]);
        *reinterpret_cast<uint4*>(dst + 128 + lane * 4) = make_uint4(cur.b0, cur.b0, cur.b0, cur.b0);
        *reinterpret_cast<uint4*>(dst + 256 + lane * 4) = make_uint4(cur.b1, cur.b1, cur.b1, cur.b1);
        fence_proxy_async();      // generic-proxy stores -> async-proxy (tcgen05.cp) reads
        __syncwarp();
        if (lane == 0) mbar_arrive(&sf_full[slot]);
        ++gcount;
        if (!more) break;
        cur = nxt;
        tile = ntile;
        g = ng;
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue: one read of the accumulator per tile, fused output math =====================
    // warp -> TMEM lane quadrant q (hardware: warp id % 4) and column group cg:
    //   gate_up: gate columns [32cg, 32cg+32) and up columns [128+32cg, +32)  (32 outputs, 64 accumulators)
    //   down:    columns [64cg, 64cg+64)                                      (64 outputs)
    const int q = warp & 3;
    const int cg = (warp - 4) >> 2;
    const uint32_t lane_base = static_cast<uint32_t>(q * 32) << 16;
    const int r = q * 32 + lane;                    // token row inside the tile held by this thread
    pdl_wait();
    for (int it = 0;; ++it) {
      const int w = next_work(it);
      if (w < 0) break;
      const WorkRef tile = get_work<kXM>(tiles, w, n_feat_tiles, n_tiles, 0);
      const bool row_ok = r < tile.nrows;
      const int row = tile.row_start + r;
      const int col0 = GATED ? 32 * cg : 64 * cg;
      const int col1 = GATED ? 128 + 32 * cg : 64 * cg + 32;
      float rw = 1.0f;
      if (!GATED && row_ok && row < n_routed_rows) rw = __ldg(row_weight + row);
      mbar_wait(&tmem_full[0], it & 1u);
      tc_fence_after();
      const uint32_t t_acc = tmem_base + lane_base;
      uint32_t a0[16], a1[16], b0[16], b1[16];
      tmem_ld16(t_acc + col0, a0);
      tmem_ld16(t_acc + col0 + 16, a1);
      tmem_ld16(t_acc + col1, b0);
      tmem_ld16(t_acc + col1 + 16, b1);
      tmem_ld_wait();
      // everything is in registers: hand the accumulator back before the output math
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[0]);
      if (GATED) {
        // a = gate, b = up of output features f*128 + 32cg + j.  SiLU*mul (activation.py:13-14), one
        // rounding to the model dtype (the reference stores the intermediate in it), then the 1x128 e4m3
        // quantiser of the down projection's input with a power-of-two scale: the group = this tile's 128
        // features of the token = the 4 column groups of the quadrant.
        float v[32];
        float amax = 0.f;
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          v[j] = Dt<T>::round(silu_fast(__uint_as_float(a0[j])) * __uint_as_float(b0[j]));
          v[16 + j] = Dt<T>::round(silu_fast(__uint_as_float(a1[j])) * __uint_as_float(b1[j]));
          amax = fmaxf(amax, fmaxf(fabsf(v[j]), fabsf(v[16 + j])));
        }
        float* ax = amax_x + (it & 1) * 4 * kXM;
        ax[cg * kXM + r] = amax;
        xepi_bar_sync();
        amax = fmaxf(fmaxf(ax[r], ax[kXM + r]), fmaxf(ax[2 * kXM + r], ax[3 * kXM + r]));
        uint32_t ebyte;
        const float sc = pow2_scale(amax, &ebyte);
        const float inv = pow2_inverse(ebyte);
        if (row_ok) {
          uint32_t p[8];
#pragma unroll
          for (int j = 0; j < 8; ++j)
            p[j] = xpack_e4m3x4(v[4 * j] * inv, v[4 * j + 1] * inv, v[4 * j + 2] * inv, v[4 * j + 3] * inv);
          uint4* dst = reinterpret_cast<uint4*>(out_q + static_cast<size_t>(row) * ld_out + tile.f * 128 + 32 * cg);
          dst[0] = make_uint4(p[0], p[1], p[2], p[3]);
          dst[1] = make_uint4(p[4], p[5], p[6], p[7]);
          if (cg == 0) {
            out_scale_t[static_cast<size_t>(tile.f) * scale_ld + row] = sc;
            out_sf[(static_cast<size_t>(tile.f >> 2) * scale_ld + row) * 4 + (tile.f & 3)] = static_cast<uint8_t>(ebyte);
          }
        }
      } else {
        // y = acc * routing weight (mlp.py:52, moe.py:149); shared-expert rows carry weight 1
        if (row_ok) {
          uint4* dst = reinterpret_cast<uint4*>(out + static_cast<size_t>(row) * ld_out + tile.f * 256 + 64 * cg);
          auto f = [&](const uint32_t* s, int j) { return __uint_as_float(s[j]) * rw; };
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const uint32_t* s = h ? a1 : a0;
            dst[2 * h] = make_uint4(xpack2<T>(f(s, 0), f(s, 1)), xpack2<T>(f(s, 2), f(s, 3)), xpack2<T>(f(s, 4), f(s, 5)),
                                    xpack2<T>(f(s, 6), f(s, 7)));
            dst[2 * h + 1] = make_uint4(xpack2<T>(f(s, 8), f(s, 9)), xpack2<T>(f(s, 10), f(s, 11)),
                                        xpack2<T>(f(s, 12), f(s, 13)), xpack2<T>(f(s, 14), f(s, 15)));
          }
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const uint32_t* s = h ? b1 : b0;
            dst[4 + 2 * h] = make_uint4(xpack2<T>(f(s, 0), f(s, 1)), xpack2<T>(f(s, 2), f(s, 3)),
                                        xpack2<T>(f(s, 4), f(s, 5)), xpack2<T>(f(s, 6), f(s, 7)));
            dst[4 + 2 * h + 1] = make_uint4(xpack2<T>(f(s, 8), f(s, 9)), xpack2<T>(f(s, 10), f(s, 11)),
                                            xpack2<T>(f(s, 12), f(s, 13)), xpack2<T>(f(s, 14), f(s, 15)));
          }
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
static int launch_mx_tm_one(const UmmaFp8MxLaunch& a, cudaStream_t st) {
  auto kern = umma_fp8_mx_tm_kernel<T, GATED>;
  static unsigned long long attr_done = 0;   // per instantiation, one bit per device
  GLM4MOE_CUDA_CHECK(ensure_dynamic_smem(kern, kXSmemBytes, &attr_done));
  GLM4MOE_CUDA_CHECK(launch_pdl(kern, dim3(a.grid), dim3(kXThreads), kXSmemBytes, st,
                                *static_cast<const CUtensorMap*>(a.tm_w), *static_cast<const CUtensorMap*>(a.tm_x),
                                a.w_sf, a.x_sf, a.scale_ld, reinterpret_cast<const int4*>(a.tiles),
                                a.num_tiles, static_cast<T*>(a.out), a.ld_out, static_cast<uint8_t*>(a.out_q),
                                a.out_scale_t, static_cast<uint8_t*>(a.out_sf), a.row_weight, a.n_routed_rows,
                                a.n_feat_tiles, a.n_kblocks, a.w_rows_per_expert, a.up_row_offset, a.sched));
  return GLM4MOE_OK;
}

int umma_fp8_mx_tm_launch(const UmmaFp8MxLaunch& a, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (!a.w_sf || !a.x_sf) return GLM4MOE_ERR_ARG;
  if (a.gated ? (!a.out_q || !a.out_scale_t || !a.out_sf) : !a.out) return GLM4MOE_ERR_ARG;
  if (a.dtype == GLM4MOE_BF16)
    return a.gated ? launch_mx_tm_one<__nv_bfloat16, true>(a, st) : launch_mx_tm_one<__nv_bfloat16, false>(a, st);
  if (a.dtype == GLM4MOE_F16)
    return a.gated ? launch_mx_tm_one<__half, true>(a, st) : launch_mx_tm_one<__half, false>(a, st);
  return GLM4MOE_ERR_UNSUPPORTED;
}

}  // namespace glm4moe

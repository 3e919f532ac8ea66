// FP8 (e4m3) grouped expert GEMMs for the prefill regime with HARDWARE block scaling, sm_100a.
//
// The 128x128 weight block scales and the per-token 1x128 activation scales are applied by the tensor
// pipe itself: `tcgen05.mma.kind::mxf8f6f4.block_scale` multiplies every K = 32 slice by one UE8M0
// (power-of-two) factor per A row and one per B row, read from tensor memory.  The whole K loop therefore
// accumulates into ONE TMEM accumulator and the epilogue reads it once per tile, instead of pulling every
// k-block's fp32 partial through `tcgen05.ld` to scale it in registers (the TMEM-read bound the
// software-scaled kernel umma_fp8_prefill.cu sits on).
//
// Tile: TWO 128-row weight slabs (MMA M = 128 each: gate rows + up rows of 128 features, or 256 consecutive
// down features) x one token tile of <= 128 rows (MMA N), one 128-byte k-block per pipeline stage (48 KB).
// Tokens are the N dimension so that a ragged last tile of an expert costs tensor time in proportion to its
// rows (N = rows rounded up to 16): with tokens as M a 14-row tail cost a full 128-row tile, 20 % of all MMA
// work at T = 4096 where an expert holds 256 +- 16 rows (profiles/r2f: pipe 69 % active, 55 % useful).
// Accumulators: slab 0 -> TMEM columns [0, 128), slab 1 -> [128, 256); lane = output feature, column = token.
//
// Requires power-of-two scales: activation scales always are (common.cuh `pow2_scale`); weight block
// scales when the checkpoint / `quantize_fp8(pow2_scales=True)` made them so
// (GLM4MOE_W_FP8_BLOCK128_POW2).  Scale operands, per 4 consecutive k-blocks (K = 512), one uint32 per
// row holding their 4 exponent bytes (byte j = k-block 4g + j; the MMA's a_sf_id / b_sf_id pick the byte):
//   SFB  x_sf [ceil(nkb/4), scale_ld]   written by the activation quantisers (token row minor)
//   SFA  w_sf [n_stack, rows/128, ceil(nkb/4)]   packed once per layer from the fp32 block scales;
//        one value per 128-row weight slab, replicated over its rows here
// staged per k-group by warp 3 into shared memory in the layout `tcgen05.cp.32x128b.warpx4` expects
// (row r of a 128-row block at word (r % 32) * 4 + r / 32), copied to TMEM by the MMA warp right in
// front of the group's MMAs (tcgen05.cp and tcgen05.mma execute in issue order).
//
// TMEM: 256 accumulator columns + 2 x (4 SFB + 2 x 4 SFA) columns.  Two accumulator pairs would fill all
// 512 columns, so there is ONE accumulator stage; the ring keeps filling while the epilogue drains it.
// Replaces mlp.py:50-52 for the quant tier at prefill (BASELINE.json configs[2]).
#include <cuda_fp8.h>

#include "umma_common.cuh"
#include "umma_gemm.cuh"

namespace glm4moe {

namespace {

constexpr int kXK = 128;                         // e4m3 elements (= bytes) per k-block = one scale block
constexpr int kXM = 128;                         // token rows per tile (MMA N, <= 128)
constexpr int kXN = 256;                         // weight rows per tile: two slabs of MMA M = 128
constexpr int kXStages = 4;
constexpr int kXABytes = kXM * kXK;              // token tile, 16 KB
constexpr int kXBBytes = kXN * kXK;              // two weight slabs, 32 KB
constexpr int kXStageBytes = kXABytes + kXBBytes;
constexpr int kXEpiWarps = 16;
constexpr int kXEpiThreads = 32 * kXEpiWarps;
constexpr int kXThreads = 128 + kXEpiThreads;    // 640
constexpr int kXSfSlotBytes = (kXM + kXN) * 4;   // 128 token words + 2 x 128 weight-slab words
constexpr int kXAmaxBytes = 2 * 4 * kXM * 4;     // [tile parity][column group][row] fp32
constexpr int kXBarBytes = 512;
constexpr int kXSmemBytes = kXStages * kXStageBytes + 1024 /*align*/ + kXBarBytes + 2 * kXSfSlotBytes + kXAmaxBytes;
constexpr int kXSchedDepth = 4;
constexpr uint32_t kXSfCol0 = 256;               // first scale-factor column in TMEM
constexpr uint32_t kXSfColStride = 16;           // per slot: tokens (SFB) at +0, slab 0 (SFA) at +4, slab 1 at +8

__device__ __forceinline__ void xepi_bar_sync() {   // the 16 epilogue warps only
  asm volatile("bar.sync 1, %0;" ::"n"(kXEpiThreads) : "memory");
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
// M = 128, N = n (multiple of 16), scale-factor byte ids of this K = 32 slice's k-block
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
__global__ void __launch_bounds__(kXThreads, 1) umma_fp8_mx_kernel(
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
    // ===================== MMA issuer: scale factors to TMEM, then 2 x 4 block-scaled MMAs per k-block =====================
    int stage = 0;
    uint32_t phase = 0;
    uint32_t gcount = 0;      // k-groups consumed so far (scale-factor slot ring)
    for (int it = 0;; ++it) {
      const int w = next_work(it);
      if (w < 0) break;
      const WorkRef tile = get_work<kXM>(tiles, w, n_feat_tiles, n_tiles, 0);
      const int n_eff = min(kXM, (tile.nrows + 15) & ~15);     // MMA N: the tile's rows, rounded up to 16
      mbar_wait(&tmem_empty[0], (it & 1u) ^ 1u);
      tc_fence_after();
      for (int kb = 0; kb < n_kblocks; ++kb) {
        const int slot = gcount & 1u;
        const uint32_t sfb_t = tmem_base + kXSfCol0 + slot * kXSfColStride;   // tokens
        const uint32_t sfa_t = sfb_t + 4;                                     // slab 0; slab 1 at +4
        if ((kb & 3) == 0) {
          mbar_wait(&sf_full[slot], (gcount >> 1) & 1u);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t ss = smem_u32(sf_smem + slot * (kXM + kXN));
            utccp_32x128b_x4(sfb_t, make_sf_desc(ss));                 // 128 token rows      -> 4 columns
            utccp_32x128b_x4(sfa_t, make_sf_desc(ss + 512));           // weight slab 0 rows  -> 4 columns
            utccp_32x128b_x4(sfa_t + 4, make_sf_desc(ss + 1024));      // weight slab 1 rows  -> 4 columns
          }
          __syncwarp();
        }
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(smem + stage * kXStageBytes);
          const uint64_t b_desc = make_kmajor_desc(sa);                           // tokens  [128 x 128 B]
          const uint64_t a0_desc = make_kmajor_desc(sa + kXABytes);               // slab 0  [128 x 128 B]
          const uint64_t a1_desc = make_kmajor_desc(sa + kXABytes + kXBBytes / 2);
          const uint32_t idesc = make_idesc_mx(n_eff, kb & 3, kb & 3);
#pragma unroll
          for (int ks = 0; ks < kXK / 32; ++ks) {
            const uint32_t acc = (kb > 0 || ks > 0) ? 1u : 0u;
            umma_f8_block_scaled(tmem_base, a0_desc + 2 * ks, b_desc + 2 * ks, idesc, sfa_t, sfb_t, acc);
            umma_f8_block_scaled(tmem_base + kXM, a1_desc + 2 * ks, b_desc + 2 * ks, idesc, sfa_t + 4, sfb_t, acc);
          }
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
    // slab words of this tile, written in the order tcgen05.cp.32x128b.warpx4 reads them
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
    // ===================== epilogue: one read of the accumulators per tile, fused output math =====================
    // warp -> TMEM lane quadrant q (hardware: warp id % 4) = output features [32q, 32q + 32) of both slabs, and
    // token group cg = token columns [32cg, 32cg + 32); a thread holds, for its feature, 32 tokens x 2 slabs.
    const int q = warp & 3;
    const int cg = (warp - 4) >> 2;
    const uint32_t lane_base = static_cast<uint32_t>(q * 32) << 16;
    const int feat = q * 32 + lane;                 // feature inside a slab
    pdl_wait();
    for (int it = 0;; ++it) {
      const int w = next_work(it);
      if (w < 0) break;
      const WorkRef tile = get_work<kXM>(tiles, w, n_feat_tiles, n_tiles, 0);
      const int n_eff = min(kXM, (tile.nrows + 15) & ~15);
      const bool have = 32 * cg < n_eff;            // this warp's token columns were computed (warp-uniform)
      const int n_here = min(32, tile.nrows - 32 * cg);   // valid tokens among them (may be <= 0)
      mbar_wait(&tmem_full[0], it & 1u);
      tc_fence_after();
      const uint32_t t_acc = tmem_base + lane_base + 32 * cg;
      uint32_t a0[16], a1[16], b0[16], b1[16];      // slab 0 (gate) / slab 1 (up), tokens 0-15 / 16-31 of the group
      if (have) {
        tmem_ld16(t_acc, a0);
        tmem_ld16(t_acc + kXM, b0);
        if (32 * cg + 16 < n_eff) {
          tmem_ld16(t_acc + 16, a1);
          tmem_ld16(t_acc + kXM + 16, b1);
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) a1[j] = b1[j] = 0u;
        }
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) a0[j] = a1[j] = b0[j] = b1[j] = 0u;
      }
      // everything is in registers: hand the accumulators back before the output math
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[0]);
      if (GATED) {
        // a = gate, b = up of feature f*128 + feat for 32 tokens.  SiLU*mul (activation.py:13-14), one rounding
        // to the model dtype (the reference stores the intermediate in it), then the 1x128 e4m3 quantiser of the
        // down projection's input with a power-of-two scale: the group = this tile's 128 features of a token =
        // the 32 lanes of the 4 quadrant warps.
        float* ax = amax_x + (it & 1) * 4 * kXM;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t* g_ = h ? a1 : a0;
          const uint32_t* u_ = h ? b1 : b0;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float v = Dt<T>::round(silu_fast(__uint_as_float(g_[j])) * __uint_as_float(u_[j]));
            g_[j] = __float_as_uint(v);
            // |v| as an unsigned integer orders like the float: one redux instead of 5 shuffles
            const uint32_t m = __reduce_max_sync(0xffffffffu, __float_as_uint(fabsf(v)));
            if (lane == 16 * h + j) ax[q * kXM + 32 * cg + lane] = __uint_as_float(m);
          }
        }
        xepi_bar_sync();
        const int tok = 32 * cg + lane;             // the token this lane publishes the scale of
        const float am = fmaxf(fmaxf(ax[tok], ax[kXM + tok]), fmaxf(ax[2 * kXM + tok], ax[3 * kXM + tok]));
        uint32_t ebyte;
        const float sc = pow2_scale(am, &ebyte);
        if (q == 0 && lane < n_here) {
          const int row = tile.row_start + tok;
          out_scale_t[static_cast<size_t>(tile.f) * scale_ld + row] = sc;
          out_sf[(static_cast<size_t>(tile.f >> 2) * scale_ld + row) * 4 + (tile.f & 3)] = static_cast<uint8_t>(ebyte);
        }
        uint8_t* dst = out_q + static_cast<size_t>(tile.row_start + 32 * cg) * ld_out + tile.f * 128 + feat;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const uint32_t* g_ = h ? a1 : a0;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int t = 16 * h + j;
            const float inv = pow2_inverse(__shfl_sync(0xffffffffu, ebyte, t));
            if (t < n_here)
              dst[static_cast<size_t>(t) * ld_out] = static_cast<uint8_t>(
                  __nv_cvt_float_to_fp8(__uint_as_float(g_[j]) * inv, __NV_SATFINITE, __NV_E4M3));
          }
        }
      } else {
        // y = acc * routing weight (mlp.py:52, moe.py:149); shared-expert rows carry weight 1
        const int row_l = tile.row_start + 32 * cg + lane;
        float rw_l = 1.0f;
        if (lane < n_here && row_l < n_routed_rows) rw_l = __ldg(row_weight + row_l);
        T* dst = out + static_cast<size_t>(tile.row_start + 32 * cg) * ld_out + tile.f * 256 + feat;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const uint32_t* s0 = h ? a1 : a0;
          const uint32_t* s1 = h ? b1 : b0;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int t = 16 * h + j;
            const float rw = __shfl_sync(0xffffffffu, rw_l, t);
            if (t < n_here) {
              dst[static_cast<size_t>(t) * ld_out] = Dt<T>::from_f(__uint_as_float(s0[j]) * rw);
              dst[static_cast<size_t>(t) * ld_out + 128] = Dt<T>::from_f(__uint_as_float(s1[j]) * rw);
            }
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<512>(tmem_base);
}

// fp32 block scales -> packed exponent bytes (layer creation).  One thread per output word.
__global__ void pack_weight_sf_kernel(const float* __restrict__ s, uint32_t* __restrict__ out, size_t n_blocks,
                                      int nkb, int n_groups, int* __restrict__ bad) {
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n_blocks * n_groups) return;
  const size_t b = i / n_groups;
  const int g = static_cast<int>(i - b * n_groups);
  uint32_t word = 0;
  for (int j = 0; j < 4; ++j) {
    const int kb = 4 * g + j;
    uint32_t e = 127;                                 // K tail of the last word: 1.0, never selected
    if (kb < nkb) {
      const uint32_t bits = __float_as_uint(s[b * nkb + kb]);
      e = bits >> 23;                                 // sign bit set -> e > 255 below
      if ((bits & 0x7fffffu) != 0u || e == 0u || e >= 255u) atomicExch(bad, 1);
    }
    word |= (e & 0xffu) << (8 * j);
  }
  out[i] = word;
}

int pack_weight_sf(const float* scales, uint32_t* out, size_t n_blocks, int nkb, int* bad, void* stream) {
  const int n_groups = (nkb + 3) / 4;
  const size_t n = n_blocks * n_groups;
  if (n == 0) return GLM4MOE_OK;
  pack_weight_sf_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      scales, out, n_blocks, nkb, n_groups, bad);
  GLM4MOE_CUDA_CHECK(cudaGetLastError());
  return GLM4MOE_OK;
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
template <typename T, bool GATED>
static int launch_mx_one(const UmmaFp8MxLaunch& a, cudaStream_t st) {
  auto kern = umma_fp8_mx_kernel<T, GATED>;
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

int umma_fp8_mx_launch(const UmmaFp8MxLaunch& a, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (!a.w_sf || !a.x_sf) return GLM4MOE_ERR_ARG;
  if (a.gated ? (!a.out_q || !a.out_scale_t || !a.out_sf) : !a.out) return GLM4MOE_ERR_ARG;
  if (a.dtype == GLM4MOE_BF16)
    return a.gated ? launch_mx_one<__nv_bfloat16, true>(a, st) : launch_mx_one<__nv_bfloat16, false>(a, st);
  if (a.dtype == GLM4MOE_F16)
    return a.gated ? launch_mx_one<__half, true>(a, st) : launch_mx_one<__half, false>(a, st);
  return GLM4MOE_ERR_UNSUPPORTED;
}

}  // namespace glm4moe

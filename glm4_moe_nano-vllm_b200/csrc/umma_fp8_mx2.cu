// FP8 (e4m3) grouped expert GEMMs for the prefill regime: hardware block scaling on a CTA PAIR
// (`tcgen05.mma.cta_group::2.kind::mxf8f6f4.block_scale`), sm_100a.
//
// Why a pair: with both operands in shared memory one CTA reads AND fills (M + N) x 128 B per k-block, and
// the tensor pipe's shared-memory read path saturates first (ncu r2f/r2i: `l1tex__data_pipe_tc_wavefronts_
// mem_shared` 51-54 % of its nominal peak on every single-CTA tile tried while the tensor pipe is 48-70 %
// active).  A pair computes a 256-token x 256-weight-row tile with each CTA holding its own 128 token rows
// and HALF of the weight rows: 32 KB instead of 48 KB filled and read per CTA per k-block for the same MMA
// cycles.
//
// Tile of one pair: tokens = MMA M = 256 (two 128-row tiles of ONE expert, CTA r holds rows [128r, 128r+128)
// and their accumulator rows in its own TMEM), weight rows = MMA N = 256 (gate 128 + up 128 of one feature
// tile, or 256 down features; CTA 0 loads the first 128 rows, CTA 1 the second), one 128-byte k-block per
// stage.  Otherwise the single-CTA token-major kernel of the first block-scaled version: a thread owns one
// token row, SiLU*mul + power-of-two re-quantisation (gate_up) or the routing weight (down) in the epilogue.
//
// Cross-CTA protocol (leader = cluster rank 0 issues every tcgen05.cp / tcgen05.mma / tcgen05.commit):
//   full[s]      leader's barrier: expect_tx = both CTAs' 32 KB; the peer's TMA completes its bytes on the
//                leader's barrier (cp.async.bulk.tensor ... cta_group::2)
//   sf_full[j]   leader's barrier, 2 arrivals: each CTA's scale-factor warp (the peer's arrives remotely)
//   empty[s], sf_empty[j], tmem_full   in BOTH CTAs: tcgen05.commit ... multicast::cluster
//   tmem_empty   leader's barrier, 2 x 16 arrivals: the epilogue warps of both CTAs
// Work is dealt statically (item = cluster id + i * clusters, expert-major order so concurrent pairs share
// weight slabs in L2): both CTAs derive the same sequence without talking to each other.
// Every barrier wait is bounded (trap with a code in the workspace instead of hanging the GPU).
//
// Scale operands as in umma_fp8_mx.cu: SFA = the CTA's own 128 token rows (4 TMEM columns), SFB = the 256
// weight rows (8 columns, one value per 128-row slab; both CTAs stage all of it).
#include <cuda_fp8.h>

#include <cstdio>

#include "umma_common.cuh"
#include "umma_gemm.cuh"

namespace glm4moe {

namespace {

constexpr int kYK = 128;                         // e4m3 elements (= bytes) per k-block = one scale block
constexpr int kYM = 128;                         // token rows per tile (MMA M)
constexpr int kYN = 256;                         // weight rows per tile (MMA N)
constexpr int kYStages = 6;
constexpr int kYABytes = kYM * kYK;              // own token rows, 16 KB
constexpr int kYBBytes = kYN / 2 * kYK;          // this CTA's half of the weight rows, 16 KB
constexpr int kYStageBytes = kYABytes + kYBBytes;
constexpr int kYEpiWarps = 16;
constexpr int kYEpiThreads = 32 * kYEpiWarps;
constexpr int kYThreads = 128 + kYEpiThreads;    // 640
constexpr int kYSfSlotBytes = (kYM + kYN) * 4;   // 128 SFA words + 256 SFB words
constexpr int kYAmaxBytes = 2 * 4 * kYM * 4;     // [tile parity][column group][row] fp32
constexpr int kYBarBytes = 512;
constexpr int kYSmemBytes = kYStages * kYStageBytes + 1024 /*align*/ + kYBarBytes + 2 * kYSfSlotBytes + kYAmaxBytes;
constexpr uint32_t kYSfCol0 = 256;               // first scale-factor column in TMEM
constexpr uint32_t kYSfColStride = 16;           // per slot: SFA at +0 (4 columns), SFB at +4 (8 columns)

__device__ __forceinline__ void xepi_bar_sync() {   // the 16 epilogue warps only
  asm volatile("bar.sync 1, %0;" ::"n"(kYEpiThreads) : "memory");
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

// ---- cluster / pair primitives ----
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// address of the same shared-memory object in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t saddr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load into OWN shared memory whose bytes complete on the LEADER CTA's barrier at the same offset
// (address with the peer bit cleared, the CUTLASS SM100_TMA_2SM_LOAD form)
__device__ __forceinline__ void tma_load_2d_pair(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1,
                                                 uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1), "l"(policy)
      : "memory");
}
// completion of all prior tcgen05 operations of the pair -> the barrier at this offset in BOTH CTAs
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)),
               "h"(static_cast<uint16_t>(3))
               : "memory");
}
// bounded wait: a protocol bug traps with a code instead of hanging the GPU
__device__ __forceinline__ void mbar_wait_b(uint64_t* bar, uint32_t parity, int code) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > (1ll << 31)) {      // ~1 s
      printf("umma_fp8_mx2: barrier wait %d timed out (block %d thread %d parity %u)\n", code, blockIdx.x, threadIdx.x, parity);
      __trap();
    }
  }
}

// D[tmem of both CTAs] (+)= (A * SFA) x (B * SFB): M = 256 over the pair, e4m3 operands, UE8M0 scale per row
// per K = 32, fp32 accumulate; issued by one thread of the leader CTA
__device__ __forceinline__ void umma_f8_block_scaled_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                                          uint32_t idesc, uint32_t sfa_tmem, uint32_t sfb_tmem,
                                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::mxf8f6f4.block_scale [%0], %1, %2, %3, [%5], [%6], p;\n\t"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate), "r"(sfa_tmem), "r"(sfb_tmem)
      : "memory");
}
// instruction descriptor, kind::mxf8f6f4.block_scale: A/B e4m3 (format 0) K-major, UE8M0 scales,
// M = 256 (pair), N = n, scale-factor byte ids of this K = 32 slice's k-block
__device__ __forceinline__ uint32_t make_idesc_mx_pair(int n, uint32_t a_sf_id, uint32_t b_sf_id) {
  return (b_sf_id << 4) | (static_cast<uint32_t>(n >> 3) << 17) | (1u << 23) | (static_cast<uint32_t>(256 >> 4) << 24) |
         (a_sf_id << 29);
}
// shared memory -> tensor memory in both CTAs (each from its own shared memory): 32 rows x 128 bits x 4 quadrants
__device__ __forceinline__ void utccp_32x128b_x4_pair(uint32_t tmem_dst, uint64_t smem_desc) {
  asm volatile("tcgen05.cp.cta_group::2.32x128b.warpx4 [%0], %1;" ::"r"(tmem_dst), "l"(smem_desc) : "memory");
}
// K-major, no swizzle, 8-row groups 128 bytes apart (one 16-byte atom along K)
__device__ __forceinline__ uint64_t make_sf_desc(uint32_t saddr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr >> 4) & 0x3fffu);
  d |= static_cast<uint64_t>(128u >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  return d;
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* dst_smem) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(512));
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t addr) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(512));
}

}  // namespace

template <typename T, bool GATED>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kYThreads, 1) umma_fp8_mx2_kernel(
    const __grid_constant__ CUtensorMap tm_w,   // e4m3 weight stack [rows_w_total, K] bytes, box {128, 128}
    const __grid_constant__ CUtensorMap tm_x,   // e4m3 activation rows [rows, K], box {128, 128}
    const uint32_t* __restrict__ w_sf,          // [n_stack, w_rows_per_expert/128, n_kgroups] packed UE8M0
    const uint32_t* __restrict__ x_sf,          // [n_kgroups, scale_ld] packed UE8M0 (k-group major, row minor)
    int scale_ld,
    const int4* __restrict__ tiles,             // 256-row tiles (permute with tile_rows = 256)
    const int32_t* __restrict__ num_tiles_ptr,
    T* __restrict__ out, int ld_out,            // down: y[row, H]
    uint8_t* __restrict__ out_q,                // gate_up: e4m3 activation rows [row, I]
    float* __restrict__ out_scale_t,            // gate_up: their fp32 scales [I/128, scale_ld]
    uint8_t* __restrict__ out_sf,               // gate_up: the same as packed UE8M0 bytes [ceil(I/512), scale_ld] x 4
    const float* __restrict__ row_weight, int n_routed_rows,
    int n_feat_tiles,                           // gate_up: I/128, down: H/256
    int n_kblocks, int w_rows_per_expert, int up_row_offset) {
  extern __shared__ unsigned char smem_dyn[];
  unsigned char* smem = reinterpret_cast<unsigned char*>(
      (reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kYStages * kYStageBytes);
  uint64_t* full_bar = bars;                           // [stages]  leader's are the live ones
  uint64_t* empty_bar = bars + kYStages;               // [stages]  both CTAs (multicast commit)
  uint64_t* tmem_full = bars + 2 * kYStages;           // [1]       both CTAs
  uint64_t* tmem_empty = tmem_full + 1;                // [1]       leader's
  uint64_t* sf_full = tmem_empty + 1;                  // [2]       leader's
  uint64_t* sf_empty = sf_full + 2;                    // [2]       both CTAs
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(sf_empty + 2);
  uint32_t* sf_smem = reinterpret_cast<uint32_t*>(smem + kYStages * kYStageBytes + kYBarBytes);   // [2][128 + 256]
  float* amax_x = reinterpret_cast<float*>(smem + kYStages * kYStageBytes + kYBarBytes + 2 * kYSfSlotBytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int n_kgroups = (n_kblocks + 3) >> 2;
  const int cluster_id = static_cast<int>(blockIdx.x) >> 1;
  const int n_clusters = static_cast<int>(gridDim.x) >> 1;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tm_w);
    prefetch_tmap(&tm_x);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kYStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&tmem_full[0], 1);
    mbar_init(&tmem_empty[0], 2 * kYEpiWarps);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&sf_full[s], 2);
      mbar_init(&sf_empty[s], 1);
    }
    fence_barrier_init();
  }
  cluster_sync_all();                 // barriers of both CTAs exist before anything remote touches them
  if (warp == 2) tmem_alloc_pair(tmem_ptr_smem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  const int n_tiles = *num_tiles_ptr;
  const int total_work = n_tiles * n_feat_tiles;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (elect_one()) {
      pdl_wait();
      pdl_launch();
      int stage = 0;
      uint32_t phase = 0;
      for (int w = cluster_id; w < total_work; w += n_clusters) {
        const WorkRef wr = get_work<256>(tiles, w, n_feat_tiles, n_tiles, 0);
        // this CTA's half of the 256 weight rows: gate rows / up rows of feature tile f, or 128 down features
        const int w_row = wr.expert * w_rows_per_expert +
                          (GATED ? wr.f * 128 + static_cast<int>(rank) * up_row_offset : wr.f * 256 + static_cast<int>(rank) * 128);
        const int x_row = wr.row_start + static_cast<int>(rank) * kYM;
        for (int kb = 0; kb < n_kblocks; ++kb) {
          mbar_wait_b(&empty_bar[stage], phase ^ 1u, 1);
          unsigned char* sa = smem + stage * kYStageBytes;
          if (leader) mbar_expect_tx(&full_bar[stage], 2 * kYStageBytes);
          tma_load_2d_pair(&tm_x, &full_bar[stage], sa, kb * kYK, x_row, kEvictLast);
          tma_load_2d_pair(&tm_w, &full_bar[stage], sa + kYABytes, kb * kYK, w_row, kEvictFirst);
          if (++stage == kYStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader) {
      int stage = 0;
      uint32_t phase = 0;
      uint32_t gcount = 0;      // k-groups consumed so far (scale-factor slot ring)
      int it = 0;
      for (int w = cluster_id; w < total_work; w += n_clusters, ++it) {
        mbar_wait_b(&tmem_empty[0], (it & 1u) ^ 1u, 2);
        tc_fence_after();
        for (int kb = 0; kb < n_kblocks; ++kb) {
          const int slot = gcount & 1u;
          const uint32_t sfa_t = tmem_base + kYSfCol0 + slot * kYSfColStride;
          const uint32_t sfb_t = sfa_t + 4;
          if ((kb & 3) == 0) {
            mbar_wait_b(&sf_full[slot], (gcount >> 1) & 1u, 3);
            tc_fence_after();
            if (elect_one()) {
              const uint32_t ss = smem_u32(sf_smem + slot * (kYM + kYN));
              utccp_32x128b_x4_pair(sfa_t, make_sf_desc(ss));                 // each CTA's 128 token rows -> 4 columns
              utccp_32x128b_x4_pair(sfb_t, make_sf_desc(ss + 512));           // weight rows 0..127   -> 4 columns
              utccp_32x128b_x4_pair(sfb_t + 4, make_sf_desc(ss + 1024));      // weight rows 128..255 -> 4 columns
            }
            __syncwarp();
          }
          mbar_wait_b(&full_bar[stage], phase, 4);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t sa = smem_u32(smem + stage * kYStageBytes);
            const uint64_t a_desc = make_kmajor_desc(sa);              // tokens  [128 x 128 B] per CTA
            const uint64_t b_desc = make_kmajor_desc(sa + kYABytes);   // weights [128 x 128 B] per CTA = N / 2
            const uint32_t idesc = make_idesc_mx_pair(kYN, kb & 3, kb & 3);
#pragma unroll
            for (int ks = 0; ks < kYK / 32; ++ks)
              umma_f8_block_scaled_pair(tmem_base, a_desc + 2 * ks, b_desc + 2 * ks, idesc, sfa_t, sfb_t,
                                        (kb > 0 || ks > 0) ? 1u : 0u);
            umma_commit_pair(&empty_bar[stage]);
            if ((kb & 3) == 3 || kb == n_kblocks - 1) umma_commit_pair(&sf_empty[slot]);
            if (kb == n_kblocks - 1) umma_commit_pair(&tmem_full[0]);
          }
          __syncwarp();
          if ((kb & 3) == 3 || kb == n_kblocks - 1) ++gcount;
          if (++stage == kYStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 3) {
    // ===================== scale-factor loader (both CTAs; signals the leader) =====================
    pdl_wait();      // x_sf is the direct predecessor's output
    // group g + 1 is fetched into registers before group g is published (see umma_fp8_mx.cu)
    const int wblocks = w_rows_per_expert / 128;
    const uint32_t sf_full_leader0 = mapa_u32(smem_u32(&sf_full[0]), 0);
    struct SfWords { uint32_t a[4], b0, b1; };
    auto fetch = [&](const WorkRef& tile, int g) {
      const uint32_t* wb0 = w_sf + (static_cast<size_t>(tile.expert) * wblocks + (GATED ? tile.f : 2 * tile.f)) * n_kgroups;
      const uint32_t* wb1 = GATED ? wb0 + static_cast<size_t>(up_row_offset / 128) * n_kgroups : wb0 + n_kgroups;
      const uint32_t* xs = x_sf + static_cast<size_t>(g) * scale_ld;
      const int row0 = tile.row_start + static_cast<int>(rank) * kYM;
      SfWords r;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int row = min(row0 + i * 32 + lane, scale_ld - 1);   // rows past the tile: any in-bounds word
        r.a[i] = __ldcg(xs + row);
      }
      r.b0 = __ldg(wb0 + g);
      r.b1 = __ldg(wb1 + g);
      return r;
    };
    uint32_t gcount = 0;
    int w = cluster_id;
    if (w < total_work) {
      WorkRef tile = get_work<256>(tiles, w, n_feat_tiles, n_tiles, 0);
      int g = 0;
      SfWords cur = fetch(tile, g);
      for (;;) {
        bool more = true;
        WorkRef ntile = tile;
        int ng = g + 1;
        if (ng == n_kgroups) {
          w += n_clusters;
          more = w < total_work;
          ng = 0;
          if (more) ntile = get_work<256>(tiles, w, n_feat_tiles, n_tiles, 0);
        }
        SfWords nxt = cur;
        if (more) nxt = fetch(ntile, ng);
        const int slot = gcount & 1u;
        mbar_wait_b(&sf_empty[slot], ((gcount >> 1) & 1u) ^ 1u, 5);
        uint32_t* dst = sf_smem + slot * (kYM + kYN);
        *reinterpret_cast<uint4*>(dst + lane * 4) = make_uint4(cur.a[0], cur.a[1], cur.a[2], cur.a[3]);
        *reinterpret_cast<uint4*>(dst + 128 + lane * 4) = make_uint4(cur.b0, cur.b0, cur.b0, cur.b0);
        *reinterpret_cast<uint4*>(dst + 256 + lane * 4) = make_uint4(cur.b1, cur.b1, cur.b1, cur.b1);
        fence_proxy_async();      // generic-proxy stores -> async-proxy (tcgen05.cp) reads
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(sf_full_leader0 + 8u * slot);
        ++gcount;
        if (!more) break;
        cur = nxt;
        tile = ntile;
        g = ng;
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue (both CTAs, own 128 token rows): one accumulator read per tile =====================
    // warp -> TMEM lane quadrant q (hardware: warp id % 4) and column group cg:
    //   gate_up: gate columns [32cg, 32cg+32) and up columns [128+32cg, +32)  (32 outputs, 64 accumulators)
    //   down:    columns [64cg, 64cg+64)                                      (64 outputs)
    const int q = warp & 3;
    const int cg = (warp - 4) >> 2;
    const uint32_t lane_base = static_cast<uint32_t>(q * 32) << 16;
    const int r = q * 32 + lane;                    // token row inside this CTA's half of the tile
    const uint32_t tmem_empty_leader = mapa_u32(smem_u32(&tmem_empty[0]), 0);
    pdl_wait();
    int it = 0;
    for (int w = cluster_id; w < total_work; w += n_clusters, ++it) {
      const WorkRef tile = get_work<256>(tiles, w, n_feat_tiles, n_tiles, 0);
      const int nrows = tile.nrows - static_cast<int>(rank) * kYM;     // valid rows of this CTA's half (may be <= 0)
      const bool row_ok = r < nrows;
      const int row = tile.row_start + static_cast<int>(rank) * kYM + r;
      const int col0 = GATED ? 32 * cg : 64 * cg;
      const int col1 = GATED ? 128 + 32 * cg : 64 * cg + 32;
      float rw = 1.0f;
      if (!GATED && row_ok && row < n_routed_rows) rw = __ldg(row_weight + row);
      mbar_wait_b(&tmem_full[0], it & 1u, 6);
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
      if (lane == 0) mbar_arrive_cluster(tmem_empty_leader);
      if (GATED) {
        float v[32];
        float amax = 0.f;
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          v[j] = Dt<T>::round(silu_fast(__uint_as_float(a0[j])) * __uint_as_float(b0[j]));
          v[16 + j] = Dt<T>::round(silu_fast(__uint_as_float(a1[j])) * __uint_as_float(b1[j]));
          amax = fmaxf(amax, fmaxf(fabsf(v[j]), fabsf(v[16 + j])));
        }
        float* ax = amax_x + (it & 1) * 4 * kYM;
        ax[cg * kYM + r] = amax;
        xepi_bar_sync();
        amax = fmaxf(fmaxf(ax[r], ax[kYM + r]), fmaxf(ax[2 * kYM + r], ax[3 * kYM + r]));
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

  // every remote arrive / multicast commit / pair TMA of this CTA has landed before either CTA leaves
  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) tmem_dealloc_pair(tmem_base);
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
template <typename T, bool GATED>
static int launch_mx2_one(const UmmaFp8MxLaunch& a, cudaStream_t st) {
  auto kern = umma_fp8_mx2_kernel<T, GATED>;
  static unsigned long long attr_done = 0;   // per instantiation, one bit per device
  GLM4MOE_CUDA_CHECK(ensure_dynamic_smem(kern, kYSmemBytes, &attr_done));
  const int grid = a.grid & ~1;              // whole pairs
  GLM4MOE_CUDA_CHECK(launch_pdl(kern, dim3(grid), dim3(kYThreads), kYSmemBytes, st,
                                *static_cast<const CUtensorMap*>(a.tm_w), *static_cast<const CUtensorMap*>(a.tm_x),
                                a.w_sf, a.x_sf, a.scale_ld, reinterpret_cast<const int4*>(a.tiles),
                                a.num_tiles, static_cast<T*>(a.out), a.ld_out, static_cast<uint8_t*>(a.out_q),
                                a.out_scale_t, static_cast<uint8_t*>(a.out_sf), a.row_weight, a.n_routed_rows,
                                a.n_feat_tiles, a.n_kblocks, a.w_rows_per_expert, a.up_row_offset));
  return GLM4MOE_OK;
}

// CTA-pair form of umma_fp8_mx_launch: `tiles` must be 256-row tiles
int umma_fp8_mx2_launch(const UmmaFp8MxLaunch& a, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (!a.w_sf || !a.x_sf || a.grid < 2) return GLM4MOE_ERR_ARG;
  if (a.gated ? (!a.out_q || !a.out_scale_t || !a.out_sf) : !a.out) return GLM4MOE_ERR_ARG;
  if (a.dtype == GLM4MOE_BF16)
    return a.gated ? launch_mx2_one<__nv_bfloat16, true>(a, st) : launch_mx2_one<__nv_bfloat16, false>(a, st);
  if (a.dtype == GLM4MOE_F16)
    return a.gated ? launch_mx2_one<__half, true>(a, st) : launch_mx2_one<__half, false>(a, st);
  return GLM4MOE_ERR_UNSUPPORTED;
}

}  // namespace glm4moe

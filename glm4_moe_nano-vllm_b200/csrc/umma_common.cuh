// PTX wrappers shared by the tcgen05 grouped-GEMM kernels (umma_gemm.cu, umma_fp8.cu): mbarrier,
// TMA, TMEM allocation, tcgen05.mma / commit / ld, shared-memory and instruction descriptors.
#pragma once
#include <cuda.h>

#include <type_traits>

#include "common.cuh"

namespace glm4moe {

// ------------------------------------------------------------------------------------------
// PTX wrappers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint64_t* bar, void* dst, int c0,
                                            int c1, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
template <int kCols>
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(dst_smem)),
               "n"(kCols));
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
}
template <int kCols>
__device__ __forceinline__ void tmem_dealloc(uint32_t addr) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "n"(kCols));
}
// D[tmem] (+)= A[smem desc] * B[smem desc], kind::f16 (bf16 or fp16 inputs, fp32 accumulate)
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                         uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier once every previously issued tcgen05.mma of this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t* r) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr));
}
template <int N>
__device__ __forceinline__ void tmem_ldn(uint32_t taddr, uint32_t* r) {
  static_assert(N == 8 || N == 16, "tmem_ldn: 8 or 16 columns");
  if constexpr (N == 8) tmem_ld8(taddr, r); else tmem_ld16(taddr, r);
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t"
      ".reg .pred P;\n\t"
      "elect.sync _|P, 0xffffffff;\n\t"
      "selp.b32 %0, 1, 0, P;\n\t"
      "}\n"
      : "=r"(pred));
  return pred != 0;
}

// K-major operand tile, rows of 64 two-byte elements = one 128-byte swizzle atom per row:
// start address >> 4 | LBO (unused for a single K atom) | SBO = 8 rows * 128 B | version 1 |
// layout SWIZZLE_128B.  Advancing K by 16 elements inside the atom = +32 bytes on the address.
__device__ __forceinline__ uint64_t make_kmajor_desc(uint32_t saddr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr >> 4) & 0x3fffu);
  d |= static_cast<uint64_t>(1024u >> 4) << 32;   // stride byte offset between 8-row groups
  d |= static_cast<uint64_t>(1) << 46;            // descriptor version (sm_100)
  d |= static_cast<uint64_t>(2) << 61;            // SWIZZLE_128B
  return d;
}

// K-major operand tile whose rows are 32 two-byte elements = one 64-byte swizzle atom per row
// (SWIZZLE_64B: 8-row groups of 512 B).  Half-size pipeline stages: finer-grained ring, more bytes
// in flight for the same shared memory (umma_gemm.cu, BK = 32).
__device__ __forceinline__ uint64_t make_kmajor_desc_sw64(uint32_t saddr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr >> 4) & 0x3fffu);
  d |= static_cast<uint64_t>(512u >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(4) << 61;            // SWIZZLE_64B
  return d;
}

// instruction descriptor, kind::f16: D fp32, A/B both K-major, M = 128, N = n
__host__ __device__ constexpr uint32_t make_idesc(bool bf16, int n) {
  return (1u << 4)                         // c_format = F32
         | ((bf16 ? 1u : 0u) << 7)         // a_format
         | ((bf16 ? 1u : 0u) << 10)        // b_format
         | (static_cast<uint32_t>(n >> 3) << 17) | (static_cast<uint32_t>(128 >> 4) << 24);
}

constexpr uint64_t kEvictFirst = 0x12F0000000000000ull;   // weights: touched once
constexpr uint64_t kEvictLast = 0x14F0000000000000ull;    // activation rows: re-read per feature tile


// kind::f8f6f4 with e4m3 x e4m3 inputs (formats 0), fp32 accumulate; K = 32 per instruction
__device__ __forceinline__ void umma_f8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                        uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// instruction descriptor, kind::f8f6f4: D fp32, A/B e4m3 (format 0) K-major, M = 128, N = n
__host__ __device__ constexpr uint32_t make_idesc_f8(int n) {
  return (1u << 4) | (static_cast<uint32_t>(n >> 3) << 17) | (static_cast<uint32_t>(128 >> 4) << 24);
}

constexpr int kBM = 128;   // weight rows (output features) per tile

struct WorkRef { int expert, row_start, nrows, f; };

// work index -> (token tile, feature tile).  Tiles of one expert are contiguous in the list and
// carry {index inside the expert | tiles of the expert << 16} in .w; the work items of an expert
// are ordered feature-tile major, token-tile minor, so the CTAs that share a weight slab
// [128 x K] run at the same time and it crosses HBM once (the second reader hits L2).
// `tiles == nullptr`: dense rows [0, dense_rows) of a single weight, BN-row tiles, same order.
template <int BN>
__device__ __forceinline__ WorkRef get_work(const int4* __restrict__ tiles, int w, int n_feat_tiles,
                                            int n_tiles, int dense_rows) {
  if (!tiles) {
    const int f = w / n_tiles, ti = w - f * n_tiles;
    return {0, ti * BN, min(BN, dense_rows - ti * BN), f};
  }
  const int t0 = w / n_feat_tiles;
  int4 t = tiles[t0];
  int f = w - t0 * n_feat_tiles;
  const int nt = t.w >> 16;
  if (nt > 1) {
    const int first = t0 - (t.w & 0xffff);
    const int local = w - first * n_feat_tiles;
    f = local / nt;
    t = tiles[first + local - f * nt];
  }
  return {t.x, t.y, t.z, f};
}

}  // namespace glm4moe

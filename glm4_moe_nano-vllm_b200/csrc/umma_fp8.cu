// FP8 (e4m3) grouped expert GEMMs with 128x128 block-scaled weights and per-token 1x128 scaled
// activations (the quant tier of BASELINE.json configs[2]), on tcgen05 `kind::f8f6f4`.
//
// Same swap-AB / TMA / TMEM structure as umma_gemm.cu with one change forced by the fp32 block
// scales: a k-block is 128 fp8 elements = exactly one scale block, so every k-block's raw partial
// product is written to a fresh TMEM accumulator (2 alternate) and the epilogue warps promote it,
//     acc[feature, token] += partial * s_w[expert][feature/128][kb] * s_x[token][kb]
// in fp32 registers (lane = feature, one register per token column).  In the decode regime the
// k-block period is set by HBM (a 32 KB weight slab per ~1400 SM cycles) so the promotion is free.
//
// Also here: the activation quantiser (rows -> e4m3 + per-128 fp32 scales, scale = amax/448),
// fused with the row gather for the first GEMM.
#include <cuda_fp8.h>

#include "umma_common.cuh"
#include "umma_gemm.cuh"

namespace glm4moe {

constexpr int kBKF8 = 128;   // fp8 elements per k-block = 128-byte swizzle atom = one scale block

template <int BN, bool GATED>
struct Fp8Cfg {
  static constexpr int kABytes = kBM * kBKF8;
  static constexpr int kBBytes = BN * kBKF8;
  static constexpr int kStageBytes = (GATED ? 2 : 1) * kABytes + kBBytes;
  static constexpr int kStages = (212 * 1024) / kStageBytes > 12 ? 12 : (212 * 1024) / kStageBytes;
  static constexpr int kAccCols = (GATED ? 2 : 1) * BN;
  static constexpr int kTmemColsRaw = 2 * kAccCols;
  static constexpr int kTmemCols = kTmemColsRaw <= 32 ? 32 : kTmemColsRaw <= 64 ? 64
                                   : kTmemColsRaw <= 128 ? 128 : kTmemColsRaw <= 256 ? 256 : 512;
  static constexpr int kEpiWarps = BN >= 64 ? 8 : 4;
  static constexpr int kColsPerWarp = BN / (kEpiWarps / 4);     // token columns one epilogue warp owns
  static constexpr int kThreads = 128 + 32 * kEpiWarps;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 512;
  static_assert(kBBytes % 1024 == 0, "operand tiles must stay 1024-byte aligned");
  static_assert(kColsPerWarp % 16 == 0 && kColsPerWarp <= 32, "column split");
};

template <typename T, int BN, bool GATED>
__global__ void __launch_bounds__(Fp8Cfg<BN, GATED>::kThreads, 1) umma_fp8_kernel(
    const __grid_constant__ CUtensorMap tm_w,   // fp8 weights [rows_w_total, K] bytes, box {128, 128}
    const __grid_constant__ CUtensorMap tm_x,   // fp8 activation rows [rows, K], box {128, BN}
    const float* __restrict__ w_scale,          // [n_stack, w_rows_per_expert/128, K/128]
    const float* __restrict__ x_scale_t,        // [K/128, x_scale_ld]   (k-block major)
    int x_scale_ld,
    const int4* __restrict__ tiles, const int32_t* __restrict__ num_tiles_ptr,
    T* __restrict__ out, int ld_out, const float* __restrict__ row_weight, int n_routed_rows,
    int n_feat_tiles, int n_kblocks, int w_rows_per_expert, int up_row_offset,
    int32_t* __restrict__ sched) {
  using Cfg = Fp8Cfg<BN, GATED>;
  constexpr int EW = Cfg::kEpiWarps;
  constexpr int NC = Cfg::kColsPerWarp;
  extern __shared__ unsigned char smem_dyn[];
  unsigned char* smem = reinterpret_cast<unsigned char*>(
      (reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + Cfg::kStages;
  uint64_t* tmem_full = bars + 2 * Cfg::kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint64_t* sched_full = tmem_empty + 2;
  uint64_t* sched_empty = sched_full + 4;
  int32_t* sched_w = reinterpret_cast<int32_t*>(sched_empty + 4);
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(sched_w + 4);

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
      mbar_init(&tmem_empty[s], EW);
    }
    for (int s = 0; s < 4; ++s) {
      mbar_init(&sched_full[s], 1);
      mbar_init(&sched_empty[s], 1 + EW);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<Cfg::kTmemCols>(tmem_ptr_smem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  const int n_tiles = *num_tiles_ptr;
  const int total_work = n_tiles * n_feat_tiles;

  auto next_work = [&](int n) -> int {
    const int q = n & 3;
    mbar_wait(&sched_full[q], (n >> 2) & 1u);
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
        const int q = n_pushed & 3;
        mbar_wait(&sched_empty[q], ((n_pushed >> 2) & 1u) ^ 1u);
        sched_w[q] = w;
        mbar_arrive(&sched_full[q]);
        ++n_pushed;
      };
      auto issue = [&](int w, int kb, int stage, bool do_w, bool do_x) {
        const WorkRef tile = get_work<BN>(tiles, w, n_feat_tiles, n_tiles, 0);
        const int w_row = tile.expert * w_rows_per_expert + tile.f * kBM;
        unsigned char* sa = smem + stage * Cfg::kStageBytes;
        if (do_w) {
          mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes);
          tma_load_2d(&tm_w, &full_bar[stage], sa, kb * kBKF8, w_row, kEvictFirst);
          if (GATED)
            tma_load_2d(&tm_w, &full_bar[stage], sa + Cfg::kABytes, kb * kBKF8, w_row + up_row_offset,
                        kEvictFirst);
        }
        if (do_x)
          tma_load_2d(&tm_x, &full_bar[stage], sa + (GATED ? 2 : 1) * Cfg::kABytes, kb * kBKF8, tile.row_start,
                      kEvictLast);
      };
      int w = static_cast<int>(blockIdx.x) < total_work ? static_cast<int>(blockIdx.x) : -1;
      push_work(w);
      int kb = 0, n_pre = 0;
      if (w >= 0)
        for (; n_pre < Cfg::kStages && n_pre < n_kblocks; ++n_pre) issue(w, n_pre, n_pre, true, false);
      pdl_wait();
      pdl_launch();
      for (int i = 0; i < n_pre; ++i) issue(w, i, i, false, true);
      kb = n_pre;
      int stage = n_pre == Cfg::kStages ? 0 : n_pre;
      uint32_t phase = n_pre == Cfg::kStages ? 1u : 0u;
      while (w >= 0) {
        if (kb == n_kblocks) {
          kb = 0;
          const int nw = atomicAdd(&sched[0], 1) + static_cast<int>(gridDim.x);
          w = nw < total_work ? nw : -1;
          push_work(w);
          continue;
        }
        mbar_wait(&empty_bar[stage], phase ^ 1u);
        issue(w, kb, stage, true, true);
        ++kb;
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1u; }
      }
      __threadfence();
      if (atomicAdd(&sched[1], 1) == static_cast<int>(gridDim.x) - 1) {
        sched[0] = 0;
        sched[1] = 0;
        __threadfence();
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer: one fresh accumulator per k-block =====================
    int stage = 0;
    uint32_t phase = 0;
    uint32_t pcount = 0;           // k-blocks issued so far (partial accumulator parity)
    for (int it = 0;; ++it) {
      const int w = next_work(it);
      if (w < 0) break;
      const WorkRef tile = get_work<BN>(tiles, w, n_feat_tiles, n_tiles, 0);
      const int n_eff = min(BN, (tile.nrows + 15) & ~15);
      const uint32_t idesc = make_idesc_f8(0) | (static_cast<uint32_t>(n_eff >> 3) << 17);
      for (int kb = 0; kb < n_kblocks; ++kb, ++pcount) {
        const int ps = pcount & 1u;
        mbar_wait(&tmem_empty[ps], ((pcount >> 1) & 1u) ^ 1u);
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(smem + stage * Cfg::kStageBytes);
          const uint64_t a_desc = make_kmajor_desc(sa);
          const uint64_t u_desc = make_kmajor_desc(sa + Cfg::kABytes);
          const uint64_t b_desc = make_kmajor_desc(sa + (GATED ? 2 : 1) * Cfg::kABytes);
          const uint32_t d_tmem = tmem_base + ps * Cfg::kAccCols;
#pragma unroll
          for (int ks = 0; ks < kBKF8 / 32; ++ks) {
            umma_f8(d_tmem, a_desc + 2 * ks, b_desc + 2 * ks, idesc, ks > 0 ? 1u : 0u);
            if (GATED) umma_f8(d_tmem + BN, u_desc + 2 * ks, b_desc + 2 * ks, idesc, ks > 0 ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);
          umma_commit(&tmem_full[ps]);
        }
        __syncwarp();
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue: promote every k-block, then the fused math =====================
    const int ew = (warp - 4) & 3;
    const int col_base = ((warp - 4) >> 2) * NC;      // this warp's token columns [col_base, col_base + NC)
    const uint32_t lane_base = static_cast<uint32_t>(ew * 32) << 16;
    const int kb_total = n_kblocks;
    uint32_t pcount = 0;
    // the activation scales are written by the direct predecessor (the quantiser): unlike the
    // bf16 epilogue, these warps read predecessor data themselves and must pass the dependency wait
    pdl_wait();
    for (int it = 0;; ++it) {
      const int w = next_work(it);
      if (w < 0) break;
      const WorkRef tile = get_work<BN>(tiles, w, n_feat_tiles, n_tiles, 0);
      const int f = tile.f;
      const int nrows = tile.nrows;
      float acc_g[NC], acc_u[GATED ? NC : 1];
#pragma unroll
      for (int j = 0; j < NC; ++j) acc_g[j] = 0.f;
#pragma unroll
      for (int j = 0; j < (GATED ? NC : 1); ++j) acc_u[j] = 0.f;
      // weight scale rows of this tile: gate/down block row f, up block row (up_row_offset/128 + f)
      const float* sw_g = w_scale + (static_cast<size_t>(tile.expert) * (w_rows_per_expert / 128) + f) * kb_total;
      const float* sw_u = sw_g + static_cast<size_t>(up_row_offset / 128) * kb_total;
      // lane j (< NC) owns the activation scale of token column col_base + j
      const int my_col = col_base + (lane % NC);
      const bool my_col_ok = my_col < nrows;
      const float* sx = x_scale_t + tile.row_start + my_col;
      for (int kb = 0; kb < kb_total; ++kb, ++pcount) {
        const int ps = pcount & 1u;
        // scales first: they do not depend on the MMA
        const float swg = __ldg(sw_g + kb);
        const float swu = GATED ? __ldg(sw_u + kb) : 0.f;
        const float sx_l = my_col_ok ? __ldg(sx + static_cast<size_t>(kb) * x_scale_ld) : 0.f;
        mbar_wait(&tmem_full[ps], (pcount >> 1) & 1u);
        tc_fence_after();
        const uint32_t t_acc = tmem_base + lane_base + ps * Cfg::kAccCols + col_base;
        if (col_base < nrows) {
#pragma unroll
          for (int c0 = 0; c0 < NC; c0 += 16) {
            uint32_t g[16], u[16];
            tmem_ld16(t_acc + c0, g);
            if (GATED) tmem_ld16(t_acc + BN + c0, u);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float s = __shfl_sync(0xffffffffu, sx_l, c0 + j);
              acc_g[c0 + j] = fmaf(__uint_as_float(g[j]), s * swg, acc_g[c0 + j]);
              if (GATED) acc_u[c0 + j] = fmaf(__uint_as_float(u[j]), s * swu, acc_u[c0 + j]);
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty[ps]);
      }
      // fused output math + store (lane = feature, coalesced 2-byte stores per token)
      const int feat = f * kBM + ew * 32 + lane;
      T* out_col = out + static_cast<size_t>(tile.row_start) * ld_out + feat;
      float rw_lane = 1.0f;
      if (!GATED) {
        const int row = tile.row_start + my_col;
        if (my_col_ok && row < n_routed_rows) rw_lane = __ldg(row_weight + row);
      }
#pragma unroll
      for (int j = 0; j < NC; ++j) {
        const float rw = GATED ? 1.0f : __shfl_sync(0xffffffffu, rw_lane, j);
        if (col_base + j < nrows) {
          const float v = GATED ? silu_fast(acc_g[j]) * acc_u[j] : acc_g[j] * rw;
          out_col[static_cast<size_t>(col_base + j) * ld_out] = Dt<T>::from_f(v);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
}

// ------------------------------------------------------------------------------------------
// activation quantiser (+ gather): one warp per row, a half-warp per 128-element group
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) quant_rows_fp8_kernel(
    const T* __restrict__ src, int src_ld, const int32_t* __restrict__ row_token /*nullable*/,
    const int32_t* __restrict__ expert_offsets, int E, int T_tokens, int k, int K, int sh_lo, int sh_hi,
    uint8_t* __restrict__ dst, float* __restrict__ scale_t, int scale_ld, int pow2, uint8_t* __restrict__ sf) {
  pdl_wait();
  pdl_launch();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + warp;
  const int Tk = T_tokens * k;
  const int n_routed = expert_offsets[E];
  int src_row;
  if (row < Tk) {
    if (row >= n_routed) return;
    src_row = row_token ? row_token[row] : row;
  } else {
    const int t = row - Tk;
    if (t < sh_lo || t >= sh_hi) return;
    src_row = row_token ? t : row;
  }
  const T* s = src + static_cast<size_t>(src_row) * src_ld;
  uint8_t* d = dst + static_cast<size_t>(row) * K;
  // 256 elements per warp iteration: lanes 0-15 one 128-group, lanes 16-31 the next
  for (int c0 = 0; c0 < K; c0 += 256) {
    const int c = c0 + lane * 8;
    const bool ok = c < K;                      // K % 128 == 0: a half-warp is in or out as a whole
    float v[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) v[q] = 0.f;
    if (ok) unpack16<T>(ld_cached16(s + c), v);
    float amax = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) amax = fmaxf(amax, fabsf(v[q]));
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    // IEEE divisions, exactly the published formula (scale = amax / 448, q = rn_e4m3(v / scale)):
    // with 8-bit-mantissa inputs v / scale lands on e4m3 rounding ties often enough that a
    // reciprocal-multiply variant re-rounds ~0.3 % of the elements differently
    // (power-of-two tier: the next power of two at or above it, and the division is exact)
    uint32_t ebyte = 0;
    const float scale = pow2 ? pow2_scale(amax, &ebyte) : __fdiv_rn(fmaxf(amax, 1e-10f), 448.0f);
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      lo |= static_cast<uint32_t>(__nv_cvt_float_to_fp8(__fdiv_rn(v[q], scale), __NV_SATFINITE, __NV_E4M3)) << (8 * q);
      hi |= static_cast<uint32_t>(__nv_cvt_float_to_fp8(__fdiv_rn(v[4 + q], scale), __NV_SATFINITE, __NV_E4M3)) << (8 * q);
    }
    if (ok) {
      *reinterpret_cast<uint2*>(d + c) = make_uint2(lo, hi);
      if ((lane & 15) == 0) {
        const int g = c / 128;
        scale_t[static_cast<size_t>(g) * scale_ld + row] = scale;
        // packed UE8M0 copy for the block-scaled MMA: [g / 4, scale_ld] words, byte g % 4
        if (sf) sf[(static_cast<size_t>(g >> 2) * scale_ld + row) * 4 + (g & 3)] = static_cast<uint8_t>(ebyte);
      }
    }
  }
}

// Token-major variant for the first GEMM's input: every token is quantised ONCE (same formula) and its
// e4m3 row + scales are written to each of its k expert-major rows and its shared-expert row
// (slot_row[t, j] = row of (token t, slot j), -1 = expert not on this rank).  The row-major kernel above
// re-quantises a token k+1 times; at prefill that made the quantiser cost as much as a GEMM stage.
template <typename T>
__global__ void __launch_bounds__(256) quant_scatter_fp8_kernel(
    const T* __restrict__ x, int K, const int32_t* __restrict__ slot_row, int T_tokens, int k, int sh_lo,
    int sh_hi, uint8_t* __restrict__ dst, float* __restrict__ scale_t, int scale_ld, int pow2,
    uint8_t* __restrict__ sf) {
  pdl_wait();
  pdl_launch();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t = blockIdx.x * 8 + warp;
  if (t >= T_tokens) return;
  // lane j < k: row of slot j; lane k: the shared-expert row
  int my_row = -1;
  if (lane < k) my_row = slot_row[static_cast<size_t>(t) * k + lane];
  else if (lane == k && t >= sh_lo && t < sh_hi) my_row = T_tokens * k + t;
  const T* s = x + static_cast<size_t>(t) * K;
  for (int c0 = 0; c0 < K; c0 += 256) {
    const int c = c0 + lane * 8;
    const bool ok = c < K;
    float v[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) v[q] = 0.f;
    if (ok) unpack16<T>(ld_cached16(s + c), v);
    float amax = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) amax = fmaxf(amax, fabsf(v[q]));
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    uint32_t ebyte = 0;
    const float scale = pow2 ? pow2_scale(amax, &ebyte) : __fdiv_rn(fmaxf(amax, 1e-10f), 448.0f);
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      lo |= static_cast<uint32_t>(__nv_cvt_float_to_fp8(__fdiv_rn(v[q], scale), __NV_SATFINITE, __NV_E4M3)) << (8 * q);
      hi |= static_cast<uint32_t>(__nv_cvt_float_to_fp8(__fdiv_rn(v[4 + q], scale), __NV_SATFINITE, __NV_E4M3)) << (8 * q);
    }
    const int g = c / 128;
    for (int j = 0; j <= k; ++j) {
      const int row = __shfl_sync(0xffffffffu, my_row, j);
      if (row >= 0 && ok) {
        *reinterpret_cast<uint2*>(dst + static_cast<size_t>(row) * K + c) = make_uint2(lo, hi);
        if ((lane & 15) == 0) {
          scale_t[static_cast<size_t>(g) * scale_ld + row] = scale;
          if (sf) sf[(static_cast<size_t>(g >> 2) * scale_ld + row) * 4 + (g & 3)] = static_cast<uint8_t>(ebyte);
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int umma_make_tensor_map_u8(void* map_out, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows) {
  static PFN_encodeTiled enc = nullptr;
  if (!enc) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return GLM4MOE_ERR_DEVICE;
    enc = reinterpret_cast<PFN_encodeTiled>(p);
  }
  if (cols % 16 != 0 || box_rows == 0 || box_rows > 256 || rows == 0) return GLM4MOE_ERR_ARG;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {cols};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(kBKF8), box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(static_cast<CUtensorMap*>(map_out), CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base),
                   dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? GLM4MOE_OK : GLM4MOE_ERR_ARG;
}

template <typename T, int BN, bool GATED>
static int launch_fp8_one(const UmmaFp8Launch& a, cudaStream_t st) {
  using Cfg = Fp8Cfg<BN, GATED>;
  auto kern = umma_fp8_kernel<T, BN, GATED>;
  static unsigned long long attr_done = 0;   // per instantiation, one bit per device
  GLM4MOE_CUDA_CHECK(ensure_dynamic_smem(kern, Cfg::kSmemBytes, &attr_done));
  GLM4MOE_CUDA_CHECK(launch_pdl(kern, dim3(a.grid), dim3(Cfg::kThreads), Cfg::kSmemBytes, st,
                                *static_cast<const CUtensorMap*>(a.tm_w), *static_cast<const CUtensorMap*>(a.tm_x),
                                a.w_scale, a.x_scale_t, a.x_scale_ld, reinterpret_cast<const int4*>(a.tiles),
                                a.num_tiles, static_cast<T*>(a.out), a.ld_out, a.row_weight, a.n_routed_rows,
                                a.n_feat_tiles, a.n_kblocks, a.w_rows_per_expert, a.up_row_offset, a.sched));
  return GLM4MOE_OK;
}

template <typename T, bool GATED>
static int launch_fp8_bn(const UmmaFp8Launch& a, cudaStream_t st) {
  switch (a.tile_rows) {
    case 16: return launch_fp8_one<T, 16, GATED>(a, st);
    case 32: return launch_fp8_one<T, 32, GATED>(a, st);
    case 64: return launch_fp8_one<T, 64, GATED>(a, st);
    default: return GLM4MOE_ERR_ARG;
  }
}

int umma_fp8_launch(const UmmaFp8Launch& a, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (a.dtype == GLM4MOE_BF16)
    return a.gated ? launch_fp8_bn<__nv_bfloat16, true>(a, st) : launch_fp8_bn<__nv_bfloat16, false>(a, st);
  if (a.dtype == GLM4MOE_F16)
    return a.gated ? launch_fp8_bn<__half, true>(a, st) : launch_fp8_bn<__half, false>(a, st);
  return GLM4MOE_ERR_UNSUPPORTED;
}

int quant_rows_fp8(const void* src, int src_ld, const int32_t* row_token, const int32_t* expert_offsets, int E,
                   int T, int k, int K, int sh_lo, int sh_hi, void* dst, float* scale_t, int scale_ld, int dtype,
                   int pow2, void* sf, void* stream) {
  if (K % 256 != 0 && K % 128 != 0) return GLM4MOE_ERR_UNSUPPORTED;
  if (K % 128 != 0) return GLM4MOE_ERR_UNSUPPORTED;
  const int rows = T * k + sh_hi;
  if (rows <= 0) return GLM4MOE_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == GLM4MOE_BF16) {
    GLM4MOE_CUDA_CHECK(launch_pdl(quant_rows_fp8_kernel<__nv_bfloat16>, dim3((rows + 7) / 8), dim3(256), 0, st,
                                  static_cast<const __nv_bfloat16*>(src), src_ld, row_token, expert_offsets, E, T, k,
                                  K, sh_lo, sh_hi, static_cast<uint8_t*>(dst), scale_t, scale_ld, pow2,
                                  static_cast<uint8_t*>(sf)));
  } else if (dtype == GLM4MOE_F16) {
    GLM4MOE_CUDA_CHECK(launch_pdl(quant_rows_fp8_kernel<__half>, dim3((rows + 7) / 8), dim3(256), 0, st,
                                  static_cast<const __half*>(src), src_ld, row_token, expert_offsets, E, T, k, K,
                                  sh_lo, sh_hi, static_cast<uint8_t*>(dst), scale_t, scale_ld, pow2,
                                  static_cast<uint8_t*>(sf)));
  } else {
    return GLM4MOE_ERR_UNSUPPORTED;
  }
  return GLM4MOE_OK;
}

int quant_scatter_fp8(const void* x, int K, const int32_t* slot_row, int T, int k, int sh_lo, int sh_hi, void* dst,
                      float* scale_t, int scale_ld, int dtype, int pow2, void* sf, void* stream) {
  if (K % 128 != 0 || k >= 32) return GLM4MOE_ERR_UNSUPPORTED;
  if (T <= 0) return GLM4MOE_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == GLM4MOE_BF16) {
    GLM4MOE_CUDA_CHECK(launch_pdl(quant_scatter_fp8_kernel<__nv_bfloat16>, dim3((T + 7) / 8), dim3(256), 0, st,
                                  static_cast<const __nv_bfloat16*>(x), K, slot_row, T, k, sh_lo, sh_hi,
                                  static_cast<uint8_t*>(dst), scale_t, scale_ld, pow2, static_cast<uint8_t*>(sf)));
  } else if (dtype == GLM4MOE_F16) {
    GLM4MOE_CUDA_CHECK(launch_pdl(quant_scatter_fp8_kernel<__half>, dim3((T + 7) / 8), dim3(256), 0, st,
                                  static_cast<const __half*>(x), K, slot_row, T, k, sh_lo, sh_hi,
                                  static_cast<uint8_t*>(dst), scale_t, scale_ld, pow2, static_cast<uint8_t*>(sf)));
  } else {
    return GLM4MOE_ERR_UNSUPPORTED;
  }
  return GLM4MOE_OK;
}

}  // namespace glm4moe

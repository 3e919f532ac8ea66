// Fused decode front-end (T <= 32): router logits + expert selection + permutation in ONE launch.
//
// Replaces GlmMoeSelectTopk.forward (reference moe.py:28-50) and the mask / where loop bookkeeping
// (moe.py:135-144) for decode-size batches, where three separate launches cost more than their
// work.  Grid = (E/EB) x (T/TT) blocks compute the logits exactly as router_logits_small_kernel
// (8 warps split K, fixed-order fp32 accumulate, one rounding to the model dtype); every block then
// takes a ticket: the last block of a token group selects the experts of its TT tokens (one warp per
// token), and the last group to finish builds the permutation / work list of the whole batch
// (single-block path of permute.cu).
// Arithmetic and tie rule are the stand-alone kernels' (same device functions): bit-identical.
#include <cuda_fp8.h>

#include "common.cuh"
#include "route_device.cuh"
#include "ep_combine.cuh"

namespace glm4moe {

template <typename T, int TT, int EB>
__global__ void __launch_bounds__(256) route_small_kernel(
    const T* __restrict__ h, const T* __restrict__ gate_w, const float* __restrict__ bias, T* logits, int Tn,
    int H, int E, int k, float scale, int e_lo, int e_hi, int tile_rows, int sh_lo, int sh_hi, int32_t* idx,
    float* w, int32_t* __restrict__ expert_offsets, int32_t* __restrict__ row_token,
    float* __restrict__ row_weight, int32_t* __restrict__ slot_row, int4* __restrict__ tiles, int tile_capacity,
    int32_t* __restrict__ num_tiles, int32_t* __restrict__ ticket, unsigned char* __restrict__ xq,
    float* __restrict__ sx_t, int sx_ld, const RouteSmallEp ep) {
  constexpr int V = Dt<T>::kVec;
  __shared__ float part[8][TT * EB];
  __shared__ PermuteSmallSmem psm;
  __shared__ int s_last;
  // first kernel of the layer: the input must be complete before anything reads it; the dependants
  // (expert GEMMs) may start their own prologue right away
  pdl_wait();
  pdl_launch();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int e0 = blockIdx.x * EB, t0 = blockIdx.y * TT;
  const int nvec = H / V;
  if constexpr (sizeof(T) == 2) {
    // e4m3 tier: the input tokens are also the MMA token tile of the fused expert kernel; quantise
    // them here (per token, per 128 elements: scale = amax / 448, q = rn_e4m3(v / scale), the formula
    // of quant_rows_fp8_kernel).  This block's share: its TT tokens x the 128-groups {bx, bx + gridDim.x, ..}.
    if (xq != nullptr) {
      const int n_groups = H / 128;
      const int per_block = (n_groups - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) /
                            static_cast<int>(gridDim.x);
      const int n_items = TT * per_block;
      for (int base = warp * 2; base < n_items; base += 16) {
        const int item = base + (lane >> 4);
        const int t = t0 + item % TT, g = static_cast<int>(blockIdx.x) + (item / TT) * static_cast<int>(gridDim.x);
        const bool ok = item < n_items && t < Tn;
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = 0.f;
        const int c = g * 128 + (lane & 15) * 8;
        if (ok) unpack16<T>(ld_cached16(h + static_cast<size_t>(t) * H + c), v);
        float amax = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) amax = fmaxf(amax, fabsf(v[i]));
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
        const float sc = act_scale(amax, ep.act_pow2);
        uint32_t lo = 0, hi = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          lo |= static_cast<uint32_t>(__nv_cvt_float_to_fp8(__fdiv_rn(v[i], sc), __NV_SATFINITE, __NV_E4M3)) << (8 * i);
          hi |= static_cast<uint32_t>(__nv_cvt_float_to_fp8(__fdiv_rn(v[4 + i], sc), __NV_SATFINITE, __NV_E4M3)) << (8 * i);
        }
        if (ok) {
          *reinterpret_cast<uint2*>(xq + static_cast<size_t>(t) * H + c) = make_uint2(lo, hi);
          if ((lane & 15) == 0) sx_t[static_cast<size_t>(g) * sx_ld + t] = sc;
        }
      }
    }
  }
  float acc[TT][EB];
#pragma unroll
  for (int i = 0; i < TT; ++i)
#pragma unroll
    for (int j = 0; j < EB; ++j) acc[i][j] = 0.f;
  for (int c = threadIdx.x; c < nvec; c += 256) {
    float wv[EB][V], xv[TT][V];
#pragma unroll
    for (int j = 0; j < EB; ++j) {
      const int e = min(e0 + j, E - 1);
      unpack16<T>(ld_cached16(reinterpret_cast<const uint4*>(gate_w + static_cast<size_t>(e) * H) + c), wv[j]);
    }
#pragma unroll
    for (int i = 0; i < TT; ++i) {
      const int t = min(t0 + i, Tn - 1);
      unpack16<T>(ld_cached16(reinterpret_cast<const uint4*>(h + static_cast<size_t>(t) * H) + c), xv[i]);
    }
#pragma unroll
    for (int i = 0; i < TT; ++i)
#pragma unroll
      for (int j = 0; j < EB; ++j)
#pragma unroll
        for (int q = 0; q < V; ++q) acc[i][j] = fmaf(xv[i][q], wv[j][q], acc[i][j]);
  }
#pragma unroll
  for (int i = 0; i < TT; ++i)
#pragma unroll
    for (int j = 0; j < EB; ++j) {
      const float s = warp_sum(acc[i][j]);
      if (lane == 0) part[warp][i * EB + j] = s;
    }
  __syncthreads();
  if (threadIdx.x < TT * EB) {
    float s = 0.f;
#pragma unroll
    for (int wdx = 0; wdx < 8; ++wdx) s += part[wdx][threadIdx.x];
    const int t = t0 + threadIdx.x / EB, e = e0 + threadIdx.x % EB;
    if (t < Tn && e < E) logits[static_cast<size_t>(t) * E + e] = Dt<T>::from_f(s);
    __threadfence();
  }
  __syncthreads();
  // ticket 1: the last of the gridDim.x blocks that share these TT tokens selects their experts
  if (threadIdx.x == 0) {
    const int prev = atomicAdd(ticket + 1 + blockIdx.y, 1);
    s_last = prev == static_cast<int>(gridDim.x) - 1;
    if (s_last) ticket[1 + blockIdx.y] = 0;   // re-armed for the next launch (graph replay safe)
  }
  for (int i = threadIdx.x; i < E * kWordsPerExpert; i += 256) psm.bitmap[i] = 0u;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (warp < TT && t0 + warp < Tn) {
    const int t = t0 + warp;
    select_token<T>(logits + static_cast<size_t>(t) * E, bias, E, k, scale, lane, idx + static_cast<size_t>(t) * k,
                    w + static_cast<size_t>(t) * k);
    __threadfence();
  }
  __syncthreads();
  // ticket 2: the last token group builds the permutation of the whole batch
  if (threadIdx.x == 0) {
    const int prev = atomicAdd(ticket, 1);
    s_last = prev == static_cast<int>(gridDim.y) - 1;
    if (s_last) *ticket = 0;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  __syncthreads();   // idx / w of this block's own global writes are visible to the block
  // expert-parallel decode fast path: the peers read slot_row of call e from the buffer of e's parity
  // (ep_combine.cu); the epoch is final here (this kernel started after the previous combine retired)
  if (ep.epoch != nullptr) slot_row = ((*reinterpret_cast<const volatile int32_t*>(ep.epoch) + 1) & 1) ? ep.slot_row_alt[1] : ep.slot_row_alt[0];
  permute_small_body(psm, idx, w, Tn, k, E, e_lo, e_hi, tile_rows, sh_lo, sh_hi, expert_offsets, row_token,
                     row_weight, slot_row, tiles, tile_capacity, num_tiles);
}

template <typename T>
static int launch_route_small(const void* h, const void* gw, const float* bias, void* logits, int T_, int H, int E,
                              int k, float scale, int e_lo, int e_hi, int tile_rows, int sh_lo, int sh_hi,
                              int32_t* idx, float* w, int32_t* offsets, int32_t* row_token, float* row_weight,
                              int32_t* slot_row, int32_t* tiles, int tile_capacity, int32_t* num_tiles,
                              int32_t* ticket, void* xq, float* sx_t, int sx_ld, const RouteSmallEp& ep,
                              cudaStream_t st) {
  constexpr int V = Dt<T>::kVec;
  if (H % V != 0) return GLM4MOE_ERR_UNSUPPORTED;
  if (T_ <= 2) {
    dim3 grid((E + 3) / 4, (T_ + 1) / 2);
    GLM4MOE_CUDA_CHECK(launch_pdl(route_small_kernel<T, 2, 4>, grid, dim3(256), 0, st, static_cast<const T*>(h),
                                  static_cast<const T*>(gw), bias, static_cast<T*>(logits), T_, H, E, k, scale, e_lo,
                                  e_hi, tile_rows, sh_lo, sh_hi, idx, w, offsets, row_token, row_weight, slot_row,
                                  reinterpret_cast<int4*>(tiles), tile_capacity, num_tiles, ticket,
                                  static_cast<unsigned char*>(xq), sx_t, sx_ld, ep));
  } else {
    dim3 grid((E + 3) / 4, (T_ + 3) / 4);
    GLM4MOE_CUDA_CHECK(launch_pdl(route_small_kernel<T, 4, 4>, grid, dim3(256), 0, st, static_cast<const T*>(h),
                                  static_cast<const T*>(gw), bias, static_cast<T*>(logits), T_, H, E, k, scale, e_lo,
                                  e_hi, tile_rows, sh_lo, sh_hi, idx, w, offsets, row_token, row_weight, slot_row,
                                  reinterpret_cast<int4*>(tiles), tile_capacity, num_tiles, ticket,
                                  static_cast<unsigned char*>(xq), sx_t, sx_ld, ep));
  }
  return GLM4MOE_OK;
}

}  // namespace glm4moe

namespace glm4moe {
int route_small_ep(const void* h, const void* gate_w, const float* bias, void* logits, int T, int H,
                   int E, int k, float scale, int dtype, int e_lo, int e_hi, int tile_rows,
                   int shared_lo, int shared_hi, int32_t* idx, float* w, int32_t* expert_offsets,
                   int32_t* row_token, float* row_weight, int32_t* slot_row, int32_t* tiles,
                   int32_t tile_capacity, int32_t* num_tiles, int32_t* ticket, void* xq,
                   float* x_scale_t, int x_scale_ld, const RouteSmallEp& ep, void* stream) {
  if (!h || !gate_w || !bias || !logits || !idx || !w || !expert_offsets || !row_token || !row_weight ||
      !slot_row || !tiles || !num_tiles || !ticket)
    return GLM4MOE_ERR_ARG;
  if (T <= 0 || T > GLM4MOE_ROUTE_SMALL_MAX_TOKENS || H <= 0) return GLM4MOE_ERR_ARG;
  if (E <= 0 || E > kMaxExperts || E > 32 * kMaxSlots || k <= 0 || k > kMaxTopK || k > E)
    return GLM4MOE_ERR_UNSUPPORTED;
  if (tile_rows <= 0 || e_lo < 0 || e_hi > E || e_lo > e_hi || shared_lo < 0 || shared_hi > T ||
      shared_lo > shared_hi)
    return GLM4MOE_ERR_ARG;
  if (tile_capacity < glm4moe_tile_capacity(T, k, E, tile_rows)) return GLM4MOE_ERR_WORKSPACE;
  if (xq && (!x_scale_t || x_scale_ld < T || H % 128 != 0 || dtype == GLM4MOE_F32)) return GLM4MOE_ERR_ARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  GLM4MOE_DISPATCH_DTYPE(dtype, T_,
                         return launch_route_small<T_>(h, gate_w, bias, logits, T, H, E, k, scale, e_lo, e_hi,
                                                       tile_rows, shared_lo, shared_hi, idx, w, expert_offsets,
                                                       row_token, row_weight, slot_row, tiles, tile_capacity,
                                                       num_tiles, ticket, xq, x_scale_t, x_scale_ld, ep, st));
  return GLM4MOE_OK;
}
}  // namespace glm4moe

using namespace glm4moe;

extern "C" int glm4moe_route_small(const void* h, const void* gate_w, const float* bias, void* logits, int T, int H,
                                   int E, int k, float scale, int dtype, int e_lo, int e_hi, int tile_rows,
                                   int shared_lo, int shared_hi, int32_t* idx, float* w, int32_t* expert_offsets,
                                   int32_t* row_token, float* row_weight, int32_t* slot_row, int32_t* tiles,
                                   int32_t tile_capacity, int32_t* num_tiles, int32_t* ticket, void* xq,
                                   float* x_scale_t, int x_scale_ld, void* stream) {
  return route_small_ep(h, gate_w, bias, logits, T, H, E, k, scale, dtype, e_lo, e_hi, tile_rows, shared_lo, shared_hi,
                        idx, w, expert_offsets, row_token, row_weight, slot_row, tiles, tile_capacity, num_tiles, ticket,
                        xq, x_scale_t, x_scale_ld, RouteSmallEp{}, stream);
}

// Router kernels: replaces GlmMoeSelectTopk.forward (reference nanovllm/models/glm4_moe/moe.py:28-50).
//
//   router_logits_kernel  logits[T,E] = h @ Wg^T, fp32 accumulate in a fixed order, one rounding to
//                         the model dtype (the reference's F.linear output dtype, moe.py:32)
//   router_select_kernel  warp per token: sigmoid (rounded to model dtype, moe.py:37), + fp32 bias
//                         (moe.py:38), k rounds of warp-shuffle arg-max with the pinned tie rule
//                         "higher score, then lower expert id" (moe.py:39), un-biased gather,
//                         renormalise and scale with the reference's roundings (moe.py:41-46).
#include "common.cuh"
#include "route_device.cuh"

namespace glm4moe {

// ---------------------------------------------------------------------------------------------
// logits: block = 8 warps, warp w owns expert (blockIdx.x*8 + w); TT tokens per block.
// Each lane walks its 16-byte chunks of the expert row once and re-uses it for TT tokens.
// ---------------------------------------------------------------------------------------------
template <typename T, int TT>
__global__ void __launch_bounds__(256) router_logits_kernel(const T* __restrict__ h,
                                                            const T* __restrict__ gate_w,
                                                            T* __restrict__ logits, int Tn, int H,
                                                            int E) {
  constexpr int V = Dt<T>::kVec;
  pdl_wait();
  pdl_launch();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int e = blockIdx.x * 8 + warp;
  const int t0 = blockIdx.y * TT;
  if (e >= E) return;
  const int nvec = H / V;
  float acc[TT];
#pragma unroll
  for (int i = 0; i < TT; ++i) acc[i] = 0.f;
  const uint4* wrow = reinterpret_cast<const uint4*>(gate_w + static_cast<size_t>(e) * H);
  for (int c = lane; c < nvec; c += 32) {
    float wv[V];
    unpack16<T>(ld_cached16(wrow + c), wv);
#pragma unroll
    for (int i = 0; i < TT; ++i) {
      const int t = t0 + i;
      if (t < Tn) {
        float xv[V];
        unpack16<T>(ld_cached16(reinterpret_cast<const uint4*>(h + static_cast<size_t>(t) * H) + c), xv);
#pragma unroll
        for (int j = 0; j < V; ++j) acc[i] = fmaf(xv[j], wv[j], acc[i]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < TT; ++i) {
    const float s = warp_sum(acc[i]);
    const int t = t0 + i;
    if (lane == 0 && t < Tn) logits[static_cast<size_t>(t) * E + e] = Dt<T>::from_f(s);
  }
}

// ---------------------------------------------------------------------------------------------
// decode-size logits (T < 64): a block owns a TT x EB tile of (token, expert) pairs and its 8 warps
// split K, so T=1 still spreads the 1 MB router matrix over E/EB blocks x 8 warps with every
// lane's loads independent (latency-bound problem: parallelism and MLP, not reuse).
// ---------------------------------------------------------------------------------------------
template <typename T, int TT, int EB>
__global__ void __launch_bounds__(256) router_logits_small_kernel(const T* __restrict__ h,
                                                                  const T* __restrict__ gate_w,
                                                                  T* __restrict__ logits, int Tn,
                                                                  int H, int E) {
  constexpr int V = Dt<T>::kVec;
  __shared__ float part[8][TT * EB];
  pdl_wait();
  pdl_launch();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int e0 = blockIdx.x * EB, t0 = blockIdx.y * TT;
  const int nvec = H / V;
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
  }
}

// ---------------------------------------------------------------------------------------------
// select: one warp per token, expert e = j*32 + lane held in register slot j (E <= 256).
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(128) router_select_kernel(const T* __restrict__ logits,
                                                            const float* __restrict__ bias, int Tn,
                                                            int E, int k, float scale,
                                                            int32_t* __restrict__ idx_out,
                                                            float* __restrict__ w_out) {
  pdl_wait();
  pdl_launch();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t = blockIdx.x * 4 + warp;
  if (t >= Tn) return;
  select_token<T>(logits + static_cast<size_t>(t) * E, bias, E, k, scale, lane,
                  idx_out + static_cast<size_t>(t) * k, w_out + static_cast<size_t>(t) * k);
}

template <typename T>
static int launch_logits(const void* h, const void* gw, void* logits, int T_, int H, int E,
                         cudaStream_t st) {
  constexpr int V = Dt<T>::kVec;
  if (H % V != 0) return GLM4MOE_ERR_UNSUPPORTED;
  if (T_ <= 2) {
    dim3 grid((E + 3) / 4, (T_ + 1) / 2);
    GLM4MOE_CUDA_CHECK(launch_pdl(router_logits_small_kernel<T, 2, 4>, grid, dim3(256), 0, st,
                                  static_cast<const T*>(h), static_cast<const T*>(gw), static_cast<T*>(logits), T_, H, E));
  } else if (T_ <= 256) {
    dim3 grid((E + 3) / 4, (T_ + 3) / 4);
    GLM4MOE_CUDA_CHECK(launch_pdl(router_logits_small_kernel<T, 4, 4>, grid, dim3(256), 0, st,
                                  static_cast<const T*>(h), static_cast<const T*>(gw), static_cast<T*>(logits), T_, H, E));
  } else {
    dim3 grid((E + 7) / 8, (T_ + 7) / 8);
    GLM4MOE_CUDA_CHECK(launch_pdl(router_logits_kernel<T, 8>, grid, dim3(256), 0, st, static_cast<const T*>(h),
                                  static_cast<const T*>(gw), static_cast<T*>(logits), T_, H, E));
  }
  GLM4MOE_LAUNCH_CHECK();
  return GLM4MOE_OK;
}

}  // namespace glm4moe

using namespace glm4moe;

extern "C" int glm4moe_router_logits(const void* h, const void* gate_w, void* logits, int T, int H,
                                     int E, int dtype, void* stream) {
  if (!h || !gate_w || !logits || T < 0 || H <= 0 || E <= 0) return GLM4MOE_ERR_ARG;
  if (T == 0) return GLM4MOE_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  GLM4MOE_DISPATCH_DTYPE(dtype, T_, return launch_logits<T_>(h, gate_w, logits, T, H, E, st));
  return GLM4MOE_OK;
}

extern "C" int glm4moe_router_select(const void* logits, const float* bias, int T, int E, int k,
                                     float scale, int dtype, int32_t* idx, float* w, void* stream) {
  if (!logits || !bias || !idx || !w || T < 0) return GLM4MOE_ERR_ARG;
  if (E <= 0 || E > 32 * kMaxSlots || k <= 0 || k > kMaxTopK || k > E) return GLM4MOE_ERR_UNSUPPORTED;
  if (T == 0) return GLM4MOE_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  GLM4MOE_DISPATCH_DTYPE(dtype, T_,
                         GLM4MOE_CUDA_CHECK(launch_pdl(router_select_kernel<T_>, dim3((T + 3) / 4), dim3(128), 0, st,
                                                       static_cast<const T_*>(logits), bias, T, E, k, scale, idx, w)));
  return GLM4MOE_OK;
}

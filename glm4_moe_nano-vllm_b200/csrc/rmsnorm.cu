// Residual add + RMSNorm, the step in front of the sparse block in the decoder layer
// (reference nanovllm/layers/layernorm.py:29-40 `add_rms_forward`, called at
// models/glm4_moe/decode_layer.py:98 as post_attention_layernorm), with the reference's dtype flow:
//   x32 = float(x) + float(residual)          residual_out = round(x32)
//   var = mean(x32^2)  (fp32, over the un-rounded sum)
//   y   = round(x32 * rsqrt(var + eps))       out = round(y * weight)      (model-dtype multiply)
// residual == nullptr is `rms_forward` (layernorm.py:17-26): x32 = float(x), no residual_out.
// One block per token row, 128-bit loads/stores, the row stays in registers between the two passes.
#include "common.cuh"

namespace glm4moe {

constexpr int kNormThreads = 256;
constexpr int kNormMaxVec = 4;     // 16-byte vectors per thread: H <= 256 * 4 * 8 = 8192 two-byte elements

template <typename T>
__global__ void __launch_bounds__(kNormThreads) add_rmsnorm_kernel(
    const T* __restrict__ x, const T* __restrict__ residual, const T* __restrict__ weight, float eps, int H,
    T* __restrict__ out, T* __restrict__ residual_out) {
  constexpr int V = Dt<T>::kVec;
  __shared__ float s_part[kNormThreads / 32];
  pdl_wait();
  pdl_launch();
  const size_t row = static_cast<size_t>(blockIdx.x) * H;
  const int nvec = H / V;
  float v[kNormMaxVec][V];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < kNormMaxVec; ++i) {
    const int c = threadIdx.x + i * kNormThreads;
    if (c < nvec) {
      unpack16<T>(ld_cached16(reinterpret_cast<const uint4*>(x + row) + c), v[i]);
      if (residual != nullptr) {
        float r[V];
        unpack16<T>(ld_cached16(reinterpret_cast<const uint4*>(residual + row) + c), r);
#pragma unroll
        for (int q = 0; q < V; ++q) v[i][q] += r[q];
        __align__(16) T o[V];
#pragma unroll
        for (int q = 0; q < V; ++q) o[q] = Dt<T>::from_f(v[i][q]);
        reinterpret_cast<uint4*>(residual_out + row)[c] = *reinterpret_cast<uint4*>(o);
      }
#pragma unroll
      for (int q = 0; q < V; ++q) ss = fmaf(v[i][q], v[i][q], ss);
    }
  }
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = ss;
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int w = 0; w < kNormThreads / 32; ++w) tot += s_part[w];
  const float rs = rsqrtf(tot / static_cast<float>(H) + eps);
#pragma unroll
  for (int i = 0; i < kNormMaxVec; ++i) {
    const int c = threadIdx.x + i * kNormThreads;
    if (c < nvec) {
      float g[V];
      unpack16<T>(ld_cached16(reinterpret_cast<const uint4*>(weight) + c), g);
      __align__(16) T o[V];
#pragma unroll
      for (int q = 0; q < V; ++q) o[q] = Dt<T>::from_f(Dt<T>::round(v[i][q] * rs) * g[q]);
      reinterpret_cast<uint4*>(out + row)[c] = *reinterpret_cast<uint4*>(o);
    }
  }
}

}  // namespace glm4moe

using namespace glm4moe;

extern "C" int glm4moe_add_rmsnorm(const void* x, const void* residual, const void* weight, float eps, int T,
                                   int H, int dtype, void* out, void* residual_out, void* stream) {
  if (!x || !weight || !out || (residual && !residual_out)) return GLM4MOE_ERR_ARG;
  if (T < 0 || H <= 0) return GLM4MOE_ERR_ARG;
  if (T == 0) return GLM4MOE_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  GLM4MOE_DISPATCH_DTYPE(dtype, T_, {
    constexpr int V = Dt<T_>::kVec;
    if (H % V != 0 || H / V > kNormThreads * kNormMaxVec) return GLM4MOE_ERR_UNSUPPORTED;
    GLM4MOE_CUDA_CHECK(launch_pdl(add_rmsnorm_kernel<T_>, dim3(T), dim3(kNormThreads), 0, st,
                                  static_cast<const T_*>(x), static_cast<const T_*>(residual),
                                  static_cast<const T_*>(weight), eps, H, static_cast<T_*>(out),
                                  static_cast<T_*>(residual_out)));
  });
  return GLM4MOE_OK;
}

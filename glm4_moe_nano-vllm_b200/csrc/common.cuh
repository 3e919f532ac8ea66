// Shared device helpers for the GLM4-MoE sm_100a kernels.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/glm4_moe_b200.h"

#define GLM4MOE_CUDA_CHECK(expr)                    \
  do {                                              \
    cudaError_t _e = (expr);                        \
    if (_e != cudaSuccess) return static_cast<int>(_e); \
  } while (0)

#define GLM4MOE_LAUNCH_CHECK()                      \
  do {                                              \
    cudaError_t _e = cudaGetLastError();            \
    if (_e != cudaSuccess) return static_cast<int>(_e); \
  } while (0)

namespace glm4moe {

constexpr int kWarp = 32;

// ---- model-dtype traits: load as fp32, store with round-to-nearest-even -----------------------
template <typename T> struct Dt;
template <> struct Dt<float> {
  static constexpr int kVec = 4;  // elements per 16-byte vector
  __device__ __forceinline__ static float to_f(float v) { return v; }
  __device__ __forceinline__ static float from_f(float v) { return v; }
  __device__ __forceinline__ static float round(float v) { return v; }
};
template <> struct Dt<__half> {
  static constexpr int kVec = 8;
  __device__ __forceinline__ static float to_f(__half v) { return __half2float(v); }
  __device__ __forceinline__ static __half from_f(float v) { return __float2half_rn(v); }
  __device__ __forceinline__ static float round(float v) { return __half2float(__float2half_rn(v)); }
};
template <> struct Dt<__nv_bfloat16> {
  static constexpr int kVec = 8;
  __device__ __forceinline__ static float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
  __device__ __forceinline__ static __nv_bfloat16 from_f(float v) { return __float2bfloat16_rn(v); }
  __device__ __forceinline__ static float round(float v) {
    return __bfloat162float(__float2bfloat16_rn(v));
  }
};

// 16-byte vector of model-dtype elements unpacked to fp32
template <typename T> struct Vec16 {
  static constexpr int N = Dt<T>::kVec;
  float v[N];
};

template <typename T>
__device__ __forceinline__ void unpack16(const uint4& raw, float* out);
template <>
__device__ __forceinline__ void unpack16<float>(const uint4& raw, float* out) {
  out[0] = __uint_as_float(raw.x); out[1] = __uint_as_float(raw.y);
  out[2] = __uint_as_float(raw.z); out[3] = __uint_as_float(raw.w);
}
template <>
__device__ __forceinline__ void unpack16<__nv_bfloat16>(const uint4& raw, float* out) {
  // bf16 -> fp32 is a 16-bit shift
  out[0] = __uint_as_float(raw.x << 16); out[1] = __uint_as_float(raw.x & 0xffff0000u);
  out[2] = __uint_as_float(raw.y << 16); out[3] = __uint_as_float(raw.y & 0xffff0000u);
  out[4] = __uint_as_float(raw.z << 16); out[5] = __uint_as_float(raw.z & 0xffff0000u);
  out[6] = __uint_as_float(raw.w << 16); out[7] = __uint_as_float(raw.w & 0xffff0000u);
}
template <>
__device__ __forceinline__ void unpack16<__half>(const uint4& raw, float* out) {
  const __half2* h = reinterpret_cast<const __half2*>(&raw);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 f = __half22float2(h[i]);
    out[2 * i] = f.x; out[2 * i + 1] = f.y;
  }
}

// streaming 128-bit load that does not pollute L1 (weights are read exactly once)
__device__ __forceinline__ uint4 ld_stream16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ uint4 ld_cached16(const void* p) {
  return __ldg(reinterpret_cast<const uint4*>(p));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ float sigmoid_f(float x) { return 1.0f / (1.0f + expf(-x)); }
__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + expf(-x)); }
// tensor-core epilogues: fast exp + fast divide (rel. error ~1e-6, far below the 2-byte output ulp)
__device__ __forceinline__ float silu_fast(float x) { return __fdividef(x, 1.0f + __expf(-x)); }

// FP8 activation scales of one 1x128 group.  Default tier: scale = amax / 448 (fp32).  Power-of-two tier
// (GLM4MOE_W_FP8_BLOCK128_POW2): the smallest power of two >= amax / 448, i.e. a UE8M0 exponent byte, the
// only scale type the tcgen05 block-scaled MMA reads (umma_fp8_mx.cu); v / scale is then exact.
__device__ __forceinline__ float pow2_scale(float amax, uint32_t* ebyte) {
  const uint32_t b = __float_as_uint(__fdiv_rn(fmaxf(amax, 1e-10f), 448.0f));
  const uint32_t e = (b >> 23) + ((b & 0x7fffffu) != 0u ? 1u : 0u);   // biased exponent of the ceiling
  *ebyte = e;
  return __uint_as_float(e << 23);
}
__device__ __forceinline__ float pow2_inverse(uint32_t ebyte) { return __uint_as_float((254u - ebyte) << 23); }
__device__ __forceinline__ float act_scale(float amax, int pow2) {
  uint32_t e;
  return pow2 ? pow2_scale(amax, &e) : __fdiv_rn(fmaxf(amax, 1e-10f), 448.0f);
}

// Programmatic dependent launch.  Every kernel on the forward path runs
//   [prologue that touches nothing a predecessor writes] -> pdl_wait() -> pdl_launch() -> body
// so a kernel only starts once its predecessor passed its own wait: everything two or more
// launches back is complete and visible, and only the direct predecessor's outputs need the wait.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool pdl_enabled();   // api.cu: false when GLM4MOE_NO_PDL=1
void note_launch();   // api.cu: per-thread count of kernels launched (glm4moe_layer_last_launches)

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  note_launch();
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// cudaFuncAttributeMaxDynamicSharedMemorySize is per device (per context): set it once per
// (kernel instantiation, device).  `done` is the instantiation's own bit mask of devices served.
template <typename K>
inline cudaError_t ensure_dynamic_smem(K kern, int bytes, unsigned long long* done) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const unsigned long long bit = 1ull << (dev & 63);
  if (*done & bit) return cudaSuccess;
  e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) *done |= bit;
  return e;
}

inline int dtype_size(int dtype) { return dtype == GLM4MOE_F32 ? 4 : 2; }

// dispatch a templated launcher on the model dtype
#define GLM4MOE_DISPATCH_DTYPE(dtype, T, ...)                     \
  switch (dtype) {                                                \
    case GLM4MOE_F32: { using T = float; __VA_ARGS__; break; }    \
    case GLM4MOE_F16: { using T = __half; __VA_ARGS__; break; }   \
    case GLM4MOE_BF16: { using T = __nv_bfloat16; __VA_ARGS__; break; } \
    default: return GLM4MOE_ERR_ARG;                              \
  }

}  // namespace glm4moe

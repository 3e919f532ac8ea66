// CUDA-core expert MLP + combine.  This is the exact-arithmetic path: fp32 model dtype (the
// tiny BASELINE config), shapes the tcgen05 tiles do not divide, and the cross-check for the
// tensor-core path.  Replaces Glm4MoeMLP.forward (reference mlp.py:49-54 -> linear.py:74-75,
// activation.py:12-14, linear.py:151-152) applied per hit expert (moe.py:147-149) and to the
// shared expert (moe.py:165), and the index_add_/final add (moe.py:151,171).
//
// Work list = tiles {expert, row_start, nrows<=RT} from the permute scan.  A CTA stages the
// tile's RT activation rows in shared memory once, then its 8 warps stream weight rows: each
// warp owns one output feature at a time and every 16-byte weight chunk it loads is re-used
// for all RT rows (weights cross HBM once per tile).
#include "common.cuh"

namespace glm4moe {

constexpr int RT = 8;        // rows per tile on this path
constexpr int kColsPerUnit = 64;

template <typename T>
__device__ __forceinline__ void stage_rows(const T* __restrict__ src, size_t ld, int row_start,
                                           int nrows, int K, T* smem) {
  constexpr int V = Dt<T>::kVec;
  const int vec_per_row = K / V;
  for (int i = threadIdx.x; i < RT * vec_per_row; i += blockDim.x) {
    const int r = i / vec_per_row, c = i - r * vec_per_row;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (r < nrows)
      v = __ldg(reinterpret_cast<const uint4*>(src + (static_cast<size_t>(row_start) + r) * ld) + c);
    reinterpret_cast<uint4*>(smem)[i] = v;
  }
}

// act[r, j] = silu(x_r . Wg[j]) * (x_r . Wu[j])
template <typename T>
__global__ void __launch_bounds__(256) simt_gate_up_kernel(
    const T* __restrict__ x_rows, const T* __restrict__ w_gate_up, const int4* __restrict__ tiles,
    const int32_t* __restrict__ num_tiles, T* __restrict__ act, int H, int I) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* xs = reinterpret_cast<T*>(smem_raw);
  constexpr int V = Dt<T>::kVec;
  pdl_wait();
  pdl_launch();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int units_per_tile = (I + kColsPerUnit - 1) / kColsPerUnit;
  const int total = *num_tiles * units_per_tile;
  const int nvec = H / V;
  for (int u = blockIdx.x; u < total; u += gridDim.x) {
    const int ti = u / units_per_tile, cb = u - ti * units_per_tile;
    const int4 tile = tiles[ti];
    __syncthreads();
    stage_rows<T>(x_rows, H, tile.y, tile.z, H, xs);
    __syncthreads();
    const T* wbase = w_gate_up + static_cast<size_t>(tile.x) * 2 * I * H;
    for (int j = cb * kColsPerUnit + warp; j < min(I, (cb + 1) * kColsPerUnit); j += 8) {
      const uint4* wg = reinterpret_cast<const uint4*>(wbase + static_cast<size_t>(j) * H);
      const uint4* wu = reinterpret_cast<const uint4*>(wbase + static_cast<size_t>(I + j) * H);
      float ag[RT], au[RT];
#pragma unroll
      for (int r = 0; r < RT; ++r) { ag[r] = 0.f; au[r] = 0.f; }
      for (int c = lane; c < nvec; c += 32) {
        float g[V], up[V];
        unpack16<T>(ld_stream16(wg + c), g);
        unpack16<T>(ld_stream16(wu + c), up);
#pragma unroll
        for (int r = 0; r < RT; ++r) {
          float xv[V];
          unpack16<T>(reinterpret_cast<const uint4*>(xs)[r * nvec + c], xv);
#pragma unroll
          for (int q = 0; q < V; ++q) {
            ag[r] = fmaf(xv[q], g[q], ag[r]);
            au[r] = fmaf(xv[q], up[q], au[r]);
          }
        }
      }
#pragma unroll
      for (int r = 0; r < RT; ++r) {
        const float gs = warp_sum(ag[r]);
        const float us = warp_sum(au[r]);
        if (lane == r && r < tile.z)
          act[(static_cast<size_t>(tile.y) + r) * I + j] = Dt<T>::from_f(silu_f(gs) * us);
      }
    }
  }
}

// y[r, h] = (act_r . Wd[h]) * row_weight[r]
template <typename T>
__global__ void __launch_bounds__(256) simt_down_kernel(
    const T* __restrict__ act, const T* __restrict__ w_down, const int4* __restrict__ tiles,
    const int32_t* __restrict__ num_tiles, const float* __restrict__ row_weight,
    T* __restrict__ y, int H, int I, int n_routed_rows) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* as = reinterpret_cast<T*>(smem_raw);
  constexpr int V = Dt<T>::kVec;
  pdl_wait();
  pdl_launch();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int units_per_tile = (H + kColsPerUnit - 1) / kColsPerUnit;
  const int total = *num_tiles * units_per_tile;
  const int nvec = I / V;
  for (int u = blockIdx.x; u < total; u += gridDim.x) {
    const int ti = u / units_per_tile, cb = u - ti * units_per_tile;
    const int4 tile = tiles[ti];
    __syncthreads();
    stage_rows<T>(act, I, tile.y, tile.z, I, as);
    __syncthreads();
    const T* wbase = w_down + static_cast<size_t>(tile.x) * H * I;
    float rw = 1.0f;
    if (lane < tile.z && tile.y + lane < n_routed_rows) rw = row_weight[tile.y + lane];
    for (int hcol = cb * kColsPerUnit + warp; hcol < min(H, (cb + 1) * kColsPerUnit); hcol += 8) {
      const uint4* wd = reinterpret_cast<const uint4*>(wbase + static_cast<size_t>(hcol) * I);
      float acc[RT];
#pragma unroll
      for (int r = 0; r < RT; ++r) acc[r] = 0.f;
      for (int c = lane; c < nvec; c += 32) {
        float wv[V];
        unpack16<T>(ld_stream16(wd + c), wv);
#pragma unroll
        for (int r = 0; r < RT; ++r) {
          float av[V];
          unpack16<T>(reinterpret_cast<const uint4*>(as)[r * nvec + c], av);
#pragma unroll
          for (int q = 0; q < V; ++q) acc[r] = fmaf(av[q], wv[q], acc[r]);
        }
      }
#pragma unroll
      for (int r = 0; r < RT; ++r) {
        const float s = warp_sum(acc[r]);
        if (lane == r && r < tile.z)
          y[(static_cast<size_t>(tile.y) + r) * H + hcol] = Dt<T>::from_f(s * rw);
      }
    }
  }
}

// out[t] = sum over slots (ascending expert id) of y[slot_row[t,s]]  (+ y[T*k + t])
template <typename T>
__global__ void __launch_bounds__(256) combine_kernel(const T* __restrict__ y,
                                                      const int32_t* __restrict__ slot_row,
                                                      T* __restrict__ out, int Tn, int k, int H,
                                                      int sh_lo, int sh_hi) {
  constexpr int V = Dt<T>::kVec;
  const int t = blockIdx.x;
  const int nvec = H / V;
  __shared__ int rows[64];
  pdl_wait();
  pdl_launch();
  if (threadIdx.x < k) rows[threadIdx.x] = slot_row[static_cast<size_t>(t) * k + threadIdx.x];
  __syncthreads();
  constexpr int KMAX = 8;
  if (k <= KMAX) {
    // All row loads of a column chunk are issued before the first add (k + 1 independent 16-byte loads in
    // flight per thread instead of one: the kernel is a pure HBM gather), then summed in slot order.
    const int sh_row = (t >= sh_lo && t < sh_hi) ? Tn * k + t : -1;
    for (int c = threadIdx.x; c < nvec; c += blockDim.x) {
      uint4 raw[KMAX + 1];
#pragma unroll
      for (int s = 0; s < KMAX; ++s)
        if (s < k && rows[s] >= 0)
          raw[s] = ld_stream16(reinterpret_cast<const uint4*>(y + static_cast<size_t>(rows[s]) * H) + c);
      if (sh_row >= 0) raw[KMAX] = ld_stream16(reinterpret_cast<const uint4*>(y + static_cast<size_t>(sh_row) * H) + c);
      float acc[V];
#pragma unroll
      for (int q = 0; q < V; ++q) acc[q] = 0.f;
#pragma unroll
      for (int s = 0; s <= KMAX; ++s) {
        const bool on = s < KMAX ? (s < k && rows[s] >= 0) : sh_row >= 0;
        if (on) {
          float v[V];
          unpack16<T>(raw[s], v);
#pragma unroll
          for (int q = 0; q < V; ++q) acc[q] += v[q];
        }
      }
      __align__(16) T o[V];
#pragma unroll
      for (int q = 0; q < V; ++q) o[q] = Dt<T>::from_f(acc[q]);
      reinterpret_cast<uint4*>(out + static_cast<size_t>(t) * H)[c] = *reinterpret_cast<uint4*>(o);
    }
    return;
  }
  for (int c = threadIdx.x; c < nvec; c += blockDim.x) {
    float acc[V];
#pragma unroll
    for (int q = 0; q < V; ++q) acc[q] = 0.f;
    for (int s = 0; s < k; ++s) {
      const int r = rows[s];
      if (r < 0) continue;
      float v[V];
      unpack16<T>(ld_stream16(reinterpret_cast<const uint4*>(y + static_cast<size_t>(r) * H) + c), v);
#pragma unroll
      for (int q = 0; q < V; ++q) acc[q] += v[q];
    }
    if (t >= sh_lo && t < sh_hi) {
      float v[V];
      const size_t r = static_cast<size_t>(Tn) * k + t;
      unpack16<T>(ld_stream16(reinterpret_cast<const uint4*>(y + r * H) + c), v);
#pragma unroll
      for (int q = 0; q < V; ++q) acc[q] += v[q];
    }
    __align__(16) T o[V];
#pragma unroll
    for (int q = 0; q < V; ++q) o[q] = Dt<T>::from_f(acc[q]);
    reinterpret_cast<uint4*>(out + static_cast<size_t>(t) * H)[c] = *reinterpret_cast<uint4*>(o);
  }
}

template <typename T>
static int launch_gate_up(const void* x_rows, const void* w, const int32_t* tiles,
                          const int32_t* num_tiles, void* act, int H, int I, cudaStream_t st) {
  constexpr int V = Dt<T>::kVec;
  if (H % V || I % V) return GLM4MOE_ERR_UNSUPPORTED;
  const size_t smem = static_cast<size_t>(RT) * H * sizeof(T);
  if (smem > 200 * 1024) return GLM4MOE_ERR_UNSUPPORTED;
  GLM4MOE_CUDA_CHECK(cudaFuncSetAttribute(simt_gate_up_kernel<T>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int per_sm = smem > 100 * 1024 ? 1 : (smem > 64 * 1024 ? 2 : 3);
  GLM4MOE_CUDA_CHECK(launch_pdl(simt_gate_up_kernel<T>, dim3(sms * per_sm), dim3(256), smem, st,
                                static_cast<const T*>(x_rows), static_cast<const T*>(w),
                                reinterpret_cast<const int4*>(tiles), num_tiles, static_cast<T*>(act), H, I));
  return GLM4MOE_OK;
}

template <typename T>
static int launch_down(const void* act, const void* w, const int32_t* tiles,
                       const int32_t* num_tiles, const float* row_weight, void* y, int H, int I,
                       int n_routed_rows, cudaStream_t st) {
  constexpr int V = Dt<T>::kVec;
  if (H % V || I % V) return GLM4MOE_ERR_UNSUPPORTED;
  const size_t smem = static_cast<size_t>(RT) * I * sizeof(T);
  if (smem > 200 * 1024) return GLM4MOE_ERR_UNSUPPORTED;
  GLM4MOE_CUDA_CHECK(cudaFuncSetAttribute(simt_down_kernel<T>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int per_sm = smem > 100 * 1024 ? 1 : (smem > 64 * 1024 ? 2 : 3);
  GLM4MOE_CUDA_CHECK(launch_pdl(simt_down_kernel<T>, dim3(sms * per_sm), dim3(256), smem, st,
                                static_cast<const T*>(act), static_cast<const T*>(w),
                                reinterpret_cast<const int4*>(tiles), num_tiles, row_weight, static_cast<T*>(y), H, I,
                                n_routed_rows));
  return GLM4MOE_OK;
}

}  // namespace glm4moe

using namespace glm4moe;

extern "C" int glm4moe_expert_gate_up_simt(const void* x_rows, const void* w_gate_up,
                                           const int32_t* tiles, const int32_t* num_tiles, void* act,
                                           int H, int I, int tile_rows, int dtype, void* stream) {
  if (!x_rows || !w_gate_up || !tiles || !num_tiles || !act) return GLM4MOE_ERR_ARG;
  if (tile_rows != RT) return GLM4MOE_ERR_ARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  GLM4MOE_DISPATCH_DTYPE(dtype, T_,
                         return launch_gate_up<T_>(x_rows, w_gate_up, tiles, num_tiles, act, H, I, st));
  return GLM4MOE_OK;
}

extern "C" int glm4moe_expert_down_simt(const void* act, const void* w_down, const int32_t* tiles,
                                        const int32_t* num_tiles, const float* row_weight, void* y,
                                        int H, int I, int T, int k, int tile_rows, int dtype,
                                        void* stream) {
  if (!act || !w_down || !tiles || !num_tiles || !row_weight || !y) return GLM4MOE_ERR_ARG;
  if (tile_rows != RT) return GLM4MOE_ERR_ARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  GLM4MOE_DISPATCH_DTYPE(
      dtype, T_, return launch_down<T_>(act, w_down, tiles, num_tiles, row_weight, y, H, I, T * k, st));
  return GLM4MOE_OK;
}

extern "C" int glm4moe_combine(const void* y, const int32_t* slot_row, void* out, int T, int k,
                               int H, int shared_lo, int shared_hi, int dtype, void* stream) {
  if (!y || !slot_row || !out || T < 0 || k <= 0 || k > 64) return GLM4MOE_ERR_ARG;
  if ((static_cast<size_t>(H) * dtype_size(dtype)) % 16 != 0) return GLM4MOE_ERR_UNSUPPORTED;
  if (T == 0) return GLM4MOE_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  GLM4MOE_DISPATCH_DTYPE(dtype, T_,
                         GLM4MOE_CUDA_CHECK(launch_pdl(combine_kernel<T_>, dim3(T), dim3(256), 0, st,
                                                       static_cast<const T_*>(y), slot_row, static_cast<T_*>(out), T,
                                                       k, H, shared_lo, shared_hi)));
  return GLM4MOE_OK;
}

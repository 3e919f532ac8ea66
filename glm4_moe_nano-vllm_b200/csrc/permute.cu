// Permutation kernels: replace the per-expert mask / where / gather loop of the reference
// (nanovllm/models/glm4_moe/moe.py:135-147) with one histogram + prefix-sum + stable scatter.
//
// The order produced is exactly the order the reference loop visits (token, slot) pairs in:
// expert-major, token-ascending inside an expert.  A token selects an expert
// at most once, so the rank of token t inside expert e is the number of smaller tokens that
// selected e — computed from a per-expert bitmap with popcounts, no atomics on the order.
//
//   rank_kernel   (grid = ceil(T/256)): per-block bitmap -> local rank of every (t,slot) and the
//                 per-block per-expert counts
//   scan_kernel   (1 block): block-exclusive prefix per expert, expert offsets, tile work list
//   scatter_kernel(grid = ceil(T/256)): row = offsets[e] + block_base[b][e] + local rank
#include <atomic>

#include "common.cuh"
#include "route_device.cuh"

namespace glm4moe {

__global__ void __launch_bounds__(kTokPerBlock) permute_rank_kernel(
    const int32_t* __restrict__ idx, int T, int k, int E, int e_lo, int e_hi,
    int32_t* __restrict__ local_rank /*[T*k]*/, int32_t* __restrict__ block_counts /*[nblk][E]*/) {
  __shared__ uint32_t bitmap[kMaxExperts * kWordsPerExpert];
  pdl_wait();
  pdl_launch();
  const int tid = threadIdx.x;
  const int t = blockIdx.x * kTokPerBlock + tid;
  for (int i = tid; i < E * kWordsPerExpert; i += kTokPerBlock) bitmap[i] = 0u;
  __syncthreads();
  const int word = tid >> 5;
  const uint32_t bit = 1u << (tid & 31);
  if (t < T) {
    for (int s = 0; s < k; ++s) {
      const int e = idx[static_cast<size_t>(t) * k + s];
      if (e >= e_lo && e < e_hi) atomicOr(&bitmap[e * kWordsPerExpert + word], bit);
    }
  }
  __syncthreads();
  if (t < T) {
    for (int s = 0; s < k; ++s) {
      const int e = idx[static_cast<size_t>(t) * k + s];
      int r = -1;
      if (e >= e_lo && e < e_hi) {
        const uint32_t* bm = &bitmap[e * kWordsPerExpert];
        r = __popc(bm[word] & (bit - 1u));
        for (int wd = 0; wd < word; ++wd) r += __popc(bm[wd]);
      }
      local_rank[static_cast<size_t>(t) * k + s] = r;
    }
  }
  for (int e = tid; e < E; e += kTokPerBlock) {
    int c = 0;
#pragma unroll
    for (int wd = 0; wd < kWordsPerExpert; ++wd) c += __popc(bitmap[e * kWordsPerExpert + wd]);
    block_counts[static_cast<size_t>(blockIdx.x) * E + e] = c;
  }
}

// One block of 256 threads; thread e owns expert e.
__global__ void __launch_bounds__(256) permute_scan_kernel(
    int32_t* __restrict__ block_counts /*[nblk][E] in: counts, out: exclusive prefix over blocks*/,
    int nblk, int E, int T, int k, int e_lo, int e_hi, int tile_rows, int sh_lo, int sh_hi,
    int32_t* __restrict__ expert_offsets /*[E+1]*/, int4* __restrict__ tiles, int tile_capacity,
    int32_t* __restrict__ num_tiles) {
  __shared__ int s_cnt[kMaxExperts];
  __shared__ int s_off[kMaxExperts + 1];
  __shared__ int s_tile_off[kMaxExperts + 1];
  pdl_wait();
  pdl_launch();
  const int e = threadIdx.x;
  int total = 0;
  if (e < E) {
    for (int b = 0; b < nblk; ++b) {
      const int c = block_counts[static_cast<size_t>(b) * E + e];
      block_counts[static_cast<size_t>(b) * E + e] = total;
      total += c;
    }
    s_cnt[e] = total;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int off = 0, toff = 0;
    for (int i = 0; i < E; ++i) {
      s_off[i] = off;
      s_tile_off[i] = toff;
      off += s_cnt[i];
      toff += (s_cnt[i] + tile_rows - 1) / tile_rows;
    }
    s_off[E] = off;
    s_tile_off[E] = toff;
    const int shared_tiles = (sh_hi - sh_lo + tile_rows - 1) / tile_rows;
    const int n = toff + shared_tiles;
    *num_tiles = n <= tile_capacity ? n : 0;   // capacity is sized for the worst case; 0 = refuse
  }
  __syncthreads();
  if (e < E) expert_offsets[e] = s_off[e];
  if (e == 0) expert_offsets[E] = s_off[E];
  if (s_tile_off[E] + (sh_hi - sh_lo + tile_rows - 1) / tile_rows > tile_capacity) return;
  if (e < E) {
    emit_tiles(tiles, s_tile_off[e], e - e_lo, s_off[e], s_cnt[e], tile_rows);
  }
  if (threadIdx.x == blockDim.x - 1)
    emit_tiles(tiles, s_tile_off[E], e_hi - e_lo, T * k + sh_lo, sh_hi - sh_lo, tile_rows);
}

__global__ void __launch_bounds__(kTokPerBlock) permute_scatter_kernel(
    const int32_t* __restrict__ idx, const float* __restrict__ w, int T, int k, int E,
    const int32_t* __restrict__ local_rank, const int32_t* __restrict__ block_base,
    const int32_t* __restrict__ expert_offsets, int32_t* __restrict__ row_token,
    float* __restrict__ row_weight, int32_t* __restrict__ slot_row) {
  pdl_wait();
  pdl_launch();
  const int t = blockIdx.x * kTokPerBlock + threadIdx.x;
  if (t >= T) return;
  for (int s = 0; s < k; ++s) {
    const size_t i = static_cast<size_t>(t) * k + s;
    const int r = local_rank[i];
    int row = -1;
    if (r >= 0) {
      const int e = idx[i];
      row = expert_offsets[e] + block_base[static_cast<size_t>(blockIdx.x) * E + e] + r;
      row_token[row] = t;
      row_weight[row] = w[i];
    }
    slot_row[i] = row;
  }
}

// Decode-size batches (T <= 256): the three phases above in ONE block, thread t owns token t.
__global__ void __launch_bounds__(kTokPerBlock) permute_small_kernel(
    const int32_t* __restrict__ idx, const float* __restrict__ w, int T, int k, int E, int e_lo,
    int e_hi, int tile_rows, int sh_lo, int sh_hi, int32_t* __restrict__ expert_offsets,
    int32_t* __restrict__ row_token, float* __restrict__ row_weight, int32_t* __restrict__ slot_row,
    int4* __restrict__ tiles, int tile_capacity, int32_t* __restrict__ num_tiles) {
  __shared__ PermuteSmallSmem sm;
  for (int i = threadIdx.x; i < E * kWordsPerExpert; i += kTokPerBlock) sm.bitmap[i] = 0u;
  pdl_wait();
  pdl_launch();
  __syncthreads();
  permute_small_body(sm, idx, w, T, k, E, e_lo, e_hi, tile_rows, sh_lo, sh_hi, expert_offsets, row_token,
                     row_weight, slot_row, tiles, tile_capacity, num_tiles);
}

// x_rows[r] = x[row_token[r]] (routed), x_rows[T*k + t] = x[t] (shared). One warp per row,
// 128-bit copies.
template <int kBytesPerElem>
__global__ void __launch_bounds__(256) gather_rows_kernel(const uint4* __restrict__ x,
                                                          const int32_t* __restrict__ row_token,
                                                          const int32_t* __restrict__ expert_offsets,
                                                          uint4* __restrict__ x_rows, int T, int k,
                                                          int E, int vec_per_row, int sh_lo, int sh_hi) {
  pdl_wait();
  pdl_launch();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + warp;
  const int n_routed = expert_offsets[E];
  const int Tk = T * k;
  int src;
  if (row < Tk) {
    if (row >= n_routed) return;
    src = row_token[row];
  } else {
    src = row - Tk;
    if (src < sh_lo || src >= sh_hi) return;
  }
  const uint4* s = x + static_cast<size_t>(src) * vec_per_row;
  uint4* d = x_rows + static_cast<size_t>(row) * vec_per_row;
  for (int c = lane; c < vec_per_row; c += 32) d[c] = __ldg(s + c);
}

}  // namespace glm4moe

using namespace glm4moe;

extern "C" size_t glm4moe_permute_scratch_elems(int T, int E) {
  const size_t nblk = (static_cast<size_t>(T) + kTokPerBlock - 1) / kTokPerBlock;
  return (nblk + 1) * static_cast<size_t>(E) + 8;
}

extern "C" int32_t glm4moe_tile_capacity(int T, int k, int E, int tile_rows) {
  if (tile_rows <= 0) return 0;
  // every expert may leave one ragged tile; the shared expert adds ceil(T/tile_rows)
  return E + (T * k) / tile_rows + (T + tile_rows - 1) / tile_rows + 1;
}

extern "C" int glm4moe_permute(const int32_t* idx, const float* w, int T, int k, int E, int e_lo,
                               int e_hi, int tile_rows, int shared_lo, int shared_hi, int32_t* expert_offsets,
                               int32_t* row_token, float* row_weight, int32_t* slot_row,
                               int32_t* tiles, int32_t tile_capacity, int32_t* num_tiles,
                               int32_t* scratch, void* stream) {
  if (!idx || !w || !expert_offsets || !row_token || !row_weight || !slot_row || !tiles ||
      !num_tiles || !scratch)
    return GLM4MOE_ERR_ARG;
  if (T < 0 || k <= 0 || E <= 0 || E > kMaxExperts || tile_rows <= 0 || e_lo < 0 || e_hi > E ||
      e_lo > e_hi || shared_lo < 0 || shared_hi > T || shared_lo > shared_hi)
    return GLM4MOE_ERR_ARG;
  if (tile_capacity < glm4moe_tile_capacity(T, k, E, tile_rows)) return GLM4MOE_ERR_WORKSPACE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nblk = (T + kTokPerBlock - 1) / kTokPerBlock;
  if (T > 0 && T <= kTokPerBlock) {
    GLM4MOE_CUDA_CHECK(launch_pdl(permute_small_kernel, dim3(1), dim3(kTokPerBlock), 0, st, idx, w, T, k, E, e_lo,
                                  e_hi, tile_rows, shared_lo, shared_hi, expert_offsets, row_token, row_weight, slot_row,
                                  reinterpret_cast<int4*>(tiles), tile_capacity, num_tiles));
    return GLM4MOE_OK;
  }
  int32_t* block_counts = scratch;                       // [nblk][E]
  // local ranks live in slot_row until the scatter overwrites them in place
  int32_t* local_rank = slot_row;
  if (nblk > 0) {
    GLM4MOE_CUDA_CHECK(launch_pdl(permute_rank_kernel, dim3(nblk), dim3(kTokPerBlock), 0, st, idx, T, k, E, e_lo,
                                  e_hi, local_rank, block_counts));
  }
  GLM4MOE_CUDA_CHECK(launch_pdl(permute_scan_kernel, dim3(1), dim3(256), 0, st, block_counts, nblk, E, T, k, e_lo,
                                e_hi, tile_rows, shared_lo, shared_hi, expert_offsets, reinterpret_cast<int4*>(tiles),
                                tile_capacity, num_tiles));
  if (nblk > 0) {
    GLM4MOE_CUDA_CHECK(launch_pdl(permute_scatter_kernel, dim3(nblk), dim3(kTokPerBlock), 0, st, idx, w, T, k, E,
                                  static_cast<const int32_t*>(local_rank), static_cast<const int32_t*>(block_counts),
                                  static_cast<const int32_t*>(expert_offsets), row_token, row_weight, slot_row));
  }
  return GLM4MOE_OK;
}

// ---- glm4moe_permute_static: the same permutation for callers that cannot hand over scratch memory ------------
// No counterpart in the reference.  Some hosts of this library have no way to obtain device memory of their own:
// a plugin that is only ever given the input and output buffers of its op by the runtime that loaded it, a
// component that must stay inside a memory budget fixed before the library was loaded (every byte of the device
// already pledged to a KV cache or another tenant), or a call site inside a region where cudaMalloc is forbidden.
// For them the [blocks x experts] count table of the three-kernel path lives in the library's own image: a
// file-scope device array, created by the CUDA runtime when the module is loaded on a GPU and kept until the
// process exits (about 65 KB on each GPU that loads the library).
namespace glm4moe {
constexpr int kPermuteStaticMaxTokens = GLM4MOE_PERMUTE_STATIC_MAX_TOKENS;
constexpr size_t kPermuteStaticElems =
    (static_cast<size_t>(kPermuteStaticMaxTokens / kTokPerBlock) + 1) * kMaxExperts + 8;
static_assert(kPermuteStaticMaxTokens % kTokPerBlock == 0, "the limit is a whole number of rank blocks");
// In host code this name is only a handle for cudaGetSymbolAddress: it is NOT a device address and must never be
// passed to a kernel.  Device code may name it directly.
static __device__ __align__(256) int32_t g_permute_static_scratch[kPermuteStaticElems];

constexpr int kMaxDevices = 64;
static std::atomic<int32_t*> g_permute_static_ptr[kMaxDevices];   // device address of the array, per GPU
}  // namespace glm4moe

extern "C" int glm4moe_permute_static(const int32_t* idx, const float* w, int T, int k, int E, int e_lo, int e_hi,
                                      int tile_rows, int shared_lo, int shared_hi, int32_t* expert_offsets,
                                      int32_t* row_token, float* row_weight, int32_t* slot_row, int32_t* tiles,
                                      int32_t tile_capacity, int32_t* num_tiles, void* stream) {
  if (T > kPermuteStaticMaxTokens) return GLM4MOE_ERR_WORKSPACE;   // the array holds no more; nothing is launched
  if (T < 0 || E <= 0 || E > kMaxExperts) return GLM4MOE_ERR_ARG;
  if (glm4moe_permute_scratch_elems(T, E) > kPermuteStaticElems) return GLM4MOE_ERR_WORKSPACE;
  int dev = -1;
  GLM4MOE_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return GLM4MOE_ERR_UNSUPPORTED;
  int32_t* scratch = g_permute_static_ptr[dev].load(std::memory_order_acquire);
  if (scratch == nullptr) {
    // Set-up, once per GPU, with that GPU current: ask the runtime where this GPU's copy of the array lives.  (Two
    // threads racing here get the same answer.)  This may load the module on the GPU, so it belongs before any
    // stream capture.
    void* p = nullptr;
    GLM4MOE_CUDA_CHECK(cudaGetSymbolAddress(&p, g_permute_static_scratch));
    if (p == nullptr) return GLM4MOE_ERR_DEVICE;
    scratch = static_cast<int32_t*>(p);
    g_permute_static_ptr[dev].store(scratch, std::memory_order_release);
  }
  // Nothing in the table has to start at zero: permute_rank_kernel writes every [block][expert] count this call
  // reads before permute_scan_kernel reads it, so what the previous call left behind is never observed.
  return glm4moe_permute(idx, w, T, k, E, e_lo, e_hi, tile_rows, shared_lo, shared_hi, expert_offsets, row_token,
                         row_weight, slot_row, tiles, tile_capacity, num_tiles, scratch, stream);
}

extern "C" int glm4moe_gather_rows(const void* x, const int32_t* row_token,
                                   const int32_t* expert_offsets, void* x_rows, int T, int k, int E,
                                   int H, int shared_lo, int shared_hi, int dtype, void* stream) {
  if (!x || !row_token || !expert_offsets || !x_rows) return GLM4MOE_ERR_ARG;
  const int es = dtype_size(dtype);
  if ((static_cast<size_t>(H) * es) % 16 != 0) return GLM4MOE_ERR_UNSUPPORTED;
  if (T == 0) return GLM4MOE_OK;
  const int rows = T * k + shared_hi;
  const int vec_per_row = H * es / 16;
  GLM4MOE_CUDA_CHECK(launch_pdl(gather_rows_kernel<2>, dim3((rows + 7) / 8), dim3(256), 0,
                                static_cast<cudaStream_t>(stream), static_cast<const uint4*>(x), row_token,
                                expert_offsets, static_cast<uint4*>(x_rows), T, k, E, vec_per_row, shared_lo, shared_hi));
  return GLM4MOE_OK;
}

// Expert-parallel combine over NVLink peer memory: the weighted un-permute / accumulate of the
// reference (moe.py:149-151,171) and the cross-GPU reduction in ONE kernel, no NCCL call.
//
// Layout "p2p" (distributed.py): tokens are replicated, rank r owns experts [r*E/R, (r+1)*E/R) and
// has just written the weighted expert outputs of ITS experts as rows of its own workspace
// (y[row, H], row = slot_row[t, j]; shared-expert rows T*k + t on the rank that owns token t's
// slice).  Every rank's workspace is one cudaMalloc'd region mapped into all peers (cudaIpc), with
// identical layout.  For each token this kernel walks the k slots in ascending expert order (=
// ascending owner rank), reads each row from its owner's memory (local or over NVLink), adds the
// shared row, accumulates in fp32 and rounds once: the same order and rounding as the single-GPU
// combine_kernel, so an expert-parallel forward is BIT-IDENTICAL to the single-GPU forward and
// every rank ends with the full replicated [T, H] (the all-reduce is implicit in the reads).
//
// Cross-GPU ordering (epoch e = calls so far + 1, kept in the workspace, CUDA-graph-replay safe):
//   ready: block 0 thread p stores e into peer p's ready[me] (st.release.sys); every block spins on
//          its own ready[0..R) >= e before the first peer read  -> the peer's rows are complete
//          (they were written by the peer's previous kernel in stream order).
//   done:  the last block of the grid (ticket) stores e into every peer's done[me] and spins on its
//          own done[0..R) >= e before the grid retires -> nobody overwrites y / slot_row (next
//          forward) while a peer still reads them.
//
// Decode fast path (T <= 32, `a.fast`): ONE rendezvous per step.  The rows and slot_row of call e live
// in buffers selected by the parity of e, the fused expert kernel itself stores the ready flags the moment
// its last down item has landed (decode_fused.cu), and the done round is gone: rank r overwrites the
// parity-p buffers again in call e + 2, after its combine of call e + 1 has seen every peer's
// ready(e + 1) — which a peer only sends after its own combine of call e (the last reader of my
// parity-p buffers) has completed.  Calls of the other modes use the plain buffers and keep their done
// round, so the modes can be mixed freely.
//
// Every cross-GPU wait is bounded (~18 s of SM clocks): a rank that never arrives, or arrives in
// another mode (`glm4moe_set_ep_sliced` set differently), records what it waited for in the process's pinned host
// error record (read it with glm4moe_ep_last_error: it survives the trap) and traps instead of hanging the GPU.
// The limit is glm4moe_ep_set_timeout_ms / GLM4MOE_EP_TIMEOUT_MS.
#include "common.cuh"
#include "ep_combine.cuh"

namespace glm4moe {

namespace {

__device__ __forceinline__ void st_release_sys(int32_t* p, int32_t v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int32_t ld_acquire_sys(const int32_t* p) {
  int32_t v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __noinline__ void ep_give_up(const EpWatch& w, int32_t* ctl, int code, int e, int observed, int peer) {
  ctl[2] = code;
  if (w.rec != nullptr) {
    // plain stores over PCIe (no reliance on host atomics): several threads may give up at once and each describes
    // a real failure, so whichever lands last is a valid report.  `code` goes last.
    volatile int32_t* r = &w.rec->code;
    r[1] = w.rank;
    r[2] = e;
    r[3] = observed;
    r[4] = w.site;
    r[5] = peer;
    __threadfence_system();
    r[0] = code;
  }
  __threadfence_system();
  __trap();
}
// wait until the flag reaches call `e` (flag = epoch * 4 + mode); wrap-safe compare
__device__ __forceinline__ void wait_flag(const int32_t* flag, int e, int mode, int32_t* ctl, const EpWatch& w,
                                          int peer) {
  const int target = e * 4;
  const long long t0 = clock64();
  int v;
  while ((v = ld_acquire_sys(flag)) - target < 0) {
    if (clock64() - t0 > w.spin_limit) ep_give_up(w, ctl, kEpErrTimeout, e, v, peer);
  }
  if ((v >> 2) == e && (v & 3) != mode) ep_give_up(w, ctl, kEpErrModeMismatch, e, v, peer);
}
__device__ __forceinline__ uint4 ld_peer16(const void* p) {   // L1-bypassing 128-bit load (peer or local)
  uint4 r;
  asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ int32_t ld_peer_s32(const int32_t* p) {
  int32_t v;
  asm volatile("ld.global.cg.s32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}

}  // namespace

template <typename T>
__global__ void __launch_bounds__(kEpThreads) ep_combine_p2p_kernel(const EpCombineArgs a, T* __restrict__ out) {
  constexpr int V = Dt<T>::kVec;   // 8 two-byte elements per 16-byte vector
  __shared__ int32_t s_row[kEpMaxTopK + 1];
  __shared__ int32_t s_owner[kEpMaxTopK + 1];
  __shared__ int s_last;
  pdl_wait();      // the local rows are the direct predecessor's output
  pdl_launch();
  unsigned char* my_ws = a.ws[a.rank];
  int32_t* ctl = reinterpret_cast<int32_t*>(my_ws + a.off_ctl);   // [0] epoch, [1] blocks done, [16..] ready, [32..] done
  const int e = *reinterpret_cast<volatile int32_t*>(ctl) + 1;
  const int mode = a.fast ? kEpModeFast : kEpModePull;
  // ---- ready: tell every peer our rows exist (fast path: the expert kernel already did), wait until theirs do ----
  if (!a.fast && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x < a.world) {
    __threadfence_system();
    int32_t* peer_ctl = reinterpret_cast<int32_t*>(a.ws[threadIdx.x] + a.off_ctl);
    st_release_sys(peer_ctl + 16 + a.rank, e * 4 + mode);
  }
  if (threadIdx.x < a.world) wait_flag(ctl + 16 + threadIdx.x, e, mode, ctl, a.watch, threadIdx.x);
  __syncthreads();

  // ---- reduce: block (x = column chunk, y strides over tokens) ----
  const int col = (blockIdx.x * kEpThreads + threadIdx.x) * V;
  const int32_t* my_idx = reinterpret_cast<const int32_t*>(my_ws + a.off_idx);
  const size_t off_y = a.fast ? a.off_y_alt[e & 1] : a.off_y;
  const size_t off_slot_row = a.fast ? a.off_slot_row_alt[e & 1] : a.off_slot_row;
  for (int t = blockIdx.y; t < a.tokens; t += gridDim.y) {
    // slot j -> (owner rank, row on the owner); the shared expert's row lives on the rank owning t's slice
    if (threadIdx.x <= a.k) {
      int owner, row;
      if (threadIdx.x < a.k) {
        const int ex = my_idx[static_cast<size_t>(t) * a.k + threadIdx.x];
        owner = ex / a.experts_per_rank;
        row = ld_peer_s32(reinterpret_cast<const int32_t*>(a.ws[owner] + off_slot_row) +
                          static_cast<size_t>(t) * a.k + threadIdx.x);
      } else {
        owner = a.with_shared ? min(t / a.shared_per_rank, a.world - 1) : 0;
        row = a.with_shared ? a.tokens * a.k + t : -1;
      }
      s_owner[threadIdx.x] = owner;
      s_row[threadIdx.x] = row;
    }
    __syncthreads();
    if (col < a.H) {
      float acc[V];
#pragma unroll
      for (int i = 0; i < V; ++i) acc[i] = 0.f;
      uint4 raw[kEpMaxTopK + 1];
      // all loads first (independent NVLink round trips in flight), then the ordered fp32 sum
#pragma unroll
      for (int j = 0; j <= kEpMaxTopK; ++j) {
        if (j <= a.k && s_row[j] >= 0)
          raw[j] = ld_peer16(a.ws[s_owner[j]] + off_y +
                             (static_cast<size_t>(s_row[j]) * a.H + col) * sizeof(T));
      }
#pragma unroll
      for (int j = 0; j <= kEpMaxTopK; ++j) {
        if (j <= a.k && s_row[j] >= 0) {
          float v[V];
          unpack16<T>(raw[j], v);
#pragma unroll
          for (int i = 0; i < V; ++i) acc[i] += v[i];
        }
      }
      __align__(16) T o[V];
#pragma unroll
      for (int i = 0; i < V; ++i) o[i] = Dt<T>::from_f(acc[i]);
      *reinterpret_cast<uint4*>(out + static_cast<size_t>(t) * a.H + col) = *reinterpret_cast<uint4*>(o);
    }
    __syncthreads();
  }

  // ---- done: the last block tells every peer we no longer read, and waits for theirs ----
  if (threadIdx.x == 0) {
    __threadfence();
    const int total = static_cast<int>(gridDim.x * gridDim.y);
    s_last = atomicAdd(ctl + 1, 1) == total - 1;
  }
  __syncthreads();
  if (!s_last) return;
  if (!a.fast && threadIdx.x < a.world) {
    int32_t* peer_ctl = reinterpret_cast<int32_t*>(a.ws[threadIdx.x] + a.off_ctl);
    st_release_sys(peer_ctl + 32 + a.rank, e * 4 + mode);
    wait_flag(ctl + 32 + threadIdx.x, e, mode, ctl, a.watch, threadIdx.x);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    ctl[1] = 0;
    *reinterpret_cast<volatile int32_t*>(ctl) = e;
    __threadfence();
  }
}

// Sliced variant for large batches (prefill): "reduce-scatter + all-gather" in one kernel.  Rank r
// reduces only ITS token slice [r*ceil(T/R), ...) exactly as above (same order, same single
// rounding: still bit-identical to one GPU) and stores each finished row into `out` locally and into
// the staging region of every peer's workspace with 128-bit NVLink stores; once every peer's done
// flag has arrived, the other slices are copied from the local staging region into `out`.  NVLink
// bytes per rank drop from T*(k+1)*H*(R-1)/R (every rank pulls every row) to
// T/R*(k+1)*H*(R-1)/R pulled + T/R*H*(R-1) pushed: ~4.7x fewer at R = 8, k = 8.
//   done (this variant): the last block to finish its slice stores e into every peer's done[me]
//          (after a system-scope fence, so the staged rows are visible first); EVERY block then spins
//          on its own done[0..R) >= e.  That one round is both "all slices have landed here" and
//          "no peer still reads my rows".  A peer cannot start overwriting the staging region for
//          call n+1 before it has seen my ready flag of call n+1, which my next kernel sends after
//          this one retired.  All blocks must be co-resident (the launch caps the grid accordingly).
template <typename T>
__global__ void __launch_bounds__(kEpThreads) ep_combine_sliced_kernel(const EpCombineArgs a, T* __restrict__ out) {
  constexpr int V = Dt<T>::kVec;
  __shared__ int32_t s_row[kEpMaxTopK + 1];
  __shared__ int32_t s_owner[kEpMaxTopK + 1];
  __shared__ int s_last;
  pdl_wait();
  pdl_launch();
  unsigned char* my_ws = a.ws[a.rank];
  int32_t* ctl = reinterpret_cast<int32_t*>(my_ws + a.off_ctl);
  const int e = *reinterpret_cast<volatile int32_t*>(ctl) + 1;
  if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x < a.world) {
    __threadfence_system();
    int32_t* peer_ctl = reinterpret_cast<int32_t*>(a.ws[threadIdx.x] + a.off_ctl);
    st_release_sys(peer_ctl + 16 + a.rank, e * 4 + kEpModeSliced);
  }
  if (threadIdx.x < a.world) wait_flag(ctl + 16 + threadIdx.x, e, kEpModeSliced, ctl, a.watch, threadIdx.x);
  __syncthreads();

  const int col = (blockIdx.x * kEpThreads + threadIdx.x) * V;
  const int32_t* my_idx = reinterpret_cast<const int32_t*>(my_ws + a.off_idx);
  const int lo = min(a.rank * a.shared_per_rank, a.tokens);
  const int hi = min(lo + a.shared_per_rank, a.tokens);
  for (int t = lo + blockIdx.y; t < hi; t += gridDim.y) {
    if (threadIdx.x <= a.k) {
      int owner, row;
      if (threadIdx.x < a.k) {
        const int ex = my_idx[static_cast<size_t>(t) * a.k + threadIdx.x];
        owner = ex / a.experts_per_rank;
        row = ld_peer_s32(reinterpret_cast<const int32_t*>(a.ws[owner] + a.off_slot_row) +
                          static_cast<size_t>(t) * a.k + threadIdx.x);
      } else {
        owner = a.with_shared ? a.rank : 0;       // the shared expert ran on the slice owner: local
        row = a.with_shared ? a.tokens * a.k + t : -1;
      }
      s_owner[threadIdx.x] = owner;
      s_row[threadIdx.x] = row;
    }
    __syncthreads();
    if (col < a.H) {
      float acc[V];
#pragma unroll
      for (int i = 0; i < V; ++i) acc[i] = 0.f;
      uint4 raw[kEpMaxTopK + 1];
#pragma unroll
      for (int j = 0; j <= kEpMaxTopK; ++j) {
        if (j <= a.k && s_row[j] >= 0)
          raw[j] = ld_peer16(a.ws[s_owner[j]] + a.off_y +
                             (static_cast<size_t>(s_row[j]) * a.H + col) * sizeof(T));
      }
#pragma unroll
      for (int j = 0; j <= kEpMaxTopK; ++j) {
        if (j <= a.k && s_row[j] >= 0) {
          float v[V];
          unpack16<T>(raw[j], v);
#pragma unroll
          for (int i = 0; i < V; ++i) acc[i] += v[i];
        }
      }
      __align__(16) T o[V];
#pragma unroll
      for (int i = 0; i < V; ++i) o[i] = Dt<T>::from_f(acc[i]);
      const uint4 ov = *reinterpret_cast<uint4*>(o);
      const size_t elem = static_cast<size_t>(t) * a.H + col;
      *reinterpret_cast<uint4*>(out + elem) = ov;
      for (int r = 0; r < a.world; ++r)
        if (r != a.rank) *reinterpret_cast<uint4*>(a.ws[r] + a.off_stage + elem * sizeof(T)) = ov;
    }
    __syncthreads();
  }

  // ---- done: slice delivered everywhere (and no more reads of peer rows) ----
  if (threadIdx.x == 0) {
    __threadfence_system();
    const int total = static_cast<int>(gridDim.x * gridDim.y);
    s_last = atomicAdd(ctl + 1, 1) == total - 1;
  }
  __syncthreads();
  if (s_last && threadIdx.x < a.world) {
    __threadfence_system();
    int32_t* peer_ctl = reinterpret_cast<int32_t*>(a.ws[threadIdx.x] + a.off_ctl);
    st_release_sys(peer_ctl + 32 + a.rank, e * 4 + kEpModeSliced);
  }
  if (threadIdx.x < a.world) wait_flag(ctl + 32 + threadIdx.x, e, kEpModeSliced, ctl, a.watch, threadIdx.x);
  __syncthreads();
  // ---- the other ranks' slices: local staging region -> out ----
  if (col < a.H) {
    const unsigned char* stage = my_ws + a.off_stage;
    for (int t = blockIdx.y; t < a.tokens; t += gridDim.y) {
      if (t >= lo && t < hi) continue;
      const size_t elem = static_cast<size_t>(t) * a.H + col;
      *reinterpret_cast<uint4*>(out + elem) = ld_peer16(stage + elem * sizeof(T));
    }
  }
  if (s_last) {
    __syncthreads();
    if (threadIdx.x == 0) {
      ctl[1] = 0;
      *reinterpret_cast<volatile int32_t*>(ctl) = e;
      __threadfence();
    }
  }
}

int ep_combine_p2p_launch(const EpCombineArgs& args, void* out, int dtype, void* stream) {
  EpCombineArgs a = args;
  a.watch = ep_watch(a.rank, kEpSiteCombine);
  if (a.world < 1 || a.world > kEpMaxRanks || a.k > kEpMaxTopK || a.rank < 0 || a.rank >= a.world)
    return GLM4MOE_ERR_ARG;
  if (dtype != GLM4MOE_BF16 && dtype != GLM4MOE_F16) return GLM4MOE_ERR_UNSUPPORTED;
  if (a.H % 8 != 0 || a.tokens <= 0) return GLM4MOE_ERR_ARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int gx = (a.H / 8 + kEpThreads - 1) / kEpThreads;
  int gy = a.tokens < 128 ? a.tokens : 128;
  if (a.fast && a.sliced) return GLM4MOE_ERR_ARG;
  if (a.sliced) {
    // every block waits for the peers inside the kernel: keep the whole grid co-resident
    // (128-thread blocks without shared-memory demand: >= 4 per SM on every register budget)
    int dev = 0, sms = 0;
    GLM4MOE_CUDA_CHECK(cudaGetDevice(&dev));
    GLM4MOE_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int cap = sms * 4 / gx;
    if (cap < 1) return GLM4MOE_ERR_UNSUPPORTED;
    if (gy > cap) gy = cap;
    dim3 grid(gx, gy);
    if (dtype == GLM4MOE_BF16)
      GLM4MOE_CUDA_CHECK(launch_pdl(ep_combine_sliced_kernel<__nv_bfloat16>, grid, dim3(kEpThreads), 0, st, a,
                                    static_cast<__nv_bfloat16*>(out)));
    else
      GLM4MOE_CUDA_CHECK(launch_pdl(ep_combine_sliced_kernel<__half>, grid, dim3(kEpThreads), 0, st, a,
                                    static_cast<__half*>(out)));
    return GLM4MOE_OK;
  }
  dim3 grid(gx, gy);
  if (dtype == GLM4MOE_BF16)
    GLM4MOE_CUDA_CHECK(launch_pdl(ep_combine_p2p_kernel<__nv_bfloat16>, grid, dim3(kEpThreads), 0, st, a,
                                  static_cast<__nv_bfloat16*>(out)));
  else
    GLM4MOE_CUDA_CHECK(launch_pdl(ep_combine_p2p_kernel<__half>, grid, dim3(kEpThreads), 0, st, a,
                                  static_cast<__half*>(out)));
  return GLM4MOE_OK;
}

}  // namespace glm4moe

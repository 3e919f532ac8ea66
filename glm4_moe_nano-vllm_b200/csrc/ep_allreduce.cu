// Tensor-parallel boundary in front of the sparse block, fused over NVLink peer memory:
//
//   reference, per decoder layer (decode_layer.py:96-99):
//     hidden = o_proj(attn)            RowParallelLinear.forward: F.linear + dist.all_reduce (linear.py:151-155)
//     hidden, residual = post_attention_layernorm(hidden, residual)      (layernorm.py:29-40, ~8 torch launches)
//     hidden = mlp(hidden)
//
//   here: every rank hands its o_proj PARTIAL sums [T, H] to ONE kernel that is the all-reduce, the residual add
//   and the RMSNorm at once; no NCCL call, CUDA-graph capturable.
//
//   y            = round_dtype( sum_r partial_r )        fp32, ranks in ascending order (identical bits on every rank:
//                                                        the routing that follows must agree everywhere)
//   x32          = float(y) + float(residual)            residual_out = round(x32)
//   out          = round(round(x32 * rsqrt(mean(x32^2) + eps)) * weight)
//   i.e. exactly `add_rmsnorm_kernel` (rmsnorm.cu, same thread mapping and summation order: bit-identical) applied
//   to an all-reduce whose sum is taken in fp32 with one rounding.  weight == nullptr: out = y (a plain all-reduce).
//
// Two forms, chosen by size (same choice on every rank: it depends on T, H only):
//   one-shot  (T*H*es <= 1 MiB, T <= 1024): every rank PUSHES its partial rows into an inbox on every peer (posted
//             NVLink writes, no read round trip) and releases a flag PER ROW; the block that owns row t waits for
//             the peers' flags of row t only, then reduces and normalises it.  (R-1)*T*H*es bytes out per rank.
//   two-shot  (prefill): reduce-scatter -> token shard -> all-gather, in one launch.  Rank r pushes, for every
//             peer s, its partial rows of s's token slice into s's inbox; after the first flag round rank s sums
//             ITS ceil(T/R) tokens and pushes y to every peer; after the second flag round every rank holds y of
//             every token and applies the residual add + norm locally (HBM traffic, which is plentiful, instead
//             of a second NVLink plane).  2*(R-1)/R*T*H*es bytes out per rank, the traffic of a ring all-reduce.
//
// Ordering (epoch e = calls completed + 1, kept in the comm buffer: graph-replay safe).  Inboxes are selected by the
// parity of e and there is NO "done" round: a rank overwrites the parity-p inboxes of its peers again in call
// e + 2, after it has seen every peer's flag of call e + 1 — which a peer sends only after its own call e has
// retired (stream order), i.e. after its last read of the parity-p inbox.  Flags carry epoch*4 + mode; every
// spin is bounded and traps with an error code (ranks that disagree on the form or on call counts fail loudly).
// The blocks of a two-shot launch wait for the peers' grid-wide flags inside the kernel: its grid is capped so
// that it is co-resident.
#include "common.cuh"
#include "ep_allreduce.cuh"

namespace glm4moe {

namespace {

constexpr int kArMaxVec = 4;     // 16-byte vectors per thread: H <= 256 * 4 * 8 = 8192 two-byte elements
static_assert(kEpArThreads * kArMaxVec * 8 == GLM4MOE_EP_COMM_MAX_HIDDEN, "glm4moe_ep_comm_bytes refuses what the kernels cannot hold");

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
// Wait until `flag` carries call e (or e + 1: a peer can be at most one call ahead, see the header).  Accepting only
// that window keeps a word that was last written ~2^29 calls ago (a row index not used since) from looking "ahead"
// once the 32-bit epoch arithmetic wraps.
__device__ __forceinline__ void wait_flag(const int32_t* flag, int e, int mode, int32_t* ctl, const EpWatch& w,
                                          int peer) {
  const int target = e * 4;
  const long long t0 = clock64();
  int d;
  while ((d = ld_acquire_sys(flag) - target) < 0 || d >= 8) {
    if (clock64() - t0 > w.spin_limit) ep_give_up(w, ctl, kEpErrTimeout, e, d + target, peer);
  }
  if (d < 4 && d != mode) ep_give_up(w, ctl, kEpErrModeMismatch, e, d + target, peer);
}
// L1-bypassing 128-bit load: inbox rows are written by peers over NVLink (L1 may hold the rows of call e - 2)
__device__ __forceinline__ uint4 ld_cg16(const void* p) {
  uint4 r;
  asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

// One ticket per block; true in the LAST block to arrive.  Release pattern of ep_combine.cu (and of NCCL's
// primitives): block barrier, then ONE thread fences at system scope before it publishes; the barrier puts every
// thread's stores before that fence in causality order, so 2-8 threads fence per block instead of 256 (the
// all-threads form cost ~4 us of a 19 us one-shot call at T = 32).  The caller's flag writers fence again after
// the second barrier, which orders the other blocks' stores (observed through the ticket) before the flag.
__device__ __forceinline__ bool block_ticket(int32_t* counter, int* s_flag) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    *s_flag = atomicAdd(counter, 1) == static_cast<int>(gridDim.x) - 1;
  }
  __syncthreads();
  return *s_flag != 0;
}

// ---- one token row, in two steps -------------------------------------------------------------------------
// Thread mapping of add_rmsnorm_kernel (rmsnorm.cu): thread `tid` owns the 16-byte vectors tid + i * 256.
template <typename T>
struct RowRegs {
  float v[kArMaxVec][Dt<T>::kVec];
};

// (1) y = round(sum_r partial_r): fp32, rank order; own partial read in place, the others from the local inbox.
//     PUSH: y also goes to every peer's all-gather inbox (two-shot).
template <typename T, bool PUSH>
__device__ __forceinline__ void sum_row(const EpArArgs& a, int t, const unsigned char* inbox, size_t slot_bytes,
                                        int inbox_row, int p, RowRegs<T>& R) {
  constexpr int V = Dt<T>::kVec;
  const int nvec = a.H / V;
  const size_t row_off = static_cast<size_t>(t) * a.H;
#pragma unroll
  for (int i = 0; i < kArMaxVec; ++i) {
    const int c = threadIdx.x + i * kEpArThreads;
    if (c < nvec) {
      uint4 raw[kEpMaxRanks];
#pragma unroll
      for (int r = 0; r < kEpMaxRanks; ++r) {
        if (r < a.world)
          raw[r] = r == a.rank
                       ? ld_cg16(reinterpret_cast<const uint4*>(static_cast<const T*>(a.partial) + row_off) + c)
                       : ld_cg16(reinterpret_cast<const uint4*>(inbox + static_cast<size_t>(r) * slot_bytes +
                                                                static_cast<size_t>(inbox_row) * a.H * sizeof(T)) + c);
      }
      float acc[V];
#pragma unroll
      for (int q = 0; q < V; ++q) acc[q] = 0.f;
#pragma unroll
      for (int r = 0; r < kEpMaxRanks; ++r) {
        if (r < a.world) {
          float f[V];
          unpack16<T>(raw[r], f);
#pragma unroll
          for (int q = 0; q < V; ++q) acc[q] += f[q];
        }
      }
      __align__(16) T o[V];
#pragma unroll
      for (int q = 0; q < V; ++q) {
        o[q] = Dt<T>::from_f(acc[q]);           // the all-reduce result in the model dtype
        R.v[i][q] = Dt<T>::to_f(o[q]);
      }
      if (PUSH) {
        const uint4 ov = *reinterpret_cast<uint4*>(o);
        for (int r = 0; r < a.world; ++r)
          if (r != a.rank) reinterpret_cast<uint4*>(a.comm[r] + a.L.ag[p] + row_off * sizeof(T))[c] = ov;
      }
    }
  }
}

// (1') y of a row another rank reduced: from the local all-gather inbox
template <typename T>
__device__ __forceinline__ void load_row(const EpArArgs& a, int t, const unsigned char* ag, RowRegs<T>& R) {
  constexpr int V = Dt<T>::kVec;
  const int nvec = a.H / V;
#pragma unroll
  for (int i = 0; i < kArMaxVec; ++i) {
    const int c = threadIdx.x + i * kEpArThreads;
    if (c < nvec)
      unpack16<T>(ld_cg16(reinterpret_cast<const uint4*>(ag + static_cast<size_t>(t) * a.H * sizeof(T)) + c), R.v[i]);
  }
}

// (2) residual add + RMSNorm of y, the arithmetic and summation order of add_rmsnorm_kernel; weight == nullptr:
//     out = y.  Contains block barriers: call it from uniform control flow.
template <typename T>
__device__ __forceinline__ void finish_row(const EpArArgs& a, int t, RowRegs<T>& R, float* s_part) {
  constexpr int V = Dt<T>::kVec;
  const int nvec = a.H / V;
  const size_t row_off = static_cast<size_t>(t) * a.H;
  const T* residual = static_cast<const T*>(a.residual);
  const T* weight = static_cast<const T*>(a.weight);
  T* out = static_cast<T*>(a.out);
  T* residual_out = static_cast<T*>(a.residual_out);
  if (weight == nullptr) {
#pragma unroll
    for (int i = 0; i < kArMaxVec; ++i) {
      const int c = threadIdx.x + i * kEpArThreads;
      if (c < nvec) {
        __align__(16) T o[V];
#pragma unroll
        for (int q = 0; q < V; ++q) o[q] = Dt<T>::from_f(R.v[i][q]);
        reinterpret_cast<uint4*>(out + row_off)[c] = *reinterpret_cast<uint4*>(o);
      }
    }
    return;
  }
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < kArMaxVec; ++i) {
    const int c = threadIdx.x + i * kEpArThreads;
    if (c < nvec) {
      if (residual != nullptr) {
        float rr[V];
        unpack16<T>(ld_cached16(reinterpret_cast<const uint4*>(residual + row_off) + c), rr);
#pragma unroll
        for (int q = 0; q < V; ++q) R.v[i][q] += rr[q];
        __align__(16) T o[V];
#pragma unroll
        for (int q = 0; q < V; ++q) o[q] = Dt<T>::from_f(R.v[i][q]);
        reinterpret_cast<uint4*>(residual_out + row_off)[c] = *reinterpret_cast<uint4*>(o);
      }
#pragma unroll
      for (int q = 0; q < V; ++q) ss = fmaf(R.v[i][q], R.v[i][q], ss);
    }
  }
  ss = warp_sum(ss);
  __syncthreads();                    // s_part of the previous row has been read by everyone
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = ss;
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int w = 0; w < kEpArThreads / 32; ++w) tot += s_part[w];
  const float rs = rsqrtf(tot / static_cast<float>(a.H) + a.eps);
#pragma unroll
  for (int i = 0; i < kArMaxVec; ++i) {
    const int c = threadIdx.x + i * kEpArThreads;
    if (c < nvec) {
      float g[V];
      unpack16<T>(ld_cached16(reinterpret_cast<const uint4*>(weight) + c), g);
      __align__(16) T o[V];
#pragma unroll
      for (int q = 0; q < V; ++q) o[q] = Dt<T>::from_f(Dt<T>::round(R.v[i][q] * rs) * g[q]);
      reinterpret_cast<uint4*>(out + row_off)[c] = *reinterpret_cast<uint4*>(o);
    }
  }
}

// retire a launch: the last block bumps the epoch and re-arms the tickets (no system-scope fence: nothing
// cross-GPU depends on it)
__device__ __forceinline__ void retire(int32_t* ctl, int e, int* s_flag) {
  __syncthreads();
  if (threadIdx.x == 0) {
    if (atomicAdd(ctl + 4, 1) == static_cast<int>(gridDim.x) - 1) {
      ctl[1] = 0;
      ctl[3] = 0;
      ctl[4] = 0;
      *reinterpret_cast<volatile int32_t*>(ctl) = e;
      __threadfence();
    }
  }
}

}  // namespace

// One-shot: ROW-granular hand-off.  The block that owns row t pushes its partial row to every peer and releases
// that row's flag there; it then waits only for the peers' flags of the same row.  No grid-wide ticket, no block
// depends on another block of its own grid (so the grid needs no co-residency), and the latency of a decode-size
// call is one push + one system-scope fence + one flag.  A flag word holds epoch*4 + mode of the last call that
// wrote the row; data live in the inbox of the call's parity, so a peer that is one call ahead passes the wait
// without having touched this call's rows (see the header comment for why it cannot be two ahead).
template <typename T>
__global__ void __launch_bounds__(kEpArThreads, 2) ep_ar_oneshot_kernel(const EpArArgs a) {
  constexpr int V = Dt<T>::kVec;
  __shared__ float s_part[kEpArThreads / 32];
  __shared__ int s_flag;
  pdl_wait();        // the partial sums are the direct predecessor's output
  pdl_launch();
  int32_t* ctl = reinterpret_cast<int32_t*>(a.comm[a.rank] + a.L.ctl);
  const int e = *reinterpret_cast<volatile int32_t*>(ctl) + 1;
  const int p = e & 1;
  const int nvec = a.H / V;
  if (a.world > 1) {
    for (int t = blockIdx.x; t < a.tokens; t += gridDim.x) {
      const size_t row_off = static_cast<size_t>(t) * a.H;
      const uint4* src = reinterpret_cast<const uint4*>(static_cast<const T*>(a.partial) + row_off);
      for (int c = threadIdx.x; c < nvec; c += kEpArThreads) {
        const uint4 val = ld_cg16(src + c);
        for (int r = 0; r < a.world; ++r)
          if (r != a.rank)
            reinterpret_cast<uint4*>(a.comm[r] + a.L.os[p] + static_cast<size_t>(a.rank) * a.L.os_slot +
                                     row_off * sizeof(T))[c] = val;
      }
      __syncthreads();
      if (threadIdx.x < a.world && threadIdx.x != a.rank) {
        int32_t* peer_flags = reinterpret_cast<int32_t*>(a.comm[threadIdx.x] + a.L.rowflag);
        __threadfence_system();      // the block's row (ordered before by the barrier) is visible before the flag
        st_release_sys(peer_flags + a.rank * kEpArOneShotRows + t, e * 4 + kEpArModeOne);
      }
    }
  }
  const unsigned char* inbox = a.comm[a.rank] + a.L.os[p];
  const int32_t* my_flags = reinterpret_cast<const int32_t*>(a.comm[a.rank] + a.L.rowflag);
  for (int t = blockIdx.x; t < a.tokens; t += gridDim.x) {
    if (threadIdx.x < a.world && threadIdx.x != a.rank)
      wait_flag(my_flags + threadIdx.x * kEpArOneShotRows + t, e, kEpArModeOne, ctl, a.watch, threadIdx.x);
    __syncthreads();
    RowRegs<T> R;
    sum_row<T, false>(a, t, inbox, a.L.os_slot, t, p, R);
    finish_row<T>(a, t, R, s_part);
  }
  retire(ctl, e, &s_flag);
}

// Two-shot: reduce-scatter -> my token shard -> all-gather of y, then the residual add + norm of every row
// locally (one NVLink plane instead of two: NVLink, not HBM, is the scarce resource at prefill sizes).
template <typename T>
__global__ void __launch_bounds__(kEpArThreads, 2) ep_ar_twoshot_kernel(const EpArArgs a) {
  constexpr int V = Dt<T>::kVec;
  __shared__ float s_part[kEpArThreads / 32];
  __shared__ int s_flag;
  pdl_wait();
  pdl_launch();
  int32_t* ctl = reinterpret_cast<int32_t*>(a.comm[a.rank] + a.L.ctl);
  const int e = *reinterpret_cast<volatile int32_t*>(ctl) + 1;
  const int p = e & 1;
  const int nvec = a.H / V;
  const int lo = min(a.rank * a.per, a.tokens);
  const int hi = min(lo + a.per, a.tokens);
  // ---- reduce-scatter, push side: my partial rows of slice s -> rank s.  The rows of a slice are contiguous on
  // both sides: R - 1 flat copies, four independent 16-byte loads in flight per thread (one load + one store per
  // iteration left the NVLink pipe latency-bound: 312 GB/s per rank at 8 ranks, profiles/r2.md section 10) ----
  {
    const size_t stride = static_cast<size_t>(gridDim.x) * kEpArThreads;
    const size_t tid = static_cast<size_t>(blockIdx.x) * kEpArThreads + threadIdx.x;
    for (int ds = 1; ds < a.world; ++ds) {
      const int s = (a.rank + ds) % a.world;            // start with the next rank: spread the first bytes over the links
      const int s_lo = min(s * a.per, a.tokens), s_hi = min(s_lo + a.per, a.tokens);
      const size_t n = static_cast<size_t>(s_hi - s_lo) * nvec;
      const uint4* src = reinterpret_cast<const uint4*>(static_cast<const T*>(a.partial) + static_cast<size_t>(s_lo) * a.H);
      uint4* dst = reinterpret_cast<uint4*>(a.comm[s] + a.L.rs[p] + static_cast<size_t>(a.rank) * a.L.rs_slot);
      for (size_t i = tid; i < n; i += 4 * stride) {
        uint4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (i + u * stride < n) v[u] = ld_cg16(src + i + u * stride);
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (i + u * stride < n) dst[i + u * stride] = v[u];
      }
    }
  }
  if (block_ticket(ctl + 1, &s_flag) && threadIdx.x < a.world) {
    int32_t* peer_ctl = reinterpret_cast<int32_t*>(a.comm[threadIdx.x] + a.L.ctl);
    __threadfence_system();
    st_release_sys(peer_ctl + 16 + a.rank, e * 4 + kEpArModeTwo);
  }
  if (threadIdx.x < a.world) wait_flag(ctl + 16 + threadIdx.x, e, kEpArModeTwo, ctl, a.watch, threadIdx.x);
  __syncthreads();
  // ---- my token slice: y to every peer's inbox; residual add + norm locally ----
  const unsigned char* inbox = a.comm[a.rank] + a.L.rs[p];
  for (int t = lo + blockIdx.x; t < hi; t += gridDim.x) {
    RowRegs<T> R;
    sum_row<T, true>(a, t, inbox, a.L.rs_slot, t - lo, p, R);
    finish_row<T>(a, t, R, s_part);
  }
  if (block_ticket(ctl + 3, &s_flag) && threadIdx.x < a.world) {
    int32_t* peer_ctl = reinterpret_cast<int32_t*>(a.comm[threadIdx.x] + a.L.ctl);
    __threadfence_system();
    st_release_sys(peer_ctl + 32 + a.rank, e * 4 + kEpArModeTwo);
  }
  if (threadIdx.x < a.world) wait_flag(ctl + 32 + threadIdx.x, e, kEpArModeTwo, ctl, a.watch, threadIdx.x);
  __syncthreads();
  // ---- the other ranks' slices: y from my inbox, same residual add + norm ----
  const unsigned char* ag = a.comm[a.rank] + a.L.ag[p];
  for (int t = blockIdx.x; t < a.tokens; t += gridDim.x) {
    if (t >= lo && t < hi) continue;
    RowRegs<T> R;
    load_row<T>(a, t, ag, R);
    finish_row<T>(a, t, R, s_part);
  }
  retire(ctl, e, &s_flag);
}

EpCommLayout ep_comm_layout(int max_tokens, int H, int dtype, int world) {
  EpCommLayout L{};
  const size_t es = dtype_size(dtype);
  const size_t row = static_cast<size_t>(H) * es;
  auto up = [](size_t v) { return (v + 1023) / 1024 * 1024; };
  size_t off = 0;
  L.ctl = off;
  off += 1024;
  const size_t full = up(static_cast<size_t>(max_tokens) * row);
  L.os_slot = full < kEpArOneShotSlot ? full : kEpArOneShotSlot;
  for (int p = 0; p < 2; ++p) {
    L.os[p] = off;
    off += L.os_slot * world;
  }
  const size_t per_cap = (static_cast<size_t>(max_tokens) + world - 1) / world;
  L.rs_slot = up(per_cap * row);
  for (int p = 0; p < 2; ++p) {
    L.rs[p] = off;
    off += L.rs_slot * world;
  }
  for (int p = 0; p < 2; ++p) {
    L.ag[p] = off;
    off += full;
  }
  L.rowflag = off;
  off += up(static_cast<size_t>(world) * kEpArOneShotRows * 4);
  L.total = off;
  return L;
}

int ep_allreduce_launch(const EpArArgs& args, int dtype, int two_shot, void* stream) {
  EpArArgs a = args;
  a.watch = ep_watch(a.rank, kEpSiteAllReduce);
  if (a.world < 1 || a.world > kEpMaxRanks || a.rank < 0 || a.rank >= a.world) return GLM4MOE_ERR_ARG;
  if (dtype != GLM4MOE_BF16 && dtype != GLM4MOE_F16) return GLM4MOE_ERR_UNSUPPORTED;
  if (a.H % 8 != 0 || a.H / 8 > kEpArThreads * kArMaxVec || a.tokens <= 0) return GLM4MOE_ERR_UNSUPPORTED;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int dev = 0, sms = 0;
  GLM4MOE_CUDA_CHECK(cudaGetDevice(&dev));
  GLM4MOE_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  // two-shot: every block waits for the peers' grid-wide flags inside the kernel, so the whole grid must be
  // co-resident (2 blocks / SM by __launch_bounds__); the one-shot form has no such need but gains nothing from more
  const int cap = sms * 2;
  int grid = a.tokens < cap ? a.tokens : cap;
  if (grid < 1) grid = 1;
  if (two_shot) {
    if (dtype == GLM4MOE_BF16)
      GLM4MOE_CUDA_CHECK(launch_pdl(ep_ar_twoshot_kernel<__nv_bfloat16>, dim3(grid), dim3(kEpArThreads), 0, st, a));
    else
      GLM4MOE_CUDA_CHECK(launch_pdl(ep_ar_twoshot_kernel<__half>, dim3(grid), dim3(kEpArThreads), 0, st, a));
  } else {
    if (dtype == GLM4MOE_BF16)
      GLM4MOE_CUDA_CHECK(launch_pdl(ep_ar_oneshot_kernel<__nv_bfloat16>, dim3(grid), dim3(kEpArThreads), 0, st, a));
    else
      GLM4MOE_CUDA_CHECK(launch_pdl(ep_ar_oneshot_kernel<__half>, dim3(grid), dim3(kEpArThreads), 0, st, a));
  }
  return GLM4MOE_OK;
}

}  // namespace glm4moe

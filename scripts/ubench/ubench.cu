// Micro-benchmarks that decide kernel structure (run on the GPU box, results in profiles/):
//   tmem  : tcgen05.ld throughput per SM with 4 / 8 reader warps  (FP8 per-k-block promotion budget)
//   stream: aggregate HBM->smem bulk-copy bandwidth vs. number of streaming CTAs and bytes in
//           flight per CTA (decode regime: is one SM's share of HBM capped below 1/148 of the chip?)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/ubench/ubench scripts/ubench/ubench.cu
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok = 0;
  while (!ok) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------
// tmem read throughput
// ---------------------------------------------------------------------------------------------
template <int X>
__device__ __forceinline__ void tmem_ld(uint32_t taddr, uint32_t* r);
template <>
__device__ __forceinline__ void tmem_ld<16>(uint32_t taddr, uint32_t* r) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                 "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]) : "r"(taddr));
}
template <>
__device__ __forceinline__ void tmem_ld<32>(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]) : "r"(taddr));
}

template <int X, int WAIT_EVERY>
__global__ void tmem_read_kernel(int iters, long long* cycles, uint32_t* sink) {
  __shared__ uint32_t tmem_ptr;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_ptr)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = tmem_ptr + (static_cast<uint32_t>((warp & 3) * 32) << 16);
  uint32_t acc = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int c = 0; c < 512; c += X * WAIT_EVERY) {
      uint32_t r[WAIT_EVERY][X];
#pragma unroll
      for (int q = 0; q < WAIT_EVERY; ++q) tmem_ld<X>(base + c + q * X, r[q]);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int q = 0; q < WAIT_EVERY; ++q)
#pragma unroll
        for (int j = 0; j < X; ++j) acc ^= r[q][j];
    }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  if (acc == 0x12345678u) sink[0] = acc;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_ptr));
}

template <int X, int WE>
static void run_tmem(int warps) {
  long long* d_cyc;
  uint32_t* d_sink;
  CK(cudaMalloc(&d_cyc, 8 * 148));
  CK(cudaMalloc(&d_sink, 4));
  const int iters = 200;
  tmem_read_kernel<X, WE><<<148, warps * 32>>>(iters, d_cyc, d_sink);
  CK(cudaDeviceSynchronize());
  tmem_read_kernel<X, WE><<<148, warps * 32>>>(iters, d_cyc, d_sink);
  CK(cudaDeviceSynchronize());
  long long c;
  CK(cudaMemcpy(&c, d_cyc, 8, cudaMemcpyDeviceToHost));
  // every warp reads its 32 lanes x 512 columns x 4 B per iteration
  const double bytes = static_cast<double>(iters) * warps * 32 * 512 * 4;
  printf("{\"bench\": \"tmem_ld\", \"shape\": \"32x32b.x%d\", \"lds_per_wait\": %d, \"warps\": %d, \"cycles\": %lld, \"bytes_per_clk_per_sm\": %.1f}\n",
         X, WE, warps, c, bytes / c);
  cudaFree(d_cyc);
  cudaFree(d_sink);
}

// ---------------------------------------------------------------------------------------------
// HBM -> smem streaming (bulk async copies, ring of `stages` chunks of `chunk` bytes per CTA)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128, 1) stream_kernel(const unsigned char* __restrict__ src, size_t bytes_per_cta,
                                                        int chunk, int stages) {
  extern __shared__ __align__(1024) unsigned char sm[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm + static_cast<size_t>(chunk) * stages);
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) mbar_init(&bars[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned char* p = src + static_cast<size_t>(blockIdx.x) * bytes_per_cta;
    const int n = static_cast<int>(bytes_per_cta / chunk);
    for (int i = 0; i < n + stages; ++i) {
      const int s = i % stages;
      if (i >= stages) mbar_wait(&bars[s], ((i / stages) - 1) & 1u);   // chunk i - stages landed
      if (i < n) {
        mbar_expect_tx(&bars[s], chunk);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         smem_u32(sm + static_cast<size_t>(s) * chunk)),
                     "l"(p + static_cast<size_t>(i) * chunk), "r"(chunk), "r"(smem_u32(&bars[s]))
                     : "memory");
      }
    }
  }
}

static void run_stream(const unsigned char* buf, size_t buf_bytes, int ctas, int chunk, int stages, size_t per_cta) {
  if (per_cta * ctas > buf_bytes) per_cta = buf_bytes / ctas / chunk * chunk;
  const size_t smem = static_cast<size_t>(chunk) * stages + 8 * stages;
  CK(cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 5; ++rep) {
    // rotate the source window so the 126 MB L2 never holds the data (buffer is 4 GB)
    const size_t off = (static_cast<size_t>(rep) * per_cta * ctas) % (buf_bytes - per_cta * ctas + 1) / 1024 * 1024;
    CK(cudaEventRecord(e0));
    stream_kernel<<<ctas, 128, smem>>>(buf + off, per_cta, chunk, stages);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    if (rep > 0 && ms < best) best = ms;
  }
  const double gb = static_cast<double>(per_cta) * ctas / 1e9;
  printf("{\"bench\": \"stream\", \"ctas\": %d, \"chunk\": %d, \"stages\": %d, \"mb_per_cta\": %.2f, \"us\": %.1f, \"GBps\": %.0f, \"GBps_per_cta\": %.1f}\n",
         ctas, chunk, stages, per_cta / 1e6, best * 1e3, gb / (best * 1e-3), gb / (best * 1e-3) / ctas);
}

int main(int argc, char** argv) {
  run_tmem<16, 1>(4);
  run_tmem<16, 1>(8);
  run_tmem<32, 1>(4);
  run_tmem<32, 1>(8);
  run_tmem<32, 2>(4);
  run_tmem<32, 2>(8);
  run_tmem<32, 4>(4);
  run_tmem<32, 4>(8);
  const size_t buf_bytes = 4ull << 30;
  unsigned char* buf;
  CK(cudaMalloc(&buf, buf_bytes));
  CK(cudaMemset(buf, 1, buf_bytes));
  const int cta_list[] = {1, 8, 37, 74, 99, 148};
  for (int ctas : cta_list) {
    run_stream(buf, buf_bytes, ctas, 16384, 12, 2u << 20);    // the T=1 gate_up work item: 2 MB per CTA
    run_stream(buf, buf_bytes, ctas, 16384, 12, 16u << 20);
    run_stream(buf, buf_bytes, ctas, 32768, 6, 16u << 20);
    run_stream(buf, buf_bytes, ctas, 8192, 24, 16u << 20);
  }
  run_stream(buf, buf_bytes, 148, 16384, 12, 1u << 20);
  run_stream(buf, buf_bytes, 148, 16384, 12, 512u << 10);
  run_stream(buf, buf_bytes, 148, 16384, 4, 16u << 20);
  run_stream(buf, buf_bytes, 148, 16384, 8, 16u << 20);
  run_stream(buf, buf_bytes, 296, 16384, 6, 8u << 20);
  return 0;
}

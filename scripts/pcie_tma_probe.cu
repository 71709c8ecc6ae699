// PCIe probe (experiment, not product): TMA bulk loads from pinned host memory, TMA bulk stores to pinned host
// memory, alone, together in one kernel, and loads concurrent with a copy-engine D2H. GB/s each.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(n)); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t ph) {
  asm volatile("{\n .reg .pred p;\n W: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra D;\n bra W;\n D:\n}" ::"r"(smem_u32(b)), "r"(ph) : "memory");
}
__device__ __forceinline__ void tma_load(void* dst, const void* src, uint32_t bytes, uint64_t* b) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void tma_store(void* dst, const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
}

constexpr int ROW = 1920, OROW = 2688, NL = 8, NS = 3;  // per iteration: 8 rows in (15360 B), 3 rows out (8064 B)

// mode bit0: loads, bit1: stores
__global__ void probe(const uint8_t* src, uint8_t* dst, long iters_total, int mode) {
  extern __shared__ __align__(128) uint8_t sm[];
  __shared__ uint64_t bar;
  uint8_t* in = sm;
  uint8_t* out = sm + NL * ROW;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;"); }
  for (int i = threadIdx.x; i < NS * OROW; i += blockDim.x) out[i] = (uint8_t)i;
  asm volatile("fence.proxy.async.shared::cta;");
  __syncthreads();
  uint32_t ph = 0;
  if (threadIdx.x == 0) {
    for (long it = blockIdx.x; it < iters_total; it += gridDim.x) {
      if (mode & 1) {
        mbar_expect(&bar, NL * ROW);
        for (int r = 0; r < NL; ++r) tma_load(in + r * ROW, src + (it * NL + r) * (long)ROW, ROW, &bar);
      }
      if (mode & 2) {
        for (int r = 0; r < NS; ++r) tma_store(dst + (it * NS + r) * (long)OROW, out + r * OROW, OROW);
        asm volatile("cp.async.bulk.commit_group;");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      }
      if (mode & 1) { mbar_wait(&bar, ph); ph ^= 1; }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}

int main() {
  const long iters = 40000;  // 614 MB in, 322 MB out
  uint8_t *hsrc, *hdst, *hdst2, *dbuf;
  CK(cudaHostAlloc(&hsrc, iters * NL * ROW, cudaHostAllocDefault));
  CK(cudaHostAlloc(&hdst, iters * NS * OROW, cudaHostAllocDefault));
  CK(cudaHostAlloc(&hdst2, iters * NS * OROW, cudaHostAllocDefault));
  CK(cudaMalloc(&dbuf, iters * NS * OROW));
  for (long i = 0; i < iters * NL * ROW; i += 4096) hsrc[i] = 1;
  cudaStream_t s1, s2; CK(cudaStreamCreate(&s1)); CK(cudaStreamCreate(&s2));
  cudaEvent_t a, b, c, d; cudaEventCreate(&a); cudaEventCreate(&b); cudaEventCreate(&c); cudaEventCreate(&d);
  int smem = NL * ROW + NS * OROW;
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  for (int grid : {148, 444, 888}) {
    for (int mode = 1; mode <= 3; ++mode) {
      for (int rep = 0; rep < 2; ++rep) {
        CK(cudaEventRecord(a, s1));
        probe<<<grid, 32, smem, s1>>>(hsrc, hdst, iters, mode);
        CK(cudaEventRecord(b, s1));
        CK(cudaStreamSynchronize(s1));
      }
      float ms; cudaEventElapsedTime(&ms, a, b);
      printf("grid %4d mode %d (%s%s): %.3f ms  in %.1f GB/s  out %.1f GB/s\n", grid, mode, mode & 1 ? "L" : "", mode & 2 ? "S" : "", ms,
             (mode & 1) ? iters * NL * ROW / ms / 1e6 : 0.0, (mode & 2) ? iters * NS * OROW / ms / 1e6 : 0.0);
    }
  }
  // loads in a kernel, D2H by the copy engine at the same time
  for (int rep = 0; rep < 2; ++rep) {
    CK(cudaEventRecord(a, s1)); CK(cudaEventRecord(c, s2));
    probe<<<444, 32, smem, s1>>>(hsrc, hdst, iters, 1);
    CK(cudaMemcpyAsync(hdst2, dbuf, iters * NS * OROW, cudaMemcpyDeviceToHost, s2));
    CK(cudaEventRecord(b, s1)); CK(cudaEventRecord(d, s2));
    CK(cudaDeviceSynchronize());
  }
  float m1, m2; cudaEventElapsedTime(&m1, a, b); cudaEventElapsedTime(&m2, c, d);
  printf("kernel loads || copy-engine D2H: kernel %.3f ms (%.1f GB/s in), memcpy %.3f ms (%.1f GB/s out)\n", m1, iters * NL * ROW / m1 / 1e6, m2, iters * NS * OROW / m2 / 1e6);
  // load kernel || store kernel on two streams
  for (int rep = 0; rep < 2; ++rep) {
    CK(cudaEventRecord(a, s1)); CK(cudaEventRecord(c, s2));
    probe<<<222, 32, smem, s1>>>(hsrc, hdst, iters, 1);
    probe<<<222, 32, smem, s2>>>(hsrc, hdst2, iters, 2);
    CK(cudaEventRecord(b, s1)); CK(cudaEventRecord(d, s2));
    CK(cudaDeviceSynchronize());
  }
  cudaEventElapsedTime(&m1, a, b); cudaEventElapsedTime(&m2, c, d);
  printf("load kernel || store kernel: loads %.3f ms (%.1f GB/s), stores %.3f ms (%.1f GB/s)\n", m1, iters * NL * ROW / m1 / 1e6, m2, iters * NS * OROW / m2 / 1e6);
  return 0;
}

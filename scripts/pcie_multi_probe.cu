// PCIe ceiling of the box with 1, 2, 4, 8 GPUs busy at once (experiment, not product). VERDICT r1: the e2e curve
// collapses at N = 4 / 8; is ~117 GB/s of aggregate host reads the box, or the kernel's access pattern?
// One host thread per GPU, pinned buffers allocated by that thread after cudaSetDevice (portable + mapped), all GPUs
// released together, device time by CUDA events, aggregate = sum over GPUs of bytes / max over GPUs of time.
//   copy H2D / D2H / both       : copy engines, 256 MiB per direction per repetition
//   kernel loads (row bytes R)  : cp.async.bulk of R-byte rows from pinned host memory into shared memory, one elected
//                                 thread per CTA, 444 CTAs (what the row-staged kernel's producer warp does);
//                                 "sampled" skips rows the way a 1080 -> 224 down-scale does (2 of every 4.8 rows)
//   kernel loads + stores       : the same with bulk stores of 2688-byte rows to pinned host memory in the same kernel
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/pcie_multi_probe scripts/pcie_multi_probe.cu -lpthread
#include <cuda_runtime.h>

#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

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

constexpr int NL = 8, NS = 3, OROW = 2688, MAXROW = 4096;
// mode bit0: loads of NL rows of `row` bytes per iteration (row pitch `pitch`, every `skip16`/16-th row), bit1: NS stores
__global__ void probe(const uint8_t* src, uint8_t* dst, long iters_total, int mode, int row, long pitch, int skip16) {
  extern __shared__ __align__(128) uint8_t sm[];
  __shared__ uint64_t bar;
  uint8_t* in = sm;
  uint8_t* out = sm + NL * MAXROW;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;"); }
  for (int i = threadIdx.x; i < NS * OROW; i += blockDim.x) out[i] = (uint8_t)i;
  asm volatile("fence.proxy.async.shared::cta;");
  __syncthreads();
  uint32_t ph = 0;
  if (threadIdx.x == 0) {
    for (long it = blockIdx.x; it < iters_total; it += gridDim.x) {
      if (mode & 1) {
        mbar_expect(&bar, NL * row);
        for (int r = 0; r < NL; ++r) {
          const long src_row = ((it * NL + r) * (long)skip16) >> 4;
          tma_load(in + r * MAXROW, src + src_row * pitch, row, &bar);
        }
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

struct Gpu {
  int dev;
  uint8_t *hsrc, *hdst, *dbuf;
  cudaStream_t s1, s2;
  cudaEvent_t a, b, c, d;
  float ms1, ms2;
};
static const size_t kCopy = 256u << 20;
static const long kIters = 16000;                                    // kernel: 16000 x 8 rows
static const size_t kSrcBytes = (size_t)kIters * NL * MAXROW * 2;     // room for pitch 4096 and skipping (x4.8/2 < 2.5) -> clamp below
static std::atomic<int> g_arrived{0};
static std::atomic<int> g_phase{0};
static void barrier(int n) {  // spin barrier between the per-GPU host threads
  const int phase = g_phase.load();
  if (g_arrived.fetch_add(1) + 1 == n) { g_arrived.store(0); g_phase.fetch_add(1); }
  else while (g_phase.load() == phase) std::this_thread::yield();
}

struct Test { const char* name; int kind; int mode; int row; long pitch; int skip16; };

static void worker(Gpu* g, int n, const std::vector<Test>* tests, std::vector<std::vector<float>>* ms, std::vector<double>* bytes_in,
                   std::vector<double>* bytes_out, int slot) {
  CK(cudaSetDevice(g->dev));
  CK(cudaHostAlloc(&g->hsrc, kSrcBytes, cudaHostAllocPortable | cudaHostAllocMapped));
  CK(cudaHostAlloc(&g->hdst, kCopy, cudaHostAllocPortable | cudaHostAllocMapped));
  CK(cudaMalloc(&g->dbuf, kCopy * 2));
  for (size_t i = 0; i < kSrcBytes; i += 4096) g->hsrc[i] = 1;
  for (size_t i = 0; i < kCopy; i += 4096) g->hdst[i] = 1;
  CK(cudaStreamCreate(&g->s1)); CK(cudaStreamCreate(&g->s2));
  cudaEventCreate(&g->a); cudaEventCreate(&g->b); cudaEventCreate(&g->c); cudaEventCreate(&g->d);
  const int smem = NL * MAXROW + NS * OROW;
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  for (size_t t = 0; t < tests->size(); ++t) {
    const Test& T = (*tests)[t];
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
      CK(cudaDeviceSynchronize());
      barrier(n);
      CK(cudaEventRecord(g->a, g->s1));
      CK(cudaEventRecord(g->c, g->s2));
      if (T.kind == 0) {
        if (T.mode & 1) CK(cudaMemcpyAsync(g->dbuf, g->hsrc, kCopy, cudaMemcpyHostToDevice, g->s1));
        if (T.mode & 2) CK(cudaMemcpyAsync(g->hdst, g->dbuf + kCopy, kCopy, cudaMemcpyDeviceToHost, g->s2));
      } else {
        probe<<<444, 32, smem, g->s1>>>(g->hsrc, g->hdst, kIters, T.mode, T.row, T.pitch, T.skip16);
      }
      CK(cudaEventRecord(g->b, g->s1));
      CK(cudaEventRecord(g->d, g->s2));
      CK(cudaDeviceSynchronize());
      float m1, m2;
      cudaEventElapsedTime(&m1, g->a, g->b);
      cudaEventElapsedTime(&m2, g->c, g->d);
      const float m = m1 > m2 ? m1 : m2;
      if (rep > 0 && m < best) best = m;
      barrier(n);
    }
    (*ms)[t][slot] = best;
    if (T.kind == 0) {
      (*bytes_in)[t] = (T.mode & 1) ? (double)kCopy : 0;
      (*bytes_out)[t] = (T.mode & 2) ? (double)kCopy : 0;
    } else {
      (*bytes_in)[t] = (T.mode & 1) ? (double)kIters * NL * T.row : 0;
      (*bytes_out)[t] = (T.mode & 2) ? (double)kIters * NS * OROW : 0;
    }
  }
  cudaFreeHost(g->hsrc); cudaFreeHost(g->hdst); cudaFree(g->dbuf);
}

int main() {
  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  // skip16 = 16: consecutive rows; 38 = 2.4 rows per step (the 1080 -> 224 luma pattern reads 2 of every 4.8 rows)
  std::vector<Test> tests = {
      {"copy engine H2D", 0, 1, 0, 0, 0},
      {"copy engine D2H", 0, 2, 0, 0, 0},
      {"copy engines both", 0, 3, 0, 0, 0},
      {"kernel loads 1920 B rows, consecutive", 1, 1, 1920, 1920, 16},
      {"kernel loads 1920 B rows, sampled (2 of 4.8)", 1, 1, 1920, 1920, 38},
      {"kernel loads 1920 B rows at 4096 B pitch", 1, 1, 1920, 4096, 16},
      {"kernel loads 3840 B (two adjacent rows)", 1, 1, 3840, 3840, 16},
      {"kernel loads 1920 B + stores 2688 B", 1, 3, 1920, 1920, 16},
      {"kernel stores 2688 B rows", 1, 2, 1920, 1920, 16},
  };
  printf("# %d GPU(s) visible; GB/s aggregate over the active GPUs (per-GPU time = best of 2 timed repetitions; aggregate uses the slowest GPU)\n", ndev);
  for (int n = 1; n <= ndev; n *= 2) {
    std::vector<Gpu> gpus(n);
    std::vector<std::vector<float>> ms(tests.size(), std::vector<float>(n, 0.f));
    std::vector<double> bin(tests.size(), 0), bout(tests.size(), 0);
    std::vector<std::thread> th;
    g_arrived = 0;
    for (int i = 0; i < n; ++i) { gpus[i].dev = i; th.emplace_back(worker, &gpus[i], n, &tests, &ms, &bin, &bout, i); }
    for (auto& t : th) t.join();
    for (size_t t = 0; t < tests.size(); ++t) {
      float worst = 0, bestg = 1e30f;
      for (int i = 0; i < n; ++i) { worst = ms[t][i] > worst ? ms[t][i] : worst; bestg = ms[t][i] < bestg ? ms[t][i] : bestg; }
      printf("N=%d  %-48s in %7.1f GB/s  out %7.1f GB/s   (per GPU: slowest %.2f ms, fastest %.2f ms)\n", n, tests[t].name,
             bin[t] * n / worst / 1e6, bout[t] * n / worst / 1e6, worst, bestg);
    }
  }
  return 0;
}

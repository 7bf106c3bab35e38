// gather_bench.cu -- how many random 4-byte gathers per clock can one SM issue, by where the table lives?
//   local   : the CTA's own shared memory                         (what the BLOCKED producer does)
//   dsmem C : the shared memory of the C CTAs of a cluster         (ld.shared::cluster through mapa)
//   global  : a 64 MB table in global memory, L2 resident at best  (what the row-major schedules do)
// Answers the question behind DESIGN.md section 3: is a cluster-wide x slice (C x 200 KB) a way past the
// one-request-per-clock L1TEX ceiling of global gathers?  Prints one JSON line per case.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o gather_bench gather_bench.cu && ./gather_bench
#include <cooperative_groups.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;

#define CK(x)                                                                     \
  do {                                                                            \
    cudaError_t e = (x);                                                          \
    if (e != cudaSuccess) {                                                       \
      printf("{\"error\": \"%s at line %d\"}\n", cudaGetErrorString(e), __LINE__); \
      exit(1);                                                                    \
    }                                                                             \
  } while (0)

constexpr int WORDS = 48 * 1024;  // 192 KB of table per CTA
constexpr int THREADS = 1024;
constexpr int ITERS = 256;
constexpr int ILP = 8;

__device__ __forceinline__ uint32_t mix(uint32_t h) {
  h ^= h >> 16; h *= 0x7feb352du; h ^= h >> 15; h *= 0x846ca68bu; h ^= h >> 16;
  return h;
}

// MODE 0 local, 1 dsmem, 2 global, 3 dsmem float add (red.shared::cluster), 4 local float add (red.shared)
template <int MODE>
__global__ void __launch_bounds__(THREADS, 1) gather_kernel(const uint32_t *__restrict__ table, uint32_t table_words,
                                                            uint32_t *__restrict__ sink, unsigned long long *__restrict__ cycles) {
  extern __shared__ uint32_t sm[];
  for (int i = threadIdx.x; i < WORDS; i += THREADS) sm[i] = mix(i + blockIdx.x * WORDS);
  uint32_t csize = 1;
  if (MODE == 1 || MODE == 3) {
    cg::cluster_group cl = cg::this_cluster();
    csize = cl.num_blocks();
    cl.sync();
  } else {
    __syncthreads();
  }
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(sm);
  uint32_t h = mix(blockIdx.x * THREADS + threadIdx.x + 1);
  uint32_t acc = 0;
  const long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) {
    uint32_t v[ILP];
#pragma unroll
    for (int k = 0; k < ILP; ++k) {
      h = h * 1664525u + 1013904223u;
      const uint32_t r = h >> 4;
      const uint32_t w = (r & 0x7fffu) + ((r >> 15) & 0x3fffu);  // < WORDS without a division
      if (MODE == 0) {
        v[k] = sm[w];
      } else if (MODE == 1) {
        const uint32_t local = sbase + 4 * w;
        const uint32_t rank = (h >> 2) & (csize - 1);
        uint32_t remote;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(rank));
        asm volatile("ld.shared::cluster.u32 %0, [%1];" : "=r"(v[k]) : "r"(remote));
      } else if (MODE == 3) {
        const uint32_t local = sbase + 4 * w;
        const uint32_t rank = (h >> 2) & (csize - 1);
        uint32_t remote;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(rank));
        asm volatile("red.relaxed.cluster.shared::cluster.add.f32 [%0], %1;" ::"r"(remote), "f"(1.0f) : "memory");
        v[k] = 0;
      } else if (MODE == 4) {
        asm volatile("red.relaxed.cta.shared::cta.add.f32 [%0], %1;" ::"r"(sbase + 4 * w), "f"(1.0f) : "memory");
        v[k] = 0;
      } else {
        v[k] = __ldg(table + (r & (table_words - 1)));
      }
    }
#pragma unroll
    for (int k = 0; k < ILP; ++k) acc += v[k];
  }
  __syncthreads();
  const long long t1 = clock64();
  if (MODE == 1 || MODE == 3) cg::this_cluster().sync();
  else __syncthreads();
  if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(t1 - t0);
  if (acc == 0x12345678u) sink[0] = acc;
}

template <int MODE>
static void run(const char *name, int cluster, const uint32_t *table, uint32_t table_words, uint32_t *sink,
                unsigned long long *cycles, int n_sm) {
  const size_t smem = (size_t)WORDS * 4;
  CK(cudaFuncSetAttribute(gather_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (cluster > 8) CK(cudaFuncSetAttribute(gather_kernel<MODE>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  int grid = n_sm - n_sm % cluster;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(THREADS);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = cluster;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = (MODE == 1 || MODE == 3) ? 1 : 0;
  if (MODE == 1 || MODE == 3) {
    int max_clusters = 0;
    cudaError_t e = cudaOccupancyMaxActiveClusters(&max_clusters, gather_kernel<MODE>, &cfg);
    if (e != cudaSuccess || max_clusters == 0) {
      printf("{\"case\": \"%s\", \"cluster\": %d, \"skipped\": \"%s, max active clusters %d\"}\n", name, cluster,
             cudaGetErrorString(e), max_clusters);
      cudaGetLastError();
      return;
    }
    if (grid > max_clusters * cluster) grid = max_clusters * cluster;
    cfg.gridDim = dim3(grid);
  }
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    CK(cudaEventRecord(e0));
    CK(cudaLaunchKernelEx(&cfg, gather_kernel<MODE>, table, table_words, sink, cycles));
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    if (rep && ms < best) best = ms;
  }
  unsigned long long h[256];
  CK(cudaMemcpy(h, cycles, sizeof(unsigned long long) * grid, cudaMemcpyDeviceToHost));
  double mean = 0;
  for (int i = 0; i < grid; ++i) mean += (double)h[i];
  mean /= grid;
  const double per_cta = (double)THREADS * ITERS * ILP;
  printf("{\"case\": \"%s\", \"cluster\": %d, \"ctas\": %d, \"gathers_per_clk_per_sm\": %.3f, \"cycles\": %.0f, \"ms\": %.4f, "
         "\"chip_gathers_per_s\": %.4g}\n",
         name, cluster, grid, per_cta / mean, mean, best, per_cta * grid / (best * 1e-3));
}


// The config-2 access pattern with nothing else around it: every thread issues G independent 4-byte gathers from a
// 64 MB table per step (distinct 128-byte lines, as 16 uniform-random columns per row give), B CTAs of 256 threads per
// SM.  What this tops out at is the ceiling any row-major SpMV schedule has on this operator.
template <int G>
__global__ void __launch_bounds__(256) pure_gather_kernel(const uint32_t *__restrict__ table, uint32_t mask, int iters,
                                                          uint32_t *__restrict__ sink) {
  uint32_t h = mix(blockIdx.x * 256 + threadIdx.x + 1);
  uint32_t acc = 0;
  for (int it = 0; it < iters; ++it) {
    uint32_t v[G];
#pragma unroll
    for (int k = 0; k < G; ++k) {
      h = h * 1664525u + 1013904223u;
      v[k] = __ldg(table + ((h >> 4) & mask));
    }
#pragma unroll
    for (int k = 0; k < G; ++k) acc += v[k];
  }
  if (acc == 0x12345678u) sink[0] = acc;
}

template <int G>
static void run_pure(const uint32_t *table, uint32_t table_words, uint32_t *sink, int n_sm, int ctas_per_sm, double sm_hz) {
  const long long total = 268435456LL;  // the gathers of one config-2 SpMV
  const int grid = n_sm * ctas_per_sm;
  const int iters = (int)(total / ((long long)grid * 256 * G));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 5; ++rep) {
    CK(cudaEventRecord(e0));
    pure_gather_kernel<G><<<grid, 256>>>(table, table_words - 1, iters, sink);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    if (rep && ms < best) best = ms;
  }
  const double n = (double)grid * 256 * G * iters;
  printf("{\"case\": \"pure gather, 64 MB table\", \"gathers_in_flight_per_thread\": %d, \"ctas_per_sm\": %d, \"gathers\": %.0f, \"ms\": %.4f, "
         "\"chip_gathers_per_s\": %.4g, \"gathers_per_clk_per_sm\": %.3f, \"ms_for_268M_gathers\": %.4f}\n",
         G, ctas_per_sm, n, best, n / (best * 1e-3), n / (best * 1e-3) / (n_sm * sm_hz), best * 268435456.0 / n);
}

int main() {
  cudaDeviceProp p;
  CK(cudaGetDeviceProperties(&p, 0));
  const uint32_t table_words = 16u << 20;  // 64 MB
  uint32_t *table, *sink;
  unsigned long long *cycles;
  CK(cudaMalloc(&table, (size_t)table_words * 4));
  CK(cudaMemset(table, 1, (size_t)table_words * 4));
  CK(cudaMalloc(&sink, 4));
  CK(cudaMalloc(&cycles, 256 * 8));
  run<0>("local smem", 1, table, table_words, sink, cycles, p.multiProcessorCount);
  for (int c : {2, 4, 8, 16}) run<1>("dsmem", c, table, table_words, sink, cycles, p.multiProcessorCount);
  run<4>("local smem red.add.f32", 1, table, table_words, sink, cycles, p.multiProcessorCount);
  for (int c : {2, 8, 16}) run<3>("dsmem red.add.f32", c, table, table_words, sink, cycles, p.multiProcessorCount);
  run<2>("global 64 MB", 1, table, table_words, sink, cycles, p.multiProcessorCount);
  const double sm_hz = (double)p.clockRate * 1e3;
  for (int b : {4, 8}) {
    run_pure<4>(table, table_words, sink, p.multiProcessorCount, b, sm_hz);
    run_pure<8>(table, table_words, sink, p.multiProcessorCount, b, sm_hz);
    run_pure<16>(table, table_words, sink, p.multiProcessorCount, b, sm_hz);
  }
  return 0;
}

// Remote shared-memory atomics inside a thread-block cluster (DSMEM): red.shared::cluster.add.u32 at random addresses
// of random CTAs of the cluster, against the same loop on local shared memory and on an L2-resident global table.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench4 tools/microbench4.cu
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;

__device__ __forceinline__ uint32_t rng(uint32_t& s) { s = s * 1664525u + 1013904223u; return s >> 8; }

template <int MODE>  // 0 local smem, 1 cluster (DSMEM) spread over all CTAs, 2 global (L2) table
__global__ void atom_kernel(uint32_t* gtable, uint32_t gmask, int iters, int words_per_cta, uint32_t* sink)
{
  extern __shared__ uint32_t sm[];
  for (int i = threadIdx.x; i < words_per_cta; i += blockDim.x) sm[i] = 0;
  cg::cluster_group cl = cg::this_cluster();
  const unsigned csz = MODE == 1 ? cl.num_blocks() : 1;
  if (MODE == 1) cl.sync(); else __syncthreads();
  uint32_t s = blockIdx.x * 9781u + threadIdx.x * 6271u + 17u;
  const uint32_t smem_base = (uint32_t)__cvta_generic_to_shared(sm);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const uint32_t r = rng(s);
      if (MODE == 0) atomicAdd(&sm[r % words_per_cta], 1u);
      else if (MODE == 1) {
        const uint32_t cta = r % csz, w = (r / csz) % words_per_cta;
        uint32_t remote;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_base + 4 * w), "r"(cta));
        asm volatile("red.shared::cluster.add.u32 [%0], %1;" ::"r"(remote), "r"(1u) : "memory");
      } else atomicAdd(&gtable[r & gmask], 1u);
    }
  }
  if (MODE == 1) cl.sync(); else __syncthreads();
  if (threadIdx.x == 0) sink[blockIdx.x] = sm[0];
}

template <int MODE>
float run(int csz, int blocks, int threads, int iters, int words, uint32_t* gtable, uint32_t gmask, uint32_t* sink)
{
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(blocks); cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = (size_t)words * 4;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = csz; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  cudaFuncSetAttribute(atom_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, words * 4);
  if (csz > 8) cudaFuncSetAttribute(atom_kernel<MODE>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float ms = 0;
  for (int rep = 0; rep < 2; ++rep) {
    cudaEventRecord(e0);
    cudaError_t e = cudaLaunchKernelEx(&cfg, atom_kernel<MODE>, gtable, gmask, iters, words, sink);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    if (e != cudaSuccess || cudaGetLastError() != cudaSuccess) { printf("launch failed: %s\n", cudaGetErrorString(e)); return -1; }
    cudaEventElapsedTime(&ms, e0, e1);
  }
  return ms;
}

int main()
{
  uint32_t *gtable, *sink;
  const uint32_t gwords = 1u << 22;  // 16 MB table: L2 resident
  cudaMalloc(&gtable, gwords * 4); cudaMemset(gtable, 0, gwords * 4);
  cudaMalloc(&sink, 4096 * 4);
  const int threads = 1024, iters = 2048, words = 200 * 1024 / 4;
  const double per_block = (double)threads * iters * 4;
  {
    float ms = run<0>(1, 148, threads, iters, words, gtable, gwords - 1, sink);
    printf("local smem     148 CTAs: %7.3f ms  %7.1f Gop/s  %.2f op/clk/SM\n", ms, 148 * per_block / ms / 1e6, per_block / (ms * 1e-3) / 1.965e9);
    ms = run<2>(1, 148, threads, iters, words, gtable, gwords - 1, sink);
    printf("global L2 16MB 148 CTAs: %7.3f ms  %7.1f Gop/s  %.2f op/clk/SM\n", ms, 148 * per_block / ms / 1e6, per_block / (ms * 1e-3) / 1.965e9);
  }
  for (int csz : {2, 4, 8, 16}) {
    const int blocks = (148 / csz) * csz;
    float ms = run<1>(csz, blocks, threads, iters, words, gtable, gwords - 1, sink);
    if (ms > 0) printf("cluster %2d     %3d CTAs: %7.3f ms  %7.1f Gop/s  %.2f op/clk/SM\n", csz, blocks, ms, blocks * per_block / ms / 1e6, per_block / (ms * 1e-3) / 1.965e9);
  }
  return 0;
}

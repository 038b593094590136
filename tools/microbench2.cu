// Microbenchmark 2: streaming-read bandwidth of one persistent 1024-thread CTA per SM as a function of how much shared
// memory the CTA reserves (the L1 data array shrinks by the same amount), for LDG.128 with different cache hints and for
// TMA bulk copies (cp.async.bulk global->shared + mbarrier) with a ring of staging buffers.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/microbench2 tools/microbench2.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint4 ld_nc_na(const uint4* p) { uint4 v; asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p)); return v; }
__device__ __forceinline__ uint4 ld_cg(const uint4* p) { uint4 v; asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p)); return v; }

// MODE 0: __ldcs, 1: ld.global.nc.L1::no_allocate, 2: ld.global.cg ; U loads of 16 B in flight per thread
template <int MODE, int U>
__global__ void __launch_bounds__(1024, 1) k_ldg(const uint4* __restrict__ a, uint64_t n16, unsigned long long* out)
{
  __shared__ unsigned char sm[16];
  if (threadIdx.x == 0) sm[0] = 1;
  uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  unsigned acc = 0;
  for (; i + (U - 1) * stride < n16; i += U * stride) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = MODE == 0 ? __ldcs(a + i + u * stride) : MODE == 1 ? ld_nc_na(a + i + u * stride) : ld_cg(a + i + u * stride);
#pragma unroll
    for (int u = 0; u < U; ++u) acc += v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
  }
  if (acc == 0x12345678u) out[0] = acc + sm[0];
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(b)), "r"(c)); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(b)) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity)
{
  asm volatile("{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" :: "r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// TMA ring: STAGES buffers of stage_bytes; thread 0 is the producer; all threads consume (sum) and arrive on the empty barrier
__global__ void __launch_bounds__(1024, 1) k_tma(const uint8_t* __restrict__ a, uint64_t bytes, int stages, uint32_t stage_bytes, uint32_t pad_bytes,
                                                 unsigned long long* out)
{
  extern __shared__ __align__(128) unsigned char sm[];
  uint64_t* full  = reinterpret_cast<uint64_t*>(sm);
  uint64_t* empty = full + 16;
  unsigned char* buf = sm + 256;
  (void)pad_bytes;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, blockDim.x / 32); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const uint64_t ntiles = bytes / stage_bytes;
  const uint64_t my0 = blockIdx.x, step = gridDim.x;
  unsigned acc = 0;
  // prologue: fill the ring
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      uint64_t t = my0 + (uint64_t)s * step;
      if (t < ntiles) { mbar_expect(full + s, stage_bytes); tma_bulk_g2s(buf + (size_t)s * stage_bytes, a + t * stage_bytes, stage_bytes, full + s); }
    }
  }
  int s = 0; uint32_t ph = 0;
  for (uint64_t t = my0; t < ntiles; t += step) {
    mbar_wait(full + s, ph);
    const uint4* p = reinterpret_cast<const uint4*>(buf + (size_t)s * stage_bytes);
    for (uint32_t i = threadIdx.x; i < stage_bytes / 16; i += blockDim.x) { uint4 v = p[i]; acc += v.x ^ v.y ^ v.z ^ v.w; }
    __syncwarp();
    if ((threadIdx.x & 31) == 0) mbar_arrive(empty + s);
    if (threadIdx.x == 0) {
      uint64_t tn = t + (uint64_t)stages * step;
      if (tn < ntiles) {
        mbar_wait(empty + s, ph);
        mbar_expect(full + s, stage_bytes);
        tma_bulk_g2s(buf + (size_t)s * stage_bytes, a + tn * stage_bytes, stage_bytes, full + s);
      }
    }
    if (++s == stages) { s = 0; ph ^= 1; }
  }
  if (acc == 0x12345678u) out[0] = acc;
}

__global__ void fill(uint64_t* a, uint64_t n) { uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, st = (uint64_t)gridDim.x * blockDim.x; for (; i < n; i += st) a[i] = i * 0x9E3779B97F4A7C15ull; }

template <typename F> static float time_ms(F f, int reps = 5)
{
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  f(); CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < reps; ++r) { CK(cudaEventRecord(e0)); f(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms; }
  CK(cudaGetLastError());
  return best;
}

int main()
{
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  int nsm = prop.multiProcessorCount;
  uint64_t bytes = 3200000000ull;
  uint8_t* a; CK(cudaMalloc(&a, bytes)); unsigned long long* out; CK(cudaMalloc(&out, 64));
  fill<<<nsm * 8, 256>>>((uint64_t*)a, bytes / 8); CK(cudaDeviceSynchronize());
  printf("== LDG.128 streaming, 1 CTA x 1024 thr per SM, GB/s vs dynamic smem reserved ==\n");
  printf("%-34s", "mode \\ smem KB");
  int smems[] = {0, 64, 128, 160, 192, 208, 220, 226};
  for (int s : smems) printf("%8d", s); printf("\n");
  auto row = [&](const char* name, auto kern) {
    printf("%-34s", name);
    for (int s : smems) {
      CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, s * 1024));
      float ms = time_ms([&] { kern<<<nsm, 1024, s * 1024>>>((const uint4*)a, bytes / 16, out); });
      printf("%8.0f", bytes / (ms * 1e6));
    }
    printf("\n");
  };
  row("ldcs U=2 (32 B/thread)", k_ldg<0, 2>); row("ldcs U=4 (64 B/thread)", k_ldg<0, 4>); row("ldcs U=8 (128 B/thread)", k_ldg<0, 8>);
  row("nc.L1::no_allocate U=4", k_ldg<1, 4>); row("nc.L1::no_allocate U=8", k_ldg<1, 8>);
  row("ld.cg U=4", k_ldg<2, 4>); row("ld.cg U=8", k_ldg<2, 8>);
  printf("\n== TMA bulk ring (cp.async.bulk + mbarrier), 1 CTA x 1024 thr per SM; table smem reserved on top of the ring ==\n");
  printf("%-34s %10s %10s\n", "stages x stage_bytes", "extraKB", "GB/s");
  CK(cudaFuncSetAttribute(k_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  for (int stages : {2, 3, 4, 6, 8}) for (uint32_t sb : {8192u, 16384u, 32768u}) {
    size_t ring = 256 + (size_t)stages * sb;
    if (ring > 200 * 1024) continue;
    for (int extra : {0, 1}) {
      size_t smem = extra ? 227 * 1024 : ring;
      float ms = time_ms([&] { k_tma<<<nsm, 1024, smem>>>(a, bytes, stages, sb, 0, out); });
      printf("%d x %-30u %10zu %10.0f\n", stages, sb, (smem - ring) / 1024, bytes / (ms * 1e6));
    }
  }
  printf("done\n");
  return 0;
}

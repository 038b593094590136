// Design microbenchmarks for the B200 hash-groupby path (round 1).
// Measures the on-chip / L2 / HBM primitives the kernel design hinges on:
//   mb_smem   : shared-memory atomic throughput on random slots (ATOMS.ADD.32, CAS-64 loops, LDS.64 probes)
//   mb_red    : global RED.ADD throughput on random slots of a table of T entries (L2- or HBM-resident)
//   mb_probe  : global random 8-byte loads (hash probe)
//   mb_stream : streaming read bandwidth (LDG.128)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/microbench tools/microbench.cu
// Not part of the product; results are recorded in profiles/.
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t lcg(uint32_t& x) { x = x * 1664525u + 1013904223u; return x; }
__device__ __forceinline__ uint64_t splitmix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// ---------------------------------------------------------------- mb_smem
// VAR 0: 1x ATOMS.ADD.32 (no return)
// VAR 1: 2x ATOMS.ADD.32 (sum_lo + count), no return
// VAR 2: LDS.64 probe + ATOMS.ADD.32 with return + ATOMS.ADD.32 (count)  [int64 SUM lo/hi split, hi rarely]
// VAR 3: LDS.64 probe + 64-bit CAS-loop add + ATOMS.ADD.32
// VAR 4: LDS.64 probe only
// VAR 5: LDS.32 only
// VAR 6: 1x ATOMS.ADD.32 with return
template <int VAR>
__global__ void __launch_bounds__(1024) mb_smem(uint32_t tmask, int iters, unsigned long long* out)
{
  extern __shared__ __align__(16) unsigned char smraw[];
  uint32_t T = tmask + 1;
  unsigned long long* keys = reinterpret_cast<unsigned long long*>(smraw);
  uint32_t* lo  = reinterpret_cast<uint32_t*>(keys + T);
  uint32_t* hi  = lo + T;
  uint32_t* cnt = hi + T;
  for (uint32_t i = threadIdx.x; i < T; i += blockDim.x) { keys[i] = i; lo[i] = 0; hi[i] = 0; cnt[i] = 0; }
  __syncthreads();
  uint32_t x = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
  unsigned long long acc = 0;
  #pragma unroll 4
  for (int it = 0; it < iters; ++it) {
    uint32_t r = lcg(x);
    uint32_t s = (r >> 7) & tmask;
    uint32_t v = r & 1023u;
    if (VAR == 0) { atomicAdd(&lo[s], v); }
    if (VAR == 1) { atomicAdd(&lo[s], v); atomicAdd(&cnt[s], 1u); }
    if (VAR == 2) {
      unsigned long long k = keys[s];
      if (k == s) {
        uint32_t old = atomicAdd(&lo[s], v);
        if (old + v < old) atomicAdd(&hi[s], 1u);
        atomicAdd(&cnt[s], 1u);
      } else acc += k;
    }
    if (VAR == 3) {
      unsigned long long k = keys[s];
      if (k == s) {
        atomicAdd(reinterpret_cast<unsigned long long*>(lo) + (s >> 1), (unsigned long long)v);
        atomicAdd(&cnt[s], 1u);
      } else acc += k;
    }
    if (VAR == 4) { acc += keys[s]; }
    if (VAR == 5) { acc += lo[s]; }
    if (VAR == 6) { acc += atomicAdd(&lo[s], v); }
  }
  __syncthreads();
  if (threadIdx.x == 0) acc += lo[1] + cnt[2] + hi[3];
  if (acc == 0x1234567812345678ull) out[0] = acc;
}

// ---------------------------------------------------------------- mb_red
// VAR 0: RED.ADD.64 to sum[s]
// VAR 1: RED.ADD.64 sum[s] + RED.ADD.32 cnt[s] (SoA: two sectors)
// VAR 2: AoS 32-byte slot {key,sum,cnt,pad}: LDG key + RED.64 + RED.32 (same sector)
// VAR 3: AoS 16-byte slot {sum,cnt,pad}: RED.64 + RED.32
// VAR 4: AoS 32-byte slot: RED only (no key load)
template <int VAR>
__global__ void __launch_bounds__(256) mb_red(unsigned long long* tab, uint64_t T, int iters, unsigned long long* out)
{
  uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t st = gid * 0x9E3779B97F4A7C15ull + 1;
  unsigned long long acc = 0;
  for (int it = 0; it < iters; ++it) {
    uint64_t r = splitmix64(st + it);
    uint64_t s = (uint64_t)(((unsigned __int128)r * T) >> 64);
    unsigned long long v = r & 1023;
    if (VAR == 0) { atomicAdd(&tab[s], v); }
    if (VAR == 1) { atomicAdd(&tab[s], v); atomicAdd(reinterpret_cast<unsigned*>(tab + T) + s, 1u); }
    if (VAR == 2) {
      unsigned long long* slot = tab + 4 * s;
      unsigned long long k = *reinterpret_cast<volatile unsigned long long*>(slot);
      if (k == 0) { atomicAdd(slot + 1, v); atomicAdd(reinterpret_cast<unsigned*>(slot + 2), 1u); } else acc += k;
    }
    if (VAR == 3) { unsigned long long* slot = tab + 2 * s; atomicAdd(slot, v); atomicAdd(reinterpret_cast<unsigned*>(slot + 1), 1u); }
    if (VAR == 4) { unsigned long long* slot = tab + 4 * s; atomicAdd(slot + 1, v); atomicAdd(reinterpret_cast<unsigned*>(slot + 2), 1u); }
  }
  if (acc == 0x1234567812345678ull) out[0] = acc;
}

// ---------------------------------------------------------------- mb_probe: random 8B loads
__global__ void __launch_bounds__(256) mb_probe(const unsigned long long* tab, uint64_t T, int iters, unsigned long long* out)
{
  uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t st = gid * 0x9E3779B97F4A7C15ull + 1;
  unsigned long long acc = 0;
  #pragma unroll 4
  for (int it = 0; it < iters; ++it) {
    uint64_t r = splitmix64(st + it);
    uint64_t s = (uint64_t)(((unsigned __int128)r * T) >> 64);
    acc += __ldg(&tab[s]);
  }
  if (acc == 0x1234567812345678ull) out[0] = acc;
}

// ---------------------------------------------------------------- mb_stream: LDG.128 streaming read
__global__ void __launch_bounds__(512) mb_stream(const uint4* __restrict__ a, uint64_t n16, unsigned long long* out)
{
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  unsigned acc = 0;
  for (; i + 3 * stride < n16; i += 4 * stride) {
    uint4 v0 = __ldcs(a + i), v1 = __ldcs(a + i + stride), v2 = __ldcs(a + i + 2 * stride), v3 = __ldcs(a + i + 3 * stride);
    acc += v0.x ^ v0.y ^ v0.z ^ v0.w ^ v1.x ^ v1.y ^ v1.z ^ v1.w ^ v2.x ^ v2.y ^ v2.z ^ v2.w ^ v3.x ^ v3.y ^ v3.z ^ v3.w;
  }
  for (; i < n16; i += stride) { uint4 v = __ldcs(a + i); acc += v.x ^ v.y ^ v.z ^ v.w; }
  if (acc == 0x12345678u) out[0] = acc;
}

// two-column streaming (keys + values), the A-config access pattern without any aggregation
__global__ void __launch_bounds__(512) mb_stream2(const uint4* __restrict__ a, const uint4* __restrict__ b, uint64_t n16, unsigned long long* out)
{
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  unsigned acc = 0;
  for (; i + stride < n16; i += 2 * stride) {
    uint4 v0 = __ldcs(a + i), v1 = __ldcs(a + i + stride), v2 = __ldcs(b + i), v3 = __ldcs(b + i + stride);
    acc += v0.x ^ v0.y ^ v0.z ^ v0.w ^ v1.x ^ v1.y ^ v1.z ^ v1.w ^ v2.x ^ v2.y ^ v2.z ^ v2.w ^ v3.x ^ v3.y ^ v3.z ^ v3.w;
  }
  for (; i < n16; i += stride) { uint4 v = __ldcs(a + i); uint4 w = __ldcs(b + i); acc += v.x ^ v.y ^ v.z ^ v.w ^ w.x; }
  if (acc == 0x12345678u) out[0] = acc;
}

__global__ void fill_rand(uint64_t* a, uint64_t n, uint64_t seed, uint64_t mod)
{
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) a[i] = mod ? splitmix64(seed * 0x1000003ull + i) % mod : splitmix64(seed * 0x1000003ull + i);
}

__global__ void clock_cal(unsigned long long* out)
{
  unsigned long long t0, c0 = clock64();
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  unsigned long long t1 = t0;
  while (t1 - t0 < 2000000ull) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
  unsigned long long c1 = clock64();
  out[0] = c1 - c0; out[1] = t1 - t0;
}

template <typename F>
static float time_ms(F f, int reps = 5)
{
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  f(); CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < reps; ++r) {
    CK(cudaEventRecord(e0)); f(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
  }
  CK(cudaGetLastError());
  return best;
}

int main()
{
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  int nsm = prop.multiProcessorCount;
  printf("device %s, %d SMs, smem/block optin %zu, L2 %d MB\n", prop.name, nsm, prop.sharedMemPerBlockOptin, prop.l2CacheSize >> 20);
  unsigned long long* out; CK(cudaMalloc(&out, 64)); CK(cudaMemset(out, 0, 64));
  double mhz = 0;
  {
    clock_cal<<<1, 1>>>(out); CK(cudaDeviceSynchronize());
    unsigned long long h[2]; CK(cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost));
    mhz = (double)h[0] / (double)h[1] * 1000.0;
    printf("SM clock (1 thread spinning): %.0f MHz\n", mhz);
  }

  // ---------------- smem atomics
  printf("\n== mb_smem: lane-ops per ns per SM (threads=1024, 1 CTA/SM) ==\n");
  printf("%-44s %8s %10s %12s\n", "variant", "T", "ms", "rows/ns/SM");
  const char* names[] = {"0: 1xATOMS.ADD32", "1: 2xATOMS.ADD32", "2: LDS64+ATOMS.ret+ATOMS (i64 sum lo/hi+cnt)", "3: LDS64+CAS64add+ATOMS", "4: LDS64", "5: LDS32", "6: 1xATOMS.ADD32 ret"};
  for (uint32_t T : {1024u, 4096u, 8192u}) {
    size_t smem = (size_t)T * 20;
    int iters = 4096;
    auto run = [&](auto kern, int var) {
      CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      for (int threads : {1024, 512}) {
        float ms = time_ms([&] { kern<<<nsm, threads, smem>>>(T - 1, iters, out); });
        double rows = (double)nsm * threads * iters;
        printf("%-44s %8u %10.4f %12.3f  (thr=%d)\n", names[var], T, ms, rows / (ms * 1e6) / nsm, threads);
      }
    };
    run(mb_smem<0>, 0); run(mb_smem<1>, 1); run(mb_smem<2>, 2); run(mb_smem<3>, 3); run(mb_smem<4>, 4); run(mb_smem<5>, 5); run(mb_smem<6>, 6);
  }

  // ---------------- global RED
  printf("\n== mb_red: random global atomics, G rows/s chip-wide ==\n");
  printf("%-44s %12s %10s %12s\n", "variant", "T", "ms", "Grows/s");
  const char* rnames[] = {"0: RED64 sum", "1: RED64 sum + RED32 cnt (SoA)", "2: AoS32 LDG key + RED64 + RED32", "3: AoS16 RED64+RED32", "4: AoS32 RED64+RED32"};
  for (uint64_t T : {10000ull, 1000000ull, 16000000ull, 134217728ull}) {
    unsigned long long* tab; size_t bytes = T * 32; CK(cudaMalloc(&tab, bytes)); CK(cudaMemset(tab, 0, bytes));
    int iters = T > 20000000ull ? 32 : 64;
    int blocks = nsm * 8 * 4, threads = 256;
    auto run = [&](auto kern, int var) {
      float ms = time_ms([&] { kern<<<blocks, threads>>>(tab, T, iters, out); }, 3);
      double rows = (double)blocks * threads * iters;
      printf("%-44s %12llu %10.4f %12.2f\n", rnames[var], (unsigned long long)T, ms, rows / (ms * 1e6));
    };
    run(mb_red<0>, 0); run(mb_red<1>, 1); run(mb_red<2>, 2); run(mb_red<3>, 3); run(mb_red<4>, 4);
    {
      float ms = time_ms([&] { mb_probe<<<blocks, threads>>>(tab, T * 4, iters, out); }, 3);
      double rows = (double)blocks * threads * iters;
      printf("%-44s %12llu %10.4f %12.2f\n", "probe: random LDG.64 over 32B*T bytes", (unsigned long long)T, ms, rows / (ms * 1e6));
    }
    CK(cudaFree(tab));
  }

  // ---------------- streaming
  printf("\n== mb_stream ==\n");
  {
    uint64_t n = 200000000ull;  // 1.6 GB
    uint64_t *a, *b; CK(cudaMalloc(&a, n * 8)); CK(cudaMalloc(&b, n * 8));
    fill_rand<<<nsm * 8, 256>>>(a, n, 1, 0); fill_rand<<<nsm * 8, 256>>>(b, n, 2, 0); CK(cudaDeviceSynchronize());
    for (int bps : {2, 4, 8}) {
      float ms = time_ms([&] { mb_stream<<<nsm * bps, 512>>>((const uint4*)a, n / 2, out); });
      printf("stream LDG.128 1.6GB  grid=%d x512: %.4f ms  %.1f GB/s\n", nsm * bps, ms, n * 8 / (ms * 1e6));
      ms = time_ms([&] { mb_stream2<<<nsm * bps, 512>>>((const uint4*)a, (const uint4*)b, n / 2, out); });
      printf("stream2 (2 cols) 3.2GB grid=%d x512: %.4f ms  %.1f GB/s\n", nsm * bps, ms, 2 * n * 8 / (ms * 1e6));
    }
    CK(cudaFree(a)); CK(cudaFree(b));
  }
  printf("done\n");
  return 0;
}

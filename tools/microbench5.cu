// One-pass radix scatter into 2^b fixed-capacity buckets with the L2 as the write combiner: every row is one returning
// atomicAdd on its bucket's cursor + one 16-byte {key, value} record store at the cursor position. Question: what does a
// single pass to 2^14..2^16 buckets cost per 1e9 rows, against two smem-staged levels of <= 9 bits (9.8 + 9.1 ms in r01)?
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench5 tools/microbench5.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <stdint.h>

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z)
{
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__device__ __forceinline__ uint32_t slot_hash32(uint64_t k)
{
  const uint64_t h = k * 0x9E3779B97F4A7C15ull;
  uint32_t x = (uint32_t)(h >> 32) ^ (uint32_t)h;
  x *= 0x85EBCA6Bu; x ^= x >> 13;
  x *= 0xC2B2AE35u; x ^= x >> 16;
  return x;
}

__global__ void gen(uint64_t* k, uint64_t* v, int64_t n, uint64_t g)
{
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    k[i] = mix64(42 * 0x1000003ull + i + 0x9E3779B97F4A7C15ull) % g;
    v[i] = mix64(43 * 0x1000003ull + i + 0x9E3779B97F4A7C15ull) % 2001 - 1000;
  }
}

// MODE 0: per-row atomics, 2 rows per thread per iteration (LDG.128 of keys and values)
// MODE 1: 4 rows per thread per iteration
template <int ROWS>
__global__ void __launch_bounds__(512) scatter_direct(const uint64_t* __restrict__ k, const uint64_t* __restrict__ v, int64_t n, int bits,
                                                      uint32_t* __restrict__ cursor, uint32_t cap, ulonglong2* __restrict__ out)
{
  const int64_t stride = (int64_t)gridDim.x * blockDim.x * ROWS;
  for (int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 2; i0 < n; i0 += stride) {
    ulonglong2 kk[ROWS / 2], vv[ROWS / 2];
#pragma unroll
    for (int j = 0; j < ROWS / 2; ++j) {
      const int64_t i = i0 + (int64_t)j * gridDim.x * blockDim.x * 2;
      if (i < n) { kk[j] = __ldcs(reinterpret_cast<const ulonglong2*>(k + i)); vv[j] = __ldcs(reinterpret_cast<const ulonglong2*>(v + i)); }
    }
#pragma unroll
    for (int j = 0; j < ROWS / 2; ++j) {
      const int64_t i = i0 + (int64_t)j * gridDim.x * blockDim.x * 2;
      if (i < n) {
        const uint32_t b0 = slot_hash32(kk[j].x) >> (32 - bits), b1 = slot_hash32(kk[j].y) >> (32 - bits);
        const uint32_t p0 = atomicAdd(cursor + b0, 1u);
        const uint32_t p1 = atomicAdd(cursor + b1, 1u);
        if (p0 < cap) out[(uint64_t)b0 * cap + p0] = make_ulonglong2(kk[j].x, vv[j].x);
        if (p1 < cap) out[(uint64_t)b1 * cap + p1] = make_ulonglong2(kk[j].y, vv[j].y);
      }
    }
  }
}

// consumer side: stream every bucket's records (LDG.128 per row) -- the read the slice kernel would do
__global__ void __launch_bounds__(512) read_buckets(const ulonglong2* __restrict__ in, const uint32_t* __restrict__ cursor, uint32_t cap, uint32_t nb,
                                                    uint64_t* sink)
{
  uint64_t acc = 0;
  for (uint32_t b = blockIdx.x; b < nb; b += gridDim.x) {
    uint32_t n = cursor[b]; n = n < cap ? n : cap;
    const ulonglong2* p = in + (uint64_t)b * cap;
    for (uint32_t r = threadIdx.x; r < n; r += blockDim.x) { ulonglong2 x = __ldcs(p + r); acc += x.x ^ x.y; }
  }
  if (acc == 0x1234567) sink[0] = acc;
}

int main(int argc, char** argv)
{
  const int64_t n = argc > 1 ? atoll(argv[1]) : 1000000000ll;
  const uint64_t g = argc > 2 ? strtoull(argv[2], 0, 10) : 100000000ull;
  uint64_t *k, *v, *sink;
  cudaMalloc(&k, n * 8); cudaMalloc(&v, n * 8); cudaMalloc(&sink, 8);
  gen<<<148 * 8, 256>>>(k, v, n, g);
  cudaDeviceSynchronize();
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int bits : {12, 14, 15, 16}) {
    const uint32_t nb = 1u << bits;
    const uint32_t cap = (uint32_t)((double)n / nb * 1.10) + 256;
    uint32_t* cursor; ulonglong2* out;
    cudaMalloc(&cursor, nb * 4);
    if (cudaMalloc(&out, (size_t)nb * cap * 16) != cudaSuccess) { printf("alloc failed\n"); return 1; }
    for (int rows : {2, 4}) for (int cps : {2, 4}) {
      float best = 1e9;
      for (int rep = 0; rep < 3; ++rep) {
        cudaMemset(cursor, 0, nb * 4);
        cudaEventRecord(e0);
        if (rows == 2) scatter_direct<2><<<148 * cps, 512>>>(k, v, n, bits, cursor, cap, out);
        else scatter_direct<4><<<148 * cps, 512>>>(k, v, n, bits, cursor, cap, out);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); best = ms < best ? ms : best;
      }
      printf("direct scatter bits=%2d rows/thread=%d ctas/sm=%d: %7.3f ms  (%.2f TB/s of 32 B/row)  err=%s\n", bits, rows, cps, best,
             32.0 * n / best / 1e9, cudaGetErrorString(cudaGetLastError()));
    }
    // check: the cursors add up to n and no bucket overflowed
    uint32_t* h = (uint32_t*)malloc(nb * 4);
    cudaMemcpy(h, cursor, nb * 4, cudaMemcpyDeviceToHost);
    uint64_t tot = 0; uint32_t mx = 0;
    for (uint32_t i = 0; i < nb; ++i) { tot += h[i]; mx = h[i] > mx ? h[i] : mx; }
    printf("   rows placed %llu of %lld, fullest bucket %u of capacity %u\n", (unsigned long long)tot, (long long)n, mx, cap);
    free(h);
    float best = 1e9;
    for (int rep = 0; rep < 3; ++rep) {
      cudaEventRecord(e0);
      read_buckets<<<148 * 4, 512>>>(out, cursor, cap, nb, sink);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1); best = ms < best ? ms : best;
    }
    printf("   read back all buckets: %7.3f ms (%.2f TB/s)\n", best, 16.0 * n / best / 1e9);
    cudaFree(cursor); cudaFree(out);
  }
  return 0;
}

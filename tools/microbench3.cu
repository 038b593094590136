// FP64 pipe rate on this GPU: independent DFMA / DADD chains per thread, many warps per SM.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench3 tools/microbench3.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void fp64_kernel(double* out, int iters, double seed)
{
  double a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = seed + i + threadIdx.x;
  const double m = 1.0000001, c = 0.5;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) a[i] = fma(a[i], m, c);
      else if (MODE == 1) a[i] = a[i] + c;
      else a[i] = a[i] * m;
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void fp32_kernel(float* out, int iters, float seed)
{
  float a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = seed + i + threadIdx.x;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = fmaf(a[i], 1.0000001f, 0.5f);
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void i64_kernel(unsigned long long* out, int iters, unsigned long long seed)
{
  unsigned long long a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = seed + i + threadIdx.x;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = a[i] * 3ull + a[(i + 1) & 7];
  }
  unsigned long long s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main()
{
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int blocks = sms * 8, threads = 256, iters = 4096;
  void* out;
  cudaMalloc(&out, (size_t)blocks * threads * 8);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const char* names[] = {"DFMA", "DADD", "DMUL", "FFMA", "IMAD64"};
  for (int mode = 0; mode < 5; ++mode) {
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      if (mode == 0) fp64_kernel<0><<<blocks, threads>>>((double*)out, iters, 1.0);
      else if (mode == 1) fp64_kernel<1><<<blocks, threads>>>((double*)out, iters, 1.0);
      else if (mode == 2) fp64_kernel<2><<<blocks, threads>>>((double*)out, iters, 1.0);
      else if (mode == 3) fp32_kernel<<<blocks, threads>>>((float*)out, iters, 1.0f);
      else i64_kernel<<<blocks, threads>>>((unsigned long long*)out, iters, 1ull);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
    }
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double ops = (double)blocks * threads * iters * 8;
    printf("%-7s %8.3f ms  %8.1f Gop/s  = %.2f lanes/clk/SM at 1.965 GHz\n", names[mode], ms, ops / ms / 1e6, ops / (ms * 1e-3) / (sms * 1.965e9));
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}

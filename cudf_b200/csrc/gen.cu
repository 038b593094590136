// Synthetic column generators: counter-based SplitMix64 x(seed,row), bit-identical to
// oracle/groupby_oracle.c orc_splitmix64 so CPU and GPU build the same columns. Bench + tests only.
#include "bgb_internal.cuh"
#include "errors.h"

namespace bgb {

template <typename T>
__global__ void __launch_bounds__(256) gen_int_kernel(T* out, int64_t n, uint64_t seed, uint64_t first, uint64_t mod, int64_t add)
{
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint64_t x = splitmix64(seed, first + (uint64_t)i);
    if (mod) x %= mod;
    out[i] = (T)((int64_t)x + add);
  }
}
template <typename T>
__global__ void __launch_bounds__(256) gen_real_kernel(T* out, int64_t n, uint64_t seed, uint64_t first, uint64_t mod, double add)
{
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint64_t x = splitmix64(seed, first + (uint64_t)i);
    if (mod) x %= mod;
    out[i] = (T)((double)x + add);
  }
}
__global__ void __launch_bounds__(256) gen_mask_kernel(uint32_t* words, int64_t n, uint64_t seed, uint64_t first, uint32_t every,
                                                       int* null_count)
{
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t padded = (n + 31) & ~31ll;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < padded; i += stride) {
    bool in    = i < n;
    bool valid = in && (splitmix64(seed, first + (uint64_t)i) % every) != 0;
    uint32_t w = __ballot_sync(0xffffffffu, valid);
    uint32_t a = __ballot_sync(0xffffffffu, in);
    if ((threadIdx.x & 31) == 0) { words[i >> 5] = w; int z = __popc(a & ~w); if (z) atomicAdd(null_count, z); }
  }
}
__global__ void __launch_bounds__(256) gen_cat_kernel(int8_t* out, int64_t n, uint64_t seed, uint64_t first, int ncat,
                                                      const int32_t* cum, const int8_t* vals)
{
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint32_t x = (uint32_t)(splitmix64(seed, first + (uint64_t)i) & 0xFFFF);
    int c = ncat - 1;
    for (int j = 0; j < ncat; ++j) if ((int32_t)x < cum[j]) { c = j; break; }
    out[i] = vals[c];
  }
}

static int grid_for(int64_t n) { int64_t b = (n + 255) / 256; return (int)(b < device_sm_count() * 16 ? (b > 0 ? b : 1) : device_sm_count() * 16); }

}  // namespace bgb

using namespace bgb;

extern "C" int bgb_gen_uniform_int(void* out, int32_t type_id, int64_t n, uint64_t seed, uint64_t first_row, uint64_t modulus,
                                   int64_t add, void* stream)
{
  cudaStream_t s = (cudaStream_t)stream;
  if (n <= 0) return BGB_OK;
  int g = grid_for(n);
  switch (type_id) {
    case BGB_INT8: case BGB_UINT8: case BGB_BOOL8: gen_int_kernel<int8_t><<<g, 256, 0, s>>>((int8_t*)out, n, seed, first_row, modulus, add); break;
    case BGB_INT16: case BGB_UINT16: gen_int_kernel<int16_t><<<g, 256, 0, s>>>((int16_t*)out, n, seed, first_row, modulus, add); break;
    case BGB_INT32: case BGB_UINT32: gen_int_kernel<int32_t><<<g, 256, 0, s>>>((int32_t*)out, n, seed, first_row, modulus, add); break;
    case BGB_INT64: case BGB_UINT64: gen_int_kernel<int64_t><<<g, 256, 0, s>>>((int64_t*)out, n, seed, first_row, modulus, add); break;
    default: return set_error(BGB_INVALID_ARGUMENT, "bgb_gen_uniform_int: not an integer type");
  }
  return check_cuda(cudaGetLastError(), "gen_int");
}

extern "C" int bgb_gen_uniform_real(void* out, int32_t type_id, int64_t n, uint64_t seed, uint64_t first_row, uint64_t modulus,
                                    double add, void* stream)
{
  cudaStream_t s = (cudaStream_t)stream;
  if (n <= 0) return BGB_OK;
  int g = grid_for(n);
  if (type_id == BGB_FLOAT64) gen_real_kernel<double><<<g, 256, 0, s>>>((double*)out, n, seed, first_row, modulus, add);
  else if (type_id == BGB_FLOAT32) gen_real_kernel<float><<<g, 256, 0, s>>>((float*)out, n, seed, first_row, modulus, add);
  else return set_error(BGB_INVALID_ARGUMENT, "bgb_gen_uniform_real: not a floating type");
  return check_cuda(cudaGetLastError(), "gen_real");
}

extern "C" int bgb_gen_null_mask(uint32_t* mask_words, int64_t n, uint64_t seed, uint64_t first_row, uint32_t null_every,
                                 int32_t* null_count, void* stream)
{
  cudaStream_t s = (cudaStream_t)stream;
  if (null_every == 0) return set_error(BGB_INVALID_ARGUMENT, "null_every must be > 0");
  int* d_cnt = nullptr;
  int rc = check_cuda(cudaMallocAsync(&d_cnt, sizeof(int), s), "alloc");
  if (rc) return rc;
  cudaMemsetAsync(d_cnt, 0, sizeof(int), s);
  if (n > 0) gen_mask_kernel<<<grid_for(n), 256, 0, s>>>(mask_words, n, seed, first_row, null_every, d_cnt);
  int h = 0;
  cudaMemcpyAsync(&h, d_cnt, sizeof(int), cudaMemcpyDeviceToHost, s);
  cudaFreeAsync(d_cnt, s);
  rc = check_cuda(cudaStreamSynchronize(s), "gen_mask");
  if (null_count) *null_count = h;
  return rc;
}

extern "C" int bgb_gen_categorical_i8(int8_t* out, int64_t n, uint64_t seed, uint64_t first_row, const int32_t* cum_host,
                                      const int8_t* values_host, int32_t num_categories, void* stream)
{
  cudaStream_t s = (cudaStream_t)stream;
  if (num_categories <= 0 || num_categories > 64) return set_error(BGB_INVALID_ARGUMENT, "1..64 categories");
  int32_t* d_cum = nullptr; int8_t* d_val = nullptr;
  int rc = check_cuda(cudaMallocAsync(&d_cum, 64 * sizeof(int32_t), s), "alloc");
  if (rc) return rc;
  rc = check_cuda(cudaMallocAsync(&d_val, 64, s), "alloc");
  if (rc) return rc;
  cudaMemcpyAsync(d_cum, cum_host, num_categories * sizeof(int32_t), cudaMemcpyHostToDevice, s);
  cudaMemcpyAsync(d_val, values_host, num_categories, cudaMemcpyHostToDevice, s);
  if (n > 0) gen_cat_kernel<<<grid_for(n), 256, 0, s>>>(out, n, seed, first_row, num_categories, d_cum, d_val);
  cudaFreeAsync(d_cum, s); cudaFreeAsync(d_val, s);
  return check_cuda(cudaStreamSynchronize(s), "gen_cat");
}

// bgb_bytes_to_bitmask_noalloc: the validity-bytes -> bitmask kernel for callers that cannot allocate (code under CUDA graph
// capture, serving loops, harnesses that forbid cudaMalloc). It has NO counterpart in the reference.
// The existing entry point (partition_api.cu: bgb_bytes_to_bitmask) takes its null counter
// from the library's memory pool; this variant keeps ALL of its scratch -- a ticket, the total, and one partial null count
// per 8192-row tile -- in a file-scope __device__ array, so it calls no cudaMalloc / cudaFree / pool routine at all.
//
// Every GPU that loads the library holds the array (kNoAllocScratchBytes = 262 400 bytes) until the process exits. All calls
// on one device share it: the header is cleared at the start of every call and the partials are overwritten before they are
// read, so back-to-back calls on one stream are safe, but two calls that OVERLAP on the GPU (different streams, no ordering
// between them) race on it.
#include <mutex>

#include "errors.h"

using namespace bgb;

namespace {

constexpr int kTileRows      = 8192;   // 256 threads x 32 rows
constexpr int kMaxTiles      = 65536;  // => at most kTileRows * kMaxTiles = 536 870 912 rows per call
constexpr int kHeaderWords   = 64;     // [0] ticket of finished tiles, [1] total null count
constexpr size_t kNoAllocScratchBytes = sizeof(int) * (size_t)(kHeaderWords + kMaxTiles);

// In HOST code this name is only a stand-in for the symbol (not a device address): the launcher below resolves it with
// cudaGetSymbolAddress once per device and passes that pointer to the kernel.
static __device__ __align__(256) int g_noalloc_scratch[kHeaderWords + kMaxTiles];

__global__ void __launch_bounds__(256) bytes_to_bitmask_tiles_kernel(const uint8_t* __restrict__ bytes, int64_t n, uint32_t* __restrict__ words,
                                                                     int* __restrict__ scratch)
{
  __shared__ int warp_nulls[8];
  __shared__ int is_last;
  const int64_t base = (int64_t)blockIdx.x * kTileRows;
  int nulls = 0;
#pragma unroll 4
  for (int j = 0; j < kTileRows / 256; ++j) {
    const int64_t i  = base + (int64_t)j * 256 + threadIdx.x;
    const bool in    = i < n;
    const bool valid = in && bytes[i] != 0;
    const uint32_t w = __ballot_sync(0xffffffffu, valid);
    const uint32_t a = __ballot_sync(0xffffffffu, in);
    if ((threadIdx.x & 31) == 0 && a) {
      words[i >> 5] = w;
      nulls += __popc(a & ~w);
    }
  }
  if ((threadIdx.x & 31) == 0) warp_nulls[threadIdx.x >> 5] = nulls;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < 8; ++w) t += warp_nulls[w];
    scratch[kHeaderWords + blockIdx.x] = t;
    __threadfence();
    is_last = atomicAdd(&scratch[0], 1) == (int)gridDim.x - 1;
  }
  __syncthreads();
  if (!is_last) return;
  // the last tile to finish adds the partials up in tile order (a fixed order: the count does not depend on scheduling)
  __threadfence();
  int t = 0;
  for (int i = threadIdx.x; i < (int)gridDim.x; i += 256) t += *reinterpret_cast<volatile int*>(&scratch[kHeaderWords + i]);
  for (int d = 16; d > 0; d >>= 1) t += __shfl_down_sync(0xffffffffu, t, d);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) warp_nulls[threadIdx.x >> 5] = t;
  __syncthreads();
  if (threadIdx.x == 0) {
    int total = 0;
    for (int w = 0; w < 8; ++w) total += warp_nulls[w];
    scratch[1] = total;
  }
}

// device address of the scratch array, resolved once per device with that device current
int* scratch_for_current_device(int& rc)
{
  constexpr int kMaxDevices = 64;
  static std::mutex mu;
  static int* ptr[kMaxDevices] = {};
  int dev = 0;
  rc = check_cuda(cudaGetDevice(&dev), "cudaGetDevice");
  if (rc) return nullptr;
  if (dev < 0 || dev >= kMaxDevices) { rc = set_error(BGB_UNSUPPORTED, "bgb_bytes_to_bitmask_noalloc: device ordinal %d", dev); return nullptr; }
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[dev]) {
    void* p = nullptr;
    rc = check_cuda(cudaGetSymbolAddress(&p, g_noalloc_scratch), "cudaGetSymbolAddress(g_noalloc_scratch)");
    if (rc) return nullptr;
    ptr[dev] = (int*)p;
  }
  return ptr[dev];
}

}  // namespace

extern "C" int64_t bgb_bytes_to_bitmask_noalloc_max_rows(void) { return (int64_t)kTileRows * kMaxTiles; }

extern "C" int bgb_bytes_to_bitmask_noalloc(const uint8_t* valid_bytes, int64_t n, uint32_t* mask_words, int32_t* null_count_host, void* stream)
{
  cudaStream_t s = (cudaStream_t)stream;
  if (null_count_host) *null_count_host = 0;
  if (n <= 0) return BGB_OK;
  if (!valid_bytes || !mask_words) return set_error(BGB_INVALID_ARGUMENT, "bgb_bytes_to_bitmask_noalloc: null pointer");
  if (n > (int64_t)kTileRows * kMaxTiles)
    return set_error(BGB_INVALID_ARGUMENT, "bgb_bytes_to_bitmask_noalloc: %lld rows exceed the %lld rows the static scratch (%zu bytes) covers; "
                     "use bgb_bytes_to_bitmask or split the call", (long long)n, (long long)kTileRows * kMaxTiles, kNoAllocScratchBytes);
  int rc = BGB_OK;
  int* scratch = scratch_for_current_device(rc);
  if (rc) return rc;
  // the array keeps the previous call's contents: ticket and total must start at zero
  rc = check_cuda(cudaMemsetAsync(scratch, 0, sizeof(int) * kHeaderWords, s), "noalloc header clear");
  if (rc) return rc;
  const int tiles = (int)((n + kTileRows - 1) / kTileRows);
  bytes_to_bitmask_tiles_kernel<<<tiles, 256, 0, s>>>(valid_bytes, n, mask_words, scratch);
  rc = check_cuda(note_launch(), "bytes_to_bitmask_noalloc");
  if (rc || !null_count_host) return rc;  // without a host count the call is fully asynchronous on `stream`
  int h = 0;
  rc = check_cuda(cudaMemcpyAsync(&h, scratch + 1, sizeof(int), cudaMemcpyDeviceToHost, s), "null count copy");
  if (!rc) rc = check_cuda(cudaStreamSynchronize(s), "bytes_to_bitmask_noalloc sync");
  *null_count_host = h;
  return rc;
}

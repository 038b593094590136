#include "pool.h"

#include <mutex>
#include <stdint.h>

namespace bgb {

static cudaMemPool_t device_pool()
{
  static std::mutex mu;
  static cudaMemPool_t pools[64] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  std::lock_guard<std::mutex> lock(mu);
  if (!pools[dev]) {
    cudaMemPoolProps props{};
    props.allocType     = cudaMemAllocationTypePinned;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id   = dev;
    if (cudaMemPoolCreate(&pools[dev], &props) != cudaSuccess) { cudaGetLastError(); pools[dev] = nullptr; return nullptr; }
    uint64_t keep = ~0ull;
    cudaMemPoolSetAttribute(pools[dev], cudaMemPoolAttrReleaseThreshold, &keep);
  }
  return pools[dev];
}

void* pool_alloc_async(size_t bytes, cudaStream_t s)
{
  void* p            = nullptr;
  cudaMemPool_t pool = device_pool();
  cudaError_t e      = pool ? cudaMallocFromPoolAsync(&p, bytes ? bytes : 16, pool, s) : cudaMallocAsync(&p, bytes ? bytes : 16, s);
  if (e != cudaSuccess) { cudaGetLastError(); return nullptr; }
  return p;
}

void pool_free_async(void* p, cudaStream_t s) { cudaFreeAsync(p, s); }

}  // namespace bgb

#include "errors.h"

#include <cstdarg>
#include <atomic>
#include <cstdio>

namespace bgb {
static thread_local char g_msg[512] = "";

int set_error(int status, const char* fmt, ...)
{
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_msg, sizeof g_msg, fmt, ap);
  va_end(ap);
  return status;
}
int check_cuda(cudaError_t e, const char* what)
{
  if (e == cudaSuccess) return BGB_OK;
  snprintf(g_msg, sizeof g_msg, "CUDA error in %s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
  return e == cudaErrorMemoryAllocation ? BGB_OUT_OF_MEMORY : BGB_CUDA_ERROR;
}
const char* last_error() { return g_msg; }
static std::atomic<long long> g_launches{0};
cudaError_t note_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); return cudaGetLastError(); }
long long launches() { return g_launches.load(); }
int device_sm_count()
{
  static std::atomic<int> cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  int n = cached[dev].load(std::memory_order_relaxed);
  if (n == 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) { cudaGetLastError(); n = 148; }
    cached[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}
}  // namespace bgb

extern "C" const char* bgb_last_error(void) { return bgb::last_error(); }
namespace bgb { long long launches(); }
extern "C" int64_t bgb_kernel_launches(void) { return bgb::launches(); }
extern "C" int bgb_version(void) { return BGB_VERSION; }
namespace bgb { void pool_trim(); }
extern "C" void bgb_release_cached_memory(void) { bgb::pool_trim(); }

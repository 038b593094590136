// NVTX ranges in the domain "cudf_b200": one per C-ABI call and one per kernel phase, so that a timeline tool shows the same
// structure the reference's CUDF_FUNC_RANGE() annotations give (/root/reference/cpp/include/cudf/utilities/nvtx/ranges.hpp; domain
// "libcudf" there). Header-only NVTX3: without a tool attached a push/pop is one load and one predictable branch.
#pragma once
#include <nvtx3/nvToolsExt.h>

namespace bgb {

inline nvtxDomainHandle_t nvtx_domain()
{
  static nvtxDomainHandle_t d = nvtxDomainCreateA("cudf_b200");
  return d;
}
inline void nvtx_push(const char* name)
{
  nvtxEventAttributes_t a{};
  a.version       = NVTX_VERSION;
  a.size          = NVTX_EVENT_ATTRIB_STRUCT_SIZE;
  a.messageType   = NVTX_MESSAGE_TYPE_ASCII;
  a.message.ascii = name;
  nvtxDomainRangePushEx(nvtx_domain(), &a);
}
inline void nvtx_pop() { nvtxDomainRangePop(nvtx_domain()); }
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtx_push(name); }
  ~NvtxRange() { nvtx_pop(); }
  NvtxRange(const NvtxRange&)            = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};

}  // namespace bgb

// Optional per-kernel CUDA-event timing on the call's stream (bgb_options.profile; read back by bgb_groupby_profile).
#pragma once
#include <cuda_runtime.h>

#include <vector>

namespace bgb {

struct Profiler {
  struct Entry { const char* name; cudaEvent_t e0, e1; float ms; };
  std::vector<Entry> entries;
  bool on = false;
  cudaStream_t stream = nullptr;

  void begin(const char* name)
  {
    if (!on) return;
    Entry p{name, nullptr, nullptr, -1.f};
    cudaEventCreate(&p.e0); cudaEventCreate(&p.e1);
    cudaEventRecord(p.e0, stream);
    entries.push_back(p);
  }
  void end() { if (on && !entries.empty()) cudaEventRecord(entries.back().e1, stream); }
  void resolve()  // call after the stream has been synchronised
  {
    for (Entry& p : entries)
      if (p.ms < 0 && p.e0) {
        if (cudaEventElapsedTime(&p.ms, p.e0, p.e1) != cudaSuccess) { cudaGetLastError(); p.ms = 0; }
        cudaEventDestroy(p.e0); cudaEventDestroy(p.e1); p.e0 = p.e1 = nullptr;
      }
  }
  ~Profiler() { for (Entry& p : entries) if (p.e0) { cudaEventDestroy(p.e0); cudaEventDestroy(p.e1); } }
};

}  // namespace bgb

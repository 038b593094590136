// Optional per-kernel CUDA-event timing on the call's stream (bgb_options.profile; read back by bgb_groupby_profile).
#pragma once
#include <cuda_runtime.h>

#include <map>
#include <mutex>
#include <vector>

#include "nvtx.h"

namespace bgb {

// Events are recycled per device: cudaEventCreate/Destroy cost several microseconds each, and a profiled call brackets
// 5-10 kernels (the bench reads its roofline from these events INSIDE the timed steps, so their cost is billed to the step).
struct EventPool {
  std::mutex mu;
  std::map<int, std::vector<cudaEvent_t>> free_events;
  static EventPool& get() { static EventPool* p = new EventPool(); return *p; }  // leaked on purpose: no teardown-order issues
  cudaEvent_t acquire(int dev)
  {
    {
      std::lock_guard<std::mutex> g(mu);
      auto& v = free_events[dev];
      if (!v.empty()) { cudaEvent_t e = v.back(); v.pop_back(); return e; }
    }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
  }
  void release(int dev, cudaEvent_t e)
  {
    if (!e) return;
    std::lock_guard<std::mutex> g(mu);
    free_events[dev].push_back(e);
  }
};

struct Profiler {
  struct Entry { const char* name; cudaEvent_t e0, e1; float ms; };
  std::vector<Entry> entries;
  bool on = false;
  cudaStream_t stream = nullptr;
  int dev = -1;

  bool ranged = false;  // an NVTX range of a kernel phase is open

  void begin(const char* name)
  {
    nvtx_push(name);
    ranged = true;
    if (!on) return;
    if (dev < 0) cudaGetDevice(&dev);
    Entry p{name, EventPool::get().acquire(dev), EventPool::get().acquire(dev), -1.f};
    cudaEventRecord(p.e0, stream);
    entries.push_back(p);
  }
  void end()
  {
    if (ranged) { nvtx_pop(); ranged = false; }
    if (on && !entries.empty()) cudaEventRecord(entries.back().e1, stream);
  }
  void resolve()  // call after the stream has been synchronised
  {
    for (Entry& p : entries)
      if (p.ms < 0 && p.e0) {
        if (cudaEventElapsedTime(&p.ms, p.e0, p.e1) != cudaSuccess) { cudaGetLastError(); p.ms = 0; }
        EventPool::get().release(dev, p.e0); EventPool::get().release(dev, p.e1); p.e0 = p.e1 = nullptr;
      }
  }
  ~Profiler() { if (ranged) nvtx_pop(); for (Entry& p : entries) if (p.e0) { EventPool::get().release(dev, p.e0); EventPool::get().release(dev, p.e1); } }
};

}  // namespace bgb

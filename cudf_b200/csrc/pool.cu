// Stream-ordered scratch memory of the library (group tables, partition copies, scan state).
//
// Two levels. (1) A library-private cudaMemPool with an unlimited release threshold: the default pool hands memory back
// to the OS at every synchronisation, which cost 100s of ms per call at 1e8 groups. (2) In front of it, a per-device
// cache of freed blocks keyed by the stream they were freed on: a repeated query re-uses exactly the blocks of the
// previous call without entering the driver. Without it, calls that interleave differently-sized multi-GB allocations
// (partition copies, then the merge table of the multi-GPU path) made the driver pool re-map physical memory on every
// call (15-290 ms per call measured on B200). Re-use on the SAME stream needs no synchronisation (stream order);
// blocks are never handed to another stream.
#include "pool.h"

#include <mutex>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <unordered_map>
#include <vector>

namespace bgb {
namespace {

struct FreeBlock { void* p; size_t bytes; };

struct DeviceCache {
  cudaMemPool_t pool = nullptr;
  bool pool_tried    = false;
  std::unordered_map<cudaStream_t, std::vector<FreeBlock>> free_blocks;
  std::unordered_map<void*, size_t> live;  // size of every block handed out
  size_t cached_bytes = 0;
  size_t limit_bytes  = 0;
};

std::mutex g_mu;
DeviceCache g_cache[64];
// diagnostics (BGB_TRACE): how often the block cache missed and how often the driver ran out of memory
size_t g_hits = 0, g_driver_allocs = 0, g_oom_flushes = 0, g_driver_bytes = 0;
struct PoolReport { ~PoolReport() { if (getenv("BGB_TRACE")) fprintf(stderr, "[bgb pool] cache hits %zu, driver allocations %zu (%.1f GB), out-of-memory flushes %zu\n", g_hits, g_driver_allocs, g_driver_bytes / 1e9, g_oom_flushes); } } g_report;

DeviceCache* current_cache()
{
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  DeviceCache& c = g_cache[dev];
  if (!c.pool_tried) {
    c.pool_tried = true;
    cudaMemPoolProps props{};
    props.allocType     = cudaMemAllocationTypePinned;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id   = dev;
    if (cudaMemPoolCreate(&c.pool, &props) != cudaSuccess) { cudaGetLastError(); c.pool = nullptr; }
    else {
      // BGB_POOL_KEEP_MB: bytes the driver pool may keep reserved across synchronisations (default: everything); a process that
      // shares the device with other allocators near its capacity sets a bound so that unused pool memory goes back to the driver
      uint64_t keep = ~0ull;
      if (const char* e = getenv("BGB_POOL_KEEP_MB")) keep = (uint64_t)atoll(e) << 20;
      cudaMemPoolSetAttribute(c.pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) { cudaGetLastError(); total_b = 64ull << 30; }
    c.limit_bytes = total_b / 10 * 4;  // keep at most 40 % of the device cached
    if (const char* e = getenv("BGB_CACHE_LIMIT_MB")) c.limit_bytes = (size_t)atoll(e) << 20;
  }
  return &c;
}

void* driver_alloc(DeviceCache& c, size_t bytes, cudaStream_t s)
{
  void* p       = nullptr;
  cudaError_t e = c.pool ? cudaMallocFromPoolAsync(&p, bytes, c.pool, s) : cudaMallocAsync(&p, bytes, s);
  if (e != cudaSuccess) { cudaGetLastError(); return nullptr; }
  return p;
}

// give every cached block back to the driver (memory pressure)
void flush_locked(DeviceCache& c)
{
  for (auto& kv : c.free_blocks) {
    for (FreeBlock& b : kv.second) cudaFreeAsync(b.p, kv.first);
    kv.second.clear();
  }
  c.cached_bytes = 0;
}

}  // namespace

void* pool_alloc_async(size_t bytes, cudaStream_t s, bool transient)
{
  bytes = ((bytes ? bytes : 16) + 255) & ~(size_t)255;
  if (bytes >= ((size_t)64 << 20)) {
    // size classes for large blocks (1/8 of the leading power of two, <= 12.5 % waste): calls whose row counts differ by a few
    // per cent (the folds of a multi-shard job) then ask for the SAME size and re-use each other's blocks; exact sizes made every
    // such call miss the cache and grow the pool until the driver ran out of memory
    size_t step = (size_t)1 << 26;
    while ((step << 4) <= bytes) step <<= 1;
    bytes = (bytes + step - 1) & ~(step - 1);
  }
  std::lock_guard<std::mutex> lock(g_mu);
  DeviceCache* c = current_cache();
  if (!c) return nullptr;
  auto it = c->free_blocks.find(s);
  if (it != c->free_blocks.end()) {
    // best fit among this stream's freed blocks, with bounded slack so a small request never pins a huge block
    // (a transient block: no bound. A 1e9-row call's 38 GB record arena then also serves the 23 GB arena of the fold that follows and
    // the 6 GB arena of the merge, instead of three blocks that together no longer fit beside the caller's data)
    const size_t slack = transient ? ~(size_t)0 : bytes + (bytes > (1u << 20) ? bytes / 4 : (1u << 18));
    int best = -1;
    std::vector<FreeBlock>& v = it->second;
    for (int i = 0; i < (int)v.size(); ++i)
      if (v[i].bytes >= bytes && v[i].bytes <= slack && (best < 0 || v[i].bytes < v[best].bytes)) best = i;
    if (best >= 0) {
      FreeBlock b = v[best];
      v[best]     = v.back();
      v.pop_back();
      c->cached_bytes -= b.bytes;
      c->live[b.p] = b.bytes;
      ++g_hits;
      return b.p;
    }
  }
  void* p = driver_alloc(*c, bytes, s);
  ++g_driver_allocs; g_driver_bytes += bytes;
  if (!p) {  // out of memory: return the cache to the driver and try once more
    ++g_oom_flushes;
    flush_locked(*c);
    cudaStreamSynchronize(s);
    p = driver_alloc(*c, bytes, s);
    if (!p) return nullptr;
  }
  c->live[p] = bytes;
  return p;
}

void pool_free_async(void* p, cudaStream_t s)
{
  if (!p) return;
  std::lock_guard<std::mutex> lock(g_mu);
  DeviceCache* c = current_cache();
  size_t bytes   = 0;
  if (c) {
    auto it = c->live.find(p);
    if (it != c->live.end()) { bytes = it->second; c->live.erase(it); }
  }
  if (!c || bytes == 0) { cudaFreeAsync(p, s); return; }  // not ours (or allocated on another device): plain free
  c->free_blocks[s].push_back({p, bytes});
  c->cached_bytes += bytes;
  while (c->cached_bytes > c->limit_bytes) {  // over budget: release the largest cached block of this stream
    std::vector<FreeBlock>& v = c->free_blocks[s];
    if (v.empty()) { flush_locked(*c); break; }
    int big = 0;
    for (int i = 1; i < (int)v.size(); ++i) if (v[i].bytes > v[big].bytes) big = i;
    cudaFreeAsync(v[big].p, s);
    c->cached_bytes -= v[big].bytes;
    v[big] = v.back();
    v.pop_back();
  }
}

void pool_release_stream(cudaStream_t s)
{
  std::lock_guard<std::mutex> lock(g_mu);
  DeviceCache* c = current_cache();
  if (!c) return;
  auto it = c->free_blocks.find(s);
  if (it == c->free_blocks.end()) return;
  for (FreeBlock& b : it->second) { cudaFreeAsync(b.p, s); c->cached_bytes -= b.bytes; }
  c->free_blocks.erase(it);  // a destroyed stream's handle must never be used again (flush_locked frees on the key)
}

void pool_trim()
{
  std::lock_guard<std::mutex> lock(g_mu);
  if (DeviceCache* c = current_cache()) flush_locked(*c);
}

}  // namespace bgb

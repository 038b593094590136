// bgb_groupby_host: the same groupby call for HOST-resident columns (end-to-end measurement and simple FFI callers).
// Stages the inputs H2D on a private stream, runs build + emit on the device, copies the results back into malloc'ed
// host buffers (pinned, from a cache: see HostCache below). Device work is identical to the device-pointer entry points; nothing is computed on the CPU.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "errors.h"
#include "plan.h"
#include "nvtx.h"
#include "pool.h"

using namespace bgb;

namespace {

void* alloc_cb(size_t bytes, void* stream, void*) { return pool_alloc_async(bytes, (cudaStream_t)stream); }
void free_cb(void* ptr, size_t, void* stream, void*) { pool_free_async(ptr, (cudaStream_t)stream); }

// Result buffers are PINNED host memory from a small cache: a D2H copy into pageable memory runs at 2-3 GB/s on this box and
// cudaHostAlloc costs ~0.5 s per GB, so blocks released by bgb_free_host_result are kept (up to 8 GB, BGB_HOST_CACHE_MB) and
// handed to the next call, which then copies its results at PCIe rate with no staging and no page faults. A block the cache
// cannot serve and cudaHostAlloc cannot pin falls back to malloc.
struct HostCache {
  struct Block { void* p; size_t bytes; };
  std::mutex mu;
  std::vector<Block> free_blocks;
  std::unordered_map<void*, size_t> live;  // pinned blocks handed out
  size_t cached = 0;
  size_t limit  = getenv("BGB_HOST_CACHE_MB") ? (size_t)atoll(getenv("BGB_HOST_CACHE_MB")) << 20 : (size_t)8 << 30;
};
HostCache& host_cache() { static HostCache* c = new HostCache(); return *c; }

void* host_alloc(size_t bytes)
{
  bytes = ((bytes ? bytes : 16) + 4095) & ~(size_t)4095;
  HostCache& c = host_cache();
  std::lock_guard<std::mutex> lock(c.mu);
  int best = -1;
  for (int i = 0; i < (int)c.free_blocks.size(); ++i)
    if (c.free_blocks[i].bytes >= bytes && c.free_blocks[i].bytes <= bytes + bytes / 4 + (1u << 20) && (best < 0 || c.free_blocks[i].bytes < c.free_blocks[best].bytes)) best = i;
  if (best >= 0) {
    HostCache::Block b = c.free_blocks[best];
    c.free_blocks[best] = c.free_blocks.back();
    c.free_blocks.pop_back();
    c.cached -= b.bytes;
    c.live[b.p] = b.bytes;
    return b.p;
  }
  void* p = nullptr;
  if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) == cudaSuccess) { c.live[p] = bytes; return p; }
  cudaGetLastError();
  return malloc(bytes);  // pageable fallback (not tracked: released with free)
}

void host_free(void* p)
{
  if (!p) return;
  HostCache& c = host_cache();
  std::lock_guard<std::mutex> lock(c.mu);
  auto it = c.live.find(p);
  if (it == c.live.end()) { free(p); return; }
  const size_t bytes = it->second;
  c.live.erase(it);
  if (c.cached + bytes <= c.limit) { c.free_blocks.push_back({p, bytes}); c.cached += bytes; }
  else cudaFreeHost(p);
}

struct DeviceStage {
  cudaStream_t s = nullptr;
  std::vector<void*> bufs;
  ~DeviceStage()
  {
    for (void* p : bufs) pool_free_async(p, s);
    if (s) {
      pool_release_stream(s);  // the pool caches freed blocks per stream: nothing may stay keyed on a stream that is about to die
      cudaStreamSynchronize(s);
      cudaStreamDestroy(s);
    }
  }
  void* alloc(size_t bytes)
  {
    void* p = pool_alloc_async(bytes, s);
    if (p) bufs.push_back(p);
    return p;
  }
};

// copies one host column to the device, keeping offset semantics (the copy starts at element 0 of the slice)
int stage_column(DeviceStage& st, const bgb_column_view& h, bgb_column_view& d)
{
  d = h;
  d.offset = 0;
  const int w = type_width(h.type_id);
  if (w == 0) return set_error(BGB_UNSUPPORTED, "host column of type %d is not fixed-width", h.type_id);
  const size_t n = (size_t)h.size;
  if (n == 0) { d.data = nullptr; d.null_mask = nullptr; return BGB_OK; }
  void* dd = st.alloc(n * w);
  if (!dd) return set_error(BGB_OUT_OF_MEMORY, "device staging buffer (%zu bytes)", n * w);
  int rc = check_cuda(cudaMemcpyAsync(dd, (const uint8_t*)h.data + (size_t)h.offset * w, n * w, cudaMemcpyHostToDevice, st.s), "H2D data");
  if (rc) return rc;
  d.data = dd;
  if (h.null_mask) {
    // whole words covering bits [offset, offset+size); the bit offset inside the first word is kept
    const size_t w0 = (size_t)h.offset / 32, w1 = ((size_t)h.offset + n + 31) / 32;
    uint32_t* dm = (uint32_t*)st.alloc((w1 - w0) * 4);
    if (!dm) return set_error(BGB_OUT_OF_MEMORY, "device staging mask");
    rc = check_cuda(cudaMemcpyAsync(dm, h.null_mask + w0, (w1 - w0) * 4, cudaMemcpyHostToDevice, st.s), "H2D mask");
    if (rc) return rc;
    d.null_mask = dm;
    d.offset    = h.offset % 32;
    d.data      = (const uint8_t*)dd - (size_t)d.offset * w;  // element i of the view is still data + (offset+i)*w
  }
  return BGB_OK;
}

}  // namespace

extern "C" void bgb_free_host_result(bgb_host_column* cols, int32_t n)
{
  if (!cols) return;
  for (int i = 0; i < n; ++i) { host_free(cols[i].data); host_free(cols[i].null_mask); }
  free(cols);
}

extern "C" int bgb_groupby_host(const bgb_column_view* host_keys, int32_t num_keys, int32_t include_null_keys,
                                const bgb_request* host_requests, int32_t num_requests, const bgb_options* options,
                                int32_t* num_groups, bgb_host_column** out_keys, bgb_host_column** out_results, int32_t* num_results)
{
  bgb::NvtxRange nvtx_call("bgb_groupby_host");
  if (!num_groups || !out_keys || !out_results || !num_results) return set_error(BGB_INVALID_ARGUMENT, "null output pointer");
  *num_groups = 0; *out_keys = nullptr; *out_results = nullptr; *num_results = 0;
  if (num_keys <= 0 || !host_keys) return set_error(BGB_INVALID_ARGUMENT, "groupby needs at least one key column");
  {
    // validate before touching the device (same checks, same order as the device entry point)
    Plan probe;
    int rc = make_plan(host_keys, num_keys, include_null_keys, host_requests, num_requests, probe, false, /*allow_wide=*/true);
    if (rc) return rc;
  }
  static const bool trace = getenv("BGB_TRACE") != nullptr;
  auto t_prev = std::chrono::steady_clock::now();
  auto mark = [&](const char* what, bool sync) {
    if (!trace) return;
    if (sync) cudaStreamSynchronize(nullptr);
    auto t = std::chrono::steady_clock::now();
    fprintf(stderr, "[bgb host] %s %.1f ms\n", what, std::chrono::duration<double, std::milli>(t - t_prev).count());
    t_prev = t;
  };
  DeviceStage st;
  int rc = check_cuda(cudaStreamCreateWithFlags(&st.s, cudaStreamNonBlocking), "stream");
  if (rc) return rc;
  bgb_allocator pool_allocator{alloc_cb, free_cb, nullptr};
  std::vector<bgb_column_view> dkeys(num_keys);
  for (int i = 0; i < num_keys; ++i) if ((rc = stage_column(st, host_keys[i], dkeys[i]))) return rc;
  std::vector<bgb_request> dreqs(num_requests > 0 ? num_requests : 0);
  for (int r = 0; r < num_requests; ++r) {
    dreqs[r] = host_requests[r];
    // the same host column in two requests is staged once (keeps the library's same-column dedup working)
    bool found = false;
    for (int q = 0; q < r && !found; ++q)
      if (host_requests[q].values.data == host_requests[r].values.data && host_requests[q].values.null_mask == host_requests[r].values.null_mask &&
          host_requests[q].values.offset == host_requests[r].values.offset && host_requests[q].values.type_id == host_requests[r].values.type_id) {
        dreqs[r].values = dreqs[q].values; found = true;
      }
    if (!found && (rc = stage_column(st, host_requests[r].values, dreqs[r].values))) return rc;
  }
  if (trace) { cudaStreamSynchronize(st.s); mark("stage H2D", false); }
  bgb_groupby* h = nullptr;
  int32_t G = 0;
  rc = bgb_groupby_build(dkeys.data(), num_keys, include_null_keys, dreqs.data(), num_requests, (void*)st.s, &pool_allocator, options, &h, &G);
  if (rc) return rc;
  mark("build", false);
  int nres = 0;
  for (int r = 0; r < num_requests; ++r) nres += host_requests[r].num_aggregations;
  std::vector<bgb_mutable_column> okeys(num_keys), ores(nres > 0 ? nres : 1);
  bgb_host_column* hk = (bgb_host_column*)calloc(num_keys, sizeof(bgb_host_column));
  bgb_host_column* hr = (bgb_host_column*)calloc(nres > 0 ? nres : 1, sizeof(bgb_host_column));
  auto fail = [&](int code) { bgb_groupby_destroy(h); bgb_free_host_result(hk, num_keys); bgb_free_host_result(hr, nres); return code; };
  if (!hk || !hr) return fail(set_error(BGB_OUT_OF_MEMORY, "host result arrays"));
  const size_t words = ((size_t)G + 31) / 32;
  auto prepare = [&](bgb_mutable_column& m, bgb_host_column& hc, int32_t type_id, int32_t nullable) -> int {
    if (type_id == BGB_STRUCT) return set_error(BGB_UNSUPPORTED, "bgb_groupby_host: struct results (SUM_OVERFLOW) are served by bgb_groupby_build / emit only");
    const int w = type_width(type_id);
    m = bgb_mutable_column{type_id, G, nullptr, nullptr, 0, 0};
    hc.type_id = type_id; hc.size = G;
    if (G == 0) return BGB_OK;
    m.data = st.alloc((size_t)G * w);
    hc.data = host_alloc((size_t)G * w);
    if (nullable) { m.null_mask = (uint32_t*)st.alloc(words * 4); hc.null_mask = (uint32_t*)host_alloc(words * 4); }
    if (!m.data || !hc.data || (nullable && (!m.null_mask || !hc.null_mask))) return set_error(BGB_OUT_OF_MEMORY, "result buffers");
    return BGB_OK;
  };
  for (int i = 0; i < num_keys; ++i) {
    int32_t t, sc, nl;
    if ((rc = bgb_groupby_key_info(h, i, &t, &sc, &nl)) || (rc = prepare(okeys[i], hk[i], t, nl))) return fail(rc);
  }
  int k = 0;
  for (int r = 0; r < num_requests; ++r)
    for (int a = 0; a < host_requests[r].num_aggregations; ++a, ++k) {
      int32_t t, sc, nl;
      if ((rc = bgb_groupby_result_info(h, r, a, &t, &sc, &nl)) || (rc = prepare(ores[k], hr[k], t, nl))) return fail(rc);
    }
  mark("result alloc", false);
  if ((rc = bgb_groupby_emit(h, okeys.data(), ores.data(), (void*)st.s))) return fail(rc);
  if (trace) { cudaStreamSynchronize(st.s); mark("emit", false); }
  auto d2h = [&](void* dst, const void* src, size_t bytes) -> int { return check_cuda(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, st.s), "D2H"); };
  auto fetch = [&](const bgb_mutable_column& m, bgb_host_column& hc) -> int {
    hc.null_count = m.null_count;
    if (G == 0) return BGB_OK;
    int rc2 = d2h(hc.data, m.data, (size_t)G * type_width(m.type_id));
    if (!rc2 && m.null_mask) rc2 = d2h(hc.null_mask, m.null_mask, words * 4);
    return rc2;
  };
  for (int i = 0; i < num_keys; ++i) if ((rc = fetch(okeys[i], hk[i]))) return fail(rc);
  for (int i = 0; i < nres; ++i) if ((rc = fetch(ores[i], hr[i]))) return fail(rc);
  if ((rc = check_cuda(cudaStreamSynchronize(st.s), "host result sync"))) return fail(rc);
  mark("D2H", false);
  bgb_groupby_destroy(h);
  *num_groups = G; *out_keys = hk; *out_results = hr; *num_results = nres;
  return BGB_OK;
}

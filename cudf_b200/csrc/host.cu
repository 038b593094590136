// bgb_groupby_host: the same groupby call for HOST-resident columns (end-to-end measurement and simple FFI callers).
// Stages the inputs H2D on a private stream, runs build + emit on the device, copies the results back into malloc'ed
// host buffers. Device work is identical to the device-pointer entry points; nothing is computed on the CPU.
#include <cstdlib>
#include <cstring>
#include <vector>

#include "errors.h"
#include "plan.h"
#include "pool.h"

using namespace bgb;

namespace {

void* alloc_cb(size_t bytes, void* stream, void*) { return pool_alloc_async(bytes, (cudaStream_t)stream); }
void free_cb(void* ptr, size_t, void* stream, void*) { pool_free_async(ptr, (cudaStream_t)stream); }

struct DeviceStage {
  cudaStream_t s = nullptr;
  std::vector<void*> bufs;
  ~DeviceStage()
  {
    for (void* p : bufs) cudaFreeAsync(p, s);
    if (s) { cudaStreamSynchronize(s); cudaStreamDestroy(s); }
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
  for (int i = 0; i < n; ++i) { free(cols[i].data); free(cols[i].null_mask); }
  free(cols);
}

extern "C" int bgb_groupby_host(const bgb_column_view* host_keys, int32_t num_keys, int32_t include_null_keys,
                                const bgb_request* host_requests, int32_t num_requests, const bgb_options* options,
                                int32_t* num_groups, bgb_host_column** out_keys, bgb_host_column** out_results, int32_t* num_results)
{
  if (!num_groups || !out_keys || !out_results || !num_results) return set_error(BGB_INVALID_ARGUMENT, "null output pointer");
  *num_groups = 0; *out_keys = nullptr; *out_results = nullptr; *num_results = 0;
  if (num_keys <= 0 || !host_keys) return set_error(BGB_INVALID_ARGUMENT, "groupby needs at least one key column");
  {
    // validate before touching the device (same checks, same order as the device entry point)
    Plan probe;
    int rc = make_plan(host_keys, num_keys, include_null_keys, host_requests, num_requests, probe);
    if (rc) return rc;
  }
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
  bgb_groupby* h = nullptr;
  int32_t G = 0;
  rc = bgb_groupby_build(dkeys.data(), num_keys, include_null_keys, dreqs.data(), num_requests, (void*)st.s, &pool_allocator, options, &h, &G);
  if (rc) return rc;
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
    hc.data = malloc((size_t)G * w);
    if (nullable) { m.null_mask = (uint32_t*)st.alloc(words * 4); hc.null_mask = (uint32_t*)malloc(words * 4); }
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
  if ((rc = bgb_groupby_emit(h, okeys.data(), ores.data(), (void*)st.s))) return fail(rc);
  auto fetch = [&](const bgb_mutable_column& m, bgb_host_column& hc) -> int {
    hc.null_count = m.null_count;
    if (G == 0) return BGB_OK;
    int rc2 = check_cuda(cudaMemcpyAsync(hc.data, m.data, (size_t)G * type_width(m.type_id), cudaMemcpyDeviceToHost, st.s), "D2H data");
    if (!rc2 && m.null_mask) rc2 = check_cuda(cudaMemcpyAsync(hc.null_mask, m.null_mask, words * 4, cudaMemcpyDeviceToHost, st.s), "D2H mask");
    return rc2;
  };
  for (int i = 0; i < num_keys; ++i) if ((rc = fetch(okeys[i], hk[i]))) return fail(rc);
  for (int i = 0; i < nres; ++i) if ((rc = fetch(ores[i], hr[i]))) return fail(rc);
  if ((rc = check_cuda(cudaStreamSynchronize(st.s), "host result sync"))) return fail(rc);
  bgb_groupby_destroy(h);
  *num_groups = G; *out_keys = hk; *out_results = hr; *num_results = nres;
  return BGB_OK;
}

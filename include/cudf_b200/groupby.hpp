// Header-only C++ host layer over the C ABI (include/bgb.h) with the shape of the reference's public groupby API:
//   cudf::groupby::groupby(table_view keys, null_policy).aggregate(span<aggregation_request>, stream)
//   (/root/reference/cpp/include/cudf/groupby.hpp:54-184), aggregation factories
//   (/root/reference/cpp/include/cudf/aggregation.hpp:293-420), error classes (cudf/utilities/error.hpp:35-97).
// Device memory comes from cudaMallocAsync on the call's stream (the role of the rmm resource). No compute here:
// every call forwards to libbgb.so; a missing GPU or library is an exception, never a fallback.
#pragma once

#include <cuda_runtime_api.h>

#include <array>
#include <cstdint>
#include <deque>
#include <memory>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

#include "../bgb.h"

namespace cudf_b200 {

using size_type = int32_t;  // cudf::size_type (cudf/types.hpp:76)

enum class type_id : int32_t {  // cudf::type_id (cudf/types.hpp:185-217)
  EMPTY = 0, INT8, INT16, INT32, INT64, UINT8, UINT16, UINT32, UINT64, FLOAT32, FLOAT64, BOOL8,
  TIMESTAMP_DAYS, TIMESTAMP_SECONDS, TIMESTAMP_MILLISECONDS, TIMESTAMP_MICROSECONDS, TIMESTAMP_NANOSECONDS,
  DURATION_DAYS, DURATION_SECONDS, DURATION_MILLISECONDS, DURATION_MICROSECONDS, DURATION_NANOSECONDS,
  DICTIONARY32, STRING, LIST, DECIMAL32, DECIMAL64, DECIMAL128, STRUCT
};
enum class null_policy : bool { EXCLUDE = false, INCLUDE = true };

struct logic_error : std::logic_error { using std::logic_error::logic_error; };     // cudf::logic_error
struct cuda_error : std::runtime_error { using std::runtime_error::runtime_error; };  // cudf::cuda_error
struct unsupported_error : std::runtime_error { using std::runtime_error::runtime_error; };  // "take the sort path"

// ---- stand-ins for the three vocabulary types of the reference signature that live outside libcudf (CCCL / rmm are not
// vendored here): same call shape as cudf/groupby.hpp:181-184
//   aggregate(std::span<aggregation_request const>, cuda::stream_ref, rmm::device_async_resource_ref)
// A build that HAS those headers passes the real objects: stream_ref converts from cudaStream_t (cuda::stream_ref::get()),
// device_async_resource_ref wraps anything with allocate_async / deallocate_async (rmm's resource_ref interface).
struct stream_ref {
  cudaStream_t s{nullptr};
  stream_ref() = default;
  stream_ref(cudaStream_t st) : s(st) {}  // NOLINT: implicit like cuda::stream_ref
  cudaStream_t get() const { return s; }
};
inline stream_ref get_default_stream() { return stream_ref{}; }  // cudf::get_default_stream()

class device_async_resource_ref {
 public:
  // (not for device_async_resource_ref itself: a non-const lvalue copy must stay a copy, not a reference to a reference)
  template <typename Resource, typename = std::enable_if_t<!std::is_same<std::decay_t<Resource>, device_async_resource_ref>::value>>
  device_async_resource_ref(Resource& r)  // NOLINT: implicit like rmm::device_async_resource_ref
    : _obj(&r),
      _alloc([](void* o, size_t b, cudaStream_t s) { return static_cast<Resource*>(o)->allocate_async(b, s); }),
      _free([](void* o, void* p, size_t b, cudaStream_t s) { static_cast<Resource*>(o)->deallocate_async(p, b, s); })
  {
  }
  void* allocate_async(size_t bytes, cudaStream_t s) const { return _alloc(_obj, bytes, s); }
  void deallocate_async(void* p, size_t bytes, cudaStream_t s) const { _free(_obj, p, bytes, s); }
  // the bgb_allocator view of this resource (scratch of the C ABI call comes from the caller's resource, like
  // cpp/src/groupby/hash/compute_groupby.cu:66,76,80 use the current device resource)
  bgb_allocator abi() const
  {
    return bgb_allocator{[](size_t b, void* s, void* ctx) { return static_cast<device_async_resource_ref const*>(ctx)->allocate_async(b, (cudaStream_t)s); },
                         [](void* p, size_t b, void* s, void* ctx) { static_cast<device_async_resource_ref const*>(ctx)->deallocate_async(p, b, (cudaStream_t)s); },
                         const_cast<device_async_resource_ref*>(this)};
  }

 private:
  void* _obj;
  void* (*_alloc)(void*, size_t, cudaStream_t);
  void (*_free)(void*, void*, size_t, cudaStream_t);
};
// rmm::mr::cuda_async_memory_resource's role: stream-ordered allocation from the device's default pool
struct cuda_async_resource {
  void* allocate_async(size_t bytes, cudaStream_t s)
  {
    void* p = nullptr;
    if (cudaMallocAsync(&p, bytes ? bytes : 1, s) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
  }
  void deallocate_async(void* p, size_t, cudaStream_t s) { cudaFreeAsync(p, s); }
};
inline device_async_resource_ref get_current_device_resource_ref()  // cudf::get_current_device_resource_ref()
{
  static cuda_async_resource r;
  return device_async_resource_ref(r);
}
// std::span<T const> of the reference signature (C++20); this header is C++17
template <typename T>
class host_span {
 public:
  host_span() = default;
  host_span(T* p, size_t n) : _p(p), _n(n) {}
  template <typename V, typename = decltype(std::declval<V const&>().data())>
  host_span(V const& v) : _p(v.data()), _n(v.size()) {}  // NOLINT: std::vector / std::array / std::span
  T* data() const { return _p; }
  size_t size() const { return _n; }
  T& operator[](size_t i) const { return _p[i]; }
  T* begin() const { return _p; }
  T* end() const { return _p + _n; }

 private:
  T* _p{nullptr};
  size_t _n{0};
};

namespace detail {
inline void throw_on(int status)
{
  if (status == BGB_OK) return;
  std::string msg = bgb_last_error();
  if (status == BGB_INVALID_ARGUMENT) throw logic_error(msg);
  if (status == BGB_UNSUPPORTED) throw unsupported_error(msg);
  if (status == BGB_OUT_OF_MEMORY) throw std::bad_alloc();
  throw cuda_error(msg);
}
inline void cuda_try(cudaError_t e)
{
  if (e != cudaSuccess) throw cuda_error(std::string(cudaGetErrorName(e)) + ": " + cudaGetErrorString(e));
}
inline size_type width_of(type_id t)
{
  switch (t) {
    case type_id::INT8: case type_id::UINT8: case type_id::BOOL8: return 1;
    case type_id::INT16: case type_id::UINT16: return 2;
    case type_id::INT32: case type_id::UINT32: case type_id::FLOAT32: case type_id::TIMESTAMP_DAYS: case type_id::DURATION_DAYS:
    case type_id::DECIMAL32: return 4;
    case type_id::EMPTY: case type_id::DICTIONARY32: case type_id::STRING: case type_id::LIST: case type_id::DECIMAL128:
    case type_id::STRUCT: return 0;
    default: return 8;
  }
}
}  // namespace detail

struct data_type {
  type_id id_{type_id::EMPTY};
  int32_t scale_{0};
  data_type() = default;
  explicit data_type(type_id id, int32_t scale = 0) : id_(id), scale_(scale) {}
  type_id id() const { return id_; }
  int32_t scale() const { return scale_; }
};

// non-owning view with cudf::column_view's fields (cudf/column/column_view.hpp:236-243)
class column_view {
 public:
  column_view() = default;
  column_view(data_type type, size_type size, void const* data, uint32_t const* null_mask = nullptr, size_type null_count = 0,
              size_type offset = 0)
    : _type(type), _size(size), _data(data), _null_mask(null_mask), _null_count(null_mask ? null_count : 0), _offset(offset) {}
  data_type type() const { return _type; }
  size_type size() const { return _size; }
  size_type null_count() const { return _null_count; }
  bool has_nulls() const { return _null_count > 0; }
  bool nullable() const { return _null_mask != nullptr; }
  size_type offset() const { return _offset; }
  void const* head() const { return _data; }
  uint32_t const* null_mask() const { return _null_mask; }
  bgb_column_view abi() const
  {
    return {static_cast<int32_t>(_type.id()), _type.scale(), _size, _null_count, _offset, 0, _data, _null_mask};
  }

 private:
  data_type _type;
  size_type _size{0};
  void const* _data{nullptr};
  uint32_t const* _null_mask{nullptr};
  size_type _null_count{0};
  size_type _offset{0};
};

// owning device column (stream-ordered allocation)
class column {
 public:
  column(data_type type, size_type size, bool with_mask, cudaStream_t stream, device_async_resource_ref mr = get_current_device_resource_ref())
    : _type(type), _size(size), _stream(stream), _mr(mr)
  {
    _bytes      = static_cast<size_t>(size) * detail::width_of(type.id());
    _mask_bytes = (with_mask && size) ? sizeof(uint32_t) * ((size + 31) / 32) : 0;
    if (_bytes && !(_data = _mr.allocate_async(_bytes, stream))) throw cuda_error("out of device memory (column data)");
    if (_mask_bytes && !(_mask = static_cast<uint32_t*>(_mr.allocate_async(_mask_bytes, stream)))) throw cuda_error("out of device memory (column mask)");
  }
  column(column const&)            = delete;
  column& operator=(column const&) = delete;
  ~column()
  {
    if (_data) _mr.deallocate_async(_data, _bytes, _stream);
    if (_mask) _mr.deallocate_async(_mask, _mask_bytes, _stream);
  }
  // host -> device convenience for tests and examples (cudf::test::fixed_width_column_wrapper's role)
  template <typename T>
  static std::unique_ptr<column> from_host(type_id id, std::vector<T> const& values, std::vector<bool> const& valid = {},
                                           cudaStream_t stream = nullptr)
  {
    auto const n = static_cast<size_type>(values.size());
    auto c       = std::make_unique<column>(data_type{id}, n, !valid.empty(), stream);
    if (n) detail::cuda_try(cudaMemcpyAsync(c->_data, values.data(), sizeof(T) * n, cudaMemcpyHostToDevice, stream));
    if (!valid.empty() && n) {
      std::vector<uint32_t> words((n + 31) / 32, 0u);
      for (size_type i = 0; i < n; ++i) {
        if (valid[i]) words[i / 32] |= 1u << (i % 32);
        else ++c->_null_count;
      }
      detail::cuda_try(cudaMemcpyAsync(c->_mask, words.data(), sizeof(uint32_t) * words.size(), cudaMemcpyHostToDevice, stream));
    }
    detail::cuda_try(cudaStreamSynchronize(stream));
    return c;
  }
  template <typename T>
  std::pair<std::vector<T>, std::vector<bool>> to_host() const
  {
    std::vector<T> v(_size);
    std::vector<bool> ok(_size, true);
    if (_size) detail::cuda_try(cudaMemcpyAsync(v.data(), _data, sizeof(T) * _size, cudaMemcpyDeviceToHost, _stream));
    std::vector<uint32_t> words((_size + 31) / 32);
    if (_mask && _size) detail::cuda_try(cudaMemcpyAsync(words.data(), _mask, sizeof(uint32_t) * words.size(), cudaMemcpyDeviceToHost, _stream));
    detail::cuda_try(cudaStreamSynchronize(_stream));
    if (_mask)
      for (size_type i = 0; i < _size; ++i) ok[i] = (words[i / 32] >> (i % 32)) & 1u;
    return {std::move(v), std::move(ok)};
  }
  column_view view() const { return column_view(_type, _size, _data, _mask, _null_count, 0); }
  data_type type() const { return _type; }
  size_type size() const { return _size; }
  size_type null_count() const { return _null_count; }
  bool nullable() const { return _mask != nullptr; }
  void set_null_count(size_type n) { _null_count = n; }
  bgb_mutable_column abi_mutable() { return {static_cast<int32_t>(_type.id()), _size, _data, _mask, 0, 0}; }
  // STRUCT columns (SUM_OVERFLOW results, column::child in cudf/column/column.hpp): children own their buffers, the parent the mask
  size_type num_children() const { return static_cast<size_type>(_children.size()); }
  column const& child(size_type i) const { return *_children.at(i); }
  column& add_child(std::unique_ptr<column> c) { _children.push_back(std::move(c)); return *_children.back(); }

 private:
  data_type _type;
  size_type _size;
  void* _data{nullptr};
  uint32_t* _mask{nullptr};
  size_type _null_count{0};
  cudaStream_t _stream;
  device_async_resource_ref _mr;
  size_t _bytes{0}, _mask_bytes{0};
  std::vector<std::unique_ptr<column>> _children;
};

class table_view {
 public:
  table_view() = default;
  table_view(std::vector<column_view> cols) : _cols(std::move(cols)) {}
  size_type num_columns() const { return static_cast<size_type>(_cols.size()); }
  size_type num_rows() const { return _cols.empty() ? 0 : _cols[0].size(); }
  column_view const& column(size_type i) const { return _cols.at(i); }
  auto begin() const { return _cols.begin(); }
  auto end() const { return _cols.end(); }

 private:
  std::vector<column_view> _cols;
};

class table {
 public:
  explicit table(std::vector<std::unique_ptr<column>> cols) : _cols(std::move(cols)) {}
  size_type num_columns() const { return static_cast<size_type>(_cols.size()); }
  size_type num_rows() const { return _cols.empty() ? 0 : _cols[0]->size(); }
  column const& get_column(size_type i) const { return *_cols.at(i); }

 private:
  std::vector<std::unique_ptr<column>> _cols;
};

// cudf::groupby_aggregation and its factories (cudf/aggregation.hpp:293-420)
struct groupby_aggregation {
  int32_t kind;
  int32_t ddof{1};
};
inline std::unique_ptr<groupby_aggregation> make_agg(int32_t kind, int32_t ddof = 1) { return std::unique_ptr<groupby_aggregation>(new groupby_aggregation{kind, ddof}); }
inline auto make_sum_aggregation() { return make_agg(BGB_SUM); }
// struct{sum, overflow} result (cudf/aggregation.hpp make_sum_overflow_aggregation): child 0 typed like the values, child 1 BOOL8
inline auto make_sum_overflow_aggregation() { return make_agg(BGB_SUM_OVERFLOW); }
inline auto make_product_aggregation() { return make_agg(BGB_PRODUCT); }
inline auto make_min_aggregation() { return make_agg(BGB_MIN); }
inline auto make_max_aggregation() { return make_agg(BGB_MAX); }
inline auto make_count_aggregation(null_policy null_handling = null_policy::EXCLUDE)
{
  return make_agg(null_handling == null_policy::EXCLUDE ? BGB_COUNT_VALID : BGB_COUNT_ALL);
}
inline auto make_sum_of_squares_aggregation() { return make_agg(BGB_SUM_OF_SQUARES); }
inline auto make_mean_aggregation() { return make_agg(BGB_MEAN); }
inline auto make_m2_aggregation() { return make_agg(BGB_M2); }
inline auto make_variance_aggregation(size_type ddof = 1) { return make_agg(BGB_VARIANCE, ddof); }
inline auto make_std_aggregation(size_type ddof = 1) { return make_agg(BGB_STD, ddof); }
inline auto make_argmax_aggregation() { return make_agg(BGB_ARGMAX); }
inline auto make_argmin_aggregation() { return make_agg(BGB_ARGMIN); }
inline auto make_median_aggregation() { return make_agg(BGB_MEDIAN); }

namespace groupby {

struct aggregation_request {  // cudf/groupby.hpp:54-57
  column_view values;
  std::vector<std::unique_ptr<groupby_aggregation>> aggregations;
};
struct aggregation_result {  // cudf/groupby.hpp:67-70
  std::vector<std::unique_ptr<column>> results;
};

class groupby {
 public:
  explicit groupby(table_view const& keys, null_policy null_handling = null_policy::EXCLUDE) : _keys(keys), _null_handling(null_handling) {}

  // detail::hash::can_use_hash_groupby (hash/groupby.cu:115-131) for this library
  bool can_use_hash_groupby(host_span<aggregation_request const> requests) const
  {
    marshalled m(_keys, requests);
    int32_t ok = 0;
    detail::throw_on(bgb_can_use_hash_groupby(m.keys.data(), (int32_t)m.keys.size(), m.reqs.data(), (int32_t)m.reqs.size(), &ok));
    return ok != 0;
  }

  // groupby::aggregate (cudf/groupby.hpp:148-184): unique keys (arbitrary order) + one aggregation_result per request
  // Same call shape as the reference (cudf/groupby.hpp:181-184): requests as a span, the stream, and the memory resource that
  // serves BOTH the returned columns and the call's scratch (group table, record buffers) through the C ABI's bgb_allocator.
  std::pair<std::unique_ptr<table>, std::vector<aggregation_result>> aggregate(host_span<aggregation_request const> requests,
                                                                               stream_ref stream_              = get_default_stream(),
                                                                               device_async_resource_ref mr    = get_current_device_resource_ref()) const
  {
    cudaStream_t stream = stream_.get();
    marshalled m(_keys, requests);
    bgb_groupby* h = nullptr;
    int32_t n      = 0;
    bgb_allocator const scratch = mr.abi();
    detail::throw_on(bgb_groupby_build(m.keys.data(), (int32_t)m.keys.size(), _null_handling == null_policy::INCLUDE, m.reqs.data(),
                                       (int32_t)m.reqs.size(), stream, &scratch, nullptr, &h, &n));
    struct guard { bgb_groupby* h; ~guard() { bgb_groupby_destroy(h); } } g{h};
    std::vector<std::unique_ptr<column>> kcols;
    std::vector<bgb_mutable_column> km, rm;
    for (int32_t i = 0; i < (int32_t)m.keys.size(); ++i) {
      int32_t t, s, nl;
      detail::throw_on(bgb_groupby_key_info(h, i, &t, &s, &nl));
      kcols.push_back(std::make_unique<column>(data_type{static_cast<type_id>(t), s}, n, nl != 0, stream, mr));
      km.push_back(kcols.back()->abi_mutable());
    }
    std::vector<aggregation_result> results(requests.size());
    std::deque<std::array<bgb_mutable_column, 2>> struct_children;  // (a deque keeps the descriptors where they are)
    for (size_t i = 0; i < requests.size(); ++i)
      for (size_t j = 0; j < requests[i].aggregations.size(); ++j) {
        int32_t t, s, nl;
        detail::throw_on(bgb_groupby_result_info(h, (int32_t)i, (int32_t)j, &t, &s, &nl));
        results[i].results.push_back(std::make_unique<column>(data_type{static_cast<type_id>(t), s}, n, nl != 0, stream, mr));
        rm.push_back(results[i].results.back()->abi_mutable());
        if (static_cast<type_id>(t) == type_id::STRUCT) {  // SUM_OVERFLOW: `data` = the two child descriptors (bgb.h)
          column& parent = *results[i].results.back();
          column& sum    = parent.add_child(std::make_unique<column>(data_type{requests[i].values.type().id(), s}, n, false, stream, mr));
          column& flag   = parent.add_child(std::make_unique<column>(data_type{type_id::BOOL8}, n, false, stream, mr));
          struct_children.push_back({sum.abi_mutable(), flag.abi_mutable()});
          rm.back().data = struct_children.back().data();
        }
      }
    detail::throw_on(bgb_groupby_emit(h, km.data(), rm.data(), stream));
    for (size_t i = 0; i < kcols.size(); ++i) kcols[i]->set_null_count(km[i].null_count);
    size_t q = 0;
    for (auto& r : results)
      for (auto& c : r.results) c->set_null_count(rm[q++].null_count);
    return {std::make_unique<table>(std::move(kcols)), std::move(results)};
  }

 private:
  struct marshalled {
    std::vector<bgb_column_view> keys;
    std::vector<std::vector<bgb_aggregation>> aggs;
    std::vector<bgb_request> reqs;
    marshalled(table_view const& k, host_span<aggregation_request const> requests) : aggs(requests.size())
    {
      for (auto const& c : k) keys.push_back(c.abi());
      for (size_t i = 0; i < requests.size(); ++i) {
        for (auto const& a : requests[i].aggregations) aggs[i].push_back({a->kind, a->ddof});
        reqs.push_back({requests[i].values.abi(), aggs[i].data(), (int32_t)aggs[i].size(), 0});
      }
    }
  };
  table_view _keys;
  null_policy _null_handling;
};

// cudf::groupby::streaming_aggregation_request / streaming_groupby (cudf/groupby.hpp:419-598)
struct streaming_aggregation_request {
  size_type column_index;
  std::unique_ptr<groupby_aggregation> aggregation;
};

inline bool is_streaming_groupby_supported(data_type values_type, int32_t kind)
{
  int32_t ok = 0;
  detail::throw_on(bgb_is_streaming_groupby_supported(static_cast<int32_t>(values_type.id()), kind, &ok));
  return ok != 0;
}

class streaming_groupby {
 public:
  streaming_groupby()                                    = delete;
  streaming_groupby(streaming_groupby const&)            = delete;
  streaming_groupby& operator=(streaming_groupby const&) = delete;
  streaming_groupby(streaming_groupby&& o) noexcept : _h(o._h), _nk(o._nk), _nreq(o._nreq) { o._h = nullptr; }
  ~streaming_groupby() { bgb_stream_destroy(_h); }

  streaming_groupby(std::vector<size_type> const& key_indices, std::vector<streaming_aggregation_request> const& requests,
                    size_type max_distinct_keys, null_policy null_handling = null_policy::EXCLUDE)
    : _nk(static_cast<int32_t>(key_indices.size())), _nreq(static_cast<int32_t>(requests.size()))
  {
    std::vector<bgb_stream_request> r;
    for (auto const& q : requests) r.push_back({q.column_index, q.aggregation->kind, q.aggregation->ddof, 0});
    detail::throw_on(bgb_stream_create(key_indices.data(), _nk, r.data(), _nreq, max_distinct_keys, null_handling == null_policy::INCLUDE, &_h));
  }

  void aggregate(table_view const& data, cudaStream_t stream = nullptr)
  {
    std::vector<bgb_column_view> cols;
    for (auto const& c : data) cols.push_back(c.abi());
    detail::throw_on(bgb_stream_aggregate(_h, cols.data(), (int32_t)cols.size(), stream, nullptr));
  }
  void merge(streaming_groupby const& other, cudaStream_t stream = nullptr) { detail::throw_on(bgb_stream_merge(_h, other._h, stream)); }
  size_type distinct_keys() const
  {
    int32_t n = 0;
    detail::throw_on(bgb_stream_distinct_keys(_h, &n));
    return n;
  }
  // does not modify the state; aggregate() may be called again afterwards
  std::pair<std::unique_ptr<table>, std::vector<aggregation_result>> finalize(stream_ref stream_           = get_default_stream(),
                                                                              device_async_resource_ref mr = get_current_device_resource_ref()) const
  {
    cudaStream_t stream = stream_.get();
    bgb_groupby* h = nullptr;
    int32_t n      = 0;
    detail::throw_on(bgb_stream_finalize(_h, stream, &h, &n));
    struct guard { bgb_groupby* h; ~guard() { bgb_groupby_destroy(h); } } g{h};
    std::vector<std::unique_ptr<column>> kcols;
    std::vector<bgb_mutable_column> km, rm;
    for (int32_t i = 0; i < _nk; ++i) {
      int32_t t, s, nl;
      detail::throw_on(bgb_groupby_key_info(h, i, &t, &s, &nl));
      kcols.push_back(std::make_unique<column>(data_type{static_cast<type_id>(t), s}, n, nl != 0, stream, mr));
      km.push_back(kcols.back()->abi_mutable());
    }
    std::vector<aggregation_result> results(_nreq);
    for (int32_t i = 0; i < _nreq; ++i) {
      int32_t t, s, nl;
      detail::throw_on(bgb_groupby_result_info(h, i, 0, &t, &s, &nl));
      results[i].results.push_back(std::make_unique<column>(data_type{static_cast<type_id>(t), s}, n, nl != 0, stream, mr));
      rm.push_back(results[i].results.back()->abi_mutable());
    }
    detail::throw_on(bgb_groupby_emit(h, km.data(), rm.data(), stream));
    for (size_t i = 0; i < kcols.size(); ++i) kcols[i]->set_null_count(km[i].null_count);
    for (int32_t i = 0; i < _nreq; ++i) results[i].results[0]->set_null_count(rm[i].null_count);
    return {std::make_unique<table>(std::move(kcols)), std::move(results)};
  }

 private:
  bgb_stream* _h{nullptr};
  int32_t _nk, _nreq;
};

}  // namespace groupby
}  // namespace cudf_b200

// C++ tests of include/cudf_b200/groupby.hpp (the cudf::groupby::groupby-shaped host layer over the C ABI).
// Cases are the reference's own known-answer tests, re-typed against this API:
//   /root/reference/cpp/tests/groupby/sum_tests.cpp (basic, zero_valid_keys, zero_valid_values, null_keys_and_values,
//   overflow_integer), count_tests.cpp, mean_tests.cpp, var_tests.cpp, keys_tests.cpp (include_null_keys, mismatch).
// Like groupby_test_util.cpp:86-94 the result is sorted by key before it is compared (group order is unspecified).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <limits>
#include <numeric>

#include "cudf_b200/groupby.hpp"

using namespace cudf_b200;
using K = int32_t;
static int g_fail = 0;
#define CHECK(cond)                                                          \
  do {                                                                       \
    if (!(cond)) { std::printf("FAIL %s:%d %s\n", __FILE__, __LINE__, #cond); ++g_fail; } \
  } while (0)

template <typename R>
struct sorted_result {
  std::vector<K> keys;
  std::vector<bool> key_valid;
  std::vector<R> vals;
  std::vector<bool> val_valid;
};

template <typename R>
sorted_result<R> run(column const& keys, column const& vals, std::unique_ptr<groupby_aggregation> agg,
                     null_policy pol = null_policy::EXCLUDE)
{
  groupby::groupby gb(table_view({keys.view()}), pol);
  std::vector<groupby::aggregation_request> reqs(1);
  reqs[0].values = vals.view();
  reqs[0].aggregations.push_back(std::move(agg));
  auto [tbl, res] = gb.aggregate(reqs);
  auto [k, kv]    = tbl->get_column(0).to_host<K>();
  auto [v, vv]    = res[0].results[0]->to_host<R>();
  std::vector<size_t> order(k.size());
  std::iota(order.begin(), order.end(), 0);
  std::sort(order.begin(), order.end(), [&](size_t a, size_t b) {  // nulls first, then ascending
    if (kv[a] != kv[b]) return !kv[a];
    return k[a] < k[b];
  });
  sorted_result<R> out;
  for (auto i : order) { out.keys.push_back(k[i]); out.key_valid.push_back(kv[i]); out.vals.push_back(v[i]); out.val_valid.push_back(vv[i]); }
  return out;
}

int main()
{
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { std::printf("no CUDA device: this layer has no CPU fallback\n"); return 2; }

  {  // sum_tests.cpp basic: keys {1,2,3,1,2,2,1,3,3,2} vals 0..9 -> {9,19,17}
    auto keys = column::from_host<K>(type_id::INT32, {1, 2, 3, 1, 2, 2, 1, 3, 3, 2});
    auto vals = column::from_host<int32_t>(type_id::INT32, {0, 1, 2, 3, 4, 5, 6, 7, 8, 9});
    auto r    = run<int64_t>(*keys, *vals, make_sum_aggregation());
    CHECK((r.keys == std::vector<K>{1, 2, 3}));
    CHECK((r.vals == std::vector<int64_t>{9, 19, 17}));
  }
  {  // zero_valid_keys -> empty result
    auto keys = column::from_host<K>(type_id::INT32, {1, 2, 3}, {false, false, false});
    auto vals = column::from_host<double>(type_id::FLOAT64, {3, 4, 5});
    auto r    = run<double>(*keys, *vals, make_sum_aggregation());
    CHECK(r.keys.empty() && r.vals.empty());
  }
  {  // zero_valid_values -> one group, null sum
    auto keys = column::from_host<K>(type_id::INT32, {1, 1, 1});
    auto vals = column::from_host<int64_t>(type_id::INT64, {3, 4, 5}, {false, false, false});
    auto r    = run<int64_t>(*keys, *vals, make_sum_aggregation());
    CHECK(r.keys == std::vector<K>{1});
    CHECK(r.val_valid == std::vector<bool>{false});
  }
  {  // null_keys_and_values -> keys {1,2,3,4}, sums {9,14,10,null}
    auto keys = column::from_host<K>(type_id::INT32, {1, 2, 3, 1, 2, 2, 1, 3, 3, 2, 4},
                                     {true, true, true, true, true, true, true, false, true, true, true});
    auto vals = column::from_host<int64_t>(type_id::INT64, {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 4},
                                           {false, true, true, true, true, false, true, true, true, true, false});
    auto r = run<int64_t>(*keys, *vals, make_sum_aggregation());
    CHECK((r.keys == std::vector<K>{1, 2, 3, 4}));
    CHECK((r.val_valid == std::vector<bool>{true, true, true, false}));
    CHECK(r.vals[0] == 9 && r.vals[1] == 14 && r.vals[2] == 10);
    auto c = run<int32_t>(*keys, *vals, make_count_aggregation());  // count_tests.cpp null_keys_and_values: {2,3,2,0}
    CHECK((c.vals == std::vector<int32_t>{2, 3, 2, 0}));
    auto ca = run<int32_t>(*keys, *vals, make_count_aggregation(null_policy::INCLUDE));  // COUNT_ALL: {3,4,2,1}
    CHECK((ca.vals == std::vector<int32_t>{3, 4, 2, 1}));
    auto inc = run<int64_t>(*keys, *vals, make_sum_aggregation(), null_policy::INCLUDE);  // keys_tests.cpp include_null_keys
    CHECK(inc.keys.size() == 5 && inc.key_valid[0] == false);
    CHECK(inc.vals[0] == 7);
  }
  {  // sum_tests.cpp overflow_integer: int32 sums accumulate in int64
    auto keys = column::from_host<K>(type_id::INT32, {0, 0, 1, 1});
    auto vals = column::from_host<int32_t>(type_id::INT32, {INT32_MAX, INT32_MAX, INT32_MIN, INT32_MIN});
    auto r    = run<int64_t>(*keys, *vals, make_sum_aggregation());
    CHECK((r.vals == std::vector<int64_t>{4294967294ll, -4294967296ll}));
  }
  {  // mean_tests.cpp basic {3, 19/4, 17/3}; var_tests.cpp basic {9, 131/12, 31/3}
    auto keys = column::from_host<K>(type_id::INT32, {1, 2, 3, 1, 2, 2, 1, 3, 3, 2});
    auto vals = column::from_host<int64_t>(type_id::INT64, {0, 1, 2, 3, 4, 5, 6, 7, 8, 9});
    auto m    = run<double>(*keys, *vals, make_mean_aggregation());
    CHECK(m.vals[0] == 3.0 && m.vals[1] == 19.0 / 4 && m.vals[2] == 17.0 / 3);
    auto v = run<double>(*keys, *vals, make_variance_aggregation());
    CHECK(std::fabs(v.vals[0] - 9.0) < 1e-12 && std::fabs(v.vals[1] - 131.0 / 12) < 1e-12 && std::fabs(v.vals[2] - 31.0 / 3) < 1e-12);
    auto s = run<double>(*keys, *vals, make_std_aggregation());
    CHECK(std::fabs(s.vals[0] - 3.0) < 1e-12);
    auto mn = run<int64_t>(*keys, *vals, make_min_aggregation());
    auto mx = run<int64_t>(*keys, *vals, make_max_aggregation());
    CHECK((mn.vals == std::vector<int64_t>{0, 1, 2}) && (mx.vals == std::vector<int64_t>{6, 9, 8}));
  }
  {  // var_tests.cpp: a single-row group has a null sample variance (n - ddof <= 0)
    auto keys = column::from_host<K>(type_id::INT32, {1, 2, 2});
    auto vals = column::from_host<double>(type_id::FLOAT64, {5, 1, 3});
    auto v    = run<double>(*keys, *vals, make_variance_aggregation(1));
    CHECK((v.val_valid == std::vector<bool>{false, true}) && v.vals[1] == 2.0);
  }
  {  // groupby.cu:226-230 "Size mismatch between request values and groupby keys."
    auto keys = column::from_host<K>(type_id::INT32, {1, 2, 3});
    auto vals = column::from_host<int32_t>(type_id::INT32, {1, 2});
    bool threw = false;
    try { run<int64_t>(*keys, *vals, make_sum_aggregation()); } catch (logic_error const& e) { threw = std::string(e.what()).find("Size mismatch") != std::string::npos; }
    CHECK(threw);
  }
  {  // median is not a hash aggregation: the dispatcher would take the sort path (groupby.cu:64-66)
    auto keys = column::from_host<K>(type_id::INT32, {1, 2, 3});
    auto vals = column::from_host<int32_t>(type_id::INT32, {1, 2, 3});
    groupby::groupby gb(table_view({keys->view()}));
    std::vector<groupby::aggregation_request> reqs(1);
    reqs[0].values = vals->view();
    reqs[0].aggregations.push_back(make_median_aggregation());
    CHECK(!gb.can_use_hash_groupby(reqs));
    reqs[0].aggregations.clear();
    reqs[0].aggregations.push_back(make_sum_aggregation());
    CHECK(gb.can_use_hash_groupby(reqs));
  }
  {  // keys_tests.cpp: duplicate aggregations in one request are answered twice; empty input gives empty typed outputs
    auto keys = column::from_host<K>(type_id::INT32, std::vector<K>{});
    auto vals = column::from_host<float>(type_id::FLOAT32, std::vector<float>{});
    groupby::groupby gb(table_view({keys->view()}));
    std::vector<groupby::aggregation_request> reqs(1);
    reqs[0].values = vals->view();
    reqs[0].aggregations.push_back(make_sum_aggregation());
    reqs[0].aggregations.push_back(make_sum_aggregation());
    auto [tbl, res] = gb.aggregate(reqs);
    CHECK(tbl->num_rows() == 0 && res[0].results.size() == 2 && res[0].results[1]->type().id() == type_id::FLOAT32);
  }
  {  // sum_overflow_tests.cpp OverflowDetection (int32): struct{sum, overflow}; flags {true,false,true,true}, group 2 sums to 100
    constexpr int32_t mx = std::numeric_limits<int32_t>::max(), mn = std::numeric_limits<int32_t>::min();
    auto keys = column::from_host<K>(type_id::INT32, {1, 2, 3, 4, 1, 2, 2, 1, 3, 3, 2, 4, 4});
    auto vals = column::from_host<int32_t>(type_id::INT32, {mx - 5, 10, 20, mn + 5, 10, 20, 30, mx - 5, mx - 5, 1, 40, -10, mn + 5});
    groupby::groupby gb(table_view({keys->view()}));
    std::vector<groupby::aggregation_request> reqs(1);
    reqs[0].values = vals->view();
    reqs[0].aggregations.push_back(make_sum_overflow_aggregation());
    auto [tbl, res] = gb.aggregate(reqs);
    column const& st = *res[0].results[0];
    CHECK(st.type().id() == type_id::STRUCT && st.num_children() == 2 && !st.nullable());
    CHECK(st.child(0).type().id() == type_id::INT32 && st.child(1).type().id() == type_id::BOOL8);
    auto [k, kv] = tbl->get_column(0).to_host<K>();
    auto [sum, sv] = st.child(0).to_host<int32_t>();
    auto [flag, fv] = st.child(1).to_host<uint8_t>();
    std::vector<std::pair<K, int>> got;
    int32_t sum2 = 0;
    for (size_t i = 0; i < k.size(); ++i) { got.push_back({k[i], (int)flag[i]}); if (k[i] == 2) sum2 = sum[i]; }
    std::sort(got.begin(), got.end());
    CHECK((got == std::vector<std::pair<K, int>>{{1, 1}, {2, 0}, {3, 1}, {4, 1}}));
    CHECK(sum2 == 100);
  }
  {  // streaming_groupby_test.cpp MergeTwoObjects: worker1 {1,2,1}/{10,20,30} + worker2 {2,3}/{40,50} -> {1,2,3} / {40,60,50}
    auto mk = [] {
      std::vector<groupby::streaming_aggregation_request> r;
      r.push_back({1, make_sum_aggregation()});
      return r;
    };
    groupby::streaming_groupby w1({0}, mk(), 1024), w2({0}, mk(), 1024);
    auto k1 = column::from_host<K>(type_id::INT32, {1, 2, 1});
    auto v1 = column::from_host<int32_t>(type_id::INT32, {10, 20, 30});
    auto k2 = column::from_host<K>(type_id::INT32, {2, 3});
    auto v2 = column::from_host<int32_t>(type_id::INT32, {40, 50});
    w1.aggregate(table_view({k1->view(), v1->view()}));
    w2.aggregate(table_view({k2->view(), v2->view()}));
    w1.merge(w2);
    auto [tbl, res] = w1.finalize();
    auto [k, kv]    = tbl->get_column(0).to_host<K>();
    auto [v, vv]    = res[0].results[0]->to_host<int64_t>();
    std::vector<std::pair<K, int64_t>> got;
    for (size_t i = 0; i < k.size(); ++i) got.push_back({k[i], v[i]});
    std::sort(got.begin(), got.end());
    CHECK((got == std::vector<std::pair<K, int64_t>>{{1, 40}, {2, 60}, {3, 50}}));
    CHECK(w1.distinct_keys() == 3 && w2.distinct_keys() == 2);
    // ExceedsDistinctKeyCapacityThrows / UnsupportedAggThrows / FinalizeBeforeAggregateThrows
    groupby::streaming_groupby small({0}, mk(), 2);
    bool threw = false;
    try { small.finalize(); } catch (logic_error const&) { threw = true; }
    CHECK(threw);
    small.aggregate(table_view({k2->view(), v2->view()}));
    threw = false;
    try { small.aggregate(table_view({k1->view(), v1->view()})); } catch (logic_error const&) { threw = true; }  // 3 rows > 2
    CHECK(threw);
    threw = false;
    try {
      std::vector<groupby::streaming_aggregation_request> bad;
      bad.push_back({1, make_argmax_aggregation()});
      groupby::streaming_groupby x({0}, bad, 16);
    } catch (logic_error const&) { threw = true; }
    CHECK(threw);
    CHECK(groupby::is_streaming_groupby_supported(data_type{type_id::FLOAT64}, BGB_STD));
  }
  // the reference's call shape (cudf/groupby.hpp:181-184): span of requests, stream, memory resource -- and the resource really
  // serves BOTH the returned columns and the call's scratch (the C ABI's bgb_allocator): every byte it hands out comes back
  {
    struct counting_resource {
      long allocs = 0, frees = 0; size_t live = 0, peak = 0;
      void* allocate_async(size_t b, cudaStream_t s) { void* p = nullptr; if (cudaMallocAsync(&p, b ? b : 1, s) != cudaSuccess) return nullptr; ++allocs; live += b; peak = std::max(peak, live); return p; }
      void deallocate_async(void* p, size_t b, cudaStream_t s) { cudaFreeAsync(p, s); ++frees; live -= b; }
    } res_mr;
    cudaStream_t st;
    cudaStreamCreate(&st);
    {
      std::vector<K> kk(50000);
      std::vector<int64_t> vv(50000);
      for (size_t i = 0; i < kk.size(); ++i) { kk[i] = (K)(i % 977); vv[i] = (int64_t)i; }
      auto keys = column::from_host<K>(type_id::INT32, kk);
      auto vals = column::from_host<int64_t>(type_id::INT64, vv);
      groupby::groupby gb(table_view({keys->view()}));
      std::vector<groupby::aggregation_request> reqs(1);
      reqs[0].values = vals->view();
      reqs[0].aggregations.push_back(make_sum_aggregation());
      reqs[0].aggregations.push_back(make_count_aggregation());
      auto [tbl, res] = gb.aggregate(reqs, stream_ref(st), device_async_resource_ref(res_mr));
      CHECK(tbl->num_rows() == 977);
      auto [s, sv] = res[0].results[0]->to_host<int64_t>();
      int64_t total = 0;
      for (auto x : s) total += x;
      CHECK(total == (int64_t)50000 * 49999 / 2);
      CHECK(res_mr.allocs >= 5);   // 3 output columns + the table and its scratch
      CHECK(res_mr.live > 0);      // the outputs are alive, the scratch is back
    }
    cudaStreamSynchronize(st);
    CHECK(res_mr.live == 0 && res_mr.allocs == res_mr.frees);
    cudaStreamDestroy(st);
  }
  std::printf(g_fail ? "%d check(s) FAILED\n" : "all C++ host-layer checks passed\n", g_fail);
  return g_fail ? 1 : 0;
}

#include "plan.h"

#include <algorithm>

#include "errors.h"

namespace bgb {

int type_width(int32_t t)
{
  switch (t) {
    case BGB_INT8: case BGB_UINT8: case BGB_BOOL8: return 1;
    case BGB_INT16: case BGB_UINT16: return 2;
    case BGB_INT32: case BGB_UINT32: case BGB_FLOAT32: case BGB_TIMESTAMP_DAYS: case BGB_DURATION_DAYS: case BGB_DECIMAL32: return 4;
    case BGB_INT64: case BGB_UINT64: case BGB_FLOAT64: case BGB_TIMESTAMP_SECONDS: case BGB_TIMESTAMP_MILLISECONDS:
    case BGB_TIMESTAMP_MICROSECONDS: case BGB_TIMESTAMP_NANOSECONDS: case BGB_DURATION_SECONDS: case BGB_DURATION_MILLISECONDS:
    case BGB_DURATION_MICROSECONDS: case BGB_DURATION_NANOSECONDS: case BGB_DECIMAL64: return 8;
    default: return 0;
  }
}

uint8_t type_src(int32_t t)
{
  switch (t) {
    case BGB_INT8: return S_I8;
    case BGB_INT16: return S_I16;
    case BGB_INT32: case BGB_TIMESTAMP_DAYS: case BGB_DURATION_DAYS: case BGB_DECIMAL32: return S_I32;
    case BGB_INT64: case BGB_TIMESTAMP_SECONDS: case BGB_TIMESTAMP_MILLISECONDS: case BGB_TIMESTAMP_MICROSECONDS:
    case BGB_TIMESTAMP_NANOSECONDS: case BGB_DURATION_SECONDS: case BGB_DURATION_MILLISECONDS: case BGB_DURATION_MICROSECONDS:
    case BGB_DURATION_NANOSECONDS: case BGB_DECIMAL64: return S_I64;
    case BGB_UINT8: return S_U8;
    case BGB_UINT16: return S_U16;
    case BGB_UINT32: return S_U32;
    case BGB_UINT64: return S_U64;
    case BGB_FLOAT32: return S_F32;
    case BGB_FLOAT64: return S_F64;
    case BGB_BOOL8: return S_BOOL;
    default: return S_BAD;
  }
}

static bool is_integral(int32_t t) { return (t >= BGB_INT8 && t <= BGB_UINT64) || t == BGB_BOOL8; }
static bool is_floating(int32_t t) { return t == BGB_FLOAT32 || t == BGB_FLOAT64; }
static bool is_duration(int32_t t) { return t >= BGB_DURATION_DAYS && t <= BGB_DURATION_NANOSECONDS; }
static bool is_timestamp(int32_t t) { return t >= BGB_TIMESTAMP_DAYS && t <= BGB_TIMESTAMP_NANOSECONDS; }
static bool is_decimal(int32_t t) { return t == BGB_DECIMAL32 || t == BGB_DECIMAL64 || t == BGB_DECIMAL128; }
static bool is_known_type(int32_t t) { return t > BGB_EMPTY && t <= BGB_STRUCT; }

// target_type_t<Source,k>: /root/reference/cpp/include/cudf/detail/aggregation/aggregation.hpp:878-1012
int32_t target_type(int32_t t, int32_t k)
{
  if (!is_known_type(t)) return -1;
  const bool fixed_width = type_width(t) > 0 || t == BGB_DECIMAL128;
  switch (k) {
    case BGB_MIN: case BGB_MAX: return t;
    case BGB_COUNT_VALID: case BGB_COUNT_ALL: case BGB_ARGMIN: case BGB_ARGMAX: return BGB_INT32;
    case BGB_MERGE_COUNT: return t == BGB_INT32 ? BGB_INT32 : -1;
    case BGB_SUM:
      if (is_integral(t)) return BGB_INT64;
      if (is_floating(t) || is_duration(t) || is_decimal(t)) return t;
      return -1;
    case BGB_PRODUCT: case BGB_SUM_OF_SQUARES:
      if (is_integral(t)) return BGB_INT64;
      if (is_floating(t)) return t;
      return -1;
    case BGB_MEAN:
      if (!fixed_width || is_timestamp(t)) return -1;
      if (is_duration(t) || is_decimal(t)) return t;
      return BGB_FLOAT64;
    case BGB_M2: case BGB_VARIANCE: case BGB_STD:
      return (is_integral(t) || is_floating(t)) ? BGB_FLOAT64 : -1;  // m2_var_std.cu:27-31
    case BGB_SUM_OVERFLOW:
      return ((is_integral(t) && t != BGB_BOOL8 && !(t >= BGB_UINT8 && t <= BGB_UINT64)) || is_decimal(t)) ? BGB_STRUCT : -1;
    case BGB_ANY: case BGB_ALL: return BGB_BOOL8;
    case BGB_MEDIAN: case BGB_QUANTILE: return (is_integral(t) || is_floating(t) || is_decimal(t)) ? BGB_FLOAT64 : -1;
    case BGB_NUNIQUE: return BGB_INT32;
    case BGB_NTH_ELEMENT: return t;
    default: return -2;  // a kind this library does not know: not "invalid", just not served (sort fallback)
  }
}

static bool kind_is_hash_capable(int32_t k)
{
  // is_hash_aggregation (/root/reference/cpp/src/groupby/common/utils.hpp:66-86)
  switch (k) {
    case BGB_SUM_OVERFLOW:
    case BGB_SUM: case BGB_PRODUCT: case BGB_MIN: case BGB_MAX: case BGB_COUNT_VALID: case BGB_COUNT_ALL: case BGB_SUM_OF_SQUARES:
    case BGB_MEAN: case BGB_M2: case BGB_VARIANCE: case BGB_STD: case BGB_ARGMIN: case BGB_ARGMAX: case BGB_MERGE_COUNT: return true;
    default: return false;
  }
}

static uint8_t sum_op(uint8_t src) { return src == S_F64 ? ACC_SUM_F64 : src == S_F32 ? ACC_SUM_F32 : ACC_SUM_I64; }
static uint8_t sumsq_op(uint8_t src) { return src == S_F64 ? ACC_SUMSQ_F64 : src == S_F32 ? ACC_SUMSQ_F32 : ACC_SUMSQ_I64; }
static uint8_t prod_op(uint8_t src) { return src == S_F64 ? ACC_PROD_F64 : src == S_F32 ? ACC_PROD_F32 : ACC_PROD_I64; }

static int get_acc(Plan& p, int value_col, uint8_t op)
{
  for (size_t i = 0; i < p.accs.size(); ++i)
    if (p.accs[i].value_col == value_col && p.accs[i].op == op) return (int)i;
  AccPlan a;
  a.value_col = value_col;
  a.op        = op;
  p.accs.push_back(a);
  return (int)p.accs.size() - 1;
}

static bool same_column(const bgb_column_view& a, const bgb_column_view& b)
{
  return a.data == b.data && a.null_mask == b.null_mask && a.offset == b.offset && a.size == b.size && a.type_id == b.type_id;
}

// DICTIONARY32 values are aggregated as their KEYS' type (hash/groupby.cu:118-120, output_utils.cu:66): the view's `data` is a HOST
// array of two child views {indices, keys}. Resolves the request's values into the view the kernels read (the indices, under the
// parent's offset and null mask, typed like the keys) + the dictionary. Non-dictionary columns pass through.
struct ResolvedValues { bgb_column_view v; const uint8_t* dict; uint8_t idx_width; int rc; };
static ResolvedValues resolve_values(const bgb_column_view& in, int r)
{
  ResolvedValues out{in, nullptr, 0, BGB_OK};
  if (in.type_id != BGB_DICTIONARY32) return out;
  const bgb_column_view* ch = static_cast<const bgb_column_view*>(in.data);
  if (ch == nullptr) { out.rc = set_error(BGB_INVALID_ARGUMENT, "request %d: a DICTIONARY32 view needs data = {indices, keys} child views", r); return out; }
  const bgb_column_view &idx = ch[0], &keys = ch[1];
  const int iw = type_width(idx.type_id), kw = type_width(keys.type_id);
  const bool idx_int = idx.type_id == BGB_INT8 || idx.type_id == BGB_INT16 || idx.type_id == BGB_INT32 || idx.type_id == BGB_UINT8 ||
                       idx.type_id == BGB_UINT16 || idx.type_id == BGB_UINT32;
  if (!idx_int || iw == 0 || kw == 0 || keys.type_id == BGB_DICTIONARY32) {
    out.rc = set_error(BGB_UNSUPPORTED, "request %d: dictionary with index type %d / key type %d is not served by the hash path", r, idx.type_id, keys.type_id);
    return out;
  }
  if (idx.size < in.offset + in.size) { out.rc = set_error(BGB_INVALID_ARGUMENT, "request %d: dictionary indices shorter than the column", r); return out; }
  out.v.type_id = keys.type_id;
  out.v.scale   = keys.scale;
  out.v.data    = idx.data;
  out.v.offset  = in.offset + idx.offset;  // (the parent's mask is indexed by in.offset: see mask_shift below -- children carry offset 0 in cudf)
  out.dict      = static_cast<const uint8_t*>(keys.data) + (size_t)keys.offset * kw;
  out.idx_width = (uint8_t)iw;
  if (idx.offset != 0 && in.null_mask != nullptr) out.rc = set_error(BGB_UNSUPPORTED, "request %d: dictionary indices child with its own offset under a null mask", r);
  return out;
}

int make_plan(const bgb_column_view* keys, int32_t num_keys, int32_t include_null_keys, const bgb_request* requests,
              int32_t num_requests, Plan& plan, bool rank_independent, bool allow_wide)
{
  if (num_keys <= 0 || keys == nullptr) return set_error(BGB_INVALID_ARGUMENT, "groupby needs at least one key column");
  if (num_requests < 0 || (num_requests > 0 && requests == nullptr)) return set_error(BGB_INVALID_ARGUMENT, "bad requests");
  plan.num_rows          = keys[0].size;
  plan.include_null_keys = include_null_keys != 0;
  for (int c = 0; c < num_keys; ++c)
    if (keys[c].size != plan.num_rows) return set_error(BGB_INVALID_ARGUMENT, "key columns differ in size");

  // ---- the reference's own checks, in its order (groupby.cu:225-231)
  for (int r = 0; r < num_requests; ++r)
    if (requests[r].values.size != plan.num_rows)
      return set_error(BGB_INVALID_ARGUMENT, "Size mismatch between request values and groupby keys.");
  for (int r = 0; r < num_requests; ++r)
    for (int a = 0; a < requests[r].num_aggregations; ++a)
      {
        int32_t vt = requests[r].values.type_id;
        if (vt == BGB_DICTIONARY32 && requests[r].values.data != nullptr) vt = static_cast<const bgb_column_view*>(requests[r].values.data)[1].type_id;
        if (target_type(vt, requests[r].aggregations[a].kind) == -1)
          return set_error(BGB_INVALID_ARGUMENT, "Invalid type/aggregation combination.");
      }

  // ---- what this library's hash path serves (everything else: BGB_UNSUPPORTED -> caller's sort fallback)
  if (num_keys > kMaxKeyCols) return set_error(BGB_UNSUPPORTED, "more than %d key columns", kMaxKeyCols);
  int total_bits = 0, null_bits = 0;
  bool any_key_mask = false;
  for (int c = 0; c < num_keys; ++c) {
    int w = type_width(keys[c].type_id);
    if (w == 0) return set_error(BGB_UNSUPPORTED, "key column %d: type %d is not a fixed-width type of <= 8 bytes", c, keys[c].type_id);
    total_bits += 8 * w;
    if (keys[c].null_mask != nullptr) any_key_mask = true;
    if (plan.include_null_keys && (keys[c].null_mask != nullptr || rank_independent)) ++null_bits;
  }
  if (total_bits + null_bits > 128 && !allow_wide) return set_error(BGB_UNSUPPORTED, "packed key wider than 128 bits (%d)", total_bits + null_bits);
  plan.key_words      = (total_bits + null_bits + 63) / 64;
  plan.skip_null_keys = any_key_mask && !plan.include_null_keys;

  // pack widest fields first so no field straddles a 64-bit word; null flags go last
  std::vector<int> order(num_keys);
  for (int c = 0; c < num_keys; ++c) order[c] = c;
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return type_width(keys[a].type_id) > type_width(keys[b].type_id); });
  plan.keys.resize(num_keys);
  int bit = 0;
  for (int c : order) {
    KeyPlan& k = plan.keys[c];
    k.v        = keys[c];
    k.width    = (uint8_t)type_width(keys[c].type_id);
    k.fclass   = keys[c].type_id == BGB_FLOAT32 ? 1 : keys[c].type_id == BGB_FLOAT64 ? 2 : keys[c].type_id == BGB_BOOL8 ? 3 : 0;
    k.shift    = (uint8_t)bit;
    k.null_bit = -1;
    k.nullable_out = plan.include_null_keys && (keys[c].null_mask != nullptr || rank_independent);
    bit += 8 * k.width;
  }
  for (int c = 0; c < num_keys; ++c)
    if (plan.keys[c].nullable_out) plan.keys[c].null_bit = (int8_t)bit++;

  // ---- value columns, accumulators, output recipes
  for (int r = 0; r < num_requests; ++r) {
    bgb_request rq = requests[r];
    plan.request_first_out.push_back((int)plan.outs.size());
    plan.request_num_aggs.push_back(rq.num_aggregations);
    if (rq.num_aggregations == 0) continue;
    const ResolvedValues rv = resolve_values(rq.values, r);
    if (rv.rc != BGB_OK) return rv.rc;
    if (rv.dict != nullptr && rank_independent) return set_error(BGB_UNSUPPORTED, "request %d: dictionary values on a distributed / streaming table", r);
    rq.values = rv.v;  // (from here on: the column the kernels read)
    const uint8_t src = type_src(rq.values.type_id);
    if (src == S_BAD) return set_error(BGB_UNSUPPORTED, "request %d: value type %d not served by the hash path", r, rq.values.type_id);
    int vc = -1;
    for (size_t i = 0; i < plan.values.size(); ++i)
      if (same_column(plan.values[i].v, rq.values) && plan.values[i].dict == rv.dict) { vc = (int)i; break; }
    if (vc < 0) {
      ValuePlan vp;
      vp.v         = rq.values;
      vp.dict      = rv.dict;
      vp.idx_width = rv.idx_width;
      if (rv.dict != nullptr) plan.has_dict = true;
      vp.src       = src;
      // rank_independent: another rank's shard of this column may hold nulls, and every rank must lay out the same cells
      vp.has_nulls = rank_independent || (rq.values.null_count > 0 && rq.values.null_mask != nullptr);
      vp.has_mask  = rq.values.null_mask != nullptr;
      plan.values.push_back(vp);
      vc = (int)plan.values.size() - 1;
    }
    const ValuePlan& V = plan.values[vc];
    for (int a = 0; a < rq.num_aggregations; ++a) {
      const int32_t kind = rq.aggregations[a].kind;
      if (!kind_is_hash_capable(kind)) return set_error(BGB_UNSUPPORTED, "aggregation kind %d is not a hash aggregation", kind);
      OutPlan o;
      o.request = r; o.agg = a;
      o.type_id = target_type(rq.values.type_id, kind);
      o.scale   = rq.values.scale;
      o.src     = src;
      o.ddof    = rq.aggregations[a].ddof;
      o.out_width = (uint8_t)type_width(o.type_id);
      // output_utils.cu:61-64: nullable iff not a COUNT and the input has nulls
      o.nullable = V.has_nulls;
      auto need_cnt = [&]() { return get_acc(plan, vc, ACC_COUNT_VALID); };
      switch (kind) {
        case BGB_SUM: o.acc_op = sum_op(src); o.acc_a = get_acc(plan, vc, o.acc_op);
          o.op = o.acc_op == ACC_SUM_F64 ? OUT_COPY_F64 : o.acc_op == ACC_SUM_F32 ? OUT_COPY_F32 : OUT_COPY_I64; break;
        case BGB_SUM_OF_SQUARES: o.acc_op = sumsq_op(src); o.acc_a = get_acc(plan, vc, o.acc_op);
          o.op = o.acc_op == ACC_SUMSQ_F64 ? OUT_COPY_F64 : o.acc_op == ACC_SUMSQ_F32 ? OUT_COPY_F32 : OUT_COPY_I64; break;
        case BGB_PRODUCT: o.acc_op = prod_op(src); o.acc_a = get_acc(plan, vc, o.acc_op);
          o.op = o.acc_op == ACC_PROD_F64 ? OUT_COPY_F64 : o.acc_op == ACC_PROD_F32 ? OUT_COPY_F32 : OUT_COPY_I64; break;
        case BGB_MIN: o.acc_op = ACC_MIN; o.acc_a = get_acc(plan, vc, ACC_MIN); o.op = OUT_MINMAX; break;
        case BGB_MAX: o.acc_op = ACC_MAX; o.acc_a = get_acc(plan, vc, ACC_MAX); o.op = OUT_MINMAX; break;
        case BGB_MERGE_COUNT:  // sum of partial counts, narrowed back to size_type (a merged group has < 2^31 rows)
          o.acc_op = ACC_SUM_I64; o.acc_a = get_acc(plan, vc, ACC_SUM_I64); o.op = OUT_COPY_I64; o.nullable = false; break;
        case BGB_COUNT_VALID: o.acc_op = ACC_COUNT_VALID; o.acc_a = need_cnt(); o.op = OUT_COUNT; o.nullable = false; break;
        case BGB_COUNT_ALL: o.acc_op = ACC_COUNT_ALL; o.acc_a = get_acc(plan, -1, ACC_COUNT_ALL); o.op = OUT_COUNT; o.nullable = false; break;
        case BGB_ARGMIN: case BGB_ARGMAX:
          if (type_width(rq.values.type_id) > 4) {
            // 64-bit values do not fit a packed (value, row) cell: MIN / MAX first, then a second pass over the rows records
            // the smallest row whose value equals the group's extremum (kernels_generic.cu: argrow_kernel)
            const int ext = get_acc(plan, vc, kind == BGB_ARGMIN ? ACC_MIN : ACC_MAX);
            o.acc_op = kind == BGB_ARGMIN ? ACC_ARGROW_MIN : ACC_ARGROW_MAX; o.acc_a = get_acc(plan, vc, o.acc_op); o.op = OUT_ARG;
            plan.accs[o.acc_a].aux_acc = ext;
            break;
          }
          o.acc_op = kind == BGB_ARGMIN ? ACC_ARGMIN : ACC_ARGMAX; o.acc_a = get_acc(plan, vc, o.acc_op); o.op = OUT_ARG; break;
        case BGB_SUM_OVERFLOW:
          // struct{sum: Source, overflow: bool} (device_aggregators.cuh:131-160, output_utils.cu:83-104). The reference raises the
          // flag when one atomic add overflows, which depends on the order the adds land in; here the EXACT total is kept (two
          // 64-bit sums) and the flag says whether it fits the Source type -- the outcome every order agrees on when it is set.
          if (rank_independent) return set_error(BGB_UNSUPPORTED, "SUM_OVERFLOW is not served by the streaming / distributed tables");
          if (o.type_id != BGB_STRUCT) return set_error(BGB_INVALID_ARGUMENT, "request %d: SUM_OVERFLOW needs a signed integer or decimal32/64 column", r);
          o.acc_op = ACC_SUM_LO32; o.acc_a = get_acc(plan, vc, ACC_SUM_LO32); o.acc_b = get_acc(plan, vc, ACC_SUM_HI32);
          o.op = OUT_SUMOV_SUM; o.out_width = (uint8_t)type_width(rq.values.type_id);
          break;
        case BGB_MEAN:
          // MEAN -> SUM + COUNT_VALID (extract_single_pass_aggs.cpp:62-73), finalised as double(sum)/count
          o.acc_op = sum_op(src); o.acc_a = get_acc(plan, vc, o.acc_op); o.acc_cnt = need_cnt();
          o.op = o.type_id == BGB_FLOAT64 ? OUT_MEAN : OUT_MEAN_INT;
          break;
        case BGB_M2: case BGB_VARIANCE: case BGB_STD:
          // -> SUM_OF_SQUARES + SUM + COUNT_VALID (extract_single_pass_aggs.cpp:75-109)
          o.acc_op = sum_op(src); o.acc_b = get_acc(plan, vc, sumsq_op(src)); o.acc_a = get_acc(plan, vc, o.acc_op); o.acc_cnt = need_cnt();
          o.op = kind == BGB_M2 ? OUT_M2 : kind == BGB_VARIANCE ? OUT_VAR : OUT_STD;
          o.nullable = kind != BGB_M2;  // m2 has no mask; var/std get one whenever a group has n - ddof <= 0
          break;
        default: return set_error(BGB_UNSUPPORTED, "aggregation kind %d", kind);
      }
      // validity of SUM/MIN/MAX/... = "saw at least one valid value" -> needs the column's COUNT_VALID
      if (o.nullable && o.acc_cnt < 0 && o.op != OUT_COUNT) o.acc_cnt = need_cnt();
      plan.outs.push_back(o);
    }
  }
  return BGB_OK;
}

}  // namespace bgb

extern "C" int32_t bgb_target_type(int32_t values_type_id, int32_t kind) { return bgb::target_type(values_type_id, kind); }

extern "C" int bgb_can_use_hash_groupby(const bgb_column_view* keys, int32_t num_keys, const bgb_request* requests,
                                        int32_t num_requests, int32_t* out)
{
  // same eligibility as bgb_groupby_build: key images wider than 128 bits are planned (build serves them by key repacking or
  // through row-index slots); the null policy does not change the answer
  bgb::Plan p;
  int rc = bgb::make_plan(keys, num_keys, 0, requests, num_requests, p, false, /*allow_wide=*/true);
  if (out) *out = (rc == BGB_OK);
  return rc == BGB_UNSUPPORTED ? BGB_OK : rc;
}

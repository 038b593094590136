#include "fastpaths.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "errors.h"
#include "kernels_smem.h"
#include "kernels_sorted.h"
#include "kernels_tiny.h"

namespace bgb {

static bool aligned_to(const void* base, int64_t offset, int width, int bytes)
{
  return (((uintptr_t)base + (uintptr_t)(offset * width)) % (uintptr_t)bytes) == 0;
}

SmemEligibility smem_eligibility(const Plan& pl)
{
  SmemEligibility e;
  if (pl.key_words != 1 || pl.num_rows == 0) return e;
  // key shape first: the estimator and the tile-sorting kernel use it whatever the accumulator set is
  const KeyPlan& k0 = pl.keys[0];
  e.integer_keys = true;
  for (const KeyPlan& k : pl.keys) if (k.fclass != 0) e.integer_keys = false;
  if (pl.keys.size() == 1 && k0.null_bit < 0) {
    const uint8_t src = type_src(k0.v.type_id);
    if (src == S_I8 || src == S_I16 || src == S_I32 || src == S_I64) e.key_flip = 1ull << (8 * k0.width - 1);
  }
  for (size_t i = 0; i < pl.accs.size(); ++i) {
    const AccPlan& a = pl.accs[i];
    if (a.op == ACC_SUM_I64) { if (e.sum_acc >= 0) return e; e.sum_acc = (int)i; }
    else if (a.op == ACC_COUNT_VALID) { if (e.cntv_acc >= 0) return e; e.cntv_acc = (int)i; }
    else if (a.op == ACC_COUNT_ALL) e.cnta_acc = (int)i;
    else return e;
    if (a.value_col >= 0) { if (e.value_col >= 0 && e.value_col != a.value_col) return e; e.value_col = a.value_col; }
  }
  if (e.value_col >= 0) {
    const ValuePlan& v = pl.values[e.value_col];
    if (e.cnta_acc >= 0 && v.has_mask) return e;  // COUNT_ALL differs from COUNT_VALID: generic path
    if (e.sum_acc >= 0) {
      if (v.src == S_I64 || v.src == S_U64) e.value_width = 8;
      else if (v.src == S_I32 || v.src == S_U32) e.value_width = 4;
      else return e;
      if (!aligned_to(v.v.data, v.v.offset, e.value_width, 2 * e.value_width)) return e;
    }
  }
  int key_bits = 0;
  for (const KeyPlan& k : pl.keys) key_bits = std::max(key_bits, std::max((int)k.shift + 8 * k.width, (int)k.null_bit + 1));
  if (pl.keys.size() == 1 && (k0.width == 4 || k0.width == 8) && k0.null_bit < 0 && aligned_to(k0.v.data, k0.v.offset, k0.width, 2 * k0.width)) {
    e.key_mode = 1; e.smem_key_bytes = k0.width;
  } else {
    e.key_mode = 0; e.smem_key_bytes = key_bits <= 32 ? 4 : 8;
  }
  e.ok = true;
  return e;
}

static int bit_length(uint64_t x) { int b = 0; while (x) { ++b; x >>= 1; } return b; }

static int run_smem_path(const BuildContext& ctx, const SmemEligibility& e, uint32_t slots, uint32_t dense, bool direct, uint64_t key_base)
{
  const Plan& pl = *ctx.plan;
  // global table init (keys + every accumulator) with the generic init kernel
  AggParams ip;
  memset(&ip, 0, sizeof ip);
  ip.table = ctx.table; ip.init_keys = 1; ip.num_accs = (int)pl.accs.size();
  for (size_t i = 0; i < pl.accs.size(); ++i) { ip.acc[i].op = pl.accs[i].op; ip.acc[i].cells = pl.accs[i].cells; }
  ctx.prof->begin("init_table");
  int rc = check_cuda(launch_init_table(ip, ctx.stream), "init_table");
  ctx.prof->end();
  if (rc) return rc;

  SmemParams sp;
  memset(&sp, 0, sizeof sp);
  sp.a.num_rows = pl.num_rows; sp.a.num_key_cols = (int)pl.keys.size(); sp.a.skip_null_keys = pl.skip_null_keys;
  for (size_t c = 0; c < pl.keys.size(); ++c) {
    const KeyPlan& k = pl.keys[c];
    sp.a.kc[c] = KeyColDesc{(const uint8_t*)k.v.data, k.v.null_mask, k.v.offset, k.width, k.fclass, k.shift, k.null_bit};
  }
  sp.a.table = ctx.table;
  if (e.value_col >= 0) {
    const ValuePlan& v = pl.values[e.value_col];
    sp.v_data = (const uint8_t*)v.v.data; sp.v_mask = v.v.null_mask; sp.v_offset = v.v.offset;
    sp.v_signed = (v.src == S_I32 || v.src == S_I64);
  }
  sp.sum_cells  = e.sum_acc >= 0 ? (unsigned long long*)pl.accs[e.sum_acc].cells : nullptr;
  sp.cntv_cells = e.cntv_acc >= 0 ? (int*)pl.accs[e.cntv_acc].cells : nullptr;
  sp.cnta_cells = e.cnta_acc >= 0 ? (int*)pl.accs[e.cnta_acc].cells : nullptr;
  sp.slots      = slots;
  sp.dense      = dense;
  sp.direct     = direct;
  sp.key_flip   = e.key_flip;
  sp.key_base   = key_base;
  // count field width c and value range m: a CTA flushes before 2^c rows, partial sums stay below 2^(c+m) <= 2^(63-c)
  int m = e.value_width == 0 ? 0 : bit_length(ctx.max_abs_value) + 1;
  if (e.value_width == 4) m = std::min(m, 31);
  int c = std::min(24, (63 - m) / 2);
  if (c < 16) { c = 16; m = 31; }
  sp.cnt_bits = c; sp.val_bits = std::min(m, 63 - 2 * c);
  sp.debug = getenv("BGB_DEBUG") != nullptr;
  if (sp.debug) {
    cudaMemsetAsync(ctx.table.status + 4, 0, 16, ctx.stream);
    fprintf(stderr, "[bgb] smem path: direct=%d slots=%u dense=%u c=%d m=%d g_raw=%.0f max_abs=%llu key_mode=%d sk=%d vw=%d\n", (int)direct, sp.slots, sp.dense,
            sp.cnt_bits, sp.val_bits, ctx.g_raw, (unsigned long long)ctx.max_abs_value, e.key_mode, e.smem_key_bytes, e.value_width);
  }
  ctx.prof->begin("aggregate_smem");
  rc = check_cuda(launch_aggregate_smem(sp, e.smem_key_bytes, e.key_mode, e.value_width, ctx.num_sms, ctx.stream), "aggregate_smem");
  ctx.prof->end();
  return rc;
}

// ---- tile-sorting kernel (kernels_sorted.cu): dense integer key range, any accumulator set without ARGMIN/ARGMAX
// (the key shape is checked by the caller: one-word integer keys with a known range)
static bool sorted_plan(const Plan& pl, SortedParams* out)
{
  if (pl.accs.empty() || pl.accs.size() > (size_t)kMaxAccs) return false;
  SortedParams local;
  SortedParams& sp = out ? *out : local;
  memset(&sp, 0, sizeof sp);
  int col_of_value[64];
  for (int& c : col_of_value) c = -1;
  for (size_t i = 0; i < pl.accs.size(); ++i) {
    const AccPlan& a = pl.accs[i];
    if (a.op == ACC_ARGMIN || a.op == ACC_ARGMAX || a.op == ACC_ARGROW_MIN || a.op == ACC_ARGROW_MAX) return false;
    if (a.op == ACC_SUM_LO32 || a.op == ACC_SUM_HI32) return false;  // SUM_OVERFLOW: generic / bucketed kernels
    int col = -1;
    if (a.value_col >= 0) {
      if (a.value_col >= 64) return false;
      if (col_of_value[a.value_col] < 0) {
        if (sp.num_cols == kSortedMaxCols) return false;
        const ValuePlan& v = pl.values[a.value_col];
        sp.col[sp.num_cols] = SortedCol{(const uint8_t*)v.v.data, v.v.null_mask, v.v.offset, (uint8_t)type_width(v.v.type_id), v.src};
        if (v.v.null_mask) sp.any_mask = 1;
        col_of_value[a.value_col] = sp.num_cols++;
      }
      col = col_of_value[a.value_col];
    }
    sp.acc[i] = SortedAcc{a.cells, a.op, (int8_t)col};
  }
  sp.num_accs = (int)pl.accs.size();
  return true;
}

uint32_t sorted_max_groups(const Plan& pl, bool forced)
{
  SortedParams sp;
  if (!sorted_plan(pl, &sp)) return 0;
  const int nc = std::max(1, sp.num_cols);
  // automatic choice: two CTAs per SM and at most 4 groups per thread -- beyond that the per-tile walk over the owned groups
  // costs more than the L2 atomics of the generic kernel (measured at 4000 groups: 5.2 ms against 4.2 ms per 1e8 rows)
  const size_t budget = forced ? smem_table_max_bytes() : (size_t)112 * 1024;
  uint32_t g = forced ? kSortedMaxGroups : kSortedMaxGroups / 2;
  while (g > 0 && sorted_smem_bytes(g, sp.num_accs, nc) > budget) g -= 64;
  return g;
}

// hashed variant: the largest power-of-two slot count whose tables fit (two CTAs per SM unless forced)
static uint32_t sorted_max_slots(const Plan& pl, bool forced)
{
  SortedParams sp;
  if (!sorted_plan(pl, &sp)) return 0;
  const int nc = std::max(1, sp.num_cols);
  const size_t budget = forced ? smem_table_max_bytes() : (size_t)112 * 1024;
  uint32_t s = forced ? kSortedMaxGroups : kSortedMaxGroups / 2;
  while (s >= 64 && sorted_smem_bytes(s, sp.num_accs, nc, true) > budget) s >>= 1;
  return s >= 64 ? s : 0;
}

static int run_sorted_path(const BuildContext& ctx, const SmemEligibility& e, SortedParams& sp, uint32_t groups, uint64_t key_base, bool hashed = false)
{
  const Plan& pl = *ctx.plan;
  AggParams ip;
  memset(&ip, 0, sizeof ip);
  ip.table = ctx.table; ip.init_keys = 1; ip.num_accs = (int)pl.accs.size();
  for (size_t i = 0; i < pl.accs.size(); ++i) { ip.acc[i].op = pl.accs[i].op; ip.acc[i].cells = pl.accs[i].cells; }
  ctx.prof->begin("init_table");
  int rc = check_cuda(launch_init_table(ip, ctx.stream), "init_table");
  ctx.prof->end();
  if (rc) return rc;
  sp.a.num_rows = pl.num_rows; sp.a.num_key_cols = (int)pl.keys.size(); sp.a.skip_null_keys = pl.skip_null_keys;
  for (size_t c = 0; c < pl.keys.size(); ++c) {
    const KeyPlan& k = pl.keys[c];
    sp.a.kc[c] = KeyColDesc{(const uint8_t*)k.v.data, k.v.null_mask, k.v.offset, k.width, k.fclass, k.shift, k.null_bit};
  }
  sp.a.table  = ctx.table;
  sp.groups   = groups;
  sp.key_flip = hashed ? 0 : e.key_flip;
  sp.key_base = key_base;
  sp.hashed   = hashed;
  sp.key_mode = pl.keys.size() == 1 && pl.keys[0].v.null_mask == nullptr && pl.keys[0].fclass == 0 && pl.keys[0].null_bit < 0;
  if (sp.key_mode == 1 && pl.keys[0].width == 8) sp.key_mode = 2;
  sp.fast8 = sp.num_cols > 0;
  for (int c = 0; c < sp.num_cols; ++c)
    sp.fast8 &= sp.col[c].width == 8 && sp.col[c].mask == nullptr && (sp.col[c].src == S_I64 || sp.col[c].src == S_U64 || sp.col[c].src == S_F64);
  ctx.prof->begin("aggregate_sorted");
  rc = check_cuda(launch_aggregate_sorted(sp, ctx.num_sms, ctx.stream), "aggregate_sorted");
  ctx.prof->end();
  return rc;
}

// ---- tiny-cardinality path (kernels_tiny.cu): per-thread private accumulators, no atomics on the row path
static bool tiny_plan(const Plan& pl, TinyParams& tp)
{
  memset(&tp, 0, sizeof tp);
  if (pl.key_words != 1 || pl.accs.empty() || pl.accs.size() > (size_t)kMaxAccs) return false;
  auto column_of = [&](int value_col) -> int {  // COUNT_ALL (value_col -1) rides on a data-less, mask-less pseudo column
    for (int j = 0; j < tp.num_cols; ++j) {
      const TinyCol& c = tp.cols[j];
      if (value_col < 0 ? (c.data == nullptr && c.width == 0)
                        : (c.width != 0 && c.data == (const uint8_t*)pl.values[value_col].v.data && c.mask == pl.values[value_col].v.null_mask &&
                           c.offset == pl.values[value_col].v.offset && c.src == pl.values[value_col].src))
        return j;
    }
    if (tp.num_cols == kTinyPlanCols) return -1;
    TinyCol& c = tp.cols[tp.num_cols];
    c.acc_sum = c.acc_sumsq = c.acc_min = c.acc_max = c.acc_count = -1;
    c.sum_op = ACC_SUM_I64;
    if (value_col >= 0) {
      const ValuePlan& v = pl.values[value_col];
      c.data = (const uint8_t*)v.v.data; c.mask = v.v.null_mask; c.offset = v.v.offset; c.src = v.src; c.width = (uint8_t)type_width(v.v.type_id);
      c.sum_op = v.src == S_F64 ? ACC_SUM_F64 : v.src == S_F32 ? ACC_SUM_F32 : ACC_SUM_I64;
    } else { c.data = nullptr; c.mask = nullptr; c.offset = 0; c.src = S_I32; c.width = 0; }
    return tp.num_cols++;
  };
  for (size_t i = 0; i < pl.accs.size(); ++i) {
    const AccPlan& a = pl.accs[i];
    const int j = column_of(a.value_col);
    if (j < 0) return false;
    TinyCol& c = tp.cols[j];
    int* where = nullptr;
    switch (a.op) {
      case ACC_SUM_I64: case ACC_SUM_F64: case ACC_SUM_F32: where = &c.acc_sum; break;
      case ACC_SUMSQ_I64: case ACC_SUMSQ_F64: case ACC_SUMSQ_F32: where = &c.acc_sumsq; c.need_sq = 1; break;
      case ACC_MIN: where = &c.acc_min; c.need_minmax = 1; break;
      case ACC_MAX: where = &c.acc_max; c.need_minmax = 1; break;
      case ACC_COUNT_VALID: case ACC_COUNT_ALL: where = &c.acc_count; break;
      default: return false;  // PRODUCT, ARGMIN/ARGMAX: general path
    }
    if (*where >= 0) return false;  // (the planner already merges identical accumulators)
    *where = (int)i;
  }
  int cells = 0;
  for (int j = 0; j < tp.num_cols; ++j) { tp.cols[j].cell_base = cells; cells += 2 + (tp.cols[j].need_sq ? 1 : 0) + (tp.cols[j].need_minmax ? 2 : 0); }
  tp.cells = cells;
  return true;
}

static int run_tiny_path(const BuildContext& ctx, TinyParams& tp, bool allow_smem_variant, bool& used_reg, bool& done)
{
  used_reg = false;
  done     = false;
  const Plan& pl = *ctx.plan;
  tp.a.num_rows = pl.num_rows; tp.a.num_key_cols = (int)pl.keys.size(); tp.a.skip_null_keys = pl.skip_null_keys;
  for (size_t c = 0; c < pl.keys.size(); ++c) {
    const KeyPlan& k = pl.keys[c];
    tp.a.kc[c] = KeyColDesc{(const uint8_t*)k.v.data, k.v.null_mask, k.v.offset, k.width, k.fclass, k.shift, k.null_bit};
  }
  tp.a.table = ctx.table; tp.a.init_keys = 1; tp.a.num_accs = (int)pl.accs.size();
  for (size_t i = 0; i < pl.accs.size(); ++i) { tp.a.acc[i].op = pl.accs[i].op; tp.a.acc[i].cells = pl.accs[i].cells; }
  TinyParams subs[kTinyPlanCols];
  const bool few     = ctx.opt.force_path == 6 || ctx.g_raw <= 8.5;
  const int nsub     = (ctx.opt.force_path != 5 && few) ? tiny_reg_split(tp, subs) : 0;
  const bool use_reg = ctx.opt.force_path != 5 && few && (nsub > 0 || tiny_reg_eligible(tp));
  if (!use_reg && ctx.opt.force_path == 6) return set_error(BGB_UNSUPPORTED, "force_path 6: this shape is not served by the register-resident tiny kernel");
  if (!use_reg && !allow_smem_variant) return BGB_OK;
  done = true;
  ctx.prof->begin("init_table");
  int rc = check_cuda(launch_init_table(tp.a, ctx.stream), "init_table");
  ctx.prof->end();
  if (rc) return rc;
  // at most 8 groups and a shape the register kernel serves: accumulators live in registers (force_path 6 insists on it,
  // 5 on the shared-memory variant)
  if (use_reg) {
    used_reg = true;
    const int gcap = ctx.g_raw <= 4.5 && ctx.opt.force_path != 6 ? 4 : 8;
    ctx.prof->begin("aggregate_tiny_reg");
    if (nsub == 0) rc = check_cuda(launch_aggregate_tiny_reg(tp, gcap, ctx.num_sms, ctx.stream), "aggregate_tiny_reg");
    for (int i = 0; i < nsub && !rc; ++i) rc = check_cuda(launch_aggregate_tiny_reg(subs[i], gcap, ctx.num_sms, ctx.stream), "aggregate_tiny_reg");
    ctx.prof->end();
    return rc;
  }
  ctx.prof->begin("aggregate_tiny");
  rc = check_cuda(launch_aggregate_tiny(tp, ctx.num_sms, ctx.stream), "aggregate_tiny");
  ctx.prof->end();
  return rc;
}

int try_fast_paths(const BuildContext& ctx, bool& done, std::string& path)
{
  done = false;
  const Plan& pl = *ctx.plan;
  if (ctx.opt.force_path == 1) return BGB_OK;  // generic kernel only
  const bool force_tiny = ctx.opt.force_path == 5 || ctx.opt.force_path == 6;
  if ((ctx.opt.force_path == 0 && !ctx.skip_tiny && ctx.g_raw <= 24 && pl.num_rows >= 65536) || force_tiny) {
    TinyParams tp;
    if (tiny_plan(pl, tp)) {
      const int cap = (int)std::min<size_t>(60, (200 * 1024) / ((size_t)tp.cells * kTinyThreads * 8));
      // (the register variant holds 8 groups whatever the cell count; run_tiny_path picks it when the shape allows)
      const bool smem_ok = tp.num_cols <= kTinyMaxCols && cap >= 2 && (force_tiny || cap >= ctx.g_raw + 1);
      if (smem_ok || (ctx.g_raw <= 8.5 && ctx.opt.force_path != 5)) {
        tp.group_cap = cap;
        bool used_reg = false;
        int rc = run_tiny_path(ctx, tp, smem_ok, used_reg, done);
        if (rc) return rc;
        if (done) { path = used_reg ? "tiny-reg" : "tiny"; return BGB_OK; }
      }
    } else if (ctx.opt.force_path == 6) return set_error(BGB_UNSUPPORTED, "force_path 6: this plan has no tiny-cardinality form");
  }
  const SmemEligibility e = smem_eligibility(pl);
  if (e.ok && (ctx.opt.force_path == 0 || ctx.opt.force_path == 2 || ctx.opt.force_path == 4)) {
    const size_t bytes = smem_table_max_bytes();
    // (a) direct addressing: integer keys whose sampled range fits -> dense id = key - min, no probing at all. Keys
    //     outside [min - pad, min - pad + rows) (the sample can miss the extremes) take the global-atomic bypass.
    const uint32_t direct_rows_max = (uint32_t)((bytes - 1024) * 32 / (8 * 32 + 4));
    const bool range_known = ctx.key_min <= ctx.key_max;
    const uint64_t range   = range_known ? ctx.key_max - ctx.key_min + 1 : ~0ull;
    if (e.integer_keys && range_known && range <= direct_rows_max && pl.num_rows >= 65536 && ctx.opt.force_path != 2) {
      const uint64_t pad  = std::min<uint64_t>({(uint64_t)64, (direct_rows_max - range) / 2, ctx.key_min});
      const uint64_t base = ctx.key_min - pad;
      const uint32_t rows = (uint32_t)std::min<uint64_t>(direct_rows_max, range + 2 * pad + 64);
      int rc = run_smem_path(ctx, e, 0, rows, true, base);
      if (rc) return rc;
      done = true; path = "smem-direct";
      return BGB_OK;
    }
    // (b) hashed: split the shared memory between dense group rows (key + 8 B of accumulators) and the 2-byte index
    //     slots: rows for the estimated groups plus slack for insert races, then as many index slots as fit (load
    //     factor <= 0.5 whenever the memory allows it)
    const int row_bytes   = e.smem_key_bytes + 8;
    const double g        = std::max(64.0, ctx.g_raw);
    const uint32_t d_max  = (uint32_t)std::min<size_t>(65000, (bytes - 2 * 1024) / row_bytes);
    const bool fits       = 1.02 * g + 256 <= d_max;
    if ((fits && pl.num_rows >= 65536) || ctx.opt.force_path == 2) {
      uint32_t dense = (uint32_t)std::min<double>(d_max, std::ceil(1.08 * g) + 384);
      uint32_t slots = (uint32_t)std::min<double>((double)((bytes - (size_t)dense * row_bytes) / 2), std::max(4096.0, 4.0 * g));
      if (slots < 1.25 * dense) {  // cardinality above what one CTA can hold: keep the index sparse enough to terminate
        dense = (uint32_t)(bytes / (row_bytes + 2.5));
        slots = (uint32_t)((bytes - (size_t)dense * row_bytes) / 2);
      }
      slots &= ~1u;
      int rc = run_smem_path(ctx, e, slots, dense, false, 0);
      if (rc) return rc;
      done = true; path = "smem";
      return BGB_OK;
    }
  }
  // (c) any other accumulator set over a dense integer key range of a few thousand ids: tile-sorting kernel
  if (ctx.opt.force_path == 0 || ctx.opt.force_path == 8) {
    SortedParams sp;
    const uint32_t gmax    = sorted_max_groups(pl, ctx.opt.force_path == 8);
    const bool range_known = ctx.key_min <= ctx.key_max;
    const uint64_t range   = range_known ? ctx.key_max - ctx.key_min + 1 : ~0ull;
    if (gmax != 0 && e.integer_keys && range_known && range <= gmax && (pl.num_rows >= 65536 || ctx.opt.force_path == 8) && sorted_plan(pl, &sp)) {
      const uint64_t pad    = std::min<uint64_t>({(uint64_t)64, (gmax - range) / 2, ctx.key_min});
      const uint32_t groups = (uint32_t)std::min<uint64_t>(gmax, range + 2 * pad + 64);
      int rc = run_sorted_path(ctx, e, sp, groups, ctx.key_min - pad);
      if (rc) return rc;
      done = true; path = "sorted";
      return BGB_OK;
    }
    // sparse integer keys, float keys, several key columns: the id is the key's slot in a small hash table of the CTA
    const uint32_t smax = sorted_max_slots(pl, ctx.opt.force_path == 8);
    if (smax != 0 && pl.key_words == 1 && (ctx.opt.force_path == 8 || (pl.num_rows >= 65536 && ctx.g_raw <= 0.6 * smax)) && sorted_plan(pl, &sp)) {
      uint32_t slots = smax;
      while (slots > 64 && ctx.g_raw <= 0.3 * slots) slots >>= 1;   // (a smaller table leaves more shared memory to the L1)
      int rc = run_sorted_path(ctx, e, sp, slots, 0, true);
      if (rc) return rc;
      done = true; path = "sorted-hashed";
      return BGB_OK;
    }
    if (ctx.opt.force_path == 8) return set_error(BGB_UNSUPPORTED, "force_path 8: the tile-sorting kernel needs a one-word key image, at most %d value columns and no ARGMIN/ARGMAX", kSortedMaxCols);
  }
  return BGB_OK;
}

}  // namespace bgb

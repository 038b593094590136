/*
 * TEST INFRASTRUCTURE ONLY -- see groupby_oracle.h for the role of this file and the reference lines it restates.
 * Plain C11 (+ optional OpenMP). Sequential per thread; threads own contiguous row ranges and their partial
 * results are merged with the associative merge rules (sum+sum, min(min), count+count, ...), the same
 * decomposition the reference's distributed callers use (python/cudf_polars/cudf_polars/streaming/groupby.py:322-435).
 */
#include "groupby_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static _Thread_local char g_err[256];
const char* orc_last_error(void) { return g_err; }
#define FAIL(code, ...) do { snprintf(g_err, sizeof g_err, __VA_ARGS__); return (code); } while (0)

uint64_t orc_splitmix64(uint64_t seed, uint64_t row)
{
  uint64_t z = seed * 0x1000003ull + row;
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

/* ---------------------------------------------------------------- type helpers */
typedef enum { CLS_SINT, CLS_UINT, CLS_BOOL, CLS_F32, CLS_F64, CLS_BAD } cls_t;

int orc_type_size(int t)
{
  switch (t) {
    case ORC_INT8: case ORC_UINT8: case ORC_BOOL8: return 1;
    case ORC_INT16: case ORC_UINT16: return 2;
    case ORC_INT32: case ORC_UINT32: case ORC_FLOAT32: case ORC_TIMESTAMP_DAYS: case ORC_DURATION_DAYS: case ORC_DECIMAL32: return 4;
    case ORC_INT64: case ORC_UINT64: case ORC_FLOAT64: case ORC_TIMESTAMP_SECONDS: case ORC_TIMESTAMP_MILLISECONDS:
    case ORC_TIMESTAMP_MICROSECONDS: case ORC_TIMESTAMP_NANOSECONDS: case ORC_DURATION_SECONDS: case ORC_DURATION_MILLISECONDS:
    case ORC_DURATION_MICROSECONDS: case ORC_DURATION_NANOSECONDS: case ORC_DECIMAL64: return 8;
    default: return 0;
  }
}
static cls_t type_class(int t)
{
  switch (t) {
    case ORC_INT8: case ORC_INT16: case ORC_INT32: case ORC_INT64: case ORC_TIMESTAMP_DAYS: case ORC_TIMESTAMP_SECONDS:
    case ORC_TIMESTAMP_MILLISECONDS: case ORC_TIMESTAMP_MICROSECONDS: case ORC_TIMESTAMP_NANOSECONDS: case ORC_DURATION_DAYS:
    case ORC_DURATION_SECONDS: case ORC_DURATION_MILLISECONDS: case ORC_DURATION_MICROSECONDS: case ORC_DURATION_NANOSECONDS:
    case ORC_DECIMAL32: case ORC_DECIMAL64: return CLS_SINT;
    case ORC_UINT8: case ORC_UINT16: case ORC_UINT32: case ORC_UINT64: return CLS_UINT;
    case ORC_BOOL8: return CLS_BOOL;
    case ORC_FLOAT32: return CLS_F32;
    case ORC_FLOAT64: return CLS_F64;
    default: return CLS_BAD;
  }
}
static int is_integral(int t) { return (t >= ORC_INT8 && t <= ORC_UINT64) || t == ORC_BOOL8; }  /* std::is_integral_v */
static int is_floating(int t) { return t == ORC_FLOAT32 || t == ORC_FLOAT64; }
static int is_numeric(int t) { return is_integral(t) || is_floating(t); }
static int is_duration(int t) { return t >= ORC_DURATION_DAYS && t <= ORC_DURATION_NANOSECONDS; }
static int is_timestamp(int t) { return t >= ORC_TIMESTAMP_DAYS && t <= ORC_TIMESTAMP_NANOSECONDS; }
static int is_decimal(int t) { return t == ORC_DECIMAL32 || t == ORC_DECIMAL64; }

/* target_type_t<Source,k>: cpp/include/cudf/detail/aggregation/aggregation.hpp:878-1012 */
int orc_target_type(int t, int k)
{
  if (type_class(t) == CLS_BAD) return -1;
  switch (k) {
    case ORC_MIN: case ORC_MAX: return t;
    case ORC_COUNT_VALID: case ORC_COUNT_ALL: case ORC_ARGMIN: case ORC_ARGMAX: return ORC_INT32;
    case ORC_SUM:
      if (is_integral(t)) return ORC_INT64;
      if (is_floating(t) || is_duration(t) || is_decimal(t)) return t;
      return -1; /* timestamps cannot be summed */
    case ORC_PRODUCT: case ORC_SUM_OF_SQUARES:
      if (is_integral(t)) return ORC_INT64;
      if (is_floating(t)) return t;
      return -1;
    /* SUM_OVERFLOW: struct{sum: Source, overflow: bool} for signed integers and decimals
     * (device_aggregators.cuh:131-160, output_utils.cu:83-104; decimal128 is outside this path) */
    case ORC_SUM_OVERFLOW:
      return (t == ORC_INT8 || t == ORC_INT16 || t == ORC_INT32 || t == ORC_INT64 || is_decimal(t)) ? t : -1;
    case ORC_SUM_OVERFLOW_FLAG:
      return (t == ORC_INT8 || t == ORC_INT16 || t == ORC_INT32 || t == ORC_INT64 || is_decimal(t)) ? ORC_BOOL8 : -1;
    case ORC_MEAN:
      if (is_timestamp(t)) return -1;
      if (is_duration(t) || is_decimal(t)) return t;
      return ORC_FLOAT64;
    case ORC_M2: case ORC_VARIANCE: case ORC_STD:
      return is_numeric(t) ? ORC_FLOAT64 : -1; /* m2_var_std.cu:27-31 */
    default: return -1;
  }
}

static inline int row_valid(const orc_column* c, int64_t i)
{
  if (!c->null_mask) return 1;
  int64_t b = (int64_t)c->offset + i;
  return (c->null_mask[b >> 5] >> (b & 31)) & 1u;
}

/* canonical 64-bit image of an element, used for key equality + hashing.
 * floats: -0 -> +0 and every NaN -> one NaN (hash_functions.cuh:19-39, equality.cuh nan_equal comparator) */
static inline uint64_t canon_bits(const orc_column* c, int64_t i)
{
  const char* p = (const char*)c->data;
  int64_t e = (int64_t)c->offset + i;
  switch (c->type_id) {
    case ORC_INT8: return (uint64_t)(int64_t)((const int8_t*)p)[e];
    case ORC_INT16: return (uint64_t)(int64_t)((const int16_t*)p)[e];
    case ORC_INT32: case ORC_TIMESTAMP_DAYS: case ORC_DURATION_DAYS: case ORC_DECIMAL32: return (uint64_t)(int64_t)((const int32_t*)p)[e];
    case ORC_UINT8: return ((const uint8_t*)p)[e];
    case ORC_BOOL8: return ((const uint8_t*)p)[e] != 0;
    case ORC_UINT16: return ((const uint16_t*)p)[e];
    case ORC_UINT32: return ((const uint32_t*)p)[e];
    case ORC_FLOAT32: {
      float f = ((const float*)p)[e];
      if (f == 0.0f) f = 0.0f;
      if (f != f) return 0x7fc00000u;
      uint32_t u; memcpy(&u, &f, 4); return u;
    }
    case ORC_FLOAT64: {
      double d = ((const double*)p)[e];
      if (d == 0.0) d = 0.0;
      if (d != d) return 0x7ff8000000000000ull;
      uint64_t u; memcpy(&u, &d, 8); return u;
    }
    default: return ((const uint64_t*)p)[e]; /* all 8-byte integer-like types */
  }
}
static inline int64_t load_i64(const orc_column* c, int64_t i) { return (int64_t)canon_bits(c, i); }
static inline double load_f64(const orc_column* c, int64_t i) { return ((const double*)c->data)[(int64_t)c->offset + i]; }
static inline float load_f32(const orc_column* c, int64_t i) { return ((const float*)c->data)[(int64_t)c->offset + i]; }

static inline uint64_t mix64(uint64_t z)
{
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

/* ---------------------------------------------------------------- partial aggregation state */
typedef struct {
  /* one accumulator record per (aggregation, group); only the fields the kind needs are touched */
  uint64_t i;   /* integer accumulator (two's complement wrap == int64 overflow behaviour of atomic_add) */
  double d;     /* FLOAT64 accumulator */
  float f;      /* FLOAT32 accumulator */
  uint64_t i2;  /* second integer accumulator (sum of squares for M2 family) */
  double d2;
  float f2;
  int32_t n;    /* COUNT_VALID of the value column within the group */
  int32_t nall; /* COUNT_ALL */
  int32_t arg;  /* ARGMIN/ARGMAX row index */
} acc_t;

typedef struct {
  int64_t cap;      /* hash slots (power of two) */
  int32_t* slots;   /* group id or -1 */
  int32_t ngroups;
  int32_t gcap;
  int64_t* rep;     /* representative (first) row of each group */
  uint64_t* hashes; /* cached hash per group */
  acc_t** acc;      /* [n_aggs][gcap] */
} partial_t;

typedef struct {
  const orc_column* keys; int n_keys; int include_null_keys;
  const orc_column* values; const orc_agg* aggs; int n_aggs;
} job_t;

static inline int rows_equal(const job_t* J, int64_t a, int64_t b)
{
  for (int c = 0; c < J->n_keys; ++c) {
    const orc_column* k = &J->keys[c];
    int va = row_valid(k, a), vb = row_valid(k, b);
    if (va != vb) return 0;
    if (va && canon_bits(k, a) != canon_bits(k, b)) return 0; /* both null -> equal (null_equality::EQUAL) */
  }
  return 1;
}
static inline uint64_t row_hash(const job_t* J, int64_t r)
{
  uint64_t h = 0x243F6A8885A308D3ull;
  for (int c = 0; c < J->n_keys; ++c) {
    const orc_column* k = &J->keys[c];
    uint64_t e = row_valid(k, r) ? canon_bits(k, r) : 0xFFFFFFFFFFFFFFFFull ^ (uint64_t)c;
    h = mix64(h ^ (e + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2)));
  }
  return h;
}

static void acc_init(acc_t* a, int kind, int vtype)
{
  memset(a, 0, sizeof *a);
  cls_t cls = type_class(vtype);
  switch (kind) {
    case ORC_MIN: /* identity = numeric_limits<T>::max() / +inf (device_operators.cuh DeviceMin::identity) */
      if (cls == CLS_F64) a->d = INFINITY; else if (cls == CLS_F32) a->f = INFINITY;
      else if (cls == CLS_SINT) a->i = (uint64_t)INT64_MAX; else a->i = UINT64_MAX;
      break;
    case ORC_MAX:
      if (cls == CLS_F64) a->d = -INFINITY; else if (cls == CLS_F32) a->f = -INFINITY;
      else if (cls == CLS_SINT) a->i = (uint64_t)INT64_MIN; else a->i = 0;
      break;
    case ORC_PRODUCT: a->i = 1; a->d = 1.0; a->f = 1.0f; break;
    case ORC_ARGMIN: case ORC_ARGMAX: a->arg = -1; break; /* ARGMIN_SENTINEL / ARGMAX_SENTINEL */
    default: break;
  }
}

/* one row into one accumulator: element_aggregator + update_target_element (device_aggregators.cuh:24-160,309-446) */
static inline void acc_update(acc_t* a, int kind, const orc_column* v, int64_t r)
{
  a->nall++;
  if (!row_valid(v, r)) return; /* null source skipped for everything but COUNT_ALL */
  a->n++;
  cls_t cls = type_class(v->type_id);
  switch (kind) {
    case ORC_SUM: case ORC_MEAN:
      if (cls == CLS_F64) a->d += load_f64(v, r); else if (cls == CLS_F32) a->f += load_f32(v, r);
      else a->i += (uint64_t)load_i64(v, r);
      break;
    case ORC_SUM_OF_SQUARES:
      if (cls == CLS_F64) { double x = load_f64(v, r); a->d += x * x; }
      else if (cls == CLS_F32) { float x = load_f32(v, r); a->f += x * x; }
      else { uint64_t x = (uint64_t)load_i64(v, r); a->i += x * x; }
      break;
    case ORC_M2: case ORC_VARIANCE: case ORC_STD: /* SUM_OF_SQUARES + SUM + COUNT_VALID (extract_single_pass_aggs.cpp:75-109) */
      if (cls == CLS_F64) { double x = load_f64(v, r); a->d += x; a->d2 += x * x; }
      else if (cls == CLS_F32) { float x = load_f32(v, r); a->f += x; a->f2 += x * x; }
      else { uint64_t x = (uint64_t)load_i64(v, r); a->i += x; a->i2 += x * x; }
      break;
    case ORC_SUM_OVERFLOW: case ORC_SUM_OVERFLOW_FLAG: {
      /* The reference adds with atomic_add and raises the flag when an add overflows the Source type, after which the group
       * ignores further rows (device_aggregators.cuh:150-158): which prefix overflows depends on the order the atomics land
       * in. The oracle keeps the EXACT total (i2:i as a 128-bit integer) and reports the order-independent outcome: the flag
       * is set iff the total does not fit the Source type -- every order overflows then, and when it fits some order of the
       * atomics sees no overflow at all. The sum child of an overflowed group is unspecified in the reference. */
      __int128 tot = (__int128)(((unsigned __int128)a->i2 << 64) | a->i);
      tot += (__int128)load_i64(v, r);
      a->i = (uint64_t)tot; a->i2 = (uint64_t)((unsigned __int128)tot >> 64);
      break;
    }
    case ORC_PRODUCT:
      if (cls == CLS_F64) a->d *= load_f64(v, r); else if (cls == CLS_F32) a->f *= load_f32(v, r);
      else a->i *= (uint64_t)load_i64(v, r);
      break;
    case ORC_MIN:
      if (cls == CLS_F64) { double x = load_f64(v, r); if (x < a->d) a->d = x; }
      else if (cls == CLS_F32) { float x = load_f32(v, r); if (x < a->f) a->f = x; }
      else if (cls == CLS_SINT) { int64_t x = load_i64(v, r); if (x < (int64_t)a->i) a->i = (uint64_t)x; }
      else { uint64_t x = canon_bits(v, r); if (x < a->i) a->i = x; }
      break;
    case ORC_MAX:
      if (cls == CLS_F64) { double x = load_f64(v, r); if (x > a->d) a->d = x; }
      else if (cls == CLS_F32) { float x = load_f32(v, r); if (x > a->f) a->f = x; }
      else if (cls == CLS_SINT) { int64_t x = load_i64(v, r); if (x > (int64_t)a->i) a->i = (uint64_t)x; }
      else { uint64_t x = canon_bits(v, r); if (x > a->i) a->i = x; }
      break;
    case ORC_ARGMIN: case ORC_ARGMAX: {
      /* first row wins among ties (the reference's CAS loop keeps the earlier winner on equality) */
      if (a->arg < 0) { a->arg = (int32_t)r; break; }
      int better;
      int64_t o = a->arg;
      if (cls == CLS_F64) better = kind == ORC_ARGMIN ? load_f64(v, r) < load_f64(v, o) : load_f64(v, r) > load_f64(v, o);
      else if (cls == CLS_F32) better = kind == ORC_ARGMIN ? load_f32(v, r) < load_f32(v, o) : load_f32(v, r) > load_f32(v, o);
      else if (cls == CLS_SINT) better = kind == ORC_ARGMIN ? load_i64(v, r) < load_i64(v, o) : load_i64(v, r) > load_i64(v, o);
      else better = kind == ORC_ARGMIN ? canon_bits(v, r) < canon_bits(v, o) : canon_bits(v, r) > canon_bits(v, o);
      if (better) a->arg = (int32_t)r;
      break;
    }
    default: break; /* COUNT_VALID / COUNT_ALL handled by n / nall */
  }
}

/* merge b into a (partials of the same aggregation over disjoint row sets) */
static void acc_merge(acc_t* a, const acc_t* b, int kind, const orc_column* v)
{
  cls_t cls = type_class(v->type_id);
  switch (kind) {
    case ORC_SUM_OVERFLOW: case ORC_SUM_OVERFLOW_FLAG: {
      __int128 ta = (__int128)(((unsigned __int128)a->i2 << 64) | a->i), tb = (__int128)(((unsigned __int128)b->i2 << 64) | b->i);
      ta += tb;
      a->i = (uint64_t)ta; a->i2 = (uint64_t)((unsigned __int128)ta >> 64);
      break;
    }
    case ORC_PRODUCT: a->i *= b->i; a->d *= b->d; a->f *= b->f; break;
    case ORC_MIN:
      if (cls == CLS_F64) { if (b->d < a->d) a->d = b->d; } else if (cls == CLS_F32) { if (b->f < a->f) a->f = b->f; }
      else if (cls == CLS_SINT) { if ((int64_t)b->i < (int64_t)a->i) a->i = b->i; } else { if (b->i < a->i) a->i = b->i; }
      break;
    case ORC_MAX:
      if (cls == CLS_F64) { if (b->d > a->d) a->d = b->d; } else if (cls == CLS_F32) { if (b->f > a->f) a->f = b->f; }
      else if (cls == CLS_SINT) { if ((int64_t)b->i > (int64_t)a->i) a->i = b->i; } else { if (b->i > a->i) a->i = b->i; }
      break;
    case ORC_ARGMIN: case ORC_ARGMAX:
      if (b->arg >= 0) {
        if (a->arg < 0) a->arg = b->arg;
        else {
          acc_t t = *a; t.n = 0; t.nall = 0; /* re-run the comparison, earlier row kept on ties */
          int64_t lo = a->arg < b->arg ? a->arg : b->arg, hi = a->arg < b->arg ? b->arg : a->arg;
          t.arg = (int32_t)lo; acc_update(&t, kind, v, hi); a->arg = t.arg;
        }
      }
      break;
    default: a->i += b->i; a->d += b->d; a->f += b->f; a->i2 += b->i2; a->d2 += b->d2; a->f2 += b->f2; break;
  }
  a->n += b->n; a->nall += b->nall;
}

static int partial_init(partial_t* P, const job_t* J, int64_t expect_rows)
{
  int64_t cap = 1024;
  while (cap < 2 * expect_rows && cap < (1ll << 22)) cap <<= 1;
  P->cap = cap; P->ngroups = 0; P->gcap = 1024;
  P->slots = (int32_t*)malloc(sizeof(int32_t) * cap);
  P->rep = (int64_t*)malloc(sizeof(int64_t) * P->gcap);
  P->hashes = (uint64_t*)malloc(sizeof(uint64_t) * P->gcap);
  P->acc = (acc_t**)calloc(J->n_aggs > 0 ? J->n_aggs : 1, sizeof(acc_t*));
  if (!P->slots || !P->rep || !P->hashes || !P->acc) return 1;
  memset(P->slots, 0xff, sizeof(int32_t) * cap);
  for (int a = 0; a < J->n_aggs; ++a) { P->acc[a] = (acc_t*)malloc(sizeof(acc_t) * P->gcap); if (!P->acc[a]) return 1; }
  return 0;
}
static void partial_free(partial_t* P, const job_t* J)
{
  free(P->slots); free(P->rep); free(P->hashes);
  if (P->acc) { for (int a = 0; a < J->n_aggs; ++a) free(P->acc[a]); free(P->acc); }
  memset(P, 0, sizeof *P);
}
static int partial_grow_groups(partial_t* P, const job_t* J)
{
  int32_t ng = P->gcap * 2;
  P->rep = (int64_t*)realloc(P->rep, sizeof(int64_t) * ng);
  P->hashes = (uint64_t*)realloc(P->hashes, sizeof(uint64_t) * ng);
  if (!P->rep || !P->hashes) return 1;
  for (int a = 0; a < J->n_aggs; ++a) { P->acc[a] = (acc_t*)realloc(P->acc[a], sizeof(acc_t) * ng); if (!P->acc[a]) return 1; }
  P->gcap = ng;
  return 0;
}
static int partial_rehash(partial_t* P)
{
  int64_t cap = P->cap * 2;
  int32_t* s = (int32_t*)malloc(sizeof(int32_t) * cap);
  if (!s) return 1;
  memset(s, 0xff, sizeof(int32_t) * cap);
  for (int32_t g = 0; g < P->ngroups; ++g) {
    int64_t p = (int64_t)(P->hashes[g] & (uint64_t)(cap - 1));
    while (s[p] >= 0) p = (p + 1) & (cap - 1);
    s[p] = g;
  }
  free(P->slots); P->slots = s; P->cap = cap;
  return 0;
}
/* find-or-insert the group of row r (hash h); returns group id or -1 on allocation failure */
static inline int32_t partial_find_or_insert(partial_t* P, const job_t* J, int64_t r, uint64_t h)
{
  int64_t mask = P->cap - 1, p = (int64_t)(h & (uint64_t)mask);
  for (;;) {
    int32_t g = P->slots[p];
    if (g < 0) break;
    if (P->hashes[g] == h && rows_equal(J, P->rep[g], r)) return g;
    p = (p + 1) & mask;
  }
  if (P->ngroups == P->gcap && partial_grow_groups(P, J)) return -1;
  int32_t g = P->ngroups++;
  P->slots[p] = g; P->rep[g] = r; P->hashes[g] = h;
  for (int a = 0; a < J->n_aggs; ++a) acc_init(&P->acc[a][g], J->aggs[a].kind, J->values[J->aggs[a].value_index].type_id);
  if ((int64_t)P->ngroups * 2 > P->cap && partial_rehash(P)) return -1;
  return g;
}

static int partial_run(partial_t* P, const job_t* J, int64_t lo, int64_t hi)
{
  for (int64_t r = lo; r < hi; ++r) {
    if (!J->include_null_keys) { /* null_policy::EXCLUDE: rows with any null key are dropped (groupby.cu:41-43) */
      int skip = 0;
      for (int c = 0; c < J->n_keys; ++c) if (!row_valid(&J->keys[c], r)) { skip = 1; break; }
      if (skip) continue;
    }
    int32_t g = partial_find_or_insert(P, J, r, row_hash(J, r));
    if (g < 0) return 1;
    for (int a = 0; a < J->n_aggs; ++a) acc_update(&P->acc[a][g], J->aggs[a].kind, &J->values[J->aggs[a].value_index], r);
  }
  return 0;
}

/* ---------------------------------------------------------------- output */
static int key_less(const job_t* J, int64_t a, int64_t b)
{
  for (int c = 0; c < J->n_keys; ++c) {
    const orc_column* k = &J->keys[c];
    int va = row_valid(k, a), vb = row_valid(k, b);
    if (va != vb) return !va; /* nulls first (null_order::BEFORE) */
    if (!va) continue;
    switch (type_class(k->type_id)) {
      case CLS_SINT: { int64_t x = load_i64(k, a), y = load_i64(k, b); if (x != y) return x < y; break; }
      case CLS_F32: case CLS_F64: {
        double x = k->type_id == ORC_FLOAT32 ? (double)load_f32(k, a) : load_f64(k, a);
        double y = k->type_id == ORC_FLOAT32 ? (double)load_f32(k, b) : load_f64(k, b);
        int nx = x != x, ny = y != y;
        if (nx != ny) return ny; /* NaN last */
        if (!nx && x != y) return x < y;
        break;
      }
      default: { uint64_t x = canon_bits(k, a), y = canon_bits(k, b); if (x != y) return x < y; break; }
    }
  }
  return 0;
}
static const job_t* g_sort_job;
static const int64_t* g_sort_rep;
static int cmp_groups(const void* pa, const void* pb)
{
  int32_t a = *(const int32_t*)pa, b = *(const int32_t*)pb;
  if (key_less(g_sort_job, g_sort_rep[a], g_sort_rep[b])) return -1;
  if (key_less(g_sort_job, g_sort_rep[b], g_sort_rep[a])) return 1;
  return 0;
}

static void store_typed(void* data, int type_id, int64_t idx, uint64_t i, double d, float f)
{
  switch (orc_type_size(type_id)) {
    case 1: ((uint8_t*)data)[idx] = (uint8_t)i; break;
    case 2: ((uint16_t*)data)[idx] = (uint16_t)i; break;
    case 4: if (type_id == ORC_FLOAT32) ((float*)data)[idx] = f; else ((uint32_t*)data)[idx] = (uint32_t)i; break;
    default: if (type_id == ORC_FLOAT64) ((double*)data)[idx] = d; else ((uint64_t*)data)[idx] = i; break;
  }
}

void orc_free_result(orc_result* r)
{
  if (!r) return;
  for (int i = 0; r->keys && i < r->n_keys; ++i) { free(r->keys[i].data); free(r->keys[i].valid); }
  for (int i = 0; r->aggs && i < r->n_aggs; ++i) { free(r->aggs[i].data); free(r->aggs[i].valid); }
  free(r->keys); free(r->aggs);
  memset(r, 0, sizeof *r);
}

int orc_groupby(const orc_column* keys, int n_keys, int include_null_keys, const orc_column* values, int n_values,
                const orc_agg* aggs, int n_aggs, int n_threads, int sort_output, orc_result* out)
{
  memset(out, 0, sizeof *out);
  if (n_keys <= 0) FAIL(1, "groupby needs at least one key column");
  int64_t n = keys[0].size;
  for (int c = 0; c < n_keys; ++c) {
    if (keys[c].size != n) FAIL(1, "key columns differ in size");
    if (type_class(keys[c].type_id) == CLS_BAD) FAIL(2, "unsupported key type %d", keys[c].type_id);
  }
  for (int a = 0; a < n_aggs; ++a) {
    if (aggs[a].value_index < 0 || aggs[a].value_index >= n_values) FAIL(1, "bad value index");
    const orc_column* v = &values[aggs[a].value_index];
    /* groupby.cu:225-229 */
    if (v->size != n) FAIL(1, "Size mismatch between request values and groupby keys.");
    /* groupby.cu:186-201 */
    if (orc_target_type(v->type_id, aggs[a].kind) < 0) FAIL(1, "Invalid type/aggregation combination.");
  }
  job_t J = {keys, n_keys, include_null_keys, values, aggs, n_aggs};

  int T = n_threads < 1 ? 1 : n_threads;
  if (n < 65536) T = 1;
  partial_t* parts = (partial_t*)calloc(T, sizeof(partial_t));
  int err = 0;
#ifdef _OPENMP
#pragma omp parallel for num_threads(T) schedule(static, 1) reduction(| : err)
#endif
  for (int t = 0; t < T; ++t) {
    int64_t lo = n * t / T, hi = n * (t + 1) / T;
    if (partial_init(&parts[t], &J, 4096)) { err |= 1; continue; }
    err |= partial_run(&parts[t], &J, lo, hi);
  }
  if (err) { for (int t = 0; t < T; ++t) partial_free(&parts[t], &J); free(parts); FAIL(1, "out of memory"); }
  /* merge partials into parts[0] in thread order (keeps first-occurrence representatives) */
  partial_t* P = &parts[0];
  for (int t = 1; t < T; ++t) {
    partial_t* Q = &parts[t];
    for (int32_t g = 0; g < Q->ngroups; ++g) {
      int32_t before = P->ngroups;
      int32_t d = partial_find_or_insert(P, &J, Q->rep[g], Q->hashes[g]);
      if (d < 0) { err = 1; break; }
      for (int a = 0; a < n_aggs; ++a) {
        if (d >= before) P->acc[a][d] = Q->acc[a][g];
        else acc_merge(&P->acc[a][d], &Q->acc[a][g], aggs[a].kind, &values[aggs[a].value_index]);
      }
    }
    partial_free(Q, &J);
  }
  if (err) { partial_free(P, &J); free(parts); FAIL(1, "out of memory"); }

  int32_t G = P->ngroups;
  int32_t* order = (int32_t*)malloc(sizeof(int32_t) * (G > 0 ? G : 1));
  for (int32_t g = 0; g < G; ++g) order[g] = g;
  if (sort_output && G > 1) { g_sort_job = &J; g_sort_rep = P->rep; qsort(order, G, sizeof(int32_t), cmp_groups); }

  out->num_groups = G; out->n_keys = n_keys; out->n_aggs = n_aggs;
  out->keys = (orc_out_column*)calloc(n_keys, sizeof(orc_out_column));
  out->aggs = (orc_out_column*)calloc(n_aggs > 0 ? n_aggs : 1, sizeof(orc_out_column));
  size_t Gs = G > 0 ? (size_t)G : 1;
  for (int c = 0; c < n_keys; ++c) {
    orc_out_column* o = &out->keys[c];
    o->type_id = keys[c].type_id; o->has_mask = include_null_keys && keys[c].null_mask != NULL;
    o->data = calloc(Gs, 8); o->valid = (uint8_t*)calloc(Gs, 1);
    int sz = orc_type_size(keys[c].type_id);
    for (int32_t j = 0; j < G; ++j) {
      int64_t r = P->rep[order[j]];
      o->valid[j] = (uint8_t)row_valid(&keys[c], r);
      memcpy((char*)o->data + (size_t)j * sz, (const char*)keys[c].data + ((int64_t)keys[c].offset + r) * sz, sz);
    }
  }
  for (int a = 0; a < n_aggs; ++a) {
    const orc_column* v = &values[aggs[a].value_index];
    int kind = aggs[a].kind, tt = orc_target_type(v->type_id, kind);
    cls_t cls = type_class(v->type_id);
    orc_out_column* o = &out->aggs[a];
    o->type_id = tt; o->data = calloc(Gs, 8); o->valid = (uint8_t*)calloc(Gs, 1);
    int has_nulls = v->null_count > 0; /* column_view::has_nulls() */
    /* output_utils.cu:61-64: nullable iff not COUNT and input has nulls (user-requested aggregation) */
    o->has_mask = (kind != ORC_COUNT_VALID && kind != ORC_COUNT_ALL && kind != ORC_M2 && has_nulls);
    for (int32_t j = 0; j < G; ++j) {
      const acc_t* A = &P->acc[a][order[j]];
      uint8_t valid = 1; uint64_t ri = 0; double rd = 0; float rf = 0;
      switch (kind) {
        case ORC_COUNT_VALID: ri = (uint64_t)(int64_t)A->n; break;
        case ORC_COUNT_ALL: ri = (uint64_t)(int64_t)A->nall; break;
        case ORC_SUM: case ORC_PRODUCT: case ORC_SUM_OF_SQUARES: case ORC_MIN: case ORC_MAX:
          ri = A->i; rd = A->d; rf = A->f;
          /* target starts null and becomes valid on the first valid contribution (device_aggregators.cuh:440-442);
           * when the input has no nulls there is no mask at all */
          valid = has_nulls ? (A->n > 0) : 1;
          break;
        case ORC_ARGMIN: case ORC_ARGMAX: ri = (uint64_t)(int64_t)A->arg; valid = has_nulls ? (A->n > 0) : 1; break;
        case ORC_SUM_OVERFLOW: case ORC_SUM_OVERFLOW_FLAG: {
          __int128 tot = (__int128)(((unsigned __int128)A->i2 << 64) | A->i);
          int bits = 8 * orc_type_size(v->type_id);
          __int128 lo = -((__int128)1 << (bits - 1)), hi = ((__int128)1 << (bits - 1)) - 1;
          if (kind == ORC_SUM_OVERFLOW) ri = (uint64_t)tot; /* low bits; store_typed keeps the Source width */
          else ri = (tot < lo || tot > hi) ? 1 : 0;
          valid = has_nulls ? (A->n > 0) : 1;
          break;
        }
        case ORC_MEAN: {
          /* hash_compound_agg_finalizer.cu:94-133: double(sum)/double(count), null where count == 0 */
          if (tt == ORC_FLOAT64) {
            double s = cls == CLS_F64 ? A->d : cls == CLS_F32 ? (double)A->f : (double)(int64_t)A->i;
            if (cls == CLS_UINT || cls == CLS_BOOL) s = (double)(int64_t)A->i; /* SUM target is int64 for every integral source */
            rd = s / (double)A->n;
          } else { /* duration / decimal MEAN keeps the source rep: integer division of the rep */
            ri = A->n ? (uint64_t)((int64_t)A->i / (int64_t)A->n) : 0;
          }
          valid = A->n > 0;
          if (!valid) { rd = 0; ri = 0; }
          break;
        }
        case ORC_M2: case ORC_VARIANCE: case ORC_STD: {
          /* m2_var_std.cu:49-59: m2 = sumsq - sum*sum/n (0 when n == 0), all in double */
          double ss = cls == CLS_F64 ? A->d2 : cls == CLS_F32 ? (double)A->f2 : (double)(int64_t)A->i2;
          double s = cls == CLS_F64 ? A->d : cls == CLS_F32 ? (double)A->f : (double)(int64_t)A->i;
          double m2 = A->n == 0 ? 0.0 : ss - s * s / A->n;
          if (kind == ORC_M2) { rd = m2; valid = 1; break; }
          /* m2_var_std.cu:160-187: null iff n == 0 or n - ddof <= 0 */
          int32_t df = A->n - aggs[a].ddof;
          if (A->n == 0 || df <= 0) { valid = 0; rd = 0; }
          else { rd = m2 / df; if (kind == ORC_STD) rd = sqrt(rd); }
          if (kind != ORC_M2) o->has_mask = 1; /* mask attached whenever a null exists; semantic nulls are what tests compare */
          break;
        }
        default: break;
      }
      o->valid[j] = valid;
      store_typed(o->data, tt, j, ri, rd, rf);
    }
  }
  free(order);
  partial_free(P, &J);
  free(parts);
  return 0;
}

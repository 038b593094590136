/*
 * TEST INFRASTRUCTURE ONLY -- CPU restatement ("oracle") of libcudf's hash groupby-aggregate semantics.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this.
 * The product path (cudf_b200/, include/bgb.h) never links, imports or calls anything under oracle/.
 *
 * Parity status: PINNED against the reference's own known-answer tests (the .cpp files under cpp/tests/groupby),
 * transcribed into tests/golden/reference_kat.json (see tests/test_oracle_golden.py).
 * libcudf itself cannot be built here (rapids-cmake/rmm/cuco/CCCL are network-fetched), so there is no oracle/_ref.
 *
 * Reference files this restates (relative to /root/reference):
 *   type rules        cpp/include/cudf/detail/aggregation/aggregation.hpp:878-1012
 *   null/validity     cpp/include/cudf/detail/aggregation/device_aggregators.cuh:428-446,
 *                     cpp/src/groupby/hash/output_utils.cu:61-87
 *   decomposition     cpp/src/groupby/hash/extract_single_pass_aggs.cpp:38-109
 *   MEAN              cpp/src/groupby/hash/hash_compound_agg_finalizer.cu:94-133
 *   M2/VAR/STD        cpp/src/groupby/common/m2_var_std.cu:41-59,152-188
 *   null-key policy   cpp/src/groupby/common/utils.cpp:13-32, cpp/src/groupby/hash/groupby.cu:41-43
 *   key equality      cpp/include/cudf/detail/row_operator/equality.cuh:59-89 (nulls equal, NaN==NaN, -0==+0)
 *   empty / no-request cpp/src/groupby/groupby.cu:233, cpp/src/groupby/hash/compute_groupby.cu:113-130
 */
#ifndef GROUPBY_ORACLE_H
#define GROUPBY_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* cudf::type_id values (cpp/include/cudf/types.hpp:185-217) */
enum {
  ORC_EMPTY = 0, ORC_INT8, ORC_INT16, ORC_INT32, ORC_INT64, ORC_UINT8, ORC_UINT16, ORC_UINT32, ORC_UINT64,
  ORC_FLOAT32, ORC_FLOAT64, ORC_BOOL8, ORC_TIMESTAMP_DAYS, ORC_TIMESTAMP_SECONDS, ORC_TIMESTAMP_MILLISECONDS,
  ORC_TIMESTAMP_MICROSECONDS, ORC_TIMESTAMP_NANOSECONDS, ORC_DURATION_DAYS, ORC_DURATION_SECONDS,
  ORC_DURATION_MILLISECONDS, ORC_DURATION_MICROSECONDS, ORC_DURATION_NANOSECONDS, ORC_DICTIONARY32, ORC_STRING,
  ORC_LIST, ORC_DECIMAL32, ORC_DECIMAL64, ORC_DECIMAL128, ORC_STRUCT
};

/* cudf::aggregation::Kind values (cpp/include/cudf/aggregation.hpp:78-121) */
enum {
  ORC_SUM = 0, ORC_SUM_OVERFLOW = 1, ORC_PRODUCT = 2, ORC_MIN = 3, ORC_MAX = 4, ORC_COUNT_VALID = 5, ORC_COUNT_ALL = 6,
  ORC_SUM_OF_SQUARES = 9, ORC_MEAN = 10, ORC_M2 = 11, ORC_VARIANCE = 12, ORC_STD = 13, ORC_ARGMAX = 16, ORC_ARGMIN = 17,
  /* oracle-internal: the BOOL8 `overflow` child of a SUM_OVERFLOW result (ORC_SUM_OVERFLOW itself yields the `sum` child,
   * typed like the values; the struct's validity is the validity of either) */
  ORC_SUM_OVERFLOW_FLAG = 101
};

/* host-memory column: same field meaning as cudf::column_view (column_view.hpp:236-243) */
typedef struct {
  int32_t type_id;
  int32_t size;
  const void* data;          /* element i at data + (offset+i)*sizeof(T) */
  const uint32_t* null_mask; /* bit (offset+i) of 32-bit words, 1 = valid; NULL = all valid */
  int32_t null_count;
  int32_t offset;
} orc_column;

typedef struct {
  int32_t value_index; /* which values column */
  int32_t kind;        /* ORC_* aggregation kind */
  int32_t ddof;        /* VARIANCE / STD only */
} orc_agg;

typedef struct {
  int32_t type_id;
  int32_t has_mask;  /* 1 when the reference would attach a null mask */
  void* data;        /* num_groups elements of type_id */
  uint8_t* valid;    /* num_groups bytes, 1 = valid (always allocated) */
} orc_out_column;

typedef struct {
  int32_t num_groups;
  int32_t n_keys;
  int32_t n_aggs;
  orc_out_column* keys; /* n_keys, rows sorted by key (nulls first, then ascending, NaN last) */
  orc_out_column* aggs; /* n_aggs, same row order */
} orc_result;

/* returns 0 on success, 1 invalid argument / invalid type-aggregation combination, 2 unsupported */
int orc_groupby(const orc_column* keys, int n_keys, int include_null_keys, const orc_column* values, int n_values,
                const orc_agg* aggs, int n_aggs, int n_threads, int sort_output, orc_result* out);
void orc_free_result(orc_result* r);
const char* orc_last_error(void);
int orc_type_size(int type_id);
/* result type of (value type, aggregation kind), or -1 when invalid */
int orc_target_type(int value_type_id, int kind);

/* counter-based SplitMix64 generator shared with the GPU generators */
uint64_t orc_splitmix64(uint64_t seed, uint64_t row);

#ifdef __cplusplus
}
#endif
#endif

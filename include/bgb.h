/*
 * bgb.h -- C ABI of the B200-native hash groupby-aggregate ("bgb" = B200 group-by).
 *
 * This is the drop-in boundary for libcudf's hash groupby path. The reference has no C ABI on this path
 * (pylibcudf binds the C++ class directly); these entry points are what a cgo / JNI / ctypes / Cython
 * binding would call in place of
 *     cudf::groupby::detail::hash::groupby(keys, requests, null_policy, stream, mr)
 *                                              /root/reference/cpp/src/groupby/hash/groupby.cu:134-147
 *     cudf::groupby::detail::hash::can_use_hash_groupby(requests)
 *                                              /root/reference/cpp/src/groupby/hash/groupby.cu:115-131
 * declared in /root/reference/cpp/include/cudf/detail/groupby.hpp:18-37 and reached from
 * cudf::groupby::groupby::aggregate (/root/reference/cpp/src/groupby/groupby.cu:219-236).
 * The C++ class with the reference's exact signatures sits on top of this ABI in include/cudf/groupby.hpp.
 *
 * Rules of the boundary
 *  - plain pointers and sizes only; every pointer in a bgb_column_view / bgb_mutable_column is DEVICE memory
 *    (except where a function name says _host);
 *  - stream-ordered on the caller's CUDA stream (passed as void* = cudaStream_t); never touches the default stream
 *    unless the caller passes it; the only host waits are on that stream (group count read-back);
 *  - output memory is allocated by the CALLER (two-phase build -> emit), scratch comes from the caller's
 *    bgb_allocator (or cudaMallocAsync on the stream when NULL); the library never frees caller memory;
 *  - no exception crosses the boundary: status codes + bgb_last_error() (thread-local message);
 *  - re-entrant: no global mutable state besides an immutable device-property cache.
 */
#ifndef BGB_H
#define BGB_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BGB_VERSION 100

typedef enum bgb_status {
  BGB_OK = 0,
  BGB_INVALID_ARGUMENT = 1, /* what the reference reports as cudf::logic_error (groupby.cu:186-229) */
  BGB_UNSUPPORTED = 2,      /* can_use_hash_groupby() == false: caller takes its sort fallback */
  BGB_CUDA_ERROR = 3,       /* cudf::cuda_error */
  BGB_NCCL_ERROR = 4,
  BGB_OUT_OF_MEMORY = 5,
  BGB_PEER_OVERFLOW = 6,    /* bgb_groupby_build_peer: a receive segment was too small; every rank reports it, the caller
                               re-creates the peer context with a larger region (or uses the NCCL exchange) */
  BGB_PEER_ABORTED = 7      /* bgb_groupby_build_peer: ANOTHER rank failed in this step and said so through the exchange header;
                               every other rank returns this in the same step (no rank is left waiting for the time-out) and
                               the context stays usable: all ranks have advanced to the same epoch */
} bgb_status;

/* cudf::type_id, same numeric values (/root/reference/cpp/include/cudf/types.hpp:185-217) */
typedef enum bgb_type_id {
  BGB_EMPTY = 0, BGB_INT8 = 1, BGB_INT16 = 2, BGB_INT32 = 3, BGB_INT64 = 4, BGB_UINT8 = 5, BGB_UINT16 = 6,
  BGB_UINT32 = 7, BGB_UINT64 = 8, BGB_FLOAT32 = 9, BGB_FLOAT64 = 10, BGB_BOOL8 = 11, BGB_TIMESTAMP_DAYS = 12,
  BGB_TIMESTAMP_SECONDS = 13, BGB_TIMESTAMP_MILLISECONDS = 14, BGB_TIMESTAMP_MICROSECONDS = 15,
  BGB_TIMESTAMP_NANOSECONDS = 16, BGB_DURATION_DAYS = 17, BGB_DURATION_SECONDS = 18, BGB_DURATION_MILLISECONDS = 19,
  BGB_DURATION_MICROSECONDS = 20, BGB_DURATION_NANOSECONDS = 21, BGB_DICTIONARY32 = 22, BGB_STRING = 23, BGB_LIST = 24,
  BGB_DECIMAL32 = 25, BGB_DECIMAL64 = 26, BGB_DECIMAL128 = 27, BGB_STRUCT = 28
} bgb_type_id;

/* cudf::aggregation::Kind, same numeric values (/root/reference/cpp/include/cudf/aggregation.hpp:78-121) */
typedef enum bgb_agg_kind {
  BGB_SUM = 0, BGB_SUM_OVERFLOW = 1, BGB_PRODUCT = 2, BGB_MIN = 3, BGB_MAX = 4, BGB_COUNT_VALID = 5, BGB_COUNT_ALL = 6,
  BGB_ANY = 7, BGB_ALL = 8, BGB_SUM_OF_SQUARES = 9, BGB_MEAN = 10, BGB_M2 = 11, BGB_VARIANCE = 12, BGB_STD = 13,
  BGB_MEDIAN = 14, BGB_QUANTILE = 15, BGB_ARGMAX = 16, BGB_ARGMIN = 17, BGB_NUNIQUE = 18, BGB_NTH_ELEMENT = 19,
  /* library extension used when MERGING partial results (streaming_groupby::merge / distributed reduce step): the sum
   * of INT32 partial counts, returned as INT32 like a COUNT (cpp/src/groupby/streaming_groupby/merge.cu:30-144) */
  BGB_MERGE_COUNT = 100
} bgb_agg_kind;

/* POD mirror of cudf::column_view's fields (/root/reference/cpp/include/cudf/column/column_view.hpp:236-243):
 * element i lives at data + (offset + i) * sizeof(T); validity bit (offset + i) of the 32-bit mask words, 1 = valid
 * (/root/reference/cpp/include/cudf/utilities/bit.hpp:47-60). null_mask == NULL means "no nulls". */
typedef struct bgb_column_view {
  int32_t type_id;    /* bgb_type_id */
  int32_t scale;      /* fixed_point scale, carried through untouched */
  int32_t size;       /* rows (cudf::size_type = int32) */
  int32_t null_count; /* as column_view::null_count(); has_nulls() == null_count > 0 */
  int32_t offset;
  int32_t reserved;
  const void* data;   /* device pointer to the elements. DICTIONARY32 (value columns only): HOST pointer to two child views
                         {indices, keys} -- cudf::dictionary_column_view::indices() / keys(), dictionary_column_view.hpp; size, offset
                         and null_mask of THIS view apply to the rows, the element of row r is keys[indices[offset + r]] and every
                         aggregation is typed by the keys (hash/groupby.cu:118-120, hash/output_utils.cu:66) */
  const uint32_t* null_mask;
} bgb_column_view;

/* one cudf::groupby_aggregation (kind + the ddof of VARIANCE/STD, cudf/aggregation.hpp:293,305) */
typedef struct bgb_aggregation {
  int32_t kind; /* bgb_agg_kind */
  int32_t ddof;
} bgb_aggregation;

/* cudf::groupby::aggregation_request (/root/reference/cpp/include/cudf/groupby.hpp:54-57) */
typedef struct bgb_request {
  bgb_column_view values;
  const bgb_aggregation* aggregations;
  int32_t num_aggregations;
  int32_t reserved;
} bgb_request;

/* A result of type BGB_STRUCT (SUM_OVERFLOW: struct{sum, overflow}, cudf/aggregation.hpp make_sum_overflow_aggregation,
 * /root/reference/cpp/src/groupby/hash/output_utils.cu:83-104) is passed to bgb_groupby_emit as a bgb_mutable_column whose `data`
 * points to an array of TWO caller-owned bgb_mutable_column children in HOST memory -- [0] the sum, typed like the values
 * column, [1] the overflow flag, BGB_BOOL8 -- and whose null_mask is the struct's mask (children carry none).
 * The overflow flag is set iff the EXACT total of the group does not fit the values type: the reference raises it when one
 * of its atomic adds overflows (device_aggregators.cuh:150-158), which depends on the order the adds land in whenever only a
 * prefix overflows; this library reports the outcome every order agrees on. The sum of an overflowed group is the wrapped
 * total (unspecified in the reference). */
/* caller-owned output column: data/null_mask sized for `size` rows (mask words = (size+31)/32, zero-filled not required).
 * null_mask may be NULL when bgb_groupby_result_info said nullable == 0. null_count is written by emit. */
typedef struct bgb_mutable_column {
  int32_t type_id;
  int32_t size;
  void* data;
  uint32_t* null_mask;
  int32_t null_count; /* out */
  int32_t reserved;
} bgb_mutable_column;

/* stream-ordered scratch allocator = the role of rmm::device_async_resource_ref on the reference path.
 * NULL allocator => cudaMallocAsync / cudaFreeAsync on the call's stream. */
typedef struct bgb_allocator {
  void* (*allocate)(size_t bytes, void* stream, void* ctx);
  void (*deallocate)(void* ptr, size_t bytes, void* stream, void* ctx);
  void* ctx;
} bgb_allocator;

/* tunables; zero-initialise for defaults. Mirrors the reference's compile-time constants
 * (/root/reference/cpp/src/groupby/hash/helpers.cuh:19-40) but selectable for tests. */
typedef struct bgb_options {
  int32_t force_path;        /* 0 auto, 1 global-table kernel only, 2 shared-memory hashed pre-aggregation, 3 partitioned,
                                4 shared-memory (direct-address variant allowed), 5 tiny-cardinality kernel (shared-memory
                                accumulators), 6 tiny-cardinality kernel with register accumulators (<= 8 groups), 7 bucketed (radix scatter
                                into fixed-capacity buckets + one shared-memory table slice per bucket), 8 tile-sorting kernel (one-word key image:
                                dense integer range of a few thousand ids, else a per-CTA hash table; any accumulators but ARGMIN/ARGMAX) */
  int32_t capacity_log2;     /* 0 auto; else global table has 2^capacity_log2 slots */
  int32_t sample_rows;       /* 0 auto (65536) rows sampled by the cardinality estimator */
  int32_t profile;           /* != 0: bracket every kernel with CUDA events on the call's stream (bgb_groupby_profile) */
  int32_t bucket_slice_log2; /* 0 auto; else the bucketed path uses table slices of 2^bucket_slice_log2 slots */
  int32_t reserved[3];
} bgb_options;

typedef struct bgb_groupby bgb_groupby; /* opaque handle between build and emit */

/* ---- tuning ------------------------------------------------------------------------------------------------------ */

/* The cardinality estimate of bgb_groupby_build (a 65 536-row sample, read back on the host) only chooses the table size and the
 * kernel family -- no result depends on it. With speculation ON (the default; BGB_NO_SPECULATE=1 in the environment turns it off) a
 * call whose SHAPE (row count, key column addresses, column types and masks, aggregations, options, device) was seen before --
 * the same table queried again, or the next batch landing in the same buffers -- starts on the remembered
 * estimate while a fresh sample runs beside the kernels, so the call synchronises with the host ONCE (the group count) instead of
 * twice; a fresh sample that disagrees with the remembered one drops it, and the next call of that shape estimates synchronously.
 * The price: which kernel family a call takes can depend on the calls before it. Returns the previous setting; a call also forgets
 * every remembered estimate. No counterpart in the reference (its table is always 2 * num_rows slots). */
int32_t bgb_set_estimate_speculation(int32_t enabled);

/* ---- capability + type rules ---------------------------------------------------------------------------- */

/* mirrors can_use_hash_groupby (hash/groupby.cu:115-131) restricted to what this library implements:
 * *out = 1 when every (value type, aggregation) pair is served by the hash path, else 0 (caller uses sort path). */
int bgb_can_use_hash_groupby(const bgb_column_view* keys, int32_t num_keys, const bgb_request* requests,
                             int32_t num_requests, int32_t* out);

/* cudf::detail::target_type(values_type, kind) (detail/aggregation/aggregation.hpp:878-1012); -1 when invalid */
int32_t bgb_target_type(int32_t values_type_id, int32_t kind);

/* ---- the groupby call, two-phase --------------------------------------------------------------------------- */

/* Phase 1: hash rows, aggregate into the group table, compact, count groups.
 * Waits on `stream` once to read *num_groups back (the reference also learns the group count on the host,
 * hash/output_utils.cu:145-147). keys/requests must stay alive until emit.
 * Keys: up to 8 fixed-width columns of any total width. Images of at most 128 bits are packed (and range-compressed into one
 * word when the exact column ranges allow it); wider ones keep the ROW NUMBER of each group's first row in the table and
 * compare rows in the caller's key columns -- the reference's scheme for every key (hash/compute_groupby.cu:79-104) -- on the
 * generic kernels only (path "global+rowkeys"; force_path other than 0 / 1 is then BGB_UNSUPPORTED; emit gathers the output keys
 * from the key columns, which is why they must still be alive). */
int bgb_groupby_build(const bgb_column_view* keys, int32_t num_keys, int32_t include_null_keys,
                      const bgb_request* requests, int32_t num_requests, void* stream, const bgb_allocator* alloc,
                      const bgb_options* options, bgb_groupby** out_handle, int32_t* num_groups);

/* result column (request r, aggregation a): type, scale and whether a null mask must be supplied */
int bgb_groupby_result_info(const bgb_groupby* handle, int32_t request, int32_t aggregation, int32_t* type_id,
                            int32_t* scale, int32_t* nullable);
/* key column k of the unique-keys table: nullable iff include_null_keys and the input key column has a mask */
int bgb_groupby_key_info(const bgb_groupby* handle, int32_t key, int32_t* type_id, int32_t* scale, int32_t* nullable);

/* Phase 2: gather unique keys and finalise every requested aggregation into caller-allocated columns.
 * out_keys[num_keys]; out_results flattened request-major (all aggregations of request 0, then request 1, ...).
 * Asynchronous on `stream` except that null_count fields are valid only after the stream is synchronised
 * by this call (it synchronises when any output is nullable). */
int bgb_groupby_emit(bgb_groupby* handle, bgb_mutable_column* out_keys, bgb_mutable_column* out_results, void* stream);

/* which kernels the last build used: "global", "smem", "partitioned", ... (diagnostics, tests) */
const char* bgb_groupby_path(const bgb_groupby* handle);

void bgb_groupby_destroy(bgb_groupby* handle);

/* Measurement hooks (bench.py): device time of kernel #index of the last build+emit on this handle, taken with CUDA
 * events on the call's stream when options->profile != 0. Returns BGB_INVALID_ARGUMENT past the last entry. */
int bgb_groupby_profile(const bgb_groupby* handle, int32_t index, const char** name, float* ms);
/* process-wide number of kernels this library has launched (every launch goes through one counter) */
int64_t bgb_kernel_launches(void);

/* ---- host-buffer convenience (end-to-end measurement and simple callers) ------------------------------------- */

/* Same call with HOST columns (pinned or pageable): copies inputs H2D in chunks overlapped with aggregation,
 * runs build+emit, copies the results D2H into host buffers the library allocates (*out_keys / *out_results: pinned host memory
 * from a cache of released result blocks, pageable memory when pinning fails; release ONLY with bgb_free_host_result). */
typedef struct bgb_host_column {
  int32_t type_id;
  int32_t size;
  void* data;          /* library-owned, size elements */
  uint32_t* null_mask; /* library-owned or NULL */
  int32_t null_count;
  int32_t reserved;
} bgb_host_column;
int bgb_groupby_host(const bgb_column_view* host_keys, int32_t num_keys, int32_t include_null_keys,
                     const bgb_request* host_requests, int32_t num_requests, const bgb_options* options,
                     int32_t* num_groups, bgb_host_column** out_keys, bgb_host_column** out_results,
                     int32_t* num_results);
void bgb_free_host_result(bgb_host_column* cols, int32_t n);

/* ---- streaming groupby: a persistent group table fed batch by batch -------------------------------------------------
 * cudf::groupby::streaming_groupby (cpp/include/cudf/groupby.hpp:428-598; src/groupby/streaming_groupby/). Key and value
 * columns are named by their index in the table passed to every aggregate() call, one aggregation per request
 * (streaming_aggregation_request, groupby.hpp:419-422). Supported kinds: SUM, SUM_OF_SQUARES, PRODUCT, MIN, MAX,
 * COUNT_VALID, COUNT_ALL, MEAN, M2, VARIANCE, STD (ARGMIN/ARGMAX are rejected: row indices are batch-local).
 * Cumulative rows are unbounded; cumulative distinct keys are bounded by max_distinct_keys, and so is a single batch
 * (the reference's limit, kept so callers see the same errors). */
typedef struct bgb_stream bgb_stream;
typedef struct bgb_stream_request {
  int32_t column_index;
  int32_t kind; /* bgb_agg_kind */
  int32_t ddof;
  int32_t reserved;
} bgb_stream_request;
int bgb_is_streaming_groupby_supported(int32_t values_type_id, int32_t kind, int32_t* out);
int bgb_stream_create(const int32_t* key_indices, int32_t num_keys, const bgb_stream_request* requests, int32_t num_requests,
                      int32_t max_distinct_keys, int32_t include_null_keys, bgb_stream** out);
/* one batch: columns[] is the data table (keys and values by index). The batch is aggregated with the stateless path into
 * a partial table and folded into the persistent one; returns after the stream is synchronised (distinct-key check). */
int bgb_stream_aggregate(bgb_stream* s, const bgb_column_view* columns, int32_t num_columns, void* stream, const bgb_options* options);
/* fold another object's accumulated state into this one (the other object is not modified) */
int bgb_stream_merge(bgb_stream* s, const bgb_stream* other, void* stream);
int bgb_stream_distinct_keys(const bgb_stream* s, int32_t* out);
/* compact the persistent table into a handle for bgb_groupby_key_info / _result_info / _emit / _destroy; the state is not
 * modified and aggregate() may be called again afterwards. The handle must be destroyed before the stream object. */
int bgb_stream_finalize(const bgb_stream* s, void* stream, bgb_groupby** out_handle, int32_t* num_groups);
void bgb_stream_destroy(bgb_stream* s);

/* ---- multi-GPU building blocks (one process per GPU; the exchange itself is an NCCL all-to-all run by the host layer,
 *      cudf_b200/distributed.py) --------------------------------------------------------------------- */

/* Permute `cols` (all with as many rows as the keys) into num_parts contiguous runs, run r holding the rows whose key has
 * hash2(key) % num_parts == r (null keys hash like any other packed image). out_cols[i].data receives the permuted
 * elements; for a column with a null mask, out_cols[i].null_mask must point at ONE BYTE PER ROW (uint8_t*, 1 = valid):
 * validity travels through the exchange as bytes and is packed again with bgb_bytes_to_bitmask on the receiving side.
 * part_offsets_host[num_parts + 1] (host) receives the run boundaries; the call synchronises `stream` once.
 * Role of cudf::hash_partition in the reference's distributed callers (cpp/include/cudf/partitioning.hpp:103-145). */
int bgb_partition_by_key_hash(const bgb_column_view* keys, int32_t num_keys, const bgb_column_view* cols, int32_t num_cols,
                              int32_t num_parts, bgb_mutable_column* out_cols, int32_t* part_offsets_host, void* stream);

/* ---- cudf::hash_partition drop-in -------------------------------------------------------------
 * Reorders the rows of `input` (num_cols fixed-width columns of equal size) so that rows with the same partition are
 * contiguous; partition = row_hash(columns_to_hash) % num_partitions, where row_hash is the REFERENCE's row hash, bit for
 * bit: MurmurHash3_x86_32 (seeded) of every hashed element's native bytes after normalising -0.0 / NaN and bool -> 0/1,
 * a null element hashing to UINT32_MAX, columns folded left to right with hash_combine
 * (cpp/include/cudf/hashing/detail/murmurhash3_x86_32.cuh:20-70, cpp/include/cudf/detail/row_operator/hashing.cuh:40-140,
 * cpp/include/cudf/hashing/detail/hashing.hpp:83-86, partitioners cpp/src/partitioning/partitioning.cu:41-77), or the
 * reference's IdentityHash (static_cast<uint32_t>(element), numeric columns only, partitioning.cu:855-874).
 * out_cols[i] (caller-allocated, >= rows of input) receives column i in partition order; when input[i] has nulls,
 * out_cols[i].null_mask (bitmask words) receives its validity and out_cols[i].null_count the count. offsets_host receives
 * num_partitions + 1 row offsets, partition p = rows [offsets[p], offsets[p + 1]). The order of rows INSIDE a partition
 * is unspecified, as in the reference (its block-level scatter ranks rows with shared-memory atomics).
 * num_partitions <= 0, zero rows or zero columns_to_hash give the reference's empty result: all offsets 0, nothing
 * written (partitioning.cu:885-888). Replaces cudf::hash_partition (cpp/include/cudf/partitioning.hpp:103-145,
 * cpp/src/partitioning/partitioning.cu:567-760,876-950). Synchronises `stream` once. Limits: <= 4096 partitions, <= 8
 * hashed columns, elements of at most 8 bytes (else BGB_UNSUPPORTED: the caller keeps the reference's own path). */
typedef enum bgb_hash_id { BGB_HASH_IDENTITY = 0, BGB_HASH_MURMUR3 = 1 } bgb_hash_id; /* cudf::hash_id, partitioning.hpp:28-31 */
#define BGB_DEFAULT_HASH_SEED 0u /* cudf::DEFAULT_HASH_SEED, cpp/include/cudf/hashing.hpp:29 */
int bgb_hash_partition(const bgb_column_view* input, int32_t num_cols, const int32_t* columns_to_hash, int32_t num_hash_cols,
                       int32_t num_partitions, int32_t hash_function, uint32_t seed, bgb_mutable_column* out_cols,
                       int32_t* offsets_host, void* stream);

/* validity bytes (1 = valid) -> cudf bitmask words ((n + 31) / 32 of them); *null_count_host = number of zero bytes */
int bgb_bytes_to_bitmask(const uint8_t* valid_bytes, int64_t n, uint32_t* mask_words, int32_t* null_count_host, void* stream);

/* The same conversion for callers that CANNOT ALLOCATE (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc). No counterpart in the reference. It calls no
 * cudaMalloc / cudaFree / memory-pool routine: its scratch (a ticket, the total, one partial null count per 8192-row tile) is
 * a file-scope __device__ array of 262 400 bytes that every GPU loading the library holds until the process exits
 * (cudf_b200/csrc/noalloc.cu). Limit: n <= bgb_bytes_to_bitmask_noalloc_max_rows() = 536 870 912 rows, beyond which the call
 * returns BGB_INVALID_ARGUMENT. All calls on one device share the array (the call clears what must start at zero), so calls
 * that OVERLAP on the GPU -- different streams without an ordering between them -- race; calls in stream order are safe.
 * null_count_host == NULL makes the call fully asynchronous on `stream` (no read-back, no synchronisation). */
int bgb_bytes_to_bitmask_noalloc(const uint8_t* valid_bytes, int64_t n, uint32_t* mask_words, int32_t* null_count_host, void* stream);
int64_t bgb_bytes_to_bitmask_noalloc_max_rows(void);

/* MEAN / M2 / VARIANCE / STD from MERGED partial columns: sum (and sumsq, unused for MEAN) are the merged SUM /
 * SUM_OF_SQUARES columns (INT64, UINT64, FLOAT32, FLOAT64, or a duration type for MEAN), count the merged INT32 count.
 * out->data receives n FLOAT64 values (the duration type for a duration MEAN); out->null_mask, when not NULL, receives
 * the validity words (MEAN: count > 0; VARIANCE/STD: count - ddof > 0) and out->null_count the number of nulls.
 * Same arithmetic as the single-GPU finalisers (cpp/src/groupby/hash/hash_compound_agg_finalizer.cu:94-133,
 * cpp/src/groupby/common/m2_var_std.cu:49-59,138-187). */
int bgb_finalize_moments(int32_t kind, int32_t ddof, const bgb_column_view* sum, const bgb_column_view* sumsq,
                         const bgb_column_view* count, bgb_mutable_column* out, void* stream);

/* ---- multi-GPU, peer-memory exchange: ONE fused path, no collective on the data path -------------------------------
 * Every rank owns an exchange buffer (bgb_peer_create) that its peers map through CUDA IPC (bgb_peer_export /
 * bgb_peer_attach; the 64-byte handles travel through whatever bootstrap channel the host layer has - torch.distributed
 * all_gather_object in cudf_b200/distributed.py). bgb_groupby_build_peer then (1) aggregates this rank's rows with the
 * single-GPU path and (2) writes one record per local group straight into the OWNER rank's buffer over NVLink (P2P stores;
 * owner = hash2(key) % world, as bgb_partition_by_key_hash) -- on the records-form path both happen in ONE kernel: every
 * table slice a CTA finishes is flushed to its owners from shared memory (path suffix "+peer-fused"), otherwise a scatter
 * kernel walks the local groups ("+peer") -- (3) publishes the per-destination counts and the step's epoch with system-scope
 * release stores, (4) waits for every source's epoch, and (5) merges the received records (SUM of sums, SUM of counts, MIN of
 * mins, MAX of maxes, PRODUCT of products - the merges of cpp/src/groupby/streaming_groupby/merge.cu:30-144): through the same
 * slice kernels with dense output when the records have one key word and at most three cells ("+slices"), else into an HBM
 * table that is then compacted. A rank whose step fails tells its peers (BGB_PEER_ABORTED on their side) so that nobody waits
 * for the timeout. The handle it returns is used exactly like
 * bgb_groupby_build's (result_info / emit / destroy) and holds the groups THIS rank owns; MEAN/M2/VARIANCE/STD are
 * finalised by emit from merged SUM/SUMSQ/COUNT cells, so integer inputs give bit-identical results on any world size.
 * The call is COLLECTIVE: every rank of the context must make it, with the same key/request shapes. Regions are
 * double-buffered by call parity, so no barrier is needed between consecutive calls. */
typedef struct bgb_peer bgb_peer;
#define BGB_PEER_HANDLE_BYTES 64
/* region_bytes: size of ONE receive region (two are allocated); a (source, destination) segment holds
 * region_bytes / (world * record_bytes) records, record_bytes = 8 * key words + accumulator cells (20 B for int64 SUM+COUNT) */
int bgb_peer_create(int32_t rank, int32_t world, int64_t region_bytes, bgb_peer** out);
int bgb_peer_export(const bgb_peer* peer, void* handle_out /* BGB_PEER_HANDLE_BYTES */);
int bgb_peer_attach(bgb_peer* peer, int32_t peer_rank, const void* handle /* from that rank's bgb_peer_export */);
/* same-process ranks (one process driving several GPUs, or several simulated ranks on one GPU in the tests) */
int bgb_peer_attach_local(bgb_peer* peer, int32_t peer_rank, const bgb_peer* other);
int64_t bgb_peer_segment_capacity(const bgb_peer* peer, int32_t record_bytes);
void bgb_peer_destroy(bgb_peer* peer);
int bgb_groupby_build_peer(bgb_peer* peer, const bgb_column_view* keys, int32_t num_keys, int32_t include_null_keys,
                           const bgb_request* requests, int32_t num_requests, void* stream, const bgb_allocator* alloc,
                           const bgb_options* options, bgb_groupby** out_handle, int32_t* num_groups);

/* ---- synthetic inputs (SURVEY.md 8d generators; bench + tests only) --------------------------------------------- */

/* out[i] = splitmix64(seed, first_row + i) % modulus (modulus 0: raw 64 bits), cast to type_id (ints) */
int bgb_gen_uniform_int(void* out, int32_t type_id, int64_t n, uint64_t seed, uint64_t first_row, uint64_t modulus,
                        int64_t add, void* stream);
/* out[i] = (double|float)(splitmix64 % modulus) + add */
int bgb_gen_uniform_real(void* out, int32_t type_id, int64_t n, uint64_t seed, uint64_t first_row, uint64_t modulus,
                         double add, void* stream);
/* validity bit i cleared when splitmix64(seed, first_row+i) % null_every == 0; returns null count in *null_count (host) */
int bgb_gen_null_mask(uint32_t* mask_words, int64_t n, uint64_t seed, uint64_t first_row, uint32_t null_every,
                      int32_t* null_count, void* stream);
/* categorical by cumulative thresholds out of 65536: out[i] = values[first j with (x & 0xFFFF) < cum[j]] (int8) */
int bgb_gen_categorical_i8(int8_t* out, int64_t n, uint64_t seed, uint64_t first_row, const int32_t* cum_host,
                           const int8_t* values_host, int32_t num_categories, void* stream);

const char* bgb_last_error(void);
int bgb_version(void);
/* The library keeps freed scratch blocks (group tables, partition copies) cached per device and stream so that a repeated
 * query does not re-enter the driver allocator (at most 40 % of the device, BGB_CACHE_LIMIT_MB overrides). This hands
 * every cached block of the current device back (the role of rmm's pool release / torch.cuda.empty_cache). */
void bgb_release_cached_memory(void);

#ifdef __cplusplus
}
#endif
#endif /* BGB_H */

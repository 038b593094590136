// Parameters of the bucketed high-cardinality path (kernels_bucket.cu): one or two radix-scatter levels that bring every
// row into the bucket of its table SLICE, then one CTA per slice aggregates in shared memory and stores the slice.
#pragma once
#include "bgb_internal.cuh"

namespace bgb {

constexpr int kMaxBucketValues = 16;

// Packed row format between the levels (structure of arrays, bucket b owns rows [b * capacity, b * capacity + count)):
//   key   : one 64-bit word per row = the packed key image; when the image needs <= 32 bits the validity word of the row
//           (bit i = value column with vbit i is valid) rides in the upper half, else it has its own column (1 or 4 bytes)
//   values: one array per value column in its native width
struct BucketValue {
  const uint8_t* in;        // level 1: the caller's column; level 2: the previous level's packed column
  const uint32_t* in_mask;  // level 1 only: validity bitmask or nullptr
  int32_t in_offset;        // level 1 only
  int32_t width;            // element bytes 1/2/4/8
  int32_t vbit;             // bit of this column in the validity word, -1 when the column has no mask
  uint8_t* out;
};

struct BucketScatterParams {
  AggParams a;               // level 1: key columns (pack_key) and num_rows; both levels: table.capacity
  int32_t level;             // 1: rows [0, num_rows) of the caller's columns; 2: packed rows in num_segs segments
  int32_t key_mode;          // level 1: 0 general pack_key, 4 / 8 one unmasked integer-like column of that width
  const uint64_t* in_key;    // level 2 input
  const uint8_t* in_vbits;
  const uint32_t* seg_cursor;   // level 2: rows written to segment s = min(seg_cursor[s], seg_capacity)
  uint32_t seg_capacity;
  uint32_t num_segs;
  int32_t log2_parts;        // partitions of this level (per segment)
  int32_t slot_shift;        // partition = ((hash & (capacity - 1)) >> slot_shift) & (P - 1)
  int32_t key_in_low32;      // validity word rides in the upper half of the key word
  int32_t vbits_width;       // separate validity column: 0 none, 1 or 4 bytes
  uint32_t* cursor;          // [num_segs << log2_parts] rows reserved per output bucket (may exceed bucket_capacity)
  uint32_t bucket_capacity;
  uint32_t* ov_cursor;       // rows that found their bucket full go to the overflow region [ov_base, ov_base + ov_capacity)
  uint32_t ov_capacity;
  uint32_t ov_base;          // = number of buckets * bucket_capacity; row ov_base + ov_capacity is a scratch row
  uint32_t* status;          // status[6] = 1 when the overflow region itself overflowed (host reruns on another path)
  uint64_t* out_key;
  uint8_t* out_vbits;
  int32_t num_vals;
  BucketValue v[kMaxBucketValues];
};

// Where an accumulator lives while a bucket is aggregated. Shared memory only pays for operations that are one
// non-returning 32-bit ATOMS (measured 22.6 lanes/ns/SM); 64-bit and floating-point shared atomics are CAS loops
// (~4 cycles per lane, slower than a RED to an L2-resident line), so those accumulators stay in the bucket's own slice of
// the HBM table, which only this CTA touches and which therefore sits in L2 while the bucket is processed.
enum BucketAccMode : uint8_t {
  BM_COUNT,    // u32 cell in shared memory
  BM_LIMBS,    // integer SUM / SUM_OF_SQUARES: three u32 arrays in shared memory holding the balanced 16-bit digits d0, d1 and
               // the rest r2 of every addend (v = d0 + 2^16 d1 + 2^32 r2); fewer than 65536 addends go into a digit sum before
               // it is folded (one fold per bucket, or one every ~64K rows of a long bucket into a 64-bit total), so no digit
               // sum leaves 32 bits and sum mod 2^64 = s0 + 2^16 s1 + 2^32 s2 exactly, without carries between the arrays
  BM_MINMAX,   // u64 ordered image in shared memory: plain load first, CAS loop only when the row improves the extremum
  BM_GLOBAL    // RED on the slice in the HBM table (float / double sums and sums of squares, products)
};

struct BucketAcc {
  const uint8_t* data;       // packed value column (nullptr: COUNT_ALL)
  void* cells;               // the accumulator's array in the HBM table
  uint32_t smem_off;         // byte offset of the accumulator's slice in shared memory
  int16_t vbit;              // validity bit of the value column, -1: always valid
  uint8_t src, op;
  uint8_t mode;              // BucketAccMode
};

struct BucketAggParams {
  TableDesc table;           // key_words == 1
  const uint64_t* key;
  const uint8_t* vbits;
  int32_t key_in_low32;
  int32_t vbits_width;
  const uint32_t* cursor;    // [num_buckets]
  uint32_t bucket_capacity;
  uint32_t num_buckets;      // = capacity >> log2_slice
  int32_t log2_slice;        // slots per bucket
  uint32_t loc_off;          // shared memory: 16-bit slot of every row of the bucket (bucket_capacity entries)
  uint32_t vb_off;           // shared memory: validity word of every row (vb_width bytes each; only when some accumulator needs it)
  int32_t vb_width;          // 0: no masked value column
  uint32_t chunk_rows;       // rows of a bucket processed per pass (= entries of the slot / validity arrays above)
  int32_t limb_total;        // buckets may exceed 65535 rows: BM_LIMBS accumulators carry a 64-bit total behind their digit sums
  int32_t num_accs;
  // overflow pass: rows [ov_base, ov_base + min(*ov_cursor, ov_capacity)) through the HBM table
  const uint32_t* ov_cursor;
  uint32_t ov_capacity;
  uint32_t ov_base;
  BucketAcc acc[kMaxAccs];
};

constexpr uint32_t kBucketTile = 4096;   // rows per scatter tile
size_t bucket_scatter_smem_bytes(int log2_parts);
cudaError_t launch_bucket_scatter(const BucketScatterParams& p, int num_sms, cudaStream_t s);
cudaError_t launch_bucket_init_reserved(const BucketAggParams& p, cudaStream_t s);
cudaError_t launch_bucket_aggregate(const BucketAggParams& p, size_t smem_bytes, cudaStream_t s);
cudaError_t launch_bucket_overflow(const BucketAggParams& p, cudaStream_t s);

}  // namespace bgb

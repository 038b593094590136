// Parameters of the radix-partitioning kernels (kernels_partition.cu).
#pragma once
#include "bgb_internal.cuh"

namespace bgb {

constexpr int kMaxPartValues = 16;

struct PartValue {
  const uint8_t* data;      // source value column
  const uint32_t* mask;     // source bitmask or nullptr
  int32_t offset;
  int32_t width;            // element bytes 1/2/4/8
  uint8_t* out_data;        // partition-ordered copy (null elements written as 0)
  uint8_t* out_valid;       // one byte per row, only when mask != nullptr
};

struct PartParams {
  AggParams a;              // key columns, num_rows, table (capacity + key_words decide the partition of a key)
  int32_t log2_parts;       // P = 1 << log2_parts, P <= 4096
  int32_t slot_shift;       // partition = slot >> slot_shift  (slot_shift = log2(capacity) - log2_parts)
  int32_t mod_parts;        // > 0: partition = hash2(key) % mod_parts instead (multi-GPU exchange: owner rank of a key)
  int32_t cudf_hash;        // != 0 (with mod_parts > 0): partition = cudf row hash % mod_parts (1 MurmurHash3_x86_32, 2 identity;
                            // columns folded with hash_combine, nulls hash to UINT32_MAX, every row kept): the assignment of
                            // cudf::hash_partition, bit for bit. The key columns are read in a.kc[] order and never packed.
  uint32_t hash_seed;
  uint32_t signed_cols;     // identity hash: bit c set = key column c is a signed integer (sign-extend before truncation)
  uint32_t* out_row;        // != nullptr: source row of every output row (gather map for tables wider than kMaxPartValues)
  uint32_t* counts;         // [P]   rows per partition (zeroed by the caller)
  uint32_t* offsets;        // [P+1] exclusive scan; offsets[P] = rows kept (rows with a null key under EXCLUDE are dropped)
  uint32_t* cursors;        // [P]   running write positions
  uint64_t* out_k0;         // packed key, low word (8-byte elements, or 4-byte when key_out_width == 4); nullptr: not written
  uint64_t* out_k1;         // high word when key_words == 2
  int32_t key_out_width;
  int32_t num_vals;
  PartValue v[kMaxPartValues];
};

cudaError_t launch_partition_hist(const PartParams& p, int num_sms, cudaStream_t s);
cudaError_t launch_partition_scan(const PartParams& p, cudaStream_t s);
cudaError_t launch_partition_scatter(const PartParams& p, int num_sms, cudaStream_t s);

}  // namespace bgb

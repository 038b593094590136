// Parameters of the tile-sorting mid-cardinality kernel (kernels_sorted.cu): up to a few thousand groups with a dense key
// range, ANY accumulator set (float sums, sums of squares, MIN/MAX, products, several value columns).
#pragma once
#include "bgb_internal.cuh"

namespace bgb {

constexpr int kSortedMaxCols   = 4;     // distinct value columns
constexpr int kSortedThreads   = 512;
constexpr int kSortedRows      = 4;     // rows per thread and tile
constexpr int kSortedTile      = kSortedThreads * kSortedRows;
constexpr int kSortedMaxGroups = kSortedThreads * 8;  // every thread owns at most 8 consecutive dense ids
constexpr int kSortedHeavy     = 64;    // a group with more rows than this in one tile is reduced by a whole warp

struct SortedCol {
  const uint8_t* data;
  const uint32_t* mask;
  int32_t offset;
  uint8_t width;   // element bytes
  uint8_t src;     // Src
};

struct SortedAcc {
  void* cells;     // accumulator array in the HBM table
  uint8_t op;      // AccOp (no ARG*)
  int8_t col;      // index into col[], -1: COUNT_ALL
};

struct SortedParams {
  AggParams a;               // key columns, table, num_rows, skip_null_keys (acc[] unused)
  SortedCol col[kSortedMaxCols];
  SortedAcc acc[kMaxAccs];
  int32_t num_cols, num_accs;
  uint32_t groups;           // dense ids [0, groups): id = (key ^ key_flip) - key_base; hashed: slots of the CTA's key table (power of two)
  int32_t hashed;            // != 0: id = slot of the key in a shared-memory open-addressing table (any one-word key image)
  int32_t key_mode;          // 2: one unmasked 8-byte integer key column, 1: one unmasked integer-like key column, 0: pack_key
  uint64_t key_flip, key_base;
  int32_t any_mask;          // some value column carries a validity mask
  int32_t fast8;             // every value column is 8 bytes wide without a mask (int64 / uint64 / float64): plain 64-bit loads
};

size_t sorted_smem_bytes(uint32_t groups, int num_accs, int num_cols, bool hashed = false);
cudaError_t launch_aggregate_sorted(const SortedParams& p, int num_sms, cudaStream_t s);

}  // namespace bgb

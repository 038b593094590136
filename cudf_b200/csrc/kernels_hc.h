// Round-2 high-cardinality path ("records" form of the bucketed path, kernels_hc.cu): rows travel between the levels as
// ARRAY-OF-STRUCTURES records of W 8-byte words -- word 0 the packed key image (validity bits of the value columns in its upper
// half when the image needs <= 32 bits), one word per value column, optionally a last word of validity bits -- so that a
// destination run is one contiguous byte range (256 B per bucket and tile at 256 buckets), every tile arrives in shared
// memory through ONE cp.async.bulk (TMA) per source, and the slice kernel reads a row as one 128-bit load. One integer value column
// without nulls under SUM / COUNT whose sampled key and value ranges fit 32 bits travels as ONE word {key32, value32} (pack32: every
// row is checked in level 1, status[7] = 1 reruns the call with two words).
//   hc_scatter (level 1)  caller's columns -> records in P1 buckets      (TMA raw column tiles, or a generic LDG packer)
//   hc_scatter (level 2)  records of every level-1 bucket -> P2 buckets  (TMA)
//   hc_slices             one CTA per final bucket: table slice in shared memory, DENSE group output (no HBM hash table,
//                         no compaction pass): groups land at [0, G) of the table arrays, group_slots = identity
//   hc_finish             group count (+ the reserved slot of the all-ones key) -> status[1]
// Replaces what the reference does with per-row probes of an HBM cuco::static_set
// (/root/reference/cpp/src/groupby/hash/compute_global_memory_aggs.cuh:123-157).
#pragma once
#include "bgb_internal.cuh"

namespace bgb {

constexpr int kHcMaxWords = 4;     // record words: key + up to 3 value / validity words
constexpr int kHcMaxParts = 1024;  // buckets per scatter level
constexpr int kHcThreads  = 512;

// Multi-GPU fusion (kernels_hc.cu, hc_slice_kernel flush): instead of appending a finished slice to the local dense table, the
// CTA writes one record per group STRAIGHT INTO THE OWNER RANK'S exchange region over NVLink (peer stores, ordered by owner in
// shared memory so that consecutive threads write consecutive positions of one region). The record layout is the one
// peer_merge / the merge-mode scatter read: key array, then one array per accumulator, each world x cap entries.
constexpr int kHcMaxPeers = 64;
struct HcPeerOut {
  int32_t world;             // 0: not fused (dense local output)
  int32_t me;
  uint64_t cap;              // records per (source, destination) segment
  uint32_t* cursors;         // [world] records produced for each destination (local memory, zeroed)
  uint8_t* region[kHcMaxPeers];
  uint64_t acc_off[kMaxAccs];  // byte offset of every accumulator's array inside a region (key array at 0)
  uint32_t perm_off;         // shared memory: u16 perm[S] then u8 own[S] behind the slice
};

// Final merge of the peer exchange as an hc pass: level 1 reads RECEIVED RECORDS (world segments of a region) instead of columns
struct HcPeerIn {
  const uint8_t* region;     // nullptr: not a merge pass
  uint64_t cap;
  int32_t world;
  int32_t num_accs;
  uint32_t prefix[kHcMaxPeers + 1];   // records received from sources [0, s)
  uint64_t acc_off[kHcMaxWords - 1];
  uint8_t cell_bytes[kHcMaxWords - 1];
};

struct HcValue {
  const uint8_t* data;
  const uint32_t* mask;
  int32_t offset;
  int32_t width;   // 1/2/4/8
  int32_t vbit;    // bit in the validity word, -1: no mask
  int32_t src;     // Src of the column (pack32: how the raw element is sign-extended)
};

struct HcScatterParams {
  AggParams a;               // level 1: key columns (generic packer), num_rows; both levels: table.capacity
  HcValue val[kHcMaxWords - 1];
  int32_t num_vals;
  int32_t level;             // 1: caller's columns, 2: records in num_segs segments
  int32_t tma_src;           // level 1: every source column is staged as a raw tile by TMA (see hc_tma_eligible)
  int32_t words;             // W
  int32_t key_in_low32;      // validity bits ride in the upper half of word 0
  int32_t vbits_word;        // index of the separate validity word, -1: none
  int32_t pack32;            // level 1, W == 1: the record is {key (low 32 bits), the ONE value column as int32 (high 32 bits)}; a row whose
                             // key or value does not fit sets status[7] and the host repeats the call with 16-byte records
  const uint64_t* in;        // level 2 input records
  const uint32_t* seg_cursor;
  uint32_t seg_cursor_stride;
  uint32_t seg_capacity;
  uint32_t num_segs;
  uint64_t* out;
  uint32_t* cursor;          // rows reserved per output bucket (may exceed bucket_capacity: status[6])
  uint32_t cursor_stride;    // in words: hot level-1 cursors sit 256 B apart (separate L2 slices)
  uint32_t bucket_capacity;
  int32_t log2_parts;        // buckets of this level (per segment)
  int32_t slot_shift;        // bucket = ((slot_hash32(key) & (capacity - 1)) >> slot_shift) & (P - 1)
  uint32_t* status;          // status[6] = 1: a bucket overflowed (skewed keys): the host reruns the call on another path
  HcPeerIn peer_in;          // level 1 of a merge pass
};

// accumulator of the slice kernel: where its cell lives in shared memory and how a record word updates it
enum HcMode : uint8_t { HM_COUNT, HM_ISUM, HM_MINMAX, HM_FSUM64, HM_FSUM32, HM_OTHER,
                        HM_ADD32 };  // merge of partial counts: the record word is ADDED to a 32-bit cell (peer exchange, final merge)
struct HcAcc {
  void* cells;               // the accumulator's array in the (dense) table
  uint32_t smem_off;         // byte offset of its cell array in shared memory
  int16_t vbit;              // validity bit, -1: always valid
  uint8_t word;              // record word holding the value (0: no value, COUNT_ALL)
  uint8_t src, op, mode;
  uint8_t cell_bytes;
};

struct HcSliceParams {
  TableDesc table;           // keys / status; capacity = allocated entries - 1 (slot `capacity` = reserved key)
  const uint64_t* rec;
  const uint32_t* cursor;
  uint32_t cursor_stride;
  uint32_t bucket_capacity;
  uint32_t num_buckets;
  int32_t log2_slice;
  int32_t words;
  int32_t key_in_low32;
  int32_t vbits_word;
  int32_t f64stats;          // every accumulator is SUM / SUMSQ / MIN / MAX / COUNT of ONE float64 value column (stat_vbit: its validity bit, -1 none)
  int32_t stat_vbit;
  int32_t val_hi32;          // pack32 records: the value of the (one) value column is the sign-extended upper half of word 0
  int32_t num_accs;
  uint32_t* group_slots;     // identity for the dense groups
  unsigned long long* gcount;
  uint64_t max_groups;       // entries the dense arrays can take
  HcPeerOut peer;
  HcAcc acc[kMaxAccs];
};

size_t hc_scatter_smem_bytes(int words, int log2_parts);
int hc_tile_rows(int words);
cudaError_t launch_hc_scatter(const HcScatterParams& p, int num_sms, cudaStream_t s);
cudaError_t launch_hc_init(const HcSliceParams& p, cudaStream_t s);
cudaError_t launch_hc_slices(const HcSliceParams& p, size_t smem_bytes, int num_sms, cudaStream_t s);
cudaError_t launch_hc_finish(const HcSliceParams& p, cudaStream_t s);

}  // namespace bgb

// Internal types shared by the host planner and the kernels of the B200 hash groupby.
// Layout of the group table in HBM (SoA):
//   keys    : capacity+1 packed keys (64-bit, or 128-bit as ulonglong2), EMPTY = all ones; slot `capacity` is the
//             reserved home of the one real key whose packed image equals EMPTY
//   cells[a]: one array per accumulator, capacity+1 cells of 4 or 8 bytes, initialised to the aggregation identity
// Accumulator semantics follow cudf::detail::update_target_element
// (/root/reference/cpp/include/cudf/detail/aggregation/device_aggregators.cuh:24-160,309-446).
#pragma once
#include <mutex>
#include <unordered_set>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/bgb.h"

namespace bgb {

constexpr int kMaxKeyCols = 8;
constexpr int kMaxAccs    = 24;   // accumulator columns per kernel launch
constexpr int kMaxOuts    = 32;   // output columns per emit launch
constexpr uint64_t kEmpty = ~0ull;

// physical source element types
enum Src : uint8_t { S_I8, S_I16, S_I32, S_I64, S_U8, S_U16, S_U32, S_U64, S_F32, S_F64, S_BOOL, S_BAD };

// accumulator ops; cell width is 8 bytes except ACC_COUNT* and *_F32 (4 bytes)
enum AccOp : uint8_t {
  ACC_SUM_I64,    // any integral source, wraps like int64 atomic_add
  ACC_SUM_F64,
  ACC_SUM_F32,
  ACC_SUMSQ_I64,
  ACC_SUMSQ_F64,
  ACC_SUMSQ_F32,
  ACC_PROD_I64,
  ACC_PROD_F64,
  ACC_PROD_F32,
  ACC_MIN,        // order-preserving unsigned 64-bit image of the source value, RED.MIN.U64
  ACC_MAX,
  ACC_COUNT_VALID,
  ACC_COUNT_ALL,
  ACC_ARGMIN,     // (ordered value image << 32 | row) for <= 32-bit sources
  ACC_ARGMAX,
  ACC_ARGROW_MIN, // ARGMIN / ARGMAX of 64-bit values: a 4-byte row-number cell filled by a SECOND pass over the rows once the
  ACC_ARGROW_MAX, // group's MIN / MAX cell (AccDesc::aux) is final: smallest row whose value equals the extremum
  // SUM_OVERFLOW keeps the EXACT total of a signed column as two wrapping 64-bit sums, total = hi * 2^32 + lo (neither can wrap
  // below 2^31 rows), so the overflow flag does not depend on the order the atomics land in
  ACC_SUM_LO32,   // sum of the low 32 bits of every value, zero-extended
  ACC_SUM_HI32    // sum of (value >> 32), arithmetic
};

__host__ __device__ inline int acc_cell_bytes(uint8_t op)
{
  return (op == ACC_SUM_F32 || op == ACC_SUMSQ_F32 || op == ACC_PROD_F32 || op == ACC_COUNT_VALID || op == ACC_COUNT_ALL ||
          op == ACC_ARGROW_MIN || op == ACC_ARGROW_MAX) ? 4 : 8;
}

struct KeyColDesc {
  const uint8_t* data;
  const uint32_t* mask;   // nullptr = no nulls
  int32_t offset;
  uint8_t width;          // bytes: 1,2,4,8
  uint8_t fclass;         // 0 integer-like, 1 float32, 2 float64 (normalise -0 / NaN), 3 bool (non-zero -> 1)
  uint8_t shift;          // bit position of the value field in the packed key (0..127)
  int8_t null_bit;        // bit position of the "is null" flag (INCLUDE policy), -1 if none
};

struct AccDesc {
  const uint8_t* data;
  const uint32_t* mask;
  int32_t offset;
  uint8_t src;            // Src
  uint8_t op;             // AccOp
  uint8_t byte_mask;      // mask points at one validity BYTE per row (partitioned copies) instead of a bitmask
  uint8_t idx_width;      // dict != nullptr: bytes of one dictionary index (`data` holds the indices)
  const uint8_t* dict;    // DICTIONARY32 values: the dictionary (keys child); the element of row r is dict[data[r]]
  void* cells;            // capacity+1 cells
  const void* aux;        // ACC_ARGROW_*: the MIN / MAX cells of the same value column
};

struct TableDesc {
  uint64_t* keys;         // 64-bit keys, or 128-bit keys as pairs when key_words == 2
  uint64_t capacity;      // power of two
  uint32_t* status;       // 16 words: [0] overflow flag, [1] number of groups (written by compaction), [2] reserved slot used,
                          // [4..5] debug counters, [8..15] cardinality estimator outputs
  int32_t key_words;      // 1 or 2
};

struct AggParams {
  KeyColDesc kc[kMaxKeyCols];
  AccDesc acc[kMaxAccs];
  TableDesc table;
  int64_t num_rows;
  int32_t num_key_cols;
  int32_t num_accs;
  int32_t skip_null_keys;  // null_policy::EXCLUDE and some key column has a mask
  int32_t init_keys;       // init kernel: also reset keys + status (first accumulator chunk)
  const uint32_t* num_rows_dev;  // when set, the row count lives on the device (partitioned copy: rows kept by the scatter)
  int32_t wide;            // key image wider than 128 bits: table slots hold the ROW NUMBER of the group's first row (find_or_insert_row)
};

// final (output) operations of the emit kernel
enum OutOp : uint8_t {
  OUT_KEY,        // unpack key column
  OUT_KEY_GATHER, // wide keys: the slot holds a row number; the key element is read from the caller's column (a = data, b = mask,
                  // key_base = offset, acc_op = float class)
  OUT_COPY_I64,   // 8-byte integer cell -> out type (int64 for sums, narrowed ordered image for MIN/MAX)
  OUT_COPY_F64,
  OUT_COPY_F32,
  OUT_MINMAX,     // decode ordered image to the source type
  OUT_COUNT,      // int32 cell
  OUT_MEAN,       // double(sum) / double(count)
  OUT_MEAN_INT,   // duration / decimal MEAN keeps the source rep: integer division
  OUT_M2,
  OUT_VAR,
  OUT_STD,
  OUT_ARG,        // low 32 bits of the packed (value,row) cell
  OUT_SUMOV_SUM,  // SUM_OVERFLOW child 0: low bits of hi * 2^32 + lo in the source width (a = lo cells, b = hi cells); carries the struct mask
  OUT_SUMOV_FLAG  // SUM_OVERFLOW child 1 (BOOL8): the exact total does not fit the source type (shift = source width in bytes)
};

struct OutDesc {
  void* out;              // output data
  uint32_t* out_mask;     // output null mask or nullptr
  int32_t* null_count;    // device counter (or nullptr)
  const void* a;          // primary cell array (sum / min / max / count ...)
  const void* b;          // sum of squares
  const int32_t* cnt;     // COUNT_VALID cells of the value column (validity + MEAN/VAR), may be nullptr
  int32_t ddof;
  uint8_t op;             // OutOp
  uint8_t src;            // Src of the value (or key) column
  uint8_t out_width;      // bytes of the output element
  uint8_t acc_op;         // AccOp that produced `a` (cell width / numeric class)
  // OUT_KEY only
  uint8_t shift;
  int8_t null_bit;
  uint8_t key_bits;       // != 0: the field is (value - key_base) in key_bits bits (repacked keys)
  uint8_t pad;
  uint64_t key_base;
};

struct EmitParams {
  OutDesc out[kMaxOuts];
  const uint64_t* keys;
  const uint32_t* group_slots;  // slot index of group j
  int32_t num_groups;
  int32_t num_outs;
  int32_t key_words;
  int32_t dense;                // group j sits at entry j (no slot list to read), except ...
  uint32_t last_slot;           // ... the last group when it is the reserved entry: its index (else 0xFFFFFFFF)
};

// ------------------------------------------------------------------ device helpers shared by kernels
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z)
{
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
// counter-based SplitMix64 x(seed,row) -- identical to oracle/groupby_oracle.c orc_splitmix64
__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t seed, uint64_t row)
{
  uint64_t z = seed * 0x1000003ull + row + 0x9E3779B97F4A7C15ull;
  return mix64(z);
}

// ---- key repacking (kernels_repack.cu): several integer key columns whose packed image would need more than 64 bits are
// range-compressed into ONE 64-bit key column, (value - column minimum) in just enough bits, so that every single-word fast
// path serves them. The reference hashes and compares the key columns row by row whatever their ranges
// (/root/reference/cpp/include/cudf/detail/row_operator/hashing.cuh:118-134, equality.cuh:59-89).
struct RepackCol {
  const uint8_t* data;
  const uint32_t* mask;
  int32_t offset;
  uint8_t width;          // 1/2/4/8
  uint8_t is_signed;      // sign-extend the element before subtracting the base
  uint8_t is_bool;        // any non-zero byte is 1
  uint8_t shift;          // bit position of the field in the packed word
  int8_t null_bit;        // INCLUDE policy: position of the column's null flag, -1: none
  uint64_t base;          // column minimum (as the sign- or zero-extended 64-bit value)
};
struct RepackParams {
  RepackCol col[kMaxKeyCols];
  int32_t num_cols;
  int32_t drop_null_rows; // EXCLUDE policy: a row with a null key loses its bit in out_mask
  int64_t num_rows;
  uint64_t* minmax;       // [2 * num_cols]: min, max per column over the valid elements (minmax kernel output)
  uint64_t* out;          // packed key column
  uint32_t* out_mask;     // validity words of the packed column (EXCLUDE + some column has a mask), else nullptr
};
cudaError_t launch_repack_minmax(const RepackParams& p, cudaStream_t s);
cudaError_t launch_repack_pack(const RepackParams& p, cudaStream_t s);

// host-side API of the kernel translation units
cudaError_t launch_init_table(const AggParams& p, cudaStream_t s);
cudaError_t launch_aggregate_global(const AggParams& p, cudaStream_t s);
// second pass for ACC_ARGROW_* accumulators (after every row has been aggregated)
cudaError_t launch_argrow(const AggParams& p, cudaStream_t s);
cudaError_t launch_estimate_cardinality(const AggParams& p, int64_t sample_rows, uint64_t* scratch_set, uint64_t set_capacity,
                                        uint32_t* out_distinct /* 8 words */, uint64_t key_flip, cudaStream_t s);
// compaction: fills group_slots[] with occupied slot ids (arbitrary order), status[1] = count
cudaError_t launch_compact(const TableDesc& t, uint32_t* group_slots, uint32_t* scan_state, cudaStream_t s);
size_t compact_scan_state_bytes(uint64_t capacity);
cudaError_t launch_emit(const EmitParams& p, cudaStream_t s);


// Opt a kernel in to the device's full dynamic shared memory (227 KB minus its static shared memory), once per device. The attribute
// is per FUNCTION, not per launch: setting it to each launch's own size let two host threads with different sizes race -- thread A
// lowers it between thread B's set and B's launch, and B's launch fails with cudaErrorInvalidValue (seen with simulated ranks whose
// merge slices differed in size). Every caller sets the same maximum, so the order of the calls no longer matters.
inline std::mutex g_smem_optin_mu;
inline std::unordered_set<unsigned long long> g_smem_optin_done;  // (kernel address, device) pairs already set up

template <typename Kernel>
inline cudaError_t allow_max_dynamic_smem(Kernel kernel)
{
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const unsigned long long key = ((unsigned long long)(uintptr_t)reinterpret_cast<const void*>(kernel) << 6) ^ (unsigned long long)(dev & 63);
  {
    std::lock_guard<std::mutex> lock(g_smem_optin_mu);
    if (g_smem_optin_done.count(key)) return cudaSuccess;
  }
  int optin = 0;
  e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  if (e != cudaSuccess) return e;
  cudaFuncAttributes fa{};
  e = cudaFuncGetAttributes(&fa, kernel);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, optin - (int)fa.sharedSizeBytes);
  if (e == cudaSuccess) {
    std::lock_guard<std::mutex> lock(g_smem_optin_mu);
    g_smem_optin_done.insert(key);
  }
  return e;
}

}  // namespace bgb

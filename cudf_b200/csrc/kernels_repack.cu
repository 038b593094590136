// Key repacking: exact per-column minimum / maximum, then one 64-bit key column holding (value - minimum) of every key
// column in just enough bits (see RepackParams in bgb_internal.cuh). HBM-bound streaming passes over the key columns.
#include "device_utils.cuh"

namespace bgb {
namespace {
constexpr int kRThreads = 256;

// raw element bits -> the 64-bit value the field encodes (bool: 0/1, signed: sign-extended, else zero-extended)
__device__ __forceinline__ uint64_t repack_extend(const RepackCol& c, uint64_t raw)
{
  if (c.is_bool) return raw != 0;
  if (c.is_signed) { const int sh = 64 - 8 * c.width; return (uint64_t)(((int64_t)(raw << sh)) >> sh); }
  return raw;
}
// order-preserving image of the extended value: signed columns compare as int64, unsigned as uint64
__device__ __forceinline__ uint64_t repack_order(const RepackCol& c, uint64_t v) { return c.is_signed ? v ^ 0x8000000000000000ull : v; }
}  // namespace

// minmax[2c] = min, minmax[2c + 1] = max of the ORDER IMAGE of column c over its valid elements (initialised by the host
// to all-ones / zero). blockIdx.y = column: a block streams ONE column, eight rows per thread in flight (the all-columns
// form kept 16 running extrema per thread: 107 registers, 25 % occupancy, 1 TB/s).
__global__ void __launch_bounds__(kRThreads) repack_minmax_kernel(const __grid_constant__ RepackParams p)
{
  constexpr int R = 8;
  const RepackCol& col = p.col[blockIdx.y];
  const int64_t stride = (int64_t)gridDim.x * kRThreads * R;
  uint64_t lo = ~0ull, hi = 0;
  for (int64_t base = (int64_t)blockIdx.x * kRThreads * R + threadIdx.x; base < p.num_rows; base += stride) {
    uint64_t v[R];
    uint32_t ok = 0;
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const int64_t row = base + (int64_t)j * kRThreads;
      v[j] = 0;
      if (row < p.num_rows) { v[j] = load_raw(col.data, col.width, (int64_t)col.offset + row); ok |= 1u << j; }
    }
    if (col.mask != nullptr) {
#pragma unroll
      for (int j = 0; j < R; ++j)
        if (((ok >> j) & 1u) && !bit_is_set(col.mask, (int64_t)col.offset + base + (int64_t)j * kRThreads)) ok &= ~(1u << j);
    }
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const uint64_t o = repack_order(col, repack_extend(col, v[j]));
      if ((ok >> j) & 1u) { lo = o < lo ? o : lo; hi = o > hi ? o : hi; }
    }
  }
  unsigned long long a = lo, b = hi;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    const unsigned long long a2 = __shfl_xor_sync(0xffffffffu, a, d), b2 = __shfl_xor_sync(0xffffffffu, b, d);
    a = a2 < a ? a2 : a; b = b2 > b ? b2 : b;
  }
  if ((threadIdx.x & 31) == 0 && a <= b) {
    atomicMin(reinterpret_cast<unsigned long long*>(p.minmax) + 2 * blockIdx.y, a);
    atomicMax(reinterpret_cast<unsigned long long*>(p.minmax) + 2 * blockIdx.y + 1, b);
  }
}

// four rows per thread; the 32 lanes of a warp hold 32 consecutive rows per j (the validity word of the packed column is
// one ballot). Columns are walked by a runtime loop: one column's loads are in flight at a time, the image stays in registers.
__global__ void __launch_bounds__(kRThreads) repack_pack_kernel(const __grid_constant__ RepackParams p)
{
  constexpr int R = 4;
  const int64_t stride = (int64_t)gridDim.x * kRThreads * R;
  const int64_t n_up   = (p.num_rows + 31) & ~31ll;  // whole warps stay together for the ballot
  for (int64_t base = (int64_t)blockIdx.x * kRThreads * R + threadIdx.x; base - threadIdx.x < n_up; base += stride) {
    uint64_t img[R];
    uint32_t in = 0, keep = 0;
#pragma unroll
    for (int j = 0; j < R; ++j) { img[j] = 0; if (base + (int64_t)j * kRThreads < p.num_rows) in |= 1u << j; }
    keep = in;
#pragma unroll 1
    for (int c = 0; c < p.num_cols; ++c) {
      const RepackCol& col = p.col[c];
      uint64_t v[R];
      uint32_t valid = in;
#pragma unroll
      for (int j = 0; j < R; ++j) { v[j] = 0; if ((in >> j) & 1u) v[j] = load_raw(col.data, col.width, (int64_t)col.offset + base + (int64_t)j * kRThreads); }
      if (col.mask != nullptr) {
#pragma unroll
        for (int j = 0; j < R; ++j)
          if (((in >> j) & 1u) && !bit_is_set(col.mask, (int64_t)col.offset + base + (int64_t)j * kRThreads)) valid &= ~(1u << j);
      }
#pragma unroll
      for (int j = 0; j < R; ++j) {
        if ((valid >> j) & 1u) img[j] |= (repack_extend(col, v[j]) - col.base) << col.shift;
        else if (p.drop_null_rows) keep &= ~(1u << j);
        else if (col.null_bit >= 0) img[j] |= 1ull << col.null_bit;  // value bits of a null key stay zero: all nulls are equal
      }
    }
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const int64_t row = base + (int64_t)j * kRThreads;
      if ((in >> j) & 1u) p.out[row] = img[j];
      if (p.out_mask != nullptr && row - (threadIdx.x & 31) < p.num_rows) {  // warp-uniform: the warp's 32 rows start inside the column
        const uint32_t word = __ballot_sync(0xffffffffu, (keep >> j) & 1u);
        if ((threadIdx.x & 31) == 0) p.out_mask[row >> 5] = word;
      }
    }
  }
}

static int repack_grid(int64_t n) { const int64_t want = (n + kRThreads * 4 - 1) / (kRThreads * 4); return (int)(want < device_sm_count() * 16 ? want : device_sm_count() * 16); }

cudaError_t launch_repack_minmax(const RepackParams& p, cudaStream_t s)
{
  if (p.num_rows == 0) return cudaSuccess;
  const int64_t want = (p.num_rows + kRThreads * 8 - 1) / (kRThreads * 8);
  const dim3 grid((unsigned)(want < device_sm_count() * 8 ? want : device_sm_count() * 8), (unsigned)p.num_cols);
  repack_minmax_kernel<<<grid, kRThreads, 0, s>>>(p);
  return note_launch();
}

cudaError_t launch_repack_pack(const RepackParams& p, cudaStream_t s)
{
  if (p.num_rows == 0) return cudaSuccess;
  repack_pack_kernel<<<repack_grid(p.num_rows), kRThreads, 0, s>>>(p);
  return note_launch();
}

}  // namespace bgb

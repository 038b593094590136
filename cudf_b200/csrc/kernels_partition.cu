// Radix partitioning of the rows by the high bits of their table slot: the high-cardinality path.
//
// With 1e8 groups the group table (5-20 GB) is far larger than the 126 MB L2, and every row of the single-pass kernel
// costs three to five random 32-byte-sector round trips to HBM (measured 17-21 G rows/s, profiles/r01_microbench.txt).
// Partitioning first makes the table accesses of the aggregation pass L2-resident:
//   pass 1  partition_hist_kernel     keys only  -> rows per partition (shared-memory histogram, one RED per bin per CTA)
//   pass 2  partition_scan_kernel     exclusive scan of the P counters (one CTA)
//   pass 3  partition_scatter_kernel  keys + values -> packed key column, value columns, validity bytes in partition order.
//           Each CTA ranks a 2048-row tile by partition in shared memory (ATOMS rank + block scan), reserves one contiguous
//           run per partition with a single RED per (tile, partition), and writes the tile out in runs (coalesced).
//   pass 4  aggregate_global_kernel   (kernels_generic.cu) over the partitioned copy: consecutive rows hit one
//           table region of capacity/P slots, which stays in L2.
// Partition id = top bits of the slot index the aggregation pass will use (hash_key & (capacity-1)), so partition p owns
// the table region [p * capacity/P, (p+1) * capacity/P).
// The reference has no such pass on this path (its cuco::static_set is probed at random for every row,
// /root/reference/cpp/src/groupby/hash/compute_global_memory_aggs.cuh:123-157); its row partitioner for distributed
// callers is cudf::hash_partition (/root/reference/cpp/src/partitioning/partitioning.cu:119-151,569-663).
#include <cstdlib>

#include "device_utils.cuh"
#include "kernels_partition.h"

namespace bgb {

namespace {
constexpr int kPThreads = 512;   // two CTAs per SM (64 registers each): one loads/ranks while the other writes out

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

__device__ __forceinline__ uint32_t partition_of(const PartParams& p, uint64_t k0, uint64_t k1)
{
  const uint64_t h = hash_key(k0, k1, p.a.table.key_words);
  // owner-rank partitioning uses a second mix so that it is independent of the table slot (hash2(key) % P)
  if (p.mod_parts > 0) return (uint32_t)(mix64(h ^ 0xA0761D6478BD642Full) % (uint32_t)p.mod_parts);
  return (uint32_t)((h & (p.a.table.capacity - 1)) >> p.slot_shift);
}
// keep + partition of one row (the general kernels); murmur mode never packs the key
__device__ __forceinline__ bool row_partition(const PartParams& p, int64_t row, uint64_t& k0, uint64_t& k1, uint32_t& pid)
{
  if (p.cudf_hash) { k0 = k1 = 0; pid = cudf_row_hash(p.a, row, p.hash_seed, p.cudf_hash, p.signed_cols) % (uint32_t)p.mod_parts; return true; }
  if (!pack_key(p.a, row, k0, k1)) return false;
  pid = partition_of(p, k0, k1);
  return true;
}
}  // namespace

__global__ void __launch_bounds__(kPThreads) partition_hist_kernel(const __grid_constant__ PartParams p)
{
  extern __shared__ uint32_t s_hist[];
  const uint32_t P = 1u << p.log2_parts;
  for (uint32_t i = threadIdx.x; i < P; i += kPThreads) s_hist[i] = 0;
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * kPThreads;
  const int64_t n      = p.a.num_rows;
  for (int64_t base = (int64_t)blockIdx.x * kPThreads + threadIdx.x; base < n; base += 4 * stride) {
    // four independent rows per thread: their key loads are all in flight before the first hash
    uint64_t k0[4], k1[4];
    uint32_t pid[4];
    bool keep[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) { const int64_t row = base + j * stride; keep[j] = row < n && row_partition(p, row, k0[j], k1[j], pid[j]); }
#pragma unroll
    for (int j = 0; j < 4; ++j) if (keep[j]) atomicAdd(&s_hist[pid[j]], 1u);
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < P; i += kPThreads)
    if (s_hist[i]) atomicAdd(p.counts + i, s_hist[i]);
}

// offsets[i] = sum of counts[0..i), offsets[P] = rows kept; cursors start at the offsets
__global__ void __launch_bounds__(1024) partition_scan_kernel(const uint32_t* counts, uint32_t* offsets, uint32_t* cursors, uint32_t P)
{
  __shared__ uint32_t s_warp[32];
  __shared__ uint32_t s_carry;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < P; base += 1024) {
    const uint32_t i = base + threadIdx.x;
    const uint32_t v = i < P ? counts[i] : 0;
    uint32_t incl    = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, d); if ((threadIdx.x & 31) >= d) incl += t; }
    if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = incl;
    __syncthreads();
    uint32_t woff = 0, total = 0;
    for (int w = 0; w < 32; ++w) { uint32_t t = s_warp[w]; if (w < (int)(threadIdx.x >> 5)) woff += t; total += t; }
    const uint32_t excl = s_carry + woff + incl - v;
    if (i < P) { offsets[i] = excl; cursors[i] = excl; }
    __syncthreads();
    if (threadIdx.x == 0) s_carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) offsets[P] = s_carry;
}

template <int kPRows>
__global__ void __launch_bounds__(kPThreads, (kPRows <= 8 ? 2 : 1)) partition_scatter_kernel(const __grid_constant__ PartParams p)
{
  constexpr int kPTile = kPThreads * kPRows;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const uint32_t P = 1u << p.log2_parts;
  // layout: stage[kPTile] (8 B) | dst[kPTile] (4 B) | cnt[P] | off[P] | gbase[P]
  uint64_t* stage  = reinterpret_cast<uint64_t*>(smem_raw);
  uint32_t* dst    = reinterpret_cast<uint32_t*>(stage + kPTile);
  uint32_t* s_cnt  = dst + kPTile;
  uint32_t* s_off  = s_cnt + P;
  uint32_t* s_gbas = s_off + P;
  __shared__ uint32_t s_warp[32];
  __shared__ uint32_t s_total;

  const int64_t n      = p.a.num_rows;
  const int64_t ntiles = (n + kPTile - 1) / kPTile;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    for (uint32_t i = threadIdx.x; i < P; i += kPThreads) s_cnt[i] = 0;
    // pull this tile's value columns (consumed a few phases from now) and the next tile's key columns towards L2
#pragma unroll
    for (int j = 0; j < kPRows; ++j) {
      const int64_t row = tile * kPTile + (int64_t)j * kPThreads + threadIdx.x;
      if (row < n) for (int c = 0; c < p.num_vals; ++c) prefetch_l2(p.v[c].data + ((int64_t)p.v[c].offset + row) * p.v[c].width);
      const int64_t nrow = row + (int64_t)gridDim.x * kPTile;
      if (nrow < n) for (int c = 0; c < p.a.num_key_cols; ++c) prefetch_l2(p.a.kc[c].data + ((int64_t)p.a.kc[c].offset + nrow) * p.a.kc[c].width);
    }
    __syncthreads();
    // ---- rank every row inside its partition (ATOMS.ADD with return)
    uint64_t k0[kPRows], k1[kPRows];
    uint32_t pid[kPRows], rank[kPRows];
    bool keep[kPRows];
#pragma unroll
    for (int j = 0; j < kPRows; ++j) {
      const int64_t row = tile * kPTile + (int64_t)j * kPThreads + threadIdx.x;
      pid[j] = 0; rank[j] = 0;
      keep[j] = row < n && row_partition(p, row, k0[j], k1[j], pid[j]);
      if (keep[j]) rank[j] = atomicAdd(&s_cnt[pid[j]], 1u);
    }
    __syncthreads();
    // ---- exclusive scan of the P counters (P <= 4096: up to 4 bins per thread) + one global reservation per partition
    {
      uint32_t carry = 0;
      for (uint32_t base = 0; base < P; base += kPThreads) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < P ? s_cnt[i] : 0;
        uint32_t incl    = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, d); if ((threadIdx.x & 31) >= d) incl += t; }
        if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = incl;
        __syncthreads();
        uint32_t woff = 0, total = 0;
        for (int w = 0; w < kPThreads / 32; ++w) { uint32_t t = s_warp[w]; if (w < (int)(threadIdx.x >> 5)) woff += t; total += t; }
        if (i < P) {
          s_off[i] = carry + woff + incl - v;
          if (v) s_gbas[i] = atomicAdd(p.cursors + i, v);
        }
        carry += total;
        __syncthreads();
      }
      if (threadIdx.x == 0) s_total = carry;
    }
    __syncthreads();
    const uint32_t total = s_total;
    // ---- destination row of every staged position, then the packed keys, then each value column through `stage`
    uint32_t pos[kPRows];
#pragma unroll
    for (int j = 0; j < kPRows; ++j) {
      pos[j] = 0;
      if (keep[j]) { pos[j] = s_off[pid[j]] + rank[j]; dst[pos[j]] = s_gbas[pid[j]] + rank[j]; stage[pos[j]] = k0[j]; }
    }
    __syncthreads();
    if (p.out_k0 == nullptr) { /* owner-rank partitioning permutes the caller's columns only */ }
    else if (p.key_out_width == 8) { for (uint32_t i = threadIdx.x; i < total; i += kPThreads) p.out_k0[dst[i]] = stage[i]; }
    else { for (uint32_t i = threadIdx.x; i < total; i += kPThreads) reinterpret_cast<uint32_t*>(p.out_k0)[dst[i]] = (uint32_t)stage[i]; }
    if (p.a.table.key_words == 2 && p.out_k1 != nullptr) {
      __syncthreads();
#pragma unroll
      for (int j = 0; j < kPRows; ++j) if (keep[j]) stage[pos[j]] = k1[j];
      __syncthreads();
      for (uint32_t i = threadIdx.x; i < total; i += kPThreads) p.out_k1[dst[i]] = stage[i];
    }
    if (p.out_row != nullptr) {
      __syncthreads();
#pragma unroll
      for (int j = 0; j < kPRows; ++j) if (keep[j]) stage[pos[j]] = (uint64_t)(tile * kPTile + (int64_t)j * kPThreads + threadIdx.x);
      __syncthreads();
      for (uint32_t i = threadIdx.x; i < total; i += kPThreads) p.out_row[dst[i]] = (uint32_t)stage[i];
    }
    for (int c = 0; c < p.num_vals; ++c) {
      const PartValue& v = p.v[c];
      __syncthreads();
#pragma unroll
      for (int j = 0; j < kPRows; ++j) {
        if (!keep[j]) continue;
        const int64_t idx = (int64_t)v.offset + tile * kPTile + (int64_t)j * kPThreads + threadIdx.x;
        const bool valid  = v.mask == nullptr || bit_is_set(v.mask, idx);
        // bit 63 of the staged word cannot carry validity for 8-byte values, so validity travels in its own pass below
        stage[pos[j]] = valid ? load_raw(v.data, v.width, idx) : 0ull;
      }
      __syncthreads();
      switch (v.width) {
        case 1: for (uint32_t i = threadIdx.x; i < total; i += kPThreads) v.out_data[dst[i]] = (uint8_t)stage[i]; break;
        case 2: for (uint32_t i = threadIdx.x; i < total; i += kPThreads) reinterpret_cast<uint16_t*>(v.out_data)[dst[i]] = (uint16_t)stage[i]; break;
        case 4: for (uint32_t i = threadIdx.x; i < total; i += kPThreads) reinterpret_cast<uint32_t*>(v.out_data)[dst[i]] = (uint32_t)stage[i]; break;
        default: for (uint32_t i = threadIdx.x; i < total; i += kPThreads) reinterpret_cast<uint64_t*>(v.out_data)[dst[i]] = stage[i]; break;
      }
      if (v.mask != nullptr) {
        __syncthreads();
#pragma unroll
        for (int j = 0; j < kPRows; ++j) {
          if (!keep[j]) continue;
          const int64_t idx = (int64_t)v.offset + tile * kPTile + (int64_t)j * kPThreads + threadIdx.x;
          reinterpret_cast<uint8_t*>(stage)[pos[j]] = bit_is_set(v.mask, idx) ? 1 : 0;
        }
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < total; i += kPThreads) v.out_valid[dst[i]] = reinterpret_cast<uint8_t*>(stage)[i];
      }
    }
    __syncthreads();
  }
}

size_t partition_scatter_smem_bytes(int log2_parts, int rows) { return (size_t)kPThreads * rows * 12 + (size_t)(3u << log2_parts) * 4; }

// ---- lean scatter for the common shape: ONE integer-like key column without a mask (4 or 8 bytes, key image = the
// element), NV value columns of 8 bytes with optional masks. The general kernel above spends 280 instructions per row
// (ncu: 30 branches, 20 descriptor loads, local-memory arrays) and is issue-bound at 17 ms per 1e9 rows; this one keeps
// the tile loop branch-free for full tiles, uses 32-bit indices inside the tile and issues all loads of a phase together.
template <int KW, int NV, bool FULL>
__device__ __forceinline__ void scatter_tile_fast(const PartParams& p, int64_t tile_base, uint32_t rows_in_tile, uint64_t* stage, uint32_t* dst,
                                                  uint32_t* s_cnt, uint32_t* s_off, uint32_t* s_gbas, uint32_t* s_warp, uint32_t P)
{
  constexpr int R = 8;
  const uint32_t t = threadIdx.x;
  for (uint32_t i = t; i < P; i += kPThreads) s_cnt[i] = 0;
  // keys of this thread's 8 rows (consecutive threads read consecutive rows)
  const uint8_t* kdata = p.a.kc[0].data + ((int64_t)p.a.kc[0].offset + tile_base) * KW;
  uint64_t key[R];
#pragma unroll
  for (int j = 0; j < R; ++j) {
    const uint32_t r = j * kPThreads + t;
    key[j] = 0;
    if (FULL || r < rows_in_tile) key[j] = KW == 8 ? __ldg(reinterpret_cast<const unsigned long long*>(kdata) + r) : __ldg(reinterpret_cast<const uint32_t*>(kdata) + r);
  }
  __syncthreads();
  uint32_t pid[R], rank[R];
#pragma unroll
  for (int j = 0; j < R; ++j) {
    pid[j] = partition_of(p, key[j], 0);
    rank[j] = 0;
    if (FULL || j * kPThreads + t < rows_in_tile) rank[j] = atomicAdd(&s_cnt[pid[j]], 1u);
  }
  __syncthreads();
  {  // exclusive scan of the P counters + one global reservation per (tile, partition)
    uint32_t carry = 0;
    for (uint32_t base = 0; base < P; base += kPThreads) {
      const uint32_t i = base + t;
      const uint32_t v = i < P ? s_cnt[i] : 0;
      uint32_t incl    = v;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) { uint32_t u = __shfl_up_sync(0xffffffffu, incl, d); if ((t & 31) >= (uint32_t)d) incl += u; }
      if ((t & 31) == 31) s_warp[t >> 5] = incl;
      __syncthreads();
      uint32_t woff = 0, total = 0;
#pragma unroll
      for (int w = 0; w < kPThreads / 32; ++w) { uint32_t u = s_warp[w]; if (w < (int)(t >> 5)) woff += u; total += u; }
      if (i < P) {
        s_off[i] = carry + woff + incl - v;
        if (v) s_gbas[i] = atomicAdd(p.cursors + i, v);
      }
      carry += total;
      __syncthreads();
    }
  }
  uint32_t pos[R];
#pragma unroll
  for (int j = 0; j < R; ++j) {
    pos[j] = s_off[pid[j]] + rank[j];
    if (FULL || j * kPThreads + t < rows_in_tile) { dst[pos[j]] = s_gbas[pid[j]] + rank[j]; stage[pos[j]] = key[j]; }
  }
  __syncthreads();
  if (p.out_k0 != nullptr) {
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const uint32_t i = j * kPThreads + t;
      if (FULL || i < rows_in_tile) {
        if (p.key_out_width == 8) p.out_k0[dst[i]] = stage[i];
        else reinterpret_cast<uint32_t*>(p.out_k0)[dst[i]] = (uint32_t)stage[i];
      }
    }
  }
#pragma unroll
  for (int c = 0; c < NV; ++c) {
    const PartValue& v   = p.v[c];
    const int64_t vbase  = (int64_t)v.offset + tile_base;
    const unsigned long long* vdata = reinterpret_cast<const unsigned long long*>(v.data) + vbase;
    uint64_t val[R];
    uint32_t vbits = 0xffu;
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const uint32_t r = j * kPThreads + t;
      val[j] = 0;
      if (FULL || r < rows_in_tile) val[j] = __ldg(vdata + r);
    }
    if (v.mask != nullptr) {
      vbits = 0;
#pragma unroll
      for (int j = 0; j < R; ++j) {
        const uint32_t r = j * kPThreads + t;
        if (FULL || r < rows_in_tile) vbits |= (uint32_t)bit_is_set(v.mask, vbase + r) << j;
      }
    }
    __syncthreads();  // the previous column's (or the keys') read-out of `stage` is complete
#pragma unroll
    for (int j = 0; j < R; ++j)
      if (FULL || j * kPThreads + t < rows_in_tile) stage[pos[j]] = ((vbits >> j) & 1u) ? val[j] : 0ull;  // null elements are written as 0
    __syncthreads();
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const uint32_t i = j * kPThreads + t;
      if (FULL || i < rows_in_tile) reinterpret_cast<uint64_t*>(v.out_data)[dst[i]] = stage[i];
    }
    if (v.mask != nullptr) {
      __syncthreads();
      uint8_t* sbytes = reinterpret_cast<uint8_t*>(stage);
#pragma unroll
      for (int j = 0; j < R; ++j)
        if (FULL || j * kPThreads + t < rows_in_tile) sbytes[pos[j]] = (vbits >> j) & 1u;
      __syncthreads();
#pragma unroll
      for (int j = 0; j < R; ++j) {
        const uint32_t i = j * kPThreads + t;
        if (FULL || i < rows_in_tile) v.out_valid[dst[i]] = sbytes[i];
      }
    }
  }
  __syncthreads();
}

template <int KW, int NV>
__global__ void __launch_bounds__(kPThreads, 2) partition_scatter_fast_kernel(const __grid_constant__ PartParams p)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr uint32_t kTile = kPThreads * 8;
  const uint32_t P = 1u << p.log2_parts;
  uint64_t* stage  = reinterpret_cast<uint64_t*>(smem_raw);
  uint32_t* dst    = reinterpret_cast<uint32_t*>(stage + kTile);
  uint32_t* s_cnt  = dst + kTile;
  uint32_t* s_off  = s_cnt + P;
  uint32_t* s_gbas = s_off + P;
  __shared__ uint32_t s_warp[32];
  const int64_t n      = p.a.num_rows;
  const int64_t ntiles = (n + kTile - 1) / kTile;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t base = tile * kTile;
    if (base + kTile <= n) scatter_tile_fast<KW, NV, true>(p, base, kTile, stage, dst, s_cnt, s_off, s_gbas, s_warp, P);
    else scatter_tile_fast<KW, NV, false>(p, base, (uint32_t)(n - base), stage, dst, s_cnt, s_off, s_gbas, s_warp, P);
  }
}

static bool scatter_fast_eligible(const PartParams& p)
{
  if (p.cudf_hash || p.out_row) return false;
  if (p.a.num_key_cols != 1 || p.a.table.key_words != 1 || p.a.kc[0].mask != nullptr || p.a.kc[0].fclass != 0) return false;
  if (p.a.kc[0].width != 4 && p.a.kc[0].width != 8) return false;
  if (p.num_vals < 1 || p.num_vals > 2) return false;
  for (int c = 0; c < p.num_vals; ++c)
    if (p.v[c].width != 8) return false;
  return p.a.num_rows < (1ll << 31);
}

template <int KW, int NV>
static cudaError_t launch_scatter_fast(const PartParams& p, int num_sms, cudaStream_t s)
{
  const size_t smem = partition_scatter_smem_bytes(p.log2_parts, 8);
  cudaError_t e = allow_max_dynamic_smem(partition_scatter_fast_kernel<KW, NV>);
  if (e != cudaSuccess) return e;
  const int64_t tile   = (int64_t)kPThreads * 8;
  const int64_t ntiles = (p.a.num_rows + tile - 1) / tile;
  const int grid       = (int)(ntiles < 2ll * num_sms ? ntiles : 2ll * num_sms);
  partition_scatter_fast_kernel<KW, NV><<<grid, kPThreads, smem, s>>>(p);
  return note_launch();
}


cudaError_t launch_partition_hist(const PartParams& p, int num_sms, cudaStream_t s)
{
  if (p.a.num_rows == 0) return cudaSuccess;
  const int64_t want = (p.a.num_rows + kPThreads - 1) / kPThreads;
  const int grid     = (int)(want < 4 * num_sms ? want : 4 * num_sms);
  partition_hist_kernel<<<grid, kPThreads, sizeof(uint32_t) << p.log2_parts, s>>>(p);
  return note_launch();
}

cudaError_t launch_partition_scan(const PartParams& p, cudaStream_t s)
{
  partition_scan_kernel<<<1, 1024, 0, s>>>(p.counts, p.offsets, p.cursors, 1u << p.log2_parts);
  return note_launch();
}

template <int ROWS>
static cudaError_t launch_scatter_rows(const PartParams& p, int num_sms, cudaStream_t s)
{
  const size_t smem = partition_scatter_smem_bytes(p.log2_parts, ROWS);
  cudaError_t e = allow_max_dynamic_smem(partition_scatter_kernel<ROWS>);
  if (e != cudaSuccess) return e;
  const int64_t tile   = (int64_t)kPThreads * ROWS;
  const int64_t ntiles = (p.a.num_rows + tile - 1) / tile;
  const int per_sm     = ROWS <= 8 ? 2 : 1;
  const int grid       = (int)(ntiles < (int64_t)per_sm * num_sms ? ntiles : (int64_t)per_sm * num_sms);
  partition_scatter_kernel<ROWS><<<grid, kPThreads, smem, s>>>(p);
  return note_launch();
}

cudaError_t launch_partition_scatter(const PartParams& p, int num_sms, cudaStream_t s)
{
  if (p.a.num_rows == 0) return cudaSuccess;
  // rows per thread per tile: bigger tiles give longer contiguous runs per partition (tile / P rows each)
  static const int rows_env = getenv("BGB_PART_ROWS") ? atoi(getenv("BGB_PART_ROWS")) : 0;
  int rows = rows_env ? rows_env : 8;  // measured at 1e9 rows, 512 partitions: 4 rows 19.1 ms, 8 rows 16.9 ms, 16 rows (1 CTA/SM) 35 ms
  if (p.a.num_rows < (int64_t)kPThreads * 64 * num_sms) rows = 4;  // small inputs: keep every SM busy
  static const bool no_fast = getenv("BGB_PART_NO_FAST") != nullptr;
  if (!no_fast && scatter_fast_eligible(p)) {
    const int kw = p.a.kc[0].width;
    if (kw == 8) return p.num_vals == 1 ? launch_scatter_fast<8, 1>(p, num_sms, s) : launch_scatter_fast<8, 2>(p, num_sms, s);
    return p.num_vals == 1 ? launch_scatter_fast<4, 1>(p, num_sms, s) : launch_scatter_fast<4, 2>(p, num_sms, s);
  }
  if (rows >= 16) return launch_scatter_rows<16>(p, num_sms, s);
  if (rows >= 8) return launch_scatter_rows<8>(p, num_sms, s);
  return launch_scatter_rows<4>(p, num_sms, s);
}

}  // namespace bgb

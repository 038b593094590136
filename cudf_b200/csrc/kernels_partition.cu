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
#include "device_utils.cuh"
#include "kernels_partition.h"

namespace bgb {

namespace {
constexpr int kPThreads = 512;   // two CTAs per SM (64 registers each): one loads/ranks while the other writes out
constexpr int kPRows    = 4;
constexpr int kPTile    = kPThreads * kPRows;

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

__device__ __forceinline__ uint32_t partition_of(const PartParams& p, uint64_t k0, uint64_t k1)
{
  const uint64_t h = hash_key(k0, k1, p.a.table.key_words);
  // owner-rank partitioning uses a second mix so that it is independent of the table slot (hash2(key) % P)
  if (p.mod_parts > 0) return (uint32_t)(mix64(h ^ 0xA0761D6478BD642Full) % (uint32_t)p.mod_parts);
  return (uint32_t)((h & (p.a.table.capacity - 1)) >> p.slot_shift);
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
    bool keep[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) { const int64_t row = base + j * stride; keep[j] = row < n && pack_key(p.a, row, k0[j], k1[j]); }
#pragma unroll
    for (int j = 0; j < 4; ++j) if (keep[j]) atomicAdd(&s_hist[partition_of(p, k0[j], k1[j])], 1u);
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

__global__ void __launch_bounds__(kPThreads, 2) partition_scatter_kernel(const __grid_constant__ PartParams p)
{
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
      keep[j] = row < n && pack_key(p.a, row, k0[j], k1[j]);
      pid[j] = 0; rank[j] = 0;
      if (keep[j]) { pid[j] = partition_of(p, k0[j], k1[j]); rank[j] = atomicAdd(&s_cnt[pid[j]], 1u); }
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

size_t partition_scatter_smem_bytes(int log2_parts) { return (size_t)kPTile * 12 + (size_t)(3u << log2_parts) * 4; }

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

cudaError_t launch_partition_scatter(const PartParams& p, int num_sms, cudaStream_t s)
{
  if (p.a.num_rows == 0) return cudaSuccess;
  const size_t smem = partition_scatter_smem_bytes(p.log2_parts);
  cudaError_t e = cudaFuncSetAttribute(partition_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const int64_t ntiles = (p.a.num_rows + kPTile - 1) / kPTile;
  const int per_sm     = 2;
  const int grid       = (int)(ntiles < (int64_t)per_sm * num_sms ? ntiles : (int64_t)per_sm * num_sms);
  partition_scatter_kernel<<<grid, kPThreads, smem, s>>>(p);
  return note_launch();
}

}  // namespace bgb

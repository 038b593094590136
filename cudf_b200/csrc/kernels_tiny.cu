// Tiny-cardinality kernel (BASELINE config Q1: 4 groups, 6e8 rows, MEAN + VAR + COUNT of two float64 columns).
//
// With a handful of groups every atomic design collapses onto a few addresses (the generic HBM-table kernel needs
// 464 ms for Q1; the reference replicates its shared-memory accumulators for the same reason:
// /root/reference/cpp/src/groupby/hash/compute_shared_memory_aggs.cu:121-353). Here each THREAD owns a private copy of
// every group's accumulator row in shared memory, laid out [group][cell][thread] so that the 32 lanes of a warp always
// touch 32 different banks whatever groups their rows belong to: a row costs one LDS + one STS per accumulator, no
// atomics, no divergence. The packed key -> group id lookup is a scan of the CTA's few published group keys (broadcast
// LDS); a key seen for the first time goes through a 64-slot CAS dictionary that hands out group ids. A CTA that meets
// more keys than it has rows for raises status[3] and the host reruns the call on the general path.
// At the end the CTA folds its threads' copies (warp shuffles + shared memory) and adds one partial per (group, cell)
// to the HBM table with global atomics.
#include "device_utils.cuh"
#include "kernels_tiny.h"

namespace bgb {

namespace {
constexpr int kDictSlots = 64;
constexpr int kMaxGroups = 64;
constexpr int kTinyRows  = 4;  // rows per thread per iteration (their loads are issued together)

enum TinyClass { TC_INT = 0, TC_F64 = 1, TC_F32 = 2 };

template <int CLS>
__device__ __forceinline__ uint64_t tiny_add(uint64_t a, uint64_t b)
{
  if (CLS == TC_F64) return (uint64_t)__double_as_longlong(__longlong_as_double((long long)a) + __longlong_as_double((long long)b));
  if (CLS == TC_F32) return (uint64_t)__float_as_uint(__uint_as_float((uint32_t)a) + __uint_as_float((uint32_t)b));
  return a + b;  // integer sums wrap like the reference's int64 atomic_add
}
template <int CLS>
__device__ __forceinline__ uint64_t tiny_value(uint64_t raw, uint8_t src)
{
  if (CLS == TC_INT) return (uint64_t)raw_to_i64(raw, src);
  return raw;
}
template <int CLS>
__device__ __forceinline__ uint64_t tiny_square(uint64_t v)
{
  if (CLS == TC_F64) { const double x = __longlong_as_double((long long)v); return (uint64_t)__double_as_longlong(x * x); }
  if (CLS == TC_F32) { const float x = __uint_as_float((uint32_t)v); return (uint64_t)__float_as_uint(x * x); }
  return v * v;
}

// private read-modify-write of one column's cells for one valid element
template <int CLS, bool SQ, bool MM>
__device__ __forceinline__ void tiny_update(uint64_t* c, uint64_t raw, uint8_t src)
{
  constexpr int T = kTinyThreads;
  const uint64_t v = tiny_value<CLS>(raw, src);
  const uint64_t n = c[0], s = c[T];
  uint64_t q = 0, lo = 0, hi = 0;
  if (SQ) q = c[2 * T];
  if (MM) { lo = c[(SQ ? 3 : 2) * T]; hi = c[(SQ ? 4 : 3) * T]; }
  c[0] = n + 1;
  c[T] = tiny_add<CLS>(s, v);
  if (SQ) c[2 * T] = tiny_add<CLS>(q, tiny_square<CLS>(v));
  if (MM) {
    const uint64_t img = ordered_image(raw, src);
    c[(SQ ? 3 : 2) * T] = img < lo ? img : lo;
    c[(SQ ? 4 : 3) * T] = img > hi ? img : hi;
  }
}

__device__ __forceinline__ int tiny_class(int sum_op) { return sum_op == ACC_SUM_F64 ? TC_F64 : sum_op == ACC_SUM_F32 ? TC_F32 : TC_INT; }

// one straight-line specialisation per (arithmetic class, cell set); kept out of line so the row loop stays small
// enough for the instruction cache (inlined into 4 rows x 4 columns it grew to 220 KB of SASS and ran 2x slower)
__device__ __noinline__ void tiny_apply(int sel, uint64_t* cc, uint64_t raw, uint8_t src)
{
  switch (sel) {
    case 0: tiny_update<TC_INT, false, false>(cc, raw, src); break;
    case 1: tiny_update<TC_INT, false, true>(cc, raw, src); break;
    case 2: tiny_update<TC_INT, true, false>(cc, raw, src); break;
    case 3: tiny_update<TC_INT, true, true>(cc, raw, src); break;
    case 4: tiny_update<TC_F64, false, false>(cc, raw, src); break;
    case 5: tiny_update<TC_F64, false, true>(cc, raw, src); break;
    case 6: tiny_update<TC_F64, true, false>(cc, raw, src); break;
    case 7: tiny_update<TC_F64, true, true>(cc, raw, src); break;
    case 8: tiny_update<TC_F32, false, false>(cc, raw, src); break;
    case 9: tiny_update<TC_F32, false, true>(cc, raw, src); break;
    case 10: tiny_update<TC_F32, true, false>(cc, raw, src); break;
    default: tiny_update<TC_F32, true, true>(cc, raw, src); break;
  }
}

// first sight of key k in this CTA: the CAS dictionary makes exactly one thread its owner, who takes the next group id
// and publishes the key in s_group_key[] (where the per-row scan finds it). Returns the id, or -2 when the CTA is full.
__device__ __noinline__ int tiny_insert_key(uint64_t k, unsigned long long* dict_key, int* dict_gid, unsigned long long* group_key,
                                            int* groups, int cap)
{
  uint32_t slot = (uint32_t)(mix64(k) >> 58);
  for (int moved = 0; moved < kDictSlots;) {  // `moved` counts slot advances only; waiting on an owner does not use them up
    const unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(&dict_key[slot]);
    if (cur == k) {
      const int g = *reinterpret_cast<volatile int*>(&dict_gid[slot]);
      if (g == -1) continue;  // the owner publishes right after its CAS: look again
      return g;
    }
    if (cur == kEmpty) {
      const unsigned long long old = atomicCAS(&dict_key[slot], (unsigned long long)kEmpty, (unsigned long long)k);
      if (old == kEmpty) {
        const int g = atomicAdd(groups, 1);
        if (g < cap) *reinterpret_cast<volatile unsigned long long*>(&group_key[g]) = k;
        __threadfence_block();
        *reinterpret_cast<volatile int*>(&dict_gid[slot]) = g < cap ? g : -2;
        return g < cap ? g : -2;
      }
      if (old == k) continue;
    }
    slot = (slot + 1) & (kDictSlots - 1);
    ++moved;
  }
  return -2;
}
}  // namespace

size_t tiny_smem_bytes(int cells, int group_cap) { return (size_t)cells * group_cap * kTinyThreads * 8; }

__global__ void __launch_bounds__(kTinyThreads, 1) agg_tiny_kernel(const __grid_constant__ TinyParams p)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint64_t* cells = reinterpret_cast<uint64_t*>(smem_raw);  // [group][cell][thread]
  __shared__ unsigned long long s_dict_key[kDictSlots];
  __shared__ int s_dict_gid[kDictSlots];
  __shared__ unsigned long long s_group_key[kMaxGroups];
  __shared__ int s_groups;
  __shared__ int s_abort;
  __shared__ uint64_t s_red[kTinyThreads / 32];

  constexpr int T = kTinyThreads;
  const int C = p.cells, GC = p.group_cap;
  for (int i = threadIdx.x; i < kDictSlots; i += T) { s_dict_key[i] = kEmpty; s_dict_gid[i] = -1; }
  for (int i = threadIdx.x; i < kMaxGroups; i += T) s_group_key[i] = kEmpty;
  if (threadIdx.x == 0) { s_groups = 0; s_abort = 0; }
  for (int g = 0; g < GC; ++g)
    for (int c = 0; c < p.num_cols; ++c) {
      const TinyCol& col = p.cols[c];
      uint64_t* cc = cells + ((size_t)g * C + col.cell_base) * T + threadIdx.x;
      const int nc = 2 + (col.need_sq ? 1 : 0) + (col.need_minmax ? 2 : 0);
      for (int i = 0; i < nc; ++i) cc[(size_t)i * T] = 0;
      if (col.need_minmax) cc[(size_t)(nc - 2) * T] = ~0ull;  // MIN identity; MAX starts at the smallest ordered image (0)
    }
  __syncthreads();

  const int64_t n      = p.a.num_rows;
  const int64_t stride = (int64_t)gridDim.x * T;
  for (int64_t base = (int64_t)blockIdx.x * T + threadIdx.x; base < n; base += kTinyRows * stride) {
    // ---- loads of kTinyRows independent rows first (keys, values, validity), then the shared-memory work
    uint64_t k0[kTinyRows], raw[kTinyRows][kTinyMaxCols];
    uint32_t valid[kTinyRows];  // bit c: value column c valid; bit 31: row present with a usable key
#pragma unroll
    for (int j = 0; j < kTinyRows; ++j) {
      const int64_t row = base + j * stride;
      uint64_t k1;
      valid[j] = 0;
      k0[j]    = 0;
      if (row < n && pack_key(p.a, row, k0[j], k1)) valid[j] = 1u << 31;
#pragma unroll
      for (int c = 0; c < kTinyMaxCols; ++c) {
        raw[j][c] = 0;
        if (c < p.num_cols && (valid[j] >> 31)) {
          const TinyCol& col = p.cols[c];
          const int64_t idx  = (int64_t)col.offset + row;
          if (col.mask == nullptr || bit_is_set(col.mask, idx)) valid[j] |= 1u << c;
          if (col.data != nullptr) raw[j][c] = load_raw(col.data, col.width, idx);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < kTinyRows; ++j) {
      if (!(valid[j] >> 31)) continue;
      // ---- packed key -> group id: scan the published group keys (all lanes read the same words: broadcast)
      int gid = -1;
      if (k0[j] != kEmpty) {
        for (int g = 0; g < GC; ++g)  // fixed trip count, no early exit: the warp stays converged
          gid = *reinterpret_cast<volatile unsigned long long*>(&s_group_key[g]) == k0[j] ? g : gid;
        if (gid < 0) gid = tiny_insert_key(k0[j], s_dict_key, s_dict_gid, s_group_key, &s_groups, GC);
      }
      __syncwarp(__activemask());
      if (gid < 0) { s_abort = 1; continue; }  // more keys than rows, or the EMPTY image: the host takes the general path
      uint64_t* row_cells = cells + (size_t)gid * C * T + threadIdx.x;
#pragma unroll
      for (int c = 0; c < kTinyMaxCols; ++c) {
        if (c >= p.num_cols) break;
        if (!((valid[j] >> c) & 1u)) continue;  // null source skipped (device_aggregators.cuh:434-436)
        const TinyCol& col = p.cols[c];
        uint64_t* cc = row_cells + (size_t)col.cell_base * T;
        tiny_apply(tiny_class(col.sum_op) * 4 + (col.need_sq ? 2 : 0) + (col.need_minmax ? 1 : 0), cc, raw[j][c], col.src);
      }
    }
  }
  __syncthreads();
  if (s_abort) { if (threadIdx.x == 0) p.a.table.status[3] = 1; return; }

  // ---- fold the threads' copies: warp shuffle, then across warps through shared memory; one global atomic per cell
  const int groups = s_groups < GC ? s_groups : GC;
  for (int g = 0; g < groups; ++g) {
    uint64_t gslot = 0;
    if (threadIdx.x == 0) gslot = find_or_insert(p.a.table, s_group_key[g], 0);
    for (int c = 0; c < p.num_cols; ++c) {
      const TinyCol& col = p.cols[c];
      const int nc  = 2 + (col.need_sq ? 1 : 0) + (col.need_minmax ? 2 : 0);
      const int cls = tiny_class(col.sum_op);
      for (int i = 0; i < nc; ++i) {
        // cell kinds: 0 count, 1 sum, 2 sumsq (if need_sq), then min, max
        const int kind = i < 2 ? i : (col.need_sq && i == 2) ? 2 : (i == nc - 2 ? 3 : 4);
        uint64_t v = cells[((size_t)g * C + col.cell_base + i) * T + threadIdx.x];
        auto merge = [&](uint64_t a, uint64_t b) -> uint64_t {
          if (kind == 0) return a + b;
          if (kind == 3) return a < b ? a : b;
          if (kind == 4) return a > b ? a : b;
          return cls == TC_F64 ? tiny_add<TC_F64>(a, b) : cls == TC_F32 ? tiny_add<TC_F32>(a, b) : a + b;
        };
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) v = merge(v, __shfl_xor_sync(0xffffffffu, v, d));
        if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
        __syncthreads();
        if (threadIdx.x == 0 && gslot != kNoSlot) {
          uint64_t t = s_red[0];
          for (int w = 1; w < T / 32; ++w) t = merge(t, s_red[w]);
          const int acc = kind == 0 ? col.acc_count : kind == 1 ? col.acc_sum : kind == 2 ? col.acc_sumsq : kind == 3 ? col.acc_min : col.acc_max;
          if (acc >= 0) {  // (identities are no-ops: MIN ~0, MAX 0, sums 0)
            void* gc = p.a.acc[acc].cells;
            if (kind == 0) atomicAdd(reinterpret_cast<int*>(gc) + gslot, (int)t);
            else if (kind == 3) atomicMin(reinterpret_cast<unsigned long long*>(gc) + gslot, (unsigned long long)t);
            else if (kind == 4) atomicMax(reinterpret_cast<unsigned long long*>(gc) + gslot, (unsigned long long)t);
            else if (cls == TC_F64) atomicAdd(reinterpret_cast<double*>(gc) + gslot, __longlong_as_double((long long)t));
            else if (cls == TC_F32) atomicAdd(reinterpret_cast<float*>(gc) + gslot, __uint_as_float((uint32_t)t));
            else atomicAdd(reinterpret_cast<unsigned long long*>(gc) + gslot, (unsigned long long)t);
          }
        }
        __syncthreads();
      }
    }
    if (threadIdx.x == 0 && gslot == kNoSlot) p.a.table.status[0] = 1;
  }
}

cudaError_t launch_aggregate_tiny(const TinyParams& p, int num_sms, cudaStream_t s)
{
  if (p.a.num_rows == 0) return cudaSuccess;
  const size_t smem = tiny_smem_bytes(p.cells, p.group_cap);
  cudaError_t e = allow_max_dynamic_smem(agg_tiny_kernel);
  if (e != cudaSuccess) return e;
  const int64_t want = (p.a.num_rows + kTinyThreads * kTinyRows - 1) / (kTinyThreads * kTinyRows);
  const int grid     = (int)(want < num_sms ? want : num_sms);
  agg_tiny_kernel<<<grid, kTinyThreads, smem, s>>>(p);
  return note_launch();
}

}  // namespace bgb

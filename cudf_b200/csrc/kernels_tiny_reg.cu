// Register-resident variant of the tiny-cardinality kernel (BASELINE config Q1: 4 groups, 6e8 rows, two int8 keys,
// MEAN + VAR + COUNT of two float64 columns = 18 B/row, 1.7 ms at the HBM roofline).
//
// With at most 8 groups every accumulator of every group fits in REGISTERS: a row is "compare its packed key with the
// G cached group keys, then G predicated adds per accumulator" -- no shared memory, no atomics, no indexed access on the
// row path. A warp reads 64 * kPairs consecutive rows per iteration as fully coalesced vectors: lane l owns rows
// {64 j + 2 l, 64 j + 2 l + 1}, j < kPairs, so a float64 column is kPairs LDG.128 per lane and an int8 key column kPairs
// LDG.U16. Group keys are learnt through the same CAS dictionary as kernels_tiny.cu and then cached in registers;
// a CTA that meets more than G keys raises status[3] and the host takes another path.
// Reference counterpart: the replicated shared-memory accumulators for low cardinalities,
// /root/reference/cpp/src/groupby/hash/compute_shared_memory_aggs.cu:260-353.
#include <algorithm>

#include "device_utils.cuh"
#include "kernels_tiny.h"

namespace bgb {
namespace {

constexpr int kRT        = 128;  // threads per CTA (111-183 registers per thread: 2-4 CTAs per SM)
constexpr int kPairs     = 2;    // row pairs per lane per iteration (warp tile = 128 rows). Measured on Q1 (r02, mask-free specialisation): 2 pairs 2.81 ms (120 registers), 4 pairs 3.11 ms (144)
constexpr int kWarpTile  = 64 * kPairs;
constexpr int kDictSlots = 64;

struct RegParams {
  AggParams a;
  const uint8_t* vdata[2];
  const uint32_t* vmask[2];
  int32_t voffset[2];
  int32_t acc_sum[2], acc_sumsq[2], acc_count[2];
  int32_t acc_count_all;
};

__device__ __noinline__ int reg_insert_key(uint64_t k, unsigned long long* dict_key, int* dict_gid, unsigned long long* group_key, int* groups, int cap)
{
  uint32_t slot = (uint32_t)(mix64(k) >> 58);
  for (int moved = 0; moved < kDictSlots;) {
    const unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(&dict_key[slot]);
    if (cur == k) {
      const int g = *reinterpret_cast<volatile int*>(&dict_gid[slot]);
      if (g == -1) continue;  // the owner publishes right after its CAS
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

template <bool F64> struct Arith;
template <> struct Arith<true> {
  using T = double;
  static __device__ __forceinline__ T from_bits(uint64_t b) { return __longlong_as_double((long long)b); }
  static __device__ __forceinline__ uint64_t to_bits(T v) { return (uint64_t)__double_as_longlong(v); }
};
template <> struct Arith<false> {
  using T = unsigned long long;  // wraps like the reference's int64 atomic add
  static __device__ __forceinline__ T from_bits(uint64_t b) { return b; }
  static __device__ __forceinline__ uint64_t to_bits(T v) { return v; }
};

// raw pair of W-byte elements as loaded (decoded later by split_pair): one 2W-byte vector, or element-wise at the tail
template <int W, bool FULL>
__device__ __forceinline__ void load_pair_raw(const uint8_t* data, int64_t idx, int64_t row, int64_t n, uint64_t (&r)[2])
{
  r[0] = r[1] = 0;
  if (FULL) {
    if (W == 1) r[0] = __ldg(reinterpret_cast<const uint16_t*>(data + idx));
    else if (W == 2) r[0] = __ldg(reinterpret_cast<const uint32_t*>(data + 2 * idx));
    else if (W == 4) { const uint2 v = __ldg(reinterpret_cast<const uint2*>(data + 4 * idx)); r[0] = (uint64_t)v.x | ((uint64_t)v.y << 32); }
    else {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(data + 8 * idx));
      r[0] = (uint64_t)v.x | ((uint64_t)v.y << 32);
      r[1] = (uint64_t)v.z | ((uint64_t)v.w << 32);
    }
  } else {
    uint64_t e0 = 0, e1 = 0;
    if (row < n) e0 = load_raw(data, W, idx);
    if (row + 1 < n) e1 = load_raw(data, W, idx + 1);
    if (W == 8) { r[0] = e0; r[1] = e1; }
    else r[0] = e0 | (e1 << (8 * (W & 7)));
  }
}
template <int W>
__device__ __forceinline__ void split_pair(const uint64_t (&r)[2], uint64_t& e0, uint64_t& e1)
{
  if (W == 8) { e0 = r[0]; e1 = r[1]; }
  else { e0 = r[0] & ((1ull << (8 * (W & 7))) - 1); e1 = r[0] >> (8 * (W & 7)); }
}

template <int NC, int W, int KCMAX, bool FULL, bool MASKS>
__device__ __forceinline__ void tile_loads(const RegParams& p, int64_t wb, int lane, int64_t n, uint64_t (&kr)[kPairs][KCMAX][2],
                                           uint32_t (&km)[kPairs][KCMAX], uint64_t (&val)[NC][kPairs][2], uint32_t (&vm)[NC][kPairs])
{
#pragma unroll
  for (int j = 0; j < kPairs; ++j) {
    const int64_t row = wb + 64 * j + 2 * lane;
#pragma unroll
    for (int c = 0; c < KCMAX; ++c) {
      kr[j][c][0] = kr[j][c][1] = 0;
      km[j][c]    = 0xffffffffu;
      if (c < p.a.num_key_cols) {
        const KeyColDesc& kc = p.a.kc[c];
        const int64_t idx    = (int64_t)kc.offset + row;
        load_pair_raw<W, FULL>(kc.data, idx, row, n, kr[j][c]);
        if (MASKS && kc.mask != nullptr && (FULL || row < n)) km[j][c] = __ldg(kc.mask + (idx >> 5)) >> (idx & 31);  // idx even: both bits in one word
      }
    }
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      vm[c][j]          = 0xffffffffu;
      const int64_t idx = (int64_t)p.voffset[c] + row;
      load_pair_raw<8, FULL>(p.vdata[c], idx, row, n, val[c][j]);
      if (MASKS && p.vmask[c] != nullptr && (FULL || row < n)) vm[c][j] = __ldg(p.vmask[c] + (idx >> 5)) >> (idx & 31);
    }
  }
}

// A full tile of the mask-free shape with a key image of fewer than 32 bits in NK columns of W bytes (BASELINE configs[3]: two int8
// keys): ~55 instructions per row instead of ~105 for the general body below (ncu r02: per-column uniform predicates for up to
// four key columns, 64-bit image assembly, 64-bit row-range tests and the gid select chain were half of the instruction stream).
// A row's group is found as G 32-bit compares that ARE the select predicates of the accumulate step; a row that hits no cached
// group key goes through the CAS dictionary (first rows of a CTA only).
template <int NC, int NK> struct LeanRaw {
  uint32_t k[kPairs][NK];
  uint4 v[NC][kPairs];
};
template <int NC, int W, int NK>
__device__ __forceinline__ void lean_load(const RegParams& p, int64_t wb, int lane, LeanRaw<NC, NK>& r)
{
  static_assert(W == 1 || W == 2, "pairs of 1- or 2-byte key elements");
#pragma unroll
  for (int j = 0; j < kPairs; ++j) {
    const int64_t row = wb + 64 * j + 2 * lane;
#pragma unroll
    for (int c = 0; c < NK; ++c) {
      const KeyColDesc& kc = p.a.kc[c];
      const int64_t idx    = (int64_t)kc.offset + row;
      r.k[j][c] = W == 1 ? (uint32_t)__ldg(reinterpret_cast<const uint16_t*>(kc.data + idx)) : __ldg(reinterpret_cast<const uint32_t*>(kc.data + 2 * idx));
    }
#pragma unroll
    for (int c = 0; c < NC; ++c) r.v[c][j] = __ldg(reinterpret_cast<const uint4*>(p.vdata[c] + 8 * ((int64_t)p.voffset[c] + row)));
  }
}
template <int G, int NC, bool SQ, bool F64, int W, int NK>
__device__ __forceinline__ void lean_compute(const RegParams& p, const LeanRaw<NC, NK>& r, unsigned long long (&gk)[G],
                                             typename Arith<F64>::T (&sum)[G][NC], typename Arith<F64>::T (&sq)[G][NC], int (&cnt_all)[G],
                                             bool& abort_seen, unsigned long long* s_dict_key, int* s_dict_gid, unsigned long long* s_group_key,
                                             int* s_groups)
{
  using A = Arith<F64>;
  using T = typename A::T;
  uint32_t gk32[G];
#pragma unroll
  for (int g = 0; g < G; ++g) gk32[g] = (uint32_t)gk[g];  // (unused entries are all ones: no image of fewer than 32 bits equals them)
  uint32_t img[kPairs][2];
  bool miss = false;
#pragma unroll
  for (int j = 0; j < kPairs; ++j) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      uint32_t m = 0;
#pragma unroll
      for (int c = 0; c < NK; ++c) m |= ((r.k[j][c] >> (8 * W * h)) & ((1u << (8 * W)) - 1)) << p.a.kc[c].shift;
      img[j][h] = m;
      bool any = false;
#pragma unroll
      for (int g = 0; g < G; ++g) any |= m == gk32[g];
      miss |= !any;
    }
  }
  // Warp-uniform split. Steady state: every row of the tile hits a cached group key, and the G compares of a row ARE the select
  // predicates of its accumulate step. A tile with a new key (first rows of a CTA) takes the second body, where a learning branch
  // assigns the group: there the compiler keeps the hits as 0/1 integers (+8 instructions per row), which is why it is kept apart.
  if (!__any_sync(0xffffffffu, miss)) {
#pragma unroll
    for (int j = 0; j < kPairs; ++j) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        bool hit[G];
#pragma unroll
        for (int g = 0; g < G; ++g) hit[g] = img[j][h] == gk32[g];
#pragma unroll
        for (int c = 0; c < NC; ++c) {
          const uint4 q = r.v[c][j];
          const T v = A::from_bits(h == 0 ? ((uint64_t)q.x | ((uint64_t)q.y << 32)) : ((uint64_t)q.z | ((uint64_t)q.w << 32)));
#pragma unroll
          for (int g = 0; g < G; ++g) {
            const T vg = hit[g] ? v : (T)0;  // select on the INPUT, unconditional adds (see the general body)
            sum[g][c] += vg;
            if (SQ) sq[g][c] += vg * vg;
          }
        }
#pragma unroll
        for (int g = 0; g < G; ++g) cnt_all[g] += hit[g] ? 1 : 0;
      }
    }
    return;
  }
#pragma unroll
  for (int j = 0; j < kPairs; ++j) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      int gid = -1;
#pragma unroll
      for (int g = 0; g < G; ++g) gid = img[j][h] == gk32[g] ? g : gid;
      if (gid < 0) {  // learn the key, refresh the cached group keys
        gid = reg_insert_key((uint64_t)img[j][h], s_dict_key, s_dict_gid, s_group_key, s_groups, G);
        if (gid < 0) abort_seen = true;
#pragma unroll
        for (int g = 0; g < G; ++g) {
          gk[g]   = *reinterpret_cast<volatile unsigned long long*>(&s_group_key[g]);
          gk32[g] = (uint32_t)gk[g];
        }
      }
#pragma unroll
      for (int c = 0; c < NC; ++c) {
        const uint4 q = r.v[c][j];
        const T v = A::from_bits(h == 0 ? ((uint64_t)q.x | ((uint64_t)q.y << 32)) : ((uint64_t)q.z | ((uint64_t)q.w << 32)));
#pragma unroll
        for (int g = 0; g < G; ++g) {
          const T vg = gid == g ? v : (T)0;
          sum[g][c] += vg;
          if (SQ) sq[g][c] += vg * vg;
        }
      }
#pragma unroll
      for (int g = 0; g < G; ++g) cnt_all[g] += gid == g ? 1 : 0;
    }
  }
}

// MASKS: some key or value column carries a null mask (else every validity test, null counter and zeroing select is compiled
// out: they were ~40 % of the instructions of the mask-free Q1 shape). K32: the packed key image fits 32 bits (one compare per
// cached group key instead of two).
// NK > 0 (implies !MASKS, K32, W <= 2): exactly NK key columns -- full tiles run lean_tile, only the last tile the general body.
template <int G, int NC, bool SQ, bool F64, int W, bool MASKS, bool K32, int NK>
__global__ void __launch_bounds__(kRT) agg_tiny_reg_kernel(const __grid_constant__ RegParams p)
{
  constexpr int KCMAX = W == 8 ? 1 : W == 4 ? 2 : 4;  // the packed key image is at most 64 bits
  using A = Arith<F64>;
  using T = typename A::T;
  __shared__ unsigned long long s_dict_key[kDictSlots];
  __shared__ int s_dict_gid[kDictSlots];
  __shared__ unsigned long long s_group_key[G];
  __shared__ int s_groups, s_abort;
  __shared__ unsigned long long s_acc[G][NC][2];  // block totals: [sum, sumsq] bit patterns (added with a CAS loop)
  __shared__ int s_cnt[G][NC + 1];                // valid counts per column, [NC] = all rows

  for (int i = threadIdx.x; i < kDictSlots; i += kRT) { s_dict_key[i] = kEmpty; s_dict_gid[i] = -1; }
  if (threadIdx.x < G) s_group_key[threadIdx.x] = kEmpty;
  for (int i = threadIdx.x; i < G * NC * 2; i += kRT) (&s_acc[0][0][0])[i] = A::to_bits((T)0);
  for (int i = threadIdx.x; i < G * (NC + 1); i += kRT) (&s_cnt[0][0])[i] = 0;
  if (threadIdx.x == 0) { s_groups = 0; s_abort = 0; }
  __syncthreads();

  T sum[G][NC], sq[G][NC];
  int nulls[G][NC], cnt_all[G];  // valid count of a column = rows of the group - its nulls
  unsigned long long gk[G];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    gk[g] = kEmpty; cnt_all[g] = 0;
#pragma unroll
    for (int c = 0; c < NC; ++c) { sum[g][c] = (T)0; sq[g][c] = (T)0; nulls[g][c] = 0; }
  }
  bool abort_seen = false;
  bool any_vmask  = false;
  if (MASKS) {
#pragma unroll
    for (int c = 0; c < NC; ++c) any_vmask |= p.vmask[c] != nullptr;
  }

  const int64_t n      = p.a.num_rows;
  const int lane       = threadIdx.x & 31;
  const int64_t warp0  = (int64_t)blockIdx.x * (kRT / 32) + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * (kRT / 32);
  int64_t wb_first = warp0 * kWarpTile;
  if constexpr (NK > 0 && W <= 2) {
    // full tiles, software-pipelined: the loads of the warp's NEXT tile are in flight while this one is accumulated (the kernel was
    // latency-bound at 16 warps per SM with one tile of loads per warp: 48 % of the cycles had no eligible warp)
    LeanRaw<NC, NK> cur, nxt;
    bool have = wb_first + kWarpTile <= n;
    if (have) lean_load<NC, W, NK>(p, wb_first, lane, cur);
    while (have) {
      const int64_t wn  = wb_first + nwarps * kWarpTile;
      const bool have_n = wn + kWarpTile <= n;
      if (have_n) lean_load<NC, W, NK>(p, wn, lane, nxt);
      lean_compute<G, NC, SQ, F64, W, NK>(p, cur, gk, sum, sq, cnt_all, abort_seen, s_dict_key, s_dict_gid, s_group_key, &s_groups);
      cur = nxt; wb_first = wn; have = have_n;
    }
  }
  for (int64_t wb = wb_first; wb < n; wb += nwarps * kWarpTile) {
    // ---- phase 1: every load of the tile is issued before anything is decoded, in ONE basic block (tile_loads): the
    //      earlier forms decoded each key right after its load, or kept each load behind its own uniform branch, and
    //      ncu showed eight serialised long-scoreboard stalls per tile (12-15 ms on Q1). Lane l owns the row pairs
    //      {wb + 64 j + 2 l, +1}; a pair of W-byte keys is one 2W-byte load, a pair of 8-byte values one LDG.128.
    uint64_t kr[kPairs][KCMAX][2], val[NC][kPairs][2];
    uint32_t km[kPairs][KCMAX], vm[NC][kPairs];
    if (wb + kWarpTile <= n) {  // warp-uniform; only the last tile takes the guarded element loads
      tile_loads<NC, W, KCMAX, true, MASKS>(p, wb, lane, n, kr, km, val, vm);
    } else {
      tile_loads<NC, W, KCMAX, false, MASKS>(p, wb, lane, n, kr, km, val, vm);
    }
    // ---- phase 2: packed key images and validity bits
    uint64_t img[kPairs][2];
    uint32_t ok[kPairs];  // bit 0/1: row usable (in range, key not dropped); bits 2c+2 / 2c+3: value column c valid
#pragma unroll
    for (int j = 0; j < kPairs; ++j) {
      const int64_t row = wb + 64 * j + 2 * lane;
      ok[j]     = row + 1 < n ? 3u : row < n ? 1u : 0u;
      img[j][0] = img[j][1] = 0;
#pragma unroll
      for (int c = 0; c < KCMAX; ++c) {
        if (c < p.a.num_key_cols) {
          const KeyColDesc& kc = p.a.kc[c];
          uint64_t e0, e1;
          split_pair<W>(kr[j][c], e0, e1);
          const uint32_t kv = MASKS ? (km[j][c] & 3u) : 3u;
          if (MASKS && kv != 3u) {
            if (p.a.skip_null_keys) ok[j] &= kv;  // EXCLUDE: the row is dropped
            else {                                 // INCLUDE: the null flag is part of the key, value bits stay zero
              if (!(kv & 1u)) { e0 = 0; img[j][0] |= 1ull << kc.null_bit; }
              if (!(kv & 2u)) { e1 = 0; img[j][1] |= 1ull << kc.null_bit; }
            }
          }
          img[j][0] |= e0 << kc.shift;
          img[j][1] |= e1 << kc.shift;
        }
      }
      if (MASKS) {
#pragma unroll
        for (int c = 0; c < NC; ++c) ok[j] |= (vm[c][j] & 3u) << (2 * c + 2);
      }
    }
    // ---- rows into the register accumulators
#pragma unroll
    for (int j = 0; j < kPairs; ++j) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const bool usable = (ok[j] >> h) & 1u;
        const unsigned long long k = img[j][h];
        int gid = -1;
#pragma unroll
        for (int g = 0; g < G; ++g) gid = (K32 ? (uint32_t)gk[g] == (uint32_t)k : gk[g] == k) ? g : gid;  // (K32: unused gk[] are all ones in the low word too)
        if (k == kEmpty) gid = -1;  // the all-ones image equals the 'unused' marker of gk[]: never a cached group (-> abort below)
        if (usable && gid < 0) {  // first rows only: learn the key, then refresh the cached group keys
          gid = (k == kEmpty) ? -2 : reg_insert_key(k, s_dict_key, s_dict_gid, s_group_key, &s_groups, G);
          if (gid < 0) abort_seen = true;
#pragma unroll
          for (int g = 0; g < G; ++g) gk[g] = *reinterpret_cast<volatile unsigned long long*>(&s_group_key[g]);
        }
        if (!usable) gid = -1;
        // G x NC updates with the select on the INPUT: vg = (gid == g) ? v : 0, then unconditional adds. Adding +0.0 is
        // exact here (a running sum that starts at +0.0 is never -0.0, and x + 0.0 == x for every other x incl. NaN/inf).
        // Earlier forms, for the record (ncu, Q1): `if`s became 582 M divergent branches (21 ms); predicated PTX adds were
        // lowered by ptxas to add + 2 FSEL on every RESULT word, 189 instructions per row (12.4 ms).
        T vz[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) {
          vz[c] = A::from_bits(val[c][j][h]);
          if (MASKS && any_vmask && !((ok[j] >> (2 * c + 2 + h)) & 1u)) vz[c] = (T)0;  // null source skipped (device_aggregators.cuh:434-436)
        }
#pragma unroll
        for (int g = 0; g < G; ++g) {
          const bool hit = gid == g;
          cnt_all[g] += hit ? 1 : 0;
#pragma unroll
          for (int c = 0; c < NC; ++c) {
            const T vg = hit ? vz[c] : (T)0;
            sum[g][c] += vg;
            if (SQ) sq[g][c] += vg * vg;
          }
        }
        if (MASKS && any_vmask) {  // warp-uniform: columns without a mask never pay for null counting
#pragma unroll
          for (int c = 0; c < NC; ++c) {
            const bool is_null = !((ok[j] >> (2 * c + 2 + h)) & 1u);
#pragma unroll
            for (int g = 0; g < G; ++g) nulls[g][c] += (is_null && gid == g) ? 1 : 0;
          }
        }
      }
    }
    __syncwarp();
  }
  if (abort_seen) s_abort = 1;
  __syncthreads();
  if (s_abort) { if (threadIdx.x == 0) p.a.table.status[3] = 1; return; }

  // ---- group ids are CTA-wide (every thread cached the same s_group_key order): fold lanes, then warps, then one global
  //      atomic per (CTA, group, accumulator)
#pragma unroll
  for (int g = 0; g < G; ++g) {
    int ca = cnt_all[g];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) ca += __shfl_xor_sync(0xffffffffu, ca, d);
    if (lane == 0 && ca) atomicAdd(&s_cnt[g][NC], ca);
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      T s = sum[g][c], q = sq[g][c];
      int m = cnt_all[g] - nulls[g][c];
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) {
        s += A::from_bits(__shfl_xor_sync(0xffffffffu, (unsigned long long)A::to_bits(s), d));
        if (SQ) q += A::from_bits(__shfl_xor_sync(0xffffffffu, (unsigned long long)A::to_bits(q), d));
        m += __shfl_xor_sync(0xffffffffu, m, d);
      }
      if (lane == 0 && m) {
        atomicAdd(&s_cnt[g][c], m);
        if (F64) {
          atomicAdd(reinterpret_cast<double*>(&s_acc[g][c][0]), __longlong_as_double((long long)A::to_bits(s)));
          if (SQ) atomicAdd(reinterpret_cast<double*>(&s_acc[g][c][1]), __longlong_as_double((long long)A::to_bits(q)));
        } else {
          atomicAdd(&s_acc[g][c][0], (unsigned long long)A::to_bits(s));
          if (SQ) atomicAdd(&s_acc[g][c][1], (unsigned long long)A::to_bits(q));
        }
      }
    }
  }
  __syncthreads();
  const int groups = s_groups < G ? s_groups : G;
  if (threadIdx.x < groups) {
    const int g = threadIdx.x;
    if (s_cnt[g][NC] > 0) {  // a group learnt by this CTA always has at least one row here
      const uint64_t slot = find_or_insert(p.a.table, s_group_key[g], 0);
      if (slot == kNoSlot) { p.a.table.status[0] = 1; return; }
      if (p.acc_count_all >= 0) atomicAdd(reinterpret_cast<int*>(p.a.acc[p.acc_count_all].cells) + slot, s_cnt[g][NC]);
      for (int c = 0; c < NC; ++c) {
        if (s_cnt[g][c] == 0) continue;
        if (p.acc_count[c] >= 0) atomicAdd(reinterpret_cast<int*>(p.a.acc[p.acc_count[c]].cells) + slot, s_cnt[g][c]);
        if (p.acc_sum[c] >= 0) {
          if (F64) atomicAdd(reinterpret_cast<double*>(p.a.acc[p.acc_sum[c]].cells) + slot, __longlong_as_double((long long)s_acc[g][c][0]));
          else atomicAdd(reinterpret_cast<unsigned long long*>(p.a.acc[p.acc_sum[c]].cells) + slot, s_acc[g][c][0]);
        }
        if (SQ && p.acc_sumsq[c] >= 0) {
          if (F64) atomicAdd(reinterpret_cast<double*>(p.a.acc[p.acc_sumsq[c]].cells) + slot, __longlong_as_double((long long)s_acc[g][c][1]));
          else atomicAdd(reinterpret_cast<unsigned long long*>(p.a.acc[p.acc_sumsq[c]].cells) + slot, s_acc[g][c][1]);
        }
      }
    }
  }
}

template <int G, int NC, bool SQ, bool F64, int W, bool MASKS, bool K32, int NK = 0>
cudaError_t launch_wmk(const RegParams& rp, int num_sms, cudaStream_t s)
{
  int per_sm = 1;
  cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, agg_tiny_reg_kernel<G, NC, SQ, F64, W, MASKS, K32, NK>, kRT, 0);
  if (e != cudaSuccess) return e;
  if (per_sm < 1) per_sm = 1;
  const int64_t tiles = (rp.a.num_rows + (int64_t)kWarpTile * (kRT / 32) - 1) / ((int64_t)kWarpTile * (kRT / 32));
  const int64_t grid  = tiles < (int64_t)num_sms * per_sm ? tiles : (int64_t)num_sms * per_sm;
  agg_tiny_reg_kernel<G, NC, SQ, F64, W, MASKS, K32, NK><<<(int)grid, kRT, 0, s>>>(rp);
  return note_launch();
}

template <int G, int NC, bool SQ, bool F64, int W>
cudaError_t launch_w(const RegParams& rp, int num_sms, cudaStream_t s)
{
  bool masks = false;
  int key_bits = 0;
  for (int c = 0; c < rp.a.num_key_cols; ++c) {
    masks |= rp.a.kc[c].mask != nullptr;
    key_bits = std::max(key_bits, std::max((int)rp.a.kc[c].shift + 8 * rp.a.kc[c].width, (int)rp.a.kc[c].null_bit + 1));
  }
  for (int c = 0; c < NC; ++c) masks |= rp.vmask[c] != nullptr;
  // (the all-ones image of a 32-bit-or-narrower key cannot occur when the image has fewer than 32 bits; at exactly 32 it could
  //  alias the 'unused' marker in the low word, so K32 wants strictly fewer)
  const bool k32 = W <= 2 && key_bits < 32;
  if (masks) return launch_wmk<G, NC, SQ, F64, W, true, false>(rp, num_sms, s);
  if constexpr (W == 1) {
    if (k32 && rp.a.num_key_cols == 1) return launch_wmk<G, NC, SQ, F64, W, false, true, 1>(rp, num_sms, s);
    if (k32 && rp.a.num_key_cols == 2) return launch_wmk<G, NC, SQ, F64, W, false, true, 2>(rp, num_sms, s);
    if (k32 && rp.a.num_key_cols == 3) return launch_wmk<G, NC, SQ, F64, W, false, true, 3>(rp, num_sms, s);
  }
  if constexpr (W == 2) {
    if (k32 && rp.a.num_key_cols == 1) return launch_wmk<G, NC, SQ, F64, W, false, true, 1>(rp, num_sms, s);
  }
  return launch_wmk<G, NC, SQ, F64, W, false, false>(rp, num_sms, s);
}

template <int G, int NC, bool SQ, bool F64>
cudaError_t launch_one(const RegParams& rp, int num_sms, cudaStream_t s)
{
  switch (rp.a.kc[0].width) {  // tiny_reg_eligible: every key column has this width
    case 1: return launch_w<G, NC, SQ, F64, 1>(rp, num_sms, s);
    case 2: return launch_w<G, NC, SQ, F64, 2>(rp, num_sms, s);
    case 4: return launch_w<G, NC, SQ, F64, 4>(rp, num_sms, s);
    default: return launch_w<G, NC, SQ, F64, 8>(rp, num_sms, s);
  }
}

template <int G, int NC>
cudaError_t launch_gn(const RegParams& rp, bool sq, bool f64, int num_sms, cudaStream_t s)
{
  if (sq) return f64 ? launch_one<G, NC, true, true>(rp, num_sms, s) : launch_one<G, NC, true, false>(rp, num_sms, s);
  return f64 ? launch_one<G, NC, false, true>(rp, num_sms, s) : launch_one<G, NC, false, false>(rp, num_sms, s);
}

bool aligned(const void* base, int64_t elem_offset, int width, int bytes)
{
  return ((uintptr_t)base + (uintptr_t)(elem_offset * width)) % (uintptr_t)bytes == 0;
}

}  // namespace

// which shapes the register kernel serves: 64-bit key image of integer-like columns of one width, one or two 8-byte value columns of
// one arithmetic class (float64, or int64/uint64), SUM / SUM_OF_SQUARES / COUNT only, pair-aligned column starts
bool tiny_reg_eligible(const TinyParams& tp)
{
  if (tp.a.table.key_words != 1) return false;
  const int w0 = tp.a.kc[0].width;
  if (tp.a.num_key_cols > (w0 == 8 ? 1 : w0 == 4 ? 2 : 4)) return false;
  for (int c = 0; c < tp.a.num_key_cols; ++c) {
    const KeyColDesc& k = tp.a.kc[c];
    if (k.width != w0) return false;  // the kernel is specialised on ONE key width (loads are issued without decoding)
    if (k.fclass != 0 || (k.offset & 1) || !aligned(k.data, k.offset, k.width, 2 * k.width)) return false;
  }
  int real = 0, cls = -1;
  for (int j = 0; j < tp.num_cols; ++j) {
    const TinyCol& col = tp.cols[j];
    if (col.need_minmax) return false;
    if (col.data == nullptr) continue;  // COUNT_ALL pseudo column
    if (++real > 2) return false;
    if (col.width != 8 || (col.offset & 1) || !aligned(col.data, col.offset, 8, 16)) return false;
    const int c = col.sum_op == ACC_SUM_F64 ? 1 : (col.src == S_I64 || col.src == S_U64) ? 0 : -2;
    if (c == -2 || (cls >= 0 && cls != c)) return false;
    cls = c;
  }
  return real >= 1;
}

int tiny_reg_split(const TinyParams& tp, TinyParams* out)
{
  if (tiny_reg_eligible(tp)) return 0;  // one launch does it
  int n = 0;
  bool used[kTinyPlanCols] = {};
  int pseudo = -1;
  for (int j = 0; j < tp.num_cols; ++j) if (tp.cols[j].data == nullptr) { pseudo = j; used[j] = true; }
  for (int j = 0; j < tp.num_cols; ++j) {
    if (used[j]) continue;
    TinyParams& sub = out[n];
    sub          = tp;
    sub.num_cols = 0;
    auto cls = [&](const TinyCol& c) { return c.sum_op == ACC_SUM_F64 ? 1 : (c.src == S_I64 || c.src == S_U64) ? 0 : -2; };
    sub.cols[sub.num_cols++] = tp.cols[j];
    used[j] = true;
    for (int k = j + 1; k < tp.num_cols; ++k)  // a partner of the same arithmetic class
      if (!used[k] && cls(tp.cols[k]) == cls(tp.cols[j])) { sub.cols[sub.num_cols++] = tp.cols[k]; used[k] = true; break; }
    if (n == 0 && pseudo >= 0) sub.cols[sub.num_cols++] = tp.cols[pseudo];  // COUNT_ALL rides with the first launch
    if (!tiny_reg_eligible(sub)) return 0;
    ++n;
  }
  if (n == 0 && pseudo >= 0) return 0;  // (COUNT_ALL alone: the register kernel wants a value column)
  return n;
}

cudaError_t launch_aggregate_tiny_reg(const TinyParams& tp, int group_cap, int num_sms, cudaStream_t s)
{
  if (tp.a.num_rows == 0) return cudaSuccess;
  RegParams rp;
  memset(&rp, 0, sizeof rp);
  rp.a = tp.a;
  rp.acc_count_all = -1;
  int nc = 0;
  bool sq = false, f64 = false;
  for (int j = 0; j < tp.num_cols; ++j) {
    const TinyCol& col = tp.cols[j];
    if (col.data == nullptr) { rp.acc_count_all = col.acc_count; continue; }
    rp.vdata[nc] = col.data; rp.vmask[nc] = col.mask; rp.voffset[nc] = col.offset;
    rp.acc_sum[nc] = col.acc_sum; rp.acc_sumsq[nc] = col.acc_sumsq; rp.acc_count[nc] = col.acc_count;
    sq |= col.need_sq != 0;
    f64 = col.sum_op == ACC_SUM_F64;
    ++nc;
  }
  if (group_cap <= 4) return nc == 1 ? launch_gn<4, 1>(rp, sq, f64, num_sms, s) : launch_gn<4, 2>(rp, sq, f64, num_sms, s);
  return nc == 1 ? launch_gn<8, 1>(rp, sq, f64, num_sms, s) : launch_gn<8, 2>(rp, sq, f64, num_sms, s);
}

}  // namespace bgb

// Bucketed high-cardinality path: radix-scatter the rows (one or two levels) until every bucket holds the rows of ONE
// shared-memory-sized SLICE of the group table, then aggregate each slice in shared memory and store it with coalesced
// writes. The HBM table is never initialised, probed or updated with atomics on this path:
//
//   bucket_scatter_kernel (level 1)   caller's columns -> packed rows in P1 buckets (by the top bits of the table slot)
//   bucket_scatter_kernel (level 2)   every level-1 bucket -> P2 sub-buckets (next bits); skipped when P1 slices suffice
//   bucket_aggregate_kernel           one CTA per bucket: slice of capacity/(P1*P2) slots built in shared memory
//                                     (ATOMS.CAS key probe + shared-memory accumulator updates), written out once
//   bucket_overflow_kernel            rows that found their fixed-capacity bucket full (skewed keys) through HBM atomics
//
// Buckets have a fixed capacity (expected rows + 12.5 % + slack) instead of exact offsets, which saves the histogram pass
// over the keys; a bucket that fills up diverts its rows to an overflow region that the last kernel folds into the same
// table (probing wraps inside a slice on this path, in both kernels). If the overflow region fills too, status[6] is set
// and the host reruns the call on the exact-offset partitioned path (kernels_partition.cu).
//
// Why: at 1e8 groups the table is 40x the L2. The L2-resident aggregation of kernels_partition.cu still pays one L2
// request per key probe and per accumulator per row (~3 requests, 19 ms per 1e9 rows at the ~210 G requests/s the L2
// serves); here a row costs shared-memory operations only and the pass streams at HBM rate whatever the number of
// accumulators. The reference probes a 2N-slot cuco::static_set in HBM for every row on this path
// (/root/reference/cpp/src/groupby/hash/compute_global_memory_aggs.cuh:123-157) and has no partitioned form.
#include "device_utils.cuh"
#include "kernels_bucket.h"

namespace bgb {

namespace {
constexpr int kBThreads = 512;
constexpr int kBCtasPerSm = 2;
constexpr int kBRows    = kBucketTile / kBThreads;  // 8
constexpr uint32_t kOvFail = 0xFFFFFFFFu;
}  // namespace

// MODE: 0 level 1 with the general key packer, 4 / 8 level 1 with one unmasked integer-like key column of that width,
//       2 level 2 (packed input)
template <int MODE>
__global__ void __launch_bounds__(kBThreads, kBCtasPerSm) bucket_scatter_kernel(const __grid_constant__ BucketScatterParams p)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const uint32_t P = 1u << p.log2_parts;
  // layout: stage[T] (8 B) | dst[T] (4 B) | cnt[P] | off[P] | gbase[P] | room[P] | ovb[P]
  uint64_t* stage  = reinterpret_cast<uint64_t*>(smem_raw);
  uint32_t* dst    = reinterpret_cast<uint32_t*>(stage + kBucketTile);
  uint32_t* s_cnt  = dst + kBucketTile;
  uint32_t* s_off  = s_cnt + P;
  uint32_t* s_gbas = s_off + P;
  uint32_t* s_room = s_gbas + P;
  uint32_t* s_ovb  = s_room + P;
  __shared__ uint32_t s_warp[kBThreads / 32];
  __shared__ uint32_t s_total, s_ovf;

  const uint32_t t        = threadIdx.x;
  const uint32_t capmask  = (uint32_t)(p.a.table.capacity - 1);
  const uint32_t pmask    = P - 1;
  const uint32_t tiles_per_seg = MODE == 2 ? (p.seg_capacity + kBucketTile - 1) / kBucketTile : 1;
  const int64_t ntiles    = MODE == 2 ? (int64_t)p.num_segs * tiles_per_seg : (p.a.num_rows + kBucketTile - 1) / kBucketTile;

  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    int64_t row0;
    uint32_t nrows, cbase = 0;
    if (MODE == 2) {
      const uint32_t seg = (uint32_t)(tile / tiles_per_seg), tt = (uint32_t)(tile % tiles_per_seg);
      uint32_t cnt = __ldg(p.seg_cursor + seg);
      cnt = cnt < p.seg_capacity ? cnt : p.seg_capacity;
      if (tt * kBucketTile >= cnt) continue;  // CTA-uniform
      row0  = (int64_t)seg * p.seg_capacity + (int64_t)tt * kBucketTile;
      nrows = cnt - tt * kBucketTile < kBucketTile ? cnt - tt * kBucketTile : kBucketTile;
      cbase = seg << p.log2_parts;
    } else {
      row0  = tile * kBucketTile;
      nrows = p.a.num_rows - row0 < (int64_t)kBucketTile ? (uint32_t)(p.a.num_rows - row0) : kBucketTile;
    }
    for (uint32_t i = t; i < P; i += kBThreads) s_cnt[i] = 0;
    if (t == 0) s_ovf = 0;

    // ---- keys (+ validity words) of this thread's rows; consecutive threads read consecutive rows
    uint64_t key[kBRows];
    uint32_t vb[kBRows];
    uint32_t keep = 0;
#pragma unroll
    for (int j = 0; j < kBRows; ++j) {
      const uint32_t r = j * kBThreads + t;
      key[j] = 0; vb[j] = 0;
      if (r < nrows) {
        const int64_t row = row0 + r;
        if (MODE == 2) {
          const uint64_t w = __ldg(reinterpret_cast<const unsigned long long*>(p.in_key) + row);
          if (p.key_in_low32) { key[j] = (uint32_t)w; vb[j] = (uint32_t)(w >> 32); } else key[j] = w;
          keep |= 1u << j;
        } else if (MODE == 8) {
          key[j] = __ldg(reinterpret_cast<const unsigned long long*>(p.a.kc[0].data) + p.a.kc[0].offset + row);
          keep |= 1u << j;
        } else if (MODE == 4) {
          key[j] = __ldg(reinterpret_cast<const uint32_t*>(p.a.kc[0].data) + p.a.kc[0].offset + row);
          keep |= 1u << j;
        } else {
          uint64_t k1;
          if (pack_key(p.a, row, key[j], k1)) keep |= 1u << j;
        }
      }
    }
    if (MODE == 2) {
      if (p.vbits_width == 1) {
#pragma unroll
        for (int j = 0; j < kBRows; ++j) if ((keep >> j) & 1u) vb[j] = __ldg(p.in_vbits + row0 + j * kBThreads + t);
      } else if (p.vbits_width == 4) {
#pragma unroll
        for (int j = 0; j < kBRows; ++j) if ((keep >> j) & 1u) vb[j] = __ldg(reinterpret_cast<const uint32_t*>(p.in_vbits) + row0 + j * kBThreads + t);
      }
    } else {
#pragma unroll 1
      for (int c = 0; c < p.num_vals; ++c) {
        const BucketValue& v = p.v[c];
        if (v.vbit < 0) continue;
#pragma unroll
        for (int j = 0; j < kBRows; ++j)
          if ((keep >> j) & 1u) vb[j] |= (uint32_t)bit_is_set(v.in_mask, (int64_t)v.in_offset + row0 + j * kBThreads + t) << v.vbit;
      }
    }
    __syncthreads();

    // ---- rank inside the partition (ATOMS.ADD with return)
    uint32_t pid[kBRows], rank[kBRows];
#pragma unroll
    for (int j = 0; j < kBRows; ++j) {
      pid[j]  = ((slot_hash32(key[j]) & capmask) >> p.slot_shift) & pmask;
      rank[j] = 0;
      if ((keep >> j) & 1u) rank[j] = atomicAdd(&s_cnt[pid[j]], 1u);
    }
    __syncthreads();

    // ---- exclusive scan of the P counters + one reservation per (tile, partition); a full bucket diverts to the overflow region
    {
      uint32_t carry = 0;
      for (uint32_t base = 0; base < P; base += kBThreads) {
        const uint32_t i = base + t;
        const uint32_t v = i < P ? s_cnt[i] : 0;
        uint32_t incl    = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { uint32_t u = __shfl_up_sync(0xffffffffu, incl, d); if ((t & 31) >= (uint32_t)d) incl += u; }
        if ((t & 31) == 31) s_warp[t >> 5] = incl;
        __syncthreads();
        uint32_t woff = 0, total = 0;
#pragma unroll
        for (int w = 0; w < kBThreads / 32; ++w) { uint32_t u = s_warp[w]; if (w < (int)(t >> 5)) woff += u; total += u; }
        if (i < P) {
          const uint32_t excl = carry + woff + incl - v;
          s_off[i] = excl;
          if (v) {
            const uint32_t g    = atomicAdd(p.cursor + cbase + i, v);
            const uint32_t room = g < p.bucket_capacity ? p.bucket_capacity - g : 0;
            s_gbas[i] = (cbase + i) * p.bucket_capacity + g - excl;  // destination row = staged position + this
            s_room[i] = excl + room;                                 // staged positions from here on do not fit the bucket
            if (v > room) {
              const uint32_t e  = v - room;
              const uint32_t ob = atomicAdd(p.ov_cursor, e);
              s_ovf = 1;
              if (ob > p.ov_capacity || e > p.ov_capacity - ob) { p.status[6] = 1; s_ovb[i] = kOvFail; }
              else s_ovb[i] = p.ov_base + ob;
            }
          }
        }
        carry += total;
        __syncthreads();
      }
      if (t == 0) s_total = carry;
    }
    __syncthreads();
    const uint32_t total = s_total;

    // ---- destination row of every staged position, then column by column through `stage`
    uint32_t pos[kBRows];
    const bool ovf = s_ovf != 0;  // CTA-uniform; only skewed inputs ever take the branch below
#pragma unroll
    for (int j = 0; j < kBRows; ++j) {
      pos[j] = 0;
      if ((keep >> j) & 1u) {
        pos[j] = s_off[pid[j]] + rank[j];
        uint32_t d = pos[j] + s_gbas[pid[j]];
        if (ovf) {
          const uint32_t lim = s_room[pid[j]];
          if (pos[j] >= lim) { const uint32_t ob = s_ovb[pid[j]]; d = ob == kOvFail ? p.ov_base + p.ov_capacity : ob + (pos[j] - lim); }
        }
        dst[pos[j]]   = d;
        stage[pos[j]] = p.key_in_low32 ? ((key[j] & 0xffffffffull) | ((uint64_t)vb[j] << 32)) : key[j];
      }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < kBRows; ++j) {
      const uint32_t i = j * kBThreads + t;
      if (i < total) p.out_key[dst[i]] = stage[i];
    }
    if (p.vbits_width != 0) {
      __syncthreads();
      uint32_t* s32 = reinterpret_cast<uint32_t*>(stage);
#pragma unroll
      for (int j = 0; j < kBRows; ++j) if ((keep >> j) & 1u) s32[pos[j]] = vb[j];
      __syncthreads();
      if (p.vbits_width == 1) {
#pragma unroll
        for (int j = 0; j < kBRows; ++j) { const uint32_t i = j * kBThreads + t; if (i < total) p.out_vbits[dst[i]] = (uint8_t)s32[i]; }
      } else {
#pragma unroll
        for (int j = 0; j < kBRows; ++j) { const uint32_t i = j * kBThreads + t; if (i < total) reinterpret_cast<uint32_t*>(p.out_vbits)[dst[i]] = s32[i]; }
      }
    }
#pragma unroll 1
    for (int c = 0; c < p.num_vals; ++c) {
      const BucketValue& v = p.v[c];
      const int64_t vbase  = (MODE == 2 ? 0 : (int64_t)v.in_offset) + row0;
      uint64_t val[kBRows];
      if (v.width == 8) {
#pragma unroll
        for (int j = 0; j < kBRows; ++j) { val[j] = 0; if ((keep >> j) & 1u) val[j] = __ldg(reinterpret_cast<const unsigned long long*>(v.in) + vbase + j * kBThreads + t); }
      } else if (v.width == 4) {
#pragma unroll
        for (int j = 0; j < kBRows; ++j) { val[j] = 0; if ((keep >> j) & 1u) val[j] = __ldg(reinterpret_cast<const uint32_t*>(v.in) + vbase + j * kBThreads + t); }
      } else {
#pragma unroll
        for (int j = 0; j < kBRows; ++j) { val[j] = 0; if ((keep >> j) & 1u) val[j] = load_raw(v.in, v.width, vbase + j * kBThreads + t); }
      }
      __syncthreads();  // the previous column's read-out of `stage` is complete
#pragma unroll
      for (int j = 0; j < kBRows; ++j) if ((keep >> j) & 1u) stage[pos[j]] = val[j];
      __syncthreads();
      switch (v.width) {
        case 8:
#pragma unroll
          for (int j = 0; j < kBRows; ++j) { const uint32_t i = j * kBThreads + t; if (i < total) reinterpret_cast<uint64_t*>(v.out)[dst[i]] = stage[i]; }
          break;
        case 4:
#pragma unroll
          for (int j = 0; j < kBRows; ++j) { const uint32_t i = j * kBThreads + t; if (i < total) reinterpret_cast<uint32_t*>(v.out)[dst[i]] = (uint32_t)stage[i]; }
          break;
        case 2:
#pragma unroll
          for (int j = 0; j < kBRows; ++j) { const uint32_t i = j * kBThreads + t; if (i < total) reinterpret_cast<uint16_t*>(v.out)[dst[i]] = (uint16_t)stage[i]; }
          break;
        default:
#pragma unroll
          for (int j = 0; j < kBRows; ++j) { const uint32_t i = j * kBThreads + t; if (i < total) v.out[dst[i]] = (uint8_t)stage[i]; }
          break;
      }
    }
    __syncthreads();
  }
}

size_t bucket_scatter_smem_bytes(int log2_parts) { return (size_t)kBucketTile * 12 + (size_t)(5u << log2_parts) * 4; }

template <int MODE>
static cudaError_t launch_scatter_mode(const BucketScatterParams& p, int64_t ntiles, int num_sms, cudaStream_t s)
{
  const size_t smem = bucket_scatter_smem_bytes(p.log2_parts);
  cudaError_t e = allow_max_dynamic_smem(bucket_scatter_kernel<MODE>);
  if (e != cudaSuccess) return e;
  const int grid = (int)(ntiles < (int64_t)kBCtasPerSm * num_sms ? ntiles : (int64_t)kBCtasPerSm * num_sms);
  bucket_scatter_kernel<MODE><<<grid, kBThreads, smem, s>>>(p);
  return note_launch();
}

cudaError_t launch_bucket_scatter(const BucketScatterParams& p, int num_sms, cudaStream_t s)
{
  if (p.level == 2) {
    const int64_t ntiles = (int64_t)p.num_segs * ((p.seg_capacity + kBucketTile - 1) / kBucketTile);
    return ntiles ? launch_scatter_mode<2>(p, ntiles, num_sms, s) : cudaSuccess;
  }
  const int64_t ntiles = (p.a.num_rows + kBucketTile - 1) / kBucketTile;
  if (ntiles == 0) return cudaSuccess;
  if (p.key_mode == 8) return launch_scatter_mode<8>(p, ntiles, num_sms, s);
  if (p.key_mode == 4) return launch_scatter_mode<4>(p, ntiles, num_sms, s);
  return launch_scatter_mode<0>(p, ntiles, num_sms, s);
}

// ------------------------------------------------------------------ the reserved slot + status words
// (slot `capacity` is the home of the one key whose image equals EMPTY; the slices never cover it)
__global__ void bucket_init_reserved_kernel(const __grid_constant__ BucketAggParams p)
{
  const int a = threadIdx.x;
  const uint64_t cap = p.table.capacity;
  if (a == 0) { p.table.keys[cap] = kEmpty; p.table.status[0] = 0; p.table.status[1] = 0; p.table.status[2] = 0; p.table.status[3] = 0; p.table.status[6] = 0; }
  if (a < p.num_accs) {
    const uint8_t op = p.acc[a].op;
    if (acc_cell_bytes(op) == 8) reinterpret_cast<uint64_t*>(p.acc[a].cells)[cap] = acc_identity64(op);
    else reinterpret_cast<uint32_t*>(p.acc[a].cells)[cap] = acc_identity32(op);
  }
}

cudaError_t launch_bucket_init_reserved(const BucketAggParams& p, cudaStream_t s)
{
  bucket_init_reserved_kernel<<<1, 32, 0, s>>>(p);
  return note_launch();
}

// ------------------------------------------------------------------ one CTA per bucket: build the slice in shared memory
namespace {
constexpr int kAThreads = 512;
constexpr int kARows    = 4;

__device__ __forceinline__ uint32_t row_vbits(const BucketAggParams& p, uint64_t kw, int64_t row)
{
  if (p.key_in_low32) return (uint32_t)(kw >> 32);
  if (p.vbits_width == 1) return __ldg(p.vbits + row);
  if (p.vbits_width == 4) return __ldg(reinterpret_cast<const uint32_t*>(p.vbits) + row);
  return 0xFFFFFFFFu;
}
// one row of the key whose image equals EMPTY: every accumulator through the reserved slot of the HBM table (cold path)
__device__ __noinline__ void reserved_row(const BucketAggParams& p, uint32_t vb, int64_t row)
{
  p.table.status[2] = 1;
  for (int a = 0; a < p.num_accs; ++a) {
    const BucketAcc& acc = p.acc[a];
    if (!(acc.op == ACC_COUNT_ALL || acc.vbit < 0 || ((vb >> acc.vbit) & 1u))) continue;
    const uint64_t raw = acc.data ? load_raw(acc.data, src_width(acc.src), row) : 0;
    apply_value(acc.op, acc.src, acc.cells, p.table.capacity, raw, 0);
  }
}
}  // namespace

// Phase 2 of a bucket, one accumulator: a tight loop over the bucket's rows. Every decision that depends on the
// accumulator (mode, source type, validity bit) is taken OUTSIDE the row loop -- the descriptor-driven form of this loop
// cost ~280 instructions per row and made the kernel issue-bound at 34 ms per 1e9 rows (profiles/r01_ncu_bucket_agg_v1).
template <typename F>
__device__ __forceinline__ void for_rows(uint32_t n, const uint16_t* s_loc, const uint8_t* s_vb, int vb_width, int vbit, F&& f)
{
  const uint32_t t = threadIdx.x;
  for (uint32_t r0 = 0; r0 < n; r0 += kAThreads * kARows) {
    uint32_t loc[kARows];
#pragma unroll
    for (int j = 0; j < kARows; ++j) {
      const uint32_t r = r0 + j * kAThreads + t;
      loc[j] = 0xFFFFu;
      if (r < n) {
        loc[j] = s_loc[r];
        if (vbit >= 0) {
          const uint32_t w = vb_width == 1 ? s_vb[r] : reinterpret_cast<const uint32_t*>(s_vb)[r];
          if (!((w >> vbit) & 1u)) loc[j] = 0xFFFFu;
        }
      }
    }
    f(r0 + t, loc);
  }
}

__global__ void __launch_bounds__(kAThreads, 3) bucket_aggregate_kernel(const __grid_constant__ BucketAggParams p)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  unsigned long long* s_keys = reinterpret_cast<unsigned long long*>(smem_raw);
  const uint32_t S     = 1u << p.log2_slice;
  const uint32_t smask = S - 1;
  const uint32_t t     = threadIdx.x;
  uint16_t* s_loc      = reinterpret_cast<uint16_t*>(smem_raw + p.loc_off);   // slot of every row (0xFFFF: skip)
  uint8_t* s_vb        = smem_raw + p.vb_off;                                 // validity word of every row (1 or 4 bytes)

  for (uint32_t fb = blockIdx.x; fb < p.num_buckets; fb += gridDim.x) {
    uint32_t ntot = __ldg(p.cursor + fb);
    ntot = ntot < p.bucket_capacity ? ntot : p.bucket_capacity;
    const int64_t base0 = (int64_t)fb * p.bucket_capacity;

    // ---- EMPTY keys + identity cells (shared-memory accumulators here, BM_GLOBAL ones in the bucket's slice of the HBM table)
    for (uint32_t i = t; i < S / 2; i += kAThreads) reinterpret_cast<uint4*>(s_keys)[i] = make_uint4(~0u, ~0u, ~0u, ~0u);
#pragma unroll 1
    for (int a = 0; a < p.num_accs; ++a) {
      const BucketAcc& acc = p.acc[a];
      const uint8_t op = acc.op;
      if (acc.mode == BM_GLOBAL) {
        const uint32_t cb = acc_cell_bytes(op);
        uint8_t* g = reinterpret_cast<uint8_t*>(acc.cells) + (uint64_t)fb * S * cb;
        if (cb == 8) { const uint64_t id = acc_identity64(op); for (uint32_t i = t; i < S / 2; i += kAThreads) reinterpret_cast<ulonglong2*>(g)[i] = make_ulonglong2(id, id); }
        else { const uint32_t id = acc_identity32(op); for (uint32_t i = t; i < S / 4; i += kAThreads) reinterpret_cast<uint4*>(g)[i] = make_uint4(id, id, id, id); }
      } else if (acc.mode == BM_MINMAX) {
        const uint64_t id = acc_identity64(op);
        ulonglong2* c = reinterpret_cast<ulonglong2*>(smem_raw + acc.smem_off);
        for (uint32_t i = t; i < S / 2; i += kAThreads) c[i] = make_ulonglong2(id, id);
      } else {
        uint4* c = reinterpret_cast<uint4*>(smem_raw + acc.smem_off);
        const uint32_t words = acc.mode == BM_LIMBS ? (p.limb_total ? 5 * S : 3 * S) : S;  // digit sums (+ the 64-bit running total)
        for (uint32_t i = t; i < words / 4; i += kAThreads) c[i] = make_uint4(0, 0, 0, 0);
      }
    }
    __syncthreads();

    // A bucket is walked in chunks of chunk_rows rows (the 16-bit slot numbers of one chunk live in shared memory); the
    // high-cardinality shapes have one chunk per bucket, mid cardinalities (few, long buckets) many.
    uint32_t since_fold = 0;
    for (uint32_t c0 = 0; c0 < ntot; c0 += p.chunk_rows) {
    const uint32_t n   = ntot - c0 < p.chunk_rows ? ntot - c0 : p.chunk_rows;
    const int64_t base = base0 + c0;

    // ---- phase 1: key -> slot of the slice for every row (LDS probe, ATOMS.CAS only on the first row of a group).
    // A lane walks its kARows rows as ONE loop (a probe step per iteration, next row when the slot is found): the warp runs
    // max-over-lanes of the SUMMED probe lengths instead of the sum over rows of the per-row maxima, and the rare insert
    // branch exists once.
    for (uint32_t r0 = 0; r0 < n; r0 += kAThreads * kARows) {
      uint64_t kq[kARows];
      uint32_t lq[kARows];
      int cnt = 0;  // this lane's rows of the iteration are a prefix: r0 + j * kAThreads + t < n
#pragma unroll
      for (int j = 0; j < kARows; ++j) {
        const uint32_t r = r0 + j * kAThreads + t;
        kq[j] = 0;
        if (r < n) { kq[j] = __ldcs(reinterpret_cast<const unsigned long long*>(p.key) + base + r); cnt = j + 1; }
      }
#pragma unroll
      for (int j = 0; j < kARows; ++j) {
        const uint32_t r = r0 + j * kAThreads + t;
        if (p.vb_width != 0 && r < n) {
          const uint32_t vb = row_vbits(p, kq[j], base + r);
          if (p.vb_width == 1) s_vb[r] = (uint8_t)vb; else reinterpret_cast<uint32_t*>(s_vb)[r] = vb;
        }
        if (p.key_in_low32) kq[j] &= 0xffffffffull;
        lq[j] = slot_hash32(kq[j]) & smask;
      }
      uint32_t r = r0 + t, probes = 0;
      while (cnt > 0) {
        const unsigned long long k = kq[0];
        uint32_t found = 0xFFFFFFFFu;  // 0xFFFF: row skipped
        if (k == kEmpty) { reserved_row(p, p.vb_width == 0 ? 0xFFFFFFFFu : (p.vb_width == 1 ? (uint32_t)s_vb[r] : reinterpret_cast<uint32_t*>(s_vb)[r]), base + r); found = 0xFFFFu; }
        else {
          const unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(s_keys + lq[0]);
          if (cur == k) found = lq[0];
          else if (cur == kEmpty) {
            const unsigned long long old = atomicCAS(s_keys + lq[0], (unsigned long long)kEmpty, k);
            if (old == kEmpty || old == k) found = lq[0];
          }
          if (found == 0xFFFFFFFFu) {
            lq[0] = (lq[0] + 1) & smask;
            if (++probes > smask) { p.table.status[0] = 1; found = 0xFFFFu; }  // more distinct keys than slots: the host retries bigger
          }
        }
        if (found != 0xFFFFFFFFu) {
          s_loc[r] = (uint16_t)found;
#pragma unroll
          for (int j = 0; j + 1 < kARows; ++j) { kq[j] = kq[j + 1]; lq[j] = lq[j + 1]; }
          r += kAThreads; --cnt; probes = 0;
        }
      }
    }
    __syncthreads();

    // ---- phase 2: one tight pass over the rows per accumulator
#pragma unroll 1
    for (int a = 0; a < p.num_accs; ++a) {
      const BucketAcc& acc = p.acc[a];
      unsigned char* cells_s = smem_raw + acc.smem_off;
      const int vbit = acc.op == ACC_COUNT_ALL ? -1 : acc.vbit;
      if (acc.mode == BM_COUNT) {
        uint32_t* c = reinterpret_cast<uint32_t*>(cells_s);
        for_rows(n, s_loc, s_vb, p.vb_width, vbit, [&](uint32_t, const uint32_t (&loc)[kARows]) {
#pragma unroll
          for (int j = 0; j < kARows; ++j) if (loc[j] != 0xFFFFu) atomicAdd(c + loc[j], 1u);
        });
        continue;
      }
      const int w = src_width(acc.src);
      const uint8_t* col = acc.data + base * w;
      if (acc.mode == BM_LIMBS) {
        // sign- or zero-extension as one shift pair: signed sources shift by 64 - 8w, unsigned by 0; BOOL8 counts non-zero
        const int sh      = (acc.src == S_I8 || acc.src == S_I16 || acc.src == S_I32) ? 64 - 8 * w : 0;
        const bool isbool = acc.src == S_BOOL, sq = acc.op == ACC_SUMSQ_I64;
        uint32_t* c = reinterpret_cast<uint32_t*>(cells_s);
        auto body = [&](uint64_t raw, uint32_t l) {
          int64_t v = ((int64_t)(raw << sh)) >> sh;
          if (isbool) v = raw != 0;
          if (sq) v = (int64_t)((uint64_t)v * (uint64_t)v);
          const int32_t d0  = (int16_t)v;
          const int64_t r1  = (v - d0) >> 16;
          const int32_t d1  = (int16_t)r1;
          const uint32_t r2 = (uint32_t)((r1 - d1) >> 16);
          if (d0) atomicAdd(c + l, (uint32_t)d0);
          if (d1) atomicAdd(c + S + l, (uint32_t)d1);
          if (r2) atomicAdd(c + 2 * S + l, r2);
        };
        if (w == 8) {
          for_rows(n, s_loc, s_vb, p.vb_width, vbit, [&](uint32_t r, const uint32_t (&loc)[kARows]) {
            uint64_t raw[kARows];
#pragma unroll
            for (int j = 0; j < kARows; ++j) { raw[j] = 0; if (loc[j] != 0xFFFFu) raw[j] = __ldcs(reinterpret_cast<const unsigned long long*>(col) + r + j * kAThreads); }
#pragma unroll
            for (int j = 0; j < kARows; ++j) if (loc[j] != 0xFFFFu) body(raw[j], loc[j]);
          });
        } else {
          for_rows(n, s_loc, s_vb, p.vb_width, vbit, [&](uint32_t r, const uint32_t (&loc)[kARows]) {
            uint64_t raw[kARows];
#pragma unroll
            for (int j = 0; j < kARows; ++j) { raw[j] = 0; if (loc[j] != 0xFFFFu) raw[j] = load_raw(col, w, r + j * kAThreads); }
#pragma unroll
            for (int j = 0; j < kARows; ++j) if (loc[j] != 0xFFFFu) body(raw[j], loc[j]);
          });
        }
      } else if (acc.mode == BM_MINMAX) {
        const bool is_min = acc.op == ACC_MIN;
        const int src     = acc.src;
        unsigned long long* c = reinterpret_cast<unsigned long long*>(cells_s);
        for_rows(n, s_loc, s_vb, p.vb_width, vbit, [&](uint32_t r, const uint32_t (&loc)[kARows]) {
          uint64_t raw[kARows];
#pragma unroll
          for (int j = 0; j < kARows; ++j) { raw[j] = 0; if (loc[j] != 0xFFFFu) raw[j] = load_raw(col, w, r + j * kAThreads); }
#pragma unroll
          for (int j = 0; j < kARows; ++j) {
            if (loc[j] == 0xFFFFu) continue;
            const unsigned long long v   = ordered_image(raw[j], src);
            const unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(c + loc[j]);
            if (is_min) { if (v < cur) atomicMin(c + loc[j], v); } else { if (v > cur) atomicMax(c + loc[j], v); }
          }
        });
      } else if (acc.op == ACC_SUM_F64) {
        double* c = reinterpret_cast<double*>(acc.cells) + (uint64_t)fb * S;
        for_rows(n, s_loc, s_vb, p.vb_width, vbit, [&](uint32_t r, const uint32_t (&loc)[kARows]) {
          uint64_t raw[kARows];
#pragma unroll
          for (int j = 0; j < kARows; ++j) { raw[j] = 0; if (loc[j] != 0xFFFFu) raw[j] = __ldcs(reinterpret_cast<const unsigned long long*>(col) + r + j * kAThreads); }
#pragma unroll
          for (int j = 0; j < kARows; ++j) if (loc[j] != 0xFFFFu) atomicAdd(c + loc[j], __longlong_as_double((long long)raw[j]));
        });
      } else {
        const uint8_t op = acc.op, src = acc.src;
        void* cells      = acc.cells;
        const uint64_t s0 = (uint64_t)fb * S;
        for_rows(n, s_loc, s_vb, p.vb_width, vbit, [&](uint32_t r, const uint32_t (&loc)[kARows]) {
#pragma unroll 1
          for (int j = 0; j < kARows; ++j) if (loc[j] != 0xFFFFu) apply_value(op, src, cells, s0 + loc[j], load_raw(col, w, r + j * kAThreads), 0);
        });
      }
    }
    __syncthreads();
    // the digit sums hold < 65536 addends each: fold them into the 64-bit totals before the next chunk could pass that
    since_fold += n;
    if (p.limb_total && c0 + n < ntot && since_fold + p.chunk_rows > 65535u) {
#pragma unroll 1
      for (int a = 0; a < p.num_accs; ++a) {
        if (p.acc[a].mode != BM_LIMBS) continue;
        uint32_t* c   = reinterpret_cast<uint32_t*>(smem_raw + p.acc[a].smem_off);
        uint64_t* tot = reinterpret_cast<uint64_t*>(c + 3 * S);
        for (uint32_t l = t; l < S; l += kAThreads) {
          tot[l] += (uint64_t)(int64_t)(int32_t)c[l] + ((uint64_t)(int64_t)(int32_t)c[S + l] << 16) + ((uint64_t)c[2 * S + l] << 32);
          c[l] = 0; c[S + l] = 0; c[2 * S + l] = 0;
        }
      }
      since_fold = 0;
      __syncthreads();
    }
    }  // chunks

    // ---- the finished slice -> HBM, 16 bytes per thread
    {
      uint4* gk = reinterpret_cast<uint4*>(p.table.keys + (uint64_t)fb * S);
      for (uint32_t i = t; i < S / 2; i += kAThreads) gk[i] = reinterpret_cast<const uint4*>(s_keys)[i];
#pragma unroll 1
      for (int a = 0; a < p.num_accs; ++a) {
        const BucketAcc& acc = p.acc[a];
        if (acc.mode == BM_GLOBAL) continue;  // already in place
        const uint32_t cb = acc_cell_bytes(acc.op);
        uint8_t* g = reinterpret_cast<uint8_t*>(acc.cells) + (uint64_t)fb * S * cb;
        if (acc.mode == BM_LIMBS) {  // sum mod 2^64 = s0 + 2^16 s1 + 2^32 s2 (s0, s1 signed digit sums)
          const uint32_t* c = reinterpret_cast<const uint32_t*>(smem_raw + acc.smem_off);
          for (uint32_t i = t; i < S / 2; i += kAThreads) {
            uint64_t o[2];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              const uint32_t l = 2 * i + h;
              o[h] = (uint64_t)(int64_t)(int32_t)c[l] + ((uint64_t)(int64_t)(int32_t)c[S + l] << 16) + ((uint64_t)c[2 * S + l] << 32);
              if (p.limb_total) o[h] += reinterpret_cast<const uint64_t*>(c + 3 * S)[l];
            }
            reinterpret_cast<ulonglong2*>(g)[i] = make_ulonglong2(o[0], o[1]);
          }
        } else {
          const uint4* c = reinterpret_cast<const uint4*>(smem_raw + acc.smem_off);
          for (uint32_t i = t; i < S * cb / 16; i += kAThreads) reinterpret_cast<uint4*>(g)[i] = c[i];
        }
      }
    }
    __syncthreads();
  }
}

cudaError_t launch_bucket_aggregate(const BucketAggParams& p, size_t smem_bytes, cudaStream_t s)
{
  cudaError_t e = allow_max_dynamic_smem(bucket_aggregate_kernel);
  if (e != cudaSuccess) return e;
  bucket_aggregate_kernel<<<p.num_buckets, kAThreads, smem_bytes, s>>>(p);
  return note_launch();
}

// ------------------------------------------------------------------ overflow rows through the HBM table
// Same slots as the slice kernel: home = hash & (capacity - 1), linear probing that wraps INSIDE the slice.
__device__ __forceinline__ uint64_t find_or_insert_sliced(const TableDesc& t, uint64_t k0, int log2_slice)
{
  const uint64_t cap = t.capacity;
  if (k0 == kEmpty) { t.status[2] = 1; return cap; }
  const uint64_t smask = (1ull << log2_slice) - 1;
  uint64_t slot = slot_hash32(k0) & (cap - 1);
  unsigned long long* keys = reinterpret_cast<unsigned long long*>(t.keys);
  for (uint64_t probes = 0; probes <= smask; ++probes) {
    unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(keys + slot);
    if (cur == k0) return slot;
    if (cur == kEmpty) {
      unsigned long long old = atomicCAS(keys + slot, (unsigned long long)kEmpty, (unsigned long long)k0);
      if (old == kEmpty || old == k0) return slot;
    }
    slot = (slot & ~smask) | ((slot + 1) & smask);
  }
  t.status[0] = 1;
  return kNoSlot;
}

__global__ void __launch_bounds__(256) bucket_overflow_kernel(const __grid_constant__ BucketAggParams p)
{
  uint32_t n = *p.ov_cursor;
  n = n < p.ov_capacity ? n : p.ov_capacity;
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n; r += stride) {
    const int64_t row = (int64_t)p.ov_base + r;
    const uint64_t kw = __ldg(reinterpret_cast<const unsigned long long*>(p.key) + row);
    const uint32_t vb = row_vbits(p, kw, row);
    const uint64_t k  = p.key_in_low32 ? (kw & 0xffffffffull) : kw;
    const uint64_t slot = find_or_insert_sliced(p.table, k, p.log2_slice);
    if (slot == kNoSlot) return;
#pragma unroll 1
    for (int a = 0; a < p.num_accs; ++a) {
      const BucketAcc& acc = p.acc[a];
      if (!(acc.op == ACC_COUNT_ALL || acc.vbit < 0 || ((vb >> acc.vbit) & 1u))) continue;
      const uint64_t raw = acc.data ? load_raw(acc.data, src_width(acc.src), row) : 0;
      apply_value(acc.op, acc.src, acc.cells, slot, raw, 0);
    }
  }
}

cudaError_t launch_bucket_overflow(const BucketAggParams& p, cudaStream_t s)
{
  if (p.ov_capacity == 0) return cudaSuccess;
  const uint32_t want = (p.ov_capacity + 255) / 256;
  bucket_overflow_kernel<<<want < (uint32_t)device_sm_count() * 8 ? want : (uint32_t)device_sm_count() * 8, 256, 0, s>>>(p);
  return note_launch();
}

}  // namespace bgb

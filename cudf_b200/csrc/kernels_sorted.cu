// Mid-cardinality kernel for accumulator sets the split-lane shared-memory kernel (kernels_smem.cu) cannot hold: float and
// double sums, sums of squares, MIN/MAX, products, several value columns -- e.g. BASELINE configs[0]
// (df.groupby(['A','B']).mean(): 1024 groups, two float64 MEANs).
//
// Shared memory has no native 64-bit or floating-point add (they compile to CAS loops, ~4 cycles per lane), and the L2 takes
// ~210 G atomic requests/s chip-wide, so neither "atomics on a shared-memory table" nor "atomics on the HBM table" reaches
// the HBM read rate when a row feeds 2-6 accumulators. This kernel uses NO atomics on accumulators at all:
//   1. a CTA reads a tile of 2048 rows; every row takes its RANK within its group from one returning 32-bit ATOMS.ADD on a
//      per-tile histogram (the one shared-memory atomic that is native),
//   2. the histogram is scanned (each thread scans the <= 8 consecutive groups it OWNS, block scan of the thread totals),
//   3. rows write their raw values to start[group] + rank: the tile is now counting-sorted by group in shared memory,
//   4. every thread folds the rows of the groups it owns, in registers, and adds the partials to the CTA's table with plain
//      loads and stores -- it is the only thread that touches those cells. COUNTs of unmasked columns are the histogram.
//      Groups with more than 64 rows in the tile (skewed keys) are left to a whole warp each (strided fold + shuffle tree).
// Group ids are (key - minimum) when the sampled key range is dense, else the slot of the key in a small open-addressing table
// of the CTA ("sorted-hashed": sparse integer keys, float keys, several key columns).
// At the end the CTA merges its table into the HBM group table (find-or-insert + one atomic per cell and CTA).
// The reference runs this shape through compute_shared_memory_aggs / global atomics
// (/root/reference/cpp/src/groupby/hash/compute_shared_memory_aggs.cu, compute_global_memory_aggs.cuh).
#include "device_utils.cuh"
#include "kernels_sorted.h"

namespace bgb {

namespace {

constexpr uint32_t kNone = 0xFFFFFFFFu;

__device__ __forceinline__ uint64_t d2u(double d) { return (uint64_t)__double_as_longlong(d); }
__device__ __forceinline__ double u2d(uint64_t u) { return __longlong_as_double((long long)u); }

__device__ __forceinline__ uint64_t sorted_identity(uint8_t op)
{
  return acc_cell_bytes(op) == 8 ? acc_identity64(op) : (uint64_t)acc_identity32(op);
}

// Values are parked in shared memory in a per-COLUMN canonical form, so the fold loops below carry no per-row type switch:
// integers sign- or zero-extended to 64 bits (raw_to_i64), floats as their raw bits.
__device__ __forceinline__ uint64_t canonical_value(uint64_t raw, uint8_t src)
{
  return (src == S_F32 || src == S_F64) ? raw : (uint64_t)raw_to_i64(raw, src);
}

template <class F>
__device__ __forceinline__ uint64_t fold_rows(const uint64_t* vals, const uint8_t* sv, uint32_t vbit, uint32_t b, uint32_t e, uint32_t stride,
                                              uint64_t part, F f)
{
  if (vbit == 0) { for (uint32_t i = b; i < e; i += stride) part = f(part, vals[i]); }
  else { for (uint32_t i = b; i < e; i += stride) if (sv[i] & vbit) part = f(part, vals[i]); }
  return part;
}

// folds the canonical values at sorted positions b, b + stride, ... < e into one partial of accumulator `op` (the switch
// is outside the loops; same arithmetic as apply_value, device_utils.cuh). vbit = 0: the column has no mask.
__device__ __forceinline__ uint64_t fold_group(uint8_t op, uint8_t src, const uint64_t* vals, const uint8_t* sv, uint32_t vbit, uint32_t b,
                                               uint32_t e, uint32_t stride)
{
  auto f32 = [](uint64_t u) { return __uint_as_float((uint32_t)u); };
  auto u32 = [](float f) { return (uint64_t)__float_as_uint(f); };
  switch (op) {
    case ACC_SUM_I64: return fold_rows(vals, sv, vbit, b, e, stride, 0ull, [](uint64_t p, uint64_t v) { return p + v; });
    case ACC_SUMSQ_I64: return fold_rows(vals, sv, vbit, b, e, stride, 0ull, [](uint64_t p, uint64_t v) { return p + v * v; });
    case ACC_PROD_I64: return fold_rows(vals, sv, vbit, b, e, stride, 1ull, [](uint64_t p, uint64_t v) { return p * v; });
    case ACC_SUM_F64: return fold_rows(vals, sv, vbit, b, e, stride, 0ull, [](uint64_t p, uint64_t v) { return d2u(u2d(p) + u2d(v)); });
    case ACC_SUMSQ_F64: return fold_rows(vals, sv, vbit, b, e, stride, 0ull, [](uint64_t p, uint64_t v) { return d2u(fma(u2d(v), u2d(v), u2d(p))); });
    case ACC_PROD_F64: return fold_rows(vals, sv, vbit, b, e, stride, d2u(1.0), [](uint64_t p, uint64_t v) { return d2u(u2d(p) * u2d(v)); });
    case ACC_SUM_F32: return fold_rows(vals, sv, vbit, b, e, stride, 0ull, [=](uint64_t p, uint64_t v) { return u32(f32(p) + f32(v)); });
    case ACC_SUMSQ_F32: return fold_rows(vals, sv, vbit, b, e, stride, 0ull, [=](uint64_t p, uint64_t v) { return u32(f32(p) + f32(v) * f32(v)); });
    case ACC_PROD_F32: return fold_rows(vals, sv, vbit, b, e, stride, u32(1.0f), [=](uint64_t p, uint64_t v) { return u32(f32(p) * f32(v)); });
    case ACC_COUNT_VALID: case ACC_COUNT_ALL: return fold_rows(vals, sv, vbit, b, e, stride, 0ull, [](uint64_t p, uint64_t) { return p + 1; });
    case ACC_MIN: case ACC_MAX: {
      const bool mn = op == ACC_MIN;
      const uint64_t id = mn ? ~0ull : 0ull;
      auto pick = [=](uint64_t p, uint64_t img) { return mn ? (img < p ? img : p) : (img > p ? img : p); };
      switch (src) {
        case S_I8: case S_I16: case S_I32: case S_I64:
          return fold_rows(vals, sv, vbit, b, e, stride, id, [=](uint64_t p, uint64_t v) { return pick(p, v ^ 0x8000000000000000ull); });
        case S_F32: return fold_rows(vals, sv, vbit, b, e, stride, id, [=](uint64_t p, uint64_t v) { return pick(p, ordered_image(v, S_F32)); });
        case S_F64: return fold_rows(vals, sv, vbit, b, e, stride, id, [=](uint64_t p, uint64_t v) { return pick(p, ordered_image(v, S_F64)); });
        default: return fold_rows(vals, sv, vbit, b, e, stride, id, [=](uint64_t p, uint64_t v) { return pick(p, v); });
      }
    }
    default: return 0;
  }
}

// associative combination of two partials of one accumulator
__device__ __forceinline__ uint64_t fold(uint8_t op, uint64_t a, uint64_t b)
{
  switch (op) {
    case ACC_SUM_F64: case ACC_SUMSQ_F64: return d2u(u2d(a) + u2d(b));
    case ACC_SUM_F32: case ACC_SUMSQ_F32: return __float_as_uint(__uint_as_float((uint32_t)a) + __uint_as_float((uint32_t)b));
    case ACC_PROD_I64: return (uint64_t)((int64_t)a * (int64_t)b);
    case ACC_PROD_F64: return d2u(u2d(a) * u2d(b));
    case ACC_PROD_F32: return __float_as_uint(__uint_as_float((uint32_t)a) * __uint_as_float((uint32_t)b));
    case ACC_MIN: return a < b ? a : b;
    case ACC_MAX: return a > b ? a : b;
    default: return a + b;  // integer sums, counts
  }
}

// a CTA's partial into the HBM cell
__device__ __forceinline__ void atomic_fold(uint8_t op, void* cells, uint64_t slot, uint64_t v)
{
  unsigned long long* c64 = reinterpret_cast<unsigned long long*>(cells) + slot;
  unsigned int* c32       = reinterpret_cast<unsigned int*>(cells) + slot;
  switch (op) {
    case ACC_COUNT_VALID: case ACC_COUNT_ALL: if ((uint32_t)v) atomicAdd(c32, (uint32_t)v); break;
    case ACC_SUM_I64: case ACC_SUMSQ_I64: atomicAdd(c64, (unsigned long long)v); break;
    case ACC_SUM_F64: case ACC_SUMSQ_F64: atomicAdd(reinterpret_cast<double*>(c64), u2d(v)); break;
    case ACC_SUM_F32: case ACC_SUMSQ_F32: atomicAdd(reinterpret_cast<float*>(c32), __uint_as_float((uint32_t)v)); break;
    case ACC_PROD_I64: atomic_mul_i64(c64, (int64_t)v); break;
    case ACC_PROD_F64: atomic_mul_f64(c64, u2d(v)); break;
    case ACC_PROD_F32: atomic_mul_f32(c32, __uint_as_float((uint32_t)v)); break;
    case ACC_MIN: atomicMin(c64, (unsigned long long)v); break;
    case ACC_MAX: atomicMax(c64, (unsigned long long)v); break;
    default: break;
  }
}

// a key outside the dense range (the sampled range missed it): straight into the HBM table
__device__ __noinline__ void bypass_row(const SortedParams& p, int64_t row)
{
  uint64_t k = 0, k1 = 0;
  if (!pack_key(p.a, row, k, k1)) return;
  const uint64_t slot = find_or_insert(p.a.table, k, 0);
  if (slot == kNoSlot) return;
  for (int a = 0; a < p.num_accs; ++a) {
    const SortedAcc& ac = p.acc[a];
    if (ac.col < 0) { apply_value(ac.op, S_I32, ac.cells, slot, 0, row); continue; }
    const SortedCol& c = p.col[ac.col];
    const int64_t idx  = (int64_t)c.offset + row;
    if (c.mask != nullptr && !bit_is_set(c.mask, idx)) continue;
    apply_value(ac.op, c.src, ac.cells, slot, load_raw(c.data, c.width, idx), row);
  }
}

// ---- owner phase of ONE accumulator: the thread walks the groups [g0, gend) it owns; their rows are consecutive in the sorted
// tile. f folds a value into the partial, c combines the partial with the CTA's cell. Heavy groups are left to the warps.
struct OwnerCtx {
  const uint32_t* start;
  uint64_t* tab;          // this accumulator's cells
  const uint64_t* vals;   // this accumulator's value column, sorted
  const uint8_t* sv;
  uint32_t vbit;          // 0: the column has no mask
  uint32_t g0, gend;
};

template <class F, class C>
__device__ __forceinline__ void owner_rows(const OwnerCtx& x, uint64_t init, F f, C c)
{
  uint32_t b = x.start[x.g0];
  for (uint32_t g = x.g0; g < x.gend; ++g) {
    const uint32_t e = x.start[g + 1];
    const uint32_t n = e - b;
    if (n != 0 && n <= (uint32_t)kSortedHeavy) {
      uint64_t part = init;
      if (x.vbit == 0) { for (uint32_t i = b; i < e; ++i) part = f(part, x.vals[i]); }
      else { for (uint32_t i = b; i < e; ++i) if (x.sv[i] & x.vbit) part = f(part, x.vals[i]); }
      x.tab[g] = c(x.tab[g], part);
    }
    b = e;
  }
}

__device__ __forceinline__ void owner_acc(uint8_t op, uint8_t src, const OwnerCtx& x)
{
  auto f32 = [](uint64_t u) { return __uint_as_float((uint32_t)u); };
  auto u32 = [](float f) { return (uint64_t)__float_as_uint(f); };
  auto addi = [](uint64_t a, uint64_t b) { return a + b; };
  auto addd = [](uint64_t a, uint64_t b) { return d2u(u2d(a) + u2d(b)); };
  auto addf = [=](uint64_t a, uint64_t b) { return u32(f32(a) + f32(b)); };
  switch (op) {
    case ACC_SUM_I64: owner_rows(x, 0ull, addi, addi); break;
    case ACC_SUMSQ_I64: owner_rows(x, 0ull, [](uint64_t p, uint64_t v) { return p + v * v; }, addi); break;
    case ACC_PROD_I64: owner_rows(x, 1ull, [](uint64_t p, uint64_t v) { return p * v; }, [](uint64_t a, uint64_t b) { return a * b; }); break;
    case ACC_SUM_F64: owner_rows(x, 0ull, addd, addd); break;
    case ACC_SUMSQ_F64: owner_rows(x, 0ull, [](uint64_t p, uint64_t v) { return d2u(u2d(p) + u2d(v) * u2d(v)); }, addd); break;
    case ACC_PROD_F64: owner_rows(x, d2u(1.0), [](uint64_t p, uint64_t v) { return d2u(u2d(p) * u2d(v)); }, [](uint64_t a, uint64_t b) { return d2u(u2d(a) * u2d(b)); }); break;
    case ACC_SUM_F32: owner_rows(x, 0ull, addf, addf); break;
    case ACC_SUMSQ_F32: owner_rows(x, 0ull, [=](uint64_t p, uint64_t v) { return u32(f32(p) + f32(v) * f32(v)); }, addf); break;
    case ACC_PROD_F32: owner_rows(x, u32(1.0f), [=](uint64_t p, uint64_t v) { return u32(f32(p) * f32(v)); }, [=](uint64_t a, uint64_t b) { return u32(f32(a) * f32(b)); }); break;
    case ACC_COUNT_VALID: case ACC_COUNT_ALL: owner_rows(x, 0ull, [](uint64_t p, uint64_t) { return p + 1; }, addi); break;
    case ACC_MIN: {
      auto mn = [](uint64_t a, uint64_t b) { return a < b ? a : b; };
      switch (src) {
        case S_I8: case S_I16: case S_I32: case S_I64: owner_rows(x, ~0ull, [=](uint64_t p, uint64_t v) { return mn(p, v ^ 0x8000000000000000ull); }, mn); break;
        case S_F32: owner_rows(x, ~0ull, [=](uint64_t p, uint64_t v) { return mn(p, ordered_image(v, S_F32)); }, mn); break;
        case S_F64: owner_rows(x, ~0ull, [=](uint64_t p, uint64_t v) { return mn(p, ordered_image(v, S_F64)); }, mn); break;
        default: owner_rows(x, ~0ull, mn, mn); break;
      }
      break;
    }
    case ACC_MAX: {
      auto mx = [](uint64_t a, uint64_t b) { return a > b ? a : b; };
      switch (src) {
        case S_I8: case S_I16: case S_I32: case S_I64: owner_rows(x, 0ull, [=](uint64_t p, uint64_t v) { return mx(p, v ^ 0x8000000000000000ull); }, mx); break;
        case S_F32: owner_rows(x, 0ull, [=](uint64_t p, uint64_t v) { return mx(p, ordered_image(v, S_F32)); }, mx); break;
        case S_F64: owner_rows(x, 0ull, [=](uint64_t p, uint64_t v) { return mx(p, ordered_image(v, S_F64)); }, mx); break;
        default: owner_rows(x, 0ull, mx, mx); break;
      }
      break;
    }
    default: break;
  }
}

template <int NC>
__global__ void __launch_bounds__(kSortedThreads, 2) aggregate_sorted_kernel(const __grid_constant__ SortedParams p)
{
  extern __shared__ __align__(16) uint8_t smem_raw[];
  constexpr int NT     = kSortedThreads;
  constexpr uint32_t T = kSortedTile;
  const uint32_t G     = p.groups;
  const int A          = p.num_accs;
  uint64_t* tab     = reinterpret_cast<uint64_t*>(smem_raw);           // [A][G] the CTA's accumulators
  uint64_t* skey    = tab + (size_t)A * G;                              // [G] hashed mode: the key that owns each slot (kEmpty: free)
  uint64_t* sval    = skey + (p.hashed ? G : 0);                        // [NC][T] canonical values of the tile, sorted by group
  uint32_t* cnt     = reinterpret_cast<uint32_t*>(sval + (size_t)NC * T);  // [G] rows per group in this tile
  uint32_t* start   = cnt + G;                                          // [G + 1] first sorted position of every group
  uint32_t* rows    = start + G + 1;                                    // [G] != 0: the CTA met the group
  uint32_t* wsum    = rows + G;                                         // [32] scan scratch; wsum[31] = number of heavy groups
  uint16_t* heavy   = reinterpret_cast<uint16_t*>(wsum + 32);           // [T / kSortedHeavy]
  uint8_t* svalid   = reinterpret_cast<uint8_t*>(heavy + T / kSortedHeavy);  // [T] validity bits of the tile's rows, sorted

  const int tid  = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  for (uint32_t i = tid; i < (uint32_t)A * G; i += NT) tab[i] = sorted_identity(p.acc[i / G].op);
  for (uint32_t g = tid; g < G; g += NT) { cnt[g] = 0; rows[g] = 0; if (p.hashed) skey[g] = kEmpty; }
  if (tid == 0) { wsum[31] = 0; start[0] = 0; }
  const uint32_t gpt  = (G + NT - 1) / NT;   // groups owned per thread
  const uint32_t g0   = min(tid * gpt, G);
  const uint32_t gend = min(g0 + gpt, G);
  __syncthreads();

  const int64_t N      = p.a.num_rows;
  const int64_t ntiles = (N + T - 1) / T;
  // keys of a tile: loaded one tile AHEAD (right after the scatter step of the previous tile), so that their HBM latency hides
  // behind the owner phase; `ok` bit r = row r exists and is not dropped (EXCLUDE policy + null key)
  uint64_t k[kSortedRows];
  uint32_t ok = 0;
  auto load_keys = [&](int64_t t) {
    ok = 0;
    const int64_t base = t * T;
#pragma unroll
    for (int r = 0; r < kSortedRows; ++r) {
      const int64_t row = base + r * NT + tid;
      k[r] = 0;
      if (t < ntiles && row < N) {
        if (p.key_mode == 2) { k[r] = __ldg(reinterpret_cast<const unsigned long long*>(p.a.kc[0].data) + p.a.kc[0].offset + row); ok |= 1u << r; }
        else if (p.key_mode == 1) { k[r] = load_raw(p.a.kc[0].data, p.a.kc[0].width, (int64_t)p.a.kc[0].offset + row); ok |= 1u << r; }
        else { uint64_t k1; if (pack_key(p.a, row, k[r], k1)) ok |= 1u << r; }
      }
    }
  };
  load_keys(blockIdx.x);
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t base = tile * T;
    uint32_t gid[kSortedRows], rank[kSortedRows], vb[kSortedRows];
    uint64_t val[kSortedRows][NC];
    // ---- 1. rows -> dense group id, canonical values, rank within the group
    uint32_t byp = 0;
#pragma unroll
    for (int r = 0; r < kSortedRows; ++r) {
      gid[r] = kNone; vb[r] = 0; rank[r] = 0;
      if ((ok >> r) & 1) {
        if (p.hashed) {
          // the group id is the SLOT of the key in the CTA's open-addressing table (64-bit ATOMS.CAS only for the first row of
          // a group); a key that finds no slot within 32 probes (more groups than estimated) goes to the HBM table
          uint32_t slot = slot_hash32(k[r]) & (G - 1);
          gid[r] = kNone;
          if (k[r] != kEmpty) {
            for (int probe = 0; probe < 32; ++probe) {
              unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(skey + slot);
              if (cur == kEmpty) cur = atomicCAS(reinterpret_cast<unsigned long long*>(skey + slot), (unsigned long long)kEmpty, (unsigned long long)k[r]);
              if (cur == k[r] || cur == kEmpty) { gid[r] = slot; break; }
              slot = (slot + 1) & (G - 1);
            }
          }
          if (gid[r] == kNone) byp |= 1u << r;
        } else {
          const uint64_t d = (k[r] ^ p.key_flip) - p.key_base;
          if (d < G) gid[r] = (uint32_t)d; else byp |= 1u << r;
        }
      }
    }
#pragma unroll
    for (int r = 0; r < kSortedRows; ++r) {
      const int64_t row = base + r * NT + tid;
#pragma unroll
      for (int c = 0; c < NC; ++c) {
        val[r][c] = 0;
        if (gid[r] != kNone) {
          const SortedCol& sc = p.col[c];
          if (p.fast8) val[r][c] = __ldg(reinterpret_cast<const unsigned long long*>(sc.data) + sc.offset + row);
          else if (c < p.num_cols) {
            const int64_t idx = (int64_t)sc.offset + row;
            if (sc.mask == nullptr || bit_is_set(sc.mask, idx)) { val[r][c] = canonical_value(load_raw(sc.data, sc.width, idx), sc.src); vb[r] |= 1u << c; }
          }
        }
      }
    }
#pragma unroll
    for (int r = 0; r < kSortedRows; ++r)
      if (gid[r] != kNone) rank[r] = atomicAdd(&cnt[gid[r]], 1u);
    if (byp != 0) {
#pragma unroll
      for (int r = 0; r < kSortedRows; ++r)
        if ((byp >> r) & 1) bypass_row(p, base + r * NT + tid);
    }
    __syncthreads();
    // ---- 2. exclusive scan of the histogram: thread totals -> warp scan -> block scan
    uint32_t mine = 0;
    for (uint32_t g = g0; g < gend; ++g) mine += cnt[g];
    uint32_t incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t n = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += n; }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      uint32_t w = lane < NT / 32 ? wsum[lane] : 0, wi = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const uint32_t n = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += n; }
      if (lane < NT / 32) wsum[lane] = wi - w;
    }
    __syncthreads();
    {
      uint32_t run = wsum[warp] + incl - mine;
      for (uint32_t g = g0; g < gend; ++g) {
        const uint32_t c = cnt[g];
        run += c;
        start[g + 1] = run;   // start[0] stays 0
        if (c != 0) {
          cnt[g] = 0; rows[g] = 1;
          if (c > (uint32_t)kSortedHeavy) heavy[atomicAdd(&wsum[31], 1u)] = (uint16_t)g;
        }
      }
    }
    __syncthreads();
    // ---- 3. counting sort of the tile's values into shared memory
#pragma unroll
    for (int r = 0; r < kSortedRows; ++r) {
      if (gid[r] != kNone) {
        const uint32_t pos = start[gid[r]] + rank[r];
#pragma unroll
        for (int c = 0; c < NC; ++c) sval[c * T + pos] = val[r][c];
        if (p.any_mask) svalid[pos] = (uint8_t)vb[r];
      }
    }
    load_keys(tile + gridDim.x);
    __syncthreads();
    // ---- 4. owner-computes, one accumulator at a time (the type switch runs once per accumulator and tile)
    if (g0 < gend) {
      for (int a = 0; a < A; ++a) {
        const uint8_t op = p.acc[a].op;
        const int col    = p.acc[a].col;
        uint64_t* cells  = tab + (size_t)a * G;
        if (col < 0 || ((op == ACC_COUNT_VALID || op == ACC_COUNT_ALL) && p.col[col].mask == nullptr)) {
          uint32_t b = start[g0];   // COUNT of an unmasked column is the histogram (heavy groups included)
          for (uint32_t g = g0; g < gend; ++g) { const uint32_t e = start[g + 1]; if (e != b) cells[g] += e - b; b = e; }
        } else {
          const OwnerCtx x{start, cells, sval + col * T, svalid, p.col[col].mask != nullptr ? 1u << col : 0u, g0, gend};
          owner_acc(op, p.col[col].src, x);
        }
      }
    }
    // ---- 5. heavy groups: one warp each
    const uint32_t hn = wsum[31];   // (complete since the barrier after the scan)
    if (hn != 0) {
      for (uint32_t h = warp; h < hn; h += NT / 32) {
        const uint32_t g = heavy[h];
        const uint32_t b = start[g], e = start[g + 1];
        for (int a = 0; a < A; ++a) {
          const uint8_t op = p.acc[a].op;
          const int col    = p.acc[a].col;
          if (col < 0 || ((op == ACC_COUNT_VALID || op == ACC_COUNT_ALL) && p.col[col].mask == nullptr)) continue;  // done above
          uint64_t part = fold_group(op, p.col[col].src, sval + col * T, svalid, p.col[col].mask != nullptr ? 1u << col : 0u, b + lane, e, 32);
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) part = fold(op, part, __shfl_xor_sync(0xffffffffu, part, o));
          if (lane == 0) tab[a * G + g] = fold(op, tab[a * G + g], part);
        }
      }
      __syncthreads();
      if (tid == 0) wsum[31] = 0;   // (the next scan phase is a barrier away)
    }
  }
  __syncthreads();
  // ---- merge the CTA's table into the HBM table; CTAs start at different groups so they do not queue on one address
  for (uint32_t i = tid; i < G; i += NT) {
    const uint32_t g = (uint32_t)(((uint64_t)i + (uint64_t)blockIdx.x * 61u) % G);
    if (rows[g] == 0) continue;
    const uint64_t key  = p.hashed ? skey[g] : (p.key_base + g) ^ p.key_flip;
    const uint64_t slot = find_or_insert(p.a.table, key, 0);
    if (slot == kNoSlot) continue;
    for (int a = 0; a < A; ++a) atomic_fold(p.acc[a].op, p.acc[a].cells, slot, tab[a * G + g]);
  }
}

}  // namespace

size_t sorted_smem_bytes(uint32_t groups, int num_accs, int num_cols, bool hashed)
{
  return (size_t)num_accs * groups * 8 + (hashed ? (size_t)groups * 8 : 0) + (size_t)num_cols * kSortedTile * 8 + (size_t)groups * 4 + ((size_t)groups + 1) * 4 +
         (size_t)groups * 4 + 32 * 4 + (kSortedTile / kSortedHeavy) * 2 + kSortedTile + 16;
}

cudaError_t launch_aggregate_sorted(const SortedParams& p, int num_sms, cudaStream_t s)
{
  void (*kern)(const SortedParams) = nullptr;
  switch (p.num_cols) {
    case 0: case 1: kern = aggregate_sorted_kernel<1>; break;
    case 2: kern = aggregate_sorted_kernel<2>; break;
    case 3: kern = aggregate_sorted_kernel<3>; break;
    case 4: kern = aggregate_sorted_kernel<4>; break;
    default: return cudaErrorInvalidValue;
  }
  const size_t bytes = sorted_smem_bytes(p.groups, p.num_accs, p.num_cols < 1 ? 1 : p.num_cols, p.hashed != 0);
  cudaError_t e = allow_max_dynamic_smem(kern);
  if (e != cudaSuccess) return e;
  int per_sm = 1;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kSortedThreads, bytes);
  if (e != cudaSuccess) return e;
  if (per_sm < 1) return cudaErrorInvalidConfiguration;
  const int64_t ntiles = (p.a.num_rows + kSortedTile - 1) / kSortedTile;
  int64_t grid = (int64_t)num_sms * per_sm;
  if (grid > ntiles) grid = ntiles;
  if (grid < 1) grid = 1;
  kern<<<(unsigned)grid, kSortedThreads, bytes, s>>>(p);
  return cudaGetLastError();
}

}  // namespace bgb

// Prototype of the round-2 high-cardinality pipeline for the A9 shape (int64 key, int64 value, SUM + COUNT):
//   scatter level 1 (caller's SoA columns -> 16-byte {key, value} records in P1 buckets), TMA-fed (cp.async.bulk + mbarrier ring)
//   scatter level 2 (records of every level-1 bucket -> P2 sub-buckets), TMA-fed
//   slice kernel    (one CTA per final bucket: slice of the table in shared memory, DENSE group output, no HBM table)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/proto_hc tools/proto_hc.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <stdint.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z)
{
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__device__ __forceinline__ uint32_t slot_hash32(uint64_t k)
{
  const uint64_t h = k * 0x9E3779B97F4A7C15ull;
  uint32_t x = (uint32_t)(h >> 32) ^ (uint32_t)h;
  x *= 0x85EBCA6Bu; x ^= x >> 13;
  x *= 0xC2B2AE35u; x ^= x >> 16;
  return x;
}
__global__ void gen(uint64_t* k, uint64_t* v, int64_t n, uint64_t g)
{
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    k[i] = mix64(42 * 0x1000003ull + i + 0x9E3779B97F4A7C15ull) % g;
    v[i] = mix64(43 * 0x1000003ull + i + 0x9E3779B97F4A7C15ull) % 2001 - 1000;
  }
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(b)), "r"(c)); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity)
{
  asm volatile("{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" :: "r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

struct ScatterParams {
  const uint64_t* keys; const uint64_t* vals; int64_t n;          // SoA input (level 1)
  const ulonglong2* in; const uint32_t* seg_cursor; uint32_t seg_capacity, num_segs;  // AoS input (level 2)
  ulonglong2* out; uint32_t* cursor; uint32_t bucket_capacity;
  int log2_parts; int slot_shift; uint32_t capmask; uint32_t* status;
  uint32_t cursor_stride, seg_cursor_stride;  // in words: hot cursors sit on separate L2 slices
};

// T rows per tile, THREADS threads, NS input stages. Shared memory: ring NS x T x 16 | sorted T x 16 | counters.
// Software pipeline over the tiles of a CTA (2 barriers per tile):
//   top      warp 0 turns the reservation atomics of tile c-1 (issued one phase ago) into s_gbase / s_end
//   A(c)     bucket id + rank of every row (ATOMS.ADD with return); barrier
//   B(c)     warp 0: scan of the counters -> s_off, reservation atomics for tile c ISSUED (results consumed at the next top)
//   D(c-1)   all threads: coalesced write-out of the previous tile from `sorted`; barrier
//   C(c)     records of tile c into their sorted positions
template <int T, int THREADS, int NS, bool AOS_IN>
__global__ void __launch_bounds__(THREADS, 2) scatter_kernel(const __grid_constant__ ScatterParams p)
{
  constexpr int R = T / THREADS;  // rows per thread
  extern __shared__ __align__(128) unsigned char smem[];
  ulonglong2* ring   = reinterpret_cast<ulonglong2*>(smem);              // [NS][T] records, or per stage keys[T] | vals[T]
  ulonglong2* sorted = ring + NS * T;
  uint32_t* s_cnt    = reinterpret_cast<uint32_t*>(sorted + T);          // [256]
  uint32_t* s_off    = s_cnt + 256;                                      // [2][256]
  uint32_t* s_gbase  = s_off + 512;                                      // [2][256]
  uint32_t* s_end    = s_gbase + 512;                                    // [2][256]
  uint64_t* full     = reinterpret_cast<uint64_t*>(s_end + 512);         // [NS]
  const uint32_t t = threadIdx.x;
  const uint32_t P = 1u << p.log2_parts, pmask = P - 1;

  const uint32_t tiles_per_seg = AOS_IN ? (p.seg_capacity + T - 1) / T : 1;
  const int64_t ntiles = AOS_IN ? (int64_t)p.num_segs * tiles_per_seg : (p.n + T - 1) / T;

  auto tile_info = [&](int64_t tile, int64_t& row0, uint32_t& nrows, uint32_t& cbase) {
    if (AOS_IN) {
      const uint32_t seg = (uint32_t)(tile / tiles_per_seg), tt = (uint32_t)(tile % tiles_per_seg);
      uint32_t cnt = __ldg(p.seg_cursor + (size_t)seg * p.seg_cursor_stride);
      cnt = cnt < p.seg_capacity ? cnt : p.seg_capacity;
      nrows = tt * T >= cnt ? 0 : (cnt - tt * T < T ? cnt - tt * T : T);
      row0  = (int64_t)seg * p.seg_capacity + (int64_t)tt * T;
      cbase = seg << p.log2_parts;
    } else {
      row0  = tile * T;
      nrows = p.n - row0 < (int64_t)T ? (uint32_t)(p.n - row0) : T;
      cbase = 0;
    }
  };
  int64_t ptile = blockIdx.x;
  uint32_t pcount = 0;
  auto issue = [&]() {  // thread 0 only: load the next non-empty tile into stage pcount % NS
    while (ptile < ntiles) {
      int64_t row0; uint32_t nrows, cbase;
      tile_info(ptile, row0, nrows, cbase);
      ptile += gridDim.x;
      if (nrows == 0) continue;
      const int s = pcount % NS;
      unsigned char* dst = reinterpret_cast<unsigned char*>(ring + (size_t)s * T);
      if (AOS_IN) {
        mbar_expect(full + s, nrows * 16);
        tma_g2s(dst, p.in + row0, nrows * 16, full + s);
      } else {
        const uint32_t nb = ((nrows + 1) & ~1u) * 8;
        mbar_expect(full + s, 2 * nb);
        tma_g2s(dst, p.keys + row0, nb, full + s);
        tma_g2s(dst + T * 8, p.vals + row0, nb, full + s);
      }
      ++pcount;
      return;
    }
  };
  if (t == 0) { for (int s = 0; s < NS; ++s) mbar_init(full + s, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (t < 256) s_cnt[t] = 0;
  __syncthreads();
  if (t == 0) for (int s = 0; s < NS; ++s) issue();

  uint32_t gres[8], cn[8], ex[8];   // warp 0: reservation of the tile in flight
  uint32_t prev_rows = 0, prev_cbase = 0;
  auto publish = [&](uint32_t buf, uint32_t cbase) {  // warp 0
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const uint32_t i = t * 8 + j;
      if (i < P) {
        const uint32_t g    = gres[j];
        const uint32_t room = g < p.bucket_capacity ? p.bucket_capacity - g : 0;
        s_gbase[buf * 256 + i] = (cbase + i) * p.bucket_capacity + g - ex[j];
        s_end[buf * 256 + i]   = ex[j] + (cn[j] < room ? cn[j] : room);
        if (cn[j] > room) p.status[0] = 1;  // prototype: overflow rows are dropped and flagged
      }
    }
  };
  auto write_out = [&](uint32_t buf, uint32_t nrows) {
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const uint32_t i = j * THREADS + t;
      if (i < nrows) {
        const ulonglong2 rec = sorted[i];
        const uint32_t b = ((slot_hash32(rec.x) & p.capmask) >> p.slot_shift) & pmask;
        if (i < s_end[buf * 256 + b]) p.out[s_gbase[buf * 256 + b] + i] = rec;
      }
    }
  };

  uint32_t ccount = 0;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    int64_t row0; uint32_t nrows, cbase;
    tile_info(tile, row0, nrows, cbase);
    if (nrows == 0) continue;
    const uint32_t buf = ccount & 1;
    if (t < 32 && ccount > 0) publish(buf ^ 1, prev_cbase);
    const int s = ccount % NS;
    mbar_wait(full + s, (ccount / NS) & 1);
    const unsigned char* st = reinterpret_cast<const unsigned char*>(ring + (size_t)s * T);

    // ---- A
    uint64_t key[R];
    uint32_t pid[R], rank[R];
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const uint32_t r = j * THREADS + t;
      key[j] = 0; pid[j] = 0; rank[j] = 0;
      if (r < nrows) {
        key[j]  = AOS_IN ? reinterpret_cast<const ulonglong2*>(st)[r].x : reinterpret_cast<const uint64_t*>(st)[r];
        pid[j]  = ((slot_hash32(key[j]) & p.capmask) >> p.slot_shift) & pmask;
        rank[j] = atomicAdd(&s_cnt[pid[j]], 1u);
      }
    }
    __syncthreads();
    // the stage of the PREVIOUS tile is free now (everybody is past its phase C)
    if (t == 0 && ccount > 0) issue();
    // ---- B (warp 0) and D of the previous tile
    if (t < 32) {
      uint32_t tot = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) { const uint32_t i = t * 8 + j; cn[j] = i < P ? s_cnt[i] : 0; ex[j] = tot; tot += cn[j]; if (i < P) s_cnt[i] = 0; }
      uint32_t incl = tot;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) { uint32_t u = __shfl_up_sync(0xffffffffu, incl, d); if (t >= (uint32_t)d) incl += u; }
      const uint32_t base = incl - tot;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint32_t i = t * 8 + j;
        ex[j] += base;
        if (i < P) s_off[buf * 256 + i] = ex[j];
        gres[j] = cn[j] ? atomicAdd(p.cursor + (size_t)(cbase + i) * p.cursor_stride, cn[j]) : 0;
      }
    }
    if (ccount > 0) write_out(buf ^ 1, prev_rows);
    __syncthreads();
    // ---- C
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const uint32_t r = j * THREADS + t;
      if (r < nrows) {
        const uint64_t v = AOS_IN ? reinterpret_cast<const ulonglong2*>(st)[r].y : reinterpret_cast<const uint64_t*>(st + T * 8)[r];
        sorted[s_off[buf * 256 + pid[j]] + rank[j]] = make_ulonglong2(key[j], v);
      }
    }
    prev_rows = nrows; prev_cbase = cbase;
    ++ccount;
  }
  if (ccount > 0) {
    if (t < 32) publish((ccount - 1) & 1, prev_cbase);
    __syncthreads();
    write_out((ccount - 1) & 1, prev_rows);
  }
}

struct SliceParams {
  const ulonglong2* rec; const uint32_t* cursor; uint32_t bucket_capacity, num_buckets;
  int log2_slice;
  uint64_t* out_key; int64_t* out_sum; uint32_t* out_cnt; unsigned long long* gcount; uint32_t* status;
};

// ---------------------------------------------------------------- scatter, variant 2: ONE 64 KB stage per CTA, 3 CTAs per SM
// The stage keeps the tile until it has been written out; phase C only stores a 16-bit permutation (position -> row), phase D
// gathers key and value of the row from the stage. No sorted copy, so T = 4096 rows fit three CTAs per SM (runs of 256 B at
// 256 buckets). The CTAs of an SM interleave: while one waits for its TMA load the others rank / write.
template <int T, int THREADS, bool AOS_IN>
__global__ void __launch_bounds__(THREADS, 3) scatter2_kernel(const __grid_constant__ ScatterParams p)
{
  constexpr int R = T / THREADS;
  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char* stage = smem;                                                    // T x 16
  uint16_t* s_idx    = reinterpret_cast<uint16_t*>(smem + (size_t)T * 16);        // [T]
  uint32_t* s_cnt    = reinterpret_cast<uint32_t*>(s_idx + T);                    // [256]
  uint32_t* s_off    = s_cnt + 256;
  uint32_t* s_gbase  = s_off + 256;
  uint32_t* s_end    = s_gbase + 256;
  uint64_t* full     = reinterpret_cast<uint64_t*>(s_end + 256);
  const uint32_t t = threadIdx.x;
  const uint32_t P = 1u << p.log2_parts, pmask = P - 1;
  const uint32_t tiles_per_seg = AOS_IN ? (p.seg_capacity + T - 1) / T : 1;
  const int64_t ntiles = AOS_IN ? (int64_t)p.num_segs * tiles_per_seg : (p.n + T - 1) / T;
  if (t == 0) { mbar_init(full, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (t < 256) s_cnt[t] = 0;
  __syncthreads();
  uint32_t phase = 0;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    int64_t row0; uint32_t nrows, cbase;
    if (AOS_IN) {
      const uint32_t seg = (uint32_t)(tile / tiles_per_seg), tt = (uint32_t)(tile % tiles_per_seg);
      uint32_t cnt = __ldg(p.seg_cursor + (size_t)seg * p.seg_cursor_stride);
      cnt = cnt < p.seg_capacity ? cnt : p.seg_capacity;
      nrows = tt * T >= cnt ? 0 : (cnt - tt * T < T ? cnt - tt * T : T);
      row0  = (int64_t)seg * p.seg_capacity + (int64_t)tt * T;
      cbase = seg << p.log2_parts;
    } else {
      row0  = tile * T;
      nrows = p.n - row0 < (int64_t)T ? (uint32_t)(p.n - row0) : T;
      cbase = 0;
    }
    if (nrows == 0) continue;
    if (t == 0) {
      if (AOS_IN) { mbar_expect(full, nrows * 16); tma_g2s(stage, p.in + row0, nrows * 16, full); }
      else {
        const uint32_t nb = ((nrows + 1) & ~1u) * 8;
        mbar_expect(full, 2 * nb);
        tma_g2s(stage, p.keys + row0, nb, full);
        tma_g2s(stage + T * 8, p.vals + row0, nb, full);
      }
    }
    mbar_wait(full, phase); phase ^= 1;
    auto key_of = [&](uint32_t r) { return AOS_IN ? reinterpret_cast<const ulonglong2*>(stage)[r].x : reinterpret_cast<const uint64_t*>(stage)[r]; };
    auto val_of = [&](uint32_t r) { return AOS_IN ? reinterpret_cast<const ulonglong2*>(stage)[r].y : reinterpret_cast<const uint64_t*>(stage + T * 8)[r]; };
    // ---- A
    uint32_t pr[R];  // pid << 16 | rank
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const uint32_t r = j * THREADS + t;
      pr[j] = 0;
      if (r < nrows) {
        const uint32_t pid = ((slot_hash32(key_of(r)) & p.capmask) >> p.slot_shift) & pmask;
        pr[j] = (pid << 16) | atomicAdd(&s_cnt[pid], 1u);
      }
    }
    __syncthreads();
    // ---- B (warp 0)
    uint32_t gres[8], cn[8], ex[8];
    if (t < 32) {
      uint32_t tot = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) { const uint32_t i = t * 8 + j; cn[j] = i < P ? s_cnt[i] : 0; ex[j] = tot; tot += cn[j]; if (i < P) s_cnt[i] = 0; }
      uint32_t incl = tot;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) { uint32_t u = __shfl_up_sync(0xffffffffu, incl, d); if (t >= (uint32_t)d) incl += u; }
      const uint32_t base = incl - tot;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint32_t i = t * 8 + j;
        ex[j] += base;
        if (i < P) s_off[i] = ex[j];
        gres[j] = cn[j] ? atomicAdd(p.cursor + (size_t)(cbase + i) * p.cursor_stride, cn[j]) : 0;
      }
    }
    __syncthreads();
    // ---- C: permutation
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const uint32_t r = j * THREADS + t;
      if (r < nrows) s_idx[s_off[pr[j] >> 16] + (pr[j] & 0xffffu)] = (uint16_t)r;
    }
    if (t < 32) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint32_t i = t * 8 + j;
        if (i < P) {
          const uint32_t g    = gres[j];
          const uint32_t room = g < p.bucket_capacity ? p.bucket_capacity - g : 0;
          s_gbase[i] = (cbase + i) * p.bucket_capacity + g - ex[j];
          s_end[i]   = ex[j] + (cn[j] < room ? cn[j] : room);
          if (cn[j] > room) p.status[0] = 1;
        }
      }
    }
    __syncthreads();
    // ---- D
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const uint32_t i = j * THREADS + t;
      if (i < nrows) {
        const uint32_t r = s_idx[i];
        const uint64_t k = key_of(r), v = val_of(r);
        const uint32_t b = ((slot_hash32(k) & p.capmask) >> p.slot_shift) & pmask;
        if (i < s_end[b]) p.out[s_gbase[b] + i] = make_ulonglong2(k, v);
      }
    }
    __syncthreads();  // the stage is free
  }
}

// ---------------------------------------------------------------- slice kernel 2: carry-propagating 2 x 32-bit sums, register prefetch
template <int THREADS>
__global__ void __launch_bounds__(THREADS, 2) slice2_kernel(const __grid_constant__ SliceParams p)
{
  extern __shared__ __align__(16) unsigned char smem[];
  const uint32_t S = 1u << p.log2_slice, smask = S - 1;
  unsigned long long* s_key = reinterpret_cast<unsigned long long*>(smem);
  uint32_t* s_cnt = reinterpret_cast<uint32_t*>(s_key + S);
  uint32_t* s_lo  = s_cnt + S;
  uint32_t* s_hi  = s_lo + S;
  __shared__ uint32_t s_warp[THREADS / 32];
  __shared__ unsigned long long s_base;
  const uint32_t t = threadIdx.x;
  constexpr unsigned long long kEmpty = ~0ull;
  constexpr int U = 4;
  ulonglong2 nxt[U];
  uint32_t fb = blockIdx.x;
  uint32_t n = 0;
  if (fb < p.num_buckets) { n = __ldg(p.cursor + fb); n = n < p.bucket_capacity ? n : p.bucket_capacity; }
  const ulonglong2* rows = p.rec + (uint64_t)fb * p.bucket_capacity;
#pragma unroll
  for (int j = 0; j < U; ++j) { const uint32_t r = j * THREADS + t; if (r < n) nxt[j] = __ldcs(rows + r); }
  for (; fb < p.num_buckets; ) {
    for (uint32_t i = t; i < S / 2; i += THREADS) reinterpret_cast<uint4*>(s_key)[i] = make_uint4(~0u, ~0u, ~0u, ~0u);
    for (uint32_t i = t; i < 3 * S / 4; i += THREADS) reinterpret_cast<uint4*>(s_cnt)[i] = make_uint4(0, 0, 0, 0);
    __syncthreads();
    const uint32_t fb_next = fb + gridDim.x;
    uint32_t n_next = 0;
    if (fb_next < p.num_buckets) { n_next = __ldg(p.cursor + fb_next); n_next = n_next < p.bucket_capacity ? n_next : p.bucket_capacity; }
    const ulonglong2* rows_next = p.rec + (uint64_t)fb_next * p.bucket_capacity;
    for (uint32_t r0 = 0; r0 < n; r0 += THREADS * U) {
      ulonglong2 rec[U];
#pragma unroll
      for (int j = 0; j < U; ++j) rec[j] = nxt[j];
      const uint32_t r1 = r0 + THREADS * U;
      if (r1 < n) {
#pragma unroll
        for (int j = 0; j < U; ++j) { const uint32_t r = r1 + j * THREADS + t; if (r < n) nxt[j] = __ldcs(rows + r); }
      } else {
#pragma unroll
        for (int j = 0; j < U; ++j) { const uint32_t r = j * THREADS + t; if (r < n_next) nxt[j] = __ldcs(rows_next + r); }
      }
#pragma unroll
      for (int j = 0; j < U; ++j) {
        const uint32_t r = r0 + j * THREADS + t;
        const bool active = r < n;
        const unsigned long long k = rec[j].x;
        uint32_t slot = slot_hash32(k) & smask;
        bool pending = active;
        uint32_t probes = 0;
        while (__ballot_sync(0xffffffffu, pending)) {  // warp-converged probing: one probe step per iteration for the lanes still looking
          if (pending) {
            const unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(s_key + slot);
            if (cur == k) pending = false;
            else {
              bool next = true;
              if (cur == kEmpty) {
                const unsigned long long old = atomicCAS(s_key + slot, kEmpty, k);
                if (old == kEmpty || old == k) { pending = false; next = false; }
              }
              if (next) { slot = (slot + 1) & smask; if (++probes > smask) { p.status[1] = 1; pending = false; } }
            }
          }
        }
        if (active) {
          atomicAdd(s_cnt + slot, 1u);
          const uint32_t vlo = (uint32_t)rec[j].y, vhi = (uint32_t)(rec[j].y >> 32);
          const uint32_t old = atomicAdd(s_lo + slot, vlo);
          const uint32_t add = vhi + ((uint32_t)(old + vlo) < old ? 1u : 0u);
          if (add) atomicAdd(s_hi + slot, add);
        }
      }
    }
    if (n == 0) {  // nothing consumed the prefetch: reload for the next bucket
#pragma unroll
      for (int j = 0; j < U; ++j) { const uint32_t r = j * THREADS + t; if (r < n_next) nxt[j] = __ldcs(rows_next + r); }
    }
    __syncthreads();
    const uint32_t own = S / THREADS;
    uint32_t mine = 0;
    for (uint32_t j = 0; j < own; ++j) mine += s_key[t + j * THREADS] != kEmpty;
    uint32_t incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { uint32_t u = __shfl_up_sync(0xffffffffu, incl, d); if ((t & 31) >= (uint32_t)d) incl += u; }
    if ((t & 31) == 31) s_warp[t >> 5] = incl;
    __syncthreads();
    uint32_t woff = 0, total = 0;
#pragma unroll
    for (int w = 0; w < THREADS / 32; ++w) { const uint32_t u = s_warp[w]; if (w < (int)(t >> 5)) woff += u; total += u; }
    if (t == 0) s_base = atomicAdd(p.gcount, (unsigned long long)total);
    __syncthreads();
    uint64_t o = s_base + woff + incl - mine;
    for (uint32_t j = 0; j < own; ++j) {
      const uint32_t l = t + j * THREADS;
      const unsigned long long k = s_key[l];
      if (k != kEmpty) {
        p.out_key[o] = k;
        p.out_sum[o] = (int64_t)(((uint64_t)s_hi[l] << 32) | s_lo[l]);
        p.out_cnt[o] = s_cnt[l];
        ++o;
      }
    }
    __syncthreads();
    fb = fb_next; n = n_next; rows = rows_next;
  }
}

// ---------------------------------------------------------------- slice kernel: SUM(int64) + COUNT, dense output

template <int THREADS>
__global__ void __launch_bounds__(THREADS, 2) slice_kernel(const __grid_constant__ SliceParams p)
{
  extern __shared__ __align__(16) unsigned char smem[];
  const uint32_t S = 1u << p.log2_slice, smask = S - 1;
  unsigned long long* s_key = reinterpret_cast<unsigned long long*>(smem);
  uint32_t* s_cnt = reinterpret_cast<uint32_t*>(s_key + S);
  uint32_t* s_l0  = s_cnt + S;
  uint32_t* s_l1  = s_l0 + S;
  uint32_t* s_l2  = s_l1 + S;
  __shared__ uint32_t s_warp[THREADS / 32];
  __shared__ unsigned long long s_base;
  const uint32_t t = threadIdx.x;
  constexpr unsigned long long kEmpty = ~0ull;

  for (uint32_t fb = blockIdx.x; fb < p.num_buckets; fb += gridDim.x) {
    uint32_t n = __ldg(p.cursor + fb);
    n = n < p.bucket_capacity ? n : p.bucket_capacity;
    const ulonglong2* rows = p.rec + (uint64_t)fb * p.bucket_capacity;
    for (uint32_t i = t; i < S / 2; i += THREADS) reinterpret_cast<uint4*>(s_key)[i] = make_uint4(~0u, ~0u, ~0u, ~0u);
    for (uint32_t i = t; i < S; i += THREADS) reinterpret_cast<uint4*>(s_cnt)[i] = make_uint4(0, 0, 0, 0);  // cnt + 3 limb arrays = 4 S words
    __syncthreads();
    for (uint32_t r0 = 0; r0 < n; r0 += THREADS * 4) {
      ulonglong2 rec[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) { const uint32_t r = r0 + j * THREADS + t; if (r < n) rec[j] = __ldcs(rows + r); }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint32_t r = r0 + j * THREADS + t;
        if (r >= n) continue;
        const unsigned long long k = rec[j].x;
        uint32_t slot = slot_hash32(k) & smask;
        for (uint32_t probes = 0;; ++probes) {
          const unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(s_key + slot);
          if (cur == k) break;
          if (cur == kEmpty) {
            const unsigned long long old = atomicCAS(s_key + slot, kEmpty, k);
            if (old == kEmpty || old == k) break;
          }
          slot = (slot + 1) & smask;
          if (probes > smask) { p.status[1] = 1; break; }
        }
        atomicAdd(s_cnt + slot, 1u);
        const int64_t v   = (int64_t)rec[j].y;
        const int32_t d0  = (int16_t)v;
        const int64_t r1  = (v - d0) >> 16;
        const int32_t d1  = (int16_t)r1;
        const uint32_t r2 = (uint32_t)((r1 - d1) >> 16);
        if (d0) atomicAdd(s_l0 + slot, (uint32_t)d0);
        if (d1) atomicAdd(s_l1 + slot, (uint32_t)d1);
        if (r2) atomicAdd(s_l2 + slot, r2);
      }
    }
    __syncthreads();
    // ---- dense output: every thread owns S / THREADS consecutive slots
    constexpr int kMaxOwn = 16;
    const uint32_t own = S / THREADS;
    uint32_t mine = 0;
    for (uint32_t j = 0; j < own && j < kMaxOwn; ++j) mine += s_key[t * own + j] != kEmpty;
    uint32_t incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { uint32_t u = __shfl_up_sync(0xffffffffu, incl, d); if ((t & 31) >= (uint32_t)d) incl += u; }
    if ((t & 31) == 31) s_warp[t >> 5] = incl;
    __syncthreads();
    uint32_t woff = 0, total = 0;
#pragma unroll
    for (int w = 0; w < THREADS / 32; ++w) { const uint32_t u = s_warp[w]; if (w < (int)(t >> 5)) woff += u; total += u; }
    if (t == 0) s_base = atomicAdd(p.gcount, (unsigned long long)total);
    __syncthreads();
    uint64_t o = s_base + woff + incl - mine;
    for (uint32_t j = 0; j < own && j < kMaxOwn; ++j) {
      const uint32_t l = t * own + j;
      const unsigned long long k = s_key[l];
      if (k != kEmpty) {
        p.out_key[o] = k;
        p.out_sum[o] = (int64_t)((uint64_t)(int64_t)(int32_t)s_l0[l] + ((uint64_t)(int64_t)(int32_t)s_l1[l] << 16) + ((uint64_t)s_l2[l] << 32));
        p.out_cnt[o] = s_cnt[l];
        ++o;
      }
    }
    __syncthreads();
  }
}

__global__ void check_kernel(const int64_t* sum, const uint32_t* cnt, uint64_t g, unsigned long long* out)
{
  unsigned long long s = 0, c = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < g; i += (uint64_t)gridDim.x * blockDim.x) { s += (unsigned long long)sum[i]; c += cnt[i]; }
  atomicAdd(out, s); atomicAdd(out + 1, c);
}
__global__ void colsum_kernel(const uint64_t* v, int64_t n, unsigned long long* out)
{
  unsigned long long s = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) s += v[i];
  atomicAdd(out, s);
}

template <typename F> static float time_ms(F f, int reps = 3)
{
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int r = 0; r < reps; ++r) { CK(cudaEventRecord(e0)); f(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms; }
  CK(cudaGetLastError());
  return best;
}

template <int T, int THREADS, int NS>
static void run(const uint64_t* k, const uint64_t* v, int64_t n, int log2cap, int b1, int b2, int log2_slice)
{
  const uint32_t P1 = 1u << b1, NB = 1u << (b1 + b2);
  const uint32_t C1 = (uint32_t)((double)n / P1 * 1.03) + 4096, C2 = ((uint32_t)((double)n / NB * 1.27) + 256 + 15) & ~15u;
  ulonglong2 *l1, *l2; uint32_t *cur1, *cur2, *status;
  CK(cudaMalloc(&l1, (size_t)P1 * C1 * 16)); CK(cudaMalloc(&l2, (size_t)NB * C2 * 16));
  CK(cudaMalloc(&cur1, P1 * 4 * 64)); CK(cudaMalloc(&cur2, NB * 4)); CK(cudaMalloc(&status, 64));
  CK(cudaMemset(status, 0, 64));
  uint64_t* out_key; int64_t* out_sum; uint32_t* out_cnt; unsigned long long* gcount;
  const uint64_t gmax = 1ull << (log2cap - 1);
  CK(cudaMalloc(&out_key, gmax * 8)); CK(cudaMalloc(&out_sum, gmax * 8)); CK(cudaMalloc(&out_cnt, gmax * 4)); CK(cudaMalloc(&gcount, 32));
  const size_t smem = (size_t)(NS + 1) * T * 16 + 7 * 256 * 4 + 64;
  CK(cudaFuncSetAttribute(scatter_kernel<T, THREADS, NS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CK(cudaFuncSetAttribute(scatter_kernel<T, THREADS, NS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  ScatterParams s1{}; s1.keys = k; s1.vals = v; s1.n = n; s1.out = l1; s1.cursor = cur1; s1.bucket_capacity = C1;
  s1.log2_parts = b1; s1.slot_shift = log2cap - b1; s1.capmask = (uint32_t)((1ull << log2cap) - 1); s1.status = status; s1.cursor_stride = 64; s1.seg_cursor_stride = 1;
  ScatterParams s2 = s1; s2.in = l1; s2.seg_cursor = cur1; s2.seg_capacity = C1; s2.num_segs = P1; s2.out = l2; s2.cursor = cur2; s2.bucket_capacity = C2;
  s2.log2_parts = b2; s2.slot_shift = log2cap - b1 - b2; s2.cursor_stride = 1; s2.seg_cursor_stride = 64;
  SliceParams sl{}; sl.rec = l2; sl.cursor = cur2; sl.bucket_capacity = C2; sl.num_buckets = NB; sl.log2_slice = log2_slice;
  sl.out_key = out_key; sl.out_sum = out_sum; sl.out_cnt = out_cnt; sl.gcount = gcount; sl.status = status;
  const size_t ssm = (size_t)24 << log2_slice;
  CK(cudaFuncSetAttribute(slice_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ssm));
  int occ = 0; CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, scatter_kernel<T, THREADS, NS, false>, THREADS, smem));
  const int grid = 148 * occ;
  int socc = 0; CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&socc, slice_kernel<512>, 512, ssm));
  printf("  occupancy: scatter %d CTAs/SM, slices %d CTAs/SM\n", occ, socc);
  float t1 = time_ms([&] { cudaMemsetAsync(cur1, 0, P1 * 4 * 64); scatter_kernel<T, THREADS, NS, false><<<grid, THREADS, smem>>>(s1); });
  float t2 = time_ms([&] { cudaMemsetAsync(cur2, 0, NB * 4); scatter_kernel<T, THREADS, NS, true><<<grid, THREADS, smem>>>(s2); });
  float t3 = time_ms([&] { cudaMemsetAsync(gcount, 0, 32); slice_kernel<512><<<NB < 148u * 8 ? NB : 148 * 8, 512, ssm>>>(sl); });
  unsigned long long h[4] = {0, 0, 0, 0}, g = 0; uint32_t st[2];
  CK(cudaMemcpy(&g, gcount, 8, cudaMemcpyDeviceToHost));
  CK(cudaMemset(gcount + 1, 0, 24));
  check_kernel<<<148 * 4, 256>>>(out_sum, out_cnt, g, gcount + 1);
  colsum_kernel<<<148 * 4, 256>>>(v, n, gcount + 3);
  CK(cudaMemcpy(h, gcount, 32, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(st, status, 8, cudaMemcpyDeviceToHost));
  printf("T=%d thr=%d NS=%d b1=%d b2=%d slice=2^%d: scatter1 %.3f ms (%.2f TB/s)  scatter2 %.3f ms  slices %.3f ms  total %.3f | groups %llu  sum(cnt) %llu %s  sum(sum) %s  status %u %u\n",
         T, THREADS, NS, b1, b2, log2_slice, t1, 32.0 * n / t1 / 1e9, t2, t3, t1 + t2 + t3, g, h[2], h[2] == (unsigned long long)n ? "ok" : "BAD", h[1] == h[3] ? "ok" : "BAD", st[0], st[1]);
  cudaFree(l1); cudaFree(l2); cudaFree(cur1); cudaFree(cur2); cudaFree(status); cudaFree(out_key); cudaFree(out_sum); cudaFree(out_cnt); cudaFree(gcount);
}

// ---------------------------------------------------------------- slice kernel 3: TMA chunk ring + per-lane row queue
// One CTA per SM (or two); the bucket's rows arrive in CH-row chunks through a ring of NS stages (cp.async.bulk + mbarrier,
// refilled across bucket boundaries). A lane walks ITS rows of a chunk (row = q * THREADS + t) as one loop with a probe step
// per iteration: the warp runs max-over-lanes of the SUMMED probe lengths, not the sum of per-row maxima.
__device__ __forceinline__ void mbar_arrive(uint64_t* b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(b)) : "memory"); }
__device__ __forceinline__ uint32_t atoms_add_ret(uint32_t addr, uint32_t v) { uint32_t o; asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(o) : "r"(addr), "r"(v) : "memory"); return o; }
__device__ __forceinline__ void reds_add(uint32_t addr, uint32_t v) { asm volatile("red.shared.add.u32 [%0], %1;" :: "r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned long long atoms_cas64(uint32_t addr, unsigned long long c, unsigned long long n)
{ unsigned long long o; asm volatile("atom.shared.cas.b64 %0, [%1], %2, %3;" : "=l"(o) : "r"(addr), "l"(c), "l"(n) : "memory"); return o; }
__device__ __forceinline__ unsigned long long lds64(uint32_t addr) { unsigned long long o; asm volatile("ld.volatile.shared.u64 %0, [%1];" : "=l"(o) : "r"(addr) : "memory"); return o; }

template <int THREADS, int CH, int NS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) slice3_kernel(const __grid_constant__ SliceParams p)
{
  constexpr int Q = CH / THREADS;
  extern __shared__ __align__(128) unsigned char smem[];
  const uint32_t S = 1u << p.log2_slice, smask = S - 1;
  ulonglong2* ring = reinterpret_cast<ulonglong2*>(smem);                       // [NS][CH]
  unsigned long long* s_key = reinterpret_cast<unsigned long long*>(ring + NS * CH);
  uint32_t* s_cnt = reinterpret_cast<uint32_t*>(s_key + S);
  uint32_t* s_lo  = s_cnt + S;
  uint32_t* s_hi  = s_lo + S;
  uint64_t* full  = reinterpret_cast<uint64_t*>(s_hi + S);
  uint64_t* empty = full + NS;
  __shared__ uint32_t s_warp[THREADS / 32];
  __shared__ unsigned long long s_base;
  const uint32_t t = threadIdx.x;
  constexpr unsigned long long kEmpty = ~0ull;
  const uint32_t a_key = smem_u32(s_key), a_cnt = smem_u32(s_cnt), a_lo = smem_u32(s_lo), a_hi = smem_u32(s_hi);

  if (t == 0) {
    for (int s = 0; s < NS; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, THREADS / 32); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  // producer state (thread 0): walks (bucket, chunk) pairs of this CTA
  uint32_t pfb = blockIdx.x, pch = 0, pn = 0, pcount = 0;
  auto pbucket = [&]() { pn = 0; if (pfb < p.num_buckets) { pn = __ldg(p.cursor + pfb); pn = pn < p.bucket_capacity ? pn : p.bucket_capacity; } };
  auto issue = [&]() {
    for (;;) {
      if (pfb >= p.num_buckets) return;
      if (pch * CH >= pn) { pfb += gridDim.x; pch = 0; pbucket(); continue; }
      const uint32_t rows = pn - pch * CH < (uint32_t)CH ? pn - pch * CH : CH;
      const int s = pcount % NS;
      if (pcount >= NS) mbar_wait(empty + s, ((pcount / NS) - 1) & 1);
      mbar_expect(full + s, rows * 16);
      tma_g2s(ring + (size_t)s * CH, p.rec + (uint64_t)pfb * p.bucket_capacity + (uint64_t)pch * CH, rows * 16, full + s);
      ++pcount; ++pch;
      return;
    }
  };
  if (t == 0) { pbucket(); for (int s = 0; s < NS - 1; ++s) issue(); }
  uint32_t ccount = 0;
  for (uint32_t fb = blockIdx.x; fb < p.num_buckets; fb += gridDim.x) {
    uint32_t n = __ldg(p.cursor + fb);
    n = n < p.bucket_capacity ? n : p.bucket_capacity;
    for (uint32_t i = t; i < S / 2; i += THREADS) reinterpret_cast<uint4*>(s_key)[i] = make_uint4(~0u, ~0u, ~0u, ~0u);
    for (uint32_t i = t; i < 3 * S / 4; i += THREADS) reinterpret_cast<uint4*>(s_cnt)[i] = make_uint4(0, 0, 0, 0);
    __syncthreads();
    for (uint32_t c0 = 0; c0 < n; c0 += CH) {
      const uint32_t rows = n - c0 < (uint32_t)CH ? n - c0 : CH;
      if (t == 0) issue();  // keeps NS - 1 chunks in flight behind the one being consumed
      const int s = ccount % NS;
      mbar_wait(full + s, (ccount / NS) & 1);
      const ulonglong2* st = ring + (size_t)s * CH;
      // my rows of the chunk: t, t + THREADS, ... (a prefix of q = 0 .. Q-1)
      uint32_t r = t;
      ulonglong2 rec = make_ulonglong2(0, 0);
      uint32_t slot = 0, probes = 0;
      bool have = r < rows;
      if (have) { rec = st[r]; slot = slot_hash32(rec.x) & smask; }
      while (__ballot_sync(0xffffffffu, have)) {
        if (have) {
          const unsigned long long cur = lds64(a_key + slot * 8);
          bool hit = cur == rec.x;
          if (!hit && cur == kEmpty) { const unsigned long long old = atoms_cas64(a_key + slot * 8, kEmpty, rec.x); hit = old == kEmpty || old == rec.x; }
          if (hit) {
            reds_add(a_cnt + slot * 4, 1u);
            const uint32_t vlo = (uint32_t)rec.y, vhi = (uint32_t)(rec.y >> 32);
            const uint32_t old = atoms_add_ret(a_lo + slot * 4, vlo);
            const uint32_t add = vhi + ((uint32_t)(old + vlo) < old ? 1u : 0u);
            if (add) reds_add(a_hi + slot * 4, add);
            r += THREADS; probes = 0;
            have = r < rows;
            if (have) { rec = st[r]; slot = slot_hash32(rec.x) & smask; }
          } else {
            slot = (slot + 1) & smask;
            if (++probes > smask) { have = false; probes = 0x80000000u; }
          }
        }
      }
      if (probes == 0x80000000u) p.status[1] = 1;
      __syncwarp();
      if ((t & 31) == 0) mbar_arrive(empty + s);
      ++ccount;
    }
    __syncthreads();
    const uint32_t own = S / THREADS;
    uint32_t mine = 0;
    for (uint32_t j = 0; j < own; ++j) mine += s_key[t + j * THREADS] != kEmpty;
    uint32_t incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { uint32_t u = __shfl_up_sync(0xffffffffu, incl, d); if ((t & 31) >= (uint32_t)d) incl += u; }
    if ((t & 31) == 31) s_warp[t >> 5] = incl;
    __syncthreads();
    uint32_t woff = 0, total = 0;
#pragma unroll
    for (int w = 0; w < THREADS / 32; ++w) { const uint32_t u = s_warp[w]; if (w < (int)(t >> 5)) woff += u; total += u; }
    if (t == 0) s_base = atomicAdd(p.gcount, (unsigned long long)total);
    __syncthreads();
    uint64_t o = s_base + woff + incl - mine;
    for (uint32_t j = 0; j < own; ++j) {
      const uint32_t l = t + j * THREADS;
      const unsigned long long k = s_key[l];
      if (k != kEmpty) {
        p.out_key[o] = k;
        p.out_sum[o] = (int64_t)(((uint64_t)s_hi[l] << 32) | s_lo[l]);
        p.out_cnt[o] = s_cnt[l];
        ++o;
      }
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------- slice kernel 4: register prefetch, unrolled first probe for 4 rows
// (ILP), one converged clean-up loop for the rows that did not resolve at their home slot, PTX shared-memory atomics
template <int THREADS>
__global__ void __launch_bounds__(THREADS, 2) slice4_kernel(const __grid_constant__ SliceParams p)
{
  extern __shared__ __align__(16) unsigned char smem[];
  const uint32_t S = 1u << p.log2_slice, smask = S - 1;
  unsigned long long* s_key = reinterpret_cast<unsigned long long*>(smem);
  uint32_t* s_cnt = reinterpret_cast<uint32_t*>(s_key + S);
  uint32_t* s_lo  = s_cnt + S;
  uint32_t* s_hi  = s_lo + S;
  __shared__ uint32_t s_warp[THREADS / 32];
  __shared__ unsigned long long s_base;
  const uint32_t t = threadIdx.x;
  const uint32_t a_key = smem_u32(s_key), a_cnt = smem_u32(s_cnt), a_lo = smem_u32(s_lo), a_hi = smem_u32(s_hi);
  constexpr unsigned long long kEmpty = ~0ull;
  constexpr int U = 4;
  ulonglong2 nxt[U];
  uint32_t fb = blockIdx.x;
  uint32_t n = 0;
  if (fb < p.num_buckets) { n = __ldg(p.cursor + fb); n = n < p.bucket_capacity ? n : p.bucket_capacity; }
  const ulonglong2* rows = p.rec + (uint64_t)fb * p.bucket_capacity;
#pragma unroll
  for (int j = 0; j < U; ++j) { const uint32_t r = j * THREADS + t; if (r < n) nxt[j] = __ldcs(rows + r); }
  bool overflow = false;
  for (; fb < p.num_buckets; ) {
    for (uint32_t i = t; i < S / 2; i += THREADS) reinterpret_cast<uint4*>(s_key)[i] = make_uint4(~0u, ~0u, ~0u, ~0u);
    for (uint32_t i = t; i < 3 * S / 4; i += THREADS) reinterpret_cast<uint4*>(s_cnt)[i] = make_uint4(0, 0, 0, 0);
    __syncthreads();
    const uint32_t fb_next = fb + gridDim.x;
    uint32_t n_next = 0;
    if (fb_next < p.num_buckets) { n_next = __ldg(p.cursor + fb_next); n_next = n_next < p.bucket_capacity ? n_next : p.bucket_capacity; }
    const ulonglong2* rows_next = p.rec + (uint64_t)fb_next * p.bucket_capacity;
    for (uint32_t r0 = 0; r0 < n; r0 += THREADS * U) {
      ulonglong2 rec[U];
#pragma unroll
      for (int j = 0; j < U; ++j) rec[j] = nxt[j];
      const uint32_t r1 = r0 + THREADS * U;
      if (r1 < n) {
#pragma unroll
        for (int j = 0; j < U; ++j) { const uint32_t r = r1 + j * THREADS + t; if (r < n) nxt[j] = __ldcs(rows + r); }
      } else {
#pragma unroll
        for (int j = 0; j < U; ++j) { const uint32_t r = j * THREADS + t; if (r < n_next) nxt[j] = __ldcs(rows_next + r); }
      }
      uint32_t slot[U];
      unsigned long long cur[U];
      uint32_t pend = 0, act = 0;
#pragma unroll
      for (int j = 0; j < U; ++j) {
        slot[j] = slot_hash32(rec[j].x) & smask;
        if (r0 + j * THREADS + t < n) act |= 1u << j;
      }
#pragma unroll
      for (int j = 0; j < U; ++j) cur[j] = lds64(a_key + slot[j] * 8);
#pragma unroll
      for (int j = 0; j < U; ++j) if (((act >> j) & 1u) && cur[j] != rec[j].x) pend |= 1u << j;
      uint32_t guard = 0;
      while (__ballot_sync(0xffffffffu, pend != 0)) {
#pragma unroll
        for (int j = 0; j < U; ++j) {
          if ((pend >> j) & 1u) {
            bool hit = false;
            if (cur[j] == kEmpty) { const unsigned long long old = atoms_cas64(a_key + slot[j] * 8, kEmpty, rec[j].x); hit = old == kEmpty || old == rec[j].x; }
            if (!hit) { slot[j] = (slot[j] + 1) & smask; cur[j] = lds64(a_key + slot[j] * 8); hit = cur[j] == rec[j].x; }
            if (hit) pend &= ~(1u << j);
          }
        }
        if (++guard > S) { overflow = true; pend = 0; }
      }
#pragma unroll
      for (int j = 0; j < U; ++j) {
        if ((act >> j) & 1u) {
          reds_add(a_cnt + slot[j] * 4, 1u);
          const uint32_t vlo = (uint32_t)rec[j].y, vhi = (uint32_t)(rec[j].y >> 32);
          const uint32_t old = atoms_add_ret(a_lo + slot[j] * 4, vlo);
          const uint32_t add = vhi + ((uint32_t)(old + vlo) < old ? 1u : 0u);
          if (add) reds_add(a_hi + slot[j] * 4, add);
        }
      }
    }
    if (n == 0) {
#pragma unroll
      for (int j = 0; j < U; ++j) { const uint32_t r = j * THREADS + t; if (r < n_next) nxt[j] = __ldcs(rows_next + r); }
    }
    __syncthreads();
    const uint32_t own = S / THREADS;
    uint32_t mine = 0;
    for (uint32_t j = 0; j < own; ++j) mine += s_key[t + j * THREADS] != kEmpty;
    uint32_t incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { uint32_t u = __shfl_up_sync(0xffffffffu, incl, d); if ((t & 31) >= (uint32_t)d) incl += u; }
    if ((t & 31) == 31) s_warp[t >> 5] = incl;
    __syncthreads();
    uint32_t woff = 0, total = 0;
#pragma unroll
    for (int w = 0; w < THREADS / 32; ++w) { const uint32_t u = s_warp[w]; if (w < (int)(t >> 5)) woff += u; total += u; }
    if (t == 0) s_base = atomicAdd(p.gcount, (unsigned long long)total);
    __syncthreads();
    uint64_t o = s_base + woff + incl - mine;
    for (uint32_t j = 0; j < own; ++j) {
      const uint32_t l = t + j * THREADS;
      const unsigned long long k = s_key[l];
      if (k != kEmpty) {
        p.out_key[o] = k;
        p.out_sum[o] = (int64_t)(((uint64_t)s_hi[l] << 32) | s_lo[l]);
        p.out_cnt[o] = s_cnt[l];
        ++o;
      }
    }
    __syncthreads();
    fb = fb_next; n = n_next; rows = rows_next;
  }
  if (overflow) p.status[1] = 1;
}

template <int THREADS>
static void run_slices4(const ulonglong2* l2, const uint32_t* cur2, uint32_t C2, uint32_t NB, int log2_slice, const uint64_t* v, int64_t n, int log2cap)
{
  uint32_t* status; CK(cudaMalloc(&status, 64)); CK(cudaMemset(status, 0, 64));
  uint64_t* out_key; int64_t* out_sum; uint32_t* out_cnt; unsigned long long* gcount;
  const uint64_t gmax = 1ull << (log2cap - 1);
  CK(cudaMalloc(&out_key, gmax * 8)); CK(cudaMalloc(&out_sum, gmax * 8)); CK(cudaMalloc(&out_cnt, gmax * 4)); CK(cudaMalloc(&gcount, 32));
  SliceParams sl{}; sl.rec = l2; sl.cursor = cur2; sl.bucket_capacity = C2; sl.num_buckets = NB; sl.log2_slice = log2_slice;
  sl.out_key = out_key; sl.out_sum = out_sum; sl.out_cnt = out_cnt; sl.gcount = gcount; sl.status = status;
  const size_t ssm = ((size_t)20 << log2_slice);
  auto kern = slice4_kernel<THREADS>;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ssm));
  int socc = 0; CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&socc, kern, THREADS, ssm));
  if (socc == 0) { printf("  slice4 thr=%d: does not fit (%zu B)\n", THREADS, ssm); return; }
  float t3 = time_ms([&] { cudaMemsetAsync(gcount, 0, 32); kern<<<148 * socc, THREADS, ssm>>>(sl); });
  unsigned long long h[4] = {0, 0, 0, 0}, g = 0; uint32_t st[2];
  CK(cudaMemcpy(&g, gcount, 8, cudaMemcpyDeviceToHost));
  CK(cudaMemset(gcount + 1, 0, 24));
  check_kernel<<<148 * 4, 256>>>(out_sum, out_cnt, g, gcount + 1);
  colsum_kernel<<<148 * 4, 256>>>(v, n, gcount + 3);
  CK(cudaMemcpy(h, gcount, 32, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(st, status, 8, cudaMemcpyDeviceToHost));
  printf("  slice4 thr=%d occ=%d slice=2^%d smem=%zu: %.3f ms | groups %llu  sum(cnt) %s  sum(sum) %s status %u %u\n", THREADS, socc, log2_slice, ssm, t3, g,
         h[2] == (unsigned long long)n ? "ok" : "BAD", h[1] == h[3] ? "ok" : "BAD", st[0], st[1]);
  cudaFree(status); cudaFree(out_key); cudaFree(out_sum); cudaFree(out_cnt); cudaFree(gcount);
}

template <int THREADS, int CH, int NS, int MINB>
static void run_slices3(const ulonglong2* l2, const uint32_t* cur2, uint32_t C2, uint32_t NB, int log2_slice, const uint64_t* v, int64_t n, int log2cap)
{
  uint32_t* status; CK(cudaMalloc(&status, 64)); CK(cudaMemset(status, 0, 64));
  uint64_t* out_key; int64_t* out_sum; uint32_t* out_cnt; unsigned long long* gcount;
  const uint64_t gmax = 1ull << (log2cap - 1);
  CK(cudaMalloc(&out_key, gmax * 8)); CK(cudaMalloc(&out_sum, gmax * 8)); CK(cudaMalloc(&out_cnt, gmax * 4)); CK(cudaMalloc(&gcount, 32));
  SliceParams sl{}; sl.rec = l2; sl.cursor = cur2; sl.bucket_capacity = C2; sl.num_buckets = NB; sl.log2_slice = log2_slice;
  sl.out_key = out_key; sl.out_sum = out_sum; sl.out_cnt = out_cnt; sl.gcount = gcount; sl.status = status;
  const size_t ssm = ((size_t)20 << log2_slice) + (size_t)NS * CH * 16 + 2 * NS * 8 + 64;
  auto kern = slice3_kernel<THREADS, CH, NS, MINB>;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ssm));
  int socc = 0; CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&socc, kern, THREADS, ssm));
  if (socc == 0) { printf("  slice3 thr=%d CH=%d NS=%d: does not fit (%zu B)\n", THREADS, CH, NS, ssm); return; }
  float t3 = time_ms([&] { cudaMemsetAsync(gcount, 0, 32); kern<<<148 * socc, THREADS, ssm>>>(sl); });
  unsigned long long h[4] = {0, 0, 0, 0}, g = 0; uint32_t st[2];
  CK(cudaMemcpy(&g, gcount, 8, cudaMemcpyDeviceToHost));
  CK(cudaMemset(gcount + 1, 0, 24));
  check_kernel<<<148 * 4, 256>>>(out_sum, out_cnt, g, gcount + 1);
  colsum_kernel<<<148 * 4, 256>>>(v, n, gcount + 3);
  CK(cudaMemcpy(h, gcount, 32, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(st, status, 8, cudaMemcpyDeviceToHost));
  printf("  slice3 thr=%d CH=%d NS=%d occ=%d slice=2^%d smem=%zu: %.3f ms | groups %llu  sum(cnt) %s  sum(sum) %s status %u %u\n", THREADS, CH, NS, socc, log2_slice, ssm, t3, g,
         h[2] == (unsigned long long)n ? "ok" : "BAD", h[1] == h[3] ? "ok" : "BAD", st[0], st[1]);
  cudaFree(status); cudaFree(out_key); cudaFree(out_sum); cudaFree(out_cnt); cudaFree(gcount);
}

template <int T, int THREADS>
static void run2(const uint64_t* k, const uint64_t* v, int64_t n, int log2cap, int b1, int b2, int log2_slice)
{
  const uint32_t P1 = 1u << b1, NB = 1u << (b1 + b2);
  const uint32_t C1 = (uint32_t)((double)n / P1 * 1.03) + 4096, C2 = ((uint32_t)((double)n / NB * 1.27) + 256 + 15) & ~15u;
  ulonglong2 *l1, *l2; uint32_t *cur1, *cur2, *status;
  CK(cudaMalloc(&l1, (size_t)P1 * C1 * 16)); CK(cudaMalloc(&l2, (size_t)NB * C2 * 16));
  CK(cudaMalloc(&cur1, P1 * 4 * 64)); CK(cudaMalloc(&cur2, NB * 4)); CK(cudaMalloc(&status, 64));
  CK(cudaMemset(status, 0, 64));
  uint64_t* out_key; int64_t* out_sum; uint32_t* out_cnt; unsigned long long* gcount;
  const uint64_t gmax = 1ull << (log2cap - 1);
  CK(cudaMalloc(&out_key, gmax * 8)); CK(cudaMalloc(&out_sum, gmax * 8)); CK(cudaMalloc(&out_cnt, gmax * 4)); CK(cudaMalloc(&gcount, 32));
  const size_t smem = (size_t)T * 18 + 4 * 256 * 4 + 64;
  CK(cudaFuncSetAttribute(scatter2_kernel<T, THREADS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CK(cudaFuncSetAttribute(scatter2_kernel<T, THREADS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  ScatterParams s1{}; s1.keys = k; s1.vals = v; s1.n = n; s1.out = l1; s1.cursor = cur1; s1.bucket_capacity = C1;
  s1.log2_parts = b1; s1.slot_shift = log2cap - b1; s1.capmask = (uint32_t)((1ull << log2cap) - 1); s1.status = status; s1.cursor_stride = 64; s1.seg_cursor_stride = 1;
  ScatterParams s2 = s1; s2.in = l1; s2.seg_cursor = cur1; s2.seg_capacity = C1; s2.num_segs = P1; s2.out = l2; s2.cursor = cur2; s2.bucket_capacity = C2;
  s2.log2_parts = b2; s2.slot_shift = log2cap - b1 - b2; s2.cursor_stride = 1; s2.seg_cursor_stride = 64;
  SliceParams sl{}; sl.rec = l2; sl.cursor = cur2; sl.bucket_capacity = C2; sl.num_buckets = NB; sl.log2_slice = log2_slice;
  sl.out_key = out_key; sl.out_sum = out_sum; sl.out_cnt = out_cnt; sl.gcount = gcount; sl.status = status;
  const size_t ssm = (size_t)20 << log2_slice;
  CK(cudaFuncSetAttribute(slice2_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ssm));
  int occ = 0; CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, scatter2_kernel<T, THREADS, false>, THREADS, smem));
  int socc = 0; CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&socc, slice2_kernel<512>, 512, ssm));
  const int grid = 148 * occ;
  float t1 = time_ms([&] { cudaMemsetAsync(cur1, 0, P1 * 4 * 64); scatter2_kernel<T, THREADS, false><<<grid, THREADS, smem>>>(s1); });
  float t2 = time_ms([&] { cudaMemsetAsync(cur2, 0, NB * 4); scatter2_kernel<T, THREADS, true><<<grid, THREADS, smem>>>(s2); });
  float t3 = time_ms([&] { cudaMemsetAsync(gcount, 0, 32); slice2_kernel<512><<<148 * (socc ? socc : 1), 512, ssm>>>(sl); });
  unsigned long long h[4] = {0, 0, 0, 0}, g = 0; uint32_t st[2];
  CK(cudaMemcpy(&g, gcount, 8, cudaMemcpyDeviceToHost));
  CK(cudaMemset(gcount + 1, 0, 24));
  check_kernel<<<148 * 4, 256>>>(out_sum, out_cnt, g, gcount + 1);
  colsum_kernel<<<148 * 4, 256>>>(v, n, gcount + 3);
  CK(cudaMemcpy(h, gcount, 32, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(st, status, 8, cudaMemcpyDeviceToHost));
  printf("V2 T=%d thr=%d occ=%d/%d b1=%d b2=%d slice=2^%d: scatter1 %.3f ms (%.2f TB/s)  scatter2 %.3f ms  slices %.3f ms  total %.3f | groups %llu  sum(cnt) %llu %s  sum(sum) %s  status %u %u\n",
         T, THREADS, occ, socc, b1, b2, log2_slice, t1, 32.0 * n / t1 / 1e9, t2, t3, t1 + t2 + t3, g, h[2], h[2] == (unsigned long long)n ? "ok" : "BAD", h[1] == h[3] ? "ok" : "BAD", st[0], st[1]);
  if (log2_slice == 12) {
    run_slices4<512>(l2, cur2, C2, NB, log2_slice, v, n, log2cap);
    run_slices4<256>(l2, cur2, C2, NB, log2_slice, v, n, log2cap);
    run_slices4<384>(l2, cur2, C2, NB, log2_slice, v, n, log2cap);
  } else {
    run_slices4<1024>(l2, cur2, C2, NB, log2_slice, v, n, log2cap);
  }
  cudaFree(l1); cudaFree(l2); cudaFree(cur1); cudaFree(cur2); cudaFree(status); cudaFree(out_key); cudaFree(out_sum); cudaFree(out_cnt); cudaFree(gcount);
}

template <int T, int THREADS, int NS>
static void sweep1(const uint64_t* k, const uint64_t* v, int64_t n, int log2cap)
{
  const size_t smem = (size_t)(NS + 1) * T * 16 + 7 * 256 * 4 + 64;
  CK(cudaFuncSetAttribute(scatter_kernel<T, THREADS, NS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0; CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, scatter_kernel<T, THREADS, NS, false>, THREADS, smem));
  for (int b1 = 3; b1 <= 8; ++b1) {
    const uint32_t P1 = 1u << b1;
    const uint32_t C1 = (uint32_t)((double)n / P1 * 1.03) + 4096;
    ulonglong2* l1; uint32_t *cur1, *status;
    CK(cudaMalloc(&l1, (size_t)P1 * C1 * 16)); CK(cudaMalloc(&cur1, P1 * 4 * 64)); CK(cudaMalloc(&status, 64));
    ScatterParams s1{}; s1.keys = k; s1.vals = v; s1.n = n; s1.out = l1; s1.cursor = cur1; s1.bucket_capacity = C1;
    s1.log2_parts = b1; s1.slot_shift = log2cap - b1; s1.capmask = (uint32_t)((1ull << log2cap) - 1); s1.status = status; s1.cursor_stride = 64; s1.seg_cursor_stride = 1;
    float t1 = time_ms([&] { cudaMemsetAsync(cur1, 0, P1 * 4 * 64); scatter_kernel<T, THREADS, NS, false><<<148 * occ, THREADS, smem>>>(s1); });
    printf("  scatter1 only T=%d thr=%d occ=%d b1=%d (run %d B): %.3f ms (%.2f TB/s)\n", T, THREADS, occ, b1, T / (int)P1 * 16, t1, 32.0 * n / t1 / 1e9);
    cudaFree(l1); cudaFree(cur1); cudaFree(status);
  }
}

int main(int argc, char** argv)
{
  const int64_t n = argc > 1 ? atoll(argv[1]) : 1000000000ll;
  const uint64_t g = argc > 2 ? strtoull(argv[2], 0, 10) : 100000000ull;
  uint64_t *k, *v;
  CK(cudaMalloc(&k, n * 8 + 64)); CK(cudaMalloc(&v, n * 8 + 64));
  gen<<<148 * 8, 256>>>(k, v, n, g);
  CK(cudaDeviceSynchronize());
  int log2cap = 10; while ((1ull << log2cap) < 2 * 1.25 * g) ++log2cap;
  printf("n=%lld g=%llu log2cap=%d\n", (long long)n, (unsigned long long)g, log2cap);
  run2<4096, 512>(k, v, n, log2cap, 8, log2cap - 12 - 8, 12);
  run2<4096, 512>(k, v, n, log2cap, 7, log2cap - 13 - 7, 13);
  if (argc > 3) return 0;
  run2<4096, 256>(k, v, n, log2cap, 8, log2cap - 12 - 8, 12);
  run2<4096, 1024>(k, v, n, log2cap, 8, log2cap - 12 - 8, 12);
  run2<2048, 256>(k, v, n, log2cap, 8, log2cap - 12 - 8, 12);
  return 0;
}

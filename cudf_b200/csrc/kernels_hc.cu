// High-cardinality path in records form: see kernels_hc.h for the pipeline. sm_100a only (cp.async.bulk + mbarrier).
//
// hc_scatter_kernel, one tile of T rows per iteration, 512 threads, two CTAs per SM (they interleave: one waits for its bulk
// copy while the other ranks / writes):
//   0  TMA: thread 0 arms the CTA's mbarrier with the tile's byte count and issues one cp.async.bulk per source (level 1:
//      a raw tile of every source column and mask; level 2: the records); everybody waits on the barrier's phase
//   A  bucket of every row from the top bits of its table slot, rank inside the bucket by a returning ATOMS.ADD
//   B  block scan of the <= 1024 bucket counters; ONE global atomicAdd per (tile, bucket) reserves the run in the bucket
//   C  permutation, one 32-bit entry per row: sorted position -> bucket << 16 | row of the tile (no sorted copy of the records:
//      T = 4096 rows fit two CTAs per SM, which makes a run 256 B at 256 buckets -- measured 6.2 ms per 1e9 rows per level
//      against 7.5 ms with 128 B runs; the bucket rides along so that D does not hash the key again)
//   D  position i gathers row idx[i] from the stage, assembles the record (level 1 with pack32: {key32, value32} in 8 bytes,
//      every row checked to fit) and stores it at run base + i: consecutive threads write consecutive records
// Bucket capacities are fixed (expected rows + margin); a bucket that fills sets status[6] and the host reruns the call on
// the exact-offset path (kernels_partition.cu) -- skewed keys are the partitioned path's job.
//
// hc_slice_kernel, one CTA per final bucket: the bucket's rows stream through registers (128-bit loads, the next U rows are
// in flight while the current ones are processed), key -> slot by open addressing in a shared-memory table of S slots
// (first probe of U rows unrolled for ILP, then one warp-converged loop for the rows whose home slot was taken: a loop with
// a per-lane break diverges the warp for good and ran at 7 of 32 lanes), every accumulator in shared memory (COUNT: one
// RED.32; integer sums: a 64-bit cell as {lo, hi} updated by one returning ATOMS.ADD on lo and a second add only when the
// carry or the high word is non-zero -- exact mod 2^64 with any number of rows per bucket; MIN / MAX: load, atomic only when
// the row improves the extremum; float sums and products: the CAS forms). The finished slice is COMPACTED in shared memory
// order and appended to the dense group arrays at a position taken from one global atomicAdd per bucket.
#include <cstdlib>

#include "device_utils.cuh"
#include "kernels_hc.h"
#include "kernels_peer.h"

namespace bgb {

namespace {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(c)); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity)
{
  asm volatile("{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
// global -> shared bulk copy (TMA), bytes a multiple of 16, both addresses 16-byte aligned; completion on the mbarrier
__device__ __forceinline__ void tma_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ uint64_t lds_raw(const unsigned char* tile, int width, uint32_t r)
{
  switch (width) {
    case 1: return tile[r];
    case 2: return reinterpret_cast<const uint16_t*>(tile)[r];
    case 4: return reinterpret_cast<const uint32_t*>(tile)[r];
    default: return reinterpret_cast<const uint64_t*>(tile)[r];
  }
}

// value word of a record: signed 1/2/4-byte integers travel SIGN-EXTENDED, so that a plain SUM over them is the 64-bit add of the
// lean slice kernel (every other consumer re-truncates the word to the source width first: ordered_image, raw_to_i64)
__device__ __forceinline__ uint64_t record_value(uint64_t raw, int src)
{
  switch (src) {
    case S_I8: return (uint64_t)(int64_t)(int8_t)raw;
    case S_I16: return (uint64_t)(int64_t)(int16_t)raw;
    case S_I32: return (uint64_t)(int64_t)(int32_t)raw;
    default: return raw;
  }
}

template <int W> __device__ __forceinline__ void store_rec(uint64_t* dst, const uint64_t (&w)[W])
{
  if (W == 1) dst[0] = w[0];
  else if (W == 2) *reinterpret_cast<ulonglong2*>(dst) = make_ulonglong2(w[0], w[1]);
  else if (W == 3) { dst[0] = w[0]; dst[1] = w[1]; dst[2] = w[2]; }
  else { reinterpret_cast<ulonglong2*>(dst)[0] = make_ulonglong2(w[0], w[1]); reinterpret_cast<ulonglong2*>(dst)[1] = make_ulonglong2(w[2 % W], w[3 % W]); }
}
template <int W> __device__ __forceinline__ void load_rec(const uint64_t* src, uint64_t (&w)[W])
{
  if (W == 1) w[0] = __ldcs(reinterpret_cast<const unsigned long long*>(src));
  else if (W == 2) { const ulonglong2 x = __ldcs(reinterpret_cast<const ulonglong2*>(src)); w[0] = x.x; w[1] = x.y; }
  else if (W == 3) { w[0] = __ldcs(reinterpret_cast<const unsigned long long*>(src)); w[1 % W] = __ldcs(reinterpret_cast<const unsigned long long*>(src) + 1); w[2 % W] = __ldcs(reinterpret_cast<const unsigned long long*>(src) + 2); }
  else { const ulonglong2 x = __ldcs(reinterpret_cast<const ulonglong2*>(src)), y = __ldcs(reinterpret_cast<const ulonglong2*>(src) + 1); w[0] = x.x; w[1 % W] = x.y; w[2 % W] = y.x; w[3 % W] = y.y; }
}

}  // namespace

int hc_tile_rows(int words) { return words <= 2 ? 4096 : 2048; }
size_t hc_scatter_smem_bytes(int words, int log2_parts)
{
  const size_t T = (size_t)hc_tile_rows(words);
  return T * words * 8 + T * 4 + (size_t)(4u << log2_parts) * 4 + 64;
}

// MODE 0: level 1, generic packer (LDG -> records in the stage); 1: level 1, raw column tiles by TMA; 2: level 2 (records by TMA);
// 3: level 1 of a MERGE pass: row r is record (source, j) of the received region (word 0 key, one word per accumulator cell)
template <int W, int MODE>
__global__ void __launch_bounds__(kHcThreads, 2) hc_scatter_kernel(const __grid_constant__ HcScatterParams p)
{
  constexpr int T = W <= 2 ? 4096 : 2048;
  constexpr int R = T / kHcThreads;
  extern __shared__ __align__(128) unsigned char smem[];
  const uint32_t P = 1u << p.log2_parts, pmask = P - 1;
  unsigned char* stage = smem;                                               // T x W x 8 bytes: records, or raw column tiles
  const int stage_words = (MODE == 1 && p.pack32) ? 2 : W;                    // (pack32: 16 bytes of raw columns become 8-byte records)
  uint32_t* s_idx   = reinterpret_cast<uint32_t*>(smem + (size_t)T * stage_words * 8);  // [T] sorted position -> bucket << 16 | row
  uint32_t* s_cnt   = reinterpret_cast<uint32_t*>(s_idx + T);                 // [P]
  uint32_t* s_off   = s_cnt + P;
  uint32_t* s_gbase = s_off + P;
  uint32_t* s_end   = s_gbase + P;
  uint64_t* full    = reinterpret_cast<uint64_t*>(s_end + P);
  __shared__ uint32_t s_warp[kHcThreads / 32];
  __shared__ uint32_t s_total;
  const uint32_t t       = threadIdx.x;
  const uint32_t capmask = (uint32_t)(p.a.table.capacity - 1);

  // raw tile offsets (MODE 1): key | values | masks
  uint32_t voff[kHcMaxWords - 1], moff[kHcMaxWords - 1];
  const int kw = p.a.kc[0].width;
  if (MODE == 1) {
    uint32_t o = T * kw;
    for (int c = 0; c < kHcMaxWords - 1; ++c) { voff[c] = o; if (c < p.num_vals) o += T * p.val[c].width; }
    for (int c = 0; c < kHcMaxWords - 1; ++c) { moff[c] = o; if (c < p.num_vals && p.val[c].vbit >= 0) o += T / 8; }
  }

  if ((MODE == 1 || MODE == 2) && t == 0) { mbar_init(full, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  for (uint32_t i = t; i < P; i += kHcThreads) s_cnt[i] = 0;
  __syncthreads();

  const uint32_t tiles_per_seg = MODE == 2 ? (p.seg_capacity + T - 1) / T : 1;
  const int64_t ntiles = MODE == 2 ? (int64_t)p.num_segs * tiles_per_seg : (p.a.num_rows + T - 1) / T;
  uint32_t phase = 0;
  (void)phase;
  bool overflow = false, misfit = false;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    int64_t row0;
    uint32_t nrows, cbase = 0;
    if (MODE == 2) {
      const uint32_t seg = (uint32_t)(tile / tiles_per_seg), tt = (uint32_t)(tile % tiles_per_seg);
      uint32_t cnt = __ldg(p.seg_cursor + (size_t)seg * p.seg_cursor_stride);
      cnt = cnt < p.seg_capacity ? cnt : p.seg_capacity;
      if (tt * T >= cnt) continue;  // CTA-uniform
      nrows = cnt - tt * T < (uint32_t)T ? cnt - tt * T : T;
      row0  = (int64_t)seg * p.seg_capacity + (int64_t)tt * T;
      cbase = seg << p.log2_parts;
    } else {
      row0  = tile * T;
      nrows = p.a.num_rows - row0 < (int64_t)T ? (uint32_t)(p.a.num_rows - row0) : T;
    }

    // ---- 0: the tile into shared memory
    if (MODE == 2) {
      if (t == 0) {
        const uint32_t bytes = (nrows * W * 8 + 15u) & ~15u;  // an odd row count of 8- / 24-byte records reads 8 bytes of slack
        mbar_expect(full, bytes);
        tma_g2s(stage, p.in + (uint64_t)row0 * W, bytes, full);
      }
      mbar_wait(full, phase); phase ^= 1;
    } else if (MODE == 1) {
      // one bulk copy per source tile for the 16-byte multiple of its bytes; a tail of < 16 bytes (last tile only) by hand
      if (t == 0) {
        uint32_t total = ((nrows * kw) & ~15u);
        for (int c = 0; c < p.num_vals; ++c) {
          total += (nrows * p.val[c].width) & ~15u;
          if (p.val[c].vbit >= 0) total += ((nrows + 7) / 8) & ~15u;
        }
        if (total) mbar_expect(full, total); else mbar_arrive(full);
        uint32_t b = (nrows * kw) & ~15u;
        if (b) tma_g2s(stage, p.a.kc[0].data + ((int64_t)p.a.kc[0].offset + row0) * kw, b, full);
        for (int c = 0; c < p.num_vals; ++c) {
          const HcValue& v = p.val[c];
          b = (nrows * v.width) & ~15u;
          if (b) tma_g2s(stage + voff[c], v.data + ((int64_t)v.offset + row0) * v.width, b, full);
          if (v.vbit >= 0) {
            b = ((nrows + 7) / 8) & ~15u;
            if (b) tma_g2s(stage + moff[c], reinterpret_cast<const uint8_t*>(v.mask) + ((int64_t)v.offset + row0) / 8, b, full);
          }
        }
      }
      if (nrows < (uint32_t)T) {  // CTA-uniform: the last tile
        auto patch = [&](unsigned char* dst, const uint8_t* src, uint32_t bytes) {
          const uint32_t done = bytes & ~15u;
          if (t < bytes - done) dst[done + t] = src[done + t];
        };
        patch(stage, p.a.kc[0].data + ((int64_t)p.a.kc[0].offset + row0) * kw, nrows * kw);
        for (int c = 0; c < p.num_vals; ++c) {
          const HcValue& v = p.val[c];
          patch(stage + voff[c], v.data + ((int64_t)v.offset + row0) * v.width, nrows * v.width);
          if (v.vbit >= 0) patch(stage + moff[c], reinterpret_cast<const uint8_t*>(v.mask) + ((int64_t)v.offset + row0) / 8, (nrows + 7) / 8);
        }
        __syncthreads();
      }
      mbar_wait(full, phase); phase ^= 1;
    }

    auto key_image = [&](uint32_t r) -> uint64_t {  // the packed key image of row r of the stage (hash input)
      if (MODE == 1) return lds_raw(stage, kw, r);
      const uint64_t w0 = reinterpret_cast<const uint64_t*>(stage)[(size_t)r * W];
      return p.key_in_low32 ? (w0 & 0xffffffffull) : w0;
    };
    auto bucket_of = [&](uint64_t k) { return ((slot_hash32(k) & capmask) >> p.slot_shift) & pmask; };

    // ---- A
    uint32_t pr[R];  // bucket << 16 | rank; ~0: the row is dropped (null key under EXCLUDE) or past the tile
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const uint32_t r = j * kHcThreads + t;
      pr[j] = 0xFFFFFFFFu;
      if (r < nrows) {
        uint64_t k;
        bool keep = true;
        if (MODE == 3) {
          const uint32_t row = (uint32_t)(row0 + r);
          int src = 0;
          while (src + 1 < p.peer_in.world && row >= p.peer_in.prefix[src + 1]) ++src;   // world <= 64 sources
          const uint64_t idx = (uint64_t)src * p.peer_in.cap + (row - p.peer_in.prefix[src]);
          uint64_t w[W];
          // .cg loads: the records were written by another GPU, never through this SM's L1
          w[0] = __ldcg(reinterpret_cast<const unsigned long long*>(p.peer_in.region) + idx);
#pragma unroll
          for (int c = 0; c < W - 1; ++c) {
            w[(c + 1) % W] = 0;
            if (c < p.peer_in.num_accs) {
              const uint8_t* base = p.peer_in.region + p.peer_in.acc_off[c];
              w[(c + 1) % W] = p.peer_in.cell_bytes[c] == 8 ? __ldcg(reinterpret_cast<const unsigned long long*>(base) + idx)
                                                            : (uint64_t)__ldcg(reinterpret_cast<const unsigned int*>(base) + idx);
            }
          }
          k = w[0];
          store_rec<W>(reinterpret_cast<uint64_t*>(stage) + (size_t)r * W, w);
        } else if (MODE == 0) {
          const int64_t row = row0 + r;
          uint64_t k1;
          keep = pack_key(p.a, row, k, k1);
          if (keep) {
            uint64_t w[W];
            uint32_t vb = 0;
            w[0] = k;
#pragma unroll
            for (int c = 0; c < W - 1; ++c) {
              w[(c + 1) % W] = 0;
              if (c < p.num_vals) {
                const HcValue& v = p.val[c];
                w[(c + 1) % W] = record_value(load_raw(v.data, v.width, (int64_t)v.offset + row), v.src);
                if (v.vbit >= 0) vb |= (uint32_t)bit_is_set(v.mask, (int64_t)v.offset + row) << v.vbit;
              }
            }
            if (p.key_in_low32) w[0] |= (uint64_t)vb << 32;
            else if (p.vbits_word > 0) {
#pragma unroll
              for (int x = 1; x < W; ++x) if (p.vbits_word == x) w[x % W] = vb;
            }
            store_rec<W>(reinterpret_cast<uint64_t*>(stage) + (size_t)r * W, w);
          }
        } else k = key_image(r);
        if (keep) {
          const uint32_t b = bucket_of(k);
          pr[j] = (b << 16) | atomicAdd(&s_cnt[b], 1u);
        }
      }
    }
    __syncthreads();

    // ---- B: block scan of the bucket counters (1 or 2 consecutive counters per thread), reservations issued
    const uint32_t cpt = P > (uint32_t)kHcThreads ? P / kHcThreads : 1;
    uint32_t cn[2] = {0, 0}, ex[2] = {0, 0}, gres[2] = {0, 0};
    {
      uint32_t mine = 0;
#pragma unroll
      for (int x = 0; x < 2; ++x) {
        const uint32_t i = t * cpt + x;
        if ((uint32_t)x < cpt && i < P) { cn[x] = s_cnt[i]; s_cnt[i] = 0; }
        ex[x] = mine; mine += cn[x];
      }
      uint32_t incl = mine;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, incl, d); if ((t & 31) >= (uint32_t)d) incl += u; }
      if ((t & 31) == 31) s_warp[t >> 5] = incl;
      __syncthreads();
      uint32_t woff = 0, total = 0;
#pragma unroll
      for (int w = 0; w < kHcThreads / 32; ++w) { const uint32_t u = s_warp[w]; if (w < (int)(t >> 5)) woff += u; total += u; }
      if (t == 0) s_total = total;
      const uint32_t base = woff + incl - mine;
#pragma unroll
      for (int x = 0; x < 2; ++x) {
        const uint32_t i = t * cpt + x;
        ex[x] += base;
        if ((uint32_t)x < cpt && i < P) {
          s_off[i] = ex[x];
          if (cn[x]) gres[x] = atomicAdd(p.cursor + (size_t)(cbase + i) * p.cursor_stride, cn[x]);
        }
      }
    }
    __syncthreads();

    // ---- C: permutation; the reservations (in flight since B) become run bases
#pragma unroll
    for (int j = 0; j < R; ++j)
      if (pr[j] != 0xFFFFFFFFu) s_idx[s_off[pr[j] >> 16] + (pr[j] & 0xffffu)] = (pr[j] & 0xffff0000u) | (uint32_t)(j * kHcThreads + t);
#pragma unroll
    for (int x = 0; x < 2; ++x) {
      const uint32_t i = t * cpt + x;
      if ((uint32_t)x < cpt && i < P) {
        const uint32_t g    = gres[x];
        const uint32_t room = g < p.bucket_capacity ? p.bucket_capacity - g : 0;
        s_gbase[i] = (cbase + i) * p.bucket_capacity + g - ex[x];
        s_end[i]   = ex[x] + (cn[x] < room ? cn[x] : room);
        if (cn[x] > room) overflow = true;
      }
    }
    __syncthreads();

    // ---- D: coalesced write-out
    const uint32_t total = s_total;
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const uint32_t i = j * kHcThreads + t;
      if (i < total) {
        const uint32_t e = s_idx[i];  // (the bucket travels with the row number: no second hash of the key)
        const uint32_t r = e & 0xffffu, b = e >> 16;
        uint64_t w[W];
        uint64_t k;
        if (MODE == 1) {
          k = lds_raw(stage, kw, r);
          uint32_t vb = 0;
          w[0] = k;
#pragma unroll
          for (int c = 0; c < W - 1; ++c) {
            w[(c + 1) % W] = 0;
            if (c < p.num_vals) {
              w[(c + 1) % W] = record_value(lds_raw(stage + voff[c], p.val[c].width, r), p.val[c].src);
              if (p.val[c].vbit >= 0) vb |= (uint32_t)((stage[moff[c] + (r >> 3)] >> (r & 7)) & 1u) << p.val[c].vbit;
            }
          }
          if (W == 1 && p.pack32) {
            const int64_t v = raw_to_i64(lds_raw(stage + voff[0], p.val[0].width, r), p.val[0].src);
            misfit |= (k >> 32) != 0 || v != (int64_t)(int32_t)v;
            w[0] = (k & 0xffffffffull) | ((uint64_t)(uint32_t)v << 32);
          } else if (p.key_in_low32) w[0] |= (uint64_t)vb << 32;
          else if (p.vbits_word > 0) {
#pragma unroll
            for (int x = 1; x < W; ++x) if (p.vbits_word == x) w[x % W] = vb;
          }
        } else {
          const uint64_t* src = reinterpret_cast<const uint64_t*>(stage) + (size_t)r * W;
#pragma unroll
          for (int x = 0; x < W; ++x) w[x] = src[x];
        }
        if (i < s_end[b]) store_rec<W>(p.out + (uint64_t)(s_gbase[b] + i) * W, w);
      }
    }
    __syncthreads();  // the stage and the counters are free
  }
  if (overflow) p.status[6] = 1;
  if (misfit) p.status[7] = 1;
}

template <int W>
static cudaError_t launch_scatter_w(const HcScatterParams& p, int num_sms, cudaStream_t s)
{
  const int T = hc_tile_rows(W);
  const size_t smem = hc_scatter_smem_bytes((p.level == 1 && p.pack32) ? 2 : W, p.log2_parts);
  int64_t ntiles;
  cudaError_t e;
  if (p.level == 2) {
    ntiles = (int64_t)p.num_segs * ((p.seg_capacity + T - 1) / T);
    if (ntiles == 0) return cudaSuccess;
    e = allow_max_dynamic_smem(hc_scatter_kernel<W, 2>);
    if (e != cudaSuccess) return e;
    const int grid = (int)(ntiles < 2ll * num_sms ? ntiles : 2ll * num_sms);
    hc_scatter_kernel<W, 2><<<grid, kHcThreads, smem, s>>>(p);
    return note_launch();
  }
  ntiles = (p.a.num_rows + T - 1) / T;
  if (ntiles == 0) return cudaSuccess;
  const int grid = (int)(ntiles < 2ll * num_sms ? ntiles : 2ll * num_sms);
  if (p.peer_in.region != nullptr) {
    e = allow_max_dynamic_smem(hc_scatter_kernel<W, 3>);
    if (e != cudaSuccess) return e;
    hc_scatter_kernel<W, 3><<<grid, kHcThreads, smem, s>>>(p);
    return note_launch();
  }
  if (p.tma_src) {
    e = allow_max_dynamic_smem(hc_scatter_kernel<W, 1>);
    if (e != cudaSuccess) return e;
    hc_scatter_kernel<W, 1><<<grid, kHcThreads, smem, s>>>(p);
  } else {
    e = allow_max_dynamic_smem(hc_scatter_kernel<W, 0>);
    if (e != cudaSuccess) return e;
    hc_scatter_kernel<W, 0><<<grid, kHcThreads, smem, s>>>(p);
  }
  return note_launch();
}

cudaError_t launch_hc_scatter(const HcScatterParams& p, int num_sms, cudaStream_t s)
{
  switch (p.words) {
    case 1: return launch_scatter_w<1>(p, num_sms, s);
    case 2: return launch_scatter_w<2>(p, num_sms, s);
    case 3: return launch_scatter_w<3>(p, num_sms, s);
    case 4: return launch_scatter_w<4>(p, num_sms, s);
    default: return cudaErrorInvalidValue;
  }
}

// ------------------------------------------------------------------ slices
namespace {

__device__ __forceinline__ uint32_t atoms_add_ret(uint32_t addr, uint32_t v) { uint32_t o; asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(o) : "r"(addr), "r"(v) : "memory"); return o; }
__device__ __forceinline__ void reds_add(uint32_t addr, uint32_t v) { asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned long long atoms_cas64(uint32_t addr, unsigned long long c, unsigned long long n)
{ unsigned long long o; asm volatile("atom.shared.cas.b64 %0, [%1], %2, %3;" : "=l"(o) : "r"(addr), "l"(c), "l"(n) : "memory"); return o; }
__device__ __forceinline__ unsigned long long lds64v(uint32_t addr) { unsigned long long o; asm volatile("ld.volatile.shared.u64 %0, [%1];" : "=l"(o) : "r"(addr) : "memory"); return o; }

template <int W> __device__ __forceinline__ uint64_t pick_word(const uint64_t (&w)[W], int word)
{
  uint64_t r = w[0];
#pragma unroll
  for (int x = 1; x < W; ++x) if (word == x) r = w[x];
  return r;
}

// one row of the key whose image equals EMPTY: every accumulator through the reserved slot of the table (cold path)
template <int W> __device__ __noinline__ void hc_reserved_row(const HcSliceParams& p, const uint64_t (&w)[W], uint32_t vb)
{
  p.table.status[2] = 1;
  for (int a = 0; a < p.num_accs; ++a) {
    const HcAcc& acc = p.acc[a];
    if (!(acc.op == ACC_COUNT_ALL || acc.vbit < 0 || ((vb >> acc.vbit) & 1u))) continue;
    if (acc.mode == HM_ADD32) { atomicAdd(reinterpret_cast<unsigned int*>(acc.cells) + p.table.capacity, (unsigned int)pick_word<W>(w, acc.word)); continue; }  // merge pass: a partial count
    apply_value(acc.op, acc.src, acc.cells, p.table.capacity, pick_word<W>(w, acc.word), 0);
  }
}

}  // namespace

// LEAN: every accumulator is a COUNT or a plain 64-bit integer SUM and no validity bit is looked at (the SUM / COUNT shapes of
// BASELINE configs[1], [4] and the north-star target): the other accumulator forms are compiled out -- the general kernel's
// code does not fit the instruction cache well (5 % of its stall samples were instruction fetches) and its descriptor-driven
// accumulator section cost ~250 of ~930 warp instructions per 128 rows.
// KIND 2 (F64STATS): every accumulator reads ONE float64 value column (one validity bit) and is a float64 SUM / SUM_OF_SQUARES, a MIN,
// a MAX or a COUNT -- BASELINE configs[2]. The general section below spends ~620 of its 1520 warp instructions per 128 rows on
// descriptor fetches, per-row `src` switches the compiler does not hoist and an if-chain over the accumulator modes; here the value,
// its order image and the validity of the U rows are computed once and every accumulator is one uniform switch around U predicated ops.
template <int W, int THREADS, int KIND>
__global__ void __launch_bounds__(THREADS, 2) hc_slice_kernel(const __grid_constant__ HcSliceParams p)
{
  constexpr bool LEAN = KIND == 1;
  constexpr int U = W <= 2 ? 4 : 2;
  extern __shared__ __align__(128) unsigned char smem[];
  const uint32_t S = 1u << p.log2_slice, smask = S - 1;
  unsigned long long* s_key = reinterpret_cast<unsigned long long*>(smem);
  __shared__ uint32_t s_warp[THREADS / 32];
  __shared__ unsigned long long s_base;
  __shared__ uint32_t s_pcnt[kHcMaxPeers], s_pfill[kHcMaxPeers], s_poff[kHcMaxPeers + 1], s_pbase[kHcMaxPeers];
  const uint32_t t = threadIdx.x;
  const uint32_t a_key = smem_u32(s_key);
  constexpr unsigned long long kEmptyK = ~0ull;

  uint64_t nxt[U][W];
  uint32_t fb = blockIdx.x;
  uint32_t n  = 0;
  auto bucket_rows = [&](uint32_t b) { uint32_t c = 0; if (b < p.num_buckets) { c = __ldg(p.cursor + (size_t)b * p.cursor_stride); c = c < p.bucket_capacity ? c : p.bucket_capacity; } return c; };
  n = bucket_rows(fb);
  const uint64_t* rows = p.rec + (uint64_t)fb * p.bucket_capacity * W;
#pragma unroll
  for (int j = 0; j < U; ++j) { const uint32_t r = j * THREADS + t; if (r < n) load_rec<W>(rows + (uint64_t)r * W, nxt[j]); }
  bool overflow = false;
  bool any_vbit = false;  // some accumulator looks at a validity bit (else the validity word of a row is never extracted)
  for (int a = 0; a < p.num_accs; ++a) any_vbit |= p.acc[a].vbit >= 0 && p.acc[a].op != ACC_COUNT_ALL;

  while (fb < p.num_buckets) {
    // ---- EMPTY keys + identity cells
    for (uint32_t i = t; i < S / 2; i += THREADS) reinterpret_cast<uint4*>(s_key)[i] = make_uint4(~0u, ~0u, ~0u, ~0u);
#pragma unroll 1
    for (int a = 0; a < p.num_accs; ++a) {
      const HcAcc& acc = p.acc[a];
      if (acc.cell_bytes == 8) { const uint64_t id = acc_identity64(acc.op); uint64_t* c = reinterpret_cast<uint64_t*>(smem + acc.smem_off); for (uint32_t i = t; i < S; i += THREADS) c[i] = id; }
      else { const uint32_t id = acc_identity32(acc.op); uint32_t* c = reinterpret_cast<uint32_t*>(smem + acc.smem_off); for (uint32_t i = t; i < S; i += THREADS) c[i] = id; }
    }
    __syncthreads();
    const uint32_t fb_next = fb + gridDim.x;
    const uint32_t n_next  = bucket_rows(fb_next);
    const uint64_t* rows_next = p.rec + (uint64_t)fb_next * p.bucket_capacity * W;

    for (uint32_t r0 = 0; r0 < n; r0 += THREADS * U) {
      uint64_t rec[U][W];
#pragma unroll
      for (int j = 0; j < U; ++j)
#pragma unroll
        for (int x = 0; x < W; ++x) rec[j][x] = nxt[j][x];
      const uint32_t r1 = r0 + THREADS * U;
      if (r1 < n) {
#pragma unroll
        for (int j = 0; j < U; ++j) { const uint32_t r = r1 + j * THREADS + t; if (r < n) load_rec<W>(rows + (uint64_t)r * W, nxt[j]); }
      } else {  // the first rows of the CTA's next bucket
#pragma unroll
        for (int j = 0; j < U; ++j) { const uint32_t r = j * THREADS + t; if (r < n_next) load_rec<W>(rows_next + (uint64_t)r * W, nxt[j]); }
      }
      // ---- key -> slot. The table is probed in PAIRS of slots (one 128-bit LDS per step; home pair = hash & (S/2 - 1), linear in
      // pairs): a step resolves a row when either slot holds its key or is EMPTY (CAS), which halves the probe steps of linear
      // probing -- the warp pays for the LONGEST of its 32 x U probe sequences, so the tail is what costs
      uint32_t slot[U], vb[U];
      unsigned long long key[U];
      uint32_t pend = 0, act = 0;
      const uint32_t pmask = (S >> 1) - 1;
#pragma unroll
      for (int j = 0; j < U; ++j) {
        key[j] = p.key_in_low32 ? (rec[j][0] & 0xffffffffull) : rec[j][0];
        vb[j]  = (LEAN || !any_vbit) ? 0xFFFFFFFFu : p.key_in_low32 ? (uint32_t)(rec[j][0] >> 32) : (uint32_t)pick_word<W>(rec[j], p.vbits_word);
        slot[j] = (slot_hash32(key[j]) & pmask) << 1;
        if (r0 + j * THREADS + t < n) {
          if (key[j] == kEmptyK) {  // cold: a copy, so that the rows themselves never have their address taken
            uint64_t tmp[W];
#pragma unroll
            for (int x = 0; x < W; ++x) tmp[x] = rec[j][x];
            hc_reserved_row<W>(p, tmp, vb[j]);
          } else act |= 1u << j;
        }
      }
      // One probe step on (k, sl): true when sl is final. Written with ONE divergent branch (the CAS): every path a lane of the warp
      // takes is paid by the whole warp, and the five-way if-chain this replaces cost ~45 instructions per call.
      auto step = [&](unsigned long long k, uint32_t& sl) -> bool {
        unsigned long long c0, c1;
        asm volatile("ld.volatile.shared.v2.u64 {%0, %1}, [%2];" : "=l"(c0), "=l"(c1) : "r"(a_key + sl * 8) : "memory");
        const bool hit0 = c0 == k, hit1 = c1 == k, e0 = c0 == kEmptyK, e1 = c1 == kEmptyK;
        bool found = hit0 || hit1;
        uint32_t add = hit0 ? 0u : 1u;                 // slot of the pair the key sits in (or would be inserted at)
        const bool try_cas = !found && (e0 || e1);     // first EMPTY slot in probe order: c0 before c1
        if (try_cas) {
          add = e0 ? 0u : 1u;
          const unsigned long long old = atoms_cas64(a_key + (sl + add) * 8, kEmptyK, k);
          found = old == kEmptyK || old == k;          // lost to another key: the pair is looked at again
        }
        if (found) sl += add;
        else if (!try_cas) sl = (sl + 2) & smask;      // both slots hold other keys: next pair
        return found;
      };
#pragma unroll
      for (int j = 0; j < U; ++j) if (((act >> j) & 1u) && !step(key[j], slot[j])) pend |= 1u << j;
      // rows still looking: each lane works on ITS lowest pending row, one probe step per iteration, so the warp pays
      // max-over-lanes of the lane's remaining steps with ONE step body per iteration (not U bodies per iteration)
      uint32_t guard = 0;
      while (__ballot_sync(0xffffffffu, pend != 0)) {
        if (pend) {
          const int j = __ffs(pend) - 1;
          unsigned long long k = key[0];
          uint32_t sl = slot[0];
#pragma unroll
          for (int x = 1; x < U; ++x) if (j == x) { k = key[x]; sl = slot[x]; }
          const bool done = step(k, sl);
#pragma unroll
          for (int x = 0; x < U; ++x) if (j == x) slot[x] = sl;
          if (done) pend &= pend - 1;
        }
        if (++guard > 4 * S + 8) { overflow = true; act &= ~pend; pend = 0; }  // more distinct keys than slots: the host retries bigger
      }
      // ---- accumulators, one at a time over the U rows
      if (KIND == 2) {
        const int word = p.acc[0].word, vbit = p.stat_vbit;
        double x[U];
        unsigned long long img[U];
        uint32_t valid = act;
#pragma unroll
        for (int j = 0; j < U; ++j) {
          const uint64_t raw = pick_word<W>(rec[j], word);
          x[j]   = __longlong_as_double((long long)raw);
          img[j] = (raw >> 63) ? ~raw : (raw | 0x8000000000000000ull);  // ordered_image(raw, S_F64)
          if (vbit >= 0 && !((vb[j] >> vbit) & 1u)) valid &= ~(1u << j);
        }
#pragma unroll
        for (int a = 0; a < 4; ++a) {
          if (a >= p.num_accs) break;
          const HcAcc& acc = p.acc[a];
          switch (acc.op) {  // (warp-uniform)
            case ACC_COUNT_ALL: {
              const uint32_t a_cells = smem_u32(smem + acc.smem_off);
#pragma unroll
              for (int j = 0; j < U; ++j) if ((act >> j) & 1u) reds_add(a_cells + slot[j] * 4, 1u);
              break;
            }
            case ACC_COUNT_VALID: {
              const uint32_t a_cells = smem_u32(smem + acc.smem_off);
#pragma unroll
              for (int j = 0; j < U; ++j) if ((valid >> j) & 1u) reds_add(a_cells + slot[j] * 4, 1u);
              break;
            }
            case ACC_SUM_F64: case ACC_SUMSQ_F64: {
              double* c = reinterpret_cast<double*>(smem + acc.smem_off);
              const bool sqr = acc.op == ACC_SUMSQ_F64;
#pragma unroll
              for (int j = 0; j < U; ++j) if ((valid >> j) & 1u) atomicAdd(c + slot[j], sqr ? x[j] * x[j] : x[j]);
              break;
            }
            case ACC_MIN: {
              unsigned long long* c = reinterpret_cast<unsigned long long*>(smem + acc.smem_off);
#pragma unroll
              for (int j = 0; j < U; ++j)
                if (((valid >> j) & 1u) && img[j] < *reinterpret_cast<volatile unsigned long long*>(c + slot[j])) atomicMin(c + slot[j], img[j]);
              break;
            }
            default: {  // ACC_MAX
              unsigned long long* c = reinterpret_cast<unsigned long long*>(smem + acc.smem_off);
#pragma unroll
              for (int j = 0; j < U; ++j)
                if (((valid >> j) & 1u) && img[j] > *reinterpret_cast<volatile unsigned long long*>(c + slot[j])) atomicMax(c + slot[j], img[j]);
              break;
            }
          }
        }
      } else if (LEAN) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
          if (a >= p.num_accs) break;
          const uint32_t a_cells = smem_u32(smem + p.acc[a].smem_off);
          if (p.acc[a].mode == HM_COUNT) {
#pragma unroll
            for (int j = 0; j < U; ++j) if ((act >> j) & 1u) reds_add(a_cells + slot[j] * 4, 1u);
          } else if (p.acc[a].mode == HM_ADD32) {
            const int word = p.acc[a].word;
#pragma unroll
            for (int j = 0; j < U; ++j) if ((act >> j) & 1u) reds_add(a_cells + slot[j] * 4, (uint32_t)pick_word<W>(rec[j], word));
          } else {
            const int word = p.acc[a].word;
#pragma unroll
            for (int j = 0; j < U; ++j) {
              if (!((act >> j) & 1u)) continue;
              const uint64_t v   = p.val_hi32 ? (uint64_t)(int64_t)(int32_t)(rec[j][0] >> 32) : pick_word<W>(rec[j], word);
              const uint32_t vlo = (uint32_t)v, vhi = (uint32_t)(v >> 32);
              const uint32_t old = atoms_add_ret(a_cells + slot[j] * 8, vlo);
              const uint32_t add = vhi + ((uint32_t)(old + vlo) < old ? 1u : 0u);
              if (add) reds_add(a_cells + slot[j] * 8 + 4, add);
            }
          }
        }
      } else
#pragma unroll 1
      for (int a = 0; a < p.num_accs; ++a) {
        const HcAcc& acc = p.acc[a];
        const int vbit = acc.op == ACC_COUNT_ALL ? -1 : acc.vbit;
        uint32_t on = act;
        if (vbit >= 0) {
#pragma unroll
          for (int j = 0; j < U; ++j) if (!((vb[j] >> vbit) & 1u)) on &= ~(1u << j);
        }
        const uint32_t a_cells = smem_u32(smem + acc.smem_off);
        if (acc.mode == HM_COUNT) {
#pragma unroll
          for (int j = 0; j < U; ++j) if ((on >> j) & 1u) reds_add(a_cells + slot[j] * 4, 1u);
        } else if (acc.mode == HM_ADD32) {
#pragma unroll
          for (int j = 0; j < U; ++j) if ((on >> j) & 1u) reds_add(a_cells + slot[j] * 4, (uint32_t)pick_word<W>(rec[j], acc.word));
        } else if (acc.mode == HM_ISUM) {
          // lo += vlo with the old value returned: the add carried iff the new low word is smaller than the old one
          auto add64 = [&](uint32_t cell, uint64_t v) {
            const uint32_t vlo = (uint32_t)v, vhi = (uint32_t)(v >> 32);
            const uint32_t old = atoms_add_ret(cell, vlo);
            const uint32_t add = vhi + ((uint32_t)(old + vlo) < old ? 1u : 0u);
            if (add) reds_add(cell + 4, add);
          };
          if (acc.op == ACC_SUM_I64 && (acc.src == S_I64 || acc.src == S_U64)) {  // the common shape: the word is the addend
#pragma unroll
            for (int j = 0; j < U; ++j) if ((on >> j) & 1u) add64(a_cells + slot[j] * 8, pick_word<W>(rec[j], acc.word));
          } else {
            const int wd      = src_width(acc.src);
            const int sh      = (acc.src == S_I8 || acc.src == S_I16 || acc.src == S_I32) ? 64 - 8 * wd : 0;
            const bool isbool = acc.src == S_BOOL;
            const uint8_t op  = acc.op;
#pragma unroll
            for (int j = 0; j < U; ++j) {
              if (!((on >> j) & 1u)) continue;
              const uint64_t raw = pick_word<W>(rec[j], acc.word);
              int64_t v = ((int64_t)(raw << sh)) >> sh;
              if (isbool) v = raw != 0;
              if (op == ACC_SUMSQ_I64) v = (int64_t)((uint64_t)v * (uint64_t)v);
              else if (op == ACC_SUM_LO32) v = (int64_t)(uint32_t)v;
              else if (op == ACC_SUM_HI32) v = v >> 32;
              add64(a_cells + slot[j] * 8, (uint64_t)v);
            }
          }
        } else if (acc.mode == HM_MINMAX) {
          const bool is_min = acc.op == ACC_MIN;
          const int src     = acc.src;
          unsigned long long* c = reinterpret_cast<unsigned long long*>(smem + acc.smem_off);
#pragma unroll
          for (int j = 0; j < U; ++j) {
            if (!((on >> j) & 1u)) continue;
            const unsigned long long v   = ordered_image(pick_word<W>(rec[j], acc.word), src);
            const unsigned long long cv  = *reinterpret_cast<volatile unsigned long long*>(c + slot[j]);
            if (is_min) { if (v < cv) atomicMin(c + slot[j], v); } else { if (v > cv) atomicMax(c + slot[j], v); }
          }
        } else {
          void* cells = smem + acc.smem_off;
#pragma unroll
          for (int j = 0; j < U; ++j) if ((on >> j) & 1u) apply_value(acc.op, acc.src, cells, slot[j], pick_word<W>(rec[j], acc.word), 0);
        }
      }
    }
    if (n == 0) {  // nothing consumed the prefetch of this bucket: fetch the next bucket's first rows now
#pragma unroll
      for (int j = 0; j < U; ++j) { const uint32_t r = j * THREADS + t; if (r < n_next) load_rec<W>(rows_next + (uint64_t)r * W, nxt[j]); }
    }
    __syncthreads();

    if (p.peer.world > 0) {
      // ---- fused exchange: the slice's groups go straight into their OWNER ranks' regions (peer stores over NVLink), ordered by
      // owner in shared memory so that consecutive threads write consecutive records of one region
      uint16_t* s_perm = reinterpret_cast<uint16_t*>(smem + p.peer.perm_off);
      uint8_t* s_own   = reinterpret_cast<uint8_t*>(s_perm + S);
      const int world  = p.peer.world;
      if (t < (uint32_t)world) { s_pcnt[t] = 0; s_pfill[t] = 0; }
      __syncthreads();
      for (uint32_t l = t; l < S; l += THREADS) {
        const unsigned long long k = s_key[l];
        if (k != kEmptyK) atomicAdd(&s_pcnt[owner_mix(hash_key(k, 0, 1), (uint32_t)world)], 1u);
      }
      __syncthreads();
      if (t == 0) { uint32_t run = 0; for (int o = 0; o < world; ++o) { s_poff[o] = run; run += s_pcnt[o]; } s_poff[world] = run; }
      if (t < (uint32_t)world) s_pbase[t] = s_pcnt[t] ? atomicAdd(p.peer.cursors + t, s_pcnt[t]) : 0u;
      __syncthreads();
      for (uint32_t l = t; l < S; l += THREADS) {
        const unsigned long long k = s_key[l];
        if (k == kEmptyK) continue;
        const uint32_t owner = owner_mix(hash_key(k, 0, 1), (uint32_t)world);
        const uint32_t i     = s_poff[owner] + atomicAdd(&s_pfill[owner], 1u);
        s_perm[i] = (uint16_t)l;
        s_own[i]  = (uint8_t)owner;
      }
      __syncthreads();
      const uint32_t total = s_poff[world];
      const uint64_t seg   = (uint64_t)p.peer.me * p.peer.cap;
      for (uint32_t i = t; i < total; i += THREADS) {
        const uint32_t l = s_perm[i], owner = s_own[i];
        const uint64_t pos = (uint64_t)s_pbase[owner] + (i - s_poff[owner]);
        if (pos >= p.peer.cap) continue;  // over capacity: counted (the signal kernel reports the overflow to every rank), not written
        uint8_t* region = p.peer.region[owner];
        reinterpret_cast<uint64_t*>(region)[seg + pos] = s_key[l];
#pragma unroll 1
        for (int a = 0; a < p.num_accs; ++a) {
          const HcAcc& acc = p.acc[a];
          if (acc.cell_bytes == 8) reinterpret_cast<uint64_t*>(region + p.peer.acc_off[a])[seg + pos] = reinterpret_cast<const uint64_t*>(smem + acc.smem_off)[l];
          else reinterpret_cast<uint32_t*>(region + p.peer.acc_off[a])[seg + pos] = reinterpret_cast<const uint32_t*>(smem + acc.smem_off)[l];
        }
      }
      __syncthreads();
    } else {
    // ---- dense output: occupied slots of the slice, appended to the group arrays
    uint32_t mine = 0;
    for (uint32_t l = t; l < S; l += THREADS) mine += s_key[l] != kEmptyK;
    uint32_t incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, incl, d); if ((t & 31) >= (uint32_t)d) incl += u; }
    if ((t & 31) == 31) s_warp[t >> 5] = incl;
    __syncthreads();
    uint32_t woff = 0, total = 0;
#pragma unroll
    for (int w = 0; w < THREADS / 32; ++w) { const uint32_t u = s_warp[w]; if (w < (int)(t >> 5)) woff += u; total += u; }
    if (t == 0) s_base = total ? atomicAdd(p.gcount, (unsigned long long)total) : 0ull;
    __syncthreads();
    uint64_t o = s_base + woff + incl - mine;
    for (uint32_t l = t; l < S; l += THREADS) {
      const unsigned long long k = s_key[l];
      if (k == kEmptyK) continue;
      p.table.keys[o]  = k;
      p.group_slots[o] = (uint32_t)o;
#pragma unroll 1
      for (int a = 0; a < p.num_accs; ++a) {
        const HcAcc& acc = p.acc[a];
        if (acc.cell_bytes == 8) reinterpret_cast<uint64_t*>(acc.cells)[o] = reinterpret_cast<const uint64_t*>(smem + acc.smem_off)[l];
        else reinterpret_cast<uint32_t*>(acc.cells)[o] = reinterpret_cast<const uint32_t*>(smem + acc.smem_off)[l];
      }
      ++o;
    }
    __syncthreads();
    }
    fb = fb_next; n = n_next; rows = rows_next;
  }
  if (overflow) p.table.status[0] = 1;
}

// status words, group counter and the reserved slot (home of the one key whose image equals EMPTY) before the slices run
__global__ void hc_init_kernel(const __grid_constant__ HcSliceParams p)
{
  const int a = threadIdx.x;
  const uint64_t cap = p.table.capacity;
  if (a == 0) {
    p.table.keys[cap] = kEmpty;
    p.table.status[0] = 0; p.table.status[1] = 0; p.table.status[2] = 0; p.table.status[3] = 0; p.table.status[6] = 0; p.table.status[7] = 0;
    *p.gcount = 0ull;
  }
  if (a < p.num_accs) {
    const uint8_t op = p.acc[a].op;
    if (p.acc[a].cell_bytes == 8) reinterpret_cast<uint64_t*>(p.acc[a].cells)[cap] = acc_identity64(op);
    else reinterpret_cast<uint32_t*>(p.acc[a].cells)[cap] = acc_identity32(op);
  }
}

__global__ void hc_finish_kernel(const __grid_constant__ HcSliceParams p)
{
  if (p.peer.world > 0) {  // fused exchange: the group of the all-ones key (reserved slot) travels like any other record
    if (p.table.status[2] != 0) {
      const uint64_t cap   = p.table.capacity;
      const uint32_t owner = owner_mix(hash_key(kEmpty, 0, 1), (uint32_t)p.peer.world);
      const uint64_t pos   = atomicAdd(p.peer.cursors + owner, 1u);
      if (pos < p.peer.cap) {
        uint8_t* region    = p.peer.region[owner];
        const uint64_t seg = (uint64_t)p.peer.me * p.peer.cap;
        reinterpret_cast<uint64_t*>(region)[seg + pos] = kEmpty;
        for (int a = 0; a < p.num_accs; ++a) {
          if (p.acc[a].cell_bytes == 8) reinterpret_cast<uint64_t*>(region + p.peer.acc_off[a])[seg + pos] = reinterpret_cast<const uint64_t*>(p.acc[a].cells)[cap];
          else reinterpret_cast<uint32_t*>(region + p.peer.acc_off[a])[seg + pos] = reinterpret_cast<const uint32_t*>(p.acc[a].cells)[cap];
        }
      }
    }
    p.table.status[1] = 0;
    return;
  }
  unsigned long long g = *p.gcount;
  if (p.table.status[2] != 0) { p.group_slots[g] = (uint32_t)p.table.capacity; ++g; }
  p.table.status[1] = (uint32_t)g;
}

cudaError_t launch_hc_init(const HcSliceParams& p, cudaStream_t s)
{
  hc_init_kernel<<<1, 32, 0, s>>>(p);
  return note_launch();
}

template <int W, int THREADS, int KIND>
static cudaError_t launch_slices_wtl(const HcSliceParams& p, size_t smem_bytes, int num_sms, cudaStream_t s)
{
  constexpr int LEAN = KIND;
  cudaError_t e = allow_max_dynamic_smem(hc_slice_kernel<W, THREADS, LEAN>);
  if (e != cudaSuccess) return e;
  int occ = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, hc_slice_kernel<W, THREADS, LEAN>, THREADS, smem_bytes);
  if (e != cudaSuccess) return e;
  if (occ < 1) return cudaErrorInvalidConfiguration;
  const uint32_t grid = p.num_buckets < (uint32_t)(occ * num_sms) ? p.num_buckets : (uint32_t)(occ * num_sms);
  hc_slice_kernel<W, THREADS, LEAN><<<grid, THREADS, smem_bytes, s>>>(p);
  return note_launch();
}
template <int W>
static cudaError_t launch_slices_w(const HcSliceParams& p, size_t smem_bytes, int num_sms, cudaStream_t s)
{
  // threads per CTA (two CTAs per SM): measured per 1e9 rows on B200 -- SUM + COUNT (lean kernel): 8.4 ms at 512, 9.5 at 384, 9.6 at
  // 256; float64 SUM + MIN + MAX (general kernel, 78 registers): 17.2-17.9 ms whatever the count (its CAS-loop accumulators bound it)
  static const int thr_env = getenv("BGB_HC_SLICE_THREADS") ? atoi(getenv("BGB_HC_SLICE_THREADS")) : 0;
  bool lean = p.num_accs <= 3 && W <= 3;  // (W == 3: key + two value words -- the fold of partial SUM + COUNT tables)
  for (int a = 0; a < p.num_accs; ++a) {
    const HcAcc& acc = p.acc[a];
    const bool count = (acc.mode == HM_COUNT || acc.mode == HM_ADD32) && (acc.vbit < 0 || acc.op == ACC_COUNT_ALL);
    const bool isum  = acc.mode == HM_ISUM && acc.op == ACC_SUM_I64 && (acc.src == S_I64 || acc.src == S_U64) && acc.vbit < 0;
    lean = lean && (count || isum);
  }
  static const bool no_lean = getenv("BGB_HC_NO_LEAN") != nullptr;
  lean = lean && !no_lean;
  const int thr = thr_env ? thr_env : (lean ? 512 : 384);
  if (lean) {
    if (thr == 256) return launch_slices_wtl<W, 256, 1>(p, smem_bytes, num_sms, s);
    if (thr == 384) return launch_slices_wtl<W, 384, 1>(p, smem_bytes, num_sms, s);
    return launch_slices_wtl<W, 512, 1>(p, smem_bytes, num_sms, s);
  }
  if (p.f64stats && !no_lean && W >= 2) {  // (set by the host: one float64 value column, SUM / SUMSQ / MIN / MAX / COUNT over it)
    if (thr == 256) return launch_slices_wtl<W, 256, 2>(p, smem_bytes, num_sms, s);
    if (thr == 512) return launch_slices_wtl<W, 512, 2>(p, smem_bytes, num_sms, s);
    return launch_slices_wtl<W, 384, 2>(p, smem_bytes, num_sms, s);
  }
  if (thr == 256) return launch_slices_wtl<W, 256, 0>(p, smem_bytes, num_sms, s);
  if (thr == 512) return launch_slices_wtl<W, 512, 0>(p, smem_bytes, num_sms, s);
  return launch_slices_wtl<W, 384, 0>(p, smem_bytes, num_sms, s);
}

cudaError_t launch_hc_slices(const HcSliceParams& p, size_t smem_bytes, int num_sms, cudaStream_t s)
{
  switch (p.words) {
    case 1: return launch_slices_w<1>(p, smem_bytes, num_sms, s);
    case 2: return launch_slices_w<2>(p, smem_bytes, num_sms, s);
    case 3: return launch_slices_w<3>(p, smem_bytes, num_sms, s);
    case 4: return launch_slices_w<4>(p, smem_bytes, num_sms, s);
    default: return cudaErrorInvalidValue;
  }
}

cudaError_t launch_hc_finish(const HcSliceParams& p, cudaStream_t s)
{
  hc_finish_kernel<<<1, 1, 0, s>>>(p);
  return note_launch();
}

}  // namespace bgb

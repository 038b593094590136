// Shared-memory pre-aggregation kernel: the low-cardinality fast path (BASELINE config A / A9 at 1e4 groups).
//
// One persistent CTA per SM owns a private group table in shared memory (up to ~227 KB), laid out for a LOW load factor
// of the probed structure and DENSE accumulators:
//   sidx[S]   16-bit open-addressing index (linear probing), entry = dense group id, 0xFFFF = empty   (2 B per slot)
//   skd[D]    packed key of dense group id (4 or 8 bytes)
//   slo[D]    low 32 bits of the partial int64 SUM
//   shc[D]    [ high bits of the partial SUM : 32-c | row count : c ]   (one ATOMS.ADD updates both)
// With 1e4 groups this gives S = 24K index slots (load 0.41) + 11K dense rows in 227 KB; storing keys and accumulators
// in the probed slots (16 B each) would cap the table at 14.5K slots (load 0.69), where the longest of a warp's 32
// linear-probe sequences is ~11 steps (measured with ncu: profiles/r01_smem_kernel_v2_notes.txt).
// Rows are streamed from HBM in one of two ways. TMA ring (round 2): when the table leaves room for at least two stages of
// 4096 rows (true for the direct-address table at 1e4 groups: 83 KB of 227) and the key / value columns are plain 16-byte
// aligned arrays, thread 0 keeps the other stages in flight with cp.async.bulk + mbarrier (full / empty barrier per stage),
// every thread takes its four rows of a stage with 128-bit LDS and the warp's lane 0 releases the stage; no load instruction,
// no address arithmetic and no load registers in the consumer loop (the streaming microbenchmark reaches 7.3 TB/s this way
// against 6.1-6.4 TB/s for LDG.128 under a large shared-memory reservation, profiles/r01_microbench2_stream_vs_smem.txt).
// (4096-row stages: 64 KB at 16 B per row.) Otherwise (hashed table filling the shared memory, masked or unaligned columns): 128-bit coalesced loads, software-
// pipelined one chunk ahead in registers. Per row: LDS.U16 index probe + LDS key compare, ATOMS.ADD(lo) with return to
// catch the carry, ATOMS.ADD(hi|count). 64-bit shared atomics do not exist in hardware (ATOMS.CAST.SPIN loop, 3x
// slower, see profiles/r01_microbench.txt), hence the split. Values outside [-2^m, 2^m) go straight to the HBM table
// with global atomics, as do keys that no longer fit once the dense rows are used up, so any input is aggregated
// exactly; the host only routes here when the sampled cardinality says the table will hold the groups.
// Partials are flushed into the global table (find-or-insert + RED.ADD) every < 2^c rows per CTA and at the end.
//
// Replaces the reference's shared-memory path: compute_mapping_indices + single_pass_shmem_aggs_kernel
// (/root/reference/cpp/src/groupby/hash/compute_mapping_indices.cuh:23-151,
//  /root/reference/cpp/src/groupby/hash/compute_shared_memory_aggs.cu:121-353), which caps out at 128 groups per block.
#include <cstdlib>
#include <type_traits>

#include "device_utils.cuh"
#include "kernels_smem.h"

namespace bgb {

namespace {

constexpr int kThreads      = 1024;  // 32 warps: the per-row dependency chain (LDS -> LDS -> ATOMS) is hidden by warps, not ILP
constexpr int kRowsPerThread = 4;
constexpr int kChunk        = kThreads * kRowsPerThread;

// 32-bit mix whose HIGH bits select the slot (slot = umulhi(h, S)): two multiplies and one xor-shift
__device__ __forceinline__ uint32_t hash32(uint64_t k)
{
  uint32_t h = ((uint32_t)k ^ ((uint32_t)(k >> 32) * 0x85EBCA6Bu)) * 0x9E3779B1u;
  h ^= h >> 15;
  return h * 0x2C1B3C6Du;
}

template <int W>
__device__ __forceinline__ void load_pair(const uint8_t* base, int64_t idx, uint64_t& a, uint64_t& b)
{
  if (W == 8) { const ulonglong2 v = __ldcs(reinterpret_cast<const ulonglong2*>(base + idx * 8)); a = v.x; b = v.y; }
  else if (W == 4) { const uint2 v = __ldcs(reinterpret_cast<const uint2*>(base + idx * 4)); a = v.x; b = v.y; }
}
template <int W>
__device__ __forceinline__ uint64_t load_one(const uint8_t* base, int64_t idx)
{
  if (W == 8) return __ldcs(reinterpret_cast<const unsigned long long*>(base + idx * 8));
  return __ldcs(reinterpret_cast<const uint32_t*>(base + idx * 4));
}

template <int SK> struct SKey;
template <> struct SKey<8> { using type = unsigned long long; };
template <> struct SKey<4> { using type = unsigned int; };

template <int ROWS>
struct RowBufT {
  uint64_t k[ROWS];
  uint64_t v[ROWS];
  uint32_t flags;  // bit i: row i present and key usable; bit 8+i: value valid
};

// TMA ring plumbing (cp.async.bulk global -> shared, completion counted in bytes on an mbarrier)
__device__ __forceinline__ void mbar_init(uint32_t b, int c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(b), "r"(c)); }
__device__ __forceinline__ void mbar_expect(uint32_t b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(b) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t b, uint32_t parity)
{
  asm volatile("{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" ::"r"(b), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
constexpr int kTmaRows = 4096;  // rows per stage: four per thread (two 128-bit LDS of keys, two of values)


// ---- shared-window accessors on 32-bit addresses (keeps the probe loop free of generic->shared address arithmetic)
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t lds_u16(uint32_t addr)
{
  uint32_t v;
  asm volatile("ld.volatile.shared.u16 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr)
{
  uint32_t v;
  asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t lds_u64(uint32_t addr)
{
  uint64_t v;
  asm volatile("ld.volatile.shared.u64 %0, [%1];" : "=l"(v) : "r"(addr) : "memory");
  return v;
}
template <int SK>
__device__ __forceinline__ uint64_t lds_key(uint32_t addr) { return SK == 8 ? lds_u64(addr) : (uint64_t)lds_u32(addr); }
__device__ __forceinline__ uint32_t atoms_add_ret(uint32_t addr, uint32_t v)
{
  uint32_t old;
  asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(old) : "r"(addr), "r"(v) : "memory");
  return old;
}
__device__ __forceinline__ void atoms_add(uint32_t addr, uint32_t v)
{
  asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

constexpr uint32_t kIdEmpty  = 0xFFFFu;   // index entry: slot free
constexpr uint32_t kIdBypass = 0x10000u;  // row result: handled through the HBM table
constexpr uint32_t kIdSkip   = 0x10001u;  // row result: nothing to do

// slow path of one row: straight into the HBM table (value out of range, table full)
__device__ __noinline__ void smem_global_row(const SmemParams& p, uint64_t k, uint64_t v, bool vvalid, bool has_sum, int* s_abort)
{
  const uint64_t g = find_or_insert(p.a.table, k, 0);
  if (g == kNoSlot) { *s_abort = 1; return; }
  if (p.debug) atomicAdd(p.a.table.status + 4, 1u);
  if (vvalid) {
    if (has_sum && p.sum_cells) atomicAdd(p.sum_cells + g, (unsigned long long)v);
    if (p.cntv_cells) atomicAdd(p.cntv_cells + g, 1);
    if (p.cnta_cells) atomicAdd(p.cnta_cells + g, 1);
  }
}

// empty index slot seen while probing for k: take a dense row, publish the key, then try to claim the slot (16-bit CAS
// on its 32-bit word). Returns the dense id, kIdBypass when the table is full, or kIdEmpty when the slot was lost to
// another thread (the caller looks at the slot again; dense row `mine` stays unused, its count is 0, flush skips it).
template <int SK>
__device__ __noinline__ uint32_t smem_insert_key(const SmemParams& p, uint64_t k, uint32_t slot_addr, uint32_t kd_a, uint32_t* s_count,
                                                 int* s_abort)
{
  const uint32_t D = p.dense;
  if (*reinterpret_cast<volatile uint32_t*>(s_count) >= D) return kIdBypass;
  const uint32_t mine = atomicAdd(s_count, 1u);
  if (mine >= D) return kIdBypass;
  if (SK == 8) asm volatile("st.shared.u64 [%0], %1;" ::"r"(kd_a + mine * 8), "l"(k) : "memory");
  else asm volatile("st.shared.u32 [%0], %1;" ::"r"(kd_a + mine * 4), "r"((uint32_t)k) : "memory");
  __threadfence_block();
  const uint32_t sh = (slot_addr & 2u) * 8u;
  const uint32_t w  = slot_addr & ~3u;
  uint32_t cur = lds_u32(w);
  while (((cur >> sh) & 0xFFFFu) == kIdEmpty) {
    const uint32_t want = (cur & ~(0xFFFFu << sh)) | (mine << sh);
    uint32_t prev;
    asm volatile("atom.shared.cas.b32 %0, [%1], %2, %3;" : "=r"(prev) : "r"(w), "r"(cur), "r"(want) : "memory");
    if (prev == cur) {
      // first sight of this key in this CTA: make the group exist in the HBM table right away, so that groups whose
      // values are all null are not lost and flush may skip idle rows
      if (find_or_insert(p.a.table, k, 0) == kNoSlot) *s_abort = 1;
      return mine;
    }
    cur = prev;
  }
  return kIdEmpty;
}

}  // namespace

template <int SK, int KEYMODE, int VW, bool DIRECT, bool TMA>
__global__ void __launch_bounds__(kThreads, 1) agg_smem_kernel(const __grid_constant__ SmemParams p)
{
  using skey_t = typename SKey<SK>::type;
  constexpr int ROWS  = kRowsPerThread;
  constexpr int CHUNK = kThreads * ROWS;
  using RowBuf = RowBufT<ROWS>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const uint32_t S = p.slots;   // index slots (even)
  const uint32_t D = p.dense;   // dense group rows
  // hashed mode: skd[D] | slo[D] | shc[D] | sidx[S];  direct mode (dense id = key - key_base): slo[D] | shc[D] | seen[D/32]
  skey_t* skd        = reinterpret_cast<skey_t*>(smem_raw);
  uint32_t* slo      = DIRECT ? reinterpret_cast<uint32_t*>(smem_raw) : reinterpret_cast<uint32_t*>(skd + D);
  uint32_t* shc      = slo + D;
  uint32_t* sidx32   = shc + D;  // hashed: S/2 words, two 16-bit entries each; direct: "group seen" bitmap
  __shared__ uint32_t s_count;  // dense rows handed out (may exceed D: those inserts bypass)
  __shared__ int s_abort;

  for (uint32_t i = threadIdx.x; i < D; i += kThreads) { slo[i] = 0; shc[i] = 0; }
  if (DIRECT) { for (uint32_t i = threadIdx.x; i < (D + 31) / 32; i += kThreads) sidx32[i] = 0; }
  else { for (uint32_t i = threadIdx.x; i < S / 2; i += kThreads) sidx32[i] = 0xFFFFFFFFu; }
  if (threadIdx.x == 0) { s_count = 0; s_abort = 0; }
  __syncthreads();

  const AggParams& a   = p.a;
  const int64_t n      = a.num_rows;
  const int64_t nchunk = (n + CHUNK - 1) / CHUNK;
  const int c_bits     = p.cnt_bits;
  const uint32_t c_mask = (1u << c_bits) - 1u;
  const int m_bits     = p.val_bits;
  const KeyColDesc kc0 = a.kc[0];

  // ---- chunk loader: thread t owns rows {2t, 2t+1} of each 2*kThreads-row slice, so every load is a 16-byte (or
  // 8-byte) vector. `full` (block-uniform) = the chunk lies entirely inside the column: no bounds logic on that path.
  auto load_chunk_impl = [&](int64_t c, RowBuf& rb, auto full_tag) {
    constexpr bool FULL = decltype(full_tag)::value;
    rb.flags = 0;
#pragma unroll
    for (int pr = 0; pr < ROWS / 2; ++pr) {
      const int64_t r = c * CHUNK + (int64_t)pr * (kThreads * 2) + (int64_t)threadIdx.x * 2;
      const int i0 = pr * 2;
      rb.k[i0] = rb.k[i0 + 1] = 0; rb.v[i0] = rb.v[i0 + 1] = 0;
      if (!FULL && r >= n) continue;
      const bool two = FULL || r + 1 < n;
      uint32_t present = two ? 3u : 1u;
      if (KEYMODE == 1) {
        // raw loads only: anything that consumes the data (canonicalisation, sign extension, null masks) waits until
        // decode_chunk() at processing time, so these loads really stay in flight while the previous chunk is processed
        const int64_t idx = (int64_t)kc0.offset + r;
        if (two) load_pair<SK>(kc0.data, idx, rb.k[i0], rb.k[i0 + 1]);
        else rb.k[i0] = load_one<SK>(kc0.data, idx);
      } else {
        // packed key of rows r, r+1 (no early exit: both rows walk the key columns together)
        uint64_t ka = 0, kb = 0;
        bool skip_a = false, skip_b = !two;
#pragma unroll 1
        for (int col = 0; col < a.num_key_cols; ++col) {
          const KeyColDesc& kc = a.kc[col];
          const int64_t idx = (int64_t)kc.offset + r;
          const bool va = kc.mask == nullptr || bit_is_set(kc.mask, idx);
          const bool vb = two && (kc.mask == nullptr || bit_is_set(kc.mask, idx + 1));
          if (va) ka |= canon_key_bits(load_raw(kc.data, kc.width, idx), kc.fclass) << kc.shift;
          else if (a.skip_null_keys) skip_a = true;
          else if (kc.null_bit >= 0) ka |= 1ull << kc.null_bit;
          if (vb) kb |= canon_key_bits(load_raw(kc.data, kc.width, idx + 1), kc.fclass) << kc.shift;
          else if (a.skip_null_keys) skip_b = true;
          else if (two && kc.null_bit >= 0) kb |= 1ull << kc.null_bit;
        }
        rb.k[i0] = ka; rb.k[i0 + 1] = kb;
        if (skip_a) present &= ~1u;
        if (skip_b) present &= ~2u;
      }
      uint32_t vvalid = present;
      if (VW != 0) {
        const int64_t vidx = (int64_t)p.v_offset + r;
        if (two) load_pair<VW>(p.v_data, vidx, rb.v[i0], rb.v[i0 + 1]);
        else rb.v[i0] = load_one<VW>(p.v_data, vidx);
      }
      if (KEYMODE == 0) {
        if (VW == 4 && p.v_signed) { rb.v[i0] = (uint64_t)(int64_t)(int32_t)rb.v[i0]; rb.v[i0 + 1] = (uint64_t)(int64_t)(int32_t)rb.v[i0 + 1]; }
        if (p.v_mask != nullptr) {  // COUNT_VALID without a SUM still honours the value column's mask
          const int64_t vidx = (int64_t)p.v_offset + r;
          if (!bit_is_set(p.v_mask, vidx)) vvalid &= ~1u;
          if (two && !bit_is_set(p.v_mask, vidx + 1)) vvalid &= ~2u;
        }
      }
      rb.flags |= (present << i0) | (vvalid << (8 + i0));
    }
  };
  // KEYMODE 1: the consuming half of the loader, run when the chunk is processed (one chunk after its loads were issued)
  auto decode_chunk = [&](int64_t c, RowBuf& rb) {
    if (KEYMODE != 1) return;
    const bool plain = kc0.fclass == 0 && kc0.mask == nullptr && p.v_mask == nullptr && !(VW == 4 && p.v_signed);
    if (plain) return;
#pragma unroll
    for (int pr = 0; pr < ROWS / 2; ++pr) {
      const int64_t r = c * CHUNK + (int64_t)pr * (kThreads * 2) + (int64_t)threadIdx.x * 2;
      const int i0 = pr * 2;
      uint32_t present = (rb.flags >> i0) & 3u, vvalid = (rb.flags >> (8 + i0)) & 3u;
      if (kc0.fclass) { rb.k[i0] = canon_key_bits(rb.k[i0], kc0.fclass); rb.k[i0 + 1] = canon_key_bits(rb.k[i0 + 1], kc0.fclass); }
      if (VW == 4 && p.v_signed) { rb.v[i0] = (uint64_t)(int64_t)(int32_t)rb.v[i0]; rb.v[i0 + 1] = (uint64_t)(int64_t)(int32_t)rb.v[i0 + 1]; }
      if (kc0.mask != nullptr && present) {  // EXCLUDE policy: rows with a null key are dropped (direct mode never sees INCLUDE + mask)
        const int64_t idx = (int64_t)kc0.offset + r;
        if ((present & 1u) && !bit_is_set(kc0.mask, idx)) present &= ~1u;
        if ((present & 2u) && !bit_is_set(kc0.mask, idx + 1)) present &= ~2u;
      }
      vvalid &= present;
      if (p.v_mask != nullptr && vvalid) {  // COUNT_VALID without a SUM still honours the value column's mask
        const int64_t vidx = (int64_t)p.v_offset + r;
        if ((vvalid & 1u) && !bit_is_set(p.v_mask, vidx)) vvalid &= ~1u;
        if ((vvalid & 2u) && !bit_is_set(p.v_mask, vidx + 1)) vvalid &= ~2u;
      }
      rb.flags = (rb.flags & ~((3u << i0) | (3u << (8 + i0)))) | (present << i0) | (vvalid << (8 + i0));
    }
  };
  auto load_chunk = [&](int64_t c, RowBuf& rb) {
    if ((c + 1) * CHUNK <= n) load_chunk_impl(c, rb, std::true_type{});
    else load_chunk_impl(c, rb, std::false_type{});
  };

  uint32_t kd_a = smem_addr(skd), lo_a = smem_addr(slo), hc_a = smem_addr(shc), idx_a = smem_addr(sidx32);
  uint32_t idx_end = idx_a + 2 * S;
  // pin the window addresses in registers (otherwise they are re-derived from SR_CgaCtaId inside the probe loop)
  asm volatile("" : "+r"(kd_a), "+r"(lo_a), "+r"(hc_a), "+r"(idx_a), "+r"(idx_end));
  const uint64_t m_bias  = 1ull << m_bits, m_range = m_bits >= 63 ? ~0ull : (2ull << m_bits);

  // ---- one chunk: (1) every lane finds the dense id of each of its rows (divergent probe loops, inserts and bypasses
  // are rare), (2) the warp re-joins, (3) the shared atomics of all rows issue as full warps
  auto process = [&](const RowBuf& rb) {
    uint32_t id[ROWS];
#pragma unroll
    for (int i = 0; i < ROWS; ++i) {
      const uint64_t k = rb.k[i], v = rb.v[i];
      const bool present = (rb.flags >> i) & 1u, vvalid = (rb.flags >> (8 + i)) & 1u;
      id[i] = kIdSkip;
      if (present) {
        const bool in_range = VW == 0 || (v + m_bias) < m_range;  // -2^m <= v < 2^m as ONE unsigned compare
        const uint64_t didx = (k ^ p.key_flip) - p.key_base;  // direct mode: order-preserving image minus the base
        if (!in_range || (DIRECT && didx >= D)) { smem_global_row(p, k, v, vvalid, VW != 0, &s_abort); }
        else if (DIRECT) {
          if (vvalid) id[i] = (uint32_t)didx;
          else atomicOr(sidx32 + (didx >> 5), 1u << (didx & 31));  // all-null group: remember that it exists
        }
        else {
          uint32_t sa = idx_a + 2 * __umulhi(hash32(k), S);
          uint32_t e;
          while (true) {
            e = lds_u16(sa);
            if (e == kIdEmpty) {
              e = smem_insert_key<SK>(p, k, sa, kd_a, &s_count, &s_abort);
              if (e == kIdEmpty) continue;  // slot taken meanwhile: look at it again
              break;
            }
            if (lds_key<SK>(kd_a + e * SK) == (SK == 8 ? k : (uint64_t)(uint32_t)k)) break;
            sa += 2;
            if (sa == idx_end) sa = idx_a;
          }
          if (e == kIdBypass) smem_global_row(p, k, v, vvalid, VW != 0, &s_abort);
          else if (vvalid) id[i] = e;  // null value: the group exists, nothing to add (device_aggregators.cuh:434-436)
        }
      }
      if (!DIRECT) __syncwarp();  // lanes leave the probe loop at different steps: re-join before the next row
    }
    if (DIRECT) __syncwarp();
#pragma unroll
    for (int i = 0; i < ROWS; ++i) {
      if (id[i] < kIdEmpty) {
        uint32_t hi_delta = 0;
        if (VW != 0) {
          const uint32_t vlo = (uint32_t)rb.v[i];
          const uint32_t old = atoms_add_ret(lo_a + id[i] * 4, vlo);
          hi_delta = (uint32_t)(rb.v[i] >> 32) + ((uint32_t)(old + vlo) < old ? 1u : 0u);
        }
        atoms_add(hc_a + id[i] * 4, (hi_delta << c_bits) + 1u);
      }
    }
  };

  // ---- flush the CTA's partials into the HBM table; CTAs start at different slots so they do not queue on one address
  auto flush = [&]() {
    __syncthreads();
    if (p.debug && threadIdx.x == 0) atomicAdd(a.table.status + 5, 1u);
    const uint32_t used = DIRECT ? D : min(*reinterpret_cast<volatile uint32_t*>(&s_count), D);
    const uint32_t rot  = used ? (uint32_t)(((uint64_t)blockIdx.x * used) / gridDim.x) : 0;
    for (uint32_t i = threadIdx.x; i < used; i += kThreads) {
      uint32_t s = i + rot; if (s >= used) s -= used;
      const uint32_t hc = shc[s], lo = slo[s];
      bool seen = false;
      if (DIRECT && hc == 0) seen = (sidx32[s >> 5] >> (s & 31)) & 1u;
      if (hc == 0 && !seen) continue;  // no row since the last flush (the count field is never 0 once a valid row was added)
      shc[s] = 0; slo[s] = 0;
      const uint64_t key = DIRECT ? ((p.key_base + s) ^ p.key_flip) : (uint64_t)skd[s];
      const uint64_t g = find_or_insert(a.table, key, 0);
      if (g == kNoSlot) { s_abort = 1; continue; }
      if (hc == 0) continue;
      const uint32_t cnt = hc & c_mask;
      if (VW != 0 && p.sum_cells) {
        const int64_t hi = (int64_t)((int32_t)hc >> c_bits);
        atomicAdd(p.sum_cells + g, (unsigned long long)(((uint64_t)hi << 32) + lo));
      }
      if (p.cntv_cells) atomicAdd(p.cntv_cells + g, (int)cnt);
      if (p.cnta_cells) atomicAdd(p.cnta_cells + g, (int)cnt);
    }
    __syncthreads();
    if (DIRECT) {
      for (uint32_t i = threadIdx.x; i < (D + 31) / 32; i += kThreads) sidx32[i] = 0;
      __syncthreads();
    }
  };

  uint32_t rows_since_flush = 0;
  if (TMA) {
    // ---- TMA ring behind the table: [full[NS] | empty[NS]] barriers, then NS stages of keys[2048] | values[2048]
    const int NS = p.tma_stages;
    constexpr uint32_t kStageBytes = kTmaRows * (SK + VW);
    const uint32_t bar_a   = smem_addr(smem_raw + p.tma_offset);
    const uint32_t stage_a = bar_a + 128;
    const int64_t nfull    = n / CHUNK;  // whole stages; the tail (< 2048 rows) goes through the register loader below
    if (threadIdx.x == 0) {
      for (int s = 0; s < NS; ++s) { mbar_init(bar_a + 8 * s, 1); mbar_init(bar_a + 64 + 8 * s, kThreads / 32); }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const uint8_t* kbase = kc0.data + (int64_t)kc0.offset * SK;
    const uint8_t* vbase = VW ? p.v_data + (int64_t)p.v_offset * VW : nullptr;
    // producer state (thread 0): next chunk of this CTA to request, its stage and the parity its stage's EMPTY barrier must show
    int64_t pc = blockIdx.x;
    int ps = 0;
    uint32_t pph = 1;  // first pass over the ring: nothing to wait for (parity 1 = "the phase before the first" is complete)
    auto issue = [&]() {
      if (pc >= nfull) return;
      mbar_wait(bar_a + 64 + 8 * ps, pph);
      mbar_expect(bar_a + 8 * ps, kStageBytes);
      const uint32_t dst = stage_a + (uint32_t)ps * kStageBytes;
      tma_g2s(dst, kbase + pc * CHUNK * SK, kTmaRows * SK, bar_a + 8 * ps);
      if (VW) tma_g2s(dst + kTmaRows * SK, vbase + pc * CHUNK * VW, kTmaRows * VW, bar_a + 8 * ps);
      pc += gridDim.x;
      if (++ps == NS) { ps = 0; pph ^= 1u; }
    };
    if (threadIdx.x == 0) for (int j = 0; j < NS; ++j) issue();
    int cs = 0;
    uint32_t cph = 0;
    bool aborted = false;
    for (int64_t c = blockIdx.x; c < nfull; c += gridDim.x) {
      mbar_wait(bar_a + 8 * cs, cph);
      RowBuf rb;
      const uint32_t src = stage_a + (uint32_t)cs * kStageBytes;
      // thread t owns rows {2t, 2t+1} and {2048 + 2t, 2048 + 2t + 1} of the stage: conflict-free 128-bit (64-bit) LDS
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const uint32_t ko = src + (uint32_t)h * (kTmaRows / 2) * SK, vo = src + kTmaRows * SK + (uint32_t)h * (kTmaRows / 2) * VW;
        if (SK == 8) { asm volatile("ld.shared.v2.u64 {%0, %1}, [%2];" : "=l"(rb.k[2 * h]), "=l"(rb.k[2 * h + 1]) : "r"(ko + threadIdx.x * 16) : "memory"); }
        else { uint32_t x, y; asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(x), "=r"(y) : "r"(ko + threadIdx.x * 8) : "memory"); rb.k[2 * h] = x; rb.k[2 * h + 1] = y; }
        rb.v[2 * h] = rb.v[2 * h + 1] = 0;
        if (VW == 8) { asm volatile("ld.shared.v2.u64 {%0, %1}, [%2];" : "=l"(rb.v[2 * h]), "=l"(rb.v[2 * h + 1]) : "r"(vo + threadIdx.x * 16) : "memory"); }
        else if (VW == 4) {
          uint32_t x, y;
          asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(x), "=r"(y) : "r"(vo + threadIdx.x * 8) : "memory");
          rb.v[2 * h] = p.v_signed ? (uint64_t)(int64_t)(int32_t)x : x; rb.v[2 * h + 1] = p.v_signed ? (uint64_t)(int64_t)(int32_t)y : y;
        }
      }
      rb.flags = 0x0F0Fu;
      __syncwarp();
      if ((threadIdx.x & 31) == 0) mbar_arrive(bar_a + 64 + 8 * cs);  // this warp is done with the stage
      if (threadIdx.x == 0) issue();                                  // refill it once every warp has released it
      if (++cs == NS) { cs = 0; cph ^= 1u; }
      if (aborted) continue;  // (table overflow: the host retries bigger; keep draining the ring -- a CTA must not exit with bulk copies in flight)
      process(rb);
      rows_since_flush += CHUNK;
      if (rows_since_flush + CHUNK > c_mask) {
        flush();
        rows_since_flush = 0;
        if (*reinterpret_cast<volatile int*>(&s_abort)) aborted = true;
      }
    }
    // tail rows (n % 4096) through the register loader, on the CTA whose turn it would be
    if (nfull * CHUNK < n && (int64_t)blockIdx.x == nfull % gridDim.x) {
      RowBuf rb;
      load_chunk(nfull, rb);
      decode_chunk(nfull, rb);
      process(rb);
    }
  } else {
  RowBuf cur, nxt;
  int64_t c = blockIdx.x;
  if (c < nchunk) load_chunk(c, cur);
  for (; c < nchunk; c += gridDim.x) {
    __syncwarp();  // the trip count is block-uniform: re-join lanes that drifted apart in the previous iteration
    const int64_t cn = c + gridDim.x;
    if (cn < nchunk) load_chunk(cn, nxt); else nxt.flags = 0;
    decode_chunk(c, cur);
    process(cur);
    cur = nxt;
    rows_since_flush += CHUNK;
    if (rows_since_flush + CHUNK > c_mask) {  // block-uniform: count field would overflow
      flush();
      rows_since_flush = 0;
      if (*reinterpret_cast<volatile int*>(&s_abort)) break;
    }
  }
  }
  flush();
  if (threadIdx.x == 0 && s_abort) a.table.status[0] = 1;
}

// ------------------------------------------------------------------ host side
size_t smem_table_bytes(uint32_t slots, uint32_t dense, int smem_key_bytes, bool direct)
{
  if (direct) return (size_t)dense * 8 + (((size_t)dense + 31) / 32) * 4;
  return (size_t)dense * (smem_key_bytes + 8) + (size_t)slots * 2;
}

size_t smem_table_max_bytes()
{
  // 227 KB opt-in maximum per CTA minus the static __shared__ words of the kernel
  return 232448 - 64;
}

template <int SK, int KEYMODE, int VW, bool DIRECT, bool TMA = false>
static cudaError_t launch_one(const SmemParams& p, int grid, size_t smem, cudaStream_t s)
{
  auto kern = agg_smem_kernel<SK, KEYMODE, VW, DIRECT, TMA>;
  cudaError_t e = allow_max_dynamic_smem(kern);
  if (e != cudaSuccess) return e;
  kern<<<grid, kThreads, smem, s>>>(p);
  return note_launch();
}

cudaError_t launch_aggregate_smem(const SmemParams& p_in, int smem_key_bytes, int key_mode, int value_width, int num_sms, cudaStream_t s)
{
  if (p_in.a.num_rows == 0) return cudaSuccess;
  SmemParams p = p_in;
  const bool direct = p.direct != 0;
  size_t smem = smem_table_bytes(p.slots, p.dense, smem_key_bytes, direct);
  // ---- TMA ring when the columns are plain, aligned arrays and the table leaves room for >= 2 stages of 2048 rows
  static const bool no_tma = getenv("BGB_SMEM_NO_TMA") != nullptr;
  const KeyColDesc& kc = p.a.kc[0];
  const size_t stage = (size_t)kTmaRows * (smem_key_bytes + value_width);
  const size_t ring0 = (smem + 127) & ~(size_t)127;
  int stages = (int)((smem_table_max_bytes() - 1024 - ring0 - 128) / stage);
  if (stages > 3) stages = 3;
  bool tma = !no_tma && key_mode == 1 && stages >= 2 && kc.mask == nullptr && kc.fclass == 0 && p.v_mask == nullptr &&
             ((uintptr_t)kc.data + (size_t)kc.offset * smem_key_bytes) % 16 == 0 &&
             (value_width == 0 || ((uintptr_t)p.v_data + (size_t)p.v_offset * value_width) % 16 == 0) && p.a.num_rows >= 4 * kTmaRows;
  p.tma_stages = 0; p.tma_offset = 0;
  const int rows_per_chunk = tma ? kTmaRows : kChunk;
  if (tma) { p.tma_stages = stages; p.tma_offset = (uint32_t)ring0; smem = ring0 + 128 + (size_t)stages * stage; }
  const int64_t nchunk = (p.a.num_rows + rows_per_chunk - 1) / rows_per_chunk;
  const int grid = (int)(nchunk < num_sms ? nchunk : num_sms);
#define BGB_SMEM_CASE(SK, KM, VW) \
  if (smem_key_bytes == SK && key_mode == KM && value_width == VW)       \
    return direct ? launch_one<SK, KM, VW, true>(p, grid, smem, s) : launch_one<SK, KM, VW, false>(p, grid, smem, s);
#define BGB_SMEM_TMA_CASE(SK, VW) \
  if (tma && smem_key_bytes == SK && value_width == VW)       \
    return direct ? launch_one<SK, 1, VW, true, true>(p, grid, smem, s) : launch_one<SK, 1, VW, false, true>(p, grid, smem, s);
  BGB_SMEM_TMA_CASE(8, 8) BGB_SMEM_TMA_CASE(8, 4) BGB_SMEM_TMA_CASE(8, 0)
  BGB_SMEM_TMA_CASE(4, 8) BGB_SMEM_TMA_CASE(4, 4) BGB_SMEM_TMA_CASE(4, 0)
  BGB_SMEM_CASE(8, 1, 8) BGB_SMEM_CASE(8, 1, 4) BGB_SMEM_CASE(8, 1, 0)
  BGB_SMEM_CASE(4, 1, 8) BGB_SMEM_CASE(4, 1, 4) BGB_SMEM_CASE(4, 1, 0)
  BGB_SMEM_CASE(8, 0, 8) BGB_SMEM_CASE(8, 0, 4) BGB_SMEM_CASE(8, 0, 0)
  BGB_SMEM_CASE(4, 0, 8) BGB_SMEM_CASE(4, 0, 4) BGB_SMEM_CASE(4, 0, 0)
#undef BGB_SMEM_CASE
#undef BGB_SMEM_TMA_CASE
  return cudaErrorInvalidValue;
}

}  // namespace bgb

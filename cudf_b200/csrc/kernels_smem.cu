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
// Rows are streamed from HBM with 128-bit coalesced loads, software-pipelined one chunk ahead in registers (the whole
// shared memory belongs to the table, so there is no room for a TMA staging ring here; the TMA-staged kernels are the
// tiny-cardinality and partition kernels). Per row: LDS.U16 index probe + LDS key compare, ATOMS.ADD(lo) with return to
// catch the carry, ATOMS.ADD(hi|count). 64-bit shared atomics do not exist in hardware (ATOMS.CAST.SPIN loop, 3x
// slower, see profiles/r01_microbench.txt), hence the split. Values outside [-2^m, 2^m) go straight to the HBM table
// with global atomics, as do keys that no longer fit once the dense rows are used up, so any input is aggregated
// exactly; the host only routes here when the sampled cardinality says the table will hold the groups.
// Partials are flushed into the global table (find-or-insert + RED.ADD) every < 2^c rows per CTA and at the end.
//
// Replaces the reference's shared-memory path: compute_mapping_indices + single_pass_shmem_aggs_kernel
// (/root/reference/cpp/src/groupby/hash/compute_mapping_indices.cuh:23-151,
//  /root/reference/cpp/src/groupby/hash/compute_shared_memory_aggs.cu:121-353), which caps out at 128 groups per block.
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

struct RowBuf {
  uint64_t k[kRowsPerThread];
  uint64_t v[kRowsPerThread];
  uint32_t flags;  // bit i: row i present and key usable; bit 8+i: value valid
};


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

template <int SK, int KEYMODE, int VW, bool DIRECT>
__global__ void __launch_bounds__(kThreads, 1) agg_smem_kernel(const __grid_constant__ SmemParams p)
{
  using skey_t = typename SKey<SK>::type;
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
  const int64_t nchunk = (n + kChunk - 1) / kChunk;
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
    for (int pr = 0; pr < kRowsPerThread / 2; ++pr) {
      const int64_t r = c * kChunk + (int64_t)pr * (kThreads * 2) + (int64_t)threadIdx.x * 2;
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
    for (int pr = 0; pr < kRowsPerThread / 2; ++pr) {
      const int64_t r = c * kChunk + (int64_t)pr * (kThreads * 2) + (int64_t)threadIdx.x * 2;
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
    if ((c + 1) * kChunk <= n) load_chunk_impl(c, rb, std::true_type{});
    else load_chunk_impl(c, rb, std::false_type{});
  };

  uint32_t kd_a = smem_addr(skd), lo_a = smem_addr(slo), hc_a = smem_addr(shc), idx_a = smem_addr(sidx32);
  uint32_t idx_end = idx_a + 2 * S;
  // pin the window addresses in registers (otherwise they are re-derived from SR_CgaCtaId inside the probe loop)
  asm volatile("" : "+r"(kd_a), "+r"(lo_a), "+r"(hc_a), "+r"(idx_a), "+r"(idx_end));
  const uint64_t m_bias  = 1ull << m_bits;

  // ---- one chunk: (1) every lane finds the dense id of each of its rows (divergent probe loops, inserts and bypasses
  // are rare), (2) the warp re-joins, (3) the shared atomics of all rows issue as full warps
  auto process = [&](const RowBuf& rb) {
    uint32_t id[kRowsPerThread];
#pragma unroll
    for (int i = 0; i < kRowsPerThread; ++i) {
      const uint64_t k = rb.k[i], v = rb.v[i];
      const bool present = (rb.flags >> i) & 1u, vvalid = (rb.flags >> (8 + i)) & 1u;
      id[i] = kIdSkip;
      if (present) {
        const bool in_range = VW == 0 || ((v + m_bias) >> (m_bits + 1)) == 0;
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
    for (int i = 0; i < kRowsPerThread; ++i) {
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

  RowBuf cur, nxt;
  int64_t c = blockIdx.x;
  uint32_t rows_since_flush = 0;
  if (c < nchunk) load_chunk(c, cur);
  for (; c < nchunk; c += gridDim.x) {
    __syncwarp();  // the trip count is block-uniform: re-join lanes that drifted apart in the previous iteration
    const int64_t cn = c + gridDim.x;
    if (cn < nchunk) load_chunk(cn, nxt); else nxt.flags = 0;
    decode_chunk(c, cur);
    process(cur);
    cur = nxt;
    rows_since_flush += kChunk;
    if (rows_since_flush + kChunk > c_mask) {  // block-uniform: count field would overflow
      flush();
      rows_since_flush = 0;
      if (*reinterpret_cast<volatile int*>(&s_abort)) break;
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

template <int SK, int KEYMODE, int VW, bool DIRECT>
static cudaError_t launch_one(const SmemParams& p, int grid, size_t smem, cudaStream_t s)
{
  auto kern = agg_smem_kernel<SK, KEYMODE, VW, DIRECT>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  kern<<<grid, kThreads, smem, s>>>(p);
  return note_launch();
}

cudaError_t launch_aggregate_smem(const SmemParams& p, int smem_key_bytes, int key_mode, int value_width, int num_sms, cudaStream_t s)
{
  if (p.a.num_rows == 0) return cudaSuccess;
  const bool direct = p.direct != 0;
  const size_t smem = smem_table_bytes(p.slots, p.dense, smem_key_bytes, direct);
  const int64_t nchunk = (p.a.num_rows + kChunk - 1) / kChunk;
  const int grid = (int)(nchunk < num_sms ? nchunk : num_sms);
#define BGB_SMEM_CASE(SK, KM, VW) \
  if (smem_key_bytes == SK && key_mode == KM && value_width == VW)       \
    return direct ? launch_one<SK, KM, VW, true>(p, grid, smem, s) : launch_one<SK, KM, VW, false>(p, grid, smem, s);
  BGB_SMEM_CASE(8, 1, 8) BGB_SMEM_CASE(8, 1, 4) BGB_SMEM_CASE(8, 1, 0)
  BGB_SMEM_CASE(4, 1, 8) BGB_SMEM_CASE(4, 1, 4) BGB_SMEM_CASE(4, 1, 0)
  BGB_SMEM_CASE(8, 0, 8) BGB_SMEM_CASE(8, 0, 4) BGB_SMEM_CASE(8, 0, 0)
  BGB_SMEM_CASE(4, 0, 8) BGB_SMEM_CASE(4, 0, 4) BGB_SMEM_CASE(4, 0, 0)
#undef BGB_SMEM_CASE
  return cudaErrorInvalidValue;
}

}  // namespace bgb

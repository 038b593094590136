// Device helpers: element loads, key packing, ordered images, open-addressing probe.
#pragma once
#include "bgb_internal.cuh"
#include "errors.h"

namespace bgb {

__device__ __forceinline__ bool bit_is_set(const uint32_t* mask, int64_t i)
{
  return (__ldg(mask + (i >> 5)) >> (i & 31)) & 1u;
}

// raw little-endian bits of one element, zero-extended
__device__ __forceinline__ uint64_t load_raw(const uint8_t* data, int width, int64_t idx)
{
  switch (width) {
    case 1: return __ldg(data + idx);
    case 2: return __ldg(reinterpret_cast<const uint16_t*>(data) + idx);
    case 4: return __ldg(reinterpret_cast<const uint32_t*>(data) + idx);
    default: return __ldg(reinterpret_cast<const unsigned long long*>(data) + idx);
  }
}

// key element canonicalisation: -0 -> +0, any NaN -> one quiet NaN
// (/root/reference/cpp/include/cudf/hashing/detail/hash_functions.cuh:19-39; equality.cuh nan_equal comparator)
__device__ __forceinline__ uint64_t canon_key_bits(uint64_t raw, int fclass)
{
  if (fclass == 1) {
    uint32_t u = (uint32_t)raw;
    if ((u & 0x7fffffffu) == 0) return 0;
    if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fc00000u;
    return u;
  }
  if (fclass == 2) {
    if ((raw & 0x7fffffffffffffffull) == 0) return 0;
    if ((raw & 0x7fffffffffffffffull) > 0x7ff0000000000000ull) return 0x7ff8000000000000ull;
  }
  if (fclass == 3) return raw != 0;  // BOOL8: element<bool>() reads any non-zero byte as true
  return raw;
}

// ---- the reference's row hash for hash_partition: MurmurHash3_x86_32 of every element's bytes (seeded), nulls hash to
// UINT32_MAX, columns folded with boost's hash_combine (cudf/hashing/detail/murmurhash3_x86_32.cuh:26-60,
// cudf/detail/row_operator/hashing.cuh:40-134, cudf/hashing/detail/hashing.hpp:83-86). The element is hashed in its native
// width after the same canonicalisation as the key image (-0 -> +0, one NaN, bool -> 0/1).
__device__ __forceinline__ uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }
__device__ __forceinline__ uint32_t murmur3_fixed(uint64_t bits, int width, uint32_t seed)
{
  constexpr uint32_t c1 = 0xcc9e2d51u, c2 = 0x1b873593u;
  uint32_t h = seed;
  auto block = [&](uint32_t k) { k *= c1; k = rotl32(k, 15); k *= c2; h ^= k; h = rotl32(h, 13); h = h * 5u + 0xe6546b64u; };
  if (width >= 4) block((uint32_t)bits);
  if (width == 8) block((uint32_t)(bits >> 32));
  if (width < 4) {  // 1- and 2-byte elements are all tail
    uint32_t k = (uint32_t)bits & (width == 1 ? 0xffu : 0xffffu);
    k *= c1; k = rotl32(k, 15); k *= c2; h ^= k;
  }
  h ^= (uint32_t)width;
  h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
  return h;
}
__device__ __forceinline__ uint32_t hash_combine32(uint32_t lhs, uint32_t rhs) { return lhs ^ (rhs + 0x9e3779b9u + (lhs << 6) + (lhs >> 2)); }

// mode 1 = MurmurHash3_x86_32, mode 2 = the reference's IdentityHash (static_cast<uint32_t>(element): sign-extended
// integers truncated to 32 bits, floats converted with the saturating cvt.rzi.u32; partitioning.cu:855-874)
__device__ __forceinline__ uint32_t cudf_row_hash(const AggParams& p, int64_t row, uint32_t seed, int mode, uint32_t signed_cols)
{
  uint32_t h = seed;
#pragma unroll 1
  for (int c = 0; c < p.num_key_cols; ++c) {
    const KeyColDesc& kc = p.kc[c];
    const int64_t idx    = (int64_t)kc.offset + row;
    uint32_t e           = 0xffffffffu;  // null element
    if (kc.mask == nullptr || bit_is_set(kc.mask, idx)) {
      const uint64_t raw = load_raw(kc.data, kc.width, idx);
      if (mode == 1) e = murmur3_fixed(canon_key_bits(raw, kc.fclass), kc.width, seed);
      else if (kc.fclass == 1) e = __float2uint_rz(__uint_as_float((uint32_t)raw));
      else if (kc.fclass == 2) e = __double2uint_rz(__longlong_as_double((long long)raw));
      else if (kc.fclass == 3) e = raw != 0;
      else if (kc.width == 1 && ((signed_cols >> c) & 1)) e = (uint32_t)(int32_t)(int8_t)raw;
      else if (kc.width == 2 && ((signed_cols >> c) & 1)) e = (uint32_t)(int32_t)(int16_t)raw;
      else e = (uint32_t)raw;
    }
    h = c == 0 ? e : hash_combine32(h, e);
  }
  return h;
}

// packs the key columns of `row` into (k0,k1); returns false when the row must be skipped (EXCLUDE + null key)
__device__ __forceinline__ bool pack_key(const AggParams& p, int64_t row, uint64_t& k0, uint64_t& k1)
{
  k0 = 0; k1 = 0;
  // the common shape -- one integer-like key column without a mask -- is a single load (warp-uniform test); the general
  // loop below costs ~30 instructions per row, which is what bounded the histogram and scatter passes
  if (p.num_key_cols == 1 && p.kc[0].mask == nullptr && p.kc[0].fclass == 0) {
    k0 = load_raw(p.kc[0].data, p.kc[0].width, (int64_t)p.kc[0].offset + row);
    return true;
  }
#pragma unroll 1
  for (int c = 0; c < p.num_key_cols; ++c) {
    const KeyColDesc& kc = p.kc[c];
    int64_t idx = (int64_t)kc.offset + row;
    bool valid = kc.mask == nullptr || bit_is_set(kc.mask, idx);
    if (!valid) {
      if (p.skip_null_keys) return false;
      if (kc.null_bit >= 0) { if (kc.null_bit < 64) k0 |= 1ull << kc.null_bit; else k1 |= 1ull << (kc.null_bit - 64); }
      continue;  // value bits of a null key stay zero: all nulls are equal (null_equality::EQUAL)
    }
    uint64_t v = canon_key_bits(load_raw(kc.data, kc.width, idx), kc.fclass);
    if (kc.shift < 64) k0 |= v << kc.shift; else k1 |= v << (kc.shift - 64);
  }
  return true;
}

__device__ __forceinline__ uint64_t hash_key(uint64_t k0, uint64_t k1, int key_words)
{
  uint64_t h = mix64(k0 + 0x9E3779B97F4A7C15ull);
  if (key_words == 2) h = mix64(h ^ (k1 * 0xD6E8FEB86659FD93ull));
  return h;
}

__device__ __forceinline__ void cas128(uint64_t* addr, uint64_t c0, uint64_t c1, uint64_t n0, uint64_t n1, uint64_t& o0, uint64_t& o1)
{
  asm volatile(
    "{\n\t.reg .b128 c, n, o;\n\tmov.b128 c, {%2, %3};\n\tmov.b128 n, {%4, %5};\n\t"
    "atom.global.relaxed.gpu.cas.b128 o, [%6], c, n;\n\tmov.b128 {%0, %1}, o;\n\t}"
    : "=l"(o0), "=l"(o1)
    : "l"(c0), "l"(c1), "l"(n0), "l"(n1), "l"(addr)
    : "memory");
}

constexpr uint64_t kNoSlot       = ~0ull;
constexpr uint32_t kMaxProbes    = 8192;  // beyond this the table is declared overloaded and the host retries bigger

// Slot hash of the bucketed path (kernels_bucket.cu) and of tables merged from its output (kernels_peer.cu). A table is
// only ever probed by the kernels that built it (compaction and emit read slots in order, whatever put the keys there), so
// those kernels may use a cheaper hash than hash_key(): the SplitMix64 finaliser costs ~35 SASS instructions (two 64-bit
// multiplies) and was a quarter of the scatter kernel. Here: Fibonacci product folded to 32 bits, then the murmur3 32-bit
// finaliser (11 instructions). 32 bits index any table the hash path can build (capacity <= 2^32 slots).
__device__ __forceinline__ uint32_t slot_hash32(uint64_t k)
{
  const uint64_t h = k * 0x9E3779B97F4A7C15ull;
  uint32_t x = (uint32_t)(h >> 32) ^ (uint32_t)h;
  x *= 0x85EBCA6Bu; x ^= x >> 13;
  x *= 0xC2B2AE35u; x ^= x >> 16;
  return x;
}

// one-word keys: linear probing from the slot a given hash names
__device__ __forceinline__ uint64_t find_or_insert_from(const TableDesc& t, uint64_t k0, uint64_t hash)
{
  const uint64_t cap = t.capacity;
  if (k0 == kEmpty) { t.status[2] = 1; return cap; }
  uint64_t slot = hash & (cap - 1);
  unsigned long long* keys = reinterpret_cast<unsigned long long*>(t.keys);
  for (uint32_t probes = 0; probes < kMaxProbes; ++probes) {
    unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(keys + slot);
    if (cur == k0) return slot;
    if (cur == kEmpty) {
      unsigned long long old = atomicCAS(keys + slot, (unsigned long long)kEmpty, (unsigned long long)k0);
      if (old == kEmpty || old == k0) return slot;
    }
    slot = (slot + 1) & (cap - 1);
  }
  t.status[0] = 1;
  return kNoSlot;
}

// linear-probing find-or-insert in the HBM table; returns slot or kNoSlot (overflow flag set)
__device__ __forceinline__ uint64_t find_or_insert(const TableDesc& t, uint64_t k0, uint64_t k1)
{
  const uint64_t cap = t.capacity;
  if (t.key_words == 1) return find_or_insert_from(t, k0, hash_key(k0, 0, 1));
  {
    if (k0 == kEmpty && k1 == kEmpty) { t.status[2] = 1; return cap; }
    uint64_t slot = hash_key(k0, k1, 2) & (cap - 1);
    for (uint32_t probes = 0; probes < kMaxProbes; ++probes) {
      const volatile uint64_t* e = t.keys + 2 * slot;
      uint64_t c0 = e[0], c1 = e[1];
      // a torn read can only mix EMPTY words with final words (slots go EMPTY -> key exactly once, atomically);
      // keys that themselves contain an all-ones word never trust the two-load fast path and always take the CAS
      const bool odd_key = (k0 == kEmpty) || (k1 == kEmpty);
      if (!odd_key && c0 == k0 && c1 == k1) return slot;
      if (odd_key || c0 == kEmpty || c1 == kEmpty) {
        uint64_t o0, o1;
        cas128(t.keys + 2 * slot, kEmpty, kEmpty, k0, k1, o0, o1);
        if ((o0 == kEmpty && o1 == kEmpty) || (o0 == k0 && o1 == k1)) return slot;
      }
      slot = (slot + 1) & (cap - 1);
    }
  }
  t.status[0] = 1;
  return kNoSlot;
}

// ---- keys wider than 128 bits: the table holds ROW NUMBERS (64-bit slots, EMPTY = all ones) and two rows are compared column by
// column in the caller's key columns -- the reference's own scheme for every key (row hasher + self-comparator over a set of
// row indices: /root/reference/cpp/src/groupby/hash/compute_groupby.cu:79-104, row_operator/equality.cuh:59-89).
__device__ __forceinline__ bool row_hash(const AggParams& p, int64_t row, uint64_t& h)
{
  h = 0x9E3779B97F4A7C15ull;
#pragma unroll 1
  for (int c = 0; c < p.num_key_cols; ++c) {
    const KeyColDesc& kc = p.kc[c];
    const int64_t idx    = (int64_t)kc.offset + row;
    uint64_t v           = 0xA5A5A5A5DEADBEEFull;  // a null element (all nulls are equal: null_equality::EQUAL)
    if (kc.mask == nullptr || bit_is_set(kc.mask, idx)) v = canon_key_bits(load_raw(kc.data, kc.width, idx), kc.fclass);
    else if (p.skip_null_keys) return false;
    h = mix64(h ^ (v + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2)));
  }
  return true;
}
__device__ __forceinline__ bool rows_equal(const AggParams& p, int64_t r1, int64_t r2)
{
#pragma unroll 1
  for (int c = 0; c < p.num_key_cols; ++c) {
    const KeyColDesc& kc = p.kc[c];
    const int64_t i1 = (int64_t)kc.offset + r1, i2 = (int64_t)kc.offset + r2;
    const bool v1 = kc.mask == nullptr || bit_is_set(kc.mask, i1), v2 = kc.mask == nullptr || bit_is_set(kc.mask, i2);
    if (v1 != v2) return false;
    if (v1 && canon_key_bits(load_raw(kc.data, kc.width, i1), kc.fclass) != canon_key_bits(load_raw(kc.data, kc.width, i2), kc.fclass)) return false;
  }
  return true;
}
// returns the slot of `row`'s key (inserting `row` as the group's representative when the key is new), kNoSlot on overflow
__device__ __forceinline__ uint64_t find_or_insert_row(const AggParams& p, int64_t row, uint64_t hash)
{
  const TableDesc& t = p.table;
  const uint64_t cap = t.capacity;
  uint64_t slot = hash & (cap - 1);
  unsigned long long* keys = reinterpret_cast<unsigned long long*>(t.keys);
  for (uint32_t probes = 0; probes < kMaxProbes; ++probes) {
    unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(keys + slot);
    if (cur == kEmpty) {
      cur = atomicCAS(keys + slot, (unsigned long long)kEmpty, (unsigned long long)row);
      if (cur == kEmpty) return slot;
    }
    if (cur == (unsigned long long)row || rows_equal(p, (int64_t)cur, row)) return slot;
    slot = (slot + 1) & (cap - 1);
  }
  t.status[0] = 1;
  return kNoSlot;
}

// ---- order-preserving unsigned images (so that MIN/MAX are one RED.MIN/MAX.U64 for every type)
__device__ __forceinline__ uint64_t ordered_image(uint64_t raw, int src)
{
  switch (src) {
    case S_I8: return (uint64_t)((int64_t)(int8_t)raw) ^ 0x8000000000000000ull;
    case S_I16: return (uint64_t)((int64_t)(int16_t)raw) ^ 0x8000000000000000ull;
    case S_I32: return (uint64_t)((int64_t)(int32_t)raw) ^ 0x8000000000000000ull;
    case S_I64: return raw ^ 0x8000000000000000ull;
    case S_F32: { uint32_t u = (uint32_t)raw; return (u & 0x80000000u) ? (uint32_t)~u : (u | 0x80000000u); }
    case S_F64: return (raw >> 63) ? ~raw : (raw | 0x8000000000000000ull);
    case S_BOOL: return raw != 0;
    default: return raw;  // unsigned
  }
}
__device__ __forceinline__ uint64_t ordered_decode(uint64_t img, int src)
{
  switch (src) {
    case S_I8: case S_I16: case S_I32: case S_I64: return img ^ 0x8000000000000000ull;
    case S_F32: { uint32_t u = (uint32_t)img; return (u & 0x80000000u) ? (u & 0x7fffffffu) : (uint32_t)~u; }
    case S_F64: return (img >> 63) ? (img & 0x7fffffffffffffffull) : ~img;
    default: return img;
  }
}

// 32-bit order image of a source of at most 32 bits (the packed (value, row) cell of ARGMIN / ARGMAX): the sign flip must
// sit in bit 31, not in bit 63 where ordered_image() puts it and where `<< 32` would drop it
__device__ __forceinline__ uint32_t ordered_image32(uint64_t raw, int src)
{
  switch (src) {
    case S_I8: return (uint32_t)(int32_t)(int8_t)raw ^ 0x80000000u;
    case S_I16: return (uint32_t)(int32_t)(int16_t)raw ^ 0x80000000u;
    case S_I32: return (uint32_t)raw ^ 0x80000000u;
    default: return (uint32_t)ordered_image(raw, src);  // F32, unsigned, bool: already 32-bit images
  }
}

__device__ __forceinline__ int src_width(int src)
{
  switch (src) {
    case S_I8: case S_U8: case S_BOOL: return 1;
    case S_I16: case S_U16: return 2;
    case S_I32: case S_U32: case S_F32: return 4;
    default: return 8;
  }
}
// integral source -> int64 (static_cast<int64_t>(Source) in update_target_element<Source, SUM>)
__device__ __forceinline__ int64_t raw_to_i64(uint64_t raw, int src)
{
  switch (src) {
    case S_I8: return (int8_t)raw;
    case S_I16: return (int16_t)raw;
    case S_I32: return (int32_t)raw;
    case S_BOOL: return raw != 0;
    default: return (int64_t)raw;  // I64, U8..U64 (uint64 reinterpreted, like the reference's static_cast)
  }
}

// CAS-loop products (no hardware atomic multiply; update_target_element<Source, PRODUCT>)
__device__ __forceinline__ void atomic_mul_i64(unsigned long long* addr, int64_t v)
{
  unsigned long long old = *addr, assumed;
  do { assumed = old; old = atomicCAS(addr, assumed, (unsigned long long)((int64_t)assumed * v)); } while (old != assumed);
}
__device__ __forceinline__ void atomic_mul_f64(unsigned long long* addr, double v)
{
  unsigned long long old = *addr, assumed;
  do {
    assumed = old;
    old     = atomicCAS(addr, assumed, (unsigned long long)__double_as_longlong(__longlong_as_double((long long)assumed) * v));
  } while (old != assumed);
}
__device__ __forceinline__ void atomic_mul_f32(unsigned int* addr, float v)
{
  unsigned int old = *addr, assumed;
  do { assumed = old; old = atomicCAS(addr, assumed, __float_as_uint(__uint_as_float(assumed) * v)); } while (old != assumed);
}

// one valid source element into one accumulator cell (update_target_element<Source, K>,
// /root/reference/cpp/include/cudf/detail/aggregation/device_aggregators.cuh:24-160). `cells` may point at the HBM table or
// at a shared-memory slice of it (kernels_bucket.cu): the atomics are the generic-address forms.
__device__ __forceinline__ void apply_value(uint8_t op, uint8_t src, void* cells, uint64_t slot, uint64_t raw, int64_t row)
{
  unsigned long long* c64 = reinterpret_cast<unsigned long long*>(cells) + slot;
  unsigned int* c32       = reinterpret_cast<unsigned int*>(cells) + slot;
  switch (op) {
    case ACC_COUNT_VALID: case ACC_COUNT_ALL: atomicAdd(c32, 1u); break;
    case ACC_SUM_I64: atomicAdd(c64, (unsigned long long)raw_to_i64(raw, src)); break;
    case ACC_SUM_F64: atomicAdd(reinterpret_cast<double*>(c64), __longlong_as_double((long long)raw)); break;
    case ACC_SUM_F32: atomicAdd(reinterpret_cast<float*>(c32), __uint_as_float((uint32_t)raw)); break;
    case ACC_SUMSQ_I64: { uint64_t v = (uint64_t)raw_to_i64(raw, src); atomicAdd(c64, (unsigned long long)(v * v)); break; }
    case ACC_SUMSQ_F64: { double v = __longlong_as_double((long long)raw); atomicAdd(reinterpret_cast<double*>(c64), v * v); break; }
    case ACC_SUMSQ_F32: { float v = __uint_as_float((uint32_t)raw); atomicAdd(reinterpret_cast<float*>(c32), v * v); break; }
    case ACC_PROD_I64: atomic_mul_i64(c64, raw_to_i64(raw, src)); break;
    case ACC_PROD_F64: atomic_mul_f64(c64, __longlong_as_double((long long)raw)); break;
    case ACC_PROD_F32: atomic_mul_f32(c32, __uint_as_float((uint32_t)raw)); break;
    case ACC_MIN: { unsigned long long v = ordered_image(raw, src); if (v < *reinterpret_cast<volatile unsigned long long*>(c64)) atomicMin(c64, v); break; }
    case ACC_MAX: { unsigned long long v = ordered_image(raw, src); if (v > *reinterpret_cast<volatile unsigned long long*>(c64)) atomicMax(c64, v); break; }
    case ACC_SUM_LO32: atomicAdd(c64, (unsigned long long)(uint32_t)raw_to_i64(raw, src)); break;
    case ACC_SUM_HI32: atomicAdd(c64, (unsigned long long)(raw_to_i64(raw, src) >> 32)); break;
    case ACC_ARGMIN: atomicMin(c64, (unsigned long long)(((uint64_t)ordered_image32(raw, src) << 32) | (uint32_t)row)); break;
    case ACC_ARGMAX: atomicMax(c64, (unsigned long long)(((uint64_t)ordered_image32(raw, src) << 32) | (0xFFFFFFFFu - (uint32_t)row))); break;
    default: break;
  }
}

// aggregation identities of the accumulator cells (cudf::detail::identity_initializer, aggregation.cu:39-97)
__device__ __forceinline__ uint64_t acc_identity64(uint8_t op)
{
  switch (op) {
    case ACC_PROD_I64: return 1ull;
    case ACC_PROD_F64: return 0x3ff0000000000000ull;  // 1.0
    case ACC_MIN: case ACC_ARGMIN: return ~0ull;
    default: return 0ull;  // sums, MAX (smallest ordered image), ARGMAX
  }
}
__device__ __forceinline__ uint32_t acc_identity32(uint8_t op)
{
  return op == ACC_PROD_F32 ? 0x3f800000u : (op == ACC_ARGROW_MIN || op == ACC_ARGROW_MAX) ? 0xFFFFFFFFu : 0u;
}

}  // namespace bgb

// Group-slot compaction (single-pass decoupled look-back prefix scan) and the fused output gather + finalise kernel.
//
// Replaces cuco::static_set::retrieve_all + key_transform_map scatter + detail::gather of keys and sparse results
// (/root/reference/cpp/src/groupby/hash/output_utils.cu:139-196, compute_groupby.cu:104-111,154,
//  compute_global_memory_aggs.cuh:150-155), the MEAN/M2/VAR/STD finalisers
// (/root/reference/cpp/src/groupby/hash/hash_compound_agg_finalizer.cu:94-186,
//  /root/reference/cpp/src/groupby/common/m2_var_std.cu:41-59,152-188) and batch_null_count
// (/root/reference/cpp/src/groupby/hash/output_utils.cu:198-224) with two launches.
#include "device_utils.cuh"

namespace bgb {

constexpr int kScanThreads = 256;
constexpr int kScanItems   = 8;
constexpr int kScanTile    = kScanThreads * kScanItems;

// tile descriptor: bits 63..62 status (1 = tile aggregate, 2 = inclusive prefix), low 62 bits value
constexpr uint64_t kTileAgg = 1ull << 62, kTileInc = 2ull << 62, kTileValMask = (1ull << 62) - 1;

size_t compact_scan_state_bytes(uint64_t capacity)
{
  uint64_t tiles = (capacity + 1 + kScanTile - 1) / kScanTile;
  return (tiles + 2) * sizeof(uint64_t);
}

__global__ void __launch_bounds__(kScanThreads) compact_kernel(TableDesc t, uint32_t* __restrict__ group_slots,
                                                               unsigned long long* state /* [0]=ticket, [1..] tiles */)
{
  __shared__ uint32_t s_tile;
  __shared__ uint32_t s_warp_sums[kScanThreads / 32];
  __shared__ uint64_t s_prefix;
  if (threadIdx.x == 0) s_tile = (uint32_t)atomicAdd(state, 1ull);  // dynamic tile id => forward progress of look-back
  __syncthreads();
  const uint32_t tile = s_tile;
  unsigned long long* desc = state + 1;
  const uint64_t n     = t.capacity + 1;  // slot `capacity` = reserved home of the all-ones key
  const uint64_t base  = (uint64_t)tile * kScanTile + (uint64_t)threadIdx.x * kScanItems;
  const bool reserved_used = t.status[2] != 0;

  uint32_t flags = 0, count = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    uint64_t slot = base + i;
    bool occ = false;
    if (slot < t.capacity) {
      if (t.key_words == 1) occ = t.keys[slot] != kEmpty;
      else { ulonglong2 k = reinterpret_cast<const ulonglong2*>(t.keys)[slot]; occ = !(k.x == kEmpty && k.y == kEmpty); }
    } else if (slot == t.capacity) occ = reserved_used;
    flags |= (uint32_t)occ << i;
    count += occ;
  }
  // block exclusive scan of per-thread counts
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t incl = count;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { uint32_t v = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += v; }
  if (lane == 31) s_warp_sums[warp] = incl;
  __syncthreads();
  uint32_t warp_off = 0, block_total = 0;
#pragma unroll
  for (int w = 0; w < kScanThreads / 32; ++w) { uint32_t v = s_warp_sums[w]; if (w < warp) warp_off += v; block_total += v; }
  const uint32_t thread_off = warp_off + incl - count;

  // decoupled look-back (warp 0)
  if (warp == 0) {
    uint64_t exclusive = 0;
    if (tile == 0) {
      if (lane == 0) atomicExch(desc + 0, kTileInc | block_total);
    } else {
      if (lane == 0) atomicExch(desc + tile, kTileAgg | block_total);
      int64_t look = (int64_t)tile - 1;
      while (true) {
        int64_t idx = look - lane;
        unsigned long long d = kTileInc;  // out-of-range lanes read as "inclusive 0"
        if (idx >= 0) { do { d = *reinterpret_cast<volatile unsigned long long*>(desc + idx); } while ((d >> 62) == 0); }
        uint32_t inc_mask = __ballot_sync(0xffffffffu, (d >> 62) == 2);
        int first_inc = inc_mask ? (__ffs(inc_mask) - 1) : 32;  // nearest preceding tile holding an inclusive prefix
        uint64_t v = (lane <= first_inc) ? (d & kTileValMask) : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        exclusive += v;
        if (inc_mask) break;
        look -= 32;
      }
      if (lane == 0) atomicExch(desc + tile, kTileInc | (exclusive + block_total));
    }
    if (lane == 0) {
      s_prefix = exclusive;
      if ((uint64_t)(tile + 1) * kScanTile >= n) t.status[1] = (uint32_t)(exclusive + block_total);  // last tile: group count
    }
  }
  __syncthreads();
  uint64_t out = s_prefix + thread_off;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i)
    if (flags & (1u << i)) group_slots[out++] = (uint32_t)(base + i);
}

cudaError_t launch_compact(const TableDesc& t, uint32_t* group_slots, uint32_t* scan_state, cudaStream_t s)
{
  cudaError_t e = cudaMemsetAsync(scan_state, 0, compact_scan_state_bytes(t.capacity), s);
  if (e != cudaSuccess) return e;
  uint64_t tiles = (t.capacity + 1 + kScanTile - 1) / kScanTile;
  compact_kernel<<<(unsigned)tiles, kScanThreads, 0, s>>>(t, group_slots, reinterpret_cast<unsigned long long*>(scan_state));
  return note_launch();
}

// ------------------------------------------------------------------ output gather + finalise
__device__ __forceinline__ void store_width(void* out, int width, int64_t j, uint64_t bits)
{
  switch (width) {
    case 1: reinterpret_cast<uint8_t*>(out)[j] = (uint8_t)bits; break;
    case 2: reinterpret_cast<uint16_t*>(out)[j] = (uint16_t)bits; break;
    case 4: reinterpret_cast<uint32_t*>(out)[j] = (uint32_t)bits; break;
    default: reinterpret_cast<uint64_t*>(out)[j] = bits; break;
  }
}

// numeric value of an 8/4-byte sum cell as double (static_cast<double>(sum) in m2_var_std.cu:52-56 and the DIV binop)
__device__ __forceinline__ double cell_as_double(const void* cells, uint64_t slot, uint8_t acc_op)
{
  switch (acc_op) {
    case ACC_SUM_F64: case ACC_SUMSQ_F64: case ACC_PROD_F64: return reinterpret_cast<const double*>(cells)[slot];
    case ACC_SUM_F32: case ACC_SUMSQ_F32: case ACC_PROD_F32: return (double)reinterpret_cast<const float*>(cells)[slot];
    default: return (double)reinterpret_cast<const long long*>(cells)[slot];
  }
}

__global__ void __launch_bounds__(256) emit_kernel(const __grid_constant__ EmitParams p)
{
  const int64_t j   = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool active = j < p.num_groups;
  const int lane    = threadIdx.x & 31;
  uint64_t slot = 0, k0 = 0, k1 = 0;
  if (active) {
    // dense handles: one dependent load less per thread (the kernel is latency-bound: ~10 bytes per thread in flight)
    slot = p.dense ? ((p.last_slot != 0xFFFFFFFFu && j == p.num_groups - 1) ? (uint64_t)p.last_slot : (uint64_t)j) : (uint64_t)p.group_slots[j];
    if (p.key_words == 1) k0 = p.keys[slot];
    else { ulonglong2 k = reinterpret_cast<const ulonglong2*>(p.keys)[slot]; k0 = k.x; k1 = k.y; }
    // (the reserved slot's key array entry is EMPTY == the all-ones key it stands for)
  }
#pragma unroll 1
  for (int o = 0; o < p.num_outs; ++o) {
    const OutDesc& d = p.out[o];
    bool valid    = true;
    uint64_t bits = 0;
    if (active) {
      const int32_t n = d.cnt ? d.cnt[slot] : 1;
      switch (d.op) {
        case OUT_KEY: {
          if (d.null_bit >= 0) valid = !((d.null_bit < 64 ? (k0 >> d.null_bit) : (k1 >> (d.null_bit - 64))) & 1ull);
          uint64_t w = d.shift < 64 ? (k0 >> d.shift) : (k1 >> (d.shift - 64));
          if (d.key_bits) w = (d.key_bits < 64 ? (w & ((1ull << d.key_bits) - 1)) : w) + d.key_base;  // repacked: value = base + field
          bits = d.out_width == 8 ? w : (w & ((1ull << (8 * d.out_width)) - 1));
          break;
        }
        case OUT_KEY_GATHER: {  // k0 = number of the group's representative row
          const int64_t idx = (int64_t)d.key_base + (int64_t)k0;
          if (d.b != nullptr) valid = bit_is_set(reinterpret_cast<const uint32_t*>(d.b), idx);
          bits = valid ? canon_key_bits(load_raw(reinterpret_cast<const uint8_t*>(d.a), d.out_width, idx), d.acc_op) : 0;
          break;
        }
        case OUT_COPY_I64: bits = reinterpret_cast<const uint64_t*>(d.a)[slot]; valid = n > 0; break;
        case OUT_COPY_F64: bits = reinterpret_cast<const uint64_t*>(d.a)[slot]; valid = n > 0; break;
        case OUT_COPY_F32: bits = reinterpret_cast<const uint32_t*>(d.a)[slot]; valid = n > 0; break;
        case OUT_MINMAX: bits = ordered_decode(reinterpret_cast<const uint64_t*>(d.a)[slot], d.src); valid = n > 0; break;
        case OUT_COUNT: bits = (uint32_t)reinterpret_cast<const int32_t*>(d.a)[slot]; break;
        case OUT_SUMOV_SUM: case OUT_SUMOV_FLAG: {
          const uint64_t lo = reinterpret_cast<const uint64_t*>(d.a)[slot];
          const int64_t hi  = (int64_t)reinterpret_cast<const uint64_t*>(d.b)[slot];
          valid = n > 0;
          if (d.op == OUT_SUMOV_SUM) { bits = ((uint64_t)hi << 32) + lo; break; }
          const __int128 total = ((__int128)hi << 32) + (__int128)lo;
          const int vbits      = 8 * (int)d.shift;
          const __int128 vmin = -((__int128)1 << (vbits - 1)), vmax = ((__int128)1 << (vbits - 1)) - 1;
          bits = (total < vmin || total > vmax) ? 1 : 0;
          break;
        }
        case OUT_ARG: {
          const bool rowcell = d.acc_op == ACC_ARGROW_MIN || d.acc_op == ACC_ARGROW_MAX;  // 64-bit values: the cell IS the row
          uint32_t lo = rowcell ? reinterpret_cast<const uint32_t*>(d.a)[slot] : (uint32_t)reinterpret_cast<const uint64_t*>(d.a)[slot];
          valid       = n > 0;
          if (rowcell) { bits = valid ? lo : 0xFFFFFFFFu; break; }
          bits        = !valid ? 0xFFFFFFFFu : (d.acc_op == ACC_ARGMAX ? 0xFFFFFFFFu - lo : lo);
          break;
        }
        case OUT_MEAN: {
          valid    = n > 0;
          double r = valid ? cell_as_double(d.a, slot, d.acc_op) / (double)n : 0.0;
          bits     = (uint64_t)__double_as_longlong(r);
          break;
        }
        case OUT_MEAN_INT: {
          valid = n > 0;
          bits  = valid ? (uint64_t)(reinterpret_cast<const long long*>(d.a)[slot] / (long long)n) : 0;
          break;
        }
        case OUT_M2: case OUT_VAR: case OUT_STD: {
          // m2 = sumsq - sum*sum/n evaluated in double, 0 when n == 0 (m2_var_std.cu:49-59)
          double m2 = 0.0;
          if (n != 0) {
            double ss = cell_as_double(d.b, slot, d.acc_op == ACC_SUM_F64 ? ACC_SUMSQ_F64 : d.acc_op == ACC_SUM_F32 ? ACC_SUMSQ_F32 : ACC_SUMSQ_I64);
            double sm = cell_as_double(d.a, slot, d.acc_op);
            m2        = ss - sm * sm / n;
          }
          double r = m2;
          if (d.op != OUT_M2) {
            const int32_t df = n - d.ddof;  // null iff n == 0 or n - ddof <= 0 (m2_var_std.cu:163-166,182-185)
            valid            = !(n == 0 || df <= 0);
            r                = valid ? m2 / df : 0.0;
            if (valid && d.op == OUT_STD) r = sqrt(r);
          }
          bits = (uint64_t)__double_as_longlong(r);
          break;
        }
        default: break;
      }
      store_width(d.out, d.out_width, j, bits);
    }
    if (d.out_mask != nullptr) {  // warp = 32 consecutive output rows = one bitmask word
      const uint32_t word = __ballot_sync(0xffffffffu, active && valid);
      const uint32_t act  = __ballot_sync(0xffffffffu, active);
      if (lane == 0 && act) {
        d.out_mask[j >> 5] = word;
        if (d.null_count) { int nulls = __popc(act & ~word); if (nulls) atomicAdd(d.null_count, nulls); }
      }
    }
  }
}

cudaError_t launch_emit(const EmitParams& p, cudaStream_t s)
{
  if (p.num_groups == 0 || p.num_outs == 0) return cudaSuccess;
  int blocks = (p.num_groups + 255) / 256;
  emit_kernel<<<blocks, 256, 0, s>>>(p);
  return note_launch();
}

}  // namespace bgb

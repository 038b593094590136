// Generic (descriptor-driven) kernels: table init, fused hash+probe+aggregate into the HBM group table,
// cardinality estimator. They serve every supported type/aggregation combination; the specialised
// shared-memory and partitioned kernels (kernels_smem.cu, kernels_partition.cu) are fast paths in front of them.
//
// Replaces, as ONE pass over the rows with no N-sized temporaries:
//   cuco::static_set insert_and_find + compute_aggs_* (/root/reference/cpp/src/groupby/hash/compute_global_memory_aggs.cuh:41-157,
//   /root/reference/cpp/src/groupby/hash/single_pass_functors.cuh:86-157) and initialize_with_identity
//   (/root/reference/cpp/src/aggregation/aggregation.cu:39-97).
#include <cstdlib>

#include "device_utils.cuh"

namespace bgb {

__global__ void __launch_bounds__(256) init_table_kernel(const __grid_constant__ AggParams p)
{
  const uint64_t n      = p.table.capacity + 1;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    if (p.init_keys) {
      if (p.table.key_words == 1) p.table.keys[i] = kEmpty;
      else { reinterpret_cast<ulonglong2*>(p.table.keys)[i] = make_ulonglong2(kEmpty, kEmpty); }
    }
    for (int a = 0; a < p.num_accs; ++a) {
      const uint8_t op = p.acc[a].op;
      if (acc_cell_bytes(op) == 8) reinterpret_cast<uint64_t*>(p.acc[a].cells)[i] = acc_identity64(op);
      else reinterpret_cast<uint32_t*>(p.acc[a].cells)[i] = acc_identity32(op);
    }
  }
  if (p.init_keys && blockIdx.x == 0 && threadIdx.x < 4) p.table.status[threadIdx.x] = 0;
}

cudaError_t launch_init_table(const AggParams& p, cudaStream_t s)
{
  uint64_t n = p.table.capacity + 1;
  int blocks = (int)((n + 255) / 256 < (uint64_t)device_sm_count() * 16 ? (n + 255) / 256 : (uint64_t)device_sm_count() * 16);
  init_table_kernel<<<blocks, 256, 0, s>>>(p);
  return note_launch();
}

// ------------------------------------------------------------------ one row into one accumulator
__device__ __forceinline__ void apply_acc(const AccDesc& a, int64_t row, uint64_t slot)
{
  const int64_t idx = (int64_t)a.offset + row;
  if (a.op == ACC_COUNT_ALL) { atomicAdd(reinterpret_cast<int*>(a.cells) + slot, 1); return; }
  if (a.mask != nullptr) {  // null source skipped (device_aggregators.cuh:434-436)
    const bool valid = a.byte_mask ? __ldg(reinterpret_cast<const uint8_t*>(a.mask) + idx) != 0 : bit_is_set(a.mask, idx);
    if (!valid) return;
  }
  if (a.op == ACC_COUNT_VALID) { atomicAdd(reinterpret_cast<int*>(a.cells) + slot, 1); return; }
  uint64_t raw;
  if (a.dict != nullptr) raw = load_raw(a.dict, src_width(a.src), (int64_t)load_raw(a.data, a.idx_width, idx));  // keys[indices[row]]
  else raw = load_raw(a.data, src_width(a.src), idx);
  apply_value(a.op, a.src, a.cells, slot, raw, row);
}

// ------------------------------------------------------------------ fused hash + probe + aggregate (HBM table)
// One row per thread per iteration, on purpose: the kernel is bound by the L2 atomic units (193 G op/s measured,
// tools/microbench4.cu; a row is one key probe + one RED per accumulator), not by latency -- keeping 2 or 4 rows in flight
// per thread (key loads and first probes issued together) made it SLOWER: 2.06 / 2.37 / 3.67 ms for 1 / 2 / 4 rows at
// 1e8 rows x 5e4 groups.
__global__ void __launch_bounds__(256) aggregate_global_kernel(const __grid_constant__ AggParams p)
{
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t nrows  = p.num_rows_dev ? (int64_t)*p.num_rows_dev : p.num_rows;
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < nrows; row += stride) {
    uint64_t k0, k1, slot;
    if (p.wide) {  // (warp-uniform) keys of more than 128 bits: slots hold row numbers
      if (!row_hash(p, row, k0)) continue;
      slot = find_or_insert_row(p, row, k0);
    } else {
      if (!pack_key(p, row, k0, k1)) continue;
      slot = find_or_insert(p.table, k0, k1);
    }
    if (slot == kNoSlot) return;
#pragma unroll 1
    for (int a = 0; a < p.num_accs; ++a) apply_acc(p.acc[a], row, slot);
  }
}

cudaError_t launch_aggregate_global(const AggParams& p, cudaStream_t s)
{
  if (p.num_rows == 0) return cudaSuccess;
  int64_t want = (p.num_rows + 255) / 256;
  int blocks   = (int)(want < device_sm_count() * 8 ? want : device_sm_count() * 8);
  aggregate_global_kernel<<<blocks, 256, 0, s>>>(p);
  return note_launch();
}

// ------------------------------------------------------------------ ARGMIN / ARGMAX of 64-bit values, second pass
// The reference keeps a row index per group and CAS-loops on it while comparing the values it points at
// (/root/reference/cpp/include/cudf/detail/aggregation/device_aggregators.cuh:131-160). Here the group's MIN / MAX cell is
// already final when this kernel runs, so a row only has to compare its own value with it: equal -> atomicMin of the row
// number (ties resolve to the smallest row, like the packed cells of the <= 32-bit case).
__global__ void __launch_bounds__(256) argrow_kernel(const __grid_constant__ AggParams p)
{
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < p.num_rows; row += stride) {
    uint64_t k0, k1, slot;
    if (p.wide) {
      if (!row_hash(p, row, k0)) continue;
      slot = find_or_insert_row(p, row, k0);
    } else {
      if (!pack_key(p, row, k0, k1)) continue;
      slot = find_or_insert(p.table, k0, k1);  // the key is there: pure lookup
    }
    if (slot == kNoSlot) return;
#pragma unroll 1
    for (int a = 0; a < p.num_accs; ++a) {
      const AccDesc& d = p.acc[a];
      if (d.op != ACC_ARGROW_MIN && d.op != ACC_ARGROW_MAX) continue;
      const int64_t idx = (int64_t)d.offset + row;
      if (d.mask != nullptr && !bit_is_set(d.mask, idx)) continue;
      const uint64_t raw = d.dict != nullptr ? load_raw(d.dict, src_width(d.src), (int64_t)load_raw(d.data, d.idx_width, idx))
                                             : load_raw(d.data, src_width(d.src), idx);
      const uint64_t img = ordered_image(raw, d.src);
      if (img == reinterpret_cast<const uint64_t*>(d.aux)[slot]) atomicMin(reinterpret_cast<unsigned int*>(d.cells) + slot, (unsigned int)row);
    }
  }
}

cudaError_t launch_argrow(const AggParams& p, cudaStream_t s)
{
  if (p.num_rows == 0) return cudaSuccess;
  int64_t want = (p.num_rows + 255) / 256;
  argrow_kernel<<<(int)(want < device_sm_count() * 8 ? want : device_sm_count() * 8), 256, 0, s>>>(p);
  return note_launch();
}

// ------------------------------------------------------------------ cardinality estimator
// Inserts the packed keys of `sample_rows` evenly spaced rows into a scratch set; out[0] = distinct, out[1] = sampled
// rows that were not skipped. Only a regime hint: correctness never depends on it (overflow -> host retries bigger).
__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { unsigned long long w = __shfl_xor_sync(0xffffffffu, v, o); v = w > v ? w : v; }
  return v;
}

__global__ void __launch_bounds__(256) estimate_kernel(const __grid_constant__ AggParams p, int64_t sample_rows, int64_t step,
                                                       uint64_t* set, uint64_t set_capacity, uint32_t* out, uint64_t key_flip)
{
  const int64_t t   = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t row = t * step;
  uint64_t k0 = 0, k1 = 0;
  // (wide keys: the 64-bit row hash stands for the key -- the count of distinct hashes is the estimate)
  const bool live = t < sample_rows && row < p.num_rows && (p.wide ? row_hash(p, row, k0) : pack_key(p, row, k0, k1));
  // per-warp reductions first: one atomic per warp instead of one per sampled row
  unsigned long long kimg = live && p.table.key_words == 1 ? (unsigned long long)(k0 ^ key_flip) : 0ull;
  unsigned long long mag  = 0;
  if (live && p.num_accs == 1 && p.acc[0].data != nullptr) {  // magnitude of the sampled integer SUM inputs (sizes the smem count field)
    const AccDesc& a  = p.acc[0];
    const int64_t idx = (int64_t)a.offset + row;
    if (a.mask == nullptr || bit_is_set(a.mask, idx)) {
      const int64_t v = raw_to_i64(load_raw(a.data, src_width(a.src), idx), a.src);
      mag             = v < 0 ? ~(unsigned long long)v : (unsigned long long)v;
    }
  }
  const unsigned long long wmax = warp_max_u64(kimg);
  const unsigned long long wmin = ~warp_max_u64(live && p.table.key_words == 1 ? ~kimg : 0ull);
  const unsigned long long wmag = warp_max_u64(mag);
  const uint32_t nlive          = __popc(__ballot_sync(0xffffffffu, live));
  if ((threadIdx.x & 31) == 0 && nlive) {
    atomicAdd(out + 1, nlive);
    if (p.table.key_words == 1) {  // range of the order-preserving key image (direct-address eligibility)
      atomicMin(reinterpret_cast<unsigned long long*>(out + 4), wmin);
      atomicMax(reinterpret_cast<unsigned long long*>(out + 6), wmax);
    }
    if (wmag) atomicMax(reinterpret_cast<unsigned long long*>(out + 2), wmag);
  }
  if (!live) return;
  uint64_t f = hash_key(k0, k1, p.table.key_words);
  if (f == kEmpty) f = 0;
  uint64_t slot = mix64(f) & (set_capacity - 1);
  unsigned long long* s = reinterpret_cast<unsigned long long*>(set);
  uint32_t fresh = 0;
  for (uint64_t probes = 0; probes < set_capacity; ++probes) {
    unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(s + slot);
    if (cur == f) break;
    if (cur == kEmpty) {
      unsigned long long old = atomicCAS(s + slot, (unsigned long long)kEmpty, (unsigned long long)f);
      if (old == kEmpty) { fresh = 1; break; }
      if (old == f) break;
    }
    slot = (slot + 1) & (set_capacity - 1);
  }
  if (fresh) atomicAdd(out, 1u);  // (ptxas aggregates this per warp)
}

cudaError_t launch_estimate_cardinality(const AggParams& p, int64_t sample_rows, uint64_t* scratch_set, uint64_t set_capacity,
                                        uint32_t* out_distinct, uint64_t key_flip, cudaStream_t s)
{
  cudaError_t e = cudaMemsetAsync(scratch_set, 0xff, set_capacity * sizeof(uint64_t), s);
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(out_distinct, 0, 8 * sizeof(uint32_t), s);
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(out_distinct + 4, 0xff, 2 * sizeof(uint32_t), s);  // running minimum starts at all ones
  if (e != cudaSuccess) return e;
  if (sample_rows > p.num_rows) sample_rows = p.num_rows;
  if (sample_rows <= 0) return cudaSuccess;
  int64_t step = p.num_rows / sample_rows;
  if (step < 1) step = 1;
  int blocks = (int)((sample_rows + 255) / 256);
  estimate_kernel<<<blocks, 256, 0, s>>>(p, sample_rows, step, scratch_set, set_capacity, out_distinct, key_flip);
  return note_launch();
}

}  // namespace bgb

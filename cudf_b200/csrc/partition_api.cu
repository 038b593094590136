// Multi-GPU building blocks behind the C ABI: permute a set of columns into per-owner-rank runs (hash2(key) % P), and
// turn the per-row validity bytes that travel through the exchange back into cudf bitmasks.
// Reference counterpart of the row partitioner: cudf::hash_partition
// (/root/reference/cpp/src/partitioning/partitioning.cu:119-151,569-663; API cpp/include/cudf/partitioning.hpp:103-145).
#include <cstring>
#include <vector>

#include "errors.h"
#include "device_utils.cuh"
#include "kernels_partition.h"
#include "plan.h"
#include "nvtx.h"
#include "pool.h"

using namespace bgb;

namespace bgb {
__global__ void __launch_bounds__(256) bytes_to_bitmask_kernel(const uint8_t* __restrict__ bytes, int64_t n, uint32_t* __restrict__ words,
                                                               int* null_count)
{
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t padded = (n + 31) & ~31ll;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < padded; i += stride) {
    const bool in    = i < n;
    const bool valid = in && bytes[i] != 0;
    const uint32_t w = __ballot_sync(0xffffffffu, valid);
    const uint32_t a = __ballot_sync(0xffffffffu, in);
    if ((threadIdx.x & 31) == 0) {
      words[i >> 5] = w;
      const int z = __popc(a & ~w);
      if (z && null_count) atomicAdd(null_count, z);
    }
  }
}
// MEAN / M2 / VARIANCE / STD from merged SUM, SUM_OF_SQUARES and COUNT columns: the same arithmetic as emit_kernel's
// finalisers, applied to columns instead of table cells (hash_compound_agg_finalizer.cu:94-133, m2_var_std.cu:49-59,138-187)
struct MomentParams {
  const void* sum; const void* sumsq; const int32_t* count;
  int32_t sum_type, kind, ddof, out_width;
  int64_t n;
  void* out; uint32_t* out_mask; int* null_count;
};

__device__ __forceinline__ double moment_as_double(const void* p, int64_t i, int32_t t)
{
  if (t == BGB_FLOAT64) return reinterpret_cast<const double*>(p)[i];
  if (t == BGB_FLOAT32) return (double)reinterpret_cast<const float*>(p)[i];
  return (double)reinterpret_cast<const long long*>(p)[i];
}

__global__ void __launch_bounds__(256) finalize_moments_kernel(const __grid_constant__ MomentParams p)
{
  const int64_t i   = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool active = i < p.n;
  bool valid        = false;
  if (active) {
    const int32_t n = p.count[i];
    if (p.kind == BGB_MEAN) {
      valid = n > 0;
      if (p.sum_type >= BGB_DURATION_DAYS && p.sum_type <= BGB_DURATION_NANOSECONDS) {
        // durations: integer mean in the column's own type
        reinterpret_cast<long long*>(p.out)[i] = valid ? reinterpret_cast<const long long*>(p.sum)[i] / (long long)n : 0;
      } else {
        reinterpret_cast<double*>(p.out)[i] = valid ? moment_as_double(p.sum, i, p.sum_type) / (double)n : 0.0;
      }
    } else {
      double m2 = 0.0;
      if (n != 0) {
        const double ss = moment_as_double(p.sumsq, i, p.sum_type);
        const double sm = moment_as_double(p.sum, i, p.sum_type);
        m2              = ss - sm * sm / n;
      }
      double r = m2;
      valid    = true;
      if (p.kind != BGB_M2) {
        const int32_t df = n - p.ddof;
        valid            = !(n == 0 || df <= 0);
        r                = valid ? m2 / df : 0.0;
        if (valid && p.kind == BGB_STD) r = sqrt(r);
      }
      reinterpret_cast<double*>(p.out)[i] = r;
    }
  }
  if (p.out_mask != nullptr) {
    const uint32_t word = __ballot_sync(0xffffffffu, active && valid);
    const uint32_t act  = __ballot_sync(0xffffffffu, active);
    if ((threadIdx.x & 31) == 0 && act) {
      p.out_mask[i >> 5] = word;
      const int z        = __popc(act & ~word);
      if (z && p.null_count) atomicAdd(p.null_count, z);
    }
  }
}
}  // namespace bgb

extern "C" int bgb_finalize_moments(int32_t kind, int32_t ddof, const bgb_column_view* sum, const bgb_column_view* sumsq,
                                    const bgb_column_view* count, bgb_mutable_column* out, void* stream)
{
  if (kind != BGB_MEAN && kind != BGB_M2 && kind != BGB_VARIANCE && kind != BGB_STD)
    return set_error(BGB_INVALID_ARGUMENT, "bgb_finalize_moments: kind %d is not MEAN/M2/VARIANCE/STD", kind);
  if (!sum || !count || !out || (kind != BGB_MEAN && !sumsq)) return set_error(BGB_INVALID_ARGUMENT, "bgb_finalize_moments: null argument");
  if (count->type_id != BGB_INT32) return set_error(BGB_INVALID_ARGUMENT, "bgb_finalize_moments: count must be INT32");
  const int64_t n = sum->size;
  if (count->size != n || (sumsq && kind != BGB_MEAN && sumsq->size != n) || out->size < n)
    return set_error(BGB_INVALID_ARGUMENT, "Size mismatch between the merged columns.");
  if (sum->offset || count->offset || (sumsq && sumsq->offset)) return set_error(BGB_UNSUPPORTED, "bgb_finalize_moments: sliced inputs");
  if (type_width(sum->type_id) != 8 && sum->type_id != BGB_FLOAT32) return set_error(BGB_UNSUPPORTED, "bgb_finalize_moments: sum type %d", sum->type_id);
  if (kind != BGB_MEAN && sumsq->type_id != sum->type_id) return set_error(BGB_INVALID_ARGUMENT, "bgb_finalize_moments: sum / sumsq types differ");
  out->null_count = 0;
  if (n == 0) return BGB_OK;
  cudaStream_t s = (cudaStream_t)stream;
  int* d_cnt     = nullptr;
  if (out->null_mask) {
    d_cnt = (int*)pool_alloc_async(sizeof(int), s);
    if (!d_cnt) return set_error(BGB_OUT_OF_MEMORY, "null counter");
    cudaMemsetAsync(d_cnt, 0, sizeof(int), s);
  }
  MomentParams mp{sum->data, sumsq ? sumsq->data : nullptr, (const int32_t*)count->data, sum->type_id, kind, ddof, 8, n,
                  out->data, (uint32_t*)out->null_mask, d_cnt};
  finalize_moments_kernel<<<(int)((n + 255) / 256), 256, 0, s>>>(mp);
  int rc = check_cuda(note_launch(), "finalize_moments");
  if (d_cnt) {
    int h = 0;
    if (!rc) rc = check_cuda(cudaMemcpyAsync(&h, d_cnt, sizeof(int), cudaMemcpyDeviceToHost, s), "null count copy");
    pool_free_async(d_cnt, s);
    if (!rc) rc = check_cuda(cudaStreamSynchronize(s), "finalize_moments sync");
    out->null_count = h;
  }
  return rc;
}

extern "C" int bgb_bytes_to_bitmask(const uint8_t* valid_bytes, int64_t n, uint32_t* mask_words, int32_t* null_count_host, void* stream)
{
  cudaStream_t s = (cudaStream_t)stream;
  if (null_count_host) *null_count_host = 0;
  if (n <= 0) return BGB_OK;
  if (!valid_bytes || !mask_words) return set_error(BGB_INVALID_ARGUMENT, "bgb_bytes_to_bitmask: null pointer");
  int* d_cnt = (int*)pool_alloc_async(sizeof(int), s);
  if (!d_cnt) return set_error(BGB_OUT_OF_MEMORY, "null counter");
  cudaMemsetAsync(d_cnt, 0, sizeof(int), s);
  const int64_t want = (n + 255) / 256;
  bytes_to_bitmask_kernel<<<(int)(want < device_sm_count() * 8 ? want : device_sm_count() * 8), 256, 0, s>>>(valid_bytes, n, mask_words, d_cnt);
  int rc = check_cuda(note_launch(), "bytes_to_bitmask");
  int h = 0;
  if (!rc) rc = check_cuda(cudaMemcpyAsync(&h, d_cnt, sizeof(int), cudaMemcpyDeviceToHost, s), "null count copy");
  pool_free_async(d_cnt, s);
  if (!rc) rc = check_cuda(cudaStreamSynchronize(s), "bytes_to_bitmask sync");
  if (null_count_host) *null_count_host = h;
  return rc;
}

extern "C" int bgb_partition_by_key_hash(const bgb_column_view* keys, int32_t num_keys, const bgb_column_view* cols, int32_t num_cols,
                                         int32_t num_parts, bgb_mutable_column* out_cols, int32_t* part_offsets_host, void* stream)
{
  bgb::NvtxRange nvtx_call("bgb_partition_by_key_hash");
  if (num_parts < 1 || num_parts > 4096) return set_error(BGB_INVALID_ARGUMENT, "num_parts must be in [1, 4096]");
  if (num_cols < 0 || num_cols > kMaxPartValues || (num_cols > 0 && (!cols || !out_cols)))
    return set_error(BGB_INVALID_ARGUMENT, "between 0 and %d columns can be permuted in one call", kMaxPartValues);
  if (!part_offsets_host) return set_error(BGB_INVALID_ARGUMENT, "null part_offsets_host");
  Plan pl;
  int rc = make_plan(keys, num_keys, /*include_null_keys=*/1, nullptr, 0, pl);  // null keys are kept: they hash like any other image
  if (rc) return rc;
  const int64_t N = pl.num_rows;
  for (int c = 0; c < num_cols; ++c) {
    if (cols[c].size != N) return set_error(BGB_INVALID_ARGUMENT, "column %d has %d rows, keys have %lld", c, cols[c].size, (long long)N);
    if (type_width(cols[c].type_id) == 0) return set_error(BGB_UNSUPPORTED, "column %d: not a fixed-width type", c);
    if (N > 0 && (!out_cols[c].data || out_cols[c].size < N)) return set_error(BGB_INVALID_ARGUMENT, "out_cols[%d] too small", c);
    if (N > 0 && cols[c].null_mask && !out_cols[c].null_mask) return set_error(BGB_INVALID_ARGUMENT, "out_cols[%d] needs validity bytes", c);
  }
  for (int i = 0; i <= num_parts; ++i) part_offsets_host[i] = 0;
  if (N == 0) return BGB_OK;
  cudaStream_t s = (cudaStream_t)stream;
  int num_sms = 148;
  { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev); }

  PartParams pp;
  memset(&pp, 0, sizeof pp);
  pp.a.num_rows = N; pp.a.num_key_cols = (int)pl.keys.size(); pp.a.skip_null_keys = 0;
  for (size_t c = 0; c < pl.keys.size(); ++c) {
    const KeyPlan& k = pl.keys[c];
    pp.a.kc[c] = KeyColDesc{(const uint8_t*)k.v.data, k.v.null_mask, k.v.offset, k.width, k.fclass, k.shift, k.null_bit};
  }
  pp.a.table.key_words = pl.key_words;
  pp.a.table.capacity  = 1;
  pp.mod_parts = num_parts;
  int log2p = 0; while ((1 << log2p) < num_parts) ++log2p;
  pp.log2_parts = log2p;
  const uint32_t P = 1u << log2p;
  uint32_t* counters = (uint32_t*)pool_alloc_async(sizeof(uint32_t) * (3 * (size_t)P + 1), s);
  if (!counters) return set_error(BGB_OUT_OF_MEMORY, "partition counters");
  pp.counts = counters; pp.cursors = counters + P; pp.offsets = counters + 2 * (size_t)P;
  pp.key_out_width = 8;
  pp.num_vals = num_cols;
  for (int c = 0; c < num_cols; ++c) {
    PartValue& o = pp.v[c];
    o.data = (const uint8_t*)cols[c].data; o.mask = cols[c].null_mask; o.offset = cols[c].offset; o.width = type_width(cols[c].type_id);
    o.out_data = (uint8_t*)out_cols[c].data; o.out_valid = cols[c].null_mask ? (uint8_t*)out_cols[c].null_mask : nullptr;
  }
  rc = check_cuda(cudaMemsetAsync(counters, 0, sizeof(uint32_t) * P, s), "partition counters");
  if (!rc) rc = check_cuda(launch_partition_hist(pp, num_sms, s), "partition_hist");
  if (!rc) rc = check_cuda(launch_partition_scan(pp, s), "partition_scan");
  if (!rc) rc = check_cuda(launch_partition_scatter(pp, num_sms, s), "partition_scatter");
  std::vector<uint32_t> off(P + 1, 0);
  if (!rc) rc = check_cuda(cudaMemcpyAsync(off.data(), pp.offsets, sizeof(uint32_t) * (P + 1), cudaMemcpyDeviceToHost, s), "offsets copy");
  pool_free_async(counters, s);
  if (!rc) rc = check_cuda(cudaStreamSynchronize(s), "partition sync");
  if (rc) return rc;
  for (int i = 0; i <= num_parts; ++i) part_offsets_host[i] = (int32_t)off[i < (int)P ? i : P];
  part_offsets_host[num_parts] = (int32_t)off[P];
  return BGB_OK;
}

// ---- cudf::hash_partition drop-in ---------------------------------------------------------------------------------------
namespace bgb {
struct GatherParams {
  const uint32_t* map; int64_t n; int32_t num_cols;
  PartValue v[kMaxPartValues];
};
// out[i] = in[map[i]] for up to kMaxPartValues columns (tables wider than one scatter launch can carry)
__global__ void __launch_bounds__(256) gather_rows_kernel(const __grid_constant__ GatherParams g)
{
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < g.n; i += stride) {
    const int64_t src = g.map[i];
    for (int c = 0; c < g.num_cols; ++c) {
      const PartValue& v = g.v[c];
      const int64_t idx  = (int64_t)v.offset + src;
      const bool valid   = v.mask == nullptr || bit_is_set(v.mask, idx);
      const uint64_t x   = valid ? load_raw(v.data, v.width, idx) : 0ull;
      switch (v.width) {
        case 1: v.out_data[i] = (uint8_t)x; break;
        case 2: reinterpret_cast<uint16_t*>(v.out_data)[i] = (uint16_t)x; break;
        case 4: reinterpret_cast<uint32_t*>(v.out_data)[i] = (uint32_t)x; break;
        default: reinterpret_cast<uint64_t*>(v.out_data)[i] = x; break;
      }
      if (v.out_valid) v.out_valid[i] = valid;
    }
  }
}
static bool type_is_signed_int(int32_t t)
{
  return t == BGB_INT8 || t == BGB_INT16 || t == BGB_INT32 || t == BGB_INT64;
}
static bool type_is_numeric(int32_t t)  // cudf::is_numeric: integers, floats, bool
{
  return (t >= BGB_INT8 && t <= BGB_FLOAT64) || t == BGB_BOOL8;
}
}  // namespace bgb

extern "C" int bgb_hash_partition(const bgb_column_view* input, int32_t num_cols, const int32_t* columns_to_hash, int32_t num_hash_cols,
                                  int32_t num_partitions, int32_t hash_function, uint32_t seed, bgb_mutable_column* out_cols,
                                  int32_t* offsets_host, void* stream)
{
  bgb::NvtxRange nvtx_call("bgb_hash_partition");
  if (hash_function != BGB_HASH_IDENTITY && hash_function != BGB_HASH_MURMUR3)
    return set_error(BGB_INVALID_ARGUMENT, "Unsupported hash function in hash_partition");
  if (num_cols < 0 || num_hash_cols < 0 || (num_cols > 0 && !input) || (num_hash_cols > 0 && !columns_to_hash))
    return set_error(BGB_INVALID_ARGUMENT, "bgb_hash_partition: null argument");
  if (num_partitions > 0 && !offsets_host) return set_error(BGB_INVALID_ARGUMENT, "bgb_hash_partition: null offsets_host");
  for (int i = 0; i <= (num_partitions > 0 ? num_partitions : -1); ++i) offsets_host[i] = 0;
  for (int c = 0; c < num_cols; ++c)
    if (out_cols) out_cols[c].null_count = 0;
  const int64_t N = num_cols > 0 ? input[0].size : 0;
  for (int c = 1; c < num_cols; ++c)
    if (input[c].size != N) return set_error(BGB_INVALID_ARGUMENT, "Column size mismatch: column %d has %d rows, column 0 has %lld", c, input[c].size, (long long)N);
  for (int h = 0; h < num_hash_cols; ++h)
    if (columns_to_hash[h] < 0 || columns_to_hash[h] >= num_cols) return set_error(BGB_INVALID_ARGUMENT, "columns_to_hash[%d] = %d is out of range", h, columns_to_hash[h]);
  // empty result: no partitions, no rows, or nothing to hash (partitioning.cu:885-888)
  if (num_partitions <= 0 || N == 0 || num_hash_cols == 0) return BGB_OK;
  if (num_partitions > 4096) return set_error(BGB_UNSUPPORTED, "bgb_hash_partition: more than 4096 partitions");
  if (num_hash_cols > kMaxKeyCols) return set_error(BGB_UNSUPPORTED, "bgb_hash_partition: more than %d columns to hash", kMaxKeyCols);
  if (!out_cols) return set_error(BGB_INVALID_ARGUMENT, "bgb_hash_partition: null out_cols");

  PartParams pp;
  memset(&pp, 0, sizeof pp);
  pp.a.num_rows = N; pp.a.num_key_cols = num_hash_cols;
  for (int h = 0; h < num_hash_cols; ++h) {
    const bgb_column_view& k = input[columns_to_hash[h]];
    const int w = type_width(k.type_id);
    if (w == 0) return set_error(BGB_UNSUPPORTED, "bgb_hash_partition: column %d (type %d) is not a fixed-width type of at most 8 bytes", columns_to_hash[h], k.type_id);
    if (hash_function == BGB_HASH_IDENTITY && !type_is_numeric(k.type_id)) return set_error(BGB_INVALID_ARGUMENT, "IdentityHash does not support this data type");
    const uint8_t fclass = k.type_id == BGB_FLOAT32 ? 1 : k.type_id == BGB_FLOAT64 ? 2 : k.type_id == BGB_BOOL8 ? 3 : 0;
    // a mask without nulls hashes like no mask; cudf checks nulls only when the hashed table has any (partitioning.cu:895-901)
    pp.a.kc[h] = KeyColDesc{(const uint8_t*)k.data, k.null_count != 0 ? k.null_mask : nullptr, k.offset, (uint8_t)w, fclass, 0, 0};
    if (type_is_signed_int(k.type_id)) pp.signed_cols |= 1u << h;
  }
  for (int c = 0; c < num_cols; ++c) {
    if (type_width(input[c].type_id) == 0) return set_error(BGB_UNSUPPORTED, "bgb_hash_partition: column %d (type %d) is not a fixed-width type of at most 8 bytes", c, input[c].type_id);
    if (!out_cols[c].data || out_cols[c].size < N) return set_error(BGB_INVALID_ARGUMENT, "out_cols[%d] holds fewer rows than the input", c);
    if (input[c].null_mask && input[c].null_count != 0 && !out_cols[c].null_mask) return set_error(BGB_INVALID_ARGUMENT, "out_cols[%d] needs a null mask", c);
  }
  cudaStream_t s = (cudaStream_t)stream;
  int num_sms = 148;
  { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev); }
  pp.a.table.key_words = 1; pp.a.table.capacity = 1;
  pp.mod_parts = num_partitions;
  pp.cudf_hash = hash_function == BGB_HASH_MURMUR3 ? 1 : 2;
  pp.hash_seed = seed;
  int log2p = 0; while ((1 << log2p) < num_partitions) ++log2p;
  pp.log2_parts = log2p;
  pp.key_out_width = 8;
  const uint32_t P = 1u << log2p;

  // scratch: counters | gather map (wide tables only) | one validity byte per row for every column that carries nulls
  std::vector<int> nullable;
  for (int c = 0; c < num_cols; ++c) if (input[c].null_mask && input[c].null_count != 0) nullable.push_back(c);
  const bool wide          = num_cols > kMaxPartValues;
  const size_t cnt_bytes   = ((sizeof(uint32_t) * (3 * (size_t)P + 1 + nullable.size()) + 255) / 256) * 256;
  const size_t map_bytes   = wide ? (((size_t)N * 4 + 255) / 256) * 256 : 0;
  const size_t bytes_each  = (((size_t)N + 255) / 256) * 256;
  uint8_t* scratch = (uint8_t*)pool_alloc_async(cnt_bytes + map_bytes + bytes_each * nullable.size(), s);
  if (!scratch) return set_error(BGB_OUT_OF_MEMORY, "hash_partition scratch");
  uint32_t* counters = (uint32_t*)scratch;
  pp.counts = counters; pp.cursors = counters + P; pp.offsets = counters + 2 * (size_t)P;
  int* d_nulls = (int*)(counters + 3 * (size_t)P + 1);
  uint32_t* map = wide ? (uint32_t*)(scratch + cnt_bytes) : nullptr;
  uint8_t* vbytes = scratch + cnt_bytes + map_bytes;
  std::vector<uint8_t*> col_bytes(num_cols, nullptr);
  for (size_t i = 0; i < nullable.size(); ++i) col_bytes[nullable[i]] = vbytes + i * bytes_each;
  auto describe = [&](int c, PartValue& o) {
    o.data = (const uint8_t*)input[c].data; o.mask = col_bytes[c] ? input[c].null_mask : nullptr; o.offset = input[c].offset;
    o.width = type_width(input[c].type_id); o.out_data = (uint8_t*)out_cols[c].data; o.out_valid = col_bytes[c];
  };
  int rc = check_cuda(cudaMemsetAsync(counters, 0, cnt_bytes, s), "partition counters");
  if (!rc) rc = check_cuda(launch_partition_hist(pp, num_sms, s), "partition_hist");
  if (!rc) rc = check_cuda(launch_partition_scan(pp, s), "partition_scan");
  if (!wide) {
    pp.num_vals = num_cols;
    for (int c = 0; c < num_cols; ++c) describe(c, pp.v[c]);
    if (!rc) rc = check_cuda(launch_partition_scatter(pp, num_sms, s), "partition_scatter");
  } else {
    pp.num_vals = 0; pp.out_row = map;
    if (!rc) rc = check_cuda(launch_partition_scatter(pp, num_sms, s), "partition_scatter");
    for (int c0 = 0; c0 < num_cols && !rc; c0 += kMaxPartValues) {
      GatherParams g;
      memset(&g, 0, sizeof g);
      g.map = map; g.n = N; g.num_cols = num_cols - c0 < kMaxPartValues ? num_cols - c0 : kMaxPartValues;
      for (int c = 0; c < g.num_cols; ++c) describe(c0 + c, g.v[c]);
      const int64_t want = (N + 255) / 256;
      gather_rows_kernel<<<(int)(want < (int64_t)num_sms * 16 ? want : (int64_t)num_sms * 16), 256, 0, s>>>(g);
      rc = check_cuda(note_launch(), "gather_rows");
    }
  }
  for (size_t i = 0; i < nullable.size() && !rc; ++i) {
    const int64_t want = (N + 255) / 256;
    bytes_to_bitmask_kernel<<<(int)(want < (int64_t)num_sms * 8 ? want : (int64_t)num_sms * 8), 256, 0, s>>>(col_bytes[nullable[i]], N, (uint32_t*)out_cols[nullable[i]].null_mask, d_nulls + i);
    rc = check_cuda(note_launch(), "bytes_to_bitmask");
  }
  // a mask supplied for a column without nulls is returned all-valid
  for (int c = 0; c < num_cols && !rc; ++c)
    if (!col_bytes[c] && out_cols[c].null_mask) rc = check_cuda(cudaMemsetAsync(out_cols[c].null_mask, 0xff, sizeof(uint32_t) * (size_t)((N + 31) / 32), s), "mask fill");
  std::vector<uint32_t> host(P + 1 + nullable.size(), 0);
  if (!rc) rc = check_cuda(cudaMemcpyAsync(host.data(), pp.offsets, sizeof(uint32_t) * (P + 1 + nullable.size()), cudaMemcpyDeviceToHost, s), "offsets copy");
  pool_free_async(scratch, s);
  if (!rc) rc = check_cuda(cudaStreamSynchronize(s), "hash_partition sync");
  if (rc) return rc;
  for (int i = 0; i < num_partitions; ++i) offsets_host[i] = (int32_t)host[i];
  offsets_host[num_partitions] = (int32_t)host[P];
  for (size_t i = 0; i < nullable.size(); ++i) out_cols[nullable[i]].null_count = (int32_t)host[P + 1 + i];
  return BGB_OK;
}

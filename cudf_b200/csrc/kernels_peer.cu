// Kernels of the peer-memory exchange (see kernels_peer.h). HBM / NVLink-bound byte movement; no tensor-core work.
#include "device_utils.cuh"
#include "kernels_peer.h"

namespace bgb {
namespace {
constexpr int kThreads = 256;

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p)
{
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t global_timer_ns()
{
  uint64_t t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
}  // namespace

// One record per local group: key words + every accumulator cell, written into the OWNER's region at
// [me * cap + position]. A block takes 2048 groups at a time, ranks them per destination in shared memory, reserves one
// run per destination with a single global atomic, and then moves the record column by column through a staging buffer
// ordered by destination: every peer store instruction writes consecutive positions of one region (NVLink wants full
// 128-byte writes; one scattered 8-byte store per thread ran at 240 GB/s per GPU, a quarter of the link).
constexpr int kSRows = 8;
constexpr int kSTile = kThreads * kSRows;

__global__ void __launch_bounds__(kThreads) peer_scatter_kernel(const __grid_constant__ PeerScatterParams p, const uint32_t tile_rows)
{
  __shared__ uint32_t s_cnt[kMaxPeers], s_off[kMaxPeers], s_base[kMaxPeers];
  __shared__ uint64_t stage[kSTile];
  __shared__ uint32_t s_pos[kSTile];   // position inside the destination segment, ~0u: over capacity (counted, not written)
  __shared__ uint8_t s_own[kSTile];
  const int64_t G  = p.num_groups;
  const uint32_t t = threadIdx.x;
  // tile_rows <= kSTile: few groups are spread over many small tiles so that every SM has work (latency-bound exchange)
  for (int64_t tile = (int64_t)blockIdx.x * tile_rows; tile < G; tile += (int64_t)gridDim.x * tile_rows) {
    if (t < (uint32_t)p.world) s_cnt[t] = 0;
    __syncthreads();
    uint32_t slot[kSRows], orank[kSRows];  // owner | rank << 8
    uint64_t k0[kSRows];
    const uint32_t nrec = G - tile < (int64_t)tile_rows ? (uint32_t)(G - tile) : tile_rows;
#pragma unroll
    for (int j = 0; j < kSRows; ++j) {
      const uint32_t r = j * kThreads + t;
      slot[j] = 0; k0[j] = 0; orank[j] = 0;
      if (r < nrec) slot[j] = p.group_slots[tile + r];
    }
#pragma unroll
    for (int j = 0; j < kSRows; ++j) {
      const uint32_t r = j * kThreads + t;
      if (r >= nrec) continue;
      uint64_t k1 = 0;
      if (p.key_words == 1) k0[j] = p.keys[slot[j]];
      else { k0[j] = p.keys[2 * (uint64_t)slot[j]]; k1 = p.keys[2 * (uint64_t)slot[j] + 1]; }
      const uint32_t owner = owner_mix(hash_key(k0[j], k1, p.key_words), (uint32_t)p.world);
      orank[j] = owner | (atomicAdd(&s_cnt[owner], 1u) << 8);
    }
    __syncthreads();
    if (t == 0) {
      uint32_t run = 0;
      for (int o = 0; o < p.world; ++o) { s_off[o] = run; run += s_cnt[o]; }
    }
    if (t < (uint32_t)p.world && s_cnt[t]) s_base[t] = atomicAdd(p.cursors + t, s_cnt[t]);
    __syncthreads();
    uint32_t spos[kSRows];
#pragma unroll
    for (int j = 0; j < kSRows; ++j) {
      const uint32_t r = j * kThreads + t;
      spos[j] = 0;
      if (r >= nrec) continue;
      const uint32_t owner = orank[j] & 0xffu, rank = orank[j] >> 8;
      spos[j] = s_off[owner] + rank;
      const uint64_t pos = (uint64_t)s_base[owner] + rank;
      s_own[spos[j]] = (uint8_t)owner;
      s_pos[spos[j]] = pos < p.cap ? (uint32_t)pos : ~0u;
      stage[spos[j]] = k0[j];
    }
    __syncthreads();
    const uint64_t mine = (uint64_t)p.me * p.cap;
#pragma unroll
    for (int j = 0; j < kSRows; ++j) {
      const uint32_t i = j * kThreads + t;
      if (i < nrec && s_pos[i] != ~0u) reinterpret_cast<uint64_t*>(p.region[s_own[i]])[mine + s_pos[i]] = stage[i];
    }
    if (p.key_words == 2) {
      __syncthreads();
#pragma unroll
      for (int j = 0; j < kSRows; ++j) if (j * kThreads + t < nrec) stage[spos[j]] = p.keys[2 * (uint64_t)slot[j] + 1];
      __syncthreads();
#pragma unroll
      for (int j = 0; j < kSRows; ++j) {
        const uint32_t i = j * kThreads + t;
        if (i < nrec && s_pos[i] != ~0u) reinterpret_cast<uint64_t*>(p.region[s_own[i]] + p.key1_off)[mine + s_pos[i]] = stage[i];
      }
    }
#pragma unroll 1
    for (int a = 0; a < p.num_accs; ++a) {
      const PeerAcc& pa = p.acc[a];
      __syncthreads();  // the previous column's read-out of `stage` is complete
      if (pa.cell_bytes == 8) {
#pragma unroll
        for (int j = 0; j < kSRows; ++j) if (j * kThreads + t < nrec) stage[spos[j]] = reinterpret_cast<const uint64_t*>(pa.cells)[slot[j]];
      } else {
#pragma unroll
        for (int j = 0; j < kSRows; ++j) if (j * kThreads + t < nrec) stage[spos[j]] = reinterpret_cast<const uint32_t*>(pa.cells)[slot[j]];
      }
      __syncthreads();
      if (pa.cell_bytes == 8) {
#pragma unroll
        for (int j = 0; j < kSRows; ++j) {
          const uint32_t i = j * kThreads + t;
          if (i < nrec && s_pos[i] != ~0u) reinterpret_cast<uint64_t*>(p.region[s_own[i]] + pa.region_off)[mine + s_pos[i]] = stage[i];
        }
      } else {
#pragma unroll
        for (int j = 0; j < kSRows; ++j) {
          const uint32_t i = j * kThreads + t;
          if (i < nrec && s_pos[i] != ~0u) reinterpret_cast<uint32_t*>(p.region[s_own[i]] + pa.region_off)[mine + s_pos[i]] = (uint32_t)stage[i];
        }
      }
    }
    __syncthreads();
  }
}

cudaError_t launch_peer_scatter(const PeerScatterParams& p, cudaStream_t s)
{
  if (p.num_groups <= 0) return cudaSuccess;
  uint32_t tile_rows = kSTile;
  while (tile_rows > (uint32_t)kThreads && (int64_t)p.num_groups < (int64_t)device_sm_count() * 2 * tile_rows) tile_rows /= 2;
  int64_t blocks = ((int64_t)p.num_groups + tile_rows - 1) / tile_rows;
  if (blocks > device_sm_count() * 6) blocks = device_sm_count() * 6;
  peer_scatter_kernel<<<(int)blocks, kThreads, 0, s>>>(p, tile_rows);
  return note_launch();
}

// Runs after the scatter kernel in stream order: thread d tells rank d how many records this rank left in its segment,
// then publishes the step's epoch with a system-scope release (cumulative over the scatter kernel's stores).
__global__ void peer_signal_kernel(const __grid_constant__ PeerSignalParams p)
{
  const int d = threadIdx.x;
  if (d >= p.world) return;
  const uint32_t produced = p.abort ? 0u : p.cursors[d];
  PeerHeader* h           = p.header[d];
  // an overflow towards ANY destination is reported to EVERY rank, so that all ranks fail the step together; so is an abort
  uint32_t overflow = p.abort ? 2u : 0u;
  if (!p.abort) for (int r = 0; r < p.world; ++r) overflow |= p.cursors[r] > p.cap ? 1u : 0u;
  h->count[p.parity][p.me]    = produced < p.cap ? produced : (uint32_t)p.cap;
  h->overflow[p.parity][p.me] = overflow;
  __threadfence_system();
  st_release_sys(&h->flag[p.parity][p.me], p.epoch);
}

cudaError_t launch_peer_signal(const PeerSignalParams& p, cudaStream_t s)
{
  peer_signal_kernel<<<1, kMaxPeers, 0, s>>>(p);
  return note_launch();
}

// One block, one thread per source: spins (bounded) on that source's flag. A single small block so that ranks sharing
// one GPU (the simulated-rank tests) cannot starve each other of SMs.
__global__ void peer_wait_kernel(const PeerHeader* mine, int parity, uint32_t epoch, int world, uint64_t timeout_ns, uint32_t* out)
{
  const int s = threadIdx.x;
  if (s == 0) out[2 * world] = 0;
  __syncthreads();
  if (s >= world) return;
  const uint64_t t0 = global_timer_ns();
  bool ok           = true;
  // flags only grow; (int32) difference tolerates wrap-around of the epoch counter
  while ((int32_t)(ld_acquire_sys(&mine->flag[parity][s]) - epoch) < 0) {
    if (global_timer_ns() - t0 > timeout_ns) { ok = false; break; }
    __nanosleep(200);
  }
  if (!ok) { out[2 * world] = 1; out[s] = 0; out[world + s] = 0; return; }
  out[s]         = *reinterpret_cast<const volatile uint32_t*>(&mine->count[parity][s]);
  out[world + s] = *reinterpret_cast<const volatile uint32_t*>(&mine->overflow[parity][s]);
}

cudaError_t launch_peer_wait(const PeerHeader* mine, int parity, uint32_t epoch, int world, uint64_t timeout_ns, uint32_t* out, cudaStream_t s)
{
  peer_wait_kernel<<<1, kMaxPeers, 0, s>>>(mine, parity, epoch, world, timeout_ns, out);
  return note_launch();
}

// merge one received record into the final table: the associative merges of streaming_groupby::merge
// (/root/reference/cpp/src/groupby/streaming_groupby/merge.cu:30-144) applied to accumulator cells
__device__ __forceinline__ void merge_cell(uint8_t op, void* cells, uint64_t slot, uint64_t v)
{
  unsigned long long* c64 = reinterpret_cast<unsigned long long*>(cells) + slot;
  unsigned int* c32       = reinterpret_cast<unsigned int*>(cells) + slot;
  switch (op) {
    case ACC_SUM_I64: case ACC_SUMSQ_I64: case ACC_SUM_LO32: case ACC_SUM_HI32: atomicAdd(c64, (unsigned long long)v); break;
    case ACC_SUM_F64: case ACC_SUMSQ_F64: atomicAdd(reinterpret_cast<double*>(c64), __longlong_as_double((long long)v)); break;
    case ACC_SUM_F32: case ACC_SUMSQ_F32: atomicAdd(reinterpret_cast<float*>(c32), __uint_as_float((uint32_t)v)); break;
    case ACC_PROD_I64: atomic_mul_i64(c64, (int64_t)v); break;
    case ACC_PROD_F64: atomic_mul_f64(c64, __longlong_as_double((long long)v)); break;
    case ACC_PROD_F32: atomic_mul_f32(c32, __uint_as_float((uint32_t)v)); break;
    case ACC_MIN: atomicMin(c64, (unsigned long long)v); break;
    case ACC_MAX: atomicMax(c64, (unsigned long long)v); break;
    case ACC_COUNT_VALID: case ACC_COUNT_ALL: atomicAdd(c32, (unsigned int)v); break;
    default: break;  // ARGMIN / ARGMAX hold rank-local row numbers: rejected before the exchange
  }
}

__global__ void __launch_bounds__(kThreads) peer_merge_kernel(const __grid_constant__ PeerMergeParams p)
{
  // Sources are walked in LOCKSTEP, 32 records of one source per warp: every source wrote its records in the slot order of
  // its local table, and the final table is indexed by the same hash, so all sources sweep the final table together and the
  // world records of one group meet in L2 (source-by-source order re-read every table sector from HBM world times).
  uint32_t longest = 0;
  for (int r = 0; r < p.world; ++r) { const uint32_t c = p.prefix[r + 1] - p.prefix[r]; longest = c > longest ? c : longest; }
  const uint64_t chunks = ((uint64_t)longest + 31) / 32 * (uint32_t)p.world;   // (chunk, source) pairs
  const uint64_t warps  = (uint64_t)gridDim.x * (kThreads / 32);
  for (uint64_t w = (uint64_t)blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5); w < chunks; w += warps) {
    const int src    = (int)(w % (uint32_t)p.world);
    const uint32_t j = (uint32_t)(w / (uint32_t)p.world) * 32 + (threadIdx.x & 31);
    if (j >= p.prefix[src + 1] - p.prefix[src]) continue;
    const uint64_t idx = (uint64_t)src * p.cap + j;
    // .cg loads: the records were written by another GPU, never through this SM's L1
    const uint64_t k0 = __ldcg(reinterpret_cast<const unsigned long long*>(p.region) + idx);
    const uint64_t k1 = p.key_words == 2 ? __ldcg(reinterpret_cast<const unsigned long long*>(p.region + p.key1_off) + idx) : 0;
    const uint64_t slot = p.slot_hash32 ? find_or_insert_from(p.table, k0, slot_hash32(k0)) : find_or_insert(p.table, k0, k1);
    if (slot == kNoSlot) continue;
#pragma unroll 1
    for (int a = 0; a < p.num_accs; ++a) {
      const PeerAcc& pa = p.acc[a];
      const uint64_t v  = pa.cell_bytes == 8 ? __ldcg(reinterpret_cast<const unsigned long long*>(p.region + pa.region_off) + idx)
                                             : __ldcg(reinterpret_cast<const unsigned int*>(p.region + pa.region_off) + idx);
      merge_cell(pa.op, pa.cells, slot, v);
    }
  }
}

// find_or_insert that also reports whether the key was new (the streaming object tracks its distinct-key count)
__device__ __forceinline__ uint64_t find_or_insert_counting(const TableDesc& t, uint64_t k0, uint64_t k1, uint32_t* new_keys)
{
  const uint64_t cap = t.capacity;
  if (t.key_words == 1) {
    if (k0 == kEmpty) { if (atomicExch(&t.status[2], 1u) == 0u) atomicAdd(new_keys, 1u); return cap; }
    uint64_t slot = hash_key(k0, 0, 1) & (cap - 1);
    unsigned long long* keys = reinterpret_cast<unsigned long long*>(t.keys);
    for (uint32_t probes = 0; probes < kMaxProbes; ++probes) {
      unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(keys + slot);
      if (cur == k0) return slot;
      if (cur == kEmpty) {
        unsigned long long old = atomicCAS(keys + slot, (unsigned long long)kEmpty, (unsigned long long)k0);
        if (old == kEmpty) { atomicAdd(new_keys, 1u); return slot; }
        if (old == k0) return slot;
      }
      slot = (slot + 1) & (cap - 1);
    }
  } else {
    if (k0 == kEmpty && k1 == kEmpty) { if (atomicExch(&t.status[2], 1u) == 0u) atomicAdd(new_keys, 1u); return cap; }
    uint64_t slot = hash_key(k0, k1, 2) & (cap - 1);
    for (uint32_t probes = 0; probes < kMaxProbes; ++probes) {
      uint64_t o0, o1;
      cas128(t.keys + 2 * slot, kEmpty, kEmpty, k0, k1, o0, o1);  // merges are rare relative to rows: always take the CAS
      if (o0 == kEmpty && o1 == kEmpty) { atomicAdd(new_keys, 1u); return slot; }
      if (o0 == k0 && o1 == k1) return slot;
      slot = (slot + 1) & (cap - 1);
    }
  }
  t.status[0] = 1;
  return kNoSlot;
}

__global__ void __launch_bounds__(kThreads) table_merge_kernel(const __grid_constant__ TableMergeParams p)
{
  const uint32_t stride = gridDim.x * kThreads;
  for (uint32_t j = blockIdx.x * kThreads + threadIdx.x; j < (uint32_t)p.num_groups; j += stride) {
    const uint64_t src = p.src_group_slots[j];
    const uint64_t k0  = p.key_words == 1 ? p.src_keys[src] : p.src_keys[2 * src];
    const uint64_t k1  = p.key_words == 1 ? 0 : p.src_keys[2 * src + 1];
    const uint64_t dst = find_or_insert_counting(p.dst, k0, k1, p.new_keys);
    if (dst == kNoSlot) continue;
#pragma unroll 1
    for (int a = 0; a < p.num_accs; ++a) {
      const PeerAcc& pa = p.src_acc[a];
      const uint64_t v  = pa.cell_bytes == 8 ? reinterpret_cast<const uint64_t*>(pa.cells)[src] : reinterpret_cast<const uint32_t*>(pa.cells)[src];
      merge_cell(pa.op, p.dst_cells[a], dst, v);
    }
  }
}

cudaError_t launch_table_merge(const TableMergeParams& p, cudaStream_t s)
{
  if (p.num_groups <= 0) return cudaSuccess;
  int64_t blocks = ((int64_t)p.num_groups + kThreads - 1) / kThreads;
  if (blocks > device_sm_count() * 8) blocks = device_sm_count() * 8;
  table_merge_kernel<<<(int)blocks, kThreads, 0, s>>>(p);
  return note_launch();
}

cudaError_t launch_peer_merge(const PeerMergeParams& p, cudaStream_t s)
{
  const uint32_t total = p.prefix[p.world];
  if (total == 0) return cudaSuccess;
  uint32_t blocks = (total + kThreads - 1) / kThreads;
  if (blocks > (uint32_t)device_sm_count() * 8) blocks = (uint32_t)device_sm_count() * 8;
  peer_merge_kernel<<<blocks, kThreads, 0, s>>>(p);
  return note_launch();
}

}  // namespace bgb

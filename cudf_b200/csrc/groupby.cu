// C ABI: bgb_groupby_build / _emit / _result_info / _destroy.  See include/bgb.h for the contract.
#include <unistd.h>

#include <chrono>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "errors.h"
#include "fastpaths.h"
#include "kernels_bucket.h"
#include "kernels_hc.h"
#include "kernels_partition.h"
#include "kernels_peer.h"
#include "plan.h"
#include "pool.h"
#include "profiler.h"

using namespace bgb;

struct bgb_groupby {
  Plan plan;
  bgb_allocator alloc{};
  bool has_alloc = false;
  cudaStream_t stream = nullptr;
  struct Block { void* p; size_t bytes; };
  std::vector<Block> blocks;
  // device state
  TableDesc table{};
  uint32_t* group_slots = nullptr;
  int32_t num_groups = 0;
  std::string path = "empty";
  Profiler prof;
  std::vector<KeyPlan> emit_keys;   // repacked keys: the caller's key columns and where each sits in the packed 64-bit image
  void* repack_buf = nullptr;       // the packed key column (+ its validity words), released when the build is done
  void* repack_mask = nullptr;
  bool dense = false;               // the build wrote the groups densely at [0, G) and filled group_slots itself (kernels_hc.cu)
  bool peer_fused = false;          // ... or wrote them straight into the owner ranks' exchange regions (no local groups at all)
  uint32_t* peer_cursors = nullptr; // records produced per destination by the fused slices
  bool dense_out = false;           // the finished handle holds its groups densely: group j sits at entry j (the reserved-key group, if any, last)
  bool reserved_used = false;       // ... and its last group is the reserved entry (index table_slots - 1)
  uint64_t table_slots = 0;         // entries of table.keys / the accumulator arrays (capacity + 1, or fewer: see build_local)

  void* dalloc(size_t bytes, bool transient = false)
  {
    if (bytes == 0) bytes = 16;
    void* p = nullptr;
    if (has_alloc) p = alloc.allocate(bytes, (void*)stream, alloc.ctx);
    else p = pool_alloc_async(bytes, stream, transient);
    if (p) blocks.push_back({p, bytes});
    return p;
  }
  void dfree(void* p)
  {
    for (size_t i = 0; i < blocks.size(); ++i)
      if (blocks[i].p == p) {
        if (has_alloc) alloc.deallocate(p, blocks[i].bytes, (void*)stream, alloc.ctx);
        else pool_free_async(p, stream);
        blocks.erase(blocks.begin() + i);
        return;
      }
  }
  void free_all()
  {
    while (!blocks.empty()) dfree(blocks.back().p);
  }
};

struct bgb_peer {
  int rank = 0, world = 1, device = 0;
  uint8_t* buffer = nullptr;          // header page + two regions (double-buffered by step parity)
  size_t region_bytes = 0;
  uint8_t* peer_buffer[kMaxPeers] = {};  // every rank's buffer as mapped into this process (own buffer at [rank])
  bool opened[kMaxPeers] = {};           // mapped through cudaIpcOpenMemHandle (must be closed)
  bool same_process = false;             // some peer was attached with bgb_peer_attach_local
  uint32_t epoch = 0;
  uint64_t timeout_ns = 20ull * 1000 * 1000 * 1000;
  bool poisoned = false;                 // a wait timed out or the abort signal could not be sent: epochs may differ between ranks
  double merge_ratio = 1.0;              // groups per received record in the last step's merge (sizes the next merge's slices)
};

// record layout of the peer exchange inside a region: key word arrays, then one array per accumulator, each world * cap entries
struct PeerLayout {
  size_t record_bytes = 0;
  uint64_t cap = 0, entries = 0, key1_off = 0;
  std::vector<uint64_t> acc_off;
};
static bool peer_layout(const bgb::Plan& pl, const bgb_peer* peer, PeerLayout& L)
{
  L.record_bytes = 8 * (size_t)pl.key_words;
  for (const bgb::AccPlan& a : pl.accs) L.record_bytes += bgb::acc_cell_bytes(a.op);
  L.cap = (uint64_t)((peer->region_bytes / ((size_t)peer->world * L.record_bytes)) & ~(size_t)31);
  if (L.cap == 0) return false;
  L.entries = (uint64_t)peer->world * L.cap;
  uint64_t off = 8 * L.entries;
  L.key1_off = off;
  if (pl.key_words == 2) off += 8 * L.entries;
  L.acc_off.assign(pl.accs.size(), 0);
  for (size_t a = 0; a < pl.accs.size(); ++a) { L.acc_off[a] = off; off += (uint64_t)bgb::acc_cell_bytes(pl.accs[a].op) * L.entries; }
  return true;
}
// build_local on behalf of bgb_groupby_build_peer: the records-form slices may write their groups straight into the owners' regions
struct PeerFuse { bgb_peer* peer; int parity; };

namespace {

// ---- speculation on the cardinality estimate: ONE host synchronisation per call.
// The estimate only picks the table capacity and the kernel family; no path's correctness depends on it (a table that fills sets
// the overflow flag and the call is repeated, a tiny-kernel abort reruns, keys outside a direct-address range bypass). A call whose
// SHAPE (row count, KEY COLUMN ADDRESSES, column types and masks, aggregations, options, device) was seen before -- the same table
// queried again, or the next batch landing in the same buffers -- therefore starts on the previous
// estimate while a fresh first-stage sample runs beside the kernels; its result is read back with the final status words and, when
// it disagrees with the remembered one, the entry is dropped and the next call estimates synchronously again. BGB_NO_SPECULATE=1
// turns this off.
struct EstimateEntry {
  uint64_t sig = 0;
  bool valid   = false;
  double d1 = 0, m1 = 0;          // first-stage sample (65 536 rows): distinct, sampled
  double d = 0, m = 0;            // the sample the capacity was computed from (second stage when it ran)
  uint64_t max_abs = 0, key_min = 1, key_max = 0;
};
std::mutex g_est_mu;
EstimateEntry g_est_cache[64];
std::atomic<int> g_speculate{-1};  // -1: not decided yet (BGB_NO_SPECULATE), 0 off, 1 on
bool speculation_enabled()
{
  int v = g_speculate.load(std::memory_order_relaxed);
  if (v < 0) { v = getenv("BGB_NO_SPECULATE") ? 0 : 1; g_speculate.store(v, std::memory_order_relaxed); }
  return v != 0;
}

uint64_t shape_signature(const Plan& pl, const bgb_options& opt)
{
  uint64_t h = mix64((uint64_t)pl.num_rows ^ 0x5bd1e995u);
  auto add   = [&](uint64_t v) { h = mix64(h ^ (v + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2))); };
  int dev = 0;
  cudaGetDevice(&dev);
  add((uint64_t)dev);
  add(pl.include_null_keys);
  add(pl.keys.size());
  for (const KeyPlan& k : pl.keys) { add((uint64_t)k.v.type_id); add(k.v.null_mask != nullptr); add((uint64_t)(uintptr_t)k.v.data); add((uint64_t)(int64_t)k.v.offset); }
  add(pl.values.size());
  for (const ValuePlan& v : pl.values) { add((uint64_t)v.v.type_id); add(v.v.null_mask != nullptr); }
  add(pl.accs.size());
  for (const AccPlan& a : pl.accs) { add((uint64_t)a.op); add((uint64_t)(int64_t)a.value_col); }
  add((uint64_t)opt.force_path); add((uint64_t)opt.sample_rows); add((uint64_t)opt.capacity_log2); add((uint64_t)opt.bucket_slice_log2);
  return h ? h : 1;
}

// BGB_TRACE=1: host wall-clock of each phase of a call, printed to stderr (diagnostics only)
struct HostTrace {
  bool on = getenv("BGB_TRACE") != nullptr;
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  std::string line;
  void mark(const char* what)
  {
    if (!on) return;
    auto t = std::chrono::steady_clock::now();
    char buf[64];
    snprintf(buf, sizeof buf, " %s=%.1fus", what, std::chrono::duration<double, std::micro>(t - t0).count());
    line += buf;
    t0 = t;
  }
  void flush(const char* tag) { if (on) fprintf(stderr, "[bgb trace %d] %s:%s\n", (int)getpid(), tag, line.c_str()); }
};
thread_local HostTrace* g_tr = nullptr;
inline void TR(const char* what) { if (g_tr) g_tr->mark(what); }

uint64_t next_pow2(uint64_t x) { uint64_t p = 1; while (p < x) p <<= 1; return p; }

void fill_agg_params(const bgb_groupby& h, AggParams& p, size_t acc_begin, size_t acc_end)
{
  const Plan& pl = h.plan;
  memset(&p, 0, sizeof p);
  p.num_rows       = pl.num_rows;
  p.num_key_cols   = (int)pl.keys.size();
  p.skip_null_keys = pl.skip_null_keys;
  p.wide           = pl.wide ? 1 : 0;
  for (size_t c = 0; c < pl.keys.size(); ++c) {
    const KeyPlan& k = pl.keys[c];
    p.kc[c] = KeyColDesc{(const uint8_t*)k.v.data, k.v.null_mask, k.v.offset, k.width, k.fclass, k.shift, k.null_bit};
  }
  p.num_accs = (int)(acc_end - acc_begin);
  for (size_t a = acc_begin; a < acc_end; ++a) {
    const AccPlan& ap = pl.accs[a];
    AccDesc& d        = p.acc[a - acc_begin];
    if (ap.value_col >= 0) {
      const ValuePlan& v = pl.values[ap.value_col];
      d.data = (const uint8_t*)v.v.data; d.mask = v.v.null_mask; d.offset = v.v.offset; d.src = v.src;
      d.dict = v.dict; d.idx_width = v.idx_width;
    } else { d.data = nullptr; d.mask = nullptr; d.offset = 0; d.src = S_I32; }
    d.op    = ap.op;
    d.cells = ap.cells;
    d.aux   = ap.aux_acc >= 0 ? pl.accs[ap.aux_acc].cells : nullptr;
  }
  p.table = h.table;
}

// solve d = G (1 - exp(-m/G)) for G: expected distinct values among m uniform draws from G groups
double solve_cardinality(double d, double m)
{
  if (d >= m) return 1e18;
  double lo = d, hi = 1e15;
  for (int it = 0; it < 200; ++it) {
    double mid = std::sqrt(lo * hi);
    double e   = mid * (1.0 - std::exp(-m / mid));
    if (e < d) lo = mid; else hi = mid;
    if (hi / lo < 1.0001) break;
  }
  return hi;
}

// High-cardinality path: radix-partition the rows by table region, then run the generic aggregation kernel over the
// partition-ordered copy so that its table traffic stays in L2 (kernels_partition.cu). `h->table` and the accumulator
// cells are allocated; returns BGB_OK with done=false when the plan does not fit the path.
int run_partitioned(bgb_groupby* h, int num_sms, bool& done)
{
  done = false;
  Plan& pl = h->plan;
  const int64_t N = pl.num_rows;
  if (pl.values.size() > (size_t)kMaxPartValues || N >= (1ll << 31)) return BGB_OK;
  for (const AccPlan& a : pl.accs)
    if (a.op == ACC_ARGMIN || a.op == ACC_ARGMAX || a.op == ACC_ARGROW_MIN || a.op == ACC_ARGROW_MAX) return BGB_OK;  // the partition-ordered copy loses the row numbers ARG* report
  cudaStream_t s = h->stream;
  // partitions: table region of capacity/P slots (keys + all cells) around 16 MB, P in [16, 4096]
  size_t slot_bytes = 8 * (size_t)pl.key_words;
  for (const AccPlan& a : pl.accs) slot_bytes += acc_cell_bytes(a.op);
  const double table_bytes = (double)h->table.capacity * slot_bytes;
  int log2cap = 0; while ((1ull << log2cap) < h->table.capacity) ++log2cap;
  static const double region_bytes = getenv("BGB_PART_REGION_MB") ? atof(getenv("BGB_PART_REGION_MB")) * 1e6 : 16e6;
  int log2p = 4; while (log2p < 12 && table_bytes / (double)(1u << log2p) > region_bytes) ++log2p;
  if (log2p > log2cap) log2p = log2cap;
  const uint32_t P = 1u << log2p;

  PartParams pp;
  memset(&pp, 0, sizeof pp);
  fill_agg_params(*h, pp.a, 0, 0);
  pp.log2_parts = log2p;
  pp.slot_shift = log2cap - log2p;
  int key_bits = 0;
  for (const KeyPlan& k : pl.keys) key_bits = std::max(key_bits, std::max((int)k.shift + 8 * k.width, (int)k.null_bit + 1));
  pp.key_out_width = (pl.key_words == 1 && key_bits <= 32) ? 4 : 8;
  std::vector<void*> scratch;
  auto grab = [&](size_t bytes) { void* p = h->dalloc(bytes); if (p) scratch.push_back(p); return p; };
  auto release = [&]() { for (void* p : scratch) h->dfree(p); scratch.clear(); };
  uint32_t* counters = (uint32_t*)grab(sizeof(uint32_t) * (3 * (size_t)P + 1));
  pp.out_k0 = (uint64_t*)grab((size_t)N * pp.key_out_width);
  if (pl.key_words == 2) pp.out_k1 = (uint64_t*)grab((size_t)N * 8);
  bool ok = counters && pp.out_k0 && (pl.key_words == 1 || pp.out_k1);
  pp.counts = counters; pp.cursors = counters + P; pp.offsets = counters + 2 * (size_t)P;
  pp.num_vals = (int)pl.values.size();
  for (size_t c = 0; c < pl.values.size() && ok; ++c) {
    const ValuePlan& v = pl.values[c];
    PartValue& o = pp.v[c];
    o.data = (const uint8_t*)v.v.data; o.mask = v.v.null_mask; o.offset = v.v.offset; o.width = type_width(v.v.type_id);
    o.out_data = (uint8_t*)grab((size_t)N * o.width);
    o.out_valid = v.has_mask ? (uint8_t*)grab((size_t)N) : nullptr;
    ok = o.out_data && (!v.has_mask || o.out_valid);
  }
  if (!ok) { release(); return set_error(BGB_OUT_OF_MEMORY, "partition buffers (%lld rows)", (long long)N); }
  TR("part alloc");
  int rc = check_cuda(cudaMemsetAsync(counters, 0, sizeof(uint32_t) * P, s), "partition counters");
  if (rc) { release(); return rc; }
  h->prof.begin("partition_hist");
  rc = check_cuda(launch_partition_hist(pp, num_sms, s), "partition_hist");
  h->prof.end();
  if (!rc) rc = check_cuda(launch_partition_scan(pp, s), "partition_scan");
  h->prof.begin("partition_scatter");
  if (!rc) rc = check_cuda(launch_partition_scatter(pp, num_sms, s), "partition_scatter");
  h->prof.end();
  if (rc) { release(); return rc; }

  // aggregation over the partitioned copy: packed key words are the key columns now
  const size_t na = pl.accs.size();
  h->prof.begin("init_table");
  for (size_t b = 0; (b == 0 || b < na) && !rc; b += kMaxAccs) {
    AggParams p;
    fill_agg_params(*h, p, b, std::min(na, b + kMaxAccs));
    p.init_keys = (b == 0);
    rc = check_cuda(launch_init_table(p, s), "init_table");
  }
  h->prof.end();
  h->prof.begin("aggregate_partitioned");
  for (size_t b = 0; (b == 0 || b < na) && !rc; b += kMaxAccs) {
    AggParams p;
    fill_agg_params(*h, p, b, std::min(na, b + kMaxAccs));
    p.num_key_cols   = pl.key_words;
    p.skip_null_keys = 0;  // dropped by the scatter; INCLUDE-policy nulls are part of the packed key
    p.kc[0] = KeyColDesc{(const uint8_t*)pp.out_k0, nullptr, 0, (uint8_t)pp.key_out_width, 0, 0, -1};
    if (pl.key_words == 2) p.kc[1] = KeyColDesc{(const uint8_t*)pp.out_k1, nullptr, 0, 8, 0, 64, -1};
    for (int a = 0; a < p.num_accs; ++a) {
      const AccPlan& ap = pl.accs[b + a];
      if (ap.value_col < 0) continue;
      const PartValue& o = pp.v[ap.value_col];
      p.acc[a].data = o.out_data; p.acc[a].mask = (const uint32_t*)o.out_valid; p.acc[a].offset = 0; p.acc[a].byte_mask = o.out_valid != nullptr;
    }
    p.num_rows_dev = pp.offsets + P;
    rc = check_cuda(launch_aggregate_global(p, s), "aggregate_partitioned");
  }
  h->prof.end();
  TR("part launches");
  release();  // stream-ordered: freed after the kernels above
  TR("part release");
  if (rc) return rc;
  done = true;
  return BGB_OK;
}

// Bucketed high-cardinality path (kernels_bucket.cu): one or two scatter levels into fixed-capacity buckets, one per
// shared-memory-sized table slice, then a slice kernel that never touches the HBM table with atomics. `h->table` and the
// accumulator cells are allocated (NOT initialised: every slice is stored whole). done=false when the plan does not fit.
int run_bucketed(bgb_groupby* h, int num_sms, const bgb_options& opt, bool one_level_only, bool& done)
{
  done = false;
  Plan& pl = h->plan;
  const int64_t N = pl.num_rows;
  if (pl.key_words != 1 || pl.values.size() > (size_t)kMaxBucketValues || pl.accs.size() > (size_t)kMaxAccs || N >= (1ll << 31)) return BGB_OK;
  // shared-memory bytes per slot: key + the accumulators that live in shared memory (kernels_bucket.h: BucketAccMode)
  auto acc_mode = [](uint8_t op) -> uint8_t {
    switch (op) {
      case ACC_COUNT_VALID: case ACC_COUNT_ALL: return BM_COUNT;
      case ACC_SUM_I64: case ACC_SUMSQ_I64: return BM_LIMBS;
      case ACC_MIN: case ACC_MAX: return BM_MINMAX;
      default: return BM_GLOBAL;
    }
  };
  auto mode_bytes = [](uint8_t m) -> size_t { return m == BM_COUNT ? 4 : m == BM_LIMBS ? 12 : m == BM_MINMAX ? 8 : 0; };
  size_t slot_bytes = 8;
  for (const AccPlan& a : pl.accs) {
    if (a.op == ACC_ARGMIN || a.op == ACC_ARGMAX || a.op == ACC_ARGROW_MIN || a.op == ACC_ARGROW_MAX) return BGB_OK;  // row numbers are lost by the scatter
    slot_bytes += mode_bytes(acc_mode(a.op));
  }
  int log2cap = 0; while ((1ull << log2cap) < h->table.capacity) ++log2cap;
  // validity word: one bit per masked value column
  int nvb = 0;
  std::vector<int> vbit(pl.values.size(), -1);
  for (size_t c = 0; c < pl.values.size(); ++c) if (pl.values[c].v.null_mask != nullptr) vbit[c] = nvb++;
  int key_bits = 0;
  for (const KeyPlan& k : pl.keys) key_bits = std::max(key_bits, std::max((int)k.shift + 8 * k.width, (int)k.null_bit + 1));
  const bool in_low32 = nvb > 0 && key_bits <= 32;
  const int vbw   = (nvb == 0 || in_low32) ? 0 : (nvb <= 8 ? 1 : 4);  // separate validity column between the levels
  const int vbw_s = nvb == 0 ? 0 : (nvb <= 8 ? 1 : 4);                // validity word per row in shared memory
  auto round16 = [](uint64_t x) { return (x + 15) & ~15ull; };
  // rows a bucket of the LAST level can hold: expected rows + 25 % (one level) or + 27 % (two levels) + slack
  constexpr int kOneLevelBits = 9;      // up to 512 buckets from one scatter level
  static const uint64_t chunk_env = getenv("BGB_BUCKET_CHUNK") ? (uint64_t)atoll(getenv("BGB_BUCKET_CHUNK")) : 0;  // experiments
  static const int slice_env = getenv("BGB_BUCKET_SLICE_LOG2") ? atoi(getenv("BGB_BUCKET_SLICE_LOG2")) : 0;
  const uint64_t kChunkMax = chunk_env ? chunk_env : 16384; // rows of a bucket whose slot numbers are held in shared memory at a time
  auto last_capacity = [&](int bits) { return round16(bits <= kOneLevelBits ? (uint64_t)((double)N / (double)(1u << bits) * 1.25) + 1024 : (uint64_t)((double)N / (double)(1u << bits) * 1.27) + 256); };
  int n_limbs = 0;
  for (const AccPlan& a : pl.accs) n_limbs += acc_mode(a.op) == BM_LIMBS;
  // long buckets (mid cardinalities: few slices, many rows each) are walked in chunks and fold their 16-bit digit sums into
  // a 64-bit total (8 more bytes per slot and integer-sum accumulator) before 65536 addends have gone into them
  auto long_bucket = [&](int bits) { return last_capacity(bits) > 60000; };
  // shared memory of one bucket CTA: the slice + a 16-bit slot and the validity word of every row of a chunk. Three CTAs per SM
  // (<= 72 KB each), or two (<= 108 KB) when that is what lets ONE scatter level reach the slices; opt.bucket_slice_log2
  // fixes the slice instead (tests force several levels on small inputs). A slice has < 65535 slots (16-bit slot numbers).
  auto cta_bytes = [&](int ls_) {
    const int bits   = log2cap - ls_;
    const uint64_t C = std::min<uint64_t>(last_capacity(bits), kChunkMax);
    return ((uint64_t)(slot_bytes + (long_bucket(bits) ? 8 * n_limbs : 0)) << ls_) + round16(C * 2) + round16(C * vbw_s);
  };
  auto best_slice = [&](uint64_t budget) {
    for (int c = std::min(15, log2cap - 1); c >= 4 && log2cap - c <= 18; --c)
      if (cta_bytes(c) <= budget) return c;
    return -1;
  };
  int ls = -1;
  if (opt.bucket_slice_log2 > 0 || slice_env > 0) {
    ls = std::min(opt.bucket_slice_log2 > 0 ? opt.bucket_slice_log2 : slice_env, log2cap - 1);
    if (ls < 4 || ls > 15 || log2cap - ls > 18 || cta_bytes(ls) > 200 * 1024) return BGB_OK;
  } else {
    ls = best_slice(72 * 1024);
    if (ls < 0 || log2cap - ls > kOneLevelBits) {
      const int l2 = best_slice(108 * 1024);
      if (l2 >= 0 && log2cap - l2 <= kOneLevelBits) ls = l2;
    }
    if (ls < 0) return BGB_OK;
  }
  const int b  = log2cap - ls;
  if (one_level_only && b > kOneLevelBits) return BGB_OK;
  const int b1 = b <= kOneLevelBits ? b : (b + 1) / 2, b2 = b - b1;
  const uint32_t P1 = 1u << b1, NB = 1u << b;
  cudaStream_t s = h->stream;

  const uint64_t C1  = b2 ? round16((uint64_t)((double)N / P1 * 1.125) + 1024) : last_capacity(b);
  const uint64_t C2  = last_capacity(b);
  const uint64_t OV  = round16((uint64_t)N / 16 + 4096);
  const uint64_t rows1 = (uint64_t)P1 * C1 + OV + 1;
  const uint64_t rows2 = b2 ? (uint64_t)NB * C2 + OV + 1 : 0;
  if (rows1 >= (1ull << 32) || rows2 >= (1ull << 32)) return BGB_OK;

  std::vector<void*> scratch;
  auto grab = [&](size_t bytes) { void* p = h->dalloc(bytes); if (p) scratch.push_back(p); return p; };
  auto release = [&]() { for (void* p : scratch) h->dfree(p); scratch.clear(); };
  struct Level { uint64_t* key = nullptr; uint8_t* vbits = nullptr; std::vector<uint8_t*> val; };
  auto alloc_level = [&](Level& L, uint64_t rows) {
    L.key = (uint64_t*)grab(rows * 8);
    bool ok = L.key != nullptr;
    if (vbw) { L.vbits = (uint8_t*)grab(rows * vbw); ok = ok && L.vbits; }
    for (size_t c = 0; c < pl.values.size() && ok; ++c) {
      L.val.push_back((uint8_t*)grab(rows * type_width(pl.values[c].v.type_id)));
      ok = L.val.back() != nullptr;
    }
    return ok;
  };
  Level L1, L2;
  const size_t ncount = (size_t)P1 + (b2 ? NB : 0) + 2;
  uint32_t* counters = (uint32_t*)grab(sizeof(uint32_t) * ncount);
  bool ok = counters && alloc_level(L1, rows1) && (!b2 || alloc_level(L2, rows2));
  if (!ok) { release(); return set_error(BGB_OUT_OF_MEMORY, "bucket buffers (%lld rows)", (long long)N); }
  uint32_t* cursor1 = counters;
  uint32_t* cursor2 = counters + P1;
  uint32_t* ov1     = counters + ncount - 2;
  uint32_t* ov2     = counters + ncount - 1;
  TR("bucket alloc");
  int rc = check_cuda(cudaMemsetAsync(counters, 0, sizeof(uint32_t) * ncount, s), "bucket counters");
  if (rc) { release(); return rc; }

  // ---- aggregation parameters (also used by the reserved-slot initialiser)
  BucketAggParams ap;
  memset(&ap, 0, sizeof ap);
  ap.table = h->table;
  ap.key_in_low32 = in_low32; ap.vbits_width = vbw;
  ap.log2_slice = ls; ap.num_buckets = NB;
  ap.num_accs = (int)pl.accs.size();
  const Level& LF = b2 ? L2 : L1;
  uint32_t off = (uint32_t)(8u << ls);
  for (int pass = 0; pass < 2; ++pass)  // 8-byte cells first, then the 4-byte ones: every slice stays 16-byte aligned
    for (size_t a = 0; a < pl.accs.size(); ++a) {
      const AccPlan& acc = pl.accs[a];
      const uint8_t mode = acc_mode(acc.op);
      if ((pass == 0) != (mode == BM_MINMAX)) continue;
      BucketAcc& d = ap.acc[a];
      d.cells = acc.cells; d.op = acc.op; d.mode = mode; d.smem_off = off;
      off += (uint32_t)((mode_bytes(mode) + (mode == BM_LIMBS && long_bucket(b) ? 8 : 0)) << ls);
      if (acc.value_col >= 0) { d.data = LF.val[acc.value_col]; d.src = pl.values[acc.value_col].src; d.vbit = (int16_t)vbit[acc.value_col]; }
      else { d.data = nullptr; d.src = S_I32; d.vbit = -1; }
    }
  const uint64_t CL = std::min<uint64_t>(b2 ? C2 : C1, kChunkMax);  // rows per chunk
  ap.loc_off = off; off += (uint32_t)round16(CL * 2);
  ap.vb_off = off; ap.vb_width = vbw_s; off += (uint32_t)round16(CL * vbw_s);
  ap.chunk_rows = (uint32_t)CL; ap.limb_total = long_bucket(b);
  const size_t slice_smem = off;
  rc = check_cuda(launch_bucket_init_reserved(ap, s), "bucket_init_reserved");

  // ---- level 1
  BucketScatterParams sp;
  memset(&sp, 0, sizeof sp);
  fill_agg_params(*h, sp.a, 0, 0);
  sp.level = 1;
  sp.key_mode = 0;
  if (pl.keys.size() == 1 && pl.keys[0].v.null_mask == nullptr && pl.keys[0].fclass == 0 && (pl.keys[0].width == 4 || pl.keys[0].width == 8)) sp.key_mode = pl.keys[0].width;
  sp.log2_parts = b1; sp.slot_shift = log2cap - b1;
  sp.key_in_low32 = in_low32; sp.vbits_width = vbw;
  sp.cursor = cursor1; sp.bucket_capacity = (uint32_t)C1;
  sp.ov_cursor = ov1; sp.ov_capacity = (uint32_t)OV; sp.ov_base = (uint32_t)((uint64_t)P1 * C1);
  sp.status = h->table.status;
  sp.out_key = L1.key; sp.out_vbits = L1.vbits;
  sp.num_vals = (int)pl.values.size();
  for (size_t c = 0; c < pl.values.size(); ++c) {
    const ValuePlan& v = pl.values[c];
    sp.v[c] = BucketValue{(const uint8_t*)v.v.data, v.v.null_mask, v.v.offset, type_width(v.v.type_id), vbit[c], L1.val[c]};
  }
  h->prof.begin("bucket_scatter1");
  if (!rc) rc = check_cuda(launch_bucket_scatter(sp, num_sms, s), "bucket_scatter1");
  h->prof.end();

  // ---- level 2: every level-1 bucket is a segment
  if (b2 && !rc) {
    BucketScatterParams s2 = sp;
    s2.level = 2;
    s2.in_key = L1.key; s2.in_vbits = L1.vbits;
    s2.seg_cursor = cursor1; s2.seg_capacity = (uint32_t)C1; s2.num_segs = P1;
    s2.log2_parts = b2; s2.slot_shift = log2cap - b;
    s2.cursor = cursor2; s2.bucket_capacity = (uint32_t)C2;
    s2.ov_cursor = ov2; s2.ov_base = (uint32_t)((uint64_t)NB * C2);
    s2.out_key = L2.key; s2.out_vbits = L2.vbits;
    for (size_t c = 0; c < pl.values.size(); ++c) { s2.v[c].in = L1.val[c]; s2.v[c].in_mask = nullptr; s2.v[c].in_offset = 0; s2.v[c].out = L2.val[c]; }
    h->prof.begin("bucket_scatter2");
    rc = check_cuda(launch_bucket_scatter(s2, num_sms, s), "bucket_scatter2");
    h->prof.end();
  }

  // ---- slices, then the overflow rows of the last level (and of level 1 when there were two)
  ap.key = LF.key; ap.vbits = LF.vbits;
  ap.cursor = b2 ? cursor2 : cursor1;
  ap.bucket_capacity = (uint32_t)(b2 ? C2 : C1);
  h->prof.begin("bucket_aggregate");
  if (!rc) rc = check_cuda(launch_bucket_aggregate(ap, slice_smem, s), "bucket_aggregate");
  h->prof.end();
  h->prof.begin("bucket_overflow");
  ap.ov_cursor = b2 ? ov2 : ov1; ap.ov_capacity = (uint32_t)OV; ap.ov_base = (uint32_t)((uint64_t)NB * (b2 ? C2 : C1));
  if (!rc) rc = check_cuda(launch_bucket_overflow(ap, s), "bucket_overflow");
  if (b2 && !rc) {
    BucketAggParams a1 = ap;
    a1.key = L1.key; a1.vbits = L1.vbits;
    for (size_t a = 0; a < pl.accs.size(); ++a) if (pl.accs[a].value_col >= 0) a1.acc[a].data = L1.val[pl.accs[a].value_col];
    a1.ov_cursor = ov1; a1.ov_base = (uint32_t)((uint64_t)P1 * C1);
    rc = check_cuda(launch_bucket_overflow(a1, s), "bucket_overflow");
  }
  h->prof.end();
  TR("bucket launches");
  release();  // stream-ordered
  if (rc) return rc;
  done = true;
  return BGB_OK;
}

// Records form of the bucketed path (kernels_hc.cu): TMA-fed scatter levels over 16- to 32-byte records, then one CTA per
// table slice with DENSE group output -- no HBM hash table, no compaction pass. `h->table` and the accumulator cells are
// allocated (capacity + 1 entries each); on success h->group_slots is allocated and filled and h->dense is set.
// done=false when the plan does not fit (the caller then tries the column-form kernels of kernels_bucket.cu).
// what the cardinality sample saw of the key and value ranges (pack32 eligibility); `unknown` when nothing was sampled
struct HcHints {
  uint64_t key_min = 1, key_max = 0;  // order-preserving key image (raw ^ key_flip); min > max: unknown
  uint64_t max_abs = ~0ull;           // largest |value| of the sampled SUM inputs; ~0: unknown
  bool no_pack32   = false;           // a previous attempt met a row that does not fit
};

int run_hc(bgb_groupby* h, int num_sms, const bgb_options& opt, double g_est, const PeerFuse* fuse, const HcHints& hints, bool& done)
{
  done = false;
  Plan& pl = h->plan;
  const int64_t N = pl.num_rows;
  static const bool off = getenv("BGB_NO_HC") != nullptr;
  if (off || pl.key_words != 1 || pl.accs.size() > (size_t)kMaxAccs || N >= (1ll << 31) || N < 1) return BGB_OK;
  size_t slot_bytes = 8;
  for (const AccPlan& a : pl.accs) {
    if (a.op == ACC_ARGMIN || a.op == ACC_ARGMAX || a.op == ACC_ARGROW_MIN || a.op == ACC_ARGROW_MAX) return BGB_OK;  // row numbers are lost by the scatter
    slot_bytes += acc_cell_bytes(a.op);
  }
  // multi-GPU: fuse the exchange into the slices' flush (every finished slice goes straight to its owners over NVLink)
  static const bool no_fuse = getenv("BGB_NO_PEER_FUSE") != nullptr;
  PeerLayout playout;
  const bool fused = fuse && !no_fuse && fuse->peer->world > 1 && fuse->peer->world <= kHcMaxPeers && peer_layout(pl, fuse->peer, playout);
  if (fused) slot_bytes += 3;  // owner-ordered permutation of the slice (u16 + u8 per slot) behind the table
  // record layout
  int nvb = 0;
  std::vector<int> vbit(pl.values.size(), -1);
  for (size_t c = 0; c < pl.values.size(); ++c) if (pl.values[c].v.null_mask != nullptr) vbit[c] = nvb++;
  int key_bits = 0;
  for (const KeyPlan& k : pl.keys) key_bits = std::max(key_bits, std::max((int)k.shift + 8 * k.width, (int)k.null_bit + 1));
  const bool in_low32 = nvb > 0 && key_bits <= 32;
  int W = 1 + (int)pl.values.size() + ((nvb > 0 && !in_low32) ? 1 : 0);
  if (W > kHcMaxWords || nvb > 32) return BGB_OK;
  const int vbits_word = (nvb > 0 && !in_low32) ? W - 1 : -1;
  // pack32: ONE integer value column without nulls, SUM / COUNT only, key and values within 32 bits -> 8-byte records {key32, value32}
  // instead of 16-byte ones: half the bytes of every pass after the first read (1e9 rows: 80 -> 48 GB). The sample only PREDICTS
  // that everything fits; level 1 checks every row and raises status[7] otherwise (the call then reruns with 16-byte records).
  bool pack32 = false;
  static const bool no_pack32_env = getenv("BGB_HC_NO_PACK32") != nullptr;
  if (!hints.no_pack32 && !no_pack32_env && W == 2 && nvb == 0 && pl.values.size() == 1 && pl.keys.size() == 1 && pl.keys[0].fclass == 0 &&
      pl.keys[0].v.null_mask == nullptr) {
    const ValuePlan& v = pl.values[0];
    const bool int_val = v.src == S_I8 || v.src == S_I16 || v.src == S_I32 || v.src == S_I64 || v.src == S_U8 || v.src == S_U16 || v.src == S_U32 || v.src == S_U64;
    bool shape = int_val && v.v.null_mask == nullptr;
    for (const AccPlan& a : pl.accs) shape = shape && (a.op == ACC_SUM_I64 || a.op == ACC_COUNT_VALID || a.op == ACC_COUNT_ALL);
    const int vw = type_width(v.v.type_id);
    const bool val_fits = vw < 4 || v.src == S_I32 || (hints.max_abs != ~0ull && hints.max_abs < (1ull << 31));
    bool key_fits = pl.keys[0].width <= 4;
    if (!key_fits && hints.key_min <= hints.key_max) {
      const uint64_t flip = smem_eligibility(pl).key_flip;
      key_fits = hints.key_min >= flip && (hints.key_max ^ flip) < (1ull << 32) && (hints.key_min ^ flip) <= (hints.key_max ^ flip);
    }
    pack32 = shape && val_fits && key_fits;
  }
  if (pack32) W = 1;
  int log2cap = 0; while ((1ull << log2cap) < h->table.capacity) ++log2cap;
  auto round16 = [](uint64_t x) { return (x + 15) & ~15ull; };

  // slice: the largest power of two of slots that leaves two CTAs per SM (measured: 8.3 ms per 1e9 rows at 2 x 4096 slots
  // against 12.6 ms at 1 x 8192); at least ~1000 buckets so that every SM has slices to work on
  int ls;
  if (opt.bucket_slice_log2 > 0) {
    ls = std::min(opt.bucket_slice_log2, log2cap - 1);
    if (ls < 4 || (slot_bytes << ls) > 220 * 1024) return BGB_OK;
  } else {
    ls = 4;
    while (ls + 1 < log2cap && (slot_bytes << (ls + 1)) <= 112 * 1024) ++ls;
    if (log2cap - ls < 10) ls = std::max(6, log2cap - 10);
    if (ls >= log2cap) return BGB_OK;
  }
  const int b = log2cap - ls;
  if (b < 1 || b > 20) return BGB_OK;
  const int b2 = b <= 10 ? 0 : b / 2, b1 = b - b2;
  const uint32_t P1 = 1u << b1, NB = 1u << b;
  // rows a final bucket can hold: its groups are a Poisson draw (relative sigma 1 / sqrt(groups per bucket)) of groups whose
  // own sizes vary, so the margin is 27 % at the ~1500 groups per bucket of the high-cardinality shapes and grows as buckets
  // hold fewer groups; beyond +100 % the plan is left to the other paths
  const double gpb    = std::max(1.0, std::min(g_est, (double)N) / (double)NB);
  const double margin = std::max(0.27, 6.0 / std::sqrt(gpb));
  if (margin > 1.0 && opt.force_path != 7) return BGB_OK;
  const uint64_t CL = round16((uint64_t)((double)N / NB * (1.0 + std::min(margin, 1.0))) + 256);  // last level
  const uint64_t C1 = b2 ? round16((uint64_t)((double)N / P1 * 1.125) + 1024) : CL;
  const uint64_t rows1 = (uint64_t)P1 * C1, rows2 = b2 ? (uint64_t)NB * CL : 0;
  if (rows1 >= (1ull << 32) || rows2 >= (1ull << 32)) return BGB_OK;
  cudaStream_t s = h->stream;

  // ONE arena for the record buffers and counters: calls of different sizes (a shard, then the fold of two partial tables) then
  // differ in one block size instead of four, which the pool's best-fit cache (<= 25 % slack) re-uses across them
  constexpr uint32_t kHotStride = 64;  // level-1 cursors 256 B apart: every (tile, bucket) reservation is an atomic on one of P1 words
  const size_t ncur = (size_t)P1 * kHotStride + (b2 ? NB : 0) + 4;
  auto up256 = [](size_t x) { return (x + 255) & ~(size_t)255; };
  const size_t cur_bytes = up256(sizeof(uint32_t) * ncur);
  const size_t l1_bytes  = up256(rows1 * W * 8 + 64);
  const size_t l2_bytes  = b2 ? up256(rows2 * W * 8 + 64) : 0;
  uint8_t* arena = (uint8_t*)h->dalloc(cur_bytes + l1_bytes + l2_bytes, /*transient=*/true);
  auto release = [&]() { if (arena) { h->dfree(arena); arena = nullptr; } };
  uint32_t* counters = (uint32_t*)arena;
  uint64_t* L1 = arena ? (uint64_t*)(arena + cur_bytes) : nullptr;
  uint64_t* L2 = (arena && b2) ? (uint64_t*)(arena + cur_bytes + l1_bytes) : nullptr;
  const uint64_t max_groups = std::min<uint64_t>(h->table_slots, (uint64_t)N + 1);
  h->group_slots = (uint32_t*)h->dalloc(max_groups * sizeof(uint32_t));
  if (!arena || !h->group_slots) {
    release();
    if (h->group_slots) { h->dfree(h->group_slots); h->group_slots = nullptr; }
    return set_error(BGB_OUT_OF_MEMORY, "record buffers of the bucketed path (%lld rows)", (long long)N);
  }
  uint32_t* cursor1 = counters;
  uint32_t* cursor2 = counters + (size_t)P1 * kHotStride;
  unsigned long long* gcount = reinterpret_cast<unsigned long long*>(counters + ((ncur - 4 + 1) & ~(size_t)1));
  TR("hc alloc");
  int rc = check_cuda(cudaMemsetAsync(counters, 0, sizeof(uint32_t) * ncur, s), "hc counters");

  // ---- slices: accumulators in shared memory, 8-byte cells first
  HcSliceParams sp;
  memset(&sp, 0, sizeof sp);
  sp.table = h->table;
  sp.table.capacity = h->table_slots - 1;  // (the slices use it only as the index of the reserved entry: the LAST allocated one)
  sp.words = W; sp.key_in_low32 = in_low32 || pack32; sp.vbits_word = vbits_word; sp.val_hi32 = pack32;
  sp.log2_slice = ls; sp.num_buckets = NB; sp.num_accs = (int)pl.accs.size();
  sp.group_slots = h->group_slots; sp.gcount = gcount; sp.max_groups = max_groups;
  uint32_t off_b = (uint32_t)(8u << ls);
  for (int pass = 0; pass < 2; ++pass)
    for (size_t a = 0; a < pl.accs.size(); ++a) {
      const AccPlan& acc = pl.accs[a];
      const int cb = acc_cell_bytes(acc.op);
      if ((pass == 0) != (cb == 8)) continue;
      HcAcc& d = sp.acc[a];
      d.cells = acc.cells; d.op = acc.op; d.cell_bytes = (uint8_t)cb; d.smem_off = off_b;
      off_b += (uint32_t)cb << ls;
      switch (acc.op) {
        case ACC_COUNT_VALID: case ACC_COUNT_ALL: d.mode = HM_COUNT; break;
        case ACC_SUM_I64: case ACC_SUMSQ_I64: case ACC_SUM_LO32: case ACC_SUM_HI32: d.mode = HM_ISUM; break;
        case ACC_MIN: case ACC_MAX: d.mode = HM_MINMAX; break;
        default: d.mode = HM_OTHER; break;
      }
      if (acc.value_col >= 0) { d.word = (uint8_t)(1 + acc.value_col); d.src = pl.values[acc.value_col].src; d.vbit = (int16_t)vbit[acc.value_col]; }
      else { d.word = 0; d.src = S_I32; d.vbit = -1; }
      if (pack32 && acc.value_col >= 0) { d.word = 0; d.src = S_I64; }  // (the value is the sign-extended upper half of word 0: val_hi32)
      // integer SUMs: the record word is the value as a 64-bit integer whatever the column's width (kernels_hc.cu: record_value;
      // unsigned and bool columns are zero-extended by the loader) -> the plain 64-bit add of the lean kernel
      if (d.mode == HM_ISUM && acc.op == ACC_SUM_I64 && acc.value_col >= 0 &&
          (d.src == S_I8 || d.src == S_I16 || d.src == S_I32 || d.src == S_U8 || d.src == S_U16 || d.src == S_U32))
        d.src = S_I64;
    }
  {  // the float64-statistics form of the slice kernel (KIND 2)
    bool stats = pl.accs.size() >= 1 && pl.accs.size() <= 4;
    int col = -1;
    for (const AccPlan& acc : pl.accs) {
      if (acc.op == ACC_COUNT_ALL) continue;
      const bool op_ok = acc.op == ACC_SUM_F64 || acc.op == ACC_SUMSQ_F64 || acc.op == ACC_MIN || acc.op == ACC_MAX || acc.op == ACC_COUNT_VALID;
      stats = stats && op_ok && acc.value_col >= 0 && (col < 0 || col == acc.value_col) && pl.values[acc.value_col].src == S_F64;
      if (acc.value_col >= 0) col = acc.value_col;
    }
    if (stats && col >= 0) {
      sp.f64stats = 1; sp.stat_vbit = vbit[col];
      for (size_t a = 0; a < pl.accs.size(); ++a) sp.acc[a].word = (uint8_t)(1 + col);  // (COUNT_ALL too: acc[0].word names the column)
    }
  }
  if (fused) {
    bgb_peer* peer = fuse->peer;
    uint32_t* cursors = (uint32_t*)h->dalloc(sizeof(uint32_t) * kMaxPeers);
    if (!cursors) { release(); h->dfree(h->group_slots); h->group_slots = nullptr; return set_error(BGB_OUT_OF_MEMORY, "peer cursors"); }
    if (!rc) rc = check_cuda(cudaMemsetAsync(cursors, 0, sizeof(uint32_t) * kMaxPeers, s), "peer cursors");
    h->peer_cursors = cursors;
    sp.peer.world = peer->world; sp.peer.me = peer->rank; sp.peer.cap = playout.cap; sp.peer.cursors = cursors;
    for (int r = 0; r < peer->world; ++r) sp.peer.region[r] = peer->peer_buffer[r] + kPeerHeaderBytes + (size_t)fuse->parity * peer->region_bytes;
    for (size_t a = 0; a < pl.accs.size(); ++a) sp.peer.acc_off[a] = playout.acc_off[a];
    off_b = (off_b + 15u) & ~15u;
    sp.peer.perm_off = off_b;
    off_b += (uint32_t)(3u << ls);
  }
  const size_t slice_smem = off_b;
  if (!rc) rc = check_cuda(launch_hc_init(sp, s), "hc_init");

  // ---- level 1
  HcScatterParams s1;
  memset(&s1, 0, sizeof s1);
  fill_agg_params(*h, s1.a, 0, 0);
  s1.level = 1; s1.words = W; s1.key_in_low32 = in_low32; s1.vbits_word = vbits_word;
  s1.num_vals = (int)pl.values.size();
  bool tma = pl.keys.size() == 1 && pl.keys[0].v.null_mask == nullptr && pl.keys[0].fclass == 0 &&
             ((uintptr_t)pl.keys[0].v.data + (size_t)pl.keys[0].v.offset * pl.keys[0].width) % 16 == 0;
  for (size_t c = 0; c < pl.values.size(); ++c) {
    const ValuePlan& v = pl.values[c];
    const int w = type_width(v.v.type_id);
    s1.val[c] = HcValue{(const uint8_t*)v.v.data, v.v.null_mask, v.v.offset, w, vbit[c], (int32_t)v.src};
    tma = tma && ((uintptr_t)v.v.data + (size_t)v.v.offset * w) % 16 == 0;
    if (v.v.null_mask) tma = tma && v.v.offset % 128 == 0 && ((uintptr_t)v.v.null_mask + (size_t)v.v.offset / 8) % 16 == 0;
  }
  static const bool no_tma = getenv("BGB_HC_NO_TMA") != nullptr;
  s1.tma_src = tma && !no_tma;
  if (pack32 && !s1.tma_src) {  // (only the TMA loader packs: decided before anything was launched that depends on W... see below)
    release(); h->dfree(h->group_slots); h->group_slots = nullptr;
    if (h->peer_cursors) { h->dfree(h->peer_cursors); h->peer_cursors = nullptr; }
    HcHints again = hints; again.no_pack32 = true;
    return run_hc(h, num_sms, opt, g_est, fuse, again, done);
  }
  s1.pack32 = pack32; s1.key_in_low32 = in_low32 || pack32;
  s1.out = L1; s1.cursor = cursor1; s1.cursor_stride = kHotStride; s1.bucket_capacity = (uint32_t)C1;
  s1.log2_parts = b1; s1.slot_shift = log2cap - b1; s1.status = h->table.status;
  h->prof.begin("hc_scatter1");
  if (!rc) rc = check_cuda(launch_hc_scatter(s1, num_sms, s), "hc_scatter1");
  h->prof.end();

  // ---- level 2: every level-1 bucket is a segment
  if (b2 && !rc) {
    HcScatterParams s2 = s1;
    s2.level = 2; s2.tma_src = 0;
    s2.in = L1; s2.seg_cursor = cursor1; s2.seg_cursor_stride = kHotStride; s2.seg_capacity = (uint32_t)C1; s2.num_segs = P1;
    s2.out = L2; s2.cursor = cursor2; s2.cursor_stride = 1; s2.bucket_capacity = (uint32_t)CL;
    s2.log2_parts = b2; s2.slot_shift = log2cap - b;
    h->prof.begin("hc_scatter2");
    rc = check_cuda(launch_hc_scatter(s2, num_sms, s), "hc_scatter2");
    h->prof.end();
  }

  // ---- slices + group count
  sp.rec = b2 ? L2 : L1;
  sp.cursor = b2 ? cursor2 : cursor1; sp.cursor_stride = b2 ? 1 : kHotStride;
  sp.bucket_capacity = (uint32_t)(b2 ? CL : C1);
  h->prof.begin("hc_slices");
  if (!rc) rc = check_cuda(launch_hc_slices(sp, slice_smem, num_sms, s), "hc_slices");
  h->prof.end();
  if (!rc) rc = check_cuda(launch_hc_finish(sp, s), "hc_finish");
  TR("hc launches");
  release();  // stream-ordered
  if (rc) { h->dfree(h->group_slots); h->group_slots = nullptr; return rc; }
  h->dense = true;
  h->peer_fused = fused;
  done = true;
  return BGB_OK;
}

int num_sms_of_current_device()
{
  int num_sms = 148, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
  return num_sms;
}

// Final merge of the peer exchange through the records-form kernels: level 1 reads the received records (MODE 3 of
// hc_scatter_kernel), the slices ADD partial cells (sum of sums, sum of counts, min of mins ... the merges of
// /root/reference/cpp/src/groupby/streaming_groupby/merge.cu:30-144) and write the owner's groups densely. `h->table` and the
// accumulator cells are allocated for `capacity`; on success group_slots is allocated + filled. merged=false: shape not taken.
int run_hc_merge(bgb_groupby* h, int num_sms, const uint8_t* region, uint64_t cap, int world, const std::vector<uint32_t>& prefix,
                 const std::vector<uint64_t>& acc_off, uint64_t total, bool& merged)
{
  merged = false;
  Plan& pl = h->plan;
  if (pl.key_words != 1 || pl.accs.size() > (size_t)(kHcMaxWords - 1) || world > kHcMaxPeers || total >= (1ull << 31)) return BGB_OK;
  const int W = 1 + (int)pl.accs.size();
  size_t slot_bytes = 8;
  for (const AccPlan& a : pl.accs) slot_bytes += acc_cell_bytes(a.op);
  int log2cap = 0; while ((1ull << log2cap) < h->table.capacity) ++log2cap;
  auto round16 = [](uint64_t x) { return (x + 15) & ~15ull; };
  int ls = 4;
  while (ls + 1 < log2cap && (slot_bytes << (ls + 1)) <= 112 * 1024) ++ls;
  if (log2cap - ls < 10) ls = std::max(6, log2cap - 10);
  if (ls >= log2cap) return BGB_OK;
  const int b = log2cap - ls;
  if (b < 1 || b > 20) return BGB_OK;
  const int b2 = b <= 10 ? 0 : b / 2, b1 = b - b2;
  const uint32_t P1 = 1u << b1, NB = 1u << b;
  const int64_t N = (int64_t)total;
  // every group arrives once per source that saw it: a bucket's rows are ~world records of each of its groups
  const double gpb    = std::max(1.0, (double)N / std::max(1, world) / (double)NB);
  const double margin = std::max(0.27, 6.0 / std::sqrt(gpb));
  if (margin > 1.0) return BGB_OK;
  const uint64_t CL = round16((uint64_t)((double)N / NB * (1.0 + margin)) + 256);
  const uint64_t C1 = b2 ? round16((uint64_t)((double)N / P1 * 1.125) + 1024) : CL;
  const uint64_t rows1 = (uint64_t)P1 * C1, rows2 = b2 ? (uint64_t)NB * CL : 0;
  if (rows1 >= (1ull << 32) || rows2 >= (1ull << 32)) return BGB_OK;
  cudaStream_t s = h->stream;
  constexpr uint32_t kHotStride = 64;
  const size_t ncur = (size_t)P1 * kHotStride + (b2 ? NB : 0) + 4;
  auto up256 = [](size_t x) { return (x + 255) & ~(size_t)255; };
  const size_t cur_bytes = up256(sizeof(uint32_t) * ncur), l1_bytes = up256(rows1 * W * 8 + 64), l2_bytes = b2 ? up256(rows2 * W * 8 + 64) : 0;
  uint8_t* arena = (uint8_t*)h->dalloc(cur_bytes + l1_bytes + l2_bytes, /*transient=*/true);
  const uint64_t max_groups = std::min<uint64_t>(h->table_slots, (uint64_t)N + 1);
  h->group_slots = (uint32_t*)h->dalloc(max_groups * sizeof(uint32_t));
  if (!arena || !h->group_slots) {
    if (arena) h->dfree(arena);
    if (h->group_slots) { h->dfree(h->group_slots); h->group_slots = nullptr; }
    return set_error(BGB_OUT_OF_MEMORY, "record buffers of the merge (%lld records)", (long long)N);
  }
  uint32_t* counters = (uint32_t*)arena;
  uint64_t* L1 = (uint64_t*)(arena + cur_bytes);
  uint64_t* L2 = b2 ? (uint64_t*)(arena + cur_bytes + l1_bytes) : nullptr;
  uint32_t* cursor1 = counters;
  uint32_t* cursor2 = counters + (size_t)P1 * kHotStride;
  unsigned long long* gcount = reinterpret_cast<unsigned long long*>(counters + ((ncur - 4 + 1) & ~(size_t)1));
  int rc = check_cuda(cudaMemsetAsync(counters, 0, sizeof(uint32_t) * ncur, s), "merge counters");

  HcSliceParams sp;
  memset(&sp, 0, sizeof sp);
  sp.table = h->table;
  sp.table.capacity = h->table_slots - 1;
  sp.words = W; sp.key_in_low32 = 0; sp.vbits_word = -1;
  sp.log2_slice = ls; sp.num_buckets = NB; sp.num_accs = (int)pl.accs.size();
  sp.group_slots = h->group_slots; sp.gcount = gcount; sp.max_groups = max_groups;
  uint32_t off_b = (uint32_t)(8u << ls);
  for (int pass = 0; pass < 2; ++pass)
    for (size_t a = 0; a < pl.accs.size(); ++a) {
      const AccPlan& acc = pl.accs[a];
      const int cb = acc_cell_bytes(acc.op);
      if ((pass == 0) != (cb == 8)) continue;
      HcAcc& d = sp.acc[a];
      d.cells = acc.cells; d.cell_bytes = (uint8_t)cb; d.smem_off = off_b; d.vbit = -1; d.word = (uint8_t)(1 + a);
      off_b += (uint32_t)cb << ls;
      switch (acc.op) {  // merge of two partial cells of this accumulator
        case ACC_COUNT_VALID: case ACC_COUNT_ALL: d.mode = HM_ADD32; d.op = acc.op; d.src = S_U32; break;
        case ACC_SUM_I64: case ACC_SUMSQ_I64: case ACC_SUM_LO32: case ACC_SUM_HI32: d.mode = HM_ISUM; d.op = ACC_SUM_I64; d.src = S_I64; break;
        case ACC_MIN: case ACC_MAX: d.mode = HM_MINMAX; d.op = acc.op; d.src = S_U64; break;   // cells are order images already
        case ACC_SUM_F64: case ACC_SUMSQ_F64: d.mode = HM_OTHER; d.op = ACC_SUM_F64; d.src = S_F64; break;
        case ACC_SUM_F32: case ACC_SUMSQ_F32: d.mode = HM_OTHER; d.op = ACC_SUM_F32; d.src = S_F32; break;
        case ACC_PROD_I64: d.mode = HM_OTHER; d.op = ACC_PROD_I64; d.src = S_I64; break;
        case ACC_PROD_F64: d.mode = HM_OTHER; d.op = ACC_PROD_F64; d.src = S_F64; break;
        case ACC_PROD_F32: d.mode = HM_OTHER; d.op = ACC_PROD_F32; d.src = S_F32; break;
        default: h->dfree(arena); h->dfree(h->group_slots); h->group_slots = nullptr; return BGB_OK;
      }
    }
  const size_t slice_smem = off_b;
  if (!rc) rc = check_cuda(launch_hc_init(sp, s), "hc_init");

  HcScatterParams s1;
  memset(&s1, 0, sizeof s1);
  s1.a.num_rows = N; s1.a.table = h->table;
  s1.level = 1; s1.words = W; s1.key_in_low32 = 0; s1.vbits_word = -1; s1.num_vals = 0;
  s1.peer_in.region = region; s1.peer_in.cap = cap; s1.peer_in.world = world; s1.peer_in.num_accs = (int)pl.accs.size();
  for (int r = 0; r <= world; ++r) s1.peer_in.prefix[r] = prefix[r];
  for (size_t a = 0; a < pl.accs.size(); ++a) { s1.peer_in.acc_off[a] = acc_off[a]; s1.peer_in.cell_bytes[a] = (uint8_t)acc_cell_bytes(pl.accs[a].op); }
  s1.out = L1; s1.cursor = cursor1; s1.cursor_stride = kHotStride; s1.bucket_capacity = (uint32_t)C1;
  s1.log2_parts = b1; s1.slot_shift = log2cap - b1; s1.status = h->table.status;
  if (!rc) rc = check_cuda(launch_hc_scatter(s1, num_sms, s), "hc_merge_scatter1");
  if (b2 && !rc) {
    HcScatterParams s2 = s1;
    s2.level = 2; s2.peer_in.region = nullptr;
    s2.in = L1; s2.seg_cursor = cursor1; s2.seg_cursor_stride = kHotStride; s2.seg_capacity = (uint32_t)C1; s2.num_segs = P1;
    s2.out = L2; s2.cursor = cursor2; s2.cursor_stride = 1; s2.bucket_capacity = (uint32_t)CL;
    s2.log2_parts = b2; s2.slot_shift = log2cap - b;
    rc = check_cuda(launch_hc_scatter(s2, num_sms, s), "hc_merge_scatter2");
  }
  sp.rec = b2 ? L2 : L1;
  sp.cursor = b2 ? cursor2 : cursor1; sp.cursor_stride = b2 ? 1 : kHotStride;
  sp.bucket_capacity = (uint32_t)(b2 ? CL : C1);
  if (!rc) rc = check_cuda(launch_hc_slices(sp, slice_smem, num_sms, s), "hc_merge_slices");
  if (!rc) rc = check_cuda(launch_hc_finish(sp, s), "hc_finish");
  h->dfree(arena);  // stream-ordered
  if (rc) { h->dfree(h->group_slots); h->group_slots = nullptr; return rc; }
  merged = true;
  return BGB_OK;
}

}  // namespace

// Key repacking: the caller's integer key columns would need a 128-bit image, but their exact value RANGES fit 64 bits
// together: build one packed UINT64 key column ((value - minimum) of every column, null flags on top) and re-plan the call
// on it, so that the single-word paths (shared-memory tables, tiny kernels, bucketed path) serve it instead of the generic
// 128-bit CAS kernel. Costs one min/max pass and one packing pass over the key columns. done=false: plan left untouched.
int try_repack_keys(bgb_groupby* h, const bgb_column_view* keys, int32_t num_keys, int32_t include_null_keys,
                    const bgb_request* requests, int32_t num_requests, bool& done)
{
  done = false;
  Plan& pl = h->plan;
  const int64_t N = pl.num_rows;
  cudaStream_t s = h->stream;
  RepackParams rp;
  memset(&rp, 0, sizeof rp);
  rp.num_cols = num_keys; rp.num_rows = N;
  bool any_mask = false;
  for (int c = 0; c < num_keys; ++c) {
    const uint8_t src = type_src(keys[c].type_id);
    if (src == S_F32 || src == S_F64 || src == S_BAD) return BGB_OK;  // float keys keep their canonicalised bit images
    RepackCol& rc = rp.col[c];
    rc.data = (const uint8_t*)keys[c].data; rc.mask = keys[c].null_mask; rc.offset = keys[c].offset;
    rc.width = (uint8_t)type_width(keys[c].type_id);
    rc.is_signed = src == S_I8 || src == S_I16 || src == S_I32 || src == S_I64;
    rc.is_bool = src == S_BOOL;
    rc.null_bit = -1;
    any_mask |= keys[c].null_mask != nullptr;
  }
  uint64_t* d_mm = (uint64_t*)h->dalloc(sizeof(uint64_t) * 2 * num_keys);
  if (!d_mm) return set_error(BGB_OUT_OF_MEMORY, "repack scratch");
  std::vector<uint64_t> mm(2 * num_keys);
  for (int c = 0; c < num_keys; ++c) { mm[2 * c] = ~0ull; mm[2 * c + 1] = 0; }
  int rc = check_cuda(cudaMemcpyAsync(d_mm, mm.data(), sizeof(uint64_t) * mm.size(), cudaMemcpyHostToDevice, s), "repack init");
  rp.minmax = d_mm;
  h->prof.begin("repack_minmax");
  if (!rc) rc = check_cuda(launch_repack_minmax(rp, s), "repack_minmax");
  h->prof.end();
  if (!rc) rc = check_cuda(cudaMemcpyAsync(mm.data(), d_mm, sizeof(uint64_t) * mm.size(), cudaMemcpyDeviceToHost, s), "repack copy");
  if (!rc) rc = check_cuda(cudaStreamSynchronize(s), "repack sync");
  h->prof.resolve();
  h->dfree(d_mm);
  if (rc) return rc;
  // field widths: the order images run from mm[2c] to mm[2c+1] (an all-null column keeps the initial min > max: one bit)
  int bit = 0;
  std::vector<KeyPlan> ek = pl.keys;
  for (int c = 0; c < num_keys; ++c) {
    const uint64_t lo = mm[2 * c], hi = mm[2 * c + 1];
    const uint64_t range = lo <= hi ? hi - lo : 0;
    int bits = 1; while (bits < 64 && (range >> bits) != 0) ++bits;
    rp.col[c].base  = lo <= hi ? (rp.col[c].is_signed ? lo ^ 0x8000000000000000ull : lo) : 0;
    rp.col[c].shift = (uint8_t)(bit & 63);
    ek[c].shift = (uint8_t)(bit & 63); ek[c].base = rp.col[c].base; ek[c].bits = (uint8_t)bits; ek[c].null_bit = -1;
    bit += bits;
  }
  for (int c = 0; c < num_keys; ++c)
    if (ek[c].nullable_out) { if (bit < 64) { rp.col[c].null_bit = (int8_t)bit; ek[c].null_bit = (int8_t)bit; } ++bit; }
  if (bit > 64) return BGB_OK;  // the ranges do not fit one word either: 128-bit path
  if (pl.key_words == 2) {
    // Two passes over the key columns only pay when the packed key opens a path the 128-bit image cannot take. The product
    // of the column ranges bounds the group count: a handful of groups (tiny kernels), a shared-memory table for the
    // SUM/COUNT shape, a few thousand dense ids for the tile-sorting kernel, or a table that may exceed L2 on a large input (bucketed path instead of the partitioned one).
    double gmax = 1.0;
    for (int c = 0; c < num_keys; ++c) {
      const uint64_t lo = mm[2 * c], hi = mm[2 * c + 1];
      gmax *= (lo <= hi ? (double)(hi - lo) + 1.0 : 1.0) + (ek[c].nullable_out ? 1.0 : 0.0);
    }
    bool sum_count_shape = pl.values.size() <= 1;
    for (const AccPlan& a : pl.accs) sum_count_shape &= a.op == ACC_SUM_I64 || a.op == ACC_COUNT_VALID || a.op == ACC_COUNT_ALL;
    // (the tile-sorting kernel indexes by the packed image itself: `bit` bits of dense ids, any accumulator set)
    const bool sorted_fits = bit <= 20 && (1u << bit) <= sorted_max_groups(pl);
    const bool worthwhile = gmax <= 24.0 || (gmax <= 16384.0 && sum_count_shape) || sorted_fits || (N >= (1 << 24) && gmax >= 4194304.0);
    if (!worthwhile) return BGB_OK;
  }

  rp.drop_null_rows = pl.skip_null_keys;
  rp.out = (uint64_t*)h->dalloc((size_t)N * 8);
  if (!rp.out) return set_error(BGB_OUT_OF_MEMORY, "packed key column (%lld rows)", (long long)N);
  if (pl.skip_null_keys && any_mask) {
    rp.out_mask = (uint32_t*)h->dalloc(sizeof(uint32_t) * (size_t)((N + 31) / 32));
    if (!rp.out_mask) return set_error(BGB_OUT_OF_MEMORY, "packed key mask");
  }
  h->prof.begin("repack_pack");
  rc = check_cuda(launch_repack_pack(rp, s), "repack_pack");
  h->prof.end();
  if (rc) return rc;
  bgb_column_view pk{};
  pk.type_id = BGB_UINT64; pk.size = (int32_t)N; pk.data = rp.out; pk.null_mask = rp.out_mask; pk.null_count = rp.out_mask ? 1 : 0;
  Plan p2;
  rc = make_plan(&pk, 1, include_null_keys, requests, num_requests, p2, false);
  if (rc) return rc;
  h->plan        = p2;
  h->emit_keys   = ek;
  h->repack_buf  = rp.out;
  h->repack_mask = rp.out_mask;
  done = true;
  return BGB_OK;
}

static int build_local(const bgb_column_view* keys, int32_t num_keys, int32_t include_null_keys,
                       const bgb_request* requests, int32_t num_requests, void* stream, const bgb_allocator* alloc,
                       const bgb_options* options, bgb_groupby** out_handle, int32_t* num_groups, bool rank_independent = false,
                       const PeerFuse* fuse = nullptr)
{
  if (!out_handle || !num_groups) return set_error(BGB_INVALID_ARGUMENT, "null output pointer");
  *out_handle = nullptr;
  *num_groups = 0;
  bgb_groupby* h = new bgb_groupby();
  // key images wider than 128 bits are planned too when the call may repack them (the exact value ranges decide below)
  const bool may_repack = !rank_independent && (!options || options->force_path == 0);
  int rc = make_plan(keys, num_keys, include_null_keys, requests, num_requests, h->plan, rank_independent, /*allow_wide=*/!rank_independent);
  if (rc != BGB_OK) { delete h; return rc; }
  h->stream = (cudaStream_t)stream;
  if (alloc && alloc->allocate && alloc->deallocate) { h->alloc = *alloc; h->has_alloc = true; }
  bgb_options opt{};
  if (options) opt = *options;
  h->prof.on = opt.profile != 0;
  h->prof.stream = (cudaStream_t)stream;
  Plan& pl = h->plan;
  const int64_t N = pl.num_rows;
  if (N == 0) { *out_handle = h; return BGB_OK; }
  HostTrace tr;
  g_tr = &tr;
  tr.mark("plan");  // groupby.cu:233: empty input -> empty outputs
  cudaStream_t s = h->stream;

#define FAIL_RC(expr)            \
  do {                           \
    int rc_ = (expr);            \
    if (rc_ != BGB_OK) { h->free_all(); delete h; return rc_; } \
  } while (0)

  if (may_repack && (pl.key_words > 2 || (pl.key_words == 2 && N >= 65536))) {
    bool repacked = false;
    FAIL_RC(try_repack_keys(h, keys, num_keys, include_null_keys, requests, num_requests, repacked));
    tr.mark("repack");
  }
  if (pl.has_dict) {
    // dictionary values: the element of a row is keys[indices[row]]; only the generic kernels read values through that indirection
    // (the reference keeps dictionaries off its shared-memory path too, hash/compute_single_pass_aggs.cu:18-21)
    if (opt.force_path != 0 && opt.force_path != 1) FAIL_RC(set_error(BGB_UNSUPPORTED, "force_path %d does not serve dictionary values", opt.force_path));
    opt.force_path = 1;
  }
  if (pl.key_words > 2) {
    // no packed image: the table holds the row number of each group's first row and rows are compared in the caller's columns
    // (the reference's scheme for every key). One-word table of row numbers, generic kernels only.
    if (opt.force_path != 0 && opt.force_path != 1) FAIL_RC(set_error(BGB_UNSUPPORTED, "force_path %d needs a packed key image of at most 128 bits", opt.force_path));
    pl.wide         = true;
    pl.key_words    = 1;
    opt.force_path  = 1;
  }
  uint32_t* status = (uint32_t*)h->dalloc(16 * sizeof(uint32_t));
  if (!status) FAIL_RC(set_error(BGB_OUT_OF_MEMORY, "status alloc"));
  h->table.status    = status;
  h->table.key_words = pl.key_words;

  // ---- capacity: explicit, or from the sampled cardinality estimate (one small kernel + one read-back)
  // (group_slots are 32-bit and the reserved slot's index is `capacity`: never more than 2^31 slots, load <= 0.5 for N < 2^30,
  //  <= 1.0 beyond -- N < 2^31 always)
  const uint64_t cap_max = std::min<uint64_t>(1ull << 31, next_pow2((uint64_t)(2 * N < 1024 ? 1024 : 2 * N)));
  uint64_t capacity      = 0;
  bool speculated        = false;   // the estimate is a remembered one; the fresh sample is checked after the build
  uint64_t est_sig       = 0;
  double est_d1 = 0, est_m1 = 0;
  double g_est           = (double)N, g_raw = (double)N;
  uint64_t max_abs       = ~0ull, key_min = 1, key_max = 0;  // key_min > key_max: range not sampled
  int num_sms            = 148;
  { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev); }
  if (opt.capacity_log2 > 0) capacity = 1ull << opt.capacity_log2;
  else if (N <= 4096) capacity = next_pow2(2 * (uint64_t)N < 1024 ? 1024 : 2 * (uint64_t)N);
  else {
    const int64_t sample  = opt.sample_rows > 0 ? opt.sample_rows : 65536;
    const uint64_t setcap = next_pow2((uint64_t)sample * 4);
    uint64_t* set   = (uint64_t*)h->dalloc(setcap * sizeof(uint64_t));
    if (!set) FAIL_RC(set_error(BGB_OUT_OF_MEMORY, "estimator scratch"));
    AggParams ep;
    fill_agg_params(*h, ep, 0, 0);
    const SmemEligibility se = smem_eligibility(pl);
    if (se.ok && se.sum_acc >= 0 && !pl.has_dict) {  // also sample the magnitude of the SUM inputs
      const ValuePlan& v = pl.values[se.value_col];
      ep.num_accs = 1;
      ep.acc[0].data = (const uint8_t*)v.v.data; ep.acc[0].mask = v.v.null_mask; ep.acc[0].offset = v.v.offset; ep.acc[0].src = v.src;
    }
    const bool no_speculate = !speculation_enabled();
    est_sig = shape_signature(pl, opt);
    EstimateEntry remembered;
    if (!no_speculate) {
      std::lock_guard<std::mutex> lock(g_est_mu);
      const EstimateEntry& e = g_est_cache[est_sig & 63];
      if (e.valid && e.sig == est_sig) remembered = e;
    }
    h->prof.begin("estimate_cardinality");
    FAIL_RC(check_cuda(launch_estimate_cardinality(ep, sample, set, setcap, status + 8, se.key_flip, s), "estimate"));
    h->prof.end();
    double d = 0, m = 0;
    if (remembered.valid) {
      // speculate: the fresh sample lands in status[8..15] and is looked at after the build's own synchronisation
      speculated = true;
      est_d1 = remembered.d1; est_m1 = remembered.m1;
      d = remembered.d; m = remembered.m;
      max_abs = remembered.max_abs; key_min = remembered.key_min; key_max = remembered.key_max;
      h->dfree(set);  // stream-ordered: after the estimator kernel
      tr.mark("estimate (speculated)");
    } else {
    uint32_t est[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    FAIL_RC(check_cuda(cudaMemcpyAsync(est, status + 8, sizeof est, cudaMemcpyDeviceToHost, s), "estimate copy"));
    FAIL_RC(check_cuda(cudaStreamSynchronize(s), "estimate sync"));
    tr.mark("estimate+sync");
    h->prof.resolve();
    h->dfree(set);
    d = est[0]; m = est[1];
    est_d1 = d; est_m1 = m;
    max_abs = (uint64_t)est[2] | ((uint64_t)est[3] << 32);
    key_min = (uint64_t)est[4] | ((uint64_t)est[5] << 32);
    key_max = (uint64_t)est[6] | ((uint64_t)est[7] << 32);
    if (m > 0 && d > 0.9 * m && N > 8 * sample) {
      // nearly all sampled keys distinct: the small sample cannot tell 1e6 groups from 1e9. A 2M-row sample sees
      // m^2/(2G) duplicate pairs (2e4 at G = 1e8), enough to size the table instead of assuming one group per row.
      const int64_t sample2  = std::min<int64_t>(N, 1 << 21);
      const uint64_t setcap2 = next_pow2((uint64_t)sample2 * 4);
      uint64_t* set2 = (uint64_t*)h->dalloc(setcap2 * sizeof(uint64_t));
      if (set2) {
        ep.num_accs = 0;
        h->prof.begin("estimate_cardinality_2");
        FAIL_RC(check_cuda(launch_estimate_cardinality(ep, sample2, set2, setcap2, status + 8, se.key_flip, s), "estimate 2"));
        h->prof.end();
        FAIL_RC(check_cuda(cudaMemcpyAsync(est, status + 8, sizeof est, cudaMemcpyDeviceToHost, s), "estimate copy"));
        FAIL_RC(check_cuda(cudaStreamSynchronize(s), "estimate sync"));
        h->prof.resolve();
        h->dfree(set2);
        d = est[0]; m = est[1];
      }
    }
    if (!no_speculate) {
      std::lock_guard<std::mutex> lock(g_est_mu);
      EstimateEntry& e = g_est_cache[est_sig & 63];
      e.sig = est_sig; e.valid = true; e.d1 = est_d1; e.m1 = est_m1; e.d = d; e.m = m;
      e.max_abs = max_abs; e.key_min = key_min; e.key_max = key_max;
    }
    }
    double safety = 1.5;
    if (m == 0) g_raw = 1;
    else if (d >= m) g_raw = (double)N;  // not a single duplicate in the sample: cardinality unknown, assume the worst
    else {
      g_raw  = std::min((double)N, solve_cardinality(d, m));
      safety = std::max(1.25, std::min(3.0, 1.0 + 4.0 / std::sqrt(m - d)));  // few duplicates -> noisy estimate -> more headroom
    }
    g_est = std::min((double)N, safety * g_raw);
    capacity = next_pow2((uint64_t)std::max(1024.0, 2.0 * g_est));
    if (capacity > cap_max) capacity = cap_max;
  }

  bool skip_tiny = false, no_bucket = false, no_hc = false, no_short = false, no_pack32 = false;
  bool hc_shape = pl.key_words == 1 && pl.accs.size() <= (size_t)kMaxAccs && N >= 1 && N < (1ll << 31) && getenv("BGB_NO_HC") == nullptr;
  for (const AccPlan& a : pl.accs) hc_shape = hc_shape && !(a.op == ACC_ARGMIN || a.op == ACC_ARGMAX || a.op == ACC_ARGROW_MIN || a.op == ACC_ARGROW_MAX);
  for (int attempt = 0;; ++attempt) {
    h->table.capacity = capacity;
    size_t slot_bytes = 8 * (size_t)pl.key_words;
    for (const AccPlan& a : pl.accs) slot_bytes += acc_cell_bytes(a.op);
    // the table will not stay in L2: partition first (measured 3-5x faster than random HBM atomics at 1e8 groups)
    const bool big = (double)capacity * slot_bytes > 96e6 && N >= (1 << 22);
    // The records-form path writes its groups DENSELY: the arrays need one entry per possible group (at most one per row) plus the
    // reserved one, not one per hash slot. That matters when rows repeat little (a fold of partial tables: capacity = 4 x rows).
    // No other kernel may then touch the arrays: the small-cardinality paths are out of reach at > 70000 estimated groups, and a
    // records-form attempt that does not complete starts over with full arrays.
    const bool short_table = !no_short && !no_hc && !no_bucket && hc_shape && (uint64_t)N < capacity &&
                             (opt.force_path == 7 || (opt.force_path == 0 && big && g_raw > 70000.0));
    const uint64_t slots = short_table ? (uint64_t)N + 1 : capacity + 1;
    h->table_slots = slots;
    h->table.keys = (uint64_t*)h->dalloc(slots * 8 * pl.key_words);
    if (!h->table.keys) FAIL_RC(set_error(BGB_OUT_OF_MEMORY, "group table keys (%llu slots)", (unsigned long long)slots));
    for (AccPlan& a : pl.accs) {
      a.cells = h->dalloc(slots * acc_cell_bytes(a.op));
      if (!a.cells) FAIL_RC(set_error(BGB_OUT_OF_MEMORY, "group table accumulators"));
    }
    // init + aggregate, kMaxAccs accumulators per launch (the first launch also initialises the keys)
    TR("table alloc");
    std::string path;
    bool done = false, was_hc_attempt = false;
    BuildContext ctx{&pl, h->table, opt, g_raw, max_abs, key_min, key_max, s, &h->prof, num_sms, skip_tiny};
    if (!short_table) FAIL_RC(try_fast_paths(ctx, done, path));
    if (!done && opt.force_path != 1 && opt.force_path != 2 && opt.force_path != 4) {
      // Mid cardinalities (past the shared-memory table, table still L2-sized) stay on the generic kernel: ONE scatter level +
      // slices was measured SLOWER there (1e9 rows: 32 / 27 / 24 ms at 3e4 / 1e5 / 3e5 groups against 19-20 ms, profiles/
      // r01_card_sweep_A9_mid.txt) -- with 64-512 buckets and one CTA per bucket most of the 444 CTA slots stay empty. It would
      // need several CTAs per bucket and a keyed merge of their partial slices; BGB_BUCKET_MID=1 enables it for experiments.
      static const bool mid_on = getenv("BGB_BUCKET_MID") != nullptr;
      const bool mid = mid_on && !big && N >= (1 << 24) && capacity >= (1u << 16);
      if ((opt.force_path == 7 || (opt.force_path == 0 && (big || mid))) && !no_bucket) {
        if (!no_hc) { HcHints hints; hints.key_min = key_min; hints.key_max = key_max; hints.max_abs = max_abs; hints.no_pack32 = no_pack32; FAIL_RC(run_hc(h, num_sms, opt, g_raw, fuse, hints, done)); }  // records form first; plans it does not take go through the column form
        was_hc_attempt = done;
        if (!done && short_table) {  // the plan has no records form after all: full arrays, every path open
          no_short = true;
          h->dfree(h->table.keys); h->table.keys = nullptr;
          for (AccPlan& a : pl.accs) { h->dfree(a.cells); a.cells = nullptr; }
          --attempt;
          continue;
        }
        if (!done) FAIL_RC(run_bucketed(h, num_sms, opt, /*one_level_only=*/opt.force_path == 0 && !big, done));
        if (done) path = "bucketed";
      }
      if (!done && (opt.force_path == 3 || opt.force_path == 7 || big)) {
        FAIL_RC(run_partitioned(h, num_sms, done));
        if (done) path = "partitioned";
      }
    }
    if (!done) {
      size_t na = pl.accs.size();
      h->prof.begin("init_table");
      for (size_t b = 0; b == 0 || b < na; b += kMaxAccs) {
        AggParams p;
        fill_agg_params(*h, p, b, std::min(na, b + kMaxAccs));
        p.init_keys = (b == 0);  // later chunks only initialise their own cells
        FAIL_RC(check_cuda(launch_init_table(p, s), "init_table"));
      }
      h->prof.end();
      h->prof.begin("aggregate_global");
      for (size_t b = 0; b == 0 || b < na; b += kMaxAccs) {
        AggParams p;
        fill_agg_params(*h, p, b, std::min(na, b + kMaxAccs));
        FAIL_RC(check_cuda(launch_aggregate_global(p, s), "aggregate_global"));
      }
      h->prof.end();
      bool argrow = false;
      for (const AccPlan& a : pl.accs) argrow |= a.op == ACC_ARGROW_MIN || a.op == ACC_ARGROW_MAX;
      if (argrow) {  // ARGMIN / ARGMAX of 64-bit values: the extrema are final now, find their rows
        h->prof.begin("argrow");
        for (size_t b = 0; b == 0 || b < na; b += kMaxAccs) {
          AggParams p;
          fill_agg_params(*h, p, b, std::min(na, b + kMaxAccs));
          FAIL_RC(check_cuda(launch_argrow(p, s), "argrow"));
        }
        h->prof.end();
      }
      path = "global";
    }
    h->path = path;
    if (pl.wide) h->path += "+rowkeys";
    // compaction: occupied slots -> dense group list (+ group count)
    const uint64_t max_groups = std::min<uint64_t>(slots, (uint64_t)N);
    uint32_t* scan_state      = nullptr;
    if (!h->dense) {  // (the records form of the bucketed path wrote dense groups and their slot list itself)
      h->group_slots = (uint32_t*)h->dalloc(max_groups * sizeof(uint32_t));
      scan_state     = (uint32_t*)h->dalloc(compact_scan_state_bytes(capacity));
      if (!h->group_slots || !scan_state) FAIL_RC(set_error(BGB_OUT_OF_MEMORY, "compaction scratch"));
      TR("compact alloc");
      h->prof.begin("compact");
      FAIL_RC(check_cuda(launch_compact(h->table, h->group_slots, scan_state, s), "compact"));
      h->prof.end();
    }
    TR("compact launch");
    uint32_t st[16] = {0};
    FAIL_RC(check_cuda(cudaMemcpyAsync(st, status, sizeof st, cudaMemcpyDeviceToHost, s), "status copy"));
    tr.mark("launches");
    FAIL_RC(check_cuda(cudaStreamSynchronize(s), "build sync"));
    tr.mark("build sync");
    if (speculated) {
      // the fresh first-stage sample against the remembered one: distinct ratio, key range, value magnitude
      speculated = false;
      const double d1 = st[8], m1 = st[9];
      const uint64_t f_abs = (uint64_t)st[10] | ((uint64_t)st[11] << 32), f_min = (uint64_t)st[12] | ((uint64_t)st[13] << 32),
                     f_max = (uint64_t)st[14] | ((uint64_t)st[15] << 32);
      const double r_now = m1 > 0 ? d1 / m1 : 0.0, r_then = est_m1 > 0 ? est_d1 / est_m1 : 0.0;
      bool drift = std::fabs(r_now - r_then) > 0.1 * std::max(r_then, 0.02) || (m1 > 0) != (est_m1 > 0);
      drift |= f_abs > max_abs && max_abs != ~0ull;
      if (key_min <= key_max && f_min <= f_max) drift |= f_min + 64 < key_min || f_max > key_max + 64;
      else drift |= (key_min <= key_max) != (f_min <= f_max);
      if (drift) {
        std::lock_guard<std::mutex> lock(g_est_mu);
        EstimateEntry& e = g_est_cache[est_sig & 63];
        if (e.sig == est_sig) e.valid = false;
      }
    }
    h->prof.resolve();
    if (scan_state) h->dfree(scan_state);
    const bool was_fused = h->peer_fused, was_dense = h->dense;
    h->dense = false;  // (a retry below starts from scratch)
    if (getenv("BGB_DEBUG")) fprintf(stderr, "[bgb] path=%s groups=%u overflow=%u bypass_rows=%u flushes=%u\n", path.c_str(), st[1], st[0], st[4], st[5]);
    if (st[7] != 0 && path == "bucketed" && was_hc_attempt && !no_pack32) {
      // a key or a value did not fit the 32-bit halves of the packed record (the sample said they would): 16-byte records
      no_pack32 = true;
      h->peer_fused = false;
      h->dfree(h->group_slots); h->group_slots = nullptr;
      h->dfree(h->table.keys); h->table.keys = nullptr;
      for (AccPlan& a : pl.accs) { h->dfree(a.cells); a.cells = nullptr; }
      --attempt;
      continue;
    }
    if (st[6] != 0 && path == "bucketed") {
      // skewed keys filled a bucket of the records form (no overflow region there): column form, whose overflow region takes
      // moderate skew; if that fills too: same table size, exact-offset partitioned path
      if (was_hc_attempt) no_hc = true; else no_bucket = true;
      h->peer_fused = false;  // (records a fused attempt already wrote are dead: the cursors that count them are replaced)
      h->dfree(h->group_slots); h->group_slots = nullptr;
      h->dfree(h->table.keys); h->table.keys = nullptr;
      for (AccPlan& a : pl.accs) { h->dfree(a.cells); a.cells = nullptr; }
      --attempt;
      continue;
    }
    if (st[0] == 0 && st[3] == 0) {
      h->num_groups = (int32_t)st[1]; h->peer_fused = was_fused;
      h->dense_out = was_dense; h->reserved_used = st[2] != 0;
      break;
    }
    h->peer_fused = false;
    if (st[0] == 0 && st[3] != 0 && !skip_tiny) {
      // the tiny-cardinality kernel met more distinct keys than it has rows for: same table, general path
      if (opt.force_path == 5 || opt.force_path == 6) opt.force_path = 0;
      skip_tiny = true;
      h->dfree(h->group_slots); h->group_slots = nullptr;
      h->dfree(h->table.keys); h->table.keys = nullptr;
      for (AccPlan& a : pl.accs) { h->dfree(a.cells); a.cells = nullptr; }
      --attempt;
      continue;
    }
    // table overloaded: free it and retry at the size that fits every row as its own group
    if (capacity >= cap_max || attempt >= 2) FAIL_RC(set_error(BGB_CUDA_ERROR, "group table overflow at maximum capacity"));
    h->dfree(h->group_slots); h->group_slots = nullptr;
    h->dfree(h->table.keys); h->table.keys = nullptr;
    for (AccPlan& a : pl.accs) { h->dfree(a.cells); a.cells = nullptr; }
    capacity = cap_max;
    g_est    = (double)N;
    g_raw    = (double)N;
    key_min = 1; key_max = 0;
  }
#undef FAIL_RC
  if (h->repack_buf) { h->dfree(h->repack_buf); h->repack_buf = nullptr; }   // stream-ordered: the kernels above are done with it
  if (h->repack_mask) { h->dfree(h->repack_mask); h->repack_mask = nullptr; }
  if (!h->emit_keys.empty()) h->path += "+repack";
  tr.mark("tail");
  tr.flush("build_local");
  g_tr = nullptr;
  *num_groups = h->num_groups;
  *out_handle = h;
  return BGB_OK;
}

extern "C" int bgb_groupby_build(const bgb_column_view* keys, int32_t num_keys, int32_t include_null_keys,
                                 const bgb_request* requests, int32_t num_requests, void* stream, const bgb_allocator* alloc,
                                 const bgb_options* options, bgb_groupby** out_handle, int32_t* num_groups)
{
  NvtxRange nvtx_call("bgb_groupby_build");
  return build_local(keys, num_keys, include_null_keys, requests, num_requests, stream, alloc, options, out_handle, num_groups);
}

extern "C" int bgb_groupby_result_info(const bgb_groupby* h, int32_t request, int32_t aggregation, int32_t* type_id,
                                       int32_t* scale, int32_t* nullable)
{
  if (!h) return set_error(BGB_INVALID_ARGUMENT, "null handle");
  const Plan& pl = h->plan;
  if (request < 0 || request >= (int)pl.request_first_out.size() || aggregation < 0 || aggregation >= pl.request_num_aggs[request])
    return set_error(BGB_INVALID_ARGUMENT, "no result (%d,%d)", request, aggregation);
  const OutPlan& o = pl.outs[pl.request_first_out[request] + aggregation];
  if (type_id) *type_id = o.type_id;
  if (scale) *scale = o.scale;
  if (nullable) *nullable = o.nullable;
  return BGB_OK;
}

extern "C" int bgb_groupby_key_info(const bgb_groupby* h, int32_t key, int32_t* type_id, int32_t* scale, int32_t* nullable)
{
  if (!h) return set_error(BGB_INVALID_ARGUMENT, "null handle");
  const std::vector<KeyPlan>& ek = !h->emit_keys.empty() ? h->emit_keys : h->plan.keys;
  if (key < 0 || key >= (int)ek.size()) return set_error(BGB_INVALID_ARGUMENT, "no key column %d", key);
  const KeyPlan& k = ek[key];
  if (type_id) *type_id = k.v.type_id;
  if (scale) *scale = k.v.scale;
  if (nullable) *nullable = k.nullable_out;
  return BGB_OK;
}

extern "C" const char* bgb_groupby_path(const bgb_groupby* h) { return h ? h->path.c_str() : ""; }

extern "C" int32_t bgb_set_estimate_speculation(int32_t enabled)
{
  const bool was = speculation_enabled();
  g_speculate.store(enabled ? 1 : 0, std::memory_order_relaxed);
  {
    std::lock_guard<std::mutex> lock(g_est_mu);
    for (EstimateEntry& e : g_est_cache) e.valid = false;  // (a toggle also forgets every remembered estimate)
  }
  return was ? 1 : 0;
}

extern "C" int bgb_groupby_emit(bgb_groupby* h, bgb_mutable_column* out_keys, bgb_mutable_column* out_results, void* stream)
{
  NvtxRange nvtx_call("bgb_groupby_emit");
  if (!h) return set_error(BGB_INVALID_ARGUMENT, "null handle");
  const Plan& pl = h->plan;
  const std::vector<KeyPlan>& ekeys = h->emit_keys.empty() ? pl.keys : h->emit_keys;  // repacked: the caller's columns
  const int nk = (int)ekeys.size(), nr = (int)pl.outs.size();
  if ((nk > 0 && !out_keys) || (nr > 0 && !out_results)) return set_error(BGB_INVALID_ARGUMENT, "null output arrays");
  for (int i = 0; i < nk; ++i) out_keys[i].null_count = 0;
  for (int i = 0; i < nr; ++i) out_results[i].null_count = 0;
  const int32_t G = h->num_groups;
  if (G == 0) return BGB_OK;
  cudaStream_t s = (cudaStream_t)stream;
  h->stream      = s;
  h->prof.stream = s;

  // validate caller buffers
  for (int i = 0; i < nk; ++i) {
    if (out_keys[i].size < G || !out_keys[i].data) return set_error(BGB_INVALID_ARGUMENT, "out_keys[%d] too small", i);
    if (ekeys[i].nullable_out && !out_keys[i].null_mask) return set_error(BGB_INVALID_ARGUMENT, "out_keys[%d] needs a null mask", i);
  }
  for (int i = 0; i < nr; ++i) {
    if (out_results[i].size < G || !out_results[i].data) return set_error(BGB_INVALID_ARGUMENT, "out_results[%d] too small", i);
    if (pl.outs[i].op == OUT_SUMOV_SUM) {  // STRUCT result: data = the caller's two child descriptors {sum, overflow}
      bgb_mutable_column* ch = (bgb_mutable_column*)out_results[i].data;
      for (int c = 0; c < 2; ++c) {
        if (ch[c].size < G || !ch[c].data) return set_error(BGB_INVALID_ARGUMENT, "out_results[%d]: struct child %d too small", i, c);
        ch[c].null_count = 0;
      }
    }
    if (pl.outs[i].nullable && !out_results[i].null_mask) return set_error(BGB_INVALID_ARGUMENT, "out_results[%d] needs a null mask", i);
  }
  // Dense handles whose every output is a plain copy of a cell array (the SUM / COUNT shapes over one 8-byte key): device-to-device
  // copies instead of the gather kernel, whose one-group-per-thread loop keeps ~10 bytes per thread in flight (2.5 TB/s of traffic: 1.75 ms
  // for 1e8 groups; the copies move the same 4.4 GB in ~0.7 ms)
  {
    bool plain = h->dense_out && !pl.wide && h->emit_keys.empty() && nk == 1 && pl.key_words == 1;
    if (plain) {
      const KeyPlan& k = ekeys[0];
      plain = k.width == 8 && k.shift == 0 && k.null_bit < 0 && !k.nullable_out && k.bits == 0 && k.fclass == 0;
    }
    for (int i = 0; i < nr && plain; ++i) {
      const OutPlan& o = pl.outs[i];
      const bool copy8 = (o.op == OUT_COPY_I64 || o.op == OUT_COPY_F64) && o.out_width == 8 && acc_cell_bytes(pl.accs[o.acc_a].op) == 8;
      const bool copy4 = o.op == OUT_COUNT && o.out_width == 4;
      plain = !o.nullable && (copy8 || copy4);
    }
    if (plain) {
      const size_t body = (size_t)G - (h->reserved_used ? 1 : 0), last = h->table_slots - 1;
      h->prof.begin("emit_copy");
      auto copy = [&](void* dst, const void* src, size_t w) -> cudaError_t {
        cudaError_t e = body ? cudaMemcpyAsync(dst, src, body * w, cudaMemcpyDeviceToDevice, s) : cudaSuccess;
        if (e == cudaSuccess && h->reserved_used)
          e = cudaMemcpyAsync((uint8_t*)dst + body * w, (const uint8_t*)src + last * w, w, cudaMemcpyDeviceToDevice, s);
        return e;
      };
      cudaError_t e = copy(out_keys[0].data, h->table.keys, 8);
      for (int i = 0; i < nr && e == cudaSuccess; ++i) {
        const OutPlan& o = pl.outs[i];
        e = copy(out_results[i].data, pl.accs[o.acc_a].cells, o.out_width);
      }
      h->prof.end();
      if (h->prof.on && e == cudaSuccess) {  // (profiled calls read the event times: they need the stream drained, like the kernel form)
        e = cudaStreamSynchronize(s);
        h->prof.resolve();
      }
      return check_cuda(e, "emit copies");
    }
  }
  const int total = nk + nr;
  int32_t* d_nulls = (int32_t*)h->dalloc(sizeof(int32_t) * total);
  if (!d_nulls) return set_error(BGB_OUT_OF_MEMORY, "null counters");
  int rc = check_cuda(cudaMemsetAsync(d_nulls, 0, sizeof(int32_t) * total, s), "memset");
  if (rc) return rc;
  bool any_nullable = false;

  std::vector<OutDesc> descs;
  for (int i = 0; i < nk; ++i) {
    const KeyPlan& k = ekeys[i];
    OutDesc d{};
    d.out = out_keys[i].data; d.out_mask = k.nullable_out ? out_keys[i].null_mask : nullptr; d.null_count = d_nulls + i;
    d.op = OUT_KEY; d.src = type_src(k.v.type_id); d.out_width = k.width; d.shift = k.shift; d.null_bit = k.null_bit;
    d.key_bits = k.bits; d.key_base = k.base;
    if (pl.wide) {  // gather from the caller's key column at the group's representative row (the columns must still be alive)
      d.op = OUT_KEY_GATHER; d.a = k.v.data; d.b = k.nullable_out ? k.v.null_mask : nullptr; d.key_base = (uint64_t)(int64_t)k.v.offset; d.acc_op = k.fclass;
    }
    any_nullable |= k.nullable_out;
    descs.push_back(d);
  }
  for (int i = 0; i < nr; ++i) {
    const OutPlan& o = pl.outs[i];
    OutDesc d{};
    d.out = out_results[i].data; d.out_mask = o.nullable ? out_results[i].null_mask : nullptr; d.null_count = d_nulls + nk + i;
    d.a   = o.acc_a >= 0 ? pl.accs[o.acc_a].cells : nullptr;
    d.b   = o.acc_b >= 0 ? pl.accs[o.acc_b].cells : nullptr;
    d.cnt = o.acc_cnt >= 0 ? (const int32_t*)pl.accs[o.acc_cnt].cells : nullptr;
    d.ddof = o.ddof; d.op = o.op; d.src = o.src; d.out_width = o.out_width; d.acc_op = o.acc_op; d.null_bit = -1;
    any_nullable |= o.nullable;
    if (o.op == OUT_SUMOV_SUM) {  // two children: the sum (it also writes the struct's mask) and the BOOL8 overflow flag
      const bgb_mutable_column* ch = (const bgb_mutable_column*)out_results[i].data;
      d.out = ch[0].data;
      descs.push_back(d);
      d.op = OUT_SUMOV_FLAG; d.out = ch[1].data; d.out_mask = nullptr; d.null_count = nullptr; d.shift = o.out_width; d.out_width = 1;
    }
    descs.push_back(d);
  }
  h->prof.begin("emit");
  for (size_t b = 0; b < descs.size(); b += kMaxOuts) {
    EmitParams ep;
    memset(&ep, 0, sizeof ep);
    ep.keys = h->table.keys; ep.group_slots = h->group_slots; ep.num_groups = G; ep.key_words = pl.key_words;
    ep.dense = h->dense_out ? 1 : 0; ep.last_slot = (h->dense_out && h->reserved_used) ? (uint32_t)(h->table_slots - 1) : 0xFFFFFFFFu;
    ep.num_outs = (int)std::min<size_t>(kMaxOuts, descs.size() - b);
    for (int i = 0; i < ep.num_outs; ++i) ep.out[i] = descs[b + i];
    rc = check_cuda(launch_emit(ep, s), "emit");
    if (rc) return rc;
  }
  h->prof.end();
  if (h->prof.on && !any_nullable) {
    rc = check_cuda(cudaStreamSynchronize(s), "emit sync");
    if (rc) return rc;
  }
  if (any_nullable) {
    std::vector<int32_t> hn(total);
    rc = check_cuda(cudaMemcpyAsync(hn.data(), d_nulls, sizeof(int32_t) * total, cudaMemcpyDeviceToHost, s), "null counts");
    if (rc) return rc;
    rc = check_cuda(cudaStreamSynchronize(s), "emit sync");
    if (rc) return rc;
    for (int i = 0; i < nk; ++i) out_keys[i].null_count = hn[i];
    for (int i = 0; i < nr; ++i) out_results[i].null_count = hn[nk + i];
  }
  h->prof.resolve();
  h->dfree(d_nulls);
  return BGB_OK;
}

extern "C" int bgb_groupby_profile(const bgb_groupby* h, int32_t index, const char** name, float* ms)
{
  if (!h || index < 0 || index >= (int)h->prof.entries.size()) return BGB_INVALID_ARGUMENT;
  if (name) *name = h->prof.entries[index].name;
  if (ms) *ms = h->prof.entries[index].ms < 0 ? 0.f : h->prof.entries[index].ms;
  return BGB_OK;
}

extern "C" void bgb_groupby_destroy(bgb_groupby* h)
{
  if (!h) return;
  h->free_all();
  delete h;
}

// =====================================================================================================================
// Peer-memory multi-GPU build: local aggregation -> records scattered into the owners' regions over NVLink -> merge.
// =====================================================================================================================


extern "C" int bgb_peer_create(int32_t rank, int32_t world, int64_t region_bytes, bgb_peer** out)
{
  if (!out) return set_error(BGB_INVALID_ARGUMENT, "null output pointer");
  *out = nullptr;
  if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world) return set_error(BGB_INVALID_ARGUMENT, "rank %d / world %d (at most %d ranks)", rank, world, kMaxPeers);
  if (region_bytes < 4096) return set_error(BGB_INVALID_ARGUMENT, "region_bytes must be at least 4096");
  bgb_peer* p = new bgb_peer();
  p->rank = rank; p->world = world;
  p->region_bytes = ((size_t)region_bytes + 255) & ~(size_t)255;
  cudaGetDevice(&p->device);
  // plain cudaMalloc: the legacy IPC handle needs a cudaMalloc'ed base (pool memory cannot be exported this way)
  int rc = check_cuda(cudaMalloc((void**)&p->buffer, kPeerHeaderBytes + 2 * p->region_bytes), "peer buffer");
  if (!rc) rc = check_cuda(cudaMemset(p->buffer, 0, kPeerHeaderBytes), "peer header");
  if (!rc) rc = check_cuda(cudaDeviceSynchronize(), "peer header");
  if (rc) { if (p->buffer) cudaFree(p->buffer); delete p; return rc; }
  p->peer_buffer[rank] = p->buffer;
  if (const char* t = getenv("BGB_PEER_TIMEOUT_MS")) p->timeout_ns = (uint64_t)atoll(t) * 1000000ull;
  *out = p;
  return BGB_OK;
}

extern "C" int bgb_peer_export(const bgb_peer* p, void* handle_out)
{
  if (!p || !handle_out) return set_error(BGB_INVALID_ARGUMENT, "null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == BGB_PEER_HANDLE_BYTES, "handle size");
  cudaIpcMemHandle_t h;
  int rc = check_cuda(cudaIpcGetMemHandle(&h, p->buffer), "cudaIpcGetMemHandle");
  if (!rc) memcpy(handle_out, &h, sizeof h);
  return rc;
}

extern "C" int bgb_peer_attach(bgb_peer* p, int32_t peer_rank, const void* handle)
{
  if (!p || !handle || peer_rank < 0 || peer_rank >= p->world) return set_error(BGB_INVALID_ARGUMENT, "bad peer rank %d", peer_rank);
  if (peer_rank == p->rank) return BGB_OK;
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof h);
  void* ptr = nullptr;
  int rc = check_cuda(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess), "cudaIpcOpenMemHandle");
  if (rc) return rc;
  p->peer_buffer[peer_rank] = (uint8_t*)ptr;
  p->opened[peer_rank] = true;
  return BGB_OK;
}

extern "C" int bgb_peer_attach_local(bgb_peer* p, int32_t peer_rank, const bgb_peer* other)
{
  if (!p || !other || peer_rank < 0 || peer_rank >= p->world) return set_error(BGB_INVALID_ARGUMENT, "bad peer rank %d", peer_rank);
  if (other->region_bytes != p->region_bytes || other->world != p->world) return set_error(BGB_INVALID_ARGUMENT, "peer contexts differ in shape");
  if (other->device != p->device) {
    int can = 0;
    cudaDeviceCanAccessPeer(&can, p->device, other->device);
    if (!can) return set_error(BGB_UNSUPPORTED, "device %d cannot access device %d", p->device, other->device);
    cudaError_t e = cudaDeviceEnablePeerAccess(other->device, 0);
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return check_cuda(e, "cudaDeviceEnablePeerAccess");
    cudaGetLastError();
  }
  p->peer_buffer[peer_rank] = other->buffer;
  p->same_process = true;
  return BGB_OK;
}

extern "C" void bgb_peer_destroy(bgb_peer* p)
{
  if (!p) return;
  for (int r = 0; r < p->world; ++r)
    if (p->opened[r]) cudaIpcCloseMemHandle(p->peer_buffer[r]);
  if (p->buffer) cudaFree(p->buffer);
  delete p;
}

extern "C" int64_t bgb_peer_segment_capacity(const bgb_peer* p, int32_t record_bytes)
{
  if (!p || record_bytes <= 0) return 0;
  return (int64_t)((p->region_bytes / ((size_t)p->world * (size_t)record_bytes)) & ~(size_t)31);
}

extern "C" int bgb_groupby_build_peer(bgb_peer* peer, const bgb_column_view* keys, int32_t num_keys, int32_t include_null_keys,
                                      const bgb_request* requests, int32_t num_requests, void* stream, const bgb_allocator* alloc,
                                      const bgb_options* options, bgb_groupby** out_handle, int32_t* num_groups)
{
  NvtxRange nvtx_call("bgb_groupby_build_peer");
  if (!peer) return set_error(BGB_INVALID_ARGUMENT, "null peer context");
  for (int r = 0; r < peer->world; ++r)
    if (!peer->peer_buffer[r]) return set_error(BGB_INVALID_ARGUMENT, "rank %d is not attached", r);
  if (peer->poisoned) return set_error(BGB_CUDA_ERROR, "this peer context is out of step with its peers (an earlier exchange timed out): create a new one");
  if (!out_handle || !num_groups) return set_error(BGB_INVALID_ARGUMENT, "null output pointer");
  bgb_groupby* h = nullptr;
  int32_t g_local = 0;
  // The step's epoch is taken BEFORE anything that can fail, and a rank that fails before it has signalled still signals --
  // with the abort word -- so every rank leaves the step at the same epoch and nobody waits for the time-out.
  const int world = peer->world, me = peer->rank;
  const uint32_t epoch = ++peer->epoch;
  const int parity     = (int)(epoch & 1u);
  bool signalled       = false;
  auto signal_abort = [&]() {
    if (signalled) return;
    signalled = true;
    PeerSignalParams gp;
    memset(&gp, 0, sizeof gp);
    for (int r = 0; r < world; ++r) gp.header[r] = (PeerHeader*)peer->peer_buffer[r];
    gp.cursors = nullptr; gp.cap = 0; gp.world = world; gp.me = me; gp.parity = parity; gp.epoch = epoch; gp.abort = 1;
    if (launch_peer_signal(gp, (cudaStream_t)stream) != cudaSuccess || cudaStreamSynchronize((cudaStream_t)stream) != cudaSuccess) {
      cudaGetLastError();
      peer->poisoned = true;  // the peers will time out on this step
    }
  };
  // ---- (1) this rank's rows through the single-GPU path
  const PeerFuse fuse{peer, parity};
  int rc = build_local(keys, num_keys, include_null_keys, requests, num_requests, stream, alloc, options, &h, &g_local, /*rank_independent=*/true, &fuse);
  if (rc) { const std::string msg = bgb_last_error(); signal_abort(); return set_error(rc, "%s", msg.c_str()); }
  HostTrace tr;
  *out_handle = nullptr;
  *num_groups = 0;
  Plan& pl       = h->plan;
  cudaStream_t s = h->stream;
#define FAIL_RC(expr)            \
  do {                           \
    int rc_ = (expr);            \
    if (rc_ != BGB_OK) { const std::string msg_ = bgb_last_error(); signal_abort(); h->free_all(); delete h; return set_error(rc_, "%s", msg_.c_str()); } \
  } while (0)
  for (const AccPlan& a : pl.accs)
    if (a.op == ACC_ARGMIN || a.op == ACC_ARGMAX || a.op == ACC_ARGROW_MIN || a.op == ACC_ARGROW_MAX) FAIL_RC(set_error(BGB_UNSUPPORTED, "ARGMIN/ARGMAX have no cross-rank merge (row numbers are rank-local)"));
  if (pl.accs.size() > (size_t)kMaxAccs) FAIL_RC(set_error(BGB_UNSUPPORTED, "more than %d accumulators in one peer exchange", kMaxAccs));

  // ---- record layout inside a region: key word arrays, then one array per accumulator, each world * cap entries
  PeerLayout playout;
  if (!peer_layout(pl, peer, playout)) FAIL_RC(set_error(BGB_INVALID_ARGUMENT, "peer region too small for %d ranks of %zu-byte records", world, playout.record_bytes));
  const uint64_t cap = playout.cap, key1_off = playout.key1_off;
  const std::vector<uint64_t>& acc_off = playout.acc_off;

  uint32_t* scratch = (uint32_t*)h->dalloc(sizeof(uint32_t) * (3 * kMaxPeers + 4));  // cursors | wait output
  if (!scratch) FAIL_RC(set_error(BGB_OUT_OF_MEMORY, "peer scratch"));
  uint32_t* cursors  = h->peer_fused ? h->peer_cursors : scratch;   // fused: the slices of the local build already counted (and wrote) the records
  uint32_t* wait_out = scratch + kMaxPeers;
  if (!h->peer_fused) FAIL_RC(check_cuda(cudaMemsetAsync(cursors, 0, sizeof(uint32_t) * kMaxPeers, s), "peer cursors"));

  // ---- (2) scatter the local groups into their owners' regions, then publish counts + epoch
  h->prof.begin(h->peer_fused ? "peer_signal" : "peer_scatter");
  if (g_local > 0 && !h->peer_fused) {
    PeerScatterParams sp;
    memset(&sp, 0, sizeof sp);
    sp.keys = h->table.keys; sp.group_slots = h->group_slots; sp.num_groups = g_local;
    sp.key_words = pl.key_words; sp.num_accs = (int)pl.accs.size(); sp.world = world; sp.me = me;
    sp.cap = cap; sp.key1_off = key1_off; sp.cursors = cursors;
    for (int r = 0; r < world; ++r) sp.region[r] = peer->peer_buffer[r] + kPeerHeaderBytes + (size_t)parity * peer->region_bytes;
    for (size_t a = 0; a < pl.accs.size(); ++a) {
      sp.acc[a].cells = pl.accs[a].cells; sp.acc[a].region_off = acc_off[a]; sp.acc[a].op = pl.accs[a].op;
      sp.acc[a].cell_bytes = (uint8_t)acc_cell_bytes(pl.accs[a].op);
    }
    FAIL_RC(check_cuda(launch_peer_scatter(sp, s), "peer_scatter"));
  }
  {
    PeerSignalParams gp;
    memset(&gp, 0, sizeof gp);
    for (int r = 0; r < world; ++r) gp.header[r] = (PeerHeader*)peer->peer_buffer[r];
    gp.cursors = cursors; gp.cap = cap; gp.world = world; gp.me = me; gp.parity = parity; gp.epoch = epoch;
    FAIL_RC(check_cuda(launch_peer_signal(gp, s), "peer_signal"));
    signalled = true;  // from here on a failure is this rank's alone: its peers have what they need from it
  }
  h->prof.end();

  tr.mark("scatter+signal launches");
  // ---- (3) wait for every source, learn how many records arrived
  h->prof.begin("peer_wait");
  std::vector<uint32_t> got(2 * world + 1, 0);
  if (!peer->same_process) {
    // one process per GPU: a one-block kernel spins (bounded) on the flags, then hands counts to the host
    FAIL_RC(check_cuda(launch_peer_wait((const PeerHeader*)peer->buffer, parity, epoch, world, peer->timeout_ns, wait_out, s), "peer_wait"));
    h->prof.end();
    FAIL_RC(check_cuda(cudaMemcpyAsync(got.data(), wait_out, sizeof(uint32_t) * got.size(), cudaMemcpyDeviceToHost, s), "peer counts"));
    FAIL_RC(check_cuda(cudaStreamSynchronize(s), "peer wait sync"));
  } else {
    // ranks sharing a process (and possibly one GPU): a spinning kernel could sit in front of another rank's scatter in
    // the same hardware queue, so the HOST polls the flags instead; counts are read only after the flags were seen
    const PeerHeader* mine = (const PeerHeader*)peer->buffer;
    std::vector<uint32_t> flags(world, 0);
    const auto t0 = std::chrono::steady_clock::now();
    for (;;) {
      FAIL_RC(check_cuda(cudaMemcpyAsync(flags.data(), &mine->flag[parity][0], sizeof(uint32_t) * world, cudaMemcpyDeviceToHost, s), "peer flags"));
      FAIL_RC(check_cuda(cudaStreamSynchronize(s), "peer flag sync"));
      bool ready = true;
      for (int r = 0; r < world; ++r) ready &= (int32_t)(flags[r] - epoch) >= 0;
      if (ready) break;
      if ((uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t0).count() > peer->timeout_ns) { got[2 * world] = 1; break; }
      std::this_thread::yield();
    }
    h->prof.end();
    if (!got[2 * world]) {
      FAIL_RC(check_cuda(cudaMemcpyAsync(got.data(), &mine->count[parity][0], sizeof(uint32_t) * world, cudaMemcpyDeviceToHost, s), "peer counts"));
      FAIL_RC(check_cuda(cudaMemcpyAsync(got.data() + world, &mine->overflow[parity][0], sizeof(uint32_t) * world, cudaMemcpyDeviceToHost, s), "peer overflow"));
      FAIL_RC(check_cuda(cudaStreamSynchronize(s), "peer wait sync"));
    }
  }
  tr.mark("wait+sync");
  h->prof.resolve();
  if (got[2 * world]) { peer->poisoned = true; FAIL_RC(set_error(BGB_CUDA_ERROR, "peer exchange timed out waiting for another rank (epoch %u)", epoch)); }
  uint64_t total = 0;
  bool overflow  = false, aborted = false;
  for (int r = 0; r < world; ++r) { total += got[r]; overflow |= got[world + r] == 1; aborted |= got[world + r] == 2; }
  if (aborted) FAIL_RC(set_error(BGB_PEER_ABORTED, "another rank failed in this step of the peer exchange (epoch %u)", epoch));
  // the local partial table is no longer needed (its records are in the owners' regions)
  if (h->group_slots) { h->dfree(h->group_slots); h->group_slots = nullptr; }
  if (h->table.keys) { h->dfree(h->table.keys); h->table.keys = nullptr; }
  for (AccPlan& a : pl.accs) if (a.cells) { h->dfree(a.cells); a.cells = nullptr; }
  if (overflow) FAIL_RC(set_error(BGB_PEER_OVERFLOW, "a peer segment (%llu records) is too small for this exchange", (unsigned long long)cap));
  if (total > 0x7fffffffull) FAIL_RC(set_error(BGB_INVALID_ARGUMENT, "a rank would own more than 2^31-1 partial groups (cudf::size_type)"));

  // ---- (4) final table sized for what arrived; merge; compact
  uint32_t* status = (uint32_t*)h->dalloc(16 * sizeof(uint32_t));
  if (!status) FAIL_RC(set_error(BGB_OUT_OF_MEMORY, "status alloc"));
  h->table.status    = status;
  h->table.key_words = pl.key_words;
  // Table size of the merge. Every record could be a group of its own (shards with disjoint keys), and then 2 x records slots are
  // needed; when every rank saw every key (the usual case) there are `world` records per group and slices sized for the records
  // are 4-8x too sparse -- the slice kernel's per-bucket work (clearing and scanning 4096 slots) then outweighs its rows (5.05 ms
  // for 1e8 records at N=8). The distinct count is not known before the merge, so the size follows the LAST step's ratio of groups
  // to records (first step: the safe size); slices that fill anyway set the overflow flag and the merge reruns at the safe size.
  const uint64_t cap_safe = next_pow2(std::max<uint64_t>(1024, 2 * total));
  uint64_t max_got = 0;
  for (int r = 0; r < world; ++r) max_got = std::max<uint64_t>(max_got, got[r]);
  const double want_groups = std::max((double)max_got, peer->merge_ratio * 1.25 * (double)total);
  const uint64_t cap_adapt = std::min<uint64_t>(cap_safe, next_pow2((uint64_t)std::max(1024.0, 2.0 * want_groups)));
  bool adaptive     = cap_adapt < cap_safe;
  uint64_t capacity = cap_safe;
  // The received records are merged like rows of a groupby whose "values" are partial cells: through the records-form slices
  // when the shape fits (one key word, <= 3 accumulators, enough records to fill the machine), else with HBM atomics
  static const bool no_hc_merge = getenv("BGB_NO_HC_MERGE") != nullptr;
  uint64_t hc_merge_min = 1u << 22;  // below this the HBM-table merge is a few microseconds anyway
  if (const char* e = getenv("BGB_HC_MERGE_MIN")) hc_merge_min = (uint64_t)atoll(e);
  bool try_hc = !no_hc_merge && total >= hc_merge_min;
  uint32_t st[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  auto free_table = [&]() {
    if (h->table.keys) { h->dfree(h->table.keys); h->table.keys = nullptr; }
    for (AccPlan& a : pl.accs) if (a.cells) { h->dfree(a.cells); a.cells = nullptr; }
  };
  for (;;) {
    capacity          = (try_hc && adaptive) ? cap_adapt : cap_safe;
    h->table.capacity = capacity;
    // (the slices write the groups densely: one entry per received record at most, not one per hash slot)
    const uint64_t slots = try_hc ? total + 1 : capacity + 1;
    h->table_slots = slots;
    h->table.keys  = (uint64_t*)h->dalloc(slots * 8 * pl.key_words);
    if (!h->table.keys) FAIL_RC(set_error(BGB_OUT_OF_MEMORY, "final table keys"));
    for (AccPlan& a : pl.accs) {
      a.cells = h->dalloc(slots * acc_cell_bytes(a.op));
      if (!a.cells) FAIL_RC(set_error(BGB_OUT_OF_MEMORY, "final table accumulators"));
    }
    bool merged = false;
    if (try_hc) {
      std::vector<uint32_t> prefix(world + 1, 0);
      for (int r = 0; r < world; ++r) prefix[r + 1] = prefix[r] + got[r];
      h->prof.begin("peer_merge_hc");
      FAIL_RC(run_hc_merge(h, num_sms_of_current_device(), peer->buffer + kPeerHeaderBytes + (size_t)parity * peer->region_bytes, cap, world, prefix, acc_off,
                           total, merged));
      h->prof.end();
      if (!merged) { free_table(); try_hc = false; continue; }  // shape not taken: full-size arrays for the HBM-table merge
    }
    uint32_t* scan_state = nullptr;
    if (!merged) {
      h->prof.begin("peer_merge");
      size_t na = pl.accs.size();
      for (size_t b = 0; b == 0 || b < na; b += kMaxAccs) {
        AggParams p;
        fill_agg_params(*h, p, b, std::min(na, b + kMaxAccs));
        p.init_keys = (b == 0);
        FAIL_RC(check_cuda(launch_init_table(p, s), "init_table"));
      }
      PeerMergeParams mp;
      memset(&mp, 0, sizeof mp);
      mp.region = peer->buffer + kPeerHeaderBytes + (size_t)parity * peer->region_bytes;
      mp.cap = cap; mp.key1_off = key1_off; mp.key_words = pl.key_words; mp.num_accs = (int)na; mp.world = world;
      mp.slot_hash32 = pl.key_words == 1 && h->path == "bucketed";
      for (int r = 0; r < world; ++r) mp.prefix[r + 1] = mp.prefix[r] + got[r];
      mp.table = h->table;
      for (size_t a = 0; a < na; ++a) {
        mp.acc[a].cells = pl.accs[a].cells; mp.acc[a].region_off = acc_off[a]; mp.acc[a].op = pl.accs[a].op;
        mp.acc[a].cell_bytes = (uint8_t)acc_cell_bytes(pl.accs[a].op);
      }
      FAIL_RC(check_cuda(launch_peer_merge(mp, s), "peer_merge"));
      h->prof.end();
      const uint64_t max_groups = std::max<uint64_t>(1, std::min<uint64_t>(slots, total));
      h->group_slots            = (uint32_t*)h->dalloc(max_groups * sizeof(uint32_t));
      scan_state                = (uint32_t*)h->dalloc(compact_scan_state_bytes(capacity));
      if (!h->group_slots || !scan_state) FAIL_RC(set_error(BGB_OUT_OF_MEMORY, "compaction scratch"));
      h->prof.begin("compact_final");
      FAIL_RC(check_cuda(launch_compact(h->table, h->group_slots, scan_state, s), "compact"));
      h->prof.end();
    }
    FAIL_RC(check_cuda(cudaMemcpyAsync(st, status, sizeof st, cudaMemcpyDeviceToHost, s), "status copy"));
    tr.mark("merge launches");
    FAIL_RC(check_cuda(cudaStreamSynchronize(s), "peer build sync"));
    tr.mark("final sync");
    h->prof.resolve();
    if (scan_state) h->dfree(scan_state);
    if (merged && (st[0] || st[6])) {
      // slices or buckets of the merge filled: more groups per record than the last step had -> the safe size; skewed keys at the
      // safe size -> the HBM-table merge
      h->dfree(h->group_slots); h->group_slots = nullptr;
      free_table();
      if (adaptive) { adaptive = false; peer->merge_ratio = 1.0; }
      else try_hc = false;
      continue;
    }
    h->dense = merged;
    break;
  }
  h->dfree(scratch);
  if (st[0]) FAIL_RC(set_error(BGB_CUDA_ERROR, "final table overflow during the merge"));
#undef FAIL_RC
  tr.flush("build_peer");
  h->num_groups = (int32_t)st[1];
  h->dense_out = h->dense; h->reserved_used = st[2] != 0;
  if (total > 0) peer->merge_ratio = std::min(1.0, (double)st[1] / (double)total);
  h->path += h->peer_fused ? "+peer-fused" : "+peer";       // how the records travelled
  if (h->dense) h->path += "+slices";                        // how they were merged (else: HBM-table atomics + compaction)
  *num_groups = h->num_groups;
  *out_handle = h;
  return BGB_OK;
}

// =====================================================================================================================
// streaming_groupby (cpp/include/cudf/groupby.hpp:428-598): a persistent table fed batch by batch. Every batch runs
// through the stateless path above (so it takes whichever kernel fits: tiny, shared-memory, partitioned, generic) into
// a batch-local partial table, which is then folded into the persistent table with the cell-level merges of
// streaming_groupby/merge.cu:30-144. merge() folds another object's table the same way; finalize() compacts the
// persistent table and returns an ordinary handle for bgb_groupby_emit without touching the state.
// =====================================================================================================================

struct bgb_stream {
  std::vector<int32_t> key_indices;
  std::vector<bgb_stream_request> requests;
  int32_t max_keys = 0;
  int32_t include_null_keys = 0;
  bool initialized = false;
  Plan master;                    // layout of the persistent table (rank/batch independent: every value nullable)
  TableDesc table{};
  std::vector<void*> cells;       // one array per accumulator of `master`
  uint32_t* new_keys = nullptr;   // device counter of distinct keys
  int32_t distinct = 0;
  cudaStream_t last_stream = nullptr;
  bool invalidated = false;       // a fold failed part-way (distinct keys beyond the limit): the table holds a partial batch, so the
                                  // object refuses every further use, like the reference (streaming_groupby/aggregate.cu:26, merge.cu:93)
};

static bool stream_kind_supported(int32_t kind)
{
  switch (kind) {  // impl.cu:40-60: the hash kinds minus ARGMIN/ARGMAX (row indices are batch-local)
    case BGB_SUM: case BGB_PRODUCT: case BGB_MIN: case BGB_MAX: case BGB_COUNT_VALID: case BGB_COUNT_ALL: case BGB_SUM_OF_SQUARES:
    case BGB_MEAN: case BGB_M2: case BGB_VARIANCE: case BGB_STD: return true;
    default: return false;
  }
}

extern "C" int bgb_is_streaming_groupby_supported(int32_t values_type_id, int32_t kind, int32_t* out)
{
  if (!out) return set_error(BGB_INVALID_ARGUMENT, "null output pointer");
  *out = stream_kind_supported(kind) && type_width(values_type_id) > 0 && type_src(values_type_id) != S_BAD &&
         target_type(values_type_id, kind) >= 0;
  return BGB_OK;
}

extern "C" int bgb_stream_create(const int32_t* key_indices, int32_t num_keys, const bgb_stream_request* requests, int32_t num_requests,
                                 int32_t max_distinct_keys, int32_t include_null_keys, bgb_stream** out)
{
  if (!out) return set_error(BGB_INVALID_ARGUMENT, "null output pointer");
  *out = nullptr;
  if (max_distinct_keys <= 0) return set_error(BGB_INVALID_ARGUMENT, "max_distinct_keys must be positive");
  if (num_keys <= 0 || !key_indices) return set_error(BGB_INVALID_ARGUMENT, "streaming groupby needs at least one key column");
  if (num_requests < 0 || (num_requests > 0 && !requests)) return set_error(BGB_INVALID_ARGUMENT, "bad requests");
  for (int i = 0; i < num_requests; ++i)
    if (!stream_kind_supported(requests[i].kind))
      return set_error(BGB_INVALID_ARGUMENT, "Unsupported aggregation kind for streaming groupby. ARGMIN/ARGMAX are not supported because row indices are batch-local.");
  bgb_stream* s = new bgb_stream();
  s->key_indices.assign(key_indices, key_indices + num_keys);
  s->requests.assign(requests, requests + num_requests);
  s->max_keys = max_distinct_keys;
  s->include_null_keys = include_null_keys != 0;
  *out = s;
  return BGB_OK;
}

extern "C" void bgb_stream_destroy(bgb_stream* s)
{
  if (!s) return;
  if (s->table.keys) pool_free_async(s->table.keys, s->last_stream);
  if (s->table.status) pool_free_async(s->table.status, s->last_stream);
  for (void* c : s->cells) if (c) pool_free_async(c, s->last_stream);
  delete s;
}

static bool plans_compatible(const Plan& a, const Plan& b)
{
  if (a.key_words != b.key_words || a.keys.size() != b.keys.size() || a.accs.size() != b.accs.size()) return false;
  for (size_t i = 0; i < a.keys.size(); ++i)
    if (a.keys[i].v.type_id != b.keys[i].v.type_id || a.keys[i].shift != b.keys[i].shift || a.keys[i].null_bit != b.keys[i].null_bit) return false;
  for (size_t i = 0; i < a.accs.size(); ++i)
    if (a.accs[i].op != b.accs[i].op) return false;
  if (a.outs.size() != b.outs.size()) return false;
  for (size_t i = 0; i < a.outs.size(); ++i)
    if (a.outs[i].type_id != b.outs[i].type_id || a.outs[i].op != b.outs[i].op) return false;
  return true;
}

// folds the groups of (src table, src_slots[0..n)) into the persistent table; reads back the distinct-key counter
static int stream_fold(bgb_stream* st, const TableDesc& src, const uint32_t* src_slots, int32_t n, const std::vector<void*>& src_cells, cudaStream_t s)
{
  TableMergeParams mp;
  memset(&mp, 0, sizeof mp);
  mp.src_keys = src.keys; mp.src_group_slots = src_slots; mp.num_groups = n;
  mp.key_words = st->master.key_words; mp.num_accs = (int)st->master.accs.size();
  mp.dst = st->table; mp.new_keys = st->new_keys;
  for (size_t a = 0; a < st->master.accs.size(); ++a) {
    mp.src_acc[a].cells = src_cells[a]; mp.src_acc[a].op = st->master.accs[a].op; mp.src_acc[a].cell_bytes = (uint8_t)acc_cell_bytes(st->master.accs[a].op);
    mp.dst_cells[a] = st->cells[a];
  }
  int rc = check_cuda(launch_table_merge(mp, s), "table_merge");
  uint32_t host[2] = {0, 0};
  if (!rc) rc = check_cuda(cudaMemcpyAsync(&host[0], st->new_keys, sizeof(uint32_t), cudaMemcpyDeviceToHost, s), "distinct copy");
  if (!rc) rc = check_cuda(cudaMemcpyAsync(&host[1], st->table.status, sizeof(uint32_t), cudaMemcpyDeviceToHost, s), "status copy");
  if (!rc) rc = check_cuda(cudaStreamSynchronize(s), "stream fold sync");
  if (rc) { st->invalidated = true; return rc; }
  st->distinct = (int32_t)host[0];
  if (host[1] != 0 || host[0] > (uint32_t)st->max_keys) {
    st->invalidated = true;
    return set_error(BGB_INVALID_ARGUMENT, "streaming groupby: cumulative distinct keys (%u) exceed max_distinct_keys (%d)", host[0], st->max_keys);
  }
  return BGB_OK;
}

extern "C" int bgb_stream_aggregate(bgb_stream* st, const bgb_column_view* columns, int32_t num_columns, void* stream, const bgb_options* options)
{
  NvtxRange nvtx_call("bgb_stream_aggregate");
  if (!st || (num_columns > 0 && !columns)) return set_error(BGB_INVALID_ARGUMENT, "null argument");
  if (st->invalidated) return set_error(BGB_INVALID_ARGUMENT, "streaming groupby: the object was invalidated by a failed aggregate() / merge() and cannot be used");
  std::vector<bgb_column_view> keys;
  for (int32_t k : st->key_indices) {
    if (k < 0 || k >= num_columns) return set_error(BGB_INVALID_ARGUMENT, "key index %d out of range (%d columns)", k, num_columns);
    keys.push_back(columns[k]);
  }
  std::vector<bgb_aggregation> aggs(st->requests.size());
  std::vector<bgb_request> reqs(st->requests.size());
  for (size_t i = 0; i < st->requests.size(); ++i) {
    const bgb_stream_request& r = st->requests[i];
    if (r.column_index < 0 || r.column_index >= num_columns) return set_error(BGB_INVALID_ARGUMENT, "value column index %d out of range", r.column_index);
    aggs[i] = bgb_aggregation{r.kind, r.ddof};
    reqs[i] = bgb_request{columns[r.column_index], &aggs[i], 1, 0};
  }
  const int64_t rows = keys[0].size;
  if (rows > st->max_keys) return set_error(BGB_INVALID_ARGUMENT, "streaming groupby: a batch of %lld rows exceeds max_distinct_keys (%d)", (long long)rows, st->max_keys);
  if (rows == 0) return BGB_OK;
  cudaStream_t s = (cudaStream_t)stream;
  bgb_groupby* hb = nullptr;
  int32_t gb = 0;
  int rc = build_local(keys.data(), (int32_t)keys.size(), st->include_null_keys, reqs.data(), (int32_t)reqs.size(), stream, nullptr, options, &hb, &gb,
                       /*rank_independent=*/true);
  if (rc) return rc;
  struct Guard { bgb_groupby* h; ~Guard() { if (h) { h->free_all(); delete h; } } } guard{hb};
  if (!st->initialized) {
    st->master = hb->plan;
    const uint64_t capacity = next_pow2(std::max<uint64_t>(1024, 2 * (uint64_t)st->max_keys));
    const uint64_t slots    = capacity + 1;
    st->table.capacity  = capacity;
    st->table.key_words = st->master.key_words;
    st->table.keys   = (uint64_t*)pool_alloc_async(slots * 8 * st->master.key_words, s);
    st->table.status = (uint32_t*)pool_alloc_async(17 * sizeof(uint32_t), s);
    bool ok = st->table.keys && st->table.status;
    st->cells.assign(st->master.accs.size(), nullptr);
    for (size_t a = 0; a < st->master.accs.size() && ok; ++a) {
      st->cells[a] = pool_alloc_async(slots * acc_cell_bytes(st->master.accs[a].op), s);
      ok = st->cells[a] != nullptr;
    }
    st->last_stream = s;
    if (!ok) return set_error(BGB_OUT_OF_MEMORY, "persistent table for %d distinct keys", st->max_keys);
    st->new_keys = st->table.status + 16;
    rc = check_cuda(cudaMemsetAsync(st->table.status, 0, 17 * sizeof(uint32_t), s), "stream status");
    const size_t na = st->master.accs.size();
    for (size_t b = 0; (b == 0 || b < na) && !rc; b += kMaxAccs) {
      AggParams p;
      memset(&p, 0, sizeof p);
      p.table = st->table; p.init_keys = (b == 0);
      p.num_accs = (int)(std::min(na, b + kMaxAccs) - b);
      for (int a = 0; a < p.num_accs; ++a) { p.acc[a].op = st->master.accs[b + a].op; p.acc[a].cells = st->cells[b + a]; }
      rc = check_cuda(launch_init_table(p, s), "init_table");
    }
    if (rc) return rc;
    st->initialized = true;
  } else if (!plans_compatible(st->master, hb->plan)) {
    return set_error(BGB_INVALID_ARGUMENT, "streaming groupby: this batch's column types differ from the first batch's");
  }
  st->last_stream = s;
  if (st->master.accs.size() > (size_t)kMaxAccs) return set_error(BGB_UNSUPPORTED, "more than %d accumulators in a streaming groupby", kMaxAccs);
  std::vector<void*> src_cells;
  for (const AccPlan& a : hb->plan.accs) src_cells.push_back(a.cells);
  return stream_fold(st, hb->table, hb->group_slots, gb, src_cells, s);
}

extern "C" int bgb_stream_merge(bgb_stream* st, const bgb_stream* other, void* stream)
{
  NvtxRange nvtx_call("bgb_stream_merge");
  if (!st || !other) return set_error(BGB_INVALID_ARGUMENT, "null argument");
  if (st->invalidated) return set_error(BGB_INVALID_ARGUMENT, "streaming groupby: the object was invalidated by a failed aggregate() / merge() and cannot be used");
  if (other->invalidated) return set_error(BGB_INVALID_ARGUMENT, "streaming groupby: the other object was invalidated by a failed aggregate() / merge()");
  if (!st->initialized) return set_error(BGB_INVALID_ARGUMENT, "streaming groupby: merge() needs an object that has seen at least one aggregate()");
  if (!other->initialized || other->distinct == 0) return BGB_OK;
  if (!plans_compatible(st->master, other->master)) return set_error(BGB_INVALID_ARGUMENT, "streaming groupby: the two objects have different aggregation requests or column types");
  if (other->distinct > st->max_keys) return set_error(BGB_INVALID_ARGUMENT, "streaming groupby: the other object holds %d distinct keys, more than max_distinct_keys (%d)", other->distinct, st->max_keys);
  cudaStream_t s = (cudaStream_t)stream;
  uint32_t* slots = (uint32_t*)pool_alloc_async(sizeof(uint32_t) * (size_t)(other->distinct + 1), s);
  uint32_t* scan  = (uint32_t*)pool_alloc_async(compact_scan_state_bytes(other->table.capacity), s);
  if (!slots || !scan) { if (slots) pool_free_async(slots, s); if (scan) pool_free_async(scan, s); return set_error(BGB_OUT_OF_MEMORY, "merge scratch"); }
  int rc = check_cuda(launch_compact(other->table, slots, scan, s), "compact");
  if (!rc) rc = stream_fold(st, other->table, slots, other->distinct, other->cells, s);
  pool_free_async(slots, s);
  pool_free_async(scan, s);
  st->last_stream = s;
  return rc;
}

extern "C" int bgb_stream_distinct_keys(const bgb_stream* st, int32_t* out)
{
  if (!st || !out) return set_error(BGB_INVALID_ARGUMENT, "null argument");
  *out = st->distinct;
  return BGB_OK;
}

extern "C" int bgb_stream_finalize(const bgb_stream* st, void* stream, bgb_groupby** out_handle, int32_t* num_groups)
{
  NvtxRange nvtx_call("bgb_stream_finalize");
  if (!st || !out_handle || !num_groups) return set_error(BGB_INVALID_ARGUMENT, "null argument");
  *out_handle = nullptr;
  *num_groups = 0;
  if (st->invalidated) return set_error(BGB_INVALID_ARGUMENT, "streaming groupby: the object was invalidated by a failed aggregate() / merge() and cannot be used");
  if (!st->initialized) return set_error(BGB_INVALID_ARGUMENT, "streaming groupby: no data has been accumulated");
  bgb_groupby* h = new bgb_groupby();
  h->plan   = st->master;
  h->stream = (cudaStream_t)stream;
  h->table  = st->table;  // borrowed: the handle owns only its compaction list, the persistent state is not modified
  for (size_t a = 0; a < h->plan.accs.size(); ++a) h->plan.accs[a].cells = st->cells[a];
  cudaStream_t s = h->stream;
  const uint64_t max_groups = (uint64_t)st->distinct + 1;
  h->group_slots       = (uint32_t*)h->dalloc(max_groups * sizeof(uint32_t));
  uint32_t* scan_state = (uint32_t*)h->dalloc(compact_scan_state_bytes(st->table.capacity));
  int rc = (!h->group_slots || !scan_state) ? set_error(BGB_OUT_OF_MEMORY, "compaction scratch") : BGB_OK;
  if (!rc) rc = check_cuda(launch_compact(h->table, h->group_slots, scan_state, s), "compact");
  uint32_t count = 0;
  if (!rc) rc = check_cuda(cudaMemcpyAsync(&count, st->table.status + 1, sizeof count, cudaMemcpyDeviceToHost, s), "group count");
  if (!rc) rc = check_cuda(cudaStreamSynchronize(s), "finalize sync");
  if (rc) { h->free_all(); delete h; return rc; }
  h->dfree(scan_state);
  h->num_groups = (int32_t)count;
  h->path       = "stream";
  *num_groups   = h->num_groups;
  *out_handle   = h;
  return BGB_OK;
}

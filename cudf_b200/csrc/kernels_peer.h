// Peer-memory exchange of partial group tables (multi-GPU, one process per GPU): every rank writes the records of the
// groups another rank owns straight into that rank's receive region over NVLink (P2P stores), signals with a
// release-store of the step's epoch, and the owner merges the received records into its final table.
#pragma once
#include "bgb_internal.cuh"

namespace bgb {

constexpr int kMaxPeers = 64;

// header at the start of every rank's exchange buffer; slots are indexed [parity][source rank]
struct PeerHeader {
  uint32_t flag[2][kMaxPeers];      // epoch of the last completed scatter from that source
  uint32_t count[2][kMaxPeers];     // records that source wrote into its segment
  uint32_t overflow[2][kMaxPeers];  // 1: source had more records for this rank than a segment holds; 2: the source aborted the step
};
constexpr size_t kPeerHeaderBytes = 4096;
static_assert(sizeof(PeerHeader) <= kPeerHeaderBytes, "header must fit its page");

struct PeerAcc {
  void* cells;          // scatter: local partial table cells; merge: final table cells
  uint64_t region_off;  // byte offset of this accumulator's array inside a region
  uint8_t op;           // AccOp
  uint8_t cell_bytes;   // 4 or 8
  uint8_t pad[6];
};

struct PeerScatterParams {
  const uint64_t* keys;         // local table keys
  const uint32_t* group_slots;  // dense list of occupied slots
  int32_t num_groups;
  int32_t key_words, num_accs, world, me;
  uint64_t cap;                 // records per (source, destination) segment
  uint64_t key1_off;            // byte offset of the second key word array inside a region
  uint8_t* region[kMaxPeers];   // destination rank's region for this parity (peer-mapped)
  uint32_t* cursors;            // [world] records produced for each destination (local, zeroed)
  PeerAcc acc[kMaxAccs];
};

struct PeerSignalParams {
  PeerHeader* header[kMaxPeers];  // every rank's header (peer-mapped)
  const uint32_t* cursors;
  uint64_t cap;
  int32_t world, me, parity;
  uint32_t epoch;
  int32_t abort;                  // != 0: this rank failed before it could scatter: counts 0, overflow word 2 to every destination
};

struct PeerMergeParams {
  const uint8_t* region;  // my region for this parity
  uint64_t cap, key1_off;
  int32_t key_words, num_accs, world;
  int32_t slot_hash32;             // != 0 (one-word keys): the final table is indexed by slot_hash32(), the hash of the bucketed
                                   // local path, so that records arriving in local slot order sweep the final table in order
  uint32_t prefix[kMaxPeers + 1];  // records received from sources [0, s)
  TableDesc table;                 // final table
  PeerAcc acc[kMaxAccs];
};

// table -> table merge (streaming_groupby: a batch's partial table, or another object's state, folded into the persistent
// table; streaming_groupby/merge.cu:30-144): every group of `src` is found-or-inserted in `dst` and its cells merged
struct TableMergeParams {
  const uint64_t* src_keys;
  const uint32_t* src_group_slots;  // dense list of occupied source slots
  int32_t num_groups;
  int32_t key_words, num_accs;
  TableDesc dst;
  uint32_t* new_keys;               // device counter: groups that did not exist in dst before
  PeerAcc src_acc[kMaxAccs];        // .cells = source cells (region_off unused)
  void* dst_cells[kMaxAccs];
};
cudaError_t launch_table_merge(const TableMergeParams& p, cudaStream_t s);

cudaError_t launch_peer_scatter(const PeerScatterParams& p, cudaStream_t s);
cudaError_t launch_peer_signal(const PeerSignalParams& p, cudaStream_t s);
// waits until every source's flag reaches `epoch` (or timeout_ns passes: out[2 * world] = 1), then copies
// count[s] -> out[s] and overflow[s] -> out[world + s]
cudaError_t launch_peer_wait(const PeerHeader* mine, int parity, uint32_t epoch, int world, uint64_t timeout_ns, uint32_t* out, cudaStream_t s);
cudaError_t launch_peer_merge(const PeerMergeParams& p, cudaStream_t s);
// owner rank of a packed key (same function as bgb_partition_by_key_hash)
__host__ __device__ inline uint32_t owner_mix(uint64_t h, uint32_t world) { return (uint32_t)(mix64(h ^ 0xA0761D6478BD642Full) % world); }

}  // namespace bgb

// Parameters of the shared-memory pre-aggregation kernel (kernels_smem.cu).
#pragma once
#include "bgb_internal.cuh"

namespace bgb {

struct SmemParams {
  AggParams a;                    // key columns + HBM table (acc[] unused here)
  const uint8_t* v_data;          // value column of the SUM (nullptr when VW == 0)
  const uint32_t* v_mask;
  int32_t v_offset;
  int32_t v_signed;               // 4-byte values: sign-extend (int32) or zero-extend (uint32)
  unsigned long long* sum_cells;  // HBM accumulators (any may be nullptr)
  int* cntv_cells;
  int* cnta_cells;
  uint32_t slots;                 // 16-bit index slots per CTA (even)
  uint32_t dense;                 // dense group rows per CTA (< 65535); new keys beyond that bypass to the HBM table
  int32_t cnt_bits;               // c: low bits of shc[] hold the row count, the rest the high part of the sum
  int32_t direct;                 // 1: dense id = (key ^ key_flip) - key_base, no probing (small integer key domains)
  uint64_t key_flip;              // sign-bit flip that makes the packed key image order-preserving (0 for unsigned)
  uint64_t key_base;
  int32_t debug;                  // count bypassed rows / flushes in status[4], status[5]
  int32_t val_bits;               // m: values in [-2^m, 2^m) take the shared-memory path; c + m <= 63 - c
  int32_t tma_stages;             // set by the launcher: > 0 = rows arrive through a TMA ring of that many 2048-row stages
  uint32_t tma_offset;            // byte offset of the ring (barriers, then stages) in the dynamic shared memory
};

size_t smem_table_max_bytes();
size_t smem_table_bytes(uint32_t slots, uint32_t dense, int smem_key_bytes, bool direct);
// smem_key_bytes 4|8, key_mode 1 = single key column loaded directly (width == smem_key_bytes), 0 = generic pack_key,
// value_width 0 (no SUM) | 4 | 8
cudaError_t launch_aggregate_smem(const SmemParams& p, int smem_key_bytes, int key_mode, int value_width, int num_sms, cudaStream_t s);

}  // namespace bgb

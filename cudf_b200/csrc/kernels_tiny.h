// Parameters of the tiny-cardinality kernel (kernels_tiny.cu): a handful of groups, any mix of SUM / SUM_OF_SQUARES /
// COUNT / MIN / MAX accumulators, accumulated in per-thread private shared-memory cells (no atomics on the row path).
#pragma once
#include "bgb_internal.cuh"

namespace bgb {

constexpr int kTinyMaxCols = 4;   // distinct value columns (+ the pseudo column that carries COUNT_ALL) of ONE launch of the shared-memory kernel
constexpr int kTinyPlanCols = 8;  // ... of a plan: the register kernel takes a wider plan as several launches of <= 2 columns (tiny_reg_split)
constexpr int kTinyThreads = 512;

// one value column and everything accumulated from it. Cells (8 bytes each) of a group row, in this order:
//   [count of valid rows] [sum] [sum of squares, if need_sq] [min, max, if need_minmax]
struct TinyCol {
  const uint8_t* data;   // nullptr for the pseudo column that counts all rows
  const uint32_t* mask;
  int32_t offset;
  uint8_t src;           // Src
  uint8_t width;
  uint8_t need_sq;
  uint8_t need_minmax;
  int32_t cell_base;     // first cell of this column inside a group row
  int32_t acc_sum, acc_sumsq, acc_min, acc_max, acc_count;  // AggParams::acc[] entries fed at the end (-1: not requested)
  int32_t sum_op;        // AccOp of the sum / sum-of-squares pair (decides integer, float64 or float32 arithmetic)
};

struct TinyParams {
  AggParams a;           // key columns, HBM table, acc[] (global cells)
  TinyCol cols[kTinyPlanCols];
  int32_t num_cols;
  int32_t cells;         // 8-byte cells per group row
  int32_t group_cap;     // groups a CTA can hold; more distinct keys -> status[3] = 1 and the host takes another path
};

size_t tiny_smem_bytes(int cells, int group_cap);
cudaError_t launch_aggregate_tiny(const TinyParams& p, int num_sms, cudaStream_t s);

// register-resident variant (kernels_tiny_reg.cu): at most 8 groups, one or two 8-byte value columns, SUM/SUMSQ/COUNT
bool tiny_reg_eligible(const TinyParams& p);  // p.a (keys, table) must be filled
cudaError_t launch_aggregate_tiny_reg(const TinyParams& p, int group_cap /* 4 or 8 */, int num_sms, cudaStream_t s);
// A plan with more value columns than one launch takes (NDS-H Q1: five float64 columns + COUNT(*)), every column of which the
// register kernel serves: split into sub-plans of <= 2 columns of one arithmetic class, each eligible on its own. The launches
// feed the same HBM table (every launch learns the group keys itself and ends in find-or-insert + atomic adds), so the key
// columns are read once per launch: 2 B/row extra per launch on Q1 against 16 B/row of values. Returns the number of sub-plans
// written to out[] (0: not splittable).
int tiny_reg_split(const TinyParams& p, TinyParams* out /* kTinyPlanCols entries */);

}  // namespace bgb

// Specialised kernels in front of the generic HBM-table kernel. try_fast_paths() initialises the table and runs the
// whole aggregation when the plan matches one of them (done = true), else leaves everything to the generic path.
#pragma once
#include <string>

#include "plan.h"
#include "profiler.h"

namespace bgb {

struct BuildContext {
  const Plan* plan;
  TableDesc table;
  bgb_options opt;
  double g_raw;             // sampled cardinality estimate (no safety factor)
  uint64_t max_abs_value;   // largest |v| among the sampled rows of the SUM column (0 when not sampled)
  uint64_t key_min, key_max;  // sampled range of the order-preserving key image (min > max: not sampled)
  cudaStream_t stream;
  Profiler* prof;
  int num_sms;
  bool skip_tiny;           // the tiny-cardinality kernel met more groups than it holds: take another path
};

// which accumulator set the shared-memory kernel can serve; value_col = -1 when there is no SUM/COUNT_VALID column
struct SmemEligibility {
  bool ok = false;
  int value_col = -1, sum_acc = -1, cntv_acc = -1, cnta_acc = -1;
  int key_mode = 0, smem_key_bytes = 8, value_width = 0;
  bool integer_keys = false;   // every key column is an integer type: the packed key has a meaningful numeric range
  uint64_t key_flip = 0;       // sign bit of a single signed key column (makes the image order-preserving)
};
SmemEligibility smem_eligibility(const Plan& plan);

// tile-sorting kernel (kernels_sorted.cu): largest dense id range it holds for this plan's accumulators, 0 = plan not served
uint32_t sorted_max_groups(const Plan& plan, bool forced = false);

int try_fast_paths(const BuildContext& ctx, bool& done, std::string& path);
}

// Specialised kernels in front of the generic HBM-table kernel. try_fast_paths() initialises the table and runs the
// whole aggregation when the plan matches one of them (done = true), else leaves everything to the generic path.
#pragma once
#include <string>

#include "plan.h"

namespace bgb {
int try_fast_paths(const Plan& plan, const TableDesc& table, const bgb_options& opt, double g_est, cudaStream_t s, bool& done,
                   std::string& path);
}

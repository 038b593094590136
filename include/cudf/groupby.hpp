// Drop-in include path: sources written against libcudf's  #include <cudf/groupby.hpp>  and  cudf::groupby::groupby
// (/root/reference/cpp/include/cudf/groupby.hpp:89-184, streaming_groupby :428-598) compile against this library's host layer
// when THIS include directory comes first and libcudf's own headers are not on the path. The namespace alias makes
// cudf::groupby::groupby, cudf::groupby::aggregation_request, cudf::table_view, cudf::column_view, cudf::null_policy,
// cudf::make_sum_aggregation ... the cudf_b200 entities of include/cudf_b200/groupby.hpp. (A build that also links libcudf uses the
// shim of INTEGRATION.md section 2 instead: there the hash path is replaced inside libcudf and this header is not needed.)
#pragma once
#include "../cudf_b200/groupby.hpp"

namespace cudf = cudf_b200;  // NOLINT

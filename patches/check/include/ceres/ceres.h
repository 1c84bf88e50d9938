// TEST INFRASTRUCTURE ONLY (patches/check): the few Ceres names the reference's
// data_structures.h / slam_util.h mention, so that csm_batch.cc of the patch can be
// syntax-checked against the reference's real headers without Ceres.
#pragma once
#include <cmath>
#include <vector>
namespace ceres {
class Problem {};
typedef void* ResidualBlockId;
inline bool IsFinite(double v) { return std::isfinite(v); }
struct CRSMatrix { int num_rows = 0, num_cols = 0; std::vector<int> cols, rows; std::vector<double> values; };
}  // namespace ceres

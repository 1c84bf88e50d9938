// TEST INFRASTRUCTURE ONLY -- the reference's lidar residual functors,
//   LIDARNormalResidual  src/optimization/slam_residuals.h:64-121
//   LIDARPointResidual   src/optimization/slam_residuals.h:123-177
// compiled UNCHANGED from /root/reference and evaluated as the solver evaluates them,
// through ::create(...)->Evaluate(parameters, residuals, jacobians) of
// ceres::AutoDiffCostFunction<., DYNAMIC, 3, 3> (solver.cc:275-296).  Ceres 1.14.0 and
// Eigen 3.3.7 are un-vendored: the AutoDiff / Jet machinery and the Eigen types come
// from oracle/shim, which restates their published algorithms (see the headers there).
// Nothing under nautilus_b200/ links or loads this.
#include <cstdint>
#include <vector>

// slam_residuals.h is not self-contained: in the reference it is included by solver.cc
// after solver.h, which brings in slam_util.h (PoseArrayToAffine) and, through
// CorrelativeScanMatcher.h:23-24, the global `using Eigen::Vector2f; using std::vector;`.
#include "util/slam_util.h"
using Eigen::Vector2f;
using std::vector;
#include "optimization/slam_residuals.h"

static std::vector<Eigen::Vector2f> slice(const float* xy, int64_t a, int64_t b) {
  std::vector<Eigen::Vector2f> c;
  c.reserve((size_t)(b - a));
  for (int64_t i = a; i < b; ++i) c.push_back(Eigen::Vector2f(xy[2 * i], xy[2 * i + 1]));
  return c;
}

// kind 0: LIDARPointResidual, 1: LIDARNormalResidual.  Layout as nautilus_residuals_eval.
extern "C" int ref_residuals_eval(int kind, const double* source_poses, const double* target_poses,
                                  const int64_t* offsets, int32_t n_blocks, const float* sp,
                                  const float* tp, const float* sn, const float* tn,
                                  double* residuals, double* jac_source, double* jac_target) {
  for (int32_t b = 0; b < n_blocks; ++b) {
    const int64_t o = offsets[b], e = offsets[b + 1];
    if (e == o) continue;  // AddPointCorrespondence skips empty correspondences (solver.cc:277,287)
    const auto p = slice(sp, o, e), q = slice(tp, o, e), ns = slice(sn, o, e), nt = slice(tn, o, e);
    const double* params[2] = {source_poses + 3 * b, target_poses + 3 * b};
    double* jac[2] = {jac_source ? jac_source + 6 * o : nullptr, jac_target ? jac_target + 6 * o : nullptr};
    double** jp = (jac_source || jac_target) ? jac : nullptr;
    bool ok;
    if (kind == 0) {
      auto* f = nautilus::LIDARPointResidual::create(p, q, ns, nt);
      ok = f->Evaluate(params, residuals + 2 * o, jp);
      delete f;
    } else {
      auto* f = nautilus::LIDARNormalResidual::create(p, q, ns, nt);
      ok = f->Evaluate(params, residuals + 2 * o, jp);
      delete f;
    }
    if (!ok) return 1;
  }
  return 0;
}

// TEST INFRASTRUCTURE ONLY -- C entry points around two more of the reference's own
// translation units, compiled UNCHANGED from /root/reference by oracle/build.py:
//   nautilus::LaserScanToPointCloud      src/input/pointcloud_helpers.cc:28-48
//     (the variant the Nautilus node itself calls, slam_type_builder.cc:56-65)
//   ComputeScatterMatrixScore            src/loop_closure/lc_candidate_filter.cc:34-52
//   LCCandidateFilter::GetLCCandidates   src/loop_closure/lc_candidate_filter.cc:66-81
// The float accumulation of the scatter matrix is the reference's; the eigenvalues come
// from oracle/shim's closed form because Eigen::EigenSolver is un-vendored.
// Nothing under nautilus_b200/ links or loads this.
#include <cstdint>
#include <memory>
#include <vector>

#include "input/pointcloud_helpers.h"
#include "loop_closure/lc_candidate_filter.h"
#include "util/slam_types.h"

namespace nautilus::loop_closure {
double ComputeScatterMatrixScore(const std::vector<Eigen::Vector2f>& pointcloud);  // lc_candidate_filter.cc:34
}

extern "C" int64_t ref_nautilus_laser_to_points(const float* ranges, int64_t n, float angle_min,
                                                float angle_max, float angle_increment,
                                                float range_min, double max_range, float* out_xy) {
  sensor_msgs::LaserScan scan;
  scan.angle_min = angle_min;
  scan.angle_max = angle_max;
  scan.angle_increment = angle_increment;
  scan.range_min = range_min;
  scan.ranges.assign(ranges, ranges + n);
  const std::vector<Eigen::Vector2f> cloud = nautilus::LaserScanToPointCloud(scan, max_range);
  for (size_t i = 0; i < cloud.size(); ++i) {
    out_xy[2 * i] = cloud[i].x();
    out_xy[2 * i + 1] = cloud[i].y();
  }
  return (int64_t)cloud.size();
}

static std::vector<Eigen::Vector2f> cloud_of(const float* xy, int64_t n) {
  std::vector<Eigen::Vector2f> c((size_t)n);
  for (int64_t i = 0; i < n; ++i) c[(size_t)i] = Eigen::Vector2f(xy[2 * i], xy[2 * i + 1]);
  return c;
}

extern "C" double ref_scatter_score(const float* xy, int64_t n) {
  return nautilus::loop_closure::ComputeScatterMatrixScore(cloud_of(xy, n));
}

// GetLCCandidates over a state made of the given clouds and (x, y, theta) poses.
extern "C" int64_t ref_lc_candidates(const float* xy, const int64_t* offsets, int32_t n_scans,
                                     const double* poses, int64_t* out) {
  using namespace nautilus::slam_types;
  SLAMProblem2D problem;
  for (int32_t k = 0; k < n_scans; ++k) {
    SLAMNode2D node;
    node.node_idx = (uint64_t)k;
    node.is_keyframe = false;
    node.pose = RobotPose2D(Eigen::Vector2f((float)poses[3 * k], (float)poses[3 * k + 1]),
                            (float)poses[3 * k + 2]);
    node.lidar_factor.pointcloud = cloud_of(xy + 2 * offsets[k], offsets[k + 1] - offsets[k]);
    problem.nodes.push_back(node);
  }
  auto state = std::make_shared<SLAMState2D>(problem);
  for (int32_t k = 0; k < n_scans; ++k) {  // solution poses in double, as the solver holds them
    state->solution[k].pose[0] = poses[3 * k];
    state->solution[k].pose[1] = poses[3 * k + 1];
    state->solution[k].pose[2] = poses[3 * k + 2];
  }
  nautilus::loop_closure::LCCandidateFilter filter(state);
  const std::vector<size_t> scans = filter.GetLCCandidates();
  for (size_t i = 0; i < scans.size(); ++i) out[i] = (int64_t)scans[i];
  return (int64_t)scans.size();
}

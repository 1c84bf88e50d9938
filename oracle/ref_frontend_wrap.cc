// TEST INFRASTRUCTURE ONLY -- C entry point around the reference's own
// pointcloud_helpers::LaserScanToPointCloud (third_party/csm/src/
// pointcloud_helpers.cc:44-66, with FilterGlancing :21-42), compiled unchanged
// from /root/reference by oracle/build.py.  Nothing under nautilus_b200/ links
// or loads this.
#include <cstdint>
#include <vector>

#include "pointcloud_helpers.h"

extern "C" int64_t ref_csm_laser_to_points(const float* ranges, int64_t n, float angle_min,
                                           float angle_max, float angle_increment,
                                           float range_min, double max_range, int truncate_ends,
                                           float* out_xy) {
  sensor_msgs::LaserScan scan;
  scan.angle_min = angle_min;
  scan.angle_max = angle_max;
  scan.angle_increment = angle_increment;
  scan.range_min = range_min;
  scan.ranges.assign(ranges, ranges + n);
  const std::vector<Eigen::Vector2f> cloud =
      pointcloud_helpers::LaserScanToPointCloud(scan, max_range, truncate_ends != 0);
  for (size_t i = 0; i < cloud.size(); ++i) {
    out_xy[2 * i] = cloud[i].x();
    out_xy[2 * i + 1] = cloud[i].y();
  }
  return (int64_t)cloud.size();
}

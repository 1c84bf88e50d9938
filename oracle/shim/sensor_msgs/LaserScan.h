// TEST INFRASTRUCTURE ONLY -- stand-in for the ROS message header so that the
// reference's third_party/csm/src/pointcloud_helpers.cc compiles unchanged.
// Field names and types follow sensor_msgs/LaserScan.msg (float32 scalars,
// float32[] ranges).
#ifndef NAUTILUS_ORACLE_SHIM_LASERSCAN_H_
#define NAUTILUS_ORACLE_SHIM_LASERSCAN_H_
#include <vector>
namespace sensor_msgs {
struct LaserScan {
  float angle_min = 0, angle_max = 0, angle_increment = 0;
  float time_increment = 0, scan_time = 0;
  float range_min = 0, range_max = 0;
  std::vector<float> ranges, intensities;
};
}  // namespace sensor_msgs
#endif

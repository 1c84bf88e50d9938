// TEST INFRASTRUCTURE ONLY -- minimal sensor_msgs::PointCloud2 (x, y floats
// interleaved) for RosCloudToPointCloud, which the checker never calls.
#ifndef NAUTILUS_ORACLE_SHIM_POINTCLOUD2_H_
#define NAUTILUS_ORACLE_SHIM_POINTCLOUD2_H_
#include <vector>
namespace sensor_msgs {
struct PointCloud2 {
  std::vector<float> xy;
};
}  // namespace sensor_msgs
#endif

// TEST INFRASTRUCTURE ONLY -- sensor_msgs::PointCloud2 as far as the reference's two
// pointcloud_helpers.cc files touch it: the message fields InitPointcloud fills
// (src/input/pointcloud_helpers.cc:50-72) and, for the csm variant's
// RosCloudToPointCloud, x / y floats interleaved (`xy`).  The checker calls neither.
#ifndef NAUTILUS_ORACLE_SHIM_POINTCLOUD2_H_
#define NAUTILUS_ORACLE_SHIM_POINTCLOUD2_H_
#include <cstdint>
#include <string>
#include <vector>
#include "ros/ros.h"
namespace sensor_msgs {
struct PointField {
  std::string name;
  uint32_t offset = 0;
  uint8_t datatype = 0;
  uint32_t count = 0;
};
struct PointCloud2 {
  struct Header { uint32_t seq = 0; ros::Time stamp; std::string frame_id; } header;
  uint32_t height = 0, width = 0;
  std::vector<PointField> fields;
  bool is_bigendian = false;
  uint32_t point_step = 0, row_step = 0;
  std::vector<uint8_t> data;
  bool is_dense = false;
  std::vector<float> xy;
};
}  // namespace sensor_msgs
#endif

// TEST INFRASTRUCTURE ONLY -- stand-in for ros/ros.h: the three names the reference's
// src/input/pointcloud_helpers.cc mentions outside LaserScanToPointCloud (a publisher
// to hand a message to, and a clock for a message header), so that the translation
// unit compiles UNCHANGED into oracle/_ref.  Nothing is ever published.
#ifndef NAUTILUS_ORACLE_SHIM_ROS_H_
#define NAUTILUS_ORACLE_SHIM_ROS_H_
namespace ros {
struct Time {
  double sec = 0;
  static Time now() { return Time(); }
};
struct Publisher {
  template <typename M>
  void publish(const M&) const {}
};
}  // namespace ros
#endif

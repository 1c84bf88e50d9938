// TEST INFRASTRUCTURE ONLY -- iterator over the shim PointCloud2.
#ifndef NAUTILUS_ORACLE_SHIM_PC2_ITER_H_
#define NAUTILUS_ORACLE_SHIM_PC2_ITER_H_
#include <string>
#include "sensor_msgs/PointCloud2.h"
namespace sensor_msgs {
template <typename T>
class PointCloud2Iterator {
 public:
  PointCloud2Iterator(PointCloud2& c, const std::string& field)
      : c_(&c), at_(field == "y" ? 1 : 0) {}
  PointCloud2Iterator end() const { PointCloud2Iterator e(*this); e.at_ = c_->xy.size() + (at_ & 1); return e; }
  bool operator!=(const PointCloud2Iterator& o) const { return at_ < o.at_; }
  PointCloud2Iterator& operator++() { at_ += 2; return *this; }
  T& operator*() { return c_->xy[at_]; }
 private:
  PointCloud2* c_;
  size_t at_;
};
}  // namespace sensor_msgs
#endif

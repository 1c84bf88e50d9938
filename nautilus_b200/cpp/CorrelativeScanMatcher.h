// Drop-in replacement for third_party/csm/src/CorrelativeScanMatcher.h of
// ut-amrl/nautilus (reference lines cited per member): same include name, same
// class and method names, argument meaning and error behaviour, but every
// method forwards to libnautilus_csm_b200.so (include/nautilus_csm.h), i.e. to
// the sm_100a kernels.  Header-only; link with -lnautilus_csm_b200.
//
// Every public member of the reference class is here (h:112-152).
#ifndef SRC_CORRELATIVESCANMATCHER_H_
#define SRC_CORRELATIVESCANMATCHER_H_

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <utility>
#include <vector>

#include "Eigen/Dense"
#include "nautilus_csm.h"

#define DEFAULT_GAUSSIAN_SIGMA 4       // h:19
#define MIN_VALUE_FOR_LOOKUP 1E-10     // h:20

namespace nautilus_b200_detail {
inline void Fatal(const char* what) {  // glog CHECK / LOG(FATAL): abort the process
  std::fprintf(stderr, "nautilus_csm_b200: %s: %s\n", what, nautilus_csm_last_error());
  std::abort();
}
inline void Check(int rc, const char* what) {
  if (rc != NAUTILUS_OK) Fatal(what);
}
// The subset of CImg<double> the reference's callers touch on a LookupTable.
struct Image {
  uint64_t w = 0, h = 0;
  std::vector<double> v;
  Image() {}
  Image(uint64_t w_, uint64_t h_) : w(w_), h(h_), v(w_ * h_, 0.0) {}
  double& operator()(uint64_t x, uint64_t y) { return v[x + y * w]; }
  const double& operator()(uint64_t x, uint64_t y) const { return v[x + y * w]; }
  double* data() { return v.data(); }
  const double* data() const { return v.data(); }
  int width() const { return (int)w; }
  int height() const { return (int)h; }
  void clear() { v.assign(v.size(), 0.0); }
};
}  // namespace nautilus_b200_detail

// struct LookupTable (h:27-110): a host copy of a table with the reference's
// world<->cell maps.  Tables used for matching never leave the device; this
// type only exists for callers that ask for a table.
struct LookupTable {
  uint64_t width;
  uint64_t height;
  double resolution;
  nautilus_b200_detail::Image values;
  LookupTable(const double range, const double resolution)
      : width(floor((range * 2.0) / resolution)),
        height(floor((range * 2.0) / resolution)),
        resolution(resolution) {  // h:32-39
    values = nautilus_b200_detail::Image(width, height);
  }
  LookupTable() : width(0), height(0), resolution(1) {}

  inline uint64_t convertX(float x) const {  // h:43-45
    return (uint64_t)nautilus_csm_cell(resolution, width, x);
  }
  inline uint64_t convertY(float y) const {  // h:47-49
    return (uint64_t)nautilus_csm_cell(resolution, height, y);
  }
  inline double GetPointValue(Eigen::Vector2f point) const {  // h:51-59
    uint64_t x = convertX(point.x());
    uint64_t y = convertY(point.y());
    if (x >= width || y >= height || values(x, y) <= MIN_VALUE_FOR_LOOKUP)
      return MIN_VALUE_FOR_LOOKUP;
    if (!(values(x, y) <= 1.0)) nautilus_b200_detail::Fatal("CHECK_LE(values(x, y), 1.0)");
    return values(x, y);
  }
  inline bool IsInside(Eigen::Vector2f point) const {  // h:61-68
    return convertX(point.x()) < width && convertY(point.y()) < height;
  }
  inline void SetPointValue(Eigen::Vector2f point, double value) {  // h:70-77
    uint64_t x = convertX(point.x());
    uint64_t y = convertY(point.y());
    if (x >= width || y >= height) return;
    values(x, y) = value;
  }
  void clear() { values.clear(); }
  nautilus_b200_detail::Image GetDebugImage() const { return values; }  // h:89
  uint64_t AbsCoords(double x, double y) const {  // h:104-109
    uint64_t row = (height / 2 + round(y / resolution)) * width;
    uint64_t col = (width / 2 + round(x / resolution));
    if (!(row + col <= width * height)) nautilus_b200_detail::Fatal("CHECK_LE(row + col, size)");
    return row + col;
  }
};

class CorrelativeScanMatcher {
 public:
  // h:114-119
  CorrelativeScanMatcher(double scanner_range, double trans_range, double low_res,
                         double high_res)
      : range_(scanner_range), trans_range_(trans_range), low_res_(low_res),
        high_res_(high_res), handle_(nullptr) {}
  CorrelativeScanMatcher(const CorrelativeScanMatcher& o)  // Solver holds it by value
      : range_(o.range_), trans_range_(o.trans_range_), low_res_(o.low_res_),
        high_res_(o.high_res_), handle_(nullptr) {}
  CorrelativeScanMatcher& operator=(const CorrelativeScanMatcher& o) {
    Release();
    range_ = o.range_; trans_range_ = o.trans_range_; low_res_ = o.low_res_; high_res_ = o.high_res_;
    return *this;
  }
  ~CorrelativeScanMatcher() { Release(); }

  // cc:145-245.  a is the query, b is rasterised.
  std::pair<double, std::pair<Eigen::Vector2f, float>> GetTransformation(
      const std::vector<Eigen::Vector2f>& pointcloud_a,
      const std::vector<Eigen::Vector2f>& pointcloud_b, const double rotation_min = 0,
      const double rotation_max = 2 * M_PI) {
    double score = 0;
    float x = 0, y = 0, theta = 0;
    nautilus_b200_detail::Check(
        nautilus_csm_match(Handle(), Data(pointcloud_a), (int64_t)pointcloud_a.size(),
                           Data(pointcloud_b), (int64_t)pointcloud_b.size(), rotation_min,
                           rotation_max, &score, &x, &y, &theta, nullptr),
        "GetTransformation");
    return std::make_pair(score, std::make_pair(Eigen::Vector2f(x, y), theta));
  }

  // cc:247-264
  std::pair<double, std::pair<Eigen::Vector2f, float>> GetTransformation(
      const std::vector<Eigen::Vector2f>& pointcloud_a,
      const std::vector<Eigen::Vector2f>& pointcloud_b, const double rotation_a,
      const double rotation_b, const double rotation_restriction) {
    double lo, hi;
    nautilus_csm_rotation_window(rotation_a, rotation_b, rotation_restriction, &lo, &hi);
    return GetTransformation(pointcloud_a, pointcloud_b, lo, hi);
  }

  // cc:50-53
  LookupTable GetLookupTableHighRes(const std::vector<Eigen::Vector2f>& pointcloud) {
    LookupTable t(range_, high_res_);
    nautilus_b200_detail::Check(
        nautilus_csm_highres_table(Handle(), Data(pointcloud), (int64_t)pointcloud.size(),
                                   t.values.data(), nullptr),
        "GetLookupTableHighRes");
    return t;
  }

  // cc:32-48
  LookupTable GetLookupTableLowRes(const LookupTable& high_res_table) {
    LookupTable t(range_, low_res_);
    nautilus_b200_detail::Check(
        nautilus_csm_lowres_from_table(Handle(), high_res_table.values.data(),
                                       high_res_table.width, high_res_table.height,
                                       high_res_table.resolution, t.values.data()),
        "GetLookupTableLowRes");
    return t;
  }

  // cc:267-319
  Eigen::Matrix3f GetUncertaintyMatrix(const std::vector<Eigen::Vector2f>& pointcloud_a,
                                       const std::vector<Eigen::Vector2f>& pointcloud_b) {
    float out[9];
    nautilus_b200_detail::Check(
        nautilus_csm_uncertainty(Handle(), Data(pointcloud_a), (int64_t)pointcloud_a.size(),
                                 Data(pointcloud_b), (int64_t)pointcloud_b.size(), 3, 0.0, out),
        "GetUncertaintyMatrix");
    Eigen::Matrix3f u;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) u(i, j) = out[i * 3 + j];
    return u;
  }
  // cc:322-373
  Eigen::Matrix2f GetUncertaintyMatrix(const std::vector<Eigen::Vector2f>& pointcloud_a,
                                       const std::vector<Eigen::Vector2f>& pointcloud_b,
                                       const double rotation) {
    float out[4];
    nautilus_b200_detail::Check(
        nautilus_csm_uncertainty(Handle(), Data(pointcloud_a), (int64_t)pointcloud_a.size(),
                                 Data(pointcloud_b), (int64_t)pointcloud_b.size(), 2, rotation, out),
        "GetUncertaintyMatrix");
    Eigen::Matrix2f u;
    for (int i = 0; i < 2; ++i)
      for (int j = 0; j < 2; ++j) u(i, j) = out[i * 2 + j];
    return u;
  }
  // cc:375-398 (the per-cloud loop runs serially here: every call already fills the GPU)
  std::pair<double, double> GetLocalUncertaintyStats(
      const std::vector<std::vector<Eigen::Vector2f>>& comparisonClouds,
      const std::vector<Eigen::Vector2f> targetCloud) {
    double condition_avg = 0.0;
    double scale_avg = 0.0;
    for (size_t i = 0; i < comparisonClouds.size(); i++) {
      std::pair<double, std::pair<Eigen::Vector2f, float>> result =
          GetTransformation(targetCloud, comparisonClouds[i]);
      Eigen::Matrix2f uncertainty =
          GetUncertaintyMatrix(targetCloud, comparisonClouds[i], result.second.second);
      Eigen::EigenSolver<Eigen::Matrix2f> es(uncertainty);
      Eigen::Vector2f eigenvalues = es.eigenvalues().real();
      condition_avg += eigenvalues.maxCoeff() / eigenvalues.minCoeff();
      scale_avg += eigenvalues.maxCoeff();
    }
    condition_avg /= (comparisonClouds.size());
    scale_avg /= (comparisonClouds.size());
    return std::make_pair(condition_avg, scale_avg);
  }

  // New, batch-first: what a batched SolveAutoLC / AddLCConstraints calls
  // (INTEGRATION.md).  The handle is created on first use.
  NautilusCsm* Handle() {
    if (!handle_) {
      NautilusCsmParams p = {};
      p.scanner_range = range_; p.trans_range = trans_range_;
      p.low_res = low_res_; p.high_res = high_res_;
      const char* dev = std::getenv("NAUTILUS_CSM_DEVICE");
      p.device = dev ? std::atoi(dev) : 0;
      nautilus_b200_detail::Check(nautilus_csm_create(&p, &handle_), "nautilus_csm_create");
    }
    return handle_;
  }

 private:
  static const float* Data(const std::vector<Eigen::Vector2f>& c) {
    // a std::vector<Eigen::Vector2f> is contiguous float (x, y) pairs
    static_assert(sizeof(Eigen::Vector2f) == 2 * sizeof(float), "Vector2f layout");
    return c.empty() ? nullptr : reinterpret_cast<const float*>(c.data());
  }
  void Release() {
    if (handle_) nautilus_csm_destroy(handle_);
    handle_ = nullptr;
  }
  double range_;
  double trans_range_;
  double low_res_;
  double high_res_;
  NautilusCsm* handle_;
};

#endif  // SRC_CORRELATIVESCANMATCHER_H_

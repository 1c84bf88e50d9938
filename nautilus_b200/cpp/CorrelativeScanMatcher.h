// Drop-in replacement for third_party/csm/src/CorrelativeScanMatcher.h of
// ut-amrl/nautilus (reference lines cited per member): same include name, same
// struct / class and member names, argument meaning and error behaviour, but every
// compute member forwards to libnautilus_csm_b200.so (include/nautilus_csm.h), i.e.
// to the sm_100a kernels.  Header-only; link with -lnautilus_csm_b200.
//
// Every public member of the reference's LookupTable (h:27-110) and
// CorrelativeScanMatcher (h:112-152) is here, together with the header's global
// `using` declarations (h:21-24), so that the reference's own unit tests
// (third_party/csm/test/correlative_scan_matcher_tests.cc) and its callers
// (src/optimization/solver.cc:56,633, bag_pair_matcher.cc, bag_avg_matcher.cc)
// compile against it unchanged (tests/test_gpu_shim.py builds the former).
#ifndef SRC_CORRELATIVESCANMATCHER_H_
#define SRC_CORRELATIVESCANMATCHER_H_

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <mutex>
#include <string>
#include <utility>
#include <vector>

#if defined(__has_include)
#if __has_include("eigen3/Eigen/Dense")
#include "eigen3/Eigen/Dense"
#else
#include "Eigen/Dense"
#endif
#else
#include "eigen3/Eigen/Dense"
#endif
#include "nautilus_csm.h"

// The reference stores a table in a CImg<double> (h:31, "./CImg.h", 79 805 lines).
// A build that has CImg.h beside this header keeps the real class (so that
// GetDebugImage() can be handed to CImgDisplay as bag_pair_matcher.cc:195-200 does);
// otherwise the few CImg members the reference's LookupTable callers touch are
// provided by the stand-in below.
#if defined(__has_include) && !defined(NAUTILUS_CSM_NO_CIMG)
#if __has_include("./CImg.h")
#include "./CImg.h"
#define NAUTILUS_CSM_HAVE_CIMG 1
#endif
#endif
#ifndef NAUTILUS_CSM_HAVE_CIMG
namespace cimg_library {
template <typename T>
struct CImg {
  unsigned int _width = 0, _height = 0;
  std::vector<T> _data;
  CImg() {}
  CImg(unsigned int w, unsigned int h, unsigned int = 1, unsigned int = 1, T fill = T(0))
      : _width(w), _height(h), _data((size_t)w * h, fill) {}
  T& operator()(unsigned int x, unsigned int y) { return _data[x + (size_t)y * _width]; }
  const T& operator()(unsigned int x, unsigned int y) const { return _data[x + (size_t)y * _width]; }
  T* data() { return _data.data(); }
  const T* data() const { return _data.data(); }
  int width() const { return (int)_width; }
  int height() const { return (int)_height; }
  size_t size() const { return _data.size(); }
  bool is_empty() const { return _data.empty(); }
  CImg<T>& fill(T v) { _data.assign(_data.size(), v); return *this; }
  CImg<T>& clear() { _width = _height = 0; _data.clear(); return *this; }  // CImg::clear empties the image
  T max() const {
    T m = _data.empty() ? T(0) : _data[0];
    for (const T& v : _data) m = v > m ? v : m;
    return m;
  }
};
}  // namespace cimg_library
#endif

#define DEFAULT_GAUSSIAN_SIGMA 4       // h:19
#define MIN_VALUE_FOR_LOOKUP 1E-10     // h:20

using cimg_library::CImg;              // h:21-24
using Eigen::Vector2f;
using std::pair;
using std::vector;

namespace nautilus_b200_detail {
inline void Fatal(const char* what) {  // glog CHECK / LOG(FATAL): abort the process
  std::fprintf(stderr, "nautilus_csm_b200: %s: %s\n", what, nautilus_csm_last_error());
  std::abort();
}
inline void Check(int rc, const char* what) {
  if (rc != NAUTILUS_OK) Fatal(what);
}
inline int Device() {  // CUDA device of handles and table operations made through this header
  const char* dev = std::getenv("NAUTILUS_CSM_DEVICE");
  return dev ? std::atoi(dev) : 0;
}
}  // namespace nautilus_b200_detail

// struct LookupTable (h:27-110): a host copy of a table with the reference's
// world<->cell maps.  Tables used for matching never leave the device; this
// type only exists for callers that ask for a table or make one by hand.
struct LookupTable {
  uint64_t width;
  uint64_t height;
  double resolution;
  CImg<double> values;
  LookupTable(const double range, const double resolution)
      : width(floor((range * 2.0) / resolution)),
        height(floor((range * 2.0) / resolution)),
        resolution(resolution) {  // h:32-39
    values = CImg<double>(width, height, 1, 1, 0);
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
  // h:79 -> CImg::normalize(0, 1) (CImg.h:39499-39517), on the device
  void normalize() {
    if (values.is_empty()) return;
    nautilus_b200_detail::Check(
        nautilus_csm_table_normalize(values.data(), (uint64_t)values.width(),
                                     (uint64_t)values.height(), nautilus_b200_detail::Device()),
        "LookupTable::normalize");
  }
  // h:81-83 -> CImg::blur(sigma, sigma, 0, true, true) (CImg.h:49337-49351), on the device
  void GaussianBlur(const double sigma) {
    if (values.is_empty()) return;
    nautilus_b200_detail::Check(
        nautilus_csm_table_blur(values.data(), (uint64_t)values.width(),
                                (uint64_t)values.height(), sigma, nautilus_b200_detail::Device()),
        "LookupTable::GaussianBlur");
  }
  void clear() { values = values.clear(); }                    // h:85
  void GaussianBlur() { GaussianBlur(DEFAULT_GAUSSIAN_SIGMA); }  // h:87
  CImg<double> GetDebugImage() const { return values; }        // h:89
  // h:91-101: get_crop(sx, sy, ex - 1, ey - 1).max() with CImg's crop rules
  // (CImg.h:45516-45564: int bounds, reversed bounds swapped, zero fill outside)
  double MaxArea(double start_x, double start_y, double end_x, double end_y) const {
    uint64_t sx = convertX(start_x);
    uint64_t sy = convertY(start_y);
    uint64_t ex = convertX(end_x);
    uint64_t ey = convertY(end_y);
    if (values.is_empty()) nautilus_b200_detail::Fatal("MaxArea: crop(): Empty instance");
    const int x0 = (int)sx, y0 = (int)sy, x1 = (int)(ex - 1), y1 = (int)(ey - 1);
    const int nx0 = x0 < x1 ? x0 : x1, nx1 = x0 < x1 ? x1 : x0;
    const int ny0 = y0 < y1 ? y0 : y1, ny1 = y0 < y1 ? y1 : y0;
    const int w = values.width(), h = values.height();
    const bool outside = nx0 < 0 || ny0 < 0 || nx1 >= w || ny1 >= h;
    bool have = outside;
    double max_val = 0.0;  // the zero fill takes part in the maximum
    for (int y = ny0 < 0 ? 0 : ny0; y <= (ny1 >= h ? h - 1 : ny1); ++y)
      for (int x = nx0 < 0 ? 0 : nx0; x <= (nx1 >= w ? w - 1 : nx1); ++x) {
        const double v = values(x, y);
        if (!have || v > max_val) { max_val = v; have = true; }
      }
    return max_val;
  }
  size_t AbsCoords(double x, double y) const {  // h:104-109
    size_t row = ((height / 2) + round(y / resolution)) * width;
    size_t col = (width / 2) + round(x / resolution);
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
        high_res_(o.high_res_), handle_(nullptr) {}  // a copy makes its own handle on first use
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
  // cc:375-398.  The reference loops over the comparison clouds (OpenMP); here all of
  // them go to the device in two batched passes (every match, then every covariance),
  // and the eigenvalue averages are taken on the host with Eigen as in cc:388-396.
  std::pair<double, double> GetLocalUncertaintyStats(
      const std::vector<std::vector<Eigen::Vector2f>>& comparisonClouds,
      const std::vector<Eigen::Vector2f> targetCloud) {
    const size_t n = comparisonClouds.size();
    std::vector<Eigen::Vector2f> all;
    std::vector<int64_t> offs(n + 1, 0);
    for (size_t i = 0; i < n; i++) {
      all.insert(all.end(), comparisonClouds[i].begin(), comparisonClouds[i].end());
      offs[i + 1] = (int64_t)all.size();
    }
    std::vector<float> u(4 * n);
    nautilus_b200_detail::Check(
        nautilus_csm_local_uncertainty(Handle(), Data(targetCloud), (int64_t)targetCloud.size(),
                                       Data(all), offs.data(), (int32_t)n, u.data(), nullptr),
        "GetLocalUncertaintyStats");
    double condition_avg = 0.0;
    double scale_avg = 0.0;
    for (size_t i = 0; i < n; i++) {
      Eigen::Matrix2f uncertainty;
      for (int a = 0; a < 2; ++a)
        for (int b = 0; b < 2; ++b) uncertainty(a, b) = u[4 * i + 2 * a + b];
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
  // (INTEGRATION.md).  The handle is created on first use, under a lock: the
  // reference object is shared by OpenMP threads (bag_avg_matcher.cc:104-111,
  // CorrelativeScanMatcher.cc:381), so two first calls may arrive together.
  NautilusCsm* Handle() {
    NautilusCsm* h = __atomic_load_n(&handle_, __ATOMIC_ACQUIRE);
    if (h) return h;
    std::lock_guard<std::mutex> lock(create_mu_);
    if (!handle_) {
      NautilusCsmParams p = {};
      p.scanner_range = range_; p.trans_range = trans_range_;
      p.low_res = low_res_; p.high_res = high_res_;
      p.device = nautilus_b200_detail::Device();
      NautilusCsm* made = nullptr;
      nautilus_b200_detail::Check(nautilus_csm_create(&p, &made), "nautilus_csm_create");
      __atomic_store_n(&handle_, made, __ATOMIC_RELEASE);
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
    std::lock_guard<std::mutex> lock(create_mu_);
    if (handle_) nautilus_csm_destroy(handle_);
    handle_ = nullptr;
  }
  double range_;
  double trans_range_;
  double low_res_;
  double high_res_;
  NautilusCsm* handle_;
  std::mutex create_mu_;
};

#endif  // SRC_CORRELATIVESCANMATCHER_H_

// TEST INFRASTRUCTURE ONLY -- never linked into, imported by, or called from the
// product path (nautilus_b200/).  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may load the library this
// file builds into.
//
// Flat C wrapper around the UNMODIFIED reference correlative scan matcher.
// oracle/build.py compiles this file together with
//   /root/reference/third_party/csm/src/CorrelativeScanMatcher.cc
// (where it lies; no reference source is copied into this repo) against the
// shim headers in oracle/shim/ into oracle/_ref/libnautilus_ref.so, using the
// reference's Release flags (CMakeLists.txt:11,16: -std=c++17 -O3 -DNDEBUG
// -fopenmp, no -march) plus -Dcimg_display=0 (no X11 here).
//
// Every entry point states the reference function it forwards to.
#include <omp.h>

#include <chrono>
#include <cstdint>
#include <cstring>
#include <vector>

#include "CorrelativeScanMatcher.h"  // from /root/reference/third_party/csm/src

using Eigen::Vector2f;
using std::vector;

namespace {
vector<Vector2f> ToCloud(const float* xy, int64_t n) {
  vector<Vector2f> c;
  c.reserve(n);
  for (int64_t i = 0; i < n; ++i) c.emplace_back(xy[2 * i], xy[2 * i + 1]);
  return c;
}
void CopyOut(const LookupTable& t, double* out) {
  if (out && t.width * t.height > 0)
    std::memcpy(out, t.values.data(), sizeof(double) * t.width * t.height);
}
}  // namespace

extern "C" {

// ---- struct LookupTable (CorrelativeScanMatcher.h:27-110) -------------------
void* ref_table_new(double range, double resolution) {
  return new LookupTable(range, resolution);
}
void ref_table_free(void* t) { delete static_cast<LookupTable*>(t); }
uint64_t ref_table_width(void* t) { return static_cast<LookupTable*>(t)->width; }
uint64_t ref_table_height(void* t) {
  return static_cast<LookupTable*>(t)->height;
}
void ref_table_set(void* t, float x, float y, double v) {
  static_cast<LookupTable*>(t)->SetPointValue(Vector2f(x, y), v);
}
double ref_table_get(void* t, float x, float y) {
  return static_cast<LookupTable*>(t)->GetPointValue(Vector2f(x, y));
}
int ref_table_is_inside(void* t, float x, float y) {
  return static_cast<LookupTable*>(t)->IsInside(Vector2f(x, y)) ? 1 : 0;
}
double ref_table_max_area(void* t, double sx, double sy, double ex, double ey) {
  return static_cast<LookupTable*>(t)->MaxArea(sx, sy, ex, ey);
}
uint64_t ref_table_abs_coords(void* t, double x, double y) {
  return static_cast<LookupTable*>(t)->AbsCoords(x, y);
}
void ref_table_values(void* t, double* out) {
  CopyOut(*static_cast<LookupTable*>(t), out);
}
void ref_table_load(void* t, const double* in) {
  LookupTable* lt = static_cast<LookupTable*>(t);
  std::memcpy(lt->values.data(), in, sizeof(double) * lt->width * lt->height);
}
void ref_table_blur(void* t) { static_cast<LookupTable*>(t)->GaussianBlur(); }
void ref_table_normalize(void* t) { static_cast<LookupTable*>(t)->normalize(); }

// CorrelativeScanMatcher::GetLookupTableLowRes (CorrelativeScanMatcher.cc:32-48)
void* ref_lowres_from_table(double range, double trans, double low, double high,
                            void* hi) {
  CorrelativeScanMatcher m(range, trans, low, high);
  return new LookupTable(
      m.GetLookupTableLowRes(*static_cast<LookupTable*>(hi)));
}

// CorrelativeScanMatcher::GetLookupTableHighRes (CorrelativeScanMatcher.cc:50-53)
void* ref_highres_table(double range, double trans, double low, double high,
                        const float* xy, int64_t n) {
  CorrelativeScanMatcher m(range, trans, low, high);
  return new LookupTable(m.GetLookupTableHighRes(ToCloud(xy, n)));
}

// GetTransformation, 4-arg form (CorrelativeScanMatcher.cc:145-245)
void ref_match(double range, double trans, double low, double high,
               const float* a, int64_t na, const float* b, int64_t nb,
               double rot_min, double rot_max, double* score, float* x,
               float* y, float* theta) {
  CorrelativeScanMatcher m(range, trans, low, high);
  auto r = m.GetTransformation(ToCloud(a, na), ToCloud(b, nb), rot_min, rot_max);
  *score = r.first;
  *x = r.second.first.x();
  *y = r.second.first.y();
  *theta = r.second.second;
}

// GetTransformation, 5-arg form (CorrelativeScanMatcher.cc:247-264)
void ref_match_restricted(double range, double trans, double low, double high,
                          const float* a, int64_t na, const float* b,
                          int64_t nb, double rot_a, double rot_b,
                          double restriction, double* score, float* x, float* y,
                          float* theta) {
  CorrelativeScanMatcher m(range, trans, low, high);
  auto r = m.GetTransformation(ToCloud(a, na), ToCloud(b, nb), rot_a, rot_b,
                               restriction);
  *score = r.first;
  *x = r.second.first.x();
  *y = r.second.first.y();
  *theta = r.second.second;
}

// GetUncertaintyMatrix, 2x2 form (CorrelativeScanMatcher.cc:322-373); row-major out
void ref_uncertainty2(double range, double trans, double low, double high,
                      const float* a, int64_t na, const float* b, int64_t nb,
                      double rotation, float* out4) {
  CorrelativeScanMatcher m(range, trans, low, high);
  Eigen::Matrix2f u =
      m.GetUncertaintyMatrix(ToCloud(a, na), ToCloud(b, nb), rotation);
  for (int i = 0; i < 2; ++i)
    for (int j = 0; j < 2; ++j) out4[i * 2 + j] = u(i, j);
}

// GetUncertaintyMatrix, 3x3 form (CorrelativeScanMatcher.cc:267-319); row-major out
void ref_uncertainty3(double range, double trans, double low, double high,
                      const float* a, int64_t na, const float* b, int64_t nb,
                      float* out9) {
  CorrelativeScanMatcher m(range, trans, low, high);
  Eigen::Matrix3f u = m.GetUncertaintyMatrix(ToCloud(a, na), ToCloud(b, nb));
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) out9[i * 3 + j] = u(i, j);
}

// Batch driver for the CPU baseline: one pair per host thread through an outer
// `omp parallel for`, the usage pattern of bag_avg_matcher.cc:104 (CImg's inner
// OpenMP regions then run serially because nested parallelism is off).
// scans: concatenated float xy; offsets[n_scans+1] in points.
// Returns wall seconds.
double ref_match_batch(double range, double trans, double low, double high,
                       const float* xy, const int64_t* offsets,
                       const int32_t* pair_a, const int32_t* pair_b,
                       const double* rot_min, const double* rot_max,
                       int64_t n_pairs, int threads, double* score, float* x,
                       float* y, float* theta) {
  if (threads > 0) omp_set_num_threads(threads);
  auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic, 1)
  for (int64_t p = 0; p < n_pairs; ++p) {
    const int32_t ia = pair_a[p], ib = pair_b[p];
    CorrelativeScanMatcher m(range, trans, low, high);
    auto r = m.GetTransformation(
        ToCloud(xy + 2 * offsets[ia], offsets[ia + 1] - offsets[ia]),
        ToCloud(xy + 2 * offsets[ib], offsets[ib + 1] - offsets[ib]),
        rot_min[p], rot_max[p]);
    score[p] = r.first;
    x[p] = r.second.first.x();
    y[p] = r.second.first.y();
    theta[p] = r.second.second;
  }
  auto t1 = std::chrono::steady_clock::now();
  return std::chrono::duration<double>(t1 - t0).count();
}

// Batch table build (hi + low) for the CPU baseline of the table-build config.
// Returns wall seconds; tables are discarded (optionally checksummed).
double ref_build_tables_batch(double range, double trans, double low,
                              double high, const float* xy,
                              const int64_t* offsets, int64_t n_scans,
                              int threads, double* checksum) {
  if (threads > 0) omp_set_num_threads(threads);
  double acc = 0;
  auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : acc)
  for (int64_t s = 0; s < n_scans; ++s) {
    CorrelativeScanMatcher m(range, trans, low, high);
    LookupTable hi = m.GetLookupTableHighRes(
        ToCloud(xy + 2 * offsets[s], offsets[s + 1] - offsets[s]));
    LookupTable lo = m.GetLookupTableLowRes(hi);
    acc += lo.values.sum();
  }
  auto t1 = std::chrono::steady_clock::now();
  if (checksum) *checksum = acc;
  return std::chrono::duration<double>(t1 - t0).count();
}

int ref_max_threads() { return omp_get_max_threads(); }

}  // extern "C"

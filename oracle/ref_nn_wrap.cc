// TEST INFRASTRUCTURE ONLY -- never linked into, imported by, or called from the
// product path (nautilus_b200/).  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may load the library this
// file builds into.
//
// Flat C wrapper around the UNMODIFIED reference KD-tree
//   /root/reference/src/util/kdtree.{h,cc}
// compiled where it lies (oracle/build.py) plus a restatement of the caller loop
//   Solver::GetPointToPointMatching (src/optimization/solver.cc:132-172)
//   FindClosestPoint / GetPointNormal (src/optimization/solver.cc:67-89)
// which itself cannot be compiled here (solver.cc needs ROS + Ceres).
#include <omp.h>

#include <chrono>
#include <cstdint>
#include <vector>

#include "util/kdtree.h"  // from /root/reference/src

using Eigen::Vector2f;
using std::vector;

// kdtree.cc:152-163 (EigenToKD) references this symbol; normals are supplied by
// the caller of this wrapper instead, so the reference's normal estimation (off
// the hot path) is stubbed.
namespace nautilus {
namespace NormalComputation {
vector<Vector2f> GetNormals(const vector<Vector2f>& points) {
  return vector<Vector2f>(points.size(), Vector2f(0.f, 0.f));
}
}  // namespace NormalComputation
}  // namespace nautilus

namespace {
typedef KDTree<float, 2> Tree;
Tree* MakeTree(const float* xy, int64_t n) {
  vector<KDNodeValue<float, 2>> values;
  values.reserve(n);
  for (int64_t i = 0; i < n; ++i)
    values.emplace_back(Vector2f(xy[2 * i], xy[2 * i + 1]), Vector2f(0.f, 0.f), (int)i);
  return new Tree(values);  // slam_types.h:57-60
}

// One source cloud against one target tree: solver.cc:146-170 with the
// source->target transform given as the row-major 2x3 the caller computed.
void MatchOne(Tree* tree, const float* src, int64_t n, const float* m, double thr, int32_t* idx,
              float* dist) {
  for (int64_t i = 0; i < n; ++i) {
    const float px = src[2 * i], py = src[2 * i + 1];
    // Affine2f * Vector2f: translation + linear * p (Eigen/src/Geometry/Transform.h)
    Vector2f q((m[0] * px + m[1] * py) + m[2], (m[3] * px + m[4] * py) + m[5]);
    KDNodeValue<float, 2> closest;
    // FindClosestPoint, solver.cc:79-89: threshold narrowed to float by the
    // const T& parameter, result compared in double
    const double d = tree->FindNearestPoint(q, static_cast<float>(thr), &closest);
    idx[i] = (d < thr) ? closest.index : -1;
    if (dist) dist[i] = static_cast<float>(d);
  }
}
}  // namespace

extern "C" {

void* ref_kd_new(const float* xy, int64_t n) { return MakeTree(xy, n); }
void ref_kd_free(void* t) { delete static_cast<Tree*>(t); }

// KDTree<float,2>::FindNearestPoint (kdtree.cc:253-305) for a list of queries.
void ref_kd_nearest(void* t, const float* q_xy, int64_t nq, float threshold, float* dist,
                    int32_t* index) {
  Tree* tree = static_cast<Tree*>(t);
  for (int64_t i = 0; i < nq; ++i) {
    KDNodeValue<float, 2> node;
    dist[i] = tree->FindNearestPoint(Vector2f(q_xy[2 * i], q_xy[2 * i + 1]), threshold, &node);
    index[i] = node.index;
  }
}

// Batch of (source, target) pairs over a scan set; trees are built first (as the
// reference does at load time) and only the matching is timed.  Returns seconds.
double ref_nn_match_batch(const float* xy, const int64_t* offsets, int32_t n_scans,
                          const int32_t* src_ids, const int32_t* tgt_ids, const float* m,
                          double threshold, int32_t n_pairs, const int64_t* out_offsets,
                          int32_t* match_idx, float* match_dist, int32_t threads) {
  vector<Tree*> trees(n_scans, nullptr);
  vector<char> need(n_scans, 0);
  for (int32_t p = 0; p < n_pairs; ++p) need[tgt_ids[p]] = 1;
#pragma omp parallel for schedule(dynamic) num_threads(threads)
  for (int32_t s = 0; s < n_scans; ++s)
    if (need[s] && offsets[s + 1] > offsets[s])
      trees[s] = MakeTree(xy + 2 * offsets[s], offsets[s + 1] - offsets[s]);
  const auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic) num_threads(threads)
  for (int32_t p = 0; p < n_pairs; ++p) {
    const int64_t n = offsets[src_ids[p] + 1] - offsets[src_ids[p]];
    int32_t* idx = match_idx + out_offsets[p];
    float* dist = match_dist ? match_dist + out_offsets[p] : nullptr;
    if (!trees[tgt_ids[p]]) {
      for (int64_t i = 0; i < n; ++i) { idx[i] = -1; if (dist) dist[i] = 0.f; }
      continue;
    }
    MatchOne(trees[tgt_ids[p]], xy + 2 * offsets[src_ids[p]], n, m + 6 * (int64_t)p, threshold,
             idx, dist);
  }
  const auto t1 = std::chrono::steady_clock::now();
  for (Tree* t : trees) delete t;
  return std::chrono::duration<double>(t1 - t0).count();
}

}  // extern "C"

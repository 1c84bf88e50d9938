// TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's point-to-point
// nearest-neighbour path; never linked into or called from the product path.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use it.
//
// Restates (file:line under /root/reference):
//   GetSplittingPlane            src/util/kdtree.cc:46-77
//   KDTree::BuildKDTree          src/util/kdtree.cc:114-150
//   KDTree::FindNearestPoint     src/util/kdtree.cc:253-305
//   FindClosestPoint             src/optimization/solver.cc:79-89
//   GetPointToPointMatching loop src/optimization/solver.cc:146-170
// C++ rather than C because the tree layout depends on std::sort's handling of
// equal keys (kdtree.cc:126); using the same libstdc++ algorithm is the only
// faithful restatement of that line.  Pinned against outputs of the compiled
// reference (tests/golden/nn_golden.npz, tests/test_oracle_golden.py).
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <vector>

namespace {

struct Value {
  float p[2];
  int index;
};

struct Node {
  int dim = 0;
  Value value{};
  Node* left = nullptr;
  Node* right = nullptr;
  ~Node() { delete left; delete right; }
};

int SplittingPlane(const std::vector<Value>& values) {  // kdtree.cc:46-77
  float mean[2] = {0.f, 0.f}, dev[2] = {0.f, 0.f};
  for (size_t i = 0; i < values.size(); ++i)
    for (int j = 0; j < 2; ++j) mean[j] = mean[j] + values[i].p[j];
  for (int j = 0; j < 2; ++j) mean[j] = mean[j] / static_cast<float>(values.size());
  for (size_t i = 0; i < values.size(); ++i)
    for (int j = 0; j < 2; ++j)
      dev[j] = dev[j] + (values[i].p[j] - mean[j]) * (values[i].p[j] - mean[j]);
  int plane = 0;
  float best = 0.0f;
  for (int j = 0; j < 2; ++j)
    if (dev[j] > best) { plane = j; best = dev[j]; }
  return plane;
}

Node* Build(std::vector<Value> values) {  // kdtree.cc:114-150
  Node* n = new Node();
  if (values.empty()) return n;
  n->dim = SplittingPlane(values);
  const int dim = n->dim;
  std::sort(values.begin(), values.end(),
            [dim](const Value& a, const Value& b) { return a.p[dim] < b.p[dim]; });
  const size_t ind = values.size() / 2;
  n->value = values[ind];
  if (ind > 0) n->left = Build(std::vector<Value>(values.begin(), values.begin() + ind));
  if (ind < values.size() - 1)
    n->right = Build(std::vector<Value>(values.begin() + ind + 1, values.end()));
  return n;
}

float Norm(const float* a, const float* b) {  // (value_.point - point).norm()
  const float dx = a[0] - b[0], dy = a[1] - b[1];
  const float xx = dx * dx, yy = dy * dy;
  return std::sqrt(xx + yy);
}

float Nearest(const Node* n, const float* q, float threshold, Value* out) {  // kdtree.cc:253-305
  float best = Norm(n->value.p, q);
  *out = n->value;
  if (best < FLT_MIN) return 0.0f;
  const float pd = q[n->dim] - n->value.p[n->dim];
  const Node* other = nullptr;
  if (pd <= 0.0 && n->left != nullptr) {
    Value v;
    const float d = Nearest(n->left, q, std::min(best, threshold), &v);
    if (d < best) { best = d; *out = v; }
    other = n->right;
  }
  if (pd >= 0.0 && n->right != nullptr) {
    Value v;
    const float d = Nearest(n->right, q, std::min(best, threshold), &v);
    if (d < best) { best = d; *out = v; }
    other = n->left;
  }
  if (other != nullptr && pd != 0.0 && std::fabs(pd) < std::min(best, threshold)) {
    Value v;
    const float d = Nearest(other, q, std::min(best, threshold), &v);
    if (d < best) { best = d; *out = v; }
  }
  return best;
}

}  // namespace

extern "C" {

void* nn_port_tree_new(const float* xy, int64_t n) {
  std::vector<Value> v(n);
  for (int64_t i = 0; i < n; ++i) { v[i].p[0] = xy[2 * i]; v[i].p[1] = xy[2 * i + 1]; v[i].index = (int)i; }
  return Build(v);
}
void nn_port_tree_free(void* t) { delete static_cast<Node*>(t); }

// solver.cc:146-170 for one (source cloud, target tree) pair; m = row-major 2x3
// source->target transform; idx = matched target index or -1.
void nn_port_match(void* tree, int64_t n_tgt, const float* src, int64_t n, const float* m,
                   double threshold, int32_t* idx, float* dist) {
  const Node* t = static_cast<const Node*>(tree);
  for (int64_t i = 0; i < n; ++i) {
    const float px = src[2 * i], py = src[2 * i + 1];
    const float a0 = m[0] * px, a1 = m[1] * py, b0 = m[3] * px, b1 = m[4] * py;
    const float q[2] = {(a0 + a1) + m[2], (b0 + b1) + m[5]};
    if (n_tgt == 0) { idx[i] = -1; if (dist) dist[i] = 0.f; continue; }
    Value v;
    const double d = Nearest(t, q, static_cast<float>(threshold), &v);
    idx[i] = (d < threshold) ? v.index : -1;
    if (dist) dist[i] = static_cast<float>(d);
  }
}

}  // extern "C"

// TEST INFRASTRUCTURE ONLY (patches/check): the Eigen geometry types the reference's
// slam_types.h / slam_util.h / data_structures.h use beyond oracle/shim's matrices,
// declared with Eigen's signatures so that patches/solver_batched.patch's new
// translation unit (src/optimization/csm_batch.cc) can be syntax-checked against the
// reference's REAL headers in an image that has no Eigen.  Bodies are the textbook
// 2-D affine algebra; nothing here is used by the product or the oracle.
#pragma once
#include "../../../oracle/shim/eigen3/Eigen/Dense"

namespace Eigen {
enum TransformTraits { Isometry = 1, Affine = 2, AffineCompact = 2 | 0x10, Projective = 0x20 };
template <typename T, int D> class Translation;
template <typename T, int D, int Mode> class Transform;

template <typename T>
class Transform<T, 2, Affine> {
 public:
  typedef Matrix<T, 2, 2> LinearMatrixType;
  typedef Matrix<T, 2, 1> VectorType;
  Transform() { l_(0, 0) = T(1); l_(1, 1) = T(1); }
  Transform(const LinearMatrixType& l, const VectorType& t) : l_(l), t_(t) {}
  const LinearMatrixType& linear() const { return l_; }
  LinearMatrixType& linear() { return l_; }
  const VectorType& translation() const { return t_; }
  VectorType& translation() { return t_; }
  Transform inverse() const {
    const T det = l_(0, 0) * l_(1, 1) - l_(0, 1) * l_(1, 0);
    LinearMatrixType inv;
    inv(0, 0) = l_(1, 1) / det; inv(0, 1) = -l_(0, 1) / det;
    inv(1, 0) = -l_(1, 0) / det; inv(1, 1) = l_(0, 0) / det;
    return Transform(inv, -(inv * t_));
  }
  Transform operator*(const Transform& o) const { return Transform(l_ * o.l_, l_ * o.t_ + t_); }
  VectorType operator*(const VectorType& p) const { return l_ * p + t_; }
  template <typename U>
  Transform<U, 2, Affine> cast() const {
    return Transform<U, 2, Affine>(l_.template cast<U>(), t_.template cast<U>());
  }
 private:
  LinearMatrixType l_;
  VectorType t_;
};

template <typename T>
class Translation<T, 2> {
 public:
  Translation(const T& x, const T& y) : v_(x, y) {}
  explicit Translation(const Matrix<T, 2, 1>& v) : v_(v) {}
  Transform<T, 2, Affine> operator*(const Matrix<T, 2, 2>& linear) const {
    return Transform<T, 2, Affine>(linear, v_);
  }
 private:
  Matrix<T, 2, 1> v_;
};
typedef Transform<float, 2, Affine> Affine2f;
typedef Transform<double, 2, Affine> Affine2d;
typedef Translation<float, 2> Translation2f;
typedef Translation<double, 2> Translation2d;

// dynamic matrix and hyperplane: only what slam_util.h's helpers name
class MatrixXd {
 public:
  MatrixXd(int r, int c) : r_(r), d_((size_t)r * c, 0.0) {}
  void setZero() { d_.assign(d_.size(), 0.0); }
  double& operator()(int i, int j) { return d_[i + (size_t)j * r_]; }
 private:
  int r_;
  std::vector<double> d_;
};
template <typename T, int D>
class Hyperplane {
 public:
  typedef Matrix<T, D, 1> V;
  static Hyperplane Through(const V& a, const V& b) {
    Hyperplane h;
    V n(-(b.y() - a.y()), b.x() - a.x());
    const T len = n.norm();
    h.n_ = n / len;
    h.o_ = -h.n_.dot(a);
    return h;
  }
  T signedDistance(const V& p) const { return n_.dot(p) + o_; }
  T absDistance(const V& p) const { return std::abs(signedDistance(p)); }
  V projection(const V& p) const { return p - n_ * signedDistance(p); }
 private:
  V n_;
  T o_ = T(0);
};
}  // namespace Eigen

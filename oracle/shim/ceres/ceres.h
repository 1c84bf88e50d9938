// TEST INFRASTRUCTURE ONLY -- stand-in for Ceres Solver 1.14.0 (un-vendored dependency
// of the reference, install_ubuntu.sh:24-45; absent from this image), restricted to
// what lets the reference's src/optimization/slam_residuals.h compile UNCHANGED into
// oracle/_ref and be evaluated the way the solver evaluates it:
//   ceres::Jet<T, N>             forward-mode dual numbers, arithmetic as published in
//                                ceres/jet.h (1.14): (f, f') op (g, g') rules below
//   ceres::AutoDiffCostFunction  Evaluate(): seeds one Jet per parameter with a unit
//                                infinitesimal part, runs the functor once, reads the
//                                residuals and the row-major Jacobian blocks off the
//                                result (ceres/internal/autodiff.h, 1.14)
// plus the few names data_structures.h / slam_util.h mention (Problem, ResidualBlockId,
// IsFinite, CRSMatrix).  No solver: nothing here minimises anything.
#pragma once
#include <cmath>
#include <vector>

namespace ceres {

class Problem {};
typedef void* ResidualBlockId;
inline bool IsFinite(double v) { return std::isfinite(v); }
struct CRSMatrix { int num_rows = 0, num_cols = 0; std::vector<int> cols, rows; std::vector<double> values; };
enum { DYNAMIC = -1 };

template <typename T, int N>
struct Jet {
  T a;
  T v[N];
  Jet() : a() { for (int i = 0; i < N; ++i) v[i] = T(); }
  explicit Jet(const T& value) : a(value) { for (int i = 0; i < N; ++i) v[i] = T(); }
  Jet(const T& value, int k) : a(value) { for (int i = 0; i < N; ++i) v[i] = T(); v[k] = T(1.0); }
  Jet& operator+=(const Jet& g) { *this = *this + g; return *this; }
  Jet& operator-=(const Jet& g) { *this = *this - g; return *this; }
};
// jet.h: unary and binary operators
template <typename T, int N> Jet<T, N> operator-(const Jet<T, N>& f) {
  Jet<T, N> h; h.a = -f.a; for (int i = 0; i < N; ++i) h.v[i] = -f.v[i]; return h;
}
template <typename T, int N> Jet<T, N> operator+(const Jet<T, N>& f, const Jet<T, N>& g) {
  Jet<T, N> h; h.a = f.a + g.a; for (int i = 0; i < N; ++i) h.v[i] = f.v[i] + g.v[i]; return h;
}
template <typename T, int N> Jet<T, N> operator-(const Jet<T, N>& f, const Jet<T, N>& g) {
  Jet<T, N> h; h.a = f.a - g.a; for (int i = 0; i < N; ++i) h.v[i] = f.v[i] - g.v[i]; return h;
}
template <typename T, int N> Jet<T, N> operator*(const Jet<T, N>& f, const Jet<T, N>& g) {
  Jet<T, N> h; h.a = f.a * g.a; for (int i = 0; i < N; ++i) h.v[i] = f.a * g.v[i] + f.v[i] * g.a; return h;
}
template <typename T, int N> Jet<T, N> operator/(const Jet<T, N>& f, const Jet<T, N>& g) {
  const T g_a_inverse = T(1.0) / g.a;
  const T f_a_by_g_a = f.a * g_a_inverse;
  Jet<T, N> h; h.a = f_a_by_g_a;
  for (int i = 0; i < N; ++i) h.v[i] = (f.v[i] - f_a_by_g_a * g.v[i]) * g_a_inverse;
  return h;
}
template <typename T, int N> Jet<T, N> operator+(const Jet<T, N>& f, T s) { Jet<T, N> h = f; h.a = f.a + s; return h; }
template <typename T, int N> Jet<T, N> operator+(T s, const Jet<T, N>& f) { Jet<T, N> h = f; h.a = f.a + s; return h; }
template <typename T, int N> Jet<T, N> operator-(const Jet<T, N>& f, T s) { Jet<T, N> h = f; h.a = f.a - s; return h; }
template <typename T, int N> Jet<T, N> operator-(T s, const Jet<T, N>& f) { Jet<T, N> h = -f; h.a = s - f.a; return h; }
template <typename T, int N> Jet<T, N> operator*(const Jet<T, N>& f, T s) {
  Jet<T, N> h; h.a = f.a * s; for (int i = 0; i < N; ++i) h.v[i] = f.v[i] * s; return h;
}
template <typename T, int N> Jet<T, N> operator*(T s, const Jet<T, N>& f) { return f * s; }
template <typename T, int N> Jet<T, N> operator/(const Jet<T, N>& f, T s) {
  const T s_inverse = T(1.0) / s;
  Jet<T, N> h; h.a = f.a * s_inverse; for (int i = 0; i < N; ++i) h.v[i] = f.v[i] * s_inverse; return h;
}
template <typename T, int N> bool operator<(const Jet<T, N>& f, const Jet<T, N>& g) { return f.a < g.a; }
template <typename T, int N> bool operator<=(const Jet<T, N>& f, const Jet<T, N>& g) { return f.a <= g.a; }
template <typename T, int N> bool operator>(const Jet<T, N>& f, const Jet<T, N>& g) { return f.a > g.a; }
template <typename T, int N> bool operator>=(const Jet<T, N>& f, const Jet<T, N>& g) { return f.a >= g.a; }
// jet.h: elementary functions (found by argument-dependent lookup from templated code)
template <typename T, int N> Jet<T, N> sin(const Jet<T, N>& f) {
  const T c = std::cos(f.a);
  Jet<T, N> h; h.a = std::sin(f.a); for (int i = 0; i < N; ++i) h.v[i] = c * f.v[i]; return h;
}
template <typename T, int N> Jet<T, N> cos(const Jet<T, N>& f) {
  const T s = -std::sin(f.a);
  Jet<T, N> h; h.a = std::cos(f.a); for (int i = 0; i < N; ++i) h.v[i] = s * f.v[i]; return h;
}
template <typename T, int N> Jet<T, N> sqrt(const Jet<T, N>& f) {
  const T r = std::sqrt(f.a), two_r_inverse = T(1.0) / (T(2.0) * r);
  Jet<T, N> h; h.a = r; for (int i = 0; i < N; ++i) h.v[i] = f.v[i] * two_r_inverse; return h;
}
template <typename T, int N> Jet<T, N> abs(const Jet<T, N>& f) { return f.a < T(0.0) ? -f : f; }
template <typename T, int N> Jet<T, N> atan2(const Jet<T, N>& g, const Jet<T, N>& f) {
  const T tmp = T(1.0) / (f.a * f.a + g.a * g.a);
  Jet<T, N> h; h.a = std::atan2(g.a, f.a);
  for (int i = 0; i < N; ++i) h.v[i] = tmp * (-g.a * f.v[i] + f.a * g.v[i]);
  return h;
}

// AutoDiffCostFunction<Functor, kNumResiduals, N0, N1>: two parameter blocks, which is
// all slam_residuals.h declares.
template <typename Functor, int kNumResiduals, int N0, int N1>
class AutoDiffCostFunction {
 public:
  explicit AutoDiffCostFunction(Functor* functor) : functor_(functor), num_residuals_(kNumResiduals) {}
  AutoDiffCostFunction(Functor* functor, int num_residuals)
      : functor_(functor), num_residuals_(num_residuals) {}
  ~AutoDiffCostFunction() { delete functor_; }
  int num_residuals() const { return num_residuals_; }
  // jacobians[b] (may be null, as may the array): num_residuals x N_b, row-major
  bool Evaluate(double const* const* parameters, double* residuals, double** jacobians) const {
    if (!jacobians) return (*functor_)(parameters[0], parameters[1], residuals);
    typedef Jet<double, N0 + N1> J;
    J x0[N0], x1[N1];
    for (int i = 0; i < N0; ++i) x0[i] = J(parameters[0][i], i);
    for (int i = 0; i < N1; ++i) x1[i] = J(parameters[1][i], N0 + i);
    std::vector<J> out((size_t)num_residuals_);
    if (!(*functor_)(x0, x1, out.data())) return false;
    for (int r = 0; r < num_residuals_; ++r) {
      residuals[r] = out[(size_t)r].a;
      if (jacobians[0]) for (int i = 0; i < N0; ++i) jacobians[0][r * N0 + i] = out[(size_t)r].v[i];
      if (jacobians[1]) for (int i = 0; i < N1; ++i) jacobians[1][r * N1 + i] = out[(size_t)r].v[N0 + i];
    }
    return true;
  }
 private:
  Functor* functor_;
  int num_residuals_;
};

}  // namespace ceres

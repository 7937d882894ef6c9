// ORACLE (test infrastructure only -- never linked into the product library).
//
// Minimal forward-mode dual number, restating what ceres::Jet<double, N> does
// for the functors in /root/reference/src/base/cost_functions.h:57-81 when they
// are wrapped in ceres::AutoDiffCostFunction (cost_functions.h:50-55).
// Ceres 2.1.0 is an un-vendored dependency of the reference
// (/root/reference/CMakeLists.txt:91, scripts/python/build.py:388-392); this
// follows its published Jet algebra: value part + N-vector of partials,
// comparisons act on the value part only.
#pragma once
#include <cmath>

namespace oracle {

template <int N>
struct Jet {
  double a;
  double v[N];
  Jet() : a(0.0) {
    for (int i = 0; i < N; ++i) v[i] = 0.0;
  }
  Jet(double s) : a(s) {  // NOLINT implicit on purpose: T(1e-4) etc.
    for (int i = 0; i < N; ++i) v[i] = 0.0;
  }
  Jet(double s, int k) : a(s) {
    for (int i = 0; i < N; ++i) v[i] = 0.0;
    v[k] = 1.0;
  }
};

template <int N>
inline Jet<N> operator+(const Jet<N>& f, const Jet<N>& g) {
  Jet<N> h;
  h.a = f.a + g.a;
  for (int i = 0; i < N; ++i) h.v[i] = f.v[i] + g.v[i];
  return h;
}
template <int N>
inline Jet<N> operator-(const Jet<N>& f, const Jet<N>& g) {
  Jet<N> h;
  h.a = f.a - g.a;
  for (int i = 0; i < N; ++i) h.v[i] = f.v[i] - g.v[i];
  return h;
}
template <int N>
inline Jet<N> operator-(const Jet<N>& f) {
  Jet<N> h;
  h.a = -f.a;
  for (int i = 0; i < N; ++i) h.v[i] = -f.v[i];
  return h;
}
template <int N>
inline Jet<N> operator*(const Jet<N>& f, const Jet<N>& g) {
  Jet<N> h;
  h.a = f.a * g.a;
  for (int i = 0; i < N; ++i) h.v[i] = f.a * g.v[i] + f.v[i] * g.a;
  return h;
}
template <int N>
inline Jet<N> operator/(const Jet<N>& f, const Jet<N>& g) {
  // Ceres: g_a_inverse = 1/g.a; f_a_by_g_a = f.a * g_a_inverse;
  //        h.v = (f.v - f_a_by_g_a * g.v) * g_a_inverse
  Jet<N> h;
  const double gi = 1.0 / g.a;
  const double fg = f.a * gi;
  h.a = fg;
  for (int i = 0; i < N; ++i) h.v[i] = (f.v[i] - fg * g.v[i]) * gi;
  return h;
}
// scalar mixes
template <int N> inline Jet<N> operator+(const Jet<N>& f, double s) { Jet<N> h = f; h.a += s; return h; }
template <int N> inline Jet<N> operator+(double s, const Jet<N>& f) { Jet<N> h = f; h.a += s; return h; }
template <int N> inline Jet<N> operator-(const Jet<N>& f, double s) { Jet<N> h = f; h.a -= s; return h; }
template <int N> inline Jet<N> operator-(double s, const Jet<N>& f) { Jet<N> h = -f; h.a += s; return h; }
template <int N> inline Jet<N> operator*(const Jet<N>& f, double s) {
  Jet<N> h; h.a = f.a * s; for (int i = 0; i < N; ++i) h.v[i] = f.v[i] * s; return h;
}
template <int N> inline Jet<N> operator*(double s, const Jet<N>& f) { return f * s; }
template <int N> inline Jet<N> operator/(const Jet<N>& f, double s) { return f * (1.0 / s); }
template <int N> inline Jet<N> operator/(double s, const Jet<N>& g) { return Jet<N>(s) / g; }
template <int N> inline Jet<N>& operator+=(Jet<N>& f, const Jet<N>& g) { f = f + g; return f; }
template <int N> inline Jet<N>& operator-=(Jet<N>& f, const Jet<N>& g) { f = f - g; return f; }
template <int N> inline Jet<N>& operator*=(Jet<N>& f, const Jet<N>& g) { f = f * g; return f; }
template <int N> inline Jet<N>& operator/=(Jet<N>& f, const Jet<N>& g) { f = f / g; return f; }

// comparisons: value part only
template <int N> inline bool operator<(const Jet<N>& f, const Jet<N>& g) { return f.a < g.a; }
template <int N> inline bool operator>(const Jet<N>& f, const Jet<N>& g) { return f.a > g.a; }
template <int N> inline bool operator<=(const Jet<N>& f, const Jet<N>& g) { return f.a <= g.a; }
template <int N> inline bool operator>=(const Jet<N>& f, const Jet<N>& g) { return f.a >= g.a; }

// elementary functions (ceres::sqrt / atan / tan / sin / cos on Jets)
template <int N> inline Jet<N> sqrt(const Jet<N>& f) {
  Jet<N> h; h.a = std::sqrt(f.a); const double d = 1.0 / (2.0 * h.a);
  for (int i = 0; i < N; ++i) h.v[i] = f.v[i] * d; return h;
}
template <int N> inline Jet<N> atan(const Jet<N>& f) {
  Jet<N> h; h.a = std::atan(f.a); const double d = 1.0 / (1.0 + f.a * f.a);
  for (int i = 0; i < N; ++i) h.v[i] = f.v[i] * d; return h;
}
template <int N> inline Jet<N> tan(const Jet<N>& f) {
  Jet<N> h; h.a = std::tan(f.a); const double d = 1.0 + h.a * h.a;
  for (int i = 0; i < N; ++i) h.v[i] = f.v[i] * d; return h;
}
template <int N> inline Jet<N> sin(const Jet<N>& f) {
  Jet<N> h; h.a = std::sin(f.a); const double d = std::cos(f.a);
  for (int i = 0; i < N; ++i) h.v[i] = f.v[i] * d; return h;
}
template <int N> inline Jet<N> cos(const Jet<N>& f) {
  Jet<N> h; h.a = std::cos(f.a); const double d = -std::sin(f.a);
  for (int i = 0; i < N; ++i) h.v[i] = f.v[i] * d; return h;
}
// double overloads so templated code can call oracle::sqrt etc. for T=double
inline double sqrt(double x) { return std::sqrt(x); }
inline double atan(double x) { return std::atan(x); }
inline double tan(double x) { return std::tan(x); }
inline double sin(double x) { return std::sin(x); }
inline double cos(double x) { return std::cos(x); }

}  // namespace oracle

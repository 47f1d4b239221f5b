// ORACLE — TEST INFRASTRUCTURE ONLY.  Nothing in the product path may include,
// link or call this file (see oracle/README.md).  PARITY UNPINNED: the reference
// ships no golden vectors and its arithmetic lives in absent third-party code
// (Eigen / IKFoM MTK), so this is a restatement of published algorithms.
//
// Minimal fixed-size f64 linear algebra + the MTK manifold helpers the IKFoM
// toolkit uses (SO3 exp/log, A_matrix, S2 charts).  Written from the published
// IKFoM description [RECALL: SURVEY.md §9.6]; no reference source was available.
#pragma once
#include <cmath>
#include <cstring>
#include <algorithm>

namespace orc {

template <int R, int C>
struct Mat {
  double a[R * C];
  Mat() { std::memset(a, 0, sizeof(a)); }
  double& operator()(int r, int c) { return a[r * C + c]; }
  double operator()(int r, int c) const { return a[r * C + c]; }
  static Mat Identity() {
    Mat m;
    for (int i = 0; i < (R < C ? R : C); ++i) m(i, i) = 1.0;
    return m;
  }
  Mat<C, R> T() const {
    Mat<C, R> t;
    for (int r = 0; r < R; ++r)
      for (int c = 0; c < C; ++c) t(c, r) = (*this)(r, c);
    return t;
  }
  Mat operator+(const Mat& o) const { Mat m; for (int i = 0; i < R * C; ++i) m.a[i] = a[i] + o.a[i]; return m; }
  Mat operator-(const Mat& o) const { Mat m; for (int i = 0; i < R * C; ++i) m.a[i] = a[i] - o.a[i]; return m; }
  Mat operator*(double s) const { Mat m; for (int i = 0; i < R * C; ++i) m.a[i] = a[i] * s; return m; }
};

template <int R, int K, int C>
Mat<R, C> operator*(const Mat<R, K>& A, const Mat<K, C>& B) {
  Mat<R, C> m;
  for (int r = 0; r < R; ++r)
    for (int c = 0; c < C; ++c) {
      double s = 0.0;
      for (int k = 0; k < K; ++k) s += A(r, k) * B(k, c);
      m(r, c) = s;
    }
  return m;
}

using V3 = Mat<3, 1>;
using M3 = Mat<3, 3>;

inline V3 vec3(double x, double y, double z) { V3 v; v.a[0] = x; v.a[1] = y; v.a[2] = z; return v; }
inline double dot(const V3& a, const V3& b) { return a.a[0] * b.a[0] + a.a[1] * b.a[1] + a.a[2] * b.a[2]; }
inline double norm(const V3& a) { return std::sqrt(dot(a, a)); }
inline M3 hat(const V3& v) {
  M3 m;
  m(0, 1) = -v.a[2]; m(0, 2) = v.a[1];
  m(1, 0) = v.a[2];  m(1, 2) = -v.a[0];
  m(2, 0) = -v.a[1]; m(2, 1) = v.a[0];
  return m;
}

// General inverse by Gauss-Jordan with partial pivoting (Eigen uses PartialPivLU
// for dynamic/large fixed sizes; same algorithm class).
template <int N>
Mat<N, N> inverse(const Mat<N, N>& Ain) {
  Mat<N, N> A = Ain, I = Mat<N, N>::Identity();
  for (int c = 0; c < N; ++c) {
    int piv = c;
    double best = std::fabs(A(c, c));
    for (int r = c + 1; r < N; ++r)
      if (std::fabs(A(r, c)) > best) { best = std::fabs(A(r, c)); piv = r; }
    if (piv != c)
      for (int k = 0; k < N; ++k) { std::swap(A(c, k), A(piv, k)); std::swap(I(c, k), I(piv, k)); }
    double d = 1.0 / A(c, c);
    for (int k = 0; k < N; ++k) { A(c, k) *= d; I(c, k) *= d; }
    for (int r = 0; r < N; ++r) {
      if (r == c) continue;
      double f = A(r, c);
      if (f == 0.0) continue;
      for (int k = 0; k < N; ++k) { A(r, k) -= f * A(c, k); I(r, k) -= f * I(c, k); }
    }
  }
  return I;
}

// ---------------------------------------------------------------- SO(3)
// Rodrigues: exp of a rotation vector -> rotation matrix.
inline M3 so3_exp(const V3& w) {
  double th = norm(w);
  M3 R = M3::Identity();
  if (th < 1e-11) {
    M3 K = hat(w);
    return R + K;  // first-order; |w| is below f64 resolution of the next term
  }
  V3 k = vec3(w.a[0] / th, w.a[1] / th, w.a[2] / th);
  M3 K = hat(k);
  return R + K * std::sin(th) + (K * K) * (1.0 - std::cos(th));
}

// log of a rotation matrix -> rotation vector.
inline V3 so3_log(const M3& R) {
  double tr = R(0, 0) + R(1, 1) + R(2, 2);
  double c = 0.5 * (tr - 1.0);
  c = std::min(1.0, std::max(-1.0, c));
  V3 v = vec3(R(2, 1) - R(1, 2), R(0, 2) - R(2, 0), R(1, 0) - R(0, 1));
  double s = 0.5 * norm(v);  // sin(theta)
  double th = std::atan2(s, c);
  if (s < 1e-11) {
    // theta ~ 0 (theta ~ pi never occurs for IEKF increments)
    return v * 0.5;
  }
  return v * (0.5 * th / s);
}

// MTK::A_matrix — the SO(3) left-Jacobian-like matrix IKFoM uses.
inline M3 A_matrix(const V3& v) {
  double sq = dot(v, v);
  double n = std::sqrt(sq);
  M3 I = M3::Identity();
  if (n < 1e-11) return I;
  M3 K = hat(v);
  return I + K * ((1.0 - std::cos(n)) / sq) + (K * K) * ((1.0 - std::sin(n) / n) / sq);
}

// ---------------------------------------------------------------- S2 (gravity, |g| = 9.809)
constexpr double kS2Len = 9.809;
constexpr double kTol = 1e-11;

inline Mat<3, 2> S2_Bx(const V3& g) {
  Mat<3, 2> B;
  if (g.a[2] + kS2Len > kTol) {
    double d = kS2Len + g.a[2];
    B(0, 0) = kS2Len - g.a[0] * g.a[0] / d; B(0, 1) = -g.a[0] * g.a[1] / d;
    B(1, 0) = -g.a[0] * g.a[1] / d;         B(1, 1) = kS2Len - g.a[1] * g.a[1] / d;
    B(2, 0) = -g.a[0];                      B(2, 1) = -g.a[1];
    B = B * (1.0 / kS2Len);
  } else {
    B(1, 1) = -1.0;
    B(2, 0) = 1.0;
  }
  return B;
}

inline V3 S2_boxplus(const V3& g, const Mat<2, 1>& d) {
  Mat<3, 2> B = S2_Bx(g);
  V3 Bu = B * d;
  return so3_exp(Bu) * g;
}

inline Mat<2, 1> S2_boxminus(const V3& g, const V3& other) {
  Mat<2, 1> res;
  V3 cr = hat(g) * other;
  double v_sin = norm(cr);
  double v_cos = dot(g, other);
  double theta = std::atan2(v_sin, v_cos);
  if (v_sin < kTol) {
    if (std::fabs(theta) > kTol) { res.a[0] = 3.1415926; res.a[1] = 0.0; }
    return res;
  }
  Mat<3, 2> B = S2_Bx(other);
  V3 t = hat(other) * g;
  Mat<2, 1> r = B.T() * t;
  return r * (theta / v_sin);
}

inline Mat<2, 3> S2_Nx_yy(const V3& g) {
  Mat<3, 2> B = S2_Bx(g);
  return (B.T() * hat(g)) * (1.0 / kS2Len / kS2Len);
}

inline Mat<3, 2> S2_Mx(const V3& g, const Mat<2, 1>& delta) {
  Mat<3, 2> B = S2_Bx(g);
  double dn = std::sqrt(delta.a[0] * delta.a[0] + delta.a[1] * delta.a[1]);
  if (dn < kTol) return (hat(g) * B) * -1.0;
  V3 Bu = B * delta;
  return (so3_exp(Bu) * hat(g) * A_matrix(Bu).T() * B) * -1.0;
}

}  // namespace orc

// Manifold algebra of the esekfom mirror (product code; FP64; host and device).
// Mirrors what IKFoM's MTK provides for state_ikfom = (pos, SO3 rot, SO3 offset_R_L_I, offset_T_L_I,
// vel, bg, ba, S2 grav): exp/log on SO(3), the A_matrix Jacobian helper and the S2 charts
// (third_party/FAST_LIO/include/IKFoM_toolkit/mtk — absent submodule; SURVEY.md §9.6, IKFoM paper).
#pragma once
#include <cmath>
#include <cstring>

// the same FP64 code runs in esekf_host.cpp and inside the device-resident IEKF step (lio_iekf.cuh)
#ifdef __CUDACC__
#define ESEKF_HD __host__ __device__
#else
#define ESEKF_HD
#endif

namespace esekf {

constexpr int kDof = 23;
constexpr double kGravLen = 9.809;  // S2 radius used by state_ikfom's grav
constexpr double kTiny = 1e-11;

struct Vec3 {
  double v[3];
  ESEKF_HD double& operator[](int i) { return v[i]; }
  ESEKF_HD double operator[](int i) const { return v[i]; }
};
struct Mat3 {
  double m[9];  // row-major
  ESEKF_HD double& operator()(int r, int c) { return m[3 * r + c]; }
  ESEKF_HD double operator()(int r, int c) const { return m[3 * r + c]; }
};

ESEKF_HD inline Mat3 eye3() { return Mat3{{1, 0, 0, 0, 1, 0, 0, 0, 1}}; }
ESEKF_HD inline Mat3 mul(const Mat3& a, const Mat3& b) {
  Mat3 r;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      double s = 0;
      for (int k = 0; k < 3; ++k) s += a(i, k) * b(k, j);
      r(i, j) = s;
    }
  return r;
}
ESEKF_HD inline Vec3 mul(const Mat3& a, const Vec3& x) {
  Vec3 r;
  for (int i = 0; i < 3; ++i) r[i] = a(i, 0) * x[0] + a(i, 1) * x[1] + a(i, 2) * x[2];
  return r;
}
ESEKF_HD inline Mat3 transpose(const Mat3& a) {
  Mat3 r;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) r(j, i) = a(i, j);
  return r;
}
ESEKF_HD inline Mat3 scale(const Mat3& a, double s) {
  Mat3 r;
  for (int i = 0; i < 9; ++i) r.m[i] = a.m[i] * s;
  return r;
}
ESEKF_HD inline Mat3 add(const Mat3& a, const Mat3& b) {
  Mat3 r;
  for (int i = 0; i < 9; ++i) r.m[i] = a.m[i] + b.m[i];
  return r;
}
ESEKF_HD inline Mat3 skew(const Vec3& w) { return Mat3{{0, -w[2], w[1], w[2], 0, -w[0], -w[1], w[0], 0}}; }
ESEKF_HD inline double dot(const Vec3& a, const Vec3& b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
ESEKF_HD inline double norm(const Vec3& a) { return std::sqrt(dot(a, a)); }

// rotation vector -> rotation matrix (Rodrigues)
ESEKF_HD inline Mat3 so3_exp(const Vec3& w) {
  const double th = norm(w);
  if (th < kTiny) return add(eye3(), skew(w));
  const Vec3 k{{w[0] / th, w[1] / th, w[2] / th}};
  const Mat3 K = skew(k);
  return add(add(eye3(), scale(K, std::sin(th))), scale(mul(K, K), 1.0 - std::cos(th)));
}

// rotation matrix -> rotation vector
ESEKF_HD inline Vec3 so3_log(const Mat3& R) {
  double c = 0.5 * ((R(0, 0) + R(1, 1) + R(2, 2)) - 1.0);
  c = c > 1.0 ? 1.0 : (c < -1.0 ? -1.0 : c);
  const Vec3 v{{R(2, 1) - R(1, 2), R(0, 2) - R(2, 0), R(1, 0) - R(0, 1)}};
  const double s = 0.5 * norm(v);
  const double th = std::atan2(s, c);
  const double f = (s < kTiny) ? 0.5 : 0.5 * th / s;
  return Vec3{{v[0] * f, v[1] * f, v[2] * f}};
}

// MTK::A_matrix
ESEKF_HD inline Mat3 a_matrix(const Vec3& v) {
  const double sq = dot(v, v), n = std::sqrt(sq);
  if (n < kTiny) return eye3();
  const Mat3 K = skew(v);
  return add(add(eye3(), scale(K, (1.0 - std::cos(n)) / sq)), scale(mul(K, K), (1.0 - std::sin(n) / n) / sq));
}

// ---- S2 charts: 3x2 basis Bx, boxplus, boxminus, and the two Jacobian helpers Nx_yy / Mx -----------
struct Mat32 {
  double m[6];  // 3 rows x 2 cols
  ESEKF_HD double& operator()(int r, int c) { return m[2 * r + c]; }
  ESEKF_HD double operator()(int r, int c) const { return m[2 * r + c]; }
};

ESEKF_HD inline Mat32 s2_basis(const Vec3& g) {
  Mat32 B{{0, 0, 0, 0, 0, 0}};
  if (g[2] + kGravLen > kTiny) {
    const double d = kGravLen + g[2];
    B(0, 0) = (kGravLen - g[0] * g[0] / d) * (1.0 / kGravLen);
    B(0, 1) = (-g[0] * g[1] / d) * (1.0 / kGravLen);
    B(1, 0) = (-g[0] * g[1] / d) * (1.0 / kGravLen);
    B(1, 1) = (kGravLen - g[1] * g[1] / d) * (1.0 / kGravLen);
    B(2, 0) = (-g[0]) * (1.0 / kGravLen);
    B(2, 1) = (-g[1]) * (1.0 / kGravLen);
  } else {
    B(1, 1) = -1.0;
    B(2, 0) = 1.0;
  }
  return B;
}

ESEKF_HD inline Vec3 s2_boxplus(const Vec3& g, const double d[2]) {
  const Mat32 B = s2_basis(g);
  const Vec3 bu{{B(0, 0) * d[0] + B(0, 1) * d[1], B(1, 0) * d[0] + B(1, 1) * d[1], B(2, 0) * d[0] + B(2, 1) * d[1]}};
  return mul(so3_exp(bu), g);
}

ESEKF_HD inline void s2_boxminus(const Vec3& g, const Vec3& other, double out[2]) {
  out[0] = out[1] = 0.0;
  const Vec3 cr = mul(skew(g), other);
  const double v_sin = norm(cr), v_cos = dot(g, other);
  const double theta = std::atan2(v_sin, v_cos);
  if (v_sin < kTiny) {
    if (std::fabs(theta) > kTiny) out[0] = 3.1415926;
    return;
  }
  const Mat32 B = s2_basis(other);
  const Vec3 t = mul(skew(other), g);
  const double f = theta / v_sin;
  out[0] = (B(0, 0) * t[0] + B(1, 0) * t[1] + B(2, 0) * t[2]) * f;
  out[1] = (B(0, 1) * t[0] + B(1, 1) * t[1] + B(2, 1) * t[2]) * f;
}

// J(2x2) = Nx_yy(g_now) * Mx(g_ref, delta)
ESEKF_HD inline void s2_jacobian(const Vec3& g_now, const Vec3& g_ref, const double delta[2], double J[4]) {
  // Nx_yy = B(g_now)^T * skew(g_now) / len^2   (2x3)
  const Mat32 Bn = s2_basis(g_now);
  const Mat3 Sn = skew(g_now);
  double N[6];
  for (int r = 0; r < 2; ++r)
    for (int c = 0; c < 3; ++c) {
      double s = 0;
      for (int k = 0; k < 3; ++k) s += Bn(k, r) * Sn(k, c);
      N[3 * r + c] = s * (1.0 / kGravLen / kGravLen);
    }
  // Mx = -exp(B u) * skew(g_ref) * A(B u)^T * B   (3x2), or -skew(g_ref) * B for u ~ 0
  const Mat32 Br = s2_basis(g_ref);
  const Mat3 Sr = skew(g_ref);
  Mat3 left;
  if (std::sqrt(delta[0] * delta[0] + delta[1] * delta[1]) < kTiny) {
    left = Sr;
  } else {
    const Vec3 bu{{Br(0, 0) * delta[0] + Br(0, 1) * delta[1], Br(1, 0) * delta[0] + Br(1, 1) * delta[1],
                   Br(2, 0) * delta[0] + Br(2, 1) * delta[1]}};
    left = mul(mul(so3_exp(bu), Sr), transpose(a_matrix(bu)));
  }
  double M[6];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 2; ++c) {
      double s = 0;
      for (int k = 0; k < 3; ++k) s += left(r, k) * Br(k, c);
      M[2 * r + c] = -s;
    }
  for (int r = 0; r < 2; ++r)
    for (int c = 0; c < 2; ++c) {
      double s = 0;
      for (int k = 0; k < 3; ++k) s += N[3 * r + k] * M[2 * k + c];
      J[2 * r + c] = s;
    }
}

// ---- dense NxN helpers on row-major arrays ---------------------------------------------------------------
// in-place inverse by Gauss-Jordan with partial pivoting; returns false when singular
template <int N>
inline bool invert(const double* A_in, double* inv) {
  double A[N * N];
  std::memcpy(A, A_in, sizeof(A));
  for (int i = 0; i < N * N; ++i) inv[i] = 0.0;
  for (int i = 0; i < N; ++i) inv[i * N + i] = 1.0;
  for (int c = 0; c < N; ++c) {
    int piv = c;
    double best = std::fabs(A[c * N + c]);
    for (int r = c + 1; r < N; ++r)
      if (std::fabs(A[r * N + c]) > best) {
        best = std::fabs(A[r * N + c]);
        piv = r;
      }
    if (best == 0.0) return false;
    if (piv != c)
      for (int k = 0; k < N; ++k) {
        double t = A[c * N + k];
        A[c * N + k] = A[piv * N + k];
        A[piv * N + k] = t;
        t = inv[c * N + k];
        inv[c * N + k] = inv[piv * N + k];
        inv[piv * N + k] = t;
      }
    const double d = 1.0 / A[c * N + c];
    for (int k = 0; k < N; ++k) {
      A[c * N + k] *= d;
      inv[c * N + k] *= d;
    }
    for (int r = 0; r < N; ++r) {
      if (r == c) continue;
      const double f = A[r * N + c];
      if (f == 0.0) continue;
      for (int k = 0; k < N; ++k) {
        A[r * N + k] -= f * A[c * N + k];
        inv[r * N + k] -= f * inv[c * N + k];
      }
    }
  }
  return true;
}

}  // namespace esekf

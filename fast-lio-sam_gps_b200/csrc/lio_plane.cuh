// esti_plane + point-to-plane residual + Jacobian row — the per-point half of kernel K4.
// Reference: third_party/FAST_LIO/include/common_lib.h esti_plane<float> and the per-point body of
// h_share_model in src/laserMapping.cpp (absent submodule; behaviour per SURVEY.md §9.1-9.2).
// The f32 operation order is that of oracle/orc_lio.cpp (column-pivoted Householder QR of the 5x3
// system, solved for b = -1); with --fmad=false the selection mask is bit-identical to the oracle's.
#pragma once
#include <cfloat>
#include "lio_device.cuh"

namespace lio {

struct StateArg {
  double s[36];  // pos3 | rot 3x3 | offR 3x3 | offT 3 | vel3 | bg3 | ba3 | grav3   (LIO_STATE_DOUBLES)
};

// p_global = rot * (offR * p_body + offT) + pos, f64 math, stored as f32
__device__ __forceinline__ void body_to_world(const StateArg& st, float bx, float by, float bz, float& wx, float& wy, float& wz) {
  const double* s = st.s;
  const double b0 = bx, b1 = by, b2 = bz;
  double l[3], g[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) l[r] = ((s[12 + 3 * r] * b0 + s[12 + 3 * r + 1] * b1) + s[12 + 3 * r + 2] * b2) + s[21 + r];
#pragma unroll
  for (int r = 0; r < 3; ++r) g[r] = ((s[3 + 3 * r] * l[0] + s[3 + 3 * r + 1] * l[1]) + s[3 + 3 * r + 2] * l[2]) + s[r];
  wx = (float)g[0];
  wy = (float)g[1];
  wz = (float)g[2];
}

__device__ __forceinline__ void swapf(float& a, float& b) {
  const float t = a;
  a = b;
  b = t;
}

// P: 5 neighbour points.  Returns validity; pabcd = unit normal and offset.
__device__ __forceinline__ bool esti_plane(float pabcd[4], const float (&P)[kK][3], float threshold) {
  constexpr int R = kK, C = 3;
  float A[R][C];
  float c[R];
#pragma unroll
  for (int i = 0; i < R; ++i) {
    A[i][0] = P[i][0];
    A[i][1] = P[i][1];
    A[i][2] = P[i][2];
    c[i] = -1.0f;
  }
  float nu[C], nd[C], hc[C];
  int perm0 = 0, perm1 = 1, perm2 = 2;
#pragma unroll
  for (int j = 0; j < C; ++j) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < R; ++i) s = s + A[i][j] * A[i][j];
    nu[j] = nd[j] = sqrtf(s);
  }
  float maxn = nu[0];
  if (nu[1] > maxn) maxn = nu[1];
  if (nu[2] > maxn) maxn = nu[2];
  const float th = maxn * FLT_EPSILON;
  const float threshold_helper = (th * th) / (float)R;
  const float downdate_thr = sqrtf(FLT_EPSILON);
  int nz = C;
#pragma unroll
  for (int k = 0; k < C; ++k) {
    // pivot: first maximum of the remaining column norms
    int jb = k;
#pragma unroll
    for (int j = k + 1; j < C; ++j)
      if (nu[j] > ((jb == 0) ? nu[0] : (jb == 1) ? nu[1] : nu[2])) jb = j;
    const float nb = (jb == 0) ? nu[0] : (jb == 1) ? nu[1] : nu[2];
    const float big = nb * nb;
    if (nz == C && big < threshold_helper * (float)(R - k)) nz = k;
#pragma unroll
    for (int j = k + 1; j < C; ++j) {
      if (jb == j) {
#pragma unroll
        for (int i = 0; i < R; ++i) swapf(A[i][k], A[i][j]);
        swapf(nu[k], nu[j]);
        swapf(nd[k], nd[j]);
        // perm[k] <-> perm[j]
        int pk = (k == 0) ? perm0 : (k == 1) ? perm1 : perm2;
        int pj = (j == 1) ? perm1 : perm2;
        if (k == 0) perm0 = pj; else if (k == 1) perm1 = pj;
        if (j == 1) perm1 = pk; else perm2 = pk;
      }
    }
    float tail = 0.f;
#pragma unroll
    for (int i = k + 1; i < R; ++i) tail = tail + A[i][k] * A[i][k];
    const float c0 = A[k][k];
    float beta, tau;
    if (tail <= FLT_MIN) {
      tau = 0.f;
      beta = c0;
#pragma unroll
      for (int i = k + 1; i < R; ++i) A[i][k] = 0.f;
    } else {
      beta = sqrtf(c0 * c0 + tail);
      if (c0 >= 0.f) beta = -beta;
      const float den = c0 - beta;
#pragma unroll
      for (int i = k + 1; i < R; ++i) A[i][k] = A[i][k] / den;
      tau = (beta - c0) / beta;
    }
    A[k][k] = beta;
    hc[k] = tau;
#pragma unroll
    for (int j = k + 1; j < C; ++j) {
      float tmp = 0.f;
#pragma unroll
      for (int i = k + 1; i < R; ++i) tmp = tmp + A[i][k] * A[i][j];
      tmp = tmp + A[k][j];
      const float tt = tau * tmp;
      A[k][j] = A[k][j] - tt;
#pragma unroll
      for (int i = k + 1; i < R; ++i) A[i][j] = A[i][j] - tt * A[i][k];
    }
#pragma unroll
    for (int j = k + 1; j < C; ++j) {
      if (nu[j] != 0.f) {
        float t = fabsf(A[k][j]) / nu[j];
        t = (1.f + t) * (1.f - t);
        if (t < 0.f) t = 0.f;
        const float ratio = nu[j] / nd[j];
        const float t2 = t * (ratio * ratio);
        if (t2 <= downdate_thr) {
          float s = 0.f;
#pragma unroll
          for (int i = k + 1; i < R; ++i) s = s + A[i][j] * A[i][j];
          nd[j] = sqrtf(s);
          nu[j] = nd[j];
        } else {
          nu[j] = nu[j] * sqrtf(t);
        }
      }
    }
  }
  // c = Q^T b
#pragma unroll
  for (int k = 0; k < C; ++k) {
    if (k < nz) {
      float tmp = 0.f;
#pragma unroll
      for (int i = k + 1; i < R; ++i) tmp = tmp + A[i][k] * c[i];
      tmp = tmp + c[k];
      const float tt = hc[k] * tmp;
      c[k] = c[k] - tt;
#pragma unroll
      for (int i = k + 1; i < R; ++i) c[i] = c[i] - tt * A[i][k];
    }
  }
  // back substitution on the leading nz x nz triangle
#pragma unroll
  for (int i = C - 1; i >= 0; --i) {
    if (i < nz) {
      float s = c[i];
#pragma unroll
      for (int j = i + 1; j < C; ++j)
        if (j < nz) s = s - A[i][j] * c[j];
      c[i] = s / A[i][i];
    }
  }
  float x0 = 0.f, x1 = 0.f, x2 = 0.f;
  {
    const float v0 = (0 < nz) ? c[0] : 0.f, v1 = (1 < nz) ? c[1] : 0.f, v2 = (2 < nz) ? c[2] : 0.f;
    const bool h0 = 0 < nz, h1 = 1 < nz, h2 = 2 < nz;
    if (h0) { if (perm0 == 0) x0 = v0; else if (perm0 == 1) x1 = v0; else x2 = v0; }
    if (h1) { if (perm1 == 0) x0 = v1; else if (perm1 == 1) x1 = v1; else x2 = v1; }
    if (h2) { if (perm2 == 0) x0 = v2; else if (perm2 == 1) x1 = v2; else x2 = v2; }
  }
  float n2 = x0 * x0 + x1 * x1;
  n2 = n2 + x2 * x2;
  const float n = sqrtf(n2);
  pabcd[0] = x0 / n;
  pabcd[1] = x1 / n;
  pabcd[2] = x2 / n;
  pabcd[3] = 1.0f / n;
  bool ok = true;
#pragma unroll
  for (int j = 0; j < R; ++j) {
    float v = pabcd[0] * P[j][0] + pabcd[1] * P[j][1];
    v = v + pabcd[2] * P[j][2];
    v = v + pabcd[3];
    if (!(fabsf(v) <= threshold)) ok = false;  // also rejects NaN from degenerate fits
  }
  return ok;
}

}  // namespace lio

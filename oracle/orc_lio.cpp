// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).  PARITY UNPINNED.
//
// CPU restatement (C++17 + OpenMP, no third-party deps) of the FAST-LIO IEKF
// measurement path.  The reference source for this half
// (third_party/FAST_LIO/src/laserMapping.cpp: h_share_model, map_incremental,
// lasermap_fov_segment; include/common_lib.h: esti_plane;
// include/IKFoM_toolkit/esekfom/esekfom.hpp: update_iterated_dyn_share_modified)
// is an un-fetched submodule (/root/reference/.SUBMODULES.json:9-15), so every
// function below follows SURVEY.md §9.1-9.3/9.6 [RECALL] and the published
// FAST-LIO2 / IKFoM / ikd-Tree papers, not a file:line.  Parameter names and
// values that ARE verified come from third_party/fastlio_config_launch
// (kitti.launch:6-12: max_iteration 3, filter_size_surf/map 0.5, cube_side_length 1000;
//  kitti.yaml:22: extrinsic_est_en false).
//
// Compile with -ffp-contract=off: the f32 arithmetic order below is mirrored
// operation-for-operation by the CUDA kernels (built with --fmad=false) so that
// kNN sets and plane-validity masks can be compared bit-exactly.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <cfloat>
#include <cmath>
#include <vector>
#include <omp.h>
#include "orc_linalg.hpp"
#include "orc_map.hpp"

using namespace orc;

namespace {

constexpr int kK = 5;  // NUM_MATCH_POINTS

// ---- esti_plane: A(5x3) n = -1 via column-pivoted Householder QR, all f32 ------------
// Restates Eigen's ColPivHouseholderQR::computeInPlace + solve with a fixed, sequential
// operation order (Eigen's own order is SIMD-dependent and cannot be pinned).
bool esti_plane(float pabcd[4], const float (*P)[3], float threshold) {
  const int R = kK, C = 3;
  float A[kK][3];
  float c[kK];
  for (int i = 0; i < R; ++i) { A[i][0] = P[i][0]; A[i][1] = P[i][1]; A[i][2] = P[i][2]; c[i] = -1.0f; }
  float nu[3], nd[3], hc[3];
  int perm[3] = {0, 1, 2};
  for (int j = 0; j < C; ++j) {
    float s = 0.f;
    for (int i = 0; i < R; ++i) s = s + A[i][j] * A[i][j];
    nu[j] = nd[j] = std::sqrt(s);
  }
  float maxn = nu[0];
  if (nu[1] > maxn) maxn = nu[1];
  if (nu[2] > maxn) maxn = nu[2];
  float th = maxn * FLT_EPSILON;
  const float threshold_helper = (th * th) / (float)R;
  const float downdate_thr = std::sqrt(FLT_EPSILON);
  int nz = C;
  for (int k = 0; k < C; ++k) {
    int jb = k;
    for (int j = k + 1; j < C; ++j)
      if (nu[j] > nu[jb]) jb = j;
    float big = nu[jb] * nu[jb];
    if (nz == C && big < threshold_helper * (float)(R - k)) nz = k;
    if (jb != k) {
      for (int i = 0; i < R; ++i) { float t = A[i][k]; A[i][k] = A[i][jb]; A[i][jb] = t; }
      float t = nu[k]; nu[k] = nu[jb]; nu[jb] = t;
      t = nd[k]; nd[k] = nd[jb]; nd[jb] = t;
      int ti = perm[k]; perm[k] = perm[jb]; perm[jb] = ti;
    }
    // Householder vector for A[k..R-1][k]
    float tail = 0.f;
    for (int i = k + 1; i < R; ++i) tail = tail + A[i][k] * A[i][k];
    float c0 = A[k][k], beta, tau;
    if (tail <= FLT_MIN) {
      tau = 0.f; beta = c0;
      for (int i = k + 1; i < R; ++i) A[i][k] = 0.f;
    } else {
      beta = std::sqrt(c0 * c0 + tail);
      if (c0 >= 0.f) beta = -beta;
      float den = c0 - beta;
      for (int i = k + 1; i < R; ++i) A[i][k] = A[i][k] / den;
      tau = (beta - c0) / beta;
    }
    A[k][k] = beta; hc[k] = tau;
    // apply H_k to the trailing columns
    for (int j = k + 1; j < C; ++j) {
      float tmp = 0.f;
      for (int i = k + 1; i < R; ++i) tmp = tmp + A[i][k] * A[i][j];
      tmp = tmp + A[k][j];
      float tt = tau * tmp;
      A[k][j] = A[k][j] - tt;
      for (int i = k + 1; i < R; ++i) A[i][j] = A[i][j] - tt * A[i][k];
    }
    // column-norm downdate
    for (int j = k + 1; j < C; ++j) {
      if (nu[j] != 0.f) {
        float t = std::fabs(A[k][j]) / nu[j];
        t = (1.f + t) * (1.f - t);
        if (t < 0.f) t = 0.f;
        float ratio = nu[j] / nd[j];
        float t2 = t * (ratio * ratio);
        if (t2 <= downdate_thr) {
          float s = 0.f;
          for (int i = k + 1; i < R; ++i) s = s + A[i][j] * A[i][j];
          nd[j] = std::sqrt(s); nu[j] = nd[j];
        } else {
          nu[j] = nu[j] * std::sqrt(t);
        }
      }
    }
  }
  // c = Q^T b
  for (int k = 0; k < nz; ++k) {
    float tmp = 0.f;
    for (int i = k + 1; i < R; ++i) tmp = tmp + A[i][k] * c[i];
    tmp = tmp + c[k];
    float tt = hc[k] * tmp;
    c[k] = c[k] - tt;
    for (int i = k + 1; i < R; ++i) c[i] = c[i] - tt * A[i][k];
  }
  // back substitution on the leading nz x nz triangle
  for (int i = nz - 1; i >= 0; --i) {
    float s = c[i];
    for (int j = i + 1; j < nz; ++j) s = s - A[i][j] * c[j];
    c[i] = s / A[i][i];
  }
  float x[3] = {0.f, 0.f, 0.f};
  for (int i = 0; i < nz; ++i) x[perm[i]] = c[i];
  float n2 = x[0] * x[0] + x[1] * x[1];
  n2 = n2 + x[2] * x[2];
  float n = std::sqrt(n2);
  pabcd[0] = x[0] / n; pabcd[1] = x[1] / n; pabcd[2] = x[2] / n; pabcd[3] = 1.0f / n;
  for (int j = 0; j < R; ++j) {
    float v = pabcd[0] * P[j][0] + pabcd[1] * P[j][1];
    v = v + pabcd[2] * P[j][2];
    v = v + pabcd[3];
    if (!(std::fabs(v) <= threshold)) return false;  // also rejects NaN from degenerate fits
  }
  return true;
}

// state36 layout: pos[0:3] rot[3:12] offR[12:21] offT[21:24] vel[24:27] bg[27:30] ba[30:33] grav[33:36]
struct St {
  V3 pos, offT, vel, bg, ba, grav;
  M3 rot, offR;
};
St unpack(const double* s) {
  St x;
  for (int i = 0; i < 3; ++i) { x.pos.a[i] = s[i]; x.offT.a[i] = s[21 + i]; x.vel.a[i] = s[24 + i]; x.bg.a[i] = s[27 + i]; x.ba.a[i] = s[30 + i]; x.grav.a[i] = s[33 + i]; }
  for (int i = 0; i < 9; ++i) { x.rot.a[i] = s[3 + i]; x.offR.a[i] = s[12 + i]; }
  return x;
}
void pack(const St& x, double* s) {
  for (int i = 0; i < 3; ++i) { s[i] = x.pos.a[i]; s[21 + i] = x.offT.a[i]; s[24 + i] = x.vel.a[i]; s[27 + i] = x.bg.a[i]; s[30 + i] = x.ba.a[i]; s[33 + i] = x.grav.a[i]; }
  for (int i = 0; i < 9; ++i) { s[3 + i] = x.rot.a[i]; s[12 + i] = x.offR.a[i]; }
}

// p_global = rot * (offR * p_body + offT) + pos, f64, stored to f32
inline void body_to_world(const St& s, const float* pb, float* pw) {
  double b0 = pb[0], b1 = pb[1], b2 = pb[2];
  double l[3], g[3];
  for (int r = 0; r < 3; ++r) l[r] = ((s.offR(r, 0) * b0 + s.offR(r, 1) * b1) + s.offR(r, 2) * b2) + s.offT.a[r];
  for (int r = 0; r < 3; ++r) g[r] = ((s.rot(r, 0) * l[0] + s.rot(r, 1) * l[1]) + s.rot(r, 2) * l[2]) + s.pos.a[r];
  pw[0] = (float)g[0]; pw[1] = (float)g[1]; pw[2] = (float)g[2];
}

struct Scan {
  int n = 0;
  std::vector<float> body;    // n x 3
  std::vector<float> world;   // n x 3
  std::vector<float> nn;      // n x 5 x 4  (xyz, id bits)
  std::vector<int> nn_cnt;    // n
  std::vector<uint8_t> sel;   // point_selected_surf
  std::vector<float> plane;   // n x 4: a b c pd2
  // Nearest_Points[i][0] of the points whose 5-NN search found nothing within max_d2: upstream's search is unbounded, so the
  // cached list still starts with the (far) nearest map point, and map_incremental looks at it.  far[i] = xyz + id bits,
  // id 0xFFFFFFFF: none (empty map, or no search ran for this scan)
  std::vector<float> far;     // n x 4
};

struct Ctx {
  KdMap map;
  Scan scan;
  int nthreads = 1;
  bool extrinsic = false;
  float max_d2 = 5.0f;
  // local-map box (lasermap_fov_segment)
  bool lm_init = false;
  float lm_min[3], lm_max[3];
  double cube_len = 1000.0;
  float det_range = 300.0f;
  explicit Ctx(float ds) : map(ds) {}
};

// One call of h_share_model.  Outputs: HTH (12x12 row-major), HTz(12), and optionally the
// compacted h_x (n_eff x 12) / h (n_eff) exactly as dyn_share_datastruct would carry them.
int h_share_model(Ctx& c, const St& s, bool converge, double* HTH, double* HTz, double* res_sum,
                  double* hx_out, double* h_out) {
  Scan& sc = c.scan;
  const int n = sc.n;
#pragma omp parallel for num_threads(c.nthreads) schedule(static)
  for (int i = 0; i < n; ++i) {
    const float* pb = &sc.body[3 * i];
    float* pw = &sc.world[3 * i];
    body_to_world(s, pb, pw);
    if (converge) {
      KnnHit hits[kK];
      int cnt = c.map.knn(pw, kK, c.max_d2, hits);
      sc.nn_cnt[i] = cnt;
      {
        float* f = &sc.far[size_t(i) * 4];
        uint32_t none = 0xFFFFFFFFu;
        std::memcpy(&f[3], &none, 4);
        if (cnt == 0) {  // the unbounded search's first neighbour (only map_incremental reads it)
          KnnHit far_hit[1];
          if (c.map.knn(pw, 1, INFINITY, far_hit) == 1) {
            f[0] = far_hit[0].x; f[1] = far_hit[0].y; f[2] = far_hit[0].z;
            std::memcpy(&f[3], &far_hit[0].id, 4);
          }
        }
      }
      for (int k = 0; k < cnt; ++k) {
        float* d = &sc.nn[(size_t(i) * kK + k) * 4];
        d[0] = hits[k].x; d[1] = hits[k].y; d[2] = hits[k].z;
        std::memcpy(&d[3], &hits[k].id, 4);
      }
      // reference: point_selected_surf[i] = (points_near.size() >= 5) && (sqdist[4] <= 5) with an
      // unbounded search; searching with max_d2 = 5 gives the identical flag.
      sc.sel[i] = cnt >= kK ? 1 : 0;
    }
    // point_selected_surf[i] is a file-scope flag upstream: it is re-derived from the
    // search only when ekfom_data.converge is set; on the calls that reuse the cached neighbours a point
    // that failed esti_plane or the weight test in the PREVIOUS call stays out ("if (!selected) continue"),
    // so the selected set can only shrink between two searches.
    if (!sc.sel[i]) continue;
    sc.sel[i] = 0;
    float P[kK][3];
    for (int k = 0; k < kK; ++k) {
      const float* d = &sc.nn[(size_t(i) * kK + k) * 4];
      P[k][0] = d[0]; P[k][1] = d[1]; P[k][2] = d[2];
    }
    float pabcd[4];
    if (esti_plane(pabcd, P, 0.1f)) {
      float pd2 = pabcd[0] * pw[0] + pabcd[1] * pw[1];
      pd2 = pd2 + pabcd[2] * pw[2];
      pd2 = pd2 + pabcd[3];
      double bx = pb[0], by = pb[1], bz = pb[2];
      double range = std::sqrt((bx * bx + by * by) + bz * bz);
      float sc_w = (float)(1.0 - 0.9 * std::fabs((double)pd2) / std::sqrt(range));
      if ((double)sc_w > 0.9) {
        sc.sel[i] = 1;
        float* pl = &sc.plane[4 * i];
        pl[0] = pabcd[0]; pl[1] = pabcd[1]; pl[2] = pabcd[2]; pl[3] = pd2;
      }
    }
  }
  // serial compaction + Jacobian (reference stages 2 and 3)
  for (int i = 0; i < 144; ++i) HTH[i] = 0.0;
  for (int i = 0; i < 12; ++i) HTz[i] = 0.0;
  double rs = 0.0;
  int n_eff = 0;
  M3 rotT = s.rot.T(), offRT = s.offR.T();
  for (int i = 0; i < n; ++i) {
    if (!sc.sel[i]) continue;
    const float* pb = &sc.body[3 * i];
    const float* pl = &sc.plane[4 * i];
    V3 pbe = vec3(pb[0], pb[1], pb[2]);
    V3 pthis;
    for (int r = 0; r < 3; ++r) pthis.a[r] = ((s.offR(r, 0) * pbe.a[0] + s.offR(r, 1) * pbe.a[1]) + s.offR(r, 2) * pbe.a[2]) + s.offT.a[r];
    V3 nv = vec3(pl[0], pl[1], pl[2]);
    V3 Cv;
    for (int r = 0; r < 3; ++r) Cv.a[r] = (rotT(r, 0) * nv.a[0] + rotT(r, 1) * nv.a[1]) + rotT(r, 2) * nv.a[2];
    // A = [point_this]x * C
    V3 Av = vec3(pthis.a[1] * Cv.a[2] - pthis.a[2] * Cv.a[1], pthis.a[2] * Cv.a[0] - pthis.a[0] * Cv.a[2],
                 pthis.a[0] * Cv.a[1] - pthis.a[1] * Cv.a[0]);
    double row[12] = {nv.a[0], nv.a[1], nv.a[2], Av.a[0], Av.a[1], Av.a[2], 0, 0, 0, 0, 0, 0};
    if (c.extrinsic) {
      V3 D;
      for (int r = 0; r < 3; ++r) D.a[r] = (offRT(r, 0) * Cv.a[0] + offRT(r, 1) * Cv.a[1]) + offRT(r, 2) * Cv.a[2];
      V3 Bv = vec3(pbe.a[1] * D.a[2] - pbe.a[2] * D.a[1], pbe.a[2] * D.a[0] - pbe.a[0] * D.a[2],
                   pbe.a[0] * D.a[1] - pbe.a[1] * D.a[0]);
      for (int r = 0; r < 3; ++r) { row[6 + r] = Bv.a[r]; row[9 + r] = Cv.a[r]; }
    }
    double hz = -(double)pl[3];
    for (int a = 0; a < 12; ++a) {
      for (int b = 0; b < 12; ++b) HTH[a * 12 + b] += row[a] * row[b];
      HTz[a] += row[a] * hz;
    }
    rs += std::fabs((double)pl[3]);
    if (hx_out) for (int a = 0; a < 12; ++a) hx_out[size_t(n_eff) * 12 + a] = row[a];
    if (h_out) h_out[n_eff] = hz;
    ++n_eff;
  }
  if (res_sum) *res_sum = rs;
  return n_eff;
}

// ---- esekfom: 23-DoF error state on (pos, SO3, SO3, vec3, vec3, vec3, vec3, S2) ---------
using V23 = Mat<23, 1>;
using M23 = Mat<23, 23>;

V23 boxminus(const St& x, const St& o) {
  V23 d;
  V3 r = so3_log(o.rot.T() * x.rot), r2 = so3_log(o.offR.T() * x.offR);
  Mat<2, 1> g = S2_boxminus(x.grav, o.grav);
  for (int i = 0; i < 3; ++i) {
    d.a[i] = x.pos.a[i] - o.pos.a[i]; d.a[3 + i] = r.a[i]; d.a[6 + i] = r2.a[i];
    d.a[9 + i] = x.offT.a[i] - o.offT.a[i]; d.a[12 + i] = x.vel.a[i] - o.vel.a[i];
    d.a[15 + i] = x.bg.a[i] - o.bg.a[i]; d.a[18 + i] = x.ba.a[i] - o.ba.a[i];
  }
  d.a[21] = g.a[0]; d.a[22] = g.a[1];
  return d;
}

void boxplus(St& x, const V23& d) {
  for (int i = 0; i < 3; ++i) {
    x.pos.a[i] += d.a[i]; x.offT.a[i] += d.a[9 + i]; x.vel.a[i] += d.a[12 + i];
    x.bg.a[i] += d.a[15 + i]; x.ba.a[i] += d.a[18 + i];
  }
  x.rot = x.rot * so3_exp(vec3(d.a[3], d.a[4], d.a[5]));
  x.offR = x.offR * so3_exp(vec3(d.a[6], d.a[7], d.a[8]));
  Mat<2, 1> g; g.a[0] = d.a[21]; g.a[1] = d.a[22];
  x.grav = S2_boxplus(x.grav, g);
}

// left-multiply rows [idx, idx+D) and right-multiply cols by J / J^T
template <int D>
void project_rows(M23& P, int idx, const Mat<D, D>& J) {
  for (int c = 0; c < 23; ++c) {
    double t[D];
    for (int r = 0; r < D; ++r) { t[r] = 0; for (int k = 0; k < D; ++k) t[r] += J(r, k) * P(idx + k, c); }
    for (int r = 0; r < D; ++r) P(idx + r, c) = t[r];
  }
}
template <int D>
void project_cols(M23& P, int idx, const Mat<D, D>& J) {
  for (int r = 0; r < 23; ++r) {
    double t[D];
    for (int c = 0; c < D; ++c) { t[c] = 0; for (int k = 0; k < D; ++k) t[c] += P(r, idx + k) * J(c, k); }
    for (int c = 0; c < D; ++c) P(r, idx + c) = t[c];
  }
}

// update_iterated_dyn_share_modified.  iter_states: (max_iter+1) x 36 trace of x_ after
// every boxplus; returns the number of iterations executed.
int iekf_update(Ctx& c, St& x, M23& P, double R, int max_iter, const double* limit, double* iter_states,
                int* n_eff_trace, int* knn_trace) {
  bool valid = true, converge = true;
  int t = 0, done = 0;
  St x_prop = x;
  M23 P_prop = P;
  V23 dx_new;
  std::vector<double> hx, hv;
  for (int i = -1; i < max_iter; ++i) {
    valid = true;
    double HTH[144], HTz[12];
    hx.assign(size_t(c.scan.n) * 12, 0.0); hv.assign(c.scan.n, 0.0);
    if (knn_trace) knn_trace[done] = converge ? 1 : 0;
    int n_eff = h_share_model(c, x, converge, HTH, HTz, nullptr, hx.data(), hv.data());
    if (n_eff < 1) valid = false;
    if (n_eff_trace) n_eff_trace[done] = n_eff;
    if (!valid) { if (iter_states) pack(x, iter_states + 36 * done); ++done; continue; }
    V23 dx = boxminus(x, x_prop);
    dx_new = dx;
    P = P_prop;
    for (int idx : {3, 6}) {
      M3 J = A_matrix(vec3(dx.a[idx], dx.a[idx + 1], dx.a[idx + 2])).T();
      V3 seg = J * vec3(dx_new.a[idx], dx_new.a[idx + 1], dx_new.a[idx + 2]);
      for (int k = 0; k < 3; ++k) dx_new.a[idx + k] = seg.a[k];
      project_rows<3>(P, idx, J);
      project_cols<3>(P, idx, J);
    }
    {
      Mat<2, 1> seg; seg.a[0] = dx.a[21]; seg.a[1] = dx.a[22];
      Mat<2, 2> J = S2_Nx_yy(x.grav) * S2_Mx(x_prop.grav, seg);
      Mat<2, 1> dn; dn.a[0] = dx_new.a[21]; dn.a[1] = dx_new.a[22];
      dn = J * dn;
      dx_new.a[21] = dn.a[0]; dx_new.a[22] = dn.a[1];
      project_rows<2>(P, 21, J);
      project_cols<2>(P, 21, J);
    }
    V23 K_h;
    M23 K_x;
    if (23 > n_eff) {
      // K = P H^T (H P H^T / R + I)^-1 / R  with H = [h_x 0], solved through the 23x23 information form,
      // which is algebraically identical and keeps a single code path for tiny scans
      M23 Pt = inverse(P * (1.0 / R));
      for (int a = 0; a < 12; ++a) for (int b = 0; b < 12; ++b) Pt(a, b) += HTH[a * 12 + b];
      M23 Pi = inverse(Pt);
      for (int r = 0; r < 23; ++r) {
        double s = 0; for (int k = 0; k < 12; ++k) s += Pi(r, k) * HTz[k];
        K_h.a[r] = s;
        for (int cc = 0; cc < 12; ++cc) { double q = 0; for (int k = 0; k < 12; ++k) q += Pi(r, k) * HTH[k * 12 + cc]; K_x(r, cc) = q; }
      }
    } else {
      M23 Pt = inverse(P * (1.0 / R));
      for (int a = 0; a < 12; ++a) for (int b = 0; b < 12; ++b) Pt(a, b) += HTH[a * 12 + b];
      M23 Pi = inverse(Pt);
      for (int r = 0; r < 23; ++r) {
        double s = 0; for (int k = 0; k < 12; ++k) s += Pi(r, k) * HTz[k];
        K_h.a[r] = s;
        for (int cc = 0; cc < 12; ++cc) { double q = 0; for (int k = 0; k < 12; ++k) q += Pi(r, k) * HTH[k * 12 + cc]; K_x(r, cc) = q; }
      }
    }
    V23 dx_;
    for (int r = 0; r < 23; ++r) {
      double s = K_h.a[r];
      for (int k = 0; k < 23; ++k) s += (K_x(r, k) - (r == k ? 1.0 : 0.0)) * dx_new.a[k];
      dx_.a[r] = s;
    }
    boxplus(x, dx_);
    if (iter_states) pack(x, iter_states + 36 * done);
    ++done;
    converge = true;
    for (int k = 0; k < 23; ++k)
      if (std::fabs(dx_.a[k]) > limit[k]) { converge = false; break; }
    if (converge) ++t;
    if (!t && i == max_iter - 2) converge = true;
    if (t > 1 || i == max_iter - 1) {
      M23 L = P;
      for (int idx : {3, 6}) {
        M3 J = A_matrix(vec3(dx_.a[idx], dx_.a[idx + 1], dx_.a[idx + 2])).T();
        project_rows<3>(L, idx, J);
        // K_x rows
        for (int cc = 0; cc < 12; ++cc) {
          double tt[3];
          for (int r = 0; r < 3; ++r) { tt[r] = 0; for (int k = 0; k < 3; ++k) tt[r] += J(r, k) * K_x(idx + k, cc); }
          for (int r = 0; r < 3; ++r) K_x(idx + r, cc) = tt[r];
        }
        project_cols<3>(L, idx, J);
        project_cols<3>(P, idx, J);
      }
      {
        Mat<2, 1> seg; seg.a[0] = dx_.a[21]; seg.a[1] = dx_.a[22];
        // [RECALL, unverified: esekfom.hpp is absent] upstream calls x_propagated.S2_Mx(...) here too; its x_before is unused
        Mat<2, 2> J = S2_Nx_yy(x.grav) * S2_Mx(x_prop.grav, seg);
        project_rows<2>(L, 21, J);
        for (int cc = 0; cc < 12; ++cc) {
          double tt[2];
          for (int r = 0; r < 2; ++r) { tt[r] = 0; for (int k = 0; k < 2; ++k) tt[r] += J(r, k) * K_x(21 + k, cc); }
          for (int r = 0; r < 2; ++r) K_x(21 + r, cc) = tt[r];
        }
        project_cols<2>(L, 21, J);
        project_cols<2>(P, 21, J);
      }
      M23 Pn;
      for (int r = 0; r < 23; ++r)
        for (int cc = 0; cc < 23; ++cc) {
          double s = 0; for (int k = 0; k < 12; ++k) s += K_x(r, k) * P(k, cc);
          Pn(r, cc) = L(r, cc) - s;
        }
      P = Pn;
      return done;
    }
  }
  return done;
}

// map_incremental: classify every scan point (at the FINAL state) into PointToAdd /
// PointNoNeedDownsample / skip using its cached neighbours, then Add_Points twice.
void map_incremental(Ctx& c, const St& s, int flg_ekf_inited, int* n_add, int* n_nodown) {
  Scan& sc = c.scan;
  const float fs = c.map.ds();
  std::vector<float> to_add, no_down;
  for (int i = 0; i < sc.n; ++i) {
    float* pw = &sc.world[3 * i];
    body_to_world(s, &sc.body[3 * i], pw);
    // Upstream: `if (!Nearest_Points[i].empty() && flg_EKF_inited)`.  Its search is unbounded, so the cached list is empty
    // only when no search ran (or the map was empty); a point with no neighbour within max_d2 still has its far nearest
    // neighbour at the head of the list (sc.far), which decides between the two lists like any other points_near[0].
    const float* n0 = nullptr;
    if (sc.nn_cnt[i] > 0) n0 = &sc.nn[size_t(i) * kK * 4];
    else {
      uint32_t fid;
      std::memcpy(&fid, &sc.far[size_t(i) * 4 + 3], 4);
      if (fid != 0xFFFFFFFFu) n0 = &sc.far[size_t(i) * 4];
    }
    if (n0 && flg_ekf_inited) {
      float mid[3];
      for (int a = 0; a < 3; ++a) mid[a] = std::floor(pw[a] / fs) * fs + 0.5f * fs;
      float dist = sqdist3(pw[0], pw[1], pw[2], mid[0], mid[1], mid[2]);
      if (std::fabs(n0[0] - mid[0]) > 0.5f * fs && std::fabs(n0[1] - mid[1]) > 0.5f * fs && std::fabs(n0[2] - mid[2]) > 0.5f * fs) {
        no_down.insert(no_down.end(), pw, pw + 3);
        continue;
      }
      bool need_add = true;
      // reference breaks out when points_near.size() < 5, which with an unbounded search only happens when the whole map
      // holds < 5 points; neighbours beyond max_d2 can never be closer to the voxel centre than the point itself, so
      // scanning the ones found within max_d2 is equivalent.
      if (c.map.size() >= kK) {
        for (int k = 0; k < sc.nn_cnt[i]; ++k) {
          const float* q = &sc.nn[(size_t(i) * kK + k) * 4];
          if (sqdist3(q[0], q[1], q[2], mid[0], mid[1], mid[2]) < dist) { need_add = false; break; }
        }
      }
      if (need_add) to_add.insert(to_add.end(), pw, pw + 3);
    } else {
      to_add.insert(to_add.end(), pw, pw + 3);
    }
  }
  c.map.add_points(to_add.data(), (int)to_add.size() / 3, true, nullptr);
  c.map.add_points(no_down.data(), (int)no_down.size() / 3, false, nullptr);
  if (n_add) *n_add = (int)to_add.size() / 3;
  if (n_nodown) *n_nodown = (int)no_down.size() / 3;
}

// lasermap_fov_segment: slide the local-map cube, emit the vacated slabs as delete boxes.
int fov_segment(Ctx& c, const double pos_lid[3], float* boxes_out /* up to 3x6 */) {
  const float MOV_THRESHOLD = 1.5f;
  if (!c.lm_init) {
    for (int i = 0; i < 3; ++i) { c.lm_min[i] = (float)(pos_lid[i] - c.cube_len / 2.0); c.lm_max[i] = (float)(pos_lid[i] + c.cube_len / 2.0); }
    c.lm_init = true;
    return 0;
  }
  float d[3][2];
  bool need_move = false;
  for (int i = 0; i < 3; ++i) {
    d[i][0] = (float)std::fabs(pos_lid[i] - c.lm_min[i]);
    d[i][1] = (float)std::fabs(pos_lid[i] - c.lm_max[i]);
    if (d[i][0] <= MOV_THRESHOLD * c.det_range || d[i][1] <= MOV_THRESHOLD * c.det_range) need_move = true;
  }
  if (!need_move) return 0;
  float nmin[3], nmax[3];
  for (int i = 0; i < 3; ++i) { nmin[i] = c.lm_min[i]; nmax[i] = c.lm_max[i]; }
  float mov = (float)std::max((c.cube_len - 2.0 * MOV_THRESHOLD * c.det_range) * 0.5 * 0.9, (double)(c.det_range * (MOV_THRESHOLD - 1)));
  int nb = 0;
  for (int i = 0; i < 3; ++i) {
    float b[6] = {c.lm_min[0], c.lm_min[1], c.lm_min[2], c.lm_max[0], c.lm_max[1], c.lm_max[2]};
    if (d[i][0] <= MOV_THRESHOLD * c.det_range) {
      nmax[i] -= mov; nmin[i] -= mov;
      b[i] = c.lm_max[i] - mov;
      std::memcpy(boxes_out + 6 * nb++, b, sizeof(b));
    } else if (d[i][1] <= MOV_THRESHOLD * c.det_range) {
      nmax[i] += mov; nmin[i] += mov;
      b[3 + i] = c.lm_min[i] + mov;
      std::memcpy(boxes_out + 6 * nb++, b, sizeof(b));
    }
  }
  for (int i = 0; i < 3; ++i) { c.lm_min[i] = nmin[i]; c.lm_max[i] = nmax[i]; }
  if (nb) c.map.delete_boxes(boxes_out, nb);
  return nb;
}

}  // namespace

// ------------------------------------------------------------------ C API (ctypes)
extern "C" {

void* orc_ctx_create(float ds, int nthreads) {
  Ctx* c = new Ctx(ds);
  c->nthreads = nthreads > 0 ? nthreads : omp_get_max_threads();
  return c;
}
void orc_ctx_destroy(void* h) { delete (Ctx*)h; }
void orc_ctx_set(void* h, int extrinsic, float max_d2, double cube_len, float det_range) {
  Ctx* c = (Ctx*)h;
  c->extrinsic = extrinsic != 0; c->max_d2 = max_d2; c->cube_len = cube_len; c->det_range = det_range;
}
int orc_max_threads() { return omp_get_max_threads(); }

void orc_map_build(void* h, const float* xyz, int n) { ((Ctx*)h)->map.build(xyz, n); }
int orc_map_add(void* h, const float* xyz, int n, int downsample, int* net_new) {
  return ((Ctx*)h)->map.add_points(xyz, n, downsample != 0, net_new);
}
int orc_map_delete_boxes(void* h, const float* boxes, int nb) { return ((Ctx*)h)->map.delete_boxes(boxes, nb); }
int orc_map_size(void* h) { return ((Ctx*)h)->map.size(); }
int orc_map_export(void* h, float* xyz, uint32_t* ids, int cap) {
  std::vector<OPt> pts;
  ((Ctx*)h)->map.flatten(pts);
  int n = std::min<int>(cap, (int)pts.size());
  for (int i = 0; i < n; ++i) { xyz[3 * i] = pts[i].x; xyz[3 * i + 1] = pts[i].y; xyz[3 * i + 2] = pts[i].z; ids[i] = pts[i].id; }
  return (int)pts.size();
}
// batched Nearest_Search; outputs are n x k, rows padded with id 0xFFFFFFFF / d2 = +inf
void orc_map_knn(void* h, const float* q, int n, int k, float max_d2, uint32_t* ids, float* d2, float* pts, int* cnt) {
  Ctx* c = (Ctx*)h;
#pragma omp parallel for num_threads(c->nthreads) schedule(static)
  for (int i = 0; i < n; ++i) {
    std::vector<KnnHit> hits(k);
    int m = c->map.knn(q + 3 * i, k, max_d2, hits.data());
    cnt[i] = m;
    for (int j = 0; j < k; ++j) {
      size_t o = size_t(i) * k + j;
      if (j < m) {
        ids[o] = hits[j].id; d2[o] = hits[j].d2;
        if (pts) { pts[3 * o] = hits[j].x; pts[3 * o + 1] = hits[j].y; pts[3 * o + 2] = hits[j].z; }
      } else {
        ids[o] = 0xFFFFFFFFu; d2[o] = INFINITY;
        if (pts) { pts[3 * o] = pts[3 * o + 1] = pts[3 * o + 2] = 0.f; }
      }
    }
  }
}

int orc_esti_plane(const float* pts5x3, float threshold, float* pabcd) {
  float P[kK][3];
  for (int i = 0; i < kK; ++i) for (int a = 0; a < 3; ++a) P[i][a] = pts5x3[3 * i + a];
  return esti_plane(pabcd, P, threshold) ? 1 : 0;
}

void orc_scan_set(void* h, const float* body_xyz, int n) {
  Scan& s = ((Ctx*)h)->scan;
  s.n = n;
  s.body.assign(body_xyz, body_xyz + size_t(n) * 3);
  s.world.assign(size_t(n) * 3, 0.f);
  s.nn.assign(size_t(n) * kK * 4, 0.f);
  s.nn_cnt.assign(n, 0);
  s.sel.assign(n, 0);
  s.plane.assign(size_t(n) * 4, 0.f);
  s.far.assign(size_t(n) * 4, 0.f);
  for (int i = 0; i < n; ++i) {  // no search has run for this scan: no far neighbour either
    const uint32_t none = 0xFFFFFFFFu;
    std::memcpy(&s.far[size_t(i) * 4 + 3], &none, 4);
  }
}

int orc_h_share(void* h, const double* state36, int converge, double* HTH, double* HTz, double* res_sum,
                double* hx_out, double* h_out) {
  Ctx* c = (Ctx*)h;
  St s = unpack(state36);
  return h_share_model(*c, s, converge != 0, HTH, HTz, res_sum, hx_out, h_out);
}

// read back the per-point products of the last h_share call
void orc_scan_get(void* h, float* world, uint32_t* nn_ids, float* nn_xyz, int* nn_cnt, uint8_t* sel, float* plane) {
  Scan& s = ((Ctx*)h)->scan;
  if (world) std::memcpy(world, s.world.data(), s.world.size() * 4);
  if (nn_cnt) std::memcpy(nn_cnt, s.nn_cnt.data(), s.nn_cnt.size() * 4);
  if (sel) std::memcpy(sel, s.sel.data(), s.sel.size());
  if (plane) std::memcpy(plane, s.plane.data(), s.plane.size() * 4);
  for (int i = 0; i < s.n; ++i)
    for (int k = 0; k < kK; ++k) {
      const float* d = &s.nn[(size_t(i) * kK + k) * 4];
      size_t o = size_t(i) * kK + k;
      bool ok = k < s.nn_cnt[i];
      if (nn_ids) { uint32_t id = 0xFFFFFFFFu; if (ok) std::memcpy(&id, &d[3], 4); nn_ids[o] = id; }
      if (nn_xyz) for (int a = 0; a < 3; ++a) nn_xyz[3 * o + a] = ok ? d[a] : 0.f;
    }
}

int orc_iekf_update(void* h, double* state36, double* P529, double R, int max_iter, const double* limit23,
                    double* iter_states, int* n_eff_trace, int* knn_trace) {
  Ctx* c = (Ctx*)h;
  St x = unpack(state36);
  M23 P;
  std::memcpy(P.a, P529, sizeof(P.a));
  int it = iekf_update(*c, x, P, R, max_iter, limit23, iter_states, n_eff_trace, knn_trace);
  pack(x, state36);
  std::memcpy(P529, P.a, sizeof(P.a));
  return it;
}

void orc_map_incremental(void* h, const double* state36, int flg_ekf_inited, int* n_add, int* n_nodown) {
  Ctx* c = (Ctx*)h;
  map_incremental(*c, unpack(state36), flg_ekf_inited, n_add, n_nodown);
}

int orc_fov_segment(void* h, const double* pos_lid, float* boxes_out) { return fov_segment(*(Ctx*)h, pos_lid, boxes_out); }

// manifold helpers exposed so the host-side C++ of the product can be checked against them
void orc_boxplus(double* state36, const double* dx23) {
  St x = unpack(state36);
  V23 d; std::memcpy(d.a, dx23, sizeof(d.a));
  boxplus(x, d);
  pack(x, state36);
}
void orc_boxminus(const double* a36, const double* b36, double* dx23) {
  V23 d = boxminus(unpack(a36), unpack(b36));
  std::memcpy(dx23, d.a, sizeof(d.a));
}

}  // extern "C"

// Host side of esekfom::update_iterated_dyn_share_modified with the GPU measurement model behind it.
// Reference: third_party/FAST_LIO/include/IKFoM_toolkit/esekfom/esekfom.hpp (absent submodule;
// behaviour per SURVEY.md §9.6 and the IKFoM paper).  The 23x23 algebra stays on the host in FP64:
// per iteration it consumes only H^T H (12x12) and H^T z (12) from the device.
#include <cmath>
#include <cstring>
#include "esekf_state.hpp"
#include "lio_ctx.h"

using namespace esekf;

namespace lio {

int iekf_update_host(lio_ctx* c, double* state, double* P_io, double R, int max_iter, const double* limit, double* iter_states,
                     int* n_eff_trace, int* knn_trace, int* iters_done) {
  State x = load(state);
  const State x_prop = x;
  double P_prop[kDof * kDof], P[kDof * kDof];
  std::memcpy(P_prop, P_io, sizeof(P_prop));
  std::memcpy(P, P_io, sizeof(P));
  bool converge = true;
  int t = 0, done = 0;
  double cur[LIO_STATE_DOUBLES];
  for (int i = -1; i < max_iter; ++i) {
    store(x, cur);
    double HTH[144], HTz[12];
    int n_eff = 0;
    if (knn_trace) knn_trace[done] = converge ? 1 : 0;
    int rc = lio_h_share(c, cur, converge ? 1 : 0, HTH, HTz, &n_eff, nullptr);
    if (rc) return rc;
    if (n_eff_trace) n_eff_trace[done] = n_eff;
    if (n_eff < 1) {  // ekfom_data.valid == false: the round is skipped
      if (iter_states) std::memcpy(iter_states + LIO_STATE_DOUBLES * done, cur, sizeof(cur));
      ++done;
      continue;
    }
    double dx[kDof], dx_new[kDof];
    state_boxminus(x, x_prop, dx);
    std::memcpy(dx_new, dx, sizeof(dx));
    std::memcpy(P, P_prop, sizeof(P));
    for (int idx : {3, 6}) {
      const Mat3 J = transpose(a_matrix(Vec3{{dx[idx], dx[idx + 1], dx[idx + 2]}}));
      const Vec3 seg = mul(J, Vec3{{dx_new[idx], dx_new[idx + 1], dx_new[idx + 2]}});
      for (int k = 0; k < 3; ++k) dx_new[idx + k] = seg[k];
      left_block(P, kDof, idx, 3, J.m);
      right_block(P, kDof, kDof, idx, 3, J.m);
    }
    {
      double J[4];
      s2_jacobian(x.grav, x_prop.grav, dx + 21, J);
      const double a = dx_new[21], b = dx_new[22];
      dx_new[21] = J[0] * a + J[1] * b;
      dx_new[22] = J[2] * a + J[3] * b;
      left_block(P, kDof, 21, 2, J);
      right_block(P, kDof, kDof, 21, 2, J);
    }
    // information form: (P/R)^-1 + [HTH 0; 0 0], then K_h = Pinv[:, :12] HTz, K_x[:, :12] = Pinv[:, :12] HTH
    double Ps[kDof * kDof], Pt[kDof * kDof], Pi[kDof * kDof];
    for (int k = 0; k < kDof * kDof; ++k) Ps[k] = P[k] * (1.0 / R);
    if (!invert<kDof>(Ps, Pt)) {
      set_error("lio_iekf_update: covariance is singular");
      return LIO_E_ARG;
    }
    for (int a = 0; a < 12; ++a)
      for (int b = 0; b < 12; ++b) Pt[a * kDof + b] += HTH[a * 12 + b];
    if (!invert<kDof>(Pt, Pi)) {
      set_error("lio_iekf_update: information matrix is singular");
      return LIO_E_ARG;
    }
    double K_h[kDof], K_x[kDof * kDof];
    std::memset(K_x, 0, sizeof(K_x));
    for (int r = 0; r < kDof; ++r) {
      double s = 0;
      for (int k = 0; k < 12; ++k) s += Pi[r * kDof + k] * HTz[k];
      K_h[r] = s;
      for (int cc = 0; cc < 12; ++cc) {
        double q = 0;
        for (int k = 0; k < 12; ++k) q += Pi[r * kDof + k] * HTH[k * 12 + cc];
        K_x[r * kDof + cc] = q;
      }
    }
    double dxu[kDof];
    for (int r = 0; r < kDof; ++r) {
      double s = K_h[r];
      for (int k = 0; k < kDof; ++k) s += (K_x[r * kDof + k] - (r == k ? 1.0 : 0.0)) * dx_new[k];
      dxu[r] = s;
    }
    const State x_before = x;
    state_boxplus(x, dxu);
    if (iter_states) store(x, iter_states + LIO_STATE_DOUBLES * done);
    ++done;
    converge = true;
    for (int k = 0; k < kDof; ++k)
      if (std::fabs(dxu[k]) > limit[k]) {
        converge = false;
        break;
      }
    if (converge) ++t;
    if (!t && i == max_iter - 2) converge = true;
    if (t > 1 || i == max_iter - 1) {
      // covariance update: P = L - K_x[:, :12] * P[:12, :] with the manifold re-projections
      double L[kDof * kDof];
      std::memcpy(L, P, sizeof(L));
      for (int idx : {3, 6}) {
        const Mat3 J = transpose(a_matrix(Vec3{{dxu[idx], dxu[idx + 1], dxu[idx + 2]}}));
        left_block(L, kDof, idx, 3, J.m);
        left_block(K_x, kDof, idx, 3, J.m);
        right_block(L, kDof, kDof, idx, 3, J.m);
        right_block(P, kDof, kDof, idx, 3, J.m);
      }
      {
        double J[4];
        s2_jacobian(x.grav, x_before.grav, dxu + 21, J);
        left_block(L, kDof, 21, 2, J);
        left_block(K_x, kDof, 21, 2, J);
        right_block(L, kDof, kDof, 21, 2, J);
        right_block(P, kDof, kDof, 21, 2, J);
      }
      for (int r = 0; r < kDof; ++r)
        for (int cc = 0; cc < kDof; ++cc) {
          double s = 0;
          for (int k = 0; k < 12; ++k) s += K_x[r * kDof + k] * P[k * kDof + cc];
          P_io[r * kDof + cc] = L[r * kDof + cc] - s;
        }
      store(x, state);
      if (iters_done) *iters_done = done;
      return LIO_OK;
    }
  }
  // every round was invalid (no effective points): state and covariance stay as propagated
  store(x, state);
  if (iters_done) *iters_done = done;
  return LIO_OK;
}

}  // namespace lio

extern "C" {

void lio_boxplus(double* state, const double* dx23) {
  State x = load(state);
  state_boxplus(x, dx23);
  store(x, state);
}

void lio_boxminus(const double* a, const double* b, double* dx23) { state_boxminus(load(a), load(b), dx23); }

}  // extern "C"

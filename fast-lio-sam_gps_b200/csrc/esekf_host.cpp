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
        // [RECALL, unverified against the absent esekfom.hpp] upstream evaluates Mx at x_propagated here as well as at the top
        // of the loop (its `x_before` copy is never read); the oracle and the device loop make the same choice
        s2_jacobian(x.grav, x_prop.grav, dxu + 21, J);
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

// esekfom::predict(dt, Q, in) with FAST-LIO's process model (use-ikfom.hpp get_f / df_dx / df_dw; absent submodule,
// SURVEY.md §8(f)-1).  The Jacobians are sparse: F_x1 = I + dt*f_x differs from the identity only in the blocks written
// below, and G = dt*f_w has four 3x3 blocks, so they are filled in place instead of going through the flat 24-row form.
int lio_imu_predict(double* state, double* P, double dt, const double* acc, const double* gyro, const double* Q) {
  if (!state || !P || !acc || !gyro || !Q) return LIO_E_ARG;
  State x = load(state);
  const Vec3 omega{{gyro[0] - x.bg[0], gyro[1] - x.bg[1], gyro[2] - x.bg[2]}};
  const Vec3 a_b{{acc[0] - x.ba[0], acc[1] - x.ba[1], acc[2] - x.ba[2]}};
  const Vec3 Ra = mul(x.rot, a_b);
  const Vec3 a_w{{Ra[0] + x.grav[0], Ra[1] + x.grav[1], Ra[2] + x.grav[2]}};
  const Mat3 R0 = x.rot;
  double F[kDof * kDof], G[kDof * 12];
  std::memset(F, 0, sizeof(F));
  std::memset(G, 0, sizeof(G));
  for (int i = 0; i < kDof; ++i) F[i * kDof + i] = 1.0;
  // rotation rows: Exp(-omega dt) on the diagonal, A(-omega dt) in front of d(omega)/d(bg) = -I and of the gyro noise
  const Vec3 seg{{omega[0] * -dt, omega[1] * -dt, omega[2] * -dt}};
  const Mat3 E = so3_exp(seg), A = a_matrix(seg);
  const Mat3 RhatA = mul(R0, skew(a_b));
  // gravity chart: Mx(g, 0) = -[g]x B(g)
  const Mat32 B = s2_basis(x.grav);
  const Mat3 Sg = skew(x.grav);
  double Mx[6];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 2; ++c) {
      double s = 0;
      for (int k = 0; k < 3; ++k) s += Sg(r, k) * B(k, c);
      Mx[2 * r + c] = -s;
    }
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) {
      F[(3 + r) * kDof + 3 + c] = E(r, c);
      F[(3 + r) * kDof + 15 + c] = (A(r, c) * -1.0) * dt;
      F[(12 + r) * kDof + 3 + c] = (RhatA(r, c) * -1.0) * dt;
      F[(12 + r) * kDof + 18 + c] = (R0(r, c) * -1.0) * dt;
      G[(3 + r) * 12 + c] = (A(r, c) * -1.0) * dt;
      G[(12 + r) * 12 + 3 + c] = (R0(r, c) * -1.0) * dt;
    }
  for (int r = 0; r < 3; ++r) {
    F[r * kDof + 12 + r] = dt;  // d(pos)/d(vel)
    G[(15 + r) * 12 + 6 + r] = dt;
    G[(18 + r) * 12 + 9 + r] = dt;
    for (int c = 0; c < 2; ++c) F[(12 + r) * kDof + 21 + c] = Mx[2 * r + c] * dt;
  }
  {
    const double zero[2] = {0.0, 0.0};
    double J[4];
    s2_jacobian(x.grav, x.grav, zero, J);  // Nx(g) * Mx(g, 0)
    F[21 * kDof + 21] = J[0], F[21 * kDof + 22] = J[1], F[22 * kDof + 21] = J[2], F[22 * kDof + 22] = J[3];
  }
  // x <- x [+] f dt
  for (int i = 0; i < 3; ++i) x.pos[i] += x.vel[i] * dt;
  x.rot = mul(x.rot, so3_exp(Vec3{{omega[0] * dt, omega[1] * dt, omega[2] * dt}}));
  for (int i = 0; i < 3; ++i) x.vel[i] += a_w[i] * dt;
  // P <- F P F^T + G Q G^T
  double T[kDof * kDof], T2[kDof * 12], Pn[kDof * kDof];
  for (int r = 0; r < kDof; ++r)
    for (int c = 0; c < kDof; ++c) {
      double s = 0;
      for (int k = 0; k < kDof; ++k) s += F[r * kDof + k] * P[k * kDof + c];
      T[r * kDof + c] = s;
    }
  for (int r = 0; r < kDof; ++r)
    for (int c = 0; c < 12; ++c) {
      double s = 0;
      for (int k = 0; k < 12; ++k) s += G[r * 12 + k] * Q[k * 12 + c];
      T2[r * 12 + c] = s;
    }
  for (int r = 0; r < kDof; ++r)
    for (int c = 0; c < kDof; ++c) {
      double s = 0, q = 0;
      for (int k = 0; k < kDof; ++k) s += T[r * kDof + k] * F[c * kDof + k];
      for (int k = 0; k < 12; ++k) q += T2[r * 12 + k] * G[c * 12 + k];
      Pn[r * kDof + c] = s + q;
    }
  std::memcpy(P, Pn, sizeof(Pn));
  store(x, state);
  return LIO_OK;
}

}  // extern "C"

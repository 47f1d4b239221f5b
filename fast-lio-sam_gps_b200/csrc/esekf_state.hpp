// state_ikfom as the 36-double block that crosses the C-ABI (LIO_STATE_DOUBLES), its boxplus / boxminus on the 23-DoF
// error state, and the block-Jacobian helpers of esekfom's iterated update.  Shared by esekf_host.cpp (host loop that
// drives lio_h_share) and lio_iekf.cuh (the device-resident loop).  Reference: IKFoM_toolkit/esekfom/esekfom.hpp and
// use-ikfom.hpp of the absent FAST_LIO submodule.
#pragma once
#include "esekf_math.hpp"

namespace esekf {

// LIO_STATE_DOUBLES layout
struct State {
  Vec3 pos, offT, vel, bg, ba, grav;
  Mat3 rot, offR;
};

ESEKF_HD inline State load(const double* s) {
  State x;
  for (int i = 0; i < 3; ++i) {
    x.pos[i] = s[i];
    x.offT[i] = s[21 + i];
    x.vel[i] = s[24 + i];
    x.bg[i] = s[27 + i];
    x.ba[i] = s[30 + i];
    x.grav[i] = s[33 + i];
  }
  for (int i = 0; i < 9; ++i) {
    x.rot.m[i] = s[3 + i];
    x.offR.m[i] = s[12 + i];
  }
  return x;
}

ESEKF_HD inline void store(const State& x, double* s) {
  for (int i = 0; i < 3; ++i) {
    s[i] = x.pos[i];
    s[21 + i] = x.offT[i];
    s[24 + i] = x.vel[i];
    s[27 + i] = x.bg[i];
    s[30 + i] = x.ba[i];
    s[33 + i] = x.grav[i];
  }
  for (int i = 0; i < 9; ++i) {
    s[3 + i] = x.rot.m[i];
    s[12 + i] = x.offR.m[i];
  }
}

// error-state order: pos 0, rot 3, offR 6, offT 9, vel 12, bg 15, ba 18, grav 21 (2 dof)
ESEKF_HD inline void state_boxminus(const State& x, const State& o, double d[kDof]) {
  const Vec3 r = so3_log(mul(transpose(o.rot), x.rot)), r2 = so3_log(mul(transpose(o.offR), x.offR));
  for (int i = 0; i < 3; ++i) {
    d[i] = x.pos[i] - o.pos[i];
    d[3 + i] = r[i];
    d[6 + i] = r2[i];
    d[9 + i] = x.offT[i] - o.offT[i];
    d[12 + i] = x.vel[i] - o.vel[i];
    d[15 + i] = x.bg[i] - o.bg[i];
    d[18 + i] = x.ba[i] - o.ba[i];
  }
  s2_boxminus(x.grav, o.grav, d + 21);
}

ESEKF_HD inline void state_boxplus(State& x, const double d[kDof]) {
  for (int i = 0; i < 3; ++i) {
    x.pos[i] += d[i];
    x.offT[i] += d[9 + i];
    x.vel[i] += d[12 + i];
    x.bg[i] += d[15 + i];
    x.ba[i] += d[18 + i];
  }
  x.rot = mul(x.rot, so3_exp(Vec3{{d[3], d[4], d[5]}}));
  x.offR = mul(x.offR, so3_exp(Vec3{{d[6], d[7], d[8]}}));
  x.grav = s2_boxplus(x.grav, d + 21);
}

// M[idx..idx+D) rows <- J * rows
ESEKF_HD inline void left_block(double* M, int ncols, int idx, int D, const double* J) {
  for (int c = 0; c < ncols; ++c) {
    double t[3];
    for (int r = 0; r < D; ++r) {
      t[r] = 0;
      for (int k = 0; k < D; ++k) t[r] += J[r * D + k] * M[(idx + k) * ncols + c];
    }
    for (int r = 0; r < D; ++r) M[(idx + r) * ncols + c] = t[r];
  }
}
// M[:, idx..idx+D) <- cols * J^T
ESEKF_HD inline void right_block(double* M, int nrows, int ncols, int idx, int D, const double* J) {
  for (int r = 0; r < nrows; ++r) {
    double t[3];
    for (int c = 0; c < D; ++c) {
      t[c] = 0;
      for (int k = 0; k < D; ++k) t[c] += M[r * ncols + idx + k] * J[c * D + k];
    }
    for (int c = 0; c < D; ++c) M[r * ncols + idx + c] = t[c];
  }
}


}  // namespace esekf

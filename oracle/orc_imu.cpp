// ORACLE — TEST INFRASTRUCTURE ONLY.  Nothing in the product path may include, link or call this file.
// PARITY UNPINNED: third_party/FAST_LIO (Kodifly/FAST_LIO@8aa7f2d, /root/reference/.SUBMODULES.json:9-15) is absent from
// the mount, so this restates, from the published FAST-LIO2 / IKFoM description [RECALL, SURVEY.md §8(f)-1], the step
// that feeds the hot path:
//   * esekfom::predict with FAST-LIO's process model (use-ikfom.hpp get_f / df_dx / df_dw)   -> orc_imu_predict
//   * ImuProcess::UndistortPcl's backward per-point motion compensation (IMU_Processing.hpp)  -> orc_undistort
// Sequential and literal on purpose (two-pointer backward sweep over the time-sorted cloud, as the reference does it).
#include <cmath>
#include <cstring>
#include "orc_linalg.hpp"

using namespace orc;

namespace {

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

template <int R, int C, int RR, int CC>
void set_block(Mat<R, C>& M, int r0, int c0, const Mat<RR, CC>& B) {
  for (int r = 0; r < RR; ++r)
    for (int c = 0; c < CC; ++c) M(r0 + r, c0 + c) = B(r, c);
}

// FAST-LIO so3_math.h Exp(ang_vel, dt): identity below 1e-7 rad/s
M3 exp_rate(const V3& w, double dt) {
  const double n = norm(w);
  M3 I = M3::Identity();
  if (n > 1e-7) {
    V3 ax = vec3(w.a[0] / n, w.a[1] / n, w.a[2] / n);
    M3 K = hat(ax);
    const double ang = n * dt;
    return I + K * std::sin(ang) + (K * K) * (1.0 - std::cos(ang));
  }
  return I;
}

}  // namespace

extern "C" {

// esekfom::predict(dt, Q, in) for state_ikfom; Q is 12x12 row-major over (ng, na, nbg, nba)
void orc_imu_predict(double* state36, double* P529, double dt, const double* acc3, const double* gyr3, const double* Q144) {
  St x = unpack(state36);
  const St xb = x;
  const V3 acc = vec3(acc3[0], acc3[1], acc3[2]), gyr = vec3(gyr3[0], gyr3[1], gyr3[2]);
  const V3 omega = gyr - x.bg, a_b = acc - x.ba;
  const V3 a_w = x.rot * a_b + x.grav;
  // flat 24-vector f and its Jacobians (get_f, df_dx, df_dw)
  Mat<24, 23> fx;
  Mat<24, 12> fw;
  set_block(fx, 0, 12, M3::Identity());
  set_block(fx, 12, 3, (x.rot * hat(a_b)) * -1.0);
  set_block(fx, 12, 18, x.rot * -1.0);
  Mat<2, 1> zero2;
  set_block(fx, 12, 21, S2_Mx(x.grav, zero2));
  set_block(fx, 3, 15, M3::Identity() * -1.0);
  set_block(fw, 12, 3, x.rot * -1.0);
  set_block(fw, 3, 0, M3::Identity() * -1.0);
  set_block(fw, 15, 6, M3::Identity());
  set_block(fw, 18, 9, M3::Identity());
  // x_.oplus(f, dt)
  x.pos = x.pos + x.vel * dt;
  x.rot = x.rot * so3_exp(omega * dt);
  x.vel = x.vel + a_w * dt;
  // F_x1, f_x_final, f_w_final
  Mat<23, 23> F = Mat<23, 23>::Identity(), fxf;
  Mat<23, 12> fwf;
  const int vect_idx[5] = {0, 9, 12, 15, 18};  // pos, offT, vel, bg, ba: error index == flat index
  for (int v = 0; v < 5; ++v)
    for (int r = 0; r < 3; ++r) {
      for (int c = 0; c < 23; ++c) fxf(vect_idx[v] + r, c) = fx(vect_idx[v] + r, c);
      for (int c = 0; c < 12; ++c) fwf(vect_idx[v] + r, c) = fw(vect_idx[v] + r, c);
    }
  const V3 f_rot[2] = {omega, vec3(0, 0, 0)};
  for (int b = 0; b < 2; ++b) {
    const int idx = 3 + 3 * b;
    const V3 seg = f_rot[b] * (-dt);
    set_block(F, idx, idx, so3_exp(seg));
    const M3 J = A_matrix(seg);
    Mat<3, 23> rx;
    Mat<3, 12> rw;
    for (int r = 0; r < 3; ++r) {
      for (int c = 0; c < 23; ++c) rx(r, c) = fx(idx + r, c);
      for (int c = 0; c < 12; ++c) rw(r, c) = fw(idx + r, c);
    }
    set_block(fxf, idx, 0, J * rx);
    set_block(fwf, idx, 0, J * rw);
  }
  {  // S2 gravity: f = 0, so the rotation is I and the f_x / f_w rows vanish
    Mat<2, 2> blk = S2_Nx_yy(x.grav) * S2_Mx(xb.grav, zero2);
    set_block(F, 21, 21, blk);
  }
  F = F + fxf * dt;
  Mat<23, 23> P;
  std::memcpy(P.a, P529, sizeof(P.a));
  Mat<12, 12> Q;
  std::memcpy(Q.a, Q144, sizeof(Q.a));
  Mat<23, 12> G = fwf * dt;
  P = F * P * F.T() + G * Q * G.T();
  std::memcpy(P529, P.a, sizeof(P.a));
  pack(x, state36);
}

// UndistortPcl's backward sweep.  xyz/t_ms: n points sorted by t_ms ascending (curvature, milliseconds from scan start);
// pose: n_pose x 22 doubles (offset_time s | acc3 | gyr3 | vel3 | pos3 | rot9 row-major), offset_time ascending, pose[0] at 0;
// end_state36: state at the scan end.  Points are rewritten in place, expressed in the LiDAR frame at the scan end.
void orc_undistort(float* xyz, const float* t_ms, int n, const double* pose, int n_pose, const double* end_state36) {
  if (n <= 0 || n_pose < 2) return;
  const St e = unpack(end_state36);
  int ip = n - 1;
  for (int kp = n_pose - 1; kp > 0; --kp) {
    const double* head = pose + 22 * (kp - 1);
    const double* tail = pose + 22 * kp;
    M3 R_imu;
    std::memcpy(R_imu.a, head + 13, sizeof(R_imu.a));
    const V3 vel = vec3(head[7], head[8], head[9]), pos = vec3(head[10], head[11], head[12]);
    const V3 acc = vec3(tail[1], tail[2], tail[3]), gyr = vec3(tail[4], tail[5], tail[6]);
    for (; t_ms[ip] / double(1000) > head[0]; --ip) {
      const double dt = t_ms[ip] / double(1000) - head[0];
      const M3 R_i = R_imu * exp_rate(gyr, dt);
      const V3 P_i = vec3(xyz[3 * ip], xyz[3 * ip + 1], xyz[3 * ip + 2]);
      const V3 T_ei = pos + vel * dt + ((acc * 0.5) * dt) * dt - e.pos;
      const V3 P_c = e.offR.T() * (e.rot.T() * (R_i * (e.offR * P_i + e.offT) + T_ei) - e.offT);
      xyz[3 * ip] = (float)P_c.a[0];
      xyz[3 * ip + 1] = (float)P_c.a[1];
      xyz[3 * ip + 2] = (float)P_c.a[2];
      if (ip == 0) return;
    }
  }
}

}  // extern "C"

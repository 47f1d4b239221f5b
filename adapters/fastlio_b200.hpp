// Reference-side binding for seam B1: FAST-LIO's measurement model, map and map update on the B200,
// behind the unchanged `h_dyn_share` callback signature and esekfom state layout.  A maintainer includes this header in
// third_party/FAST_LIO/src/laserMapping.cpp (absent submodule: the names below are FAST-LIO's public ones),
// links liblio_b200.so and replaces the bodies of
//     h_share_model(state_ikfom&, esekfom::dyn_share_datastruct<double>&)   -> g_b200.h_share_model(s, ekfom_data)
//     ikdtree.Build(feats_down_world->points)                               -> g_b200.build(*feats_down_world)
//     map_incremental()                                                     -> g_b200.map_incremental(state_point)
//     lasermap_fov_segment()                                                -> g_b200.lasermap_fov_segment(pos_LiD)
// and calls g_b200.set_scan(*feats_down_body) once per scan before kf.update_iterated_dyn_share_modified(...).
// `kf.init_dyn_share(get_f, df_dx, df_dw, h_share_model, NUM_MAX_ITERATIONS, epsi)` stays as it is.
//
// Templates instead of concrete IKFoM / PCL types so the header compiles here (adapters/check/) without those libraries:
//   State    : state_ikfom        — .pos, .offset_T_L_I (vect3: operator()(i)), .rot, .offset_R_L_I (SO3: toRotationMatrix()(r,c)),
//                                    .vel, .bg, .ba (vect3), .grav (S2: .vec(i) or operator[](i))
//   DynShare : esekfom::dyn_share_datastruct<double> — bool valid, converge; Matrix<double,Dynamic,Dynamic> h_x; Matrix<double,Dynamic,1> h
//   Cloud    : pcl::PointCloud<PointType> — .points[i].x/.y/.z, .size()
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>
#include "../include/lio_b200.h"

class FastLioB200 {
 public:
  // filter_size_map / cube_side_length / det_range / extrinsic_est_en: the launch parameters of kitti.launch:10-11, kitti.yaml:22
  FastLioB200(float filter_size_map, double cube_side_length, float det_range, bool extrinsic_est_en, uint32_t max_map_points = 1u << 24,
              uint32_t max_scan_points = 1u << 18, int device = 0) {
    lio_config c{};
    c.device = device;
    c.filter_size_map = filter_size_map;
    c.cell_mult = 2;
    c.max_points = max_map_points;
    c.max_scan = max_scan_points;
    c.max_sqdist = 5.0f;  // the `pointSearchSqDis[NUM_MATCH_POINTS - 1] > 5` gate of h_share_model
    c.extrinsic_est_en = extrinsic_est_en ? 1 : 0;
    c.cube_side_length = cube_side_length;
    c.det_range = det_range;
    ncol_ = 12;
    check(lio_ctx_create(&c, &ctx_), "lio_ctx_create");  // throws without a B200: there is no CPU fallback
  }
  ~FastLioB200() { lio_ctx_destroy(ctx_); }
  FastLioB200(const FastLioB200 &) = delete;
  FastLioB200 &operator=(const FastLioB200 &) = delete;

  template <class State>
  static void pack_state(const State &s, double out[LIO_STATE_DOUBLES]) {
    const auto R = s.rot.toRotationMatrix();
    const auto RL = s.offset_R_L_I.toRotationMatrix();
    for (int i = 0; i < 3; ++i) {
      out[i] = s.pos(i);
      out[21 + i] = s.offset_T_L_I(i);
      out[24 + i] = s.vel(i);
      out[27 + i] = s.bg(i);
      out[30 + i] = s.ba(i);
      out[33 + i] = s.grav[i];
      for (int j = 0; j < 3; ++j) {
        out[3 + 3 * i + j] = R(i, j);
        out[12 + 3 * i + j] = RL(i, j);
      }
    }
  }

  // ikdtree.Build(points)
  template <class Cloud>
  void build(const Cloud &world) {
    flatten(world);
    check(lio_map_build(ctx_, xyz_.data(), (int)world.size()), "lio_map_build");
  }
  // feats_down_body of the current scan
  template <class Cloud>
  void set_scan(const Cloud &body) {
    flatten(body);
    n_ = (int)body.size();
    check(lio_scan_upload(ctx_, xyz_.data(), n_), "lio_scan_upload");
  }

  // The h_dyn_share callback.  Fills ekfom_data.h_x (effct_feat_num x 12) and ekfom_data.h (effct_feat_num) exactly like the
  // reference's stage 2/3 (rows of the selected points in scan order); ekfom_data.valid = false when no point is selected.
  template <class State, class DynShare>
  void h_share_model(State &s, DynShare &ekfom_data) {
    double st[LIO_STATE_DOUBLES];
    pack_state(s, st);
    sel_.resize(n_);
    rows_.resize((size_t)n_ * 12);
    z_.resize(n_);
    int n_eff = 0;
    check(lio_h_share_rows(ctx_, st, ekfom_data.converge ? 1 : 0, sel_.data(), rows_.data(), z_.data(), &n_eff), "lio_h_share_rows");
    if (n_eff < 1) {
      ekfom_data.valid = false;  // "No Effective Points!"
      return;
    }
    ekfom_data.h_x.resize(n_eff, 12);
    ekfom_data.h.resize(n_eff);
    int r = 0;
    for (int i = 0; i < n_; ++i) {
      if (!sel_[i]) continue;
      for (int j = 0; j < 12; ++j) ekfom_data.h_x(r, j) = rows_[(size_t)i * 12 + j];
      ekfom_data.h(r) = z_[i];
      ++r;
    }
  }

  // Faster seam for a patched esekfom (12x12 block step): the normal equations instead of the N x 12 Jacobian.
  template <class State>
  bool h_share_normal(const State &s, bool converge, double HTH[144], double HTz[12], int *n_eff) {
    double st[LIO_STATE_DOUBLES];
    pack_state(s, st);
    check(lio_h_share(ctx_, st, converge ? 1 : 0, HTH, HTz, n_eff, nullptr), "lio_h_share");
    return *n_eff >= 1;
  }

  // The per-scan body of the main loop in one call: kf.update_iterated_dyn_share_modified(LASER_POINT_COV, ...) followed by
  // map_incremental() at the updated state (lio_iekf_update_map).  P is the 23x23 covariance (row-major), limit = epsi.
  // state_out receives the updated state in the LIO_STATE_DOUBLES layout (rotations as 3x3 row-major matrices): the caller
  // writes it back with kf.change_x(...) the way it already does after set_extrinsic / init.
  template <class State>
  void update_and_map(const State &s, double *P23x23, double laser_point_cov, int max_iter, const double limit[23],
                      double state_out[LIO_STATE_DOUBLES], bool flg_ekf_inited = true) {
    pack_state(s, state_out);
    check(lio_iekf_update_map(ctx_, state_out, P23x23, laser_point_cov, max_iter, limit, flg_ekf_inited ? 1 : 0, nullptr, nullptr, nullptr,
                              nullptr),
          "lio_iekf_update_map");
  }
  // ImuProcess::UndistortPcl's backward loop on the scan uploaded with set_scan(): t_ms[i] = points[i].curvature,
  // pose6d = IMUpose flattened to LIO_POSE6D_DOUBLES per entry, end_state = imu_state after the last predict.
  template <class State>
  void undistort(const float *t_ms, const double *pose6d, int n_pose, const State &end_state) {
    double st[LIO_STATE_DOUBLES];
    pack_state(end_state, st);
    check(lio_scan_undistort(ctx_, t_ms, pose6d, n_pose, st), "lio_scan_undistort");
  }

  // map_incremental() at the updated state (flg_EKF_inited as in the reference)
  template <class State>
  void map_incremental(const State &s, bool flg_ekf_inited = true) {
    double st[LIO_STATE_DOUBLES];
    pack_state(s, st);
    check(lio_map_incremental(ctx_, st, flg_ekf_inited ? 1 : 0, nullptr, nullptr), "lio_map_incremental");
  }
  // lasermap_fov_segment(): pos_LiD = state.pos + state.rot * state.offset_T_L_I
  void lasermap_fov_segment(const double pos_lid[3]) { check(lio_fov_segment(ctx_, pos_lid, nullptr, nullptr), "lio_fov_segment"); }
  int map_size() {
    int n = 0;
    check(lio_map_size(ctx_, &n), "lio_map_size");
    return n;
  }
  lio_ctx *handle() { return ctx_; }

 private:
  template <class Cloud>
  void flatten(const Cloud &c) {
    xyz_.resize(c.size() * 3);
    for (size_t i = 0; i < c.size(); ++i) {
      xyz_[3 * i] = c.points[i].x, xyz_[3 * i + 1] = c.points[i].y, xyz_[3 * i + 2] = c.points[i].z;
    }
  }
  void check(int rc, const char *what) const {
    if (rc != LIO_OK) throw std::runtime_error(std::string(what) + ": " + lio_last_error());
  }
  lio_ctx *ctx_ = nullptr;
  int n_ = 0, ncol_ = 12;
  std::vector<float> xyz_;
  std::vector<uint8_t> sel_;
  std::vector<double> rows_, z_;
};

// Minimal stand-ins for the PCL / Eigen / IKFoM types the adapters are written against, ONLY so that `make -C adapters`
// can type-check adapters/*.hpp in an image without those libraries.  Not used by the product or the tests' compute.
#pragma once
#include <cstddef>
#include <limits>
#include <tuple>
#include <vector>

namespace pcl {
struct PointXYZI {
  float x, y, z, pad0, intensity, pad1, pad2, pad3;  // 32-byte record like PCL's
};
template <class P>
struct PointCloud {
  std::vector<P> points;
  unsigned width = 0, height = 1;
  bool is_dense = true;
  size_t size() const { return points.size(); }
};
}  // namespace pcl
using PointType = pcl::PointXYZI;  // utilities.hpp:29

namespace Eigen {
struct Matrix4d {
  double m[16];
  static Matrix4d Identity() {
    Matrix4d r{};
    for (int i = 0; i < 4; ++i) r.m[5 * i] = 1.0;
    return r;
  }
  double &operator()(int r, int c) { return m[4 * r + c]; }
  double operator()(int r, int c) const { return m[4 * r + c]; }
};
}  // namespace Eigen

struct PosePcd {  // pose_pcd.hpp:7-19
  pcl::PointCloud<pcl::PointXYZI> pcd_;
  Eigen::Matrix4d pose_eig_ = Eigen::Matrix4d::Identity();
  Eigen::Matrix4d pose_corrected_eig_ = Eigen::Matrix4d::Identity();
  double timestamp_ = 0;
  int idx_ = 0;
  bool processed_ = false;
};
using PcdPair = std::tuple<pcl::PointCloud<PointType>, pcl::PointCloud<PointType>>;
struct LoopClosureConfig {  // loop_closure.h:21-29
  int num_submap_keyframes_ = 10;
  double voxel_res_ = 0.1;
  double loop_detection_radius_ = 0, loop_detection_timediff_threshold_ = 0, icp_score_threshold_ = 0, icp_max_corr_dist_ = 0;
};
struct RegistrationOutput {  // loop_closure.h:31-37
  bool is_valid_ = false;
  bool is_converged_ = false;
  double score_ = std::numeric_limits<double>::max();
  Eigen::Matrix4d pose_between_eig_ = Eigen::Matrix4d::Identity();
};

// IKFoM look-alikes for fastlio_b200.hpp
namespace mock {
struct Vec3 {
  double v[3] = {0, 0, 0};
  double operator()(int i) const { return v[i]; }
  double operator[](int i) const { return v[i]; }
};
struct Mat3 {
  double m[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  double operator()(int r, int c) const { return m[3 * r + c]; }
};
struct SO3 {
  Mat3 R;
  Mat3 toRotationMatrix() const { return R; }
};
struct state_ikfom {
  Vec3 pos;
  SO3 rot, offset_R_L_I;
  Vec3 offset_T_L_I, vel, bg, ba, grav;
};
struct MatX {
  std::vector<double> d;
  int r = 0, c = 0;
  void resize(int rr, int cc) { r = rr, c = cc, d.assign((size_t)rr * cc, 0.0); }
  double &operator()(int i, int j) { return d[(size_t)j * r + i]; }
};
struct VecX {
  std::vector<double> d;
  void resize(int n) { d.assign(n, 0.0); }
  double &operator()(int i) { return d[i]; }
};
struct dyn_share_datastruct {
  bool valid = true, converge = true;
  MatX h_x;
  VecX h;
};
}  // namespace mock

// Reference-side binding for seam B2: a drop-in LoopClosure whose public methods keep the signatures of
// fast_lio_sam/include/loop_closure.h:39-70 and forward to the C-ABI of include/loop_b200.h.  A maintainer adds this header
// next to loop_closure.h, links liblio_b200.so, and replaces `LoopClosure` by `LoopClosureB200` in fast_lio_sam.h; the
// factor code at fast_lio_sam.cpp:699-717 is untouched (RegistrationOutput is the reference's own struct).
//
// Requires the reference's own headers first (PointType, PosePcd, LoopClosureConfig, RegistrationOutput, PcdPair);
// adapters/check/ compiles it against minimal stand-ins because PCL/Eigen are not installed in this image.
#pragma once
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>
#include "../include/loop_b200.h"

class LoopClosureB200 {
 private:
  loop_ctx *ctx_ = nullptr;
  LoopClosureConfig config_;
  int closest_keyframe_idx_ = -1;
  size_t uploaded_ = 0;  // keyframes already resident on the device
  std::vector<double> sent_pose_;  // 16 doubles per resident keyframe: the pose_corrected_eig_ the device holds
  pcl::PointCloud<PointType> src_cloud_, dst_cloud_, aligned_;

  static void pack(const pcl::PointCloud<PointType> &c, std::vector<float> &out) {  // PointXYZI (32 B) -> 4 floats
    out.resize(c.size() * 4);
    for (size_t i = 0; i < c.size(); ++i) {
      out[4 * i] = c.points[i].x, out[4 * i + 1] = c.points[i].y, out[4 * i + 2] = c.points[i].z, out[4 * i + 3] = c.points[i].intensity;
    }
  }
  static void unpack(const std::vector<float> &in, int n, pcl::PointCloud<PointType> &c) {
    c.points.resize(n);
    for (int i = 0; i < n; ++i) {
      c.points[i].x = in[4 * i], c.points[i].y = in[4 * i + 1], c.points[i].z = in[4 * i + 2], c.points[i].intensity = in[4 * i + 3];
    }
    c.width = n, c.height = 1, c.is_dense = true;
  }
  static void pose16(const Eigen::Matrix4d &m, double *o) {
    for (int r = 0; r < 4; ++r)
      for (int c = 0; c < 4; ++c) o[4 * r + c] = m(r, c);  // row-major across the ABI
  }
  void check(int rc, const char *what) const {
    if (rc != LOOP_OK) throw std::runtime_error(std::string(what) + ": " + loop_last_error());
  }
  // keyframes_ only grows and only pose_corrected_eig_ changes (fast_lio_sam.cpp pose-graph update): mirror both.  Poses are
  // compared with what the device already holds and only the changed ones are sent, in ONE call (a tick without a pose-graph
  // update sends nothing; round 1 re-sent every pose on every call, O(K) C calls per loop check).
  void sync(const std::vector<PosePcd> &keyframes) {
    std::vector<float> buf;
    double T[16];
    if (keyframes.size() < uploaded_) {
      check(loop_keyframe_clear(ctx_), "loop_keyframe_clear");
      uploaded_ = 0;
      sent_pose_.clear();
    }
    std::vector<int> changed;
    std::vector<double> poses;
    for (size_t i = 0; i < uploaded_; ++i) {
      pose16(keyframes[i].pose_corrected_eig_, T);
      if (std::memcmp(T, &sent_pose_[16 * i], sizeof(T)) != 0) {
        changed.push_back((int)i);
        poses.insert(poses.end(), T, T + 16);
        std::memcpy(&sent_pose_[16 * i], T, sizeof(T));
      }
    }
    if (!changed.empty()) check(loop_keyframe_set_poses(ctx_, changed.data(), poses.data(), (int)changed.size()), "loop_keyframe_set_poses");
    for (; uploaded_ < keyframes.size(); ++uploaded_) {
      const PosePcd &k = keyframes[uploaded_];
      pack(k.pcd_, buf);
      pose16(k.pose_corrected_eig_, T);
      check(loop_keyframe_add(ctx_, buf.data(), (int)k.pcd_.size(), T, k.timestamp_, nullptr), "loop_keyframe_add");
      sent_pose_.insert(sent_pose_.end(), T, T + 16);
    }
  }
  RegistrationOutput to_output(const loop_result &r) const {  // loop_closure.cpp:83-90
    RegistrationOutput out;
    out.score_ = r.score;  // always written (:83)
    if (r.valid) {         // hasConverged() && score < icp_score_threshold_: the reference sets BOTH flags only here (:85-90),
      out.is_valid_ = true;  // is_converged_ stays at its default false otherwise (loop_closure.h:33-34)
      out.is_converged_ = true;
      for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) out.pose_between_eig_(i, j) = (double)r.T[4 * i + j];
    }
    return out;
  }

 public:
  explicit LoopClosureB200(const LoopClosureConfig &config, int device = 0) : config_(config) {
    loop_config c{};
    c.device = device;
    c.max_corr_dist = config.icp_max_corr_dist_;  // loop_closure.cpp:7
    c.transformation_epsilon = 0.01;              // :8
    c.euclidean_fitness_epsilon = 0.01;           // :9
    c.max_iterations = 50;                        // :10
    check(loop_ctx_create(&c, &ctx_), "loop_ctx_create");  // throws without a B200: there is no CPU fallback
  }
  ~LoopClosureB200() { loop_ctx_destroy(ctx_); }
  LoopClosureB200(const LoopClosureB200 &) = delete;
  LoopClosureB200 &operator=(const LoopClosureB200 &) = delete;

  int fetchClosestKeyframeIdx(const PosePcd &query_keyframe, const std::vector<PosePcd> &keyframes) {
    sync(keyframes);
    int out = -1;
    check(loop_fetch_closest(ctx_, query_keyframe.idx_, config_.loop_detection_radius_, config_.loop_detection_timediff_threshold_, &out),
          "loop_fetch_closest");
    return out;
  }
  PcdPair setSrcAndDstCloud(const std::vector<PosePcd> &keyframes, const int src_idx, const int dst_idx, const int submap_range,
                            const double voxel_res) {
    sync(keyframes);
    // a voxelised submap never has more points than the keyframes it is made of: one call with buffers sized from those
    size_t cap = 0;
    for (const int centre : {src_idx, dst_idx}) {
      size_t n = 0;
      for (int i = centre - submap_range; i < centre + submap_range + 1; ++i)
        if (i >= 0 && i < static_cast<int>(keyframes.size()) - 1) n += keyframes[i].pcd_.size();
      cap = n > cap ? n : cap;
    }
    int ns = 0, nd = 0;
    std::vector<float> s(cap * 4 + 4), d(cap * 4 + 4);
    check(loop_set_src_dst(ctx_, src_idx, dst_idx, submap_range, (float)voxel_res, s.data(), &ns, d.data(), &nd, (int)cap), "loop_set_src_dst");
    pcl::PointCloud<PointType> sc, dc;
    unpack(s, ns, sc);
    unpack(d, nd, dc);
    return {sc, dc};
  }
  RegistrationOutput icpAlignment(const pcl::PointCloud<PointType> &src, const pcl::PointCloud<PointType> &dst) {
    std::vector<float> s, d, al(src.size() * 4 + 4);
    pack(src, s);
    pack(dst, d);
    loop_result r{};
    check(loop_icp_align(ctx_, s.data(), (int)src.size(), d.data(), (int)dst.size(), config_.icp_score_threshold_, &r, al.data(), nullptr),
          "loop_icp_align");
    src_cloud_ = src, dst_cloud_ = dst;
    unpack(al, (int)src.size(), aligned_);
    return to_output(r);
  }
  RegistrationOutput performLoopClosure(const PosePcd &query_keyframe, const std::vector<PosePcd> &keyframes) {
    closest_keyframe_idx_ = fetchClosestKeyframeIdx(query_keyframe, keyframes);
    return performLoopClosure(query_keyframe, keyframes, closest_keyframe_idx_);
  }
  RegistrationOutput performLoopClosure(const PosePcd &query_keyframe, const std::vector<PosePcd> &keyframes, const int closest_keyframe_idx) {
    RegistrationOutput none;
    closest_keyframe_idx_ = closest_keyframe_idx;
    if (closest_keyframe_idx < 0) return none;  // loop_closure.cpp:106-124
    sync(keyframes);
    loop_pair p{query_keyframe.idx_, closest_keyframe_idx};
    loop_result r{};
    check(loop_verify_batch(ctx_, &p, 1, config_.num_submap_keyframes_, (float)config_.voxel_res_, config_.icp_score_threshold_, &r),
          "loop_verify_batch");
    return to_output(r);
  }
  // NEW (no reference counterpart): fetchClosestKeyframeIdx for several query keyframes in one device launch
  std::vector<int> fetchClosestKeyframeIdxBatch(const std::vector<int> &query_indices, const std::vector<PosePcd> &keyframes) {
    sync(keyframes);
    std::vector<int> out(query_indices.size(), -1);
    check(loop_fetch_closest_batch(ctx_, query_indices.data(), (int)query_indices.size(), config_.loop_detection_radius_,
                                   config_.loop_detection_timediff_threshold_, out.data()),
          "loop_fetch_closest_batch");
    return out;
  }
  // NEW: ingest of a keyframe straight from the WORLD-frame cloud FAST-LIO publishes (what PosePcd's constructor receives,
  // pose_pcd.hpp:25-43): the pose^-1 re-expression runs on the device, the host never transforms or copies the cloud.  Must be
  // called in keyframe order, for the keyframe that is about to be appended to keyframes_ (index == keyframes_.size()).
  int addKeyframeFromWorldCloud(const pcl::PointCloud<PointType> &cloud_world, const Eigen::Matrix4d &pose, double stamp) {
    std::vector<float> buf;
    double T[16];
    pack(cloud_world, buf);
    pose16(pose, T);
    int idx = -1;
    check(loop_keyframe_add_world(ctx_, buf.data(), (int)cloud_world.size(), T, stamp, &idx), "loop_keyframe_add_world");
    uploaded_ = (size_t)idx + 1;
    sent_pose_.insert(sent_pose_.end(), T, T + 16);
    return idx;
  }
  // NEW (no reference counterpart): all candidates of a timer tick in one device batch
  std::vector<RegistrationOutput> performLoopClosureBatch(const std::vector<PosePcd> &keyframes, const std::vector<loop_pair> &pairs) {
    sync(keyframes);
    std::vector<loop_result> r(pairs.size());
    check(loop_verify_batch(ctx_, pairs.data(), (int)pairs.size(), config_.num_submap_keyframes_, (float)config_.voxel_res_,
                            config_.icp_score_threshold_, r.data()),
          "loop_verify_batch");
    std::vector<RegistrationOutput> out;
    for (const loop_result &x : r) out.push_back(to_output(x));
    return out;
  }
  pcl::PointCloud<PointType> getCloud(int which) {  // of pair 0 of the last batch / alignment
    int n = 0;
    check(loop_get_cloud(ctx_, 0, which, nullptr, 0, &n), "loop_get_cloud");
    std::vector<float> b((size_t)n * 4 + 4);
    check(loop_get_cloud(ctx_, 0, which, b.data(), n, &n), "loop_get_cloud");
    pcl::PointCloud<PointType> c;
    unpack(b, n, c);
    return c;
  }
  pcl::PointCloud<PointType> getSourceCloud() { return getCloud(0); }
  pcl::PointCloud<PointType> getTargetCloud() { return getCloud(1); }
  pcl::PointCloud<PointType> getFinalAlignedCloud() { return getCloud(2); }
  int getClosestKeyframeidx() { return closest_keyframe_idx_; }
};

// Type-checks the reference-side adapters against stand-in types (see ref_types_mock.hpp) and, when a device is present,
// runs them once.  Built by `make -C adapters`; linked against the in-tree liblio_b200.so.
#include <cstdio>
#include "ref_types_mock.hpp"
#include "../fastlio_b200.hpp"
#include "../loop_closure_b200.hpp"

int main() {
  if (lio_device_count() <= 0) {
    std::printf("adapters: compiled and linked; no CUDA device, nothing run\n");
    return 0;
  }
  try {
    FastLioB200 lio(0.5f, 1000.0, 300.0f, false, 1u << 16, 1u << 12);
    pcl::PointCloud<PointType> world, body;
    for (int i = 0; i < 40; ++i)
      for (int j = 0; j < 40; ++j) {
        PointType p{};
        p.x = 0.25f * i, p.y = 0.25f * j, p.z = 1.f;  // (a plane through the origin has no n.p = -1 form)
        world.points.push_back(p);
        if ((i + j) % 3 == 0) {
          p.z = 1.01f;
          body.points.push_back(p);
        }
      }
    lio.build(world);
    lio.set_scan(body);
    mock::state_ikfom s;
    mock::dyn_share_datastruct d;
    lio.h_share_model(s, d);
    std::printf("adapters: h_share_model valid=%d rows=%d, map=%d\n", (int)d.valid, d.h_x.r, lio.map_size());
    lio.map_incremental(s);

    LoopClosureConfig cfg;
    cfg.num_submap_keyframes_ = 1, cfg.voxel_res_ = 0.3, cfg.loop_detection_radius_ = 35, cfg.loop_detection_timediff_threshold_ = 0.5;
    cfg.icp_score_threshold_ = 1.5, cfg.icp_max_corr_dist_ = 52.5;
    LoopClosureB200 lc(cfg);
    std::vector<PosePcd> kfs(4);
    for (int k = 0; k < 4; ++k) {
      kfs[k].pcd_ = world, kfs[k].idx_ = k, kfs[k].timestamp_ = k;
      kfs[k].pose_corrected_eig_(0, 3) = 0.05 * k;
    }
    RegistrationOutput r = lc.performLoopClosure(kfs[2], kfs, 0);
    std::printf("adapters: loop closure converged=%d valid=%d score=%g\n", (int)r.is_converged_, (int)r.is_valid_, r.score_);
    if (!r.is_valid_ || !r.is_converged_) throw std::runtime_error("identical clouds 10 cm apart must verify");
    // loop_closure.cpp:85-90: both flags are set only inside the gate; a pair that converges but scores above the threshold
    // comes back with is_valid_ == is_converged_ == false and score_ still written (loop_closure.h:33-35 defaults otherwise)
    cfg.icp_score_threshold_ = 0.0;
    LoopClosureB200 strict(cfg);
    RegistrationOutput r2 = strict.performLoopClosure(kfs[2], kfs, 0);
    std::printf("adapters: gated out: converged=%d valid=%d score=%g\n", (int)r2.is_converged_, (int)r2.is_valid_, r2.score_);
    if (r2.is_valid_ || r2.is_converged_ || !(r2.score_ < 1e300)) throw std::runtime_error("RegistrationOutput flags differ from loop_closure.cpp:85-90");
    // pose sync sends only what changed; batched candidate generation; ingest from the world-frame cloud
    kfs[1].pose_corrected_eig_(1, 3) = 0.02;
    std::vector<int> cl = lc.fetchClosestKeyframeIdxBatch({3, 2, 0}, kfs);
    if (cl.size() != 3 || cl[0] != lc.fetchClosestKeyframeIdx(kfs[3], kfs) || cl[2] != -1) throw std::runtime_error("fetchClosestKeyframeIdxBatch differs");
    LoopClosureB200 ingest(cfg);
    Eigen::Matrix4d pose = Eigen::Matrix4d::Identity();
    pose(0, 3) = 1.5, pose(1, 3) = -0.5;
    pcl::PointCloud<PointType> shifted = world;
    for (auto &p : shifted.points) p.x += 1.5f, p.y -= 0.5f;  // the same cloud as published in the world frame
    if (ingest.addKeyframeFromWorldCloud(shifted, pose, 0.0) != 0) throw std::runtime_error("addKeyframeFromWorldCloud index");
    std::printf("adapters: batch candidates %d %d %d, world ingest ok\n", cl[0], cl[1], cl[2]);
  } catch (const std::exception &e) {
    std::printf("adapters: FAILED: %s\n", e.what());
    return 1;
  }
  return 0;
}

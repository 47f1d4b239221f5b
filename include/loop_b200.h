/*
 * loop_b200.h — C-ABI of the B200-native loop-closure verification path (seam B2).
 *
 * Replaces, behind the unchanged LoopClosure / RegistrationOutput / BetweenFactor interface:
 *   fast_lio_sam/src/loop_closure.cpp:18-40    LoopClosure::fetchClosestKeyframeIdx   -> loop_fetch_closest
 *   fast_lio_sam/src/loop_closure.cpp:42-67    LoopClosure::setSrcAndDstCloud         -> loop_set_src_dst
 *   fast_lio_sam/src/loop_closure.cpp:69-92    LoopClosure::icpAlignment (+ :7-11)    -> loop_icp_align
 *   fast_lio_sam/src/loop_closure.cpp:101-126  LoopClosure::performLoopClosure        -> loop_verify_batch (n pairs at once)
 *   fast_lio_sam/include/utilities.hpp:132-142 transformPcd                           -> loop_transform_pcd
 *   fast_lio_sam/include/utilities.hpp:158-171 voxelizePcd                            -> loop_voxelize
 *   fast_lio_sam/include/pose_pcd.hpp:7-19     PosePcd (pcd_ in the LiDAR frame, pose_corrected_eig_, timestamp_)
 *                                                                                     -> loop_keyframe_add / _set_pose
 *   fast_lio_sam/include/pose_pcd.hpp:25-43    PosePcd constructor (world-frame cloud -> LiDAR frame by pose^-1)
 *                                                                                     -> loop_keyframe_add_world
 *   fast_lio_sam/include/loop_closure.h:31-37  RegistrationOutput                     -> loop_result
 *   fast_lio_sam/include/loop_closure.h:21-29  LoopClosureConfig                      -> loop_config
 * The PCL 1.10 arithmetic behind those calls (VoxelGrid, IterativeClosestPoint, TransformationEstimationSVD,
 * DefaultConvergenceCriteria, getFitnessScore) is not in the reference mount; it is restated per SURVEY.md §9.4-9.5.
 *
 * Plain C, pointers and sizes only; every entry point returns LOOP_OK or a negative code (codes shared with lio_b200.h),
 * nothing throws.  All point clouds are HOST arrays of pcl::PointXYZI payloads: 4 floats (x, y, z, intensity) per point.
 * 4x4 matrices are row-major.  One host thread at a time per handle; each handle owns a CUDA stream.
 * There is NO CPU fallback: creation fails with LOOP_E_CUDA when no sm_100 device is usable.
 */
#ifndef LOOP_B200_H
#define LOOP_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LOOP_OK 0
#define LOOP_E_CUDA (-1)
#define LOOP_E_ARG (-2)
#define LOOP_E_CAPACITY (-3)
#define LOOP_E_STATE (-4)

typedef struct loop_ctx loop_ctx; /* opaque: device keyframe store + batch workspace + stream */

typedef struct loop_config {
  int device;                     /* CUDA ordinal */
  double max_corr_dist;           /* setMaxCorrespondenceDistance: loop_detection_radius * 1.5 (fast_lio_sam.cpp:73) */
  double transformation_epsilon;  /* loop_closure.cpp:8  (0.01) */
  double euclidean_fitness_epsilon; /* loop_closure.cpp:9  (0.01) */
  int max_iterations;             /* loop_closure.cpp:10 (50) */
  double rotation_epsilon;        /* DefaultConvergenceCriteria rotation threshold; 0 -> 1 - transformation_epsilon */
  double mse_threshold_absolute;  /* DefaultConvergenceCriteria absolute MSE threshold; 0 -> 1e-12 */
  int max_similar_iterations;     /* max_iterations_similar_transforms_ (0) */
  int min_correspondences;        /* Registration::min_number_correspondences_; 0 -> 3 */
  void* stream;                   /* optional external cudaStream_t; NULL -> the context creates its own */
} loop_config;

/* RegistrationOutput (loop_closure.h:31-37) plus the counters the tests and the bench need */
typedef struct loop_result {
  float T[16];      /* getFinalTransformation(), Matrix4f; pose_between_eig_ is its cast to double when is_valid */
  double score;     /* getFitnessScore(); always written (loop_closure.cpp:83) */
  int32_t converged;  /* hasConverged() */
  int32_t valid;      /* converged && score < icp_score_threshold (loop_closure.cpp:85) */
  int32_t iterations; /* ICP iterations executed */
  int32_t state;      /* 0 none, 1 iterations, 2 transform, 3 abs mse, 4 rel mse, 5 too few correspondences */
  int32_t n_src;      /* points of the voxelised source submap */
  int32_t n_dst;      /* points of the voxelised target submap */
} loop_result;

typedef struct loop_pair {
  int32_t src_idx; /* query keyframe (latest) */
  int32_t dst_idx; /* closest keyframe returned by fetchClosestKeyframeIdx */
} loop_pair;

int loop_ctx_create(const loop_config* cfg, loop_ctx** out);
void loop_ctx_destroy(loop_ctx* ctx);
const char* loop_last_error(void);

/* ---- keyframe store: std::vector<PosePcd> keyframes_ ---------------------------------------------------------------- */
/* Appends a keyframe: pcd_ (n points, LiDAR frame), pose_corrected_eig_ (16 doubles), timestamp_.  Returns its index
 * in *idx_out (optional).  The cloud is copied to the device once and stays resident. */
int loop_keyframe_add(loop_ctx* ctx, const float* xyzi, int n, const double* pose16, double stamp, int* idx_out);
/* PosePcd's constructor (pose_pcd.hpp:25-43): the cloud is the WORLD-frame cloud FAST-LIO publishes; the store keeps
 * pcd_ = transformPcd(cloud, pose.inverse()) (pose_pcd.hpp:38-39).  The re-expression runs on the device on the uploaded
 * copy.  pose16 = pose_eig_ (= pose_corrected_eig_ at construction).  Fails with LOOP_E_ARG on a singular pose. */
int loop_keyframe_add_world(loop_ctx* ctx, const float* xyzi_world, int n, const double* pose16, double stamp, int* idx_out);
/* pose_corrected_eig_ = ... (fast_lio_sam.cpp pose-graph corrections) */
int loop_keyframe_set_pose(loop_ctx* ctx, int idx, const double* pose16);
/* the same for n keyframes in one call (idx[n], pose16s[16 n]): what a caller sends after a pose-graph update - only the
 * poses that changed, not all of them */
int loop_keyframe_set_poses(loop_ctx* ctx, const int* idx, const double* pose16s, int n);
int loop_keyframe_count(loop_ctx* ctx);
int loop_keyframe_clear(loop_ctx* ctx);

/* fetchClosestKeyframeIdx: nearest keyframe among [0, size-2] within `radius` and older than `time_diff`; -1 if none.
 * The query's pose and stamp are those stored under query_idx, and the answer is the POSITION of the keyframe in the store;
 * the reference returns keyframes[idx].idx_ and reads the pose of the PosePcd it is handed (loop_closure.cpp:18-40) - the
 * same thing as long as idx_ equals the position and the query is the stored keyframe, which is how fast_lio_sam.cpp:684-698
 * uses it. */
int loop_fetch_closest(loop_ctx* ctx, int query_idx, double radius, double time_diff, int* closest_out);
/* the same for n query keyframes in one device launch (closest_out[n]); translations and stamps are kept on the device and
 * re-sent only after keyframes were added or poses changed */
int loop_fetch_closest_batch(loop_ctx* ctx, const int* query_idx, int n, double radius, double time_diff, int* closest_out);

/* ---- utilities.hpp ---------------------------------------------------------------------------------------------------- */
/* transformPcd: out[i] = T * in[i] (f64 matrix applied to f32 points), intensity copied. */
int loop_transform_pcd(loop_ctx* ctx, const float* xyzi, int n, const double* T16, float* out_xyzi);
/* voxelizePcd: pcl::VoxelGrid with a cubic leaf; out must hold n points; *n_out = points written
 * (ascending voxel index; the unfiltered input when PCL's index-overflow guard triggers). */
int loop_voxelize(loop_ctx* ctx, const float* xyzi, int n, float leaf, float* out_xyzi, int* n_out);

/* ---- LoopClosure ------------------------------------------------------------------------------------------------------ */
/* setSrcAndDstCloud: world-frame submaps of keyframes [idx-range, idx+range] ∩ [0, size-2], voxelised at `leaf`.
 * Copies min(cap, n) points into each optional output; sizes always reported. */
int loop_set_src_dst(loop_ctx* ctx, int src_idx, int dst_idx, int range, float leaf, float* src_out, int* n_src, float* dst_out,
                     int* n_dst, int cap);
/* icpAlignment on explicit clouds.  aligned_out (optional, ns points) = aligned_; corr_out (optional, ns ints) = target index
 * matched in the last correspondence search (-1: none). `score_thr` gates `valid`. */
int loop_icp_align(loop_ctx* ctx, const float* src_xyzi, int ns, const float* dst_xyzi, int nd, double score_thr, loop_result* out,
                   float* aligned_out, int* corr_out);
/* performLoopClosure for n_pairs candidate pairs in one call: submap build + VoxelGrid + ICP + score gate per pair, all
 * pairs advanced together on the device.  out[n_pairs]. */
int loop_verify_batch(loop_ctx* ctx, const loop_pair* pairs, int n_pairs, int range, float leaf, double score_thr, loop_result* out);
/* aligned_ / src_cloud_ / dst_cloud_ debug getters (loop_closure.h:66-68) of pair `pair` of the last batch:
 * which = 0 src, 1 dst, 2 aligned.  Copies min(cap, n) points; *n_out = n. */
int loop_get_cloud(loop_ctx* ctx, int pair, int which, float* out_xyzi, int cap, int* n_out);

/* ---- measurement hooks (bench / profiling only) ----------------------------------------------------------------------- */
int loop_timing_enable(loop_ctx* ctx, int on);
/* names: "submap", "voxel", "tree", "nn", "cov", "solve", "score" */
int loop_timing_get(loop_ctx* ctx, const char* name, double* total_ms, int* launches);
int loop_timing_reset(loop_ctx* ctx);
int64_t loop_launch_count(loop_ctx* ctx);
/* totals of the last batch: ICP iterations summed over pairs, source points summed over pairs, 1-NN queries issued */
int loop_batch_stats(loop_ctx* ctx, int64_t* iterations, int64_t* src_points, int64_t* nn_queries);

#ifdef __cplusplus
}
#endif
#endif /* LOOP_B200_H */

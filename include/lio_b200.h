/*
 * lio_b200.h — C-ABI of the B200-native FAST-LIO measurement path (seam B1).
 *
 * Replaces, behind the unchanged `h_dyn_share` callback / esekfom state layout:
 *   third_party/FAST_LIO/src/laserMapping.cpp        h_share_model, map_incremental, lasermap_fov_segment
 *   third_party/FAST_LIO/include/common_lib.h        esti_plane
 *   third_party/FAST_LIO/include/ikd-Tree/ikd_Tree.h Build, Nearest_Search, Add_Points, Delete_Point_Boxes
 *   third_party/FAST_LIO/include/IKFoM_toolkit/esekfom/esekfom.hpp  update_iterated_dyn_share_modified
 * The FAST_LIO submodule is absent from the reference mount (/root/reference/.SUBMODULES.json:9-15),
 * so no file:line can be cited for these; the verified anchors are the parameter names/values in
 * /root/reference/third_party/fastlio_config_launch/kitti.launch:6-12 and kitti.yaml:16-22.
 *
 * Plain C: pointers and sizes only.  All entry points return LIO_OK (0) or a negative LIO_E_* code;
 * nothing throws.  One host thread at a time per handle; each handle owns a CUDA stream.
 * There is NO CPU fallback: every compute entry point fails with LIO_E_CUDA when no sm_100 device is usable.
 */
#ifndef LIO_B200_H
#define LIO_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LIO_OK 0
#define LIO_E_CUDA (-1)     /* CUDA runtime error (no device, launch failure, ...) */
#define LIO_E_ARG (-2)      /* bad argument */
#define LIO_E_CAPACITY (-3) /* map hash table or bucket pool exhausted */
#define LIO_E_STATE (-4)    /* call order violated (e.g. h_share before scan upload) */

#define LIO_NUM_MATCH 5       /* NUM_MATCH_POINTS */
#define LIO_STATE_DOUBLES 36  /* pos3 | rot 3x3 row-major | offset_R_L_I 3x3 | offset_T_L_I 3 | vel3 | bg3 | ba3 | grav3 */
#define LIO_ERR_DOF 23        /* esekfom error-state dimension of state_ikfom */

typedef struct lio_ctx lio_ctx; /* opaque: GPU map + current scan + stream */

typedef struct lio_config {
  int device;            /* CUDA ordinal */
  float filter_size_map; /* ikd-Tree downsample size (kitti.launch:10: 0.5) */
  int cell_mult;         /* hash cell = cell_mult x filter_size_map; power of two; 0 -> 2 */
  uint32_t max_points;   /* capacity of the point store (bucket pool sized from it) */
  uint32_t max_cells;    /* expected number of occupied hash cells; 0 -> max_points */
  uint32_t max_scan;     /* largest scan (points) that will be uploaded */
  float max_sqdist;      /* neighbour acceptance bound on the 5th squared distance (reference: 5) */
  int extrinsic_est_en;  /* 0 -> Jacobian cols 6..11 are zero (kitti.yaml:22) */
  double cube_side_length; /* kitti.launch:11 */
  float det_range;         /* mapping/det_range */
  void* stream;          /* optional external cudaStream_t; NULL -> the context creates its own */
} lio_config;

const char* lio_version(void);
const char* lio_last_error(void); /* thread-local text of the last failure */
int lio_device_count(void);

int lio_ctx_create(const lio_config* cfg, lio_ctx** out);
void lio_ctx_destroy(lio_ctx* ctx);
int lio_ctx_sync(lio_ctx* ctx); /* wait for the stream; surfaces deferred LIO_E_CAPACITY */

/* ---- map: replaces KD_TREE<PointType> ------------------------------------------------------------ */
/* Build(points): clears the map, then plain-inserts n points (xyz, stride 3 floats, HOST memory).
 * Every point handed to build/add gets id = running counter, in list order. */
int lio_map_build(lio_ctx* ctx, const float* xyz, int n);
/* Add_Points(points, downsample_on).  *net_new (optional) = points of this batch alive afterwards. */
int lio_map_add(lio_ctx* ctx, const float* xyz, int n, int downsample_on, int* net_new);
/* Delete_Point_Boxes: nb boxes of 6 floats (min xyz, max xyz), half-open.  *n_deleted optional. */
int lio_map_delete_boxes(lio_ctx* ctx, const float* boxes, int nb, int* n_deleted);
int lio_map_size(lio_ctx* ctx, int* n_live);
/* Rebuilds the store from its live points (ikd-Tree reclaims deleted nodes when it rebuilds a sub-tree; here cells, buckets
 * and overflow chains emptied by Delete_Point_Boxes are handed back in one pass).  Ids are kept, so every search and
 * lio_map_export returns exactly what it returned before; renumber_ids != 0 replaces them by their rank (same order and
 * tie-breaks) and restarts the insertion counter at the live count.  Runs by itself after a box delete once a pool is three
 * quarters used and a quarter of the live count has been deleted since the last compaction (LIO_MAP_COMPACT=0 disables
 * that), and with renumbering when the 32-bit insertion counter passes 0xC0000000. */
int lio_map_compact(lio_ctx* ctx, int renumber_ids);
/* out8: live points, insertion counter, buckets handed out, bucket capacity, cells in the table, cell limit, sticky error
 * bits (1 hash table full, 2 bucket pool full), points deleted since the last compaction. */
int lio_map_stats(lio_ctx* ctx, uint64_t* out8);
/* flatten(): live points in ascending id; returns total live count in *n_total, copies min(cap, total). */
int lio_map_export(lio_ctx* ctx, float* xyz, uint32_t* ids, int cap, int* n_total);
/* Nearest_Search, batched: n queries (host xyz) -> k<=5 neighbours each within max_sqdist, ascending
 * (sqdist, id).  Rows are padded with id 0xFFFFFFFF / d2 +inf.  pts (n*k*3) and cnt (n) optional. */
int lio_map_knn(lio_ctx* ctx, const float* q_xyz, int n, int k, float max_sqdist, uint32_t* ids, float* d2, float* pts,
                int* cnt);

/* ---- scan + measurement model: replaces h_share_model -------------------------------------------- */
/* feats_down_body: n points (xyz, stride 3 floats, HOST memory) -> device. Invalidates cached neighbours.
 * The copy is ASYNCHRONOUS on the context's stream: a pageable buffer has been staged by the time the call returns, but a
 * PINNED buffer is still being read by the device afterwards - keep it unchanged until the next call that waits for the
 * stream (lio_h_share, lio_iekf_update*, lio_ctx_sync) has returned. */
int lio_scan_upload(lio_ctx* ctx, const float* body_xyz, int n);
/* Optional: start the host->device copy of the NEXT scan on a separate copy stream so that it overlaps the kernels of the
 * scan in flight (the LiDAR driver hands over scan i+1 while scan i is in the filter).  A following lio_scan_upload with
 * the same pointer and count consumes the prefetched copy; any other upload ignores it.  The buffer must stay unchanged
 * until then, and should be pinned. */
int lio_scan_prefetch(lio_ctx* ctx, const float* body_xyz, int n);
/* Same, from a DEVICE pointer to n float4 (x,y,z,unused). */
int lio_scan_set_device(lio_ctx* ctx, const void* body_xyzw_dev, int n);

/* One call of h_share_model at `state` (LIO_STATE_DOUBLES).  redo_knn = ekfom_data.converge.
 * Outputs (HOST): HTH[144] row-major = h_x^T h_x, HTz[12] = h_x^T h, *n_eff = effct_feat_num,
 * *res_sum (optional) = sum |pd2| over selected points.  ekfom_data.valid == (*n_eff >= 1). */
int lio_h_share(lio_ctx* ctx, const double* state, int redo_knn, double* HTH, double* HTz, int* n_eff, double* res_sum);
/* Same evaluation, returning the dense per-point products so an adapter can fill the unchanged
 * dyn_share_datastruct: sel[n], rows[n*12] (h_x row, zero where !sel), z[n] (h = -pd2).  All HOST, all optional. */
int lio_h_share_rows(lio_ctx* ctx, const double* state, int redo_knn, uint8_t* sel, double* rows, double* z, int* n_eff);
/* Per-point products of the last h_share call (HOST, each optional): world[n*3], nn_ids[n*5],
 * nn_xyz[n*5*3], nn_cnt[n], sel[n], plane[n*4] = (a,b,c,pd2). */
int lio_scan_get(lio_ctx* ctx, float* world, uint32_t* nn_ids, float* nn_xyz, int* nn_cnt, uint8_t* sel, float* plane);

/* ---- esekfom::update_iterated_dyn_share_modified on (state, P) with the GPU measurement model ------ */
/* state: in/out LIO_STATE_DOUBLES; P: in/out 23x23 row-major; R = LASER_POINT_COV; limit[23] = epsi.
 * Optional traces (HOST): iter_states[(max_iter+1)*36], n_eff_trace[max_iter+1], knn_trace[max_iter+1].
 * *iters_done = number of measurement-model calls made. */
int lio_iekf_update(lio_ctx* ctx, double* state, double* P, double R, int max_iter, const double* limit,
                    double* iter_states, int* n_eff_trace, int* knn_trace, int* iters_done);
/* The per-scan body of the reference's main loop in one call: update_iterated_dyn_share_modified followed
 * by map_incremental() at the updated state (laserMapping.cpp main loop: `kf.update_iterated_dyn_share_modified(...)`
 * ... `map_incremental()`; absent submodule).  By default the iterations run as a HOST loop: one
 * measurement-model launch per iteration, the 156 doubles of H^T H / H^T z read from mapped pinned memory (a ~10 us poll), the
 * 23x23 update on the host (measured faster than the device form: 0.53 vs 0.79 ms per scan).  LIO_IEKF_MODE=device in the
 * environment at context creation enqueues the whole loop - measurement calls, the 23x23 update of every iteration and the map
 * update - without a host round trip.  Either way the call returns when the updated (state, P) are in host memory; the map
 * insert may still be in flight on the context's stream, and a capacity error it runs into is reported by the NEXT
 * lio_iekf_update* / lio_map_add / lio_ctx_sync call (LIO_E_CAPACITY). */
int lio_iekf_update_map(lio_ctx* ctx, double* state, double* P, double R, int max_iter, const double* limit,
                        int flg_ekf_inited, double* iter_states, int* n_eff_trace, int* knn_trace, int* iters_done);

/* ---- the step in front of the path: IMU forward propagation + per-point undistortion ---------- */
/* esekfom::predict(dt, Q, in) with FAST-LIO's process model (IKFoM_toolkit/esekfom/esekfom.hpp predict();
 * include/use-ikfom.hpp get_f / df_dx / df_dw — absent submodule).  HOST FP64 (23x23, serial, ~20 calls per scan):
 * state in/out LIO_STATE_DOUBLES, P in/out 23x23 row-major, acc/gyro = averaged IMU sample, Q = 12x12 row-major process
 * noise over (gyro, acc, gyro-bias, acc-bias).  No device work. */
int lio_imu_predict(double* state, double* P, double dt, const double* acc, const double* gyro, const double* Q);
#define LIO_POSE6D_DOUBLES 22 /* Pose6D: offset_time [s] | acc3 | gyr3 | vel3 | pos3 | rot 3x3 row-major */
/* ImuProcess::UndistortPcl's backward compensation (src/IMU_Processing.hpp) applied IN PLACE to the uploaded scan:
 * t_ms[n] = per-point `curvature` (milliseconds from the scan start, HOST); imu_pose = n_pose x LIO_POSE6D_DOUBLES (HOST,
 * offset_time ascending, first entry at 0); end_state = state at the scan end.  Invalidates cached neighbours. */
int lio_scan_undistort(lio_ctx* ctx, const float* t_ms, const double* imu_pose, int n_pose, const double* end_state);
/* The scan as it currently sits on the device (after upload / undistortion): body_xyz[n*3] (HOST). */
int lio_scan_get_body(lio_ctx* ctx, float* body_xyz);

/* map_incremental() at the final state, entirely on the device (classification from the cached
 * neighbours, then Add_Points(PointToAdd,true) and Add_Points(PointNoNeedDownsample,false)).
 * points_near[0] of a point whose search found nothing within max_sqdist is its FAR nearest map point (upstream's search is
 * unbounded; looked up here, from the point's position at the last search); a scan that was never searched has empty
 * Nearest_Points and every point goes to PointToAdd.  Counts are optional; asking for them synchronises. */
int lio_map_incremental(lio_ctx* ctx, const double* state, int flg_ekf_inited, int* n_add, int* n_nodown);
/* lasermap_fov_segment(): slides the local-map cube around pos_lid[3]; boxes_out optional (3x6 floats). */
int lio_fov_segment(lio_ctx* ctx, const double* pos_lid, float* boxes_out, int* n_boxes);

/* manifold helpers of the host-side esekfom mirror (exported for tests and adapters) */
void lio_boxplus(double* state, const double* dx23);
void lio_boxminus(const double* a, const double* b, double* dx23);

/* ---- measurement hooks (bench / profiling only) ------------------------------------------------------ */
/* Mean device time (ms) of the named kernel family since the last reset, measured with CUDA events on the
 * context stream when timing is enabled: "knn", "plane_reduce", "insert", "classify". */
int lio_timing_enable(lio_ctx* ctx, int on);
int lio_timing_get(lio_ctx* ctx, const char* name, double* mean_ms, int* launches);
int lio_timing_reset(lio_ctx* ctx);
/* queries of all searches so far that left the shared-memory tile path for the per-query fallback search */
int lio_knn_stats(lio_ctx* ctx, uint64_t* fallback_queries);
/* out4: [0] queries finished by the per-query kernel, [1] queries handed to it because a single query overflowed the
 * staging area, [2] queries served through a pushed sub-tile task (their tile was not coherent), [3] reserved.  Running
 * totals since context creation. */
int lio_knn_stats_detail(lio_ctx* ctx, uint64_t* out4);
/* Per-task counters of the last scan search (context created with LIO_TILE_PROF=1 in the environment).  A task is a tile of
 * 32 curve-ordered queries or a (tile, lane mask) sub-tile pushed to the search's work queue; records are in ticket order,
 * the tiles first.  8 words per task = {SM cycles, staging passes | queries served << 8 | tasks pushed << 16 | pushed task
 * << 24, staged points, cells probed, start ns (low 32 bits of %globaltimer), end ns, SM id, SM cycles spent staging}.
 * Copies min(cap_tiles, tasks) records; *n_tiles = tasks of the last search. */
int lio_tile_profile(lio_ctx* ctx, uint32_t* out8, int cap_tiles, int* n_tiles);
/* Achievable random-gather bandwidth of the device: random 16-byte loads from a table of `table_bytes` (use >= 320 MB so
 * it exceeds L2).  *gbps_16B counts the 16 requested bytes per load, *gbps_sector the 32-byte sectors they drag. */
int lio_gather_probe(lio_ctx* ctx, uint64_t table_bytes, int reps, double* gbps_16B, double* gbps_sector);
/* number of kernels this context has launched since creation */
int64_t lio_launch_count(lio_ctx* ctx);

/* ---- allocation-free entry point -------------------------------------------------------------------- */
/* No counterpart in the reference; added for callers that may not call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
 * The FP64 reduction that ends h_share_model, on its own: HTH[144] (row-major) = rows^T rows and HTz[12] = rows^T z
 * for rows[n*12] / z[n] laid out as lio_h_share_rows returns them (rows of unselected points are zero).  ALL FOUR
 * pointers are DEVICE pointers supplied by the caller, on the current device; `stream` is a cudaStream_t (NULL = the
 * default stream) and the outputs are ready when the stream reaches this point.  The call makes no cudaMalloc /
 * cudaFree and uses no allocating library routine: its per-block partial sums live in a static __device__ array of
 * 768.25 KiB that every GPU loading the library holds until exit.  n > LIO_STATIC_REDUCE_MAX_ROWS (or n < 0, or a
 * null pointer) -> LIO_E_ARG with the limit in lio_last_error(); split the rows and add the results instead.
 * Every call on a device shares that array (it keeps the last call's contents; the call clears what must start at
 * zero), so calls that overlap on one GPU - two streams, two host threads - race: serialise them.
 * Sums are taken in a fixed order (rows in order within 1024-row blocks, blocks in order): bit-reproducible, and equal
 * to lio_h_share's HTH / HTz up to FP64 summation order (tests: 1e-11 relative). */
#define LIO_STATIC_REDUCE_MAX_ROWS (1 << 20)
int lio_hth_reduce_static(const double* rows_dev, const double* z_dev, int n, double* HTH_dev, double* HTz_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* LIO_B200_H */

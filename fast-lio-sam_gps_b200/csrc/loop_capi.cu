// C-ABI of include/loop_b200.h: context, keyframe store, LoopClosure entry points.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include <new>
#include "loop_ctx.h"
#include "../../include/lio_b200.h"

namespace loopb {

void tbegin(loop_ctx* c, int id) {
  if (!c->timing) return;
  Ev ev;
  if (!c->ev_free.empty()) {
    ev = c->ev_free.back();
    c->ev_free.pop_back();
  } else {
    cudaEventCreate(&ev.a);
    cudaEventCreate(&ev.b);
  }
  ev.id = id;
  cudaEventRecord(ev.a, c->stream);
  c->ev_used.push_back(ev);
}

void tend(loop_ctx* c) {
  if (!c->timing || c->ev_used.empty()) return;
  cudaEventRecord(c->ev_used.back().b, c->stream);
}

static void tcollect(loop_ctx* c) {
  for (Ev& ev : c->ev_used) {
    if (cudaEventSynchronize(ev.b) == cudaSuccess) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, ev.a, ev.b) == cudaSuccess) {
        c->t_sum[ev.id] += ms;
        c->t_n[ev.id] += 1;
      }
    }
    c->ev_free.push_back(ev);
  }
  c->ev_used.clear();
}

static int check(loop_ctx* c) {
  if (!c) {
    lio::set_error("null loop context");
    return LOOP_E_ARG;
  }
  if (!lio::cuda_ok(cudaSetDevice(c->cfg.device), "cudaSetDevice")) return LOOP_E_CUDA;
  return LOOP_OK;
}

// keyframes of the submap around `center`: [center - range, center + range] ∩ [0, size - 2]  (loop_closure.cpp:52-65)
static JobHost submap_job(const loop_ctx* c, int center, int range) {
  JobHost j;
  const int K = (int)c->kf_n.size();
  for (int i = center - range; i < center + range + 1; ++i)
    if (i >= 0 && i < K - 1) j.kf.push_back(i);
  return j;
}

static int copy_cloud_out(loop_ctx* c, uint32_t off, uint32_t n, float* out, int cap) {
  const int m = std::min<int>(cap, (int)n);
  if (m > 0 && out) {
    LOOP_CUDA(cudaMemcpyAsync(out, c->vox.p + off, sizeof(float4) * (size_t)m, cudaMemcpyDeviceToHost, c->stream));
    LOOP_CUDA(cudaStreamSynchronize(c->stream));
  }
  return LOOP_OK;
}

// General 4x4 inverse by cofactors in f64 (what `pose_eig_.inverse()` of pose_pcd.hpp:38 asks Eigen for; Eigen's fixed-size
// 4x4 path is a cofactor expansion too, its exact operation order is not in the mount: [RECALL]).  Row-major in, row-major out.
static bool inv4(const double* m, double* o) {
  // 2x2 minors of rows (0,1) and rows (2,3); adjugate entries are sums of three minor products; same operation order as the
  // oracle's restatement so that the re-expressed clouds compare bit for bit
  const double a0 = m[0] * m[5] - m[4] * m[1], a1 = m[0] * m[6] - m[4] * m[2], a2 = m[0] * m[7] - m[4] * m[3];
  const double a3 = m[1] * m[6] - m[5] * m[2], a4 = m[1] * m[7] - m[5] * m[3], a5 = m[2] * m[7] - m[6] * m[3];
  const double b5 = m[10] * m[15] - m[14] * m[11], b4 = m[9] * m[15] - m[13] * m[11], b3 = m[9] * m[14] - m[13] * m[10];
  const double b2 = m[8] * m[15] - m[12] * m[11], b1 = m[8] * m[14] - m[12] * m[10], b0 = m[8] * m[13] - m[12] * m[9];
  const double det = a0 * b5 - a1 * b4 + a2 * b3 + a3 * b2 - a4 * b1 + a5 * b0;
  if (det == 0.0 || !std::isfinite(det)) return false;
  const double r = 1.0 / det;
  const double adj[16] = {
      m[5] * b5 - m[6] * b4 + m[7] * b3,    -m[1] * b5 + m[2] * b4 - m[3] * b3,  m[13] * a5 - m[14] * a4 + m[15] * a3,  -m[9] * a5 + m[10] * a4 - m[11] * a3,
      -m[4] * b5 + m[6] * b2 - m[7] * b1,   m[0] * b5 - m[2] * b2 + m[3] * b1,   -m[12] * a5 + m[14] * a2 - m[15] * a1, m[8] * a5 - m[10] * a2 + m[11] * a1,
      m[4] * b4 - m[5] * b2 + m[7] * b0,    -m[0] * b4 + m[1] * b2 - m[3] * b0,  m[12] * a4 - m[13] * a2 + m[15] * a0,  -m[8] * a4 + m[9] * a2 - m[11] * a0,
      -m[4] * b3 + m[5] * b1 - m[6] * b0,   m[0] * b3 - m[1] * b1 + m[2] * b0,   -m[12] * a3 + m[13] * a1 - m[14] * a0, m[8] * a3 - m[9] * a1 + m[10] * a0};
  for (int i = 0; i < 16; ++i) o[i] = adj[i] * r;
  return true;
}

// fetchClosestKeyframeIdx (loop_closure.cpp:18-40) for many query keyframes at once: one warp per query strides over the
// candidates [0, K-2]; the winner is the smallest (distance, index) among those within `radius` and older than `time_diff`,
// i.e. the first minimum of the reference's ascending scan.  pos = x y z stamp per keyframe.
__global__ void __launch_bounds__(256) k_fetch_closest(const double* __restrict__ pos, int K, const int* __restrict__ query, int n_query,
                                                       double radius, double time_diff, int* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (w >= n_query) return;
  const int q = query[w];
  double best = radius * 3.0;
  int bi = -1;
  if (q >= 0 && q < K) {
    const double qx = pos[4 * q], qy = pos[4 * q + 1], qz = pos[4 * q + 2], qt = pos[4 * q + 3];
    for (int i = lane; i + 1 < K; i += 32) {
      const double dx = pos[4 * i] - qx, dy = pos[4 * i + 1] - qy, dz = pos[4 * i + 2] - qz;
      const double d = sqrt(dx * dx + dy * dy + dz * dz);
      if (radius > d && time_diff < (qt - pos[4 * i + 3]) && d < best) {  // (ascending i per lane: the first minimum is kept)
        best = d;
        bi = i;
      }
    }
  }
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const double ob = __shfl_xor_sync(0xFFFFFFFFu, best, off);
    const int oi = __shfl_xor_sync(0xFFFFFFFFu, bi, off);
    if (oi >= 0 && (bi < 0 || ob < best || (ob == best && oi < bi))) {
      best = ob;
      bi = oi;
    }
  }
  if (lane == 0) out[w] = bi;
}

}  // namespace loopb

using namespace loopb;

extern "C" {

const char* loop_last_error(void) { return lio_last_error(); }

int loop_ctx_create(const loop_config* cfg, loop_ctx** out) {
  if (!cfg || !out) {
    lio::set_error("null argument");
    return LOOP_E_ARG;
  }
  *out = nullptr;
  if (!(cfg->max_corr_dist > 0) || cfg->max_iterations < 1 || !(cfg->transformation_epsilon >= 0) || !(cfg->euclidean_fitness_epsilon >= 0)) {
    lio::set_error("bad loop_config: max_corr_dist > 0, max_iterations >= 1, epsilons >= 0 required");
    return LOOP_E_ARG;
  }
  int ndev = 0;
  if (!lio::cuda_ok(cudaGetDeviceCount(&ndev), "cudaGetDeviceCount") || cfg->device < 0 || cfg->device >= ndev) {
    if (ndev > 0) lio::set_error("device %d out of range (%d devices)", cfg->device, ndev);
    return LOOP_E_CUDA;  // no CPU fallback
  }
  LOOP_CUDA(cudaSetDevice(cfg->device));
  cudaDeviceProp prop;
  LOOP_CUDA(cudaGetDeviceProperties(&prop, cfg->device));
  if (prop.major < 10) {
    lio::set_error("device %d is sm_%d%d; this library carries sm_100a code only", cfg->device, prop.major, prop.minor);
    return LOOP_E_CUDA;
  }
  loop_ctx* c = new (std::nothrow) loop_ctx();
  if (!c) return LOOP_E_ARG;
  c->cfg = *cfg;
  c->n_sms = prop.multiProcessorCount;
  if (const char* e = getenv("LOOP_TREE_ORDER")) c->tree_order = (e[0] == 'm' || e[0] == '0') ? 0 : 1;
  if (const char* e = getenv("LOOP_TREE")) c->tree_kind = (e[0] == 'b' || e[0] == '0') ? 0 : 1;
  if (const char* e = getenv("LOOP_GRID")) c->use_grid = e[0] != '0';
  if (const char* e = getenv("LOOP_SAMPLE_FIRST")) c->sample_first = e[0] != '0';
  c->prm.max_d2 = cfg->max_corr_dist * cfg->max_corr_dist;
  c->prm.trans_eps = cfg->transformation_epsilon;
  c->prm.fit_eps = cfg->euclidean_fitness_epsilon;
  c->prm.rot_thr = cfg->rotation_epsilon > 0 ? cfg->rotation_epsilon : 1.0 - cfg->transformation_epsilon;
  c->prm.mse_abs = cfg->mse_threshold_absolute > 0 ? cfg->mse_threshold_absolute : 1e-12;
  c->prm.max_iter = cfg->max_iterations;
  c->prm.max_similar = cfg->max_similar_iterations;
  c->prm.min_corr = cfg->min_correspondences > 0 ? cfg->min_correspondences : 3;
  if (cfg->stream) {
    c->stream = (cudaStream_t)cfg->stream;
  } else {
    if (!lio::cuda_ok(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking), "cudaStreamCreate")) {
      delete c;
      return LOOP_E_CUDA;
    }
    c->own_stream = true;
  }
  if (!lio::cuda_ok(cudaHostAlloc(&c->active_host, sizeof(int), cudaHostAllocDefault), "cudaHostAlloc") ||
      !lio::cuda_ok(cudaMalloc(&c->active_dev, sizeof(int)), "cudaMalloc")) {
    loop_ctx_destroy(c);
    return LOOP_E_CUDA;
  }
  *out = c;
  return LOOP_OK;
}

void loop_ctx_destroy(loop_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->cfg.device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  tcollect(c);
  for (Ev& ev : c->ev_free) {
    cudaEventDestroy(ev.a);
    cudaEventDestroy(ev.b);
  }
  c->kf_pts.release(); c->segs.release(); c->jobs.release(); c->grids.release(); c->raw.release();
  c->keys_a.release(); c->keys_b.release(); c->vals_a.release(); c->vals_b.release(); c->head_scan.release();
  c->blk2job.release(); c->cub_tmp.release(); c->vox.release(); c->trees.release(); c->tree_pts.release();
  c->nodes.release(); c->pairs.release(); c->cur.release(); c->match.release(); c->md2.release(); c->blk2pair.release();
  if (c->active_host) cudaFreeHost(c->active_host);
  c->kf_pos_dev.release(); c->gtab.release(); c->gnext.release();
  c->leaf_code.release(); c->node_flag.release();
  if (c->active_dev) cudaFree(c->active_dev);
  if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

// ---- keyframe store ------------------------------------------------------------------------------------------------
int loop_keyframe_add(loop_ctx* c, const float* xyzi, int n, const double* pose16, double stamp, int* idx_out) {
  int rc = check(c);
  if (rc) return rc;
  if (n < 0 || (n > 0 && !xyzi) || !pose16) {
    lio::set_error("loop_keyframe_add: bad cloud or pose");
    return LOOP_E_ARG;
  }
  if (c->kf_used + (size_t)n >= 0xFFFFFFFFull) {
    lio::set_error("keyframe store full (2^32 points)");
    return LOOP_E_CAPACITY;
  }
  if ((rc = c->kf_pts.ensure(std::max<size_t>(c->kf_used + (size_t)n, 1), c->stream, true))) return rc;
  if (n) {
    LOOP_CUDA(cudaMemcpyAsync(c->kf_pts.p + c->kf_used, xyzi, sizeof(float4) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    LOOP_CUDA(cudaStreamSynchronize(c->stream));  // the caller's buffer may go away
  }
  c->kf_off.push_back((uint32_t)c->kf_used);
  c->kf_n.push_back((uint32_t)n);
  c->kf_pose.insert(c->kf_pose.end(), pose16, pose16 + 16);
  c->kf_stamp.push_back(stamp);
  c->kf_pos_dirty = true;
  c->kf_used += (size_t)n;
  if (idx_out) *idx_out = (int)c->kf_n.size() - 1;
  return LOOP_OK;
}

int loop_keyframe_set_pose(loop_ctx* c, int idx, const double* pose16) {
  if (!c || !pose16 || idx < 0 || idx >= (int)c->kf_n.size()) {
    lio::set_error("loop_keyframe_set_pose: bad index %d", idx);
    return LOOP_E_ARG;
  }
  std::memcpy(&c->kf_pose[16 * (size_t)idx], pose16, sizeof(double) * 16);
  c->kf_pos_dirty = true;
  return LOOP_OK;
}

// PosePcd's constructor (pose_pcd.hpp:25-43): the cloud arrives in the WORLD frame (FAST-LIO publishes /cloud_registered
// there) and is kept in the LiDAR frame, pcd_ = transformPcd(tmp, pose_eig_.inverse()).  The re-expression runs on the
// device, on the copy that has just landed in the store: no host pass over the points, no second H2D.
int loop_keyframe_add_world(loop_ctx* c, const float* xyzi_world, int n, const double* pose16, double stamp, int* idx_out) {
  int rc = check(c);
  if (rc) return rc;
  if (!pose16) {
    lio::set_error("loop_keyframe_add_world: null pose");
    return LOOP_E_ARG;
  }
  double inv[16];
  if (!inv4(pose16, inv)) {
    lio::set_error("loop_keyframe_add_world: singular pose");
    return LOOP_E_ARG;
  }
  const size_t off = c->kf_used;
  int idx = -1;
  if ((rc = loop_keyframe_add(c, xyzi_world, n, pose16, stamp, &idx))) return rc;
  if ((rc = transform_store_inplace(c, off, n, inv))) return rc;
  if (idx_out) *idx_out = idx;
  return LOOP_OK;
}

int loop_keyframe_set_poses(loop_ctx* c, const int* idx, const double* pose16s, int n) {
  if (!c || n < 0 || (n > 0 && (!idx || !pose16s))) {
    lio::set_error("loop_keyframe_set_poses: bad arguments");
    return LOOP_E_ARG;
  }
  for (int k = 0; k < n; ++k)
    if (idx[k] < 0 || idx[k] >= (int)c->kf_n.size()) {
      lio::set_error("loop_keyframe_set_poses: bad index %d", idx[k]);
      return LOOP_E_ARG;
    }
  for (int k = 0; k < n; ++k) std::memcpy(&c->kf_pose[16 * (size_t)idx[k]], pose16s + 16 * (size_t)k, sizeof(double) * 16);
  if (n) c->kf_pos_dirty = true;
  return LOOP_OK;
}

int loop_keyframe_count(loop_ctx* c) { return c ? (int)c->kf_n.size() : 0; }

int loop_keyframe_clear(loop_ctx* c) {
  if (!c) return LOOP_E_ARG;
  c->kf_off.clear();
  c->kf_n.clear();
  c->kf_pose.clear();
  c->kf_stamp.clear();
  c->kf_pos_dirty = true;
  c->kf_used = 0;
  return LOOP_OK;
}

// loop_closure.cpp:18-40 — O(K) scalar scan on the host, as in the reference
int loop_fetch_closest(loop_ctx* c, int query_idx, double radius, double time_diff, int* closest_out) {
  if (!c || !closest_out || query_idx < 0 || query_idx >= (int)c->kf_n.size()) {
    lio::set_error("loop_fetch_closest: bad query index %d", query_idx);
    return LOOP_E_ARG;
  }
  const double* q = &c->kf_pose[16 * (size_t)query_idx];
  double shortest = radius * 3.0;
  int best = -1;
  for (size_t i = 0; i + 1 < c->kf_n.size(); ++i) {
    const double* k = &c->kf_pose[16 * i];
    const double dx = k[3] - q[3], dy = k[7] - q[7], dz = k[11] - q[11];
    const double d = std::sqrt(dx * dx + dy * dy + dz * dz);
    if (radius > d && time_diff < (c->kf_stamp[query_idx] - c->kf_stamp[i])) {
      if (d < shortest) {
        shortest = d;
        best = (int)i;
      }
    }
  }
  *closest_out = best;
  return LOOP_OK;
}

// the same for n queries in one launch (every keyframe of a tick, or all of them when the pose graph is re-checked): the
// translations and stamps live on the device and are refreshed only after keyframes were added or poses changed
int loop_fetch_closest_batch(loop_ctx* c, const int* query_idx, int n, double radius, double time_diff, int* closest_out) {
  int rc = check(c);
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!query_idx || !closest_out))) {
    lio::set_error("loop_fetch_closest_batch: bad arguments");
    return LOOP_E_ARG;
  }
  const int K = (int)c->kf_n.size();
  for (int k = 0; k < n; ++k)
    if (query_idx[k] < 0 || query_idx[k] >= K) {
      lio::set_error("loop_fetch_closest_batch: bad query index %d", query_idx[k]);
      return LOOP_E_ARG;
    }
  if (n == 0) return LOOP_OK;
  cudaStream_t s = c->stream;
  if (c->kf_pos_dirty) {
    std::vector<double> pos(4 * (size_t)K);
    for (int i = 0; i < K; ++i) {
      const double* p = &c->kf_pose[16 * (size_t)i];
      pos[4 * i] = p[3], pos[4 * i + 1] = p[7], pos[4 * i + 2] = p[11], pos[4 * i + 3] = c->kf_stamp[i];
    }
    if ((rc = c->kf_pos_dev.ensure(std::max<size_t>(pos.size(), 4)))) return rc;
    LOOP_CUDA(cudaMemcpyAsync(c->kf_pos_dev.p, pos.data(), sizeof(double) * pos.size(), cudaMemcpyHostToDevice, s));
    LOOP_CUDA(cudaStreamSynchronize(s));  // `pos` is a host temporary
    c->kf_pos_dirty = false;
  }
  // queries in, answers out: the tail of the 32-bit scratch buffers
  if ((rc = c->vals_a.ensure((size_t)n)) || (rc = c->vals_b.ensure((size_t)n))) return rc;
  int* q_dev = reinterpret_cast<int*>(c->vals_a.p);
  int* o_dev = reinterpret_cast<int*>(c->vals_b.p);
  LOOP_CUDA(cudaMemcpyAsync(q_dev, query_idx, sizeof(int) * (size_t)n, cudaMemcpyHostToDevice, s));
  k_fetch_closest<<<(n * 32 + 255) / 256, 256, 0, s>>>(c->kf_pos_dev.p, K, q_dev, n, radius, time_diff, o_dev);
  c->launches += 1;
  LOOP_CUDA(cudaGetLastError());
  LOOP_CUDA(cudaMemcpyAsync(closest_out, o_dev, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, s));
  LOOP_CUDA(cudaStreamSynchronize(s));
  return LOOP_OK;
}

// ---- utilities.hpp ---------------------------------------------------------------------------------------------------
int loop_transform_pcd(loop_ctx* c, const float* xyzi, int n, const double* T16, float* out_xyzi) {
  int rc = check(c);
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!xyzi || !out_xyzi)) || !T16) {
    lio::set_error("loop_transform_pcd: bad arguments");
    return LOOP_E_ARG;
  }
  return transform_host_cloud(c, xyzi, n, T16, out_xyzi);
}

int loop_voxelize(loop_ctx* c, const float* xyzi, int n, float leaf, float* out_xyzi, int* n_out) {
  int rc = check(c);
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!xyzi || !out_xyzi)) || !(leaf > 0.f) || !n_out) {
    lio::set_error("loop_voxelize: bad arguments");
    return LOOP_E_ARG;
  }
  *n_out = 0;
  if (n == 0) return LOOP_OK;
  if ((rc = voxelize_host_cloud(c, xyzi, n, leaf))) return rc;
  const GridDev& g = c->grids_host[0];
  *n_out = (int)g.vox_n;
  return copy_cloud_out(c, g.vox_off, g.vox_n, out_xyzi, n);
}

// ---- LoopClosure -------------------------------------------------------------------------------------------------------
int loop_set_src_dst(loop_ctx* c, int src_idx, int dst_idx, int range, float leaf, float* src_out, int* n_src, float* dst_out, int* n_dst,
                     int cap) {
  int rc = check(c);
  if (rc) return rc;
  if (range < 0 || !(leaf > 0.f)) {
    lio::set_error("loop_set_src_dst: bad range / leaf");
    return LOOP_E_ARG;
  }
  std::vector<JobHost> jobs = {submap_job(c, src_idx, range), submap_job(c, dst_idx, range)};
  if ((rc = voxelize_jobs(c, jobs, leaf))) return rc;
  if (n_src) *n_src = (int)c->grids_host[0].vox_n;
  if (n_dst) *n_dst = (int)c->grids_host[1].vox_n;
  if ((rc = copy_cloud_out(c, c->grids_host[0].vox_off, c->grids_host[0].vox_n, src_out, cap))) return rc;
  return copy_cloud_out(c, c->grids_host[1].vox_off, c->grids_host[1].vox_n, dst_out, cap);
}

int loop_icp_align(loop_ctx* c, const float* src_xyzi, int ns, const float* dst_xyzi, int nd, double score_thr, loop_result* out,
                   float* aligned_out, int* corr_out) {
  int rc = check(c);
  if (rc) return rc;
  if (ns < 0 || nd < 0 || (ns > 0 && !src_xyzi) || (nd > 0 && !dst_xyzi) || !out) {
    lio::set_error("loop_icp_align: bad arguments");
    return LOOP_E_ARG;
  }
  const float* clouds[2] = {src_xyzi, dst_xyzi};
  const int ns2[2] = {ns, nd};
  if ((rc = load_clouds(c, clouds, ns2, 2))) return rc;
  std::vector<PairHost> ph = {PairHost{0, 1}};
  if ((rc = run_icp(c, ph, score_thr, out))) return rc;
  if (aligned_out && (rc = fetch_aligned(c, 0, aligned_out, ns, nullptr))) return rc;
  if (corr_out && (rc = fetch_corr(c, 0, corr_out))) return rc;
  return LOOP_OK;
}

int loop_verify_batch(loop_ctx* c, const loop_pair* pairs, int n_pairs, int range, float leaf, double score_thr, loop_result* out) {
  int rc = check(c);
  if (rc) return rc;
  if (n_pairs < 0 || (n_pairs > 0 && (!pairs || !out)) || range < 0 || !(leaf > 0.f)) {
    lio::set_error("loop_verify_batch: bad arguments");
    return LOOP_E_ARG;
  }
  const int K = (int)c->kf_n.size();
  for (int i = 0; i < n_pairs; ++i)
    if (pairs[i].src_idx < 0 || pairs[i].src_idx >= K || pairs[i].dst_idx < 0 || pairs[i].dst_idx >= K) {
      lio::set_error("loop_verify_batch: pair %d names keyframe %d/%d, store holds %d", i, pairs[i].src_idx, pairs[i].dst_idx, K);
      return LOOP_E_ARG;
    }
  // chunks bounded by raw submap points so the workspace stays a few GB; identical submaps are built once per chunk
  const uint64_t kChunkRaw = 128ull << 20;
  int64_t it_sum = 0, src_sum = 0, q_sum = 0;
  int done = 0;
  while (done < n_pairs) {
    std::vector<JobHost> jobs;
    std::vector<PairHost> ph;
    std::map<int, int> job_of_center;
    uint64_t raw = 0;
    int i = done;
    for (; i < n_pairs; ++i) {
      int ids[2];
      const int centers[2] = {pairs[i].src_idx, pairs[i].dst_idx};
      uint64_t add = 0;
      std::vector<std::pair<int, JobHost>> fresh;  // (centre, job) not built yet in this chunk
      for (int s = 0; s < 2; ++s) {
        auto f = job_of_center.find(centers[s]);
        if (f != job_of_center.end()) {
          ids[s] = f->second;
          continue;
        }
        int found = -1;
        for (size_t k = 0; k < fresh.size(); ++k)
          if (fresh[k].first == centers[s]) found = (int)(jobs.size() + k);
        if (found >= 0) {
          ids[s] = found;
          continue;
        }
        JobHost j = submap_job(c, centers[s], range);
        for (int k : j.kf) add += c->kf_n[k];
        ids[s] = (int)(jobs.size() + fresh.size());
        fresh.emplace_back(centers[s], std::move(j));
      }
      if (i > done && raw + add > kChunkRaw) break;
      for (auto& f : fresh) {
        job_of_center[f.first] = (int)jobs.size();
        jobs.push_back(std::move(f.second));
      }
      raw += add;
      ph.push_back(PairHost{ids[0], ids[1]});
    }
    if ((rc = voxelize_jobs(c, jobs, leaf))) return rc;
    c->grid_cell = leaf;  // the targets are VoxelGrid outputs at this leaf: about one point per grid cell
    if ((rc = run_icp(c, ph, score_thr, out + done))) return rc;
    it_sum += c->st_iters;
    src_sum += c->st_src;
    q_sum += c->st_queries;
    done = i;
  }
  c->st_iters = it_sum;
  c->st_src = src_sum;
  c->st_queries = q_sum;
  return LOOP_OK;
}

int loop_get_cloud(loop_ctx* c, int pair, int which, float* out_xyzi, int cap, int* n_out) {
  int rc = check(c);
  if (rc) return rc;
  if (pair < 0 || pair >= c->n_pairs || which < 0 || which > 2) {
    lio::set_error("loop_get_cloud: pair %d is not part of the last batch chunk (%d pairs)", pair, c->n_pairs);
    return LOOP_E_STATE;
  }
  const PairDev& pr = c->pairs_host[pair];
  if (which == 2) return fetch_aligned(c, pair, out_xyzi, cap, n_out);
  uint32_t off = pr.src_off, n = pr.src_n;
  if (which == 1) {
    off = c->trees_host[pr.dst_job].src_off;
    n = c->trees_host[pr.dst_job].n;
  }
  if (n_out) *n_out = (int)n;
  return copy_cloud_out(c, off, n, out_xyzi, cap);
}

// ---- measurement hooks ---------------------------------------------------------------------------------------------------
int loop_timing_enable(loop_ctx* c, int on) {
  if (!c) return LOOP_E_ARG;
  c->timing = on != 0;
  return LOOP_OK;
}

int loop_timing_get(loop_ctx* c, const char* name, double* total_ms, int* launches) {
  if (!c || !name) return LOOP_E_ARG;
  cudaSetDevice(c->cfg.device);
  tcollect(c);
  static const char* names[LT_COUNT] = {"submap", "voxel", "tree", "nn", "cov", "solve", "score"};
  int id = -1;
  for (int i = 0; i < LT_COUNT; ++i)
    if (!std::strcmp(name, names[i])) id = i;
  if (id < 0) {
    lio::set_error("unknown loop timer '%s'", name);
    return LOOP_E_ARG;
  }
  if (total_ms) *total_ms = c->t_sum[id];
  if (launches) *launches = c->t_n[id];
  return LOOP_OK;
}

int loop_timing_reset(loop_ctx* c) {
  if (!c) return LOOP_E_ARG;
  cudaSetDevice(c->cfg.device);
  tcollect(c);
  for (int i = 0; i < LT_COUNT; ++i) {
    c->t_sum[i] = 0;
    c->t_n[i] = 0;
  }
  return LOOP_OK;
}

int64_t loop_launch_count(loop_ctx* c) { return c ? c->launches : 0; }

int loop_batch_stats(loop_ctx* c, int64_t* iterations, int64_t* src_points, int64_t* nn_queries) {
  if (!c) return LOOP_E_ARG;
  if (iterations) *iterations = c->st_iters;
  if (src_points) *src_points = c->st_src;
  if (nn_queries) *nn_queries = c->st_queries;
  return LOOP_OK;
}

}  // extern "C"

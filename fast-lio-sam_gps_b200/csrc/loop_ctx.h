// Host-side context of the C-ABI in include/loop_b200.h: keyframe store, batch workspace, stream.
// Internal header (not part of the boundary).
#pragma once
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>
#include <cuda_runtime.h>
#include "../../include/loop_b200.h"

namespace lio {
void set_error(const char* fmt, ...);
bool cuda_ok(cudaError_t e, const char* what);
}  // namespace lio

namespace loopb {

#define LOOP_CUDA(call)                                      \
  do {                                                       \
    if (!::lio::cuda_ok((call), #call)) return LOOP_E_CUDA;  \
  } while (0)

// growable device array; contents are NOT preserved by ensure() unless keep == true
template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t cap = 0;
  int ensure(size_t n, cudaStream_t s = nullptr, bool keep = false) {
    if (n <= cap) return LOOP_OK;
    size_t want = n + n / 4 + 256;
    T* q = nullptr;
    if (!::lio::cuda_ok(cudaMalloc(&q, want * sizeof(T)), "cudaMalloc(workspace)")) return LOOP_E_CUDA;
    if (keep && p && cap) {
      if (!::lio::cuda_ok(cudaMemcpyAsync(q, p, cap * sizeof(T), cudaMemcpyDeviceToDevice, s), "grow copy")) return LOOP_E_CUDA;
      cudaStreamSynchronize(s);
    }
    if (p) cudaFree(p);
    p = q;
    cap = want;
    return LOOP_OK;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
};

constexpr int kLeafPts = 8;  // points per tree leaf = one 128-byte line

// one cloud to voxelise: a list of keyframe segments (device ranges in the keyframe store) with their poses
struct JobDev {
  uint32_t raw_off;   // first point of the job in the raw (transformed, concatenated) buffer
  uint32_t raw_n;     // raw points
  uint32_t seg_off;   // first segment descriptor
  uint32_t seg_n;     // segments (keyframes)
};

struct SegDev {
  uint32_t src_off;  // offset in the keyframe store
  uint32_t n;        // points
  uint32_t dst_off;  // offset in the raw buffer
  uint32_t job;
  double T[12];      // pose_corrected, first three rows
};

// per-job voxel grid parameters, produced on the device
struct GridDev {
  int minb[3];
  int mul1, mul2;
  int overflow;      // PCL's index-overflow guard: the cloud is passed through unfiltered
  uint32_t n_finite; // finite raw points
  uint32_t sorted_off;  // first sorted position of the job
  uint32_t vox_off;     // first output point of the job in the voxelised buffer
  uint32_t vox_n;       // output points
  float lo[3], hi[3];   // bounding box of the raw cloud (finite points)
};

// per-pair ICP state, device resident
struct PairDev {
  uint32_t src_off, src_n;  // voxelised source cloud
  uint32_t dst_job;         // index of the target cloud's tree
  uint32_t cur_off;         // offset of the pair's working arrays (cur, match, md2), indexed by Hilbert position
  uint32_t src_pts_off;     // the source cloud in Hilbert order (tree point buffer: xyz + original index bits, leaf padded)
  uint32_t src_pad_n;       // its padded length (pad entries carry kPadIdx and are skipped)
  // accumulators (exact fixed point)
  long long ss[3], sd[3], cov[9], mse_fix, cnt;
  long long score_fix, score_n;
  float Tr[16], Tfinal[16];
  double prev_mse;
  int similar, iters, state, converged, active;
  int pad;
  int pad2[2];
};

struct TreeDev {
  uint32_t pts_off;   // first point (float4: xyz + original index bits) in the tree point buffer; leaves padded to kLeafPts
  uint32_t n;         // real points
  uint32_t nleaf;     // leaves holding real points
  uint32_t nleaf_pad; // power of two >= nleaf
  uint32_t node_off;  // first node (pairs of child boxes) in the node buffer; nleaf_pad entries, node 0 unused
  uint32_t src_off;   // the cloud's points in the voxelised buffer (original order)
  // uniform grid over the cloud (targets only): open-addressing table of (cell key -> first point of the cell), points of a
  // cell chained through gnext[] (indexed like the compact point space).  grid_mask == 0: no grid (the tree alone is used).
  uint32_t grid_off, grid_mask, cmp_off;
  uint32_t lbvh;      // 1: the nodes form a radix tree over the leaves' curve keys (explicit child references), 0: implicit balanced tree
};

struct __align__(16) GCell {
  unsigned long long key;  // packed cell coordinates, ~0: empty
  uint32_t head;           // original index of the first point of the cell
  uint32_t pad;
};

struct __align__(16) Node2 {  // boxes of the two children of a node
  float4 lo0, hi0, lo1, hi1;
};

struct IcpParams {
  double max_d2, trans_eps, fit_eps, rot_thr, mse_abs;
  int max_iter, max_similar, min_corr;
};

enum LoopTimer { LT_SUBMAP = 0, LT_VOXEL, LT_TREE, LT_NN, LT_COV, LT_SOLVE, LT_SCORE, LT_COUNT };

struct Ev {
  cudaEvent_t a, b;
  int id;
};

}  // namespace loopb

struct loop_ctx {
  loop_config cfg;
  loopb::IcpParams prm;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  int n_sms = 0;
  int tree_order = 1;  // 0: Morton, 1: Hilbert order of the target trees (LOOP_TREE_ORDER=morton|hilbert)

  // keyframe store
  loopb::DevBuf<float4> kf_pts;
  size_t kf_used = 0;
  std::vector<uint32_t> kf_off, kf_n;
  std::vector<double> kf_pose;  // 16 per keyframe
  std::vector<double> kf_stamp;
  loopb::DevBuf<double> kf_pos_dev;   // x y z stamp per keyframe, for loop_fetch_closest_batch (refreshed when kf_pos_dirty)
  bool kf_pos_dirty = true;

  // voxelisation workspace
  loopb::DevBuf<loopb::SegDev> segs;
  loopb::DevBuf<loopb::JobDev> jobs;
  loopb::DevBuf<loopb::GridDev> grids;
  loopb::DevBuf<float4> raw;
  loopb::DevBuf<unsigned long long> keys_a, keys_b;
  loopb::DevBuf<uint32_t> vals_a, vals_b, head_scan, blk2job;
  loopb::DevBuf<unsigned char> cub_tmp;
  loopb::DevBuf<float4> vox;  // voxelised clouds, all jobs of the chunk
  std::vector<loopb::GridDev> grids_host;
  int n_jobs = 0;

  // ICP workspace
  loopb::DevBuf<loopb::TreeDev> trees;
  loopb::DevBuf<float4> tree_pts;
  loopb::DevBuf<loopb::Node2> nodes;
  loopb::DevBuf<uint32_t> leaf_code, node_flag;  // radix-tree build scratch (curve key per leaf / fill state per node, all trees)
  int tree_kind = 1;                  // LOOP_TREE=balanced: implicit balanced tree (A/B); default: radix tree (LBVH)
  loopb::DevBuf<loopb::GCell> gtab;   // grids of the target clouds
  loopb::DevBuf<uint32_t> gnext;      // next point of the same cell (original index inside the cloud), 0xFFFFFFFF: end
  float grid_cell = 0.3f;             // cell edge of the target grids (the VoxelGrid leaf when the clouds come from it)
  int use_grid = 0;                   // LOOP_GRID=1: seeded searches probe the target grid first (measured no faster than the seeded tree walk: off)
  int sample_first = 1;               // LOOP_SAMPLE_FIRST=0: the first iteration searches every point cold (A/B)
  loopb::DevBuf<loopb::PairDev> pairs;
  loopb::DevBuf<float4> cur;
  loopb::DevBuf<int> match;
  loopb::DevBuf<float> md2;
  loopb::DevBuf<uint32_t> blk2pair;  // 2 x blocks: pair, first point of the block
  std::vector<loopb::PairDev> pairs_host;
  std::vector<loopb::TreeDev> trees_host;
  int n_pairs = 0;
  int* active_host = nullptr;  // mapped pinned: number of pairs still iterating
  int* active_dev = nullptr;

  // stats of the last batch
  int64_t st_iters = 0, st_src = 0, st_queries = 0;

  // timing
  bool timing = false;
  std::vector<loopb::Ev> ev_used, ev_free;
  double t_sum[loopb::LT_COUNT]{};
  int t_n[loopb::LT_COUNT]{};
  int64_t launches = 0;
};

namespace loopb {

void tbegin(loop_ctx* c, int id);
void tend(loop_ctx* c);

// loop_voxel.cu ------------------------------------------------------------------------------------------------
// Fills c->vox / c->grids_host for `jobs_host` (each a list of segments).  raw_host (optional) supplies the raw cloud of
// every job directly from HOST memory (segments ignored); used by loop_voxelize.
struct JobHost {
  std::vector<int> kf;  // keyframe indices, in concatenation order
};
int voxelize_jobs(loop_ctx* c, const std::vector<JobHost>& jobs_host, float leaf);
int voxelize_host_cloud(loop_ctx* c, const float* xyzi, int n, float leaf);
int transform_host_cloud(loop_ctx* c, const float* xyzi, int n, const double* T16, float* out);
int transform_store_inplace(loop_ctx* c, size_t off, int n, const double* T16);  // keyframe store points [off, off + n) <- T * p
// loads explicit clouds as "already voxelised" jobs (loop_icp_align)
int load_clouds(loop_ctx* c, const float* const* clouds, const int* ns, int n_clouds);

// loop_icp.cu --------------------------------------------------------------------------------------------------
struct PairHost {
  int src_job, dst_job;
};
int run_icp(loop_ctx* c, const std::vector<PairHost>& ph, double score_thr, loop_result* out);
int fetch_aligned(loop_ctx* c, int pair, float* out, int cap, int* n_out);
int fetch_corr(loop_ctx* c, int pair, int* out);

}  // namespace loopb

// Host-side context of the C-ABI in include/lio_b200.h: owns the stream, the map storage, the scan
// buffers and the pinned result block.  Internal header (not part of the boundary).
#pragma once
#include <cstdint>
#include <cstdio>
#include <string>
#include <utility>
#include <vector>
#include <cuda_runtime.h>
#include "../../include/lio_b200.h"
#include "lio_device.cuh"
#include "lio_iekf.cuh"

namespace lio {

void set_error(const char* fmt, ...);
bool cuda_ok(cudaError_t e, const char* what);

#define LIO_CUDA(call)                                  \
  do {                                                  \
    if (!::lio::cuda_ok((call), #call)) return LIO_E_CUDA; \
  } while (0)

// result block written by the reduction kernel straight into mapped pinned host memory
struct HShareOut {
  double HTH[144];
  double HTz[12];
  double res_sum;
  double n_eff;
  unsigned long long seq;  // bumped by the kernel last, so the host can tell the block is complete
  unsigned long long pad;
};

enum TimerId { T_KNN = 0, T_PLANE, T_INSERT, T_CLASSIFY, T_SORT, T_COUNT };

struct EventPair {
  cudaEvent_t a, b;
  int id;
};

}  // namespace lio

struct lio_ctx {
  lio_config cfg;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  int n_sms = 0;

  // map
  lio::MapView map{};
  uint32_t table_cap = 0;
  // scratch for insert batches
  float4* batch_pts = nullptr;      // staged batch (device)
  uint32_t* batch_ent = nullptr;    // hash slot of each batch point's cell
  uint32_t* batch_tslot = nullptr;  // slot in the per-batch voxel table
  unsigned long long* batch_ebest = nullptr;  // best existing (d,id) of the point's voxel
  uint32_t* batch_ecnt = nullptr;             // existing points in the voxel
  uint32_t batch_cap = 0;
  unsigned long long* vox_keys = nullptr;  // per-batch voxel arbitration table
  unsigned long long* vox_best = nullptr;
  uint32_t vox_cap = 0;
  float* boxes_dev = nullptr;  // up to 64 boxes
  uint32_t* map_err_host = nullptr;  // mapped pinned word: MapCounters::error | kWarnIdsHigh, written by the last kernel of every add
  uint32_t* map_err_dev = nullptr;
  uint64_t dead_since_compact = 0;   // points removed by delete_boxes since the last compaction
  uint32_t last_deleted = 0;         // points removed by the last delete_boxes
  bool auto_compact = true;          // LIO_MAP_COMPACT=0 disables the automatic compaction after a box delete

  // scan
  int scan_n = 0;
  bool nn_valid = false;
  float4* body = nullptr;   // max_scan
  float4* world = nullptr;  // max_scan (w = unused)
  float* scan_t = nullptr;    // max_scan: per-point time offsets (ms), allocated on first lio_scan_undistort
  double* imu_pose = nullptr; // IMU poses of the scan (22 doubles each)
  float4* nn = nullptr;     // max_scan * 5 (xyz, id bits)
  float* nn_d2 = nullptr;   // max_scan * 5
  int* nn_cnt = nullptr;    // max_scan
  uint8_t* sel = nullptr;   // max_scan
  float4* plane = nullptr;  // max_scan: a b c pd2
  float4* far_nn = nullptr;       // max_scan: unbounded nearest map point of the scan points whose search found none within max_sqdist (k_far_nn)
  double* knn_state_dev = nullptr;  // 36 doubles: the state of the last search of the scan (written by the search kernel)
  float4* pfit = nullptr;   // max_scan: a b c d of the plane fitted at the last search (a = NaN: no valid plane)
  double* rows = nullptr;   // max_scan * 13, allocated on first lio_h_share_rows
  uint8_t* cls = nullptr;   // map_incremental class per point
  uint32_t* blk_cnt = nullptr;  // per-block counts for the ordered compaction (2 x blocks)
  float4* list_add = nullptr;
  float4* list_nodown = nullptr;
  // tile kNN: curve order of the scan (lio_knn_tile.cuh)
  uint32_t* tile_keys[2] = {nullptr, nullptr};
  uint32_t* tile_vals[2] = {nullptr, nullptr};
  float4* tile_sorted = nullptr;             // the query points in curve order (k_sort_place)
  unsigned long long* tile_queue = nullptr;  // 2 * tile_cap pushed (tile << 32 | lane mask) tasks; a slot is zeroed by its taker
  uint32_t tile_prof_cap = 0;                // tickets recorded by LIO_TILE_PROF
  void* sort_scratch = nullptr;  // SortScratch (lio_scan.cu): coarse-bin histogram / offsets / cursors of the scan ordering
  uint32_t tile_cap = 0;
  bool order_valid = false;      // tile_vals[1] holds the order of the current scan
  double sort_state[36]{};         // the state the current order was computed at
  bool resort_each_search = true;  // the curve order is recomputed at the state of every search (LIO_RESORT=0: once per scan)
  std::vector<std::pair<int, void*>> sort_graphs;  // (padded size, cudaGraphExec_t) of the captured keys + sort passes
  bool sort_graph_ok = true;     // false after a failed capture (LIO_SORT_GRAPH=0 disables it up front)
  void* keys_args_host = nullptr;  // pinned KeysArgs staged by the graph's memcpy node
  void* keys_args_dev = nullptr;
  std::vector<std::pair<int, void*>> map_graphs;  // (2 * padded size + flg_ekf_inited, cudaGraphExec_t) of map_incremental
  bool map_graph_ok = true;      // LIO_MAP_GRAPH=0 disables the captured form
  int map_graph_launches = 0;    // kernels inside one map_incremental graph
  void* map_args_host = nullptr;
  void* map_args_dev = nullptr;
  uint4* tile_prof = nullptr;      // LIO_TILE_PROF=1: per-tile {cycles, passes | deferred<<8 | overflow<<16, staged points, cells probed}
  uint32_t* tile_stats = nullptr;  // [0] queries that went to the per-query fallback (debug/bench counter)
  int tile_flip = 0;             // which counter pair the next search uses
  int knn_mode = 1;              // 1 tile (default), 0 per-query groups (LIO_KNN_MODE=group)
  int batch_left_warp = 0;  // LIO_BATCH_LEFT=warp: generic query batches also use the warp-per-query leftover kernel
  double* state_dev = nullptr;  // 36 doubles
  double* partials = nullptr;   // reduction partials
  int partial_blocks = 0;

  // device-resident IEKF loop (lio_iekf.cuh)
  lio::IekfCtl* iekf_ctl = nullptr;  // device
  lio::IekfIO* io_host = nullptr;    // pinned + mapped
  lio::IekfIO* io_dev = nullptr;     // device alias of io_host
  unsigned long long io_seq = 0;
  int iekf_mode = 0;                 // 0 host loop over lio_h_share (default: measured faster), 1 device loop (LIO_IEKF_MODE=device)

  lio::HShareOut* out_host = nullptr;  // pinned + mapped
  lio::HShareOut* out_dev = nullptr;   // device alias of out_host
  unsigned long long out_seq = 0;

  // pinned staging for uploads
  float* stage_host = nullptr;
  size_t stage_bytes = 0;
  // lio_scan_prefetch: the next scan's H2D copy on its own stream, double-buffered, consumed by lio_scan_upload
  cudaStream_t copy_stream = nullptr;
  float* pf_dev[2] = {nullptr, nullptr};
  size_t pf_cap[2] = {0, 0};
  cudaEvent_t pf_copied[2] = {nullptr, nullptr};    // H2D of the slot finished (copy stream)
  cudaEvent_t pf_consumed[2] = {nullptr, nullptr};  // the slot was expanded into body (main stream)
  bool pf_used[2] = {false, false};
  int pf_next = 0, pf_slot = -1, pf_n = 0;
  const float* pf_ptr = nullptr;

  // fov_segment state
  bool lm_init = false;
  float lm_min[3]{}, lm_max[3]{};

  // timing
  bool timing = false;
  std::vector<lio::EventPair> ev_used;
  std::vector<lio::EventPair> ev_free;
  double t_sum[lio::T_COUNT]{};
  int t_n[lio::T_COUNT]{};
  int64_t launches = 0;
};

namespace lio {

// map (lio_map.cu)
int map_alloc(lio_ctx* c);
void map_free(lio_ctx* c);
int map_clear(lio_ctx* c);
int map_ensure_batch(lio_ctx* c, uint32_t n);
int map_add_device(lio_ctx* c, const float4* pts, uint32_t n_max, const uint32_t* n_dev, bool downsample);
int map_delete_boxes(lio_ctx* c, const float* boxes_host, int nb);
int map_read_counters(lio_ctx* c, MapCounters* out);
int map_export(lio_ctx* c, std::vector<float4>& out);
int map_compact(lio_ctx* c, bool renumber);
int map_check_health(lio_ctx* c);  // sticky capacity error -> LIO_E_CAPACITY; ids running out -> compaction with renumbering

// scan / measurement (lio_scan.cu)
int scan_alloc(lio_ctx* c);
void scan_free(lio_ctx* c);
int scan_knn(lio_ctx* c, const double* state_host, bool dev_loop = false);
int scan_plane_reduce(lio_ctx* c, const double* state_host, bool want_rows, bool refit, bool dev_loop = false, int iter_i = 0,
                      unsigned long long io_seq = 0);
int scan_classify_insert(lio_ctx* c, const double* state_host, int flg_ekf_inited, bool dev_loop = false);
int iekf_update_device(lio_ctx* c, double* state, double* P, double R, int max_iter, const double* limit, double* iter_states,
                       int* n_eff_trace, int* knn_trace, int* iters_done, bool with_map, int flg_ekf_inited);
int scan_undistort(lio_ctx* c, const float* t_ms_host, const double* pose_host, int n_pose, const double* end_state);
int knn_batch(lio_ctx* c, const float4* q_dev, int n, int k, float max_d2, float4* out_nn, float* out_d2, int* out_cnt);
int tile_ensure(lio_ctx* c, uint32_t n);
int tile_tasks_of_last_search(lio_ctx* c, uint32_t* n_tasks);

// timing helpers (lio_capi.cu)
void timer_begin(lio_ctx* c, int id);
void timer_end(lio_ctx* c);

// host esekf (esekf_host.cpp)
int iekf_update_host(lio_ctx* c, double* state, double* P, double R, int max_iter, const double* limit, double* iter_states,
                     int* n_eff_trace, int* knn_trace, int* iters_done);

}  // namespace lio

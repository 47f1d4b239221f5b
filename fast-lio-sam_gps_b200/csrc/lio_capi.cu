// C-ABI of include/lio_b200.h: context, map entry points, scan / measurement-model entry points.
#include <cmath>
#include <cstdarg>
#include <cstring>
#include <new>
#include "lio_ctx.h"

namespace lio {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

bool cuda_ok(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return true;
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return false;
}

void launch_expand_xyz(lio_ctx* c, const float* xyz_dev, int n, float4* out);

// ---- timing -----------------------------------------------------------------------------------------
void timer_begin(lio_ctx* c, int id) {
  if (!c->timing) return;
  EventPair ev;
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

void timer_end(lio_ctx* c) {
  if (!c->timing || c->ev_used.empty()) return;
  cudaEventRecord(c->ev_used.back().b, c->stream);
}

static void timer_collect(lio_ctx* c) {
  for (EventPair& ev : c->ev_used) {
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

static int check_ctx(lio_ctx* c) {
  if (!c) {
    set_error("null context");
    return LIO_E_ARG;
  }
  if (!cuda_ok(cudaSetDevice(c->cfg.device), "cudaSetDevice")) return LIO_E_CUDA;
  return LIO_OK;
}

// stage host xyz (stride 3) into a device float4 array through the raw upload buffer
static int upload_xyz(lio_ctx* c, const float* xyz, int n, float4* dst) {
  const size_t bytes = sizeof(float) * 3 * (size_t)n;
  if (bytes > c->stage_bytes) {
    if (c->stage_host) cudaFree(c->stage_host);
    c->stage_host = nullptr;
    c->stage_bytes = 0;
    LIO_CUDA(cudaMalloc(&c->stage_host, bytes));
    c->stage_bytes = bytes;
  }
  LIO_CUDA(cudaMemcpyAsync(c->stage_host, xyz, bytes, cudaMemcpyHostToDevice, c->stream));
  launch_expand_xyz(c, c->stage_host, n, dst);
  LIO_CUDA(cudaGetLastError());
  return LIO_OK;
}

}  // namespace lio

using namespace lio;

extern "C" {

const char* lio_version(void) { return "lio_b200 0.1 (sm_100a)"; }
const char* lio_last_error(void) { return lio::g_err; }

int lio_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
  return n;
}

int lio_ctx_create(const lio_config* cfg, lio_ctx** out) {
  if (!cfg || !out) {
    set_error("null argument");
    return LIO_E_ARG;
  }
  *out = nullptr;
  if (!(cfg->filter_size_map > 0.f) || cfg->max_points == 0 || cfg->max_scan == 0 || !(cfg->max_sqdist > 0.f) ||
      !std::isfinite(cfg->max_sqdist)) {
    set_error("bad config: filter_size_map, max_points, max_scan, max_sqdist must be positive and finite");
    return LIO_E_ARG;
  }
  int ndev = 0;
  if (!cuda_ok(cudaGetDeviceCount(&ndev), "cudaGetDeviceCount") || cfg->device < 0 || cfg->device >= ndev) {
    if (ndev > 0) set_error("device %d out of range (%d devices)", cfg->device, ndev);
    return LIO_E_CUDA;  // no CPU fallback: fail loudly
  }
  LIO_CUDA(cudaSetDevice(cfg->device));
  cudaDeviceProp prop;
  LIO_CUDA(cudaGetDeviceProperties(&prop, cfg->device));
  if (prop.major < 10) {
    set_error("device %d is sm_%d%d; this library carries sm_100a code only", cfg->device, prop.major, prop.minor);
    return LIO_E_CUDA;
  }
  lio_ctx* c = new (std::nothrow) lio_ctx();
  if (!c) return LIO_E_ARG;
  c->cfg = *cfg;
  c->n_sms = prop.multiProcessorCount;
  if (cfg->stream) {
    c->stream = (cudaStream_t)cfg->stream;
  } else {
    if (!cuda_ok(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking), "cudaStreamCreate")) {
      delete c;
      return LIO_E_CUDA;
    }
    c->own_stream = true;
  }
  int rc = map_alloc(c);
  if (rc == LIO_OK) rc = scan_alloc(c);
  if (rc == LIO_OK && !cuda_ok(cudaStreamSynchronize(c->stream), "ctx_create sync")) rc = LIO_E_CUDA;
  if (rc != LIO_OK) {
    lio_ctx_destroy(c);
    return rc;
  }
  *out = c;
  return LIO_OK;
}

void lio_ctx_destroy(lio_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->cfg.device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  timer_collect(c);
  for (EventPair& ev : c->ev_free) {
    cudaEventDestroy(ev.a);
    cudaEventDestroy(ev.b);
  }
  scan_free(c);
  map_free(c);
  if (c->stage_host) cudaFree(c->stage_host);
  for (int k = 0; k < 2; ++k) {
    if (c->pf_dev[k]) cudaFree(c->pf_dev[k]);
    if (c->pf_copied[k]) cudaEventDestroy(c->pf_copied[k]);
    if (c->pf_consumed[k]) cudaEventDestroy(c->pf_consumed[k]);
  }
  if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
  if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

int lio_ctx_sync(lio_ctx* c) {
  int rc = check_ctx(c);
  if (rc) return rc;
  MapCounters ctr;
  return map_read_counters(c, &ctr);
}

// ---- map ------------------------------------------------------------------------------------------
static int add_host_batch(lio_ctx* c, const float* xyz, int n, bool downsample) {
  int rc = map_ensure_batch(c, (uint32_t)n);
  if (rc) return rc;
  rc = upload_xyz(c, xyz, n, c->batch_pts);
  if (rc) return rc;
  return map_add_device(c, c->batch_pts, (uint32_t)n, nullptr, downsample);
}

int lio_map_build(lio_ctx* c, const float* xyz, int n) {
  int rc = check_ctx(c);
  if (rc) return rc;
  if (n < 0 || (n > 0 && !xyz)) {
    set_error("bad point batch");
    return LIO_E_ARG;
  }
  rc = map_clear(c);
  if (rc) return rc;
  c->nn_valid = false;
  if (n == 0) return LIO_OK;
  rc = add_host_batch(c, xyz, n, false);
  if (rc) return rc;
  return lio_ctx_sync(c);
}

int lio_map_add(lio_ctx* c, const float* xyz, int n, int downsample_on, int* net_new) {
  int rc = check_ctx(c);
  if (rc) return rc;
  if (n < 0 || (n > 0 && !xyz)) {
    set_error("bad point batch");
    return LIO_E_ARG;
  }
  if (n == 0) {
    if (net_new) *net_new = 0;
    return LIO_OK;
  }
  rc = add_host_batch(c, xyz, n, downsample_on != 0);
  if (rc) return rc;
  MapCounters ctr;
  rc = map_read_counters(c, &ctr);  // also keeps `xyz` alive until the copy has happened
  if (rc) return rc;
  if (net_new) *net_new = downsample_on ? (int)ctr.batch_alive : n;
  return LIO_OK;
}

int lio_map_delete_boxes(lio_ctx* c, const float* boxes, int nb, int* n_deleted) {
  int rc = check_ctx(c);
  if (rc) return rc;
  if (nb < 0 || (nb > 0 && !boxes)) {
    set_error("bad box list");
    return LIO_E_ARG;
  }
  if (n_deleted) *n_deleted = 0;
  if (nb == 0) return LIO_OK;
  rc = map_delete_boxes(c, boxes, nb);
  if (rc) return rc;
  if (n_deleted) *n_deleted = (int)c->last_deleted;
  return LIO_OK;
}

int lio_map_compact(lio_ctx* c, int renumber_ids) {
  int rc = check_ctx(c);
  if (rc) return rc;
  return map_compact(c, renumber_ids != 0);
}

int lio_map_stats(lio_ctx* c, uint64_t* out8) {
  int rc = check_ctx(c);
  if (rc) return rc;
  if (!out8) {
    set_error("null output");
    return LIO_E_ARG;
  }
  MapCounters ctr;
  LIO_CUDA(cudaMemcpyAsync(&ctr, c->map.ctr, sizeof(MapCounters), cudaMemcpyDeviceToHost, c->stream));
  LIO_CUDA(cudaStreamSynchronize(c->stream));
  out8[0] = ctr.live > 0 ? (uint64_t)ctr.live : 0;
  out8[1] = ctr.next_id;
  out8[2] = ctr.pool_top;
  out8[3] = c->map.pool_cap;
  out8[4] = ctr.n_cells;
  out8[5] = c->map.cell_limit;
  out8[6] = ctr.error;
  out8[7] = c->dead_since_compact;
  return LIO_OK;
}

int lio_map_size(lio_ctx* c, int* n_live) {
  int rc = check_ctx(c);
  if (rc) return rc;
  MapCounters ctr;
  rc = map_read_counters(c, &ctr);
  if (rc) return rc;
  if (n_live) *n_live = ctr.live;
  return LIO_OK;
}

int lio_map_export(lio_ctx* c, float* xyz, uint32_t* ids, int cap, int* n_total) {
  int rc = check_ctx(c);
  if (rc) return rc;
  std::vector<float4> pts;
  rc = map_export(c, pts);
  if (rc) return rc;
  if (n_total) *n_total = (int)pts.size();
  const int m = std::min<int>(cap, (int)pts.size());
  for (int i = 0; i < m; ++i) {
    if (xyz) {
      xyz[3 * i] = pts[i].x;
      xyz[3 * i + 1] = pts[i].y;
      xyz[3 * i + 2] = pts[i].z;
    }
    if (ids) std::memcpy(&ids[i], &pts[i].w, 4);
  }
  return LIO_OK;
}

int lio_map_knn(lio_ctx* c, const float* q_xyz, int n, int k, float max_sqdist, uint32_t* ids, float* d2, float* pts, int* cnt) {
  int rc = check_ctx(c);
  if (rc) return rc;
  if (n < 0 || k < 1 || k > LIO_NUM_MATCH || !(max_sqdist > 0.f) || !std::isfinite(max_sqdist) || !ids || !d2 || (n > 0 && !q_xyz)) {
    set_error("lio_map_knn: need 1<=k<=5, finite max_sqdist > 0, ids and d2 outputs");
    return LIO_E_ARG;
  }
  if (std::sqrt(max_sqdist) / c->map.cell_size > 64.f) {
    set_error("lio_map_knn: max_sqdist spans more than 64 hash cells; use a coarser cell_mult");
    return LIO_E_ARG;
  }
  if (n == 0) return LIO_OK;
  float4 *q_dev = nullptr, *nn_dev = nullptr;
  float* d2_dev = nullptr;
  int* cnt_dev = nullptr;
  std::vector<float4> nn_host((size_t)n * k);
  std::vector<int> cnt_host(n);
  bool ok = cuda_ok(cudaMalloc(&q_dev, sizeof(float4) * (size_t)n), "malloc q") &&
            cuda_ok(cudaMalloc(&nn_dev, sizeof(float4) * (size_t)n * k), "malloc nn") &&
            cuda_ok(cudaMalloc(&d2_dev, sizeof(float) * (size_t)n * k), "malloc d2") &&
            cuda_ok(cudaMalloc(&cnt_dev, sizeof(int) * (size_t)n), "malloc cnt");
  rc = ok ? upload_xyz(c, q_xyz, n, q_dev) : LIO_E_CUDA;
  if (rc == LIO_OK) rc = knn_batch(c, q_dev, n, k, max_sqdist, nn_dev, d2_dev, cnt_dev);
  if (rc == LIO_OK) {
    ok = cuda_ok(cudaMemcpyAsync(nn_host.data(), nn_dev, sizeof(float4) * (size_t)n * k, cudaMemcpyDeviceToHost, c->stream), "d2h nn") &&
         cuda_ok(cudaMemcpyAsync(d2, d2_dev, sizeof(float) * (size_t)n * k, cudaMemcpyDeviceToHost, c->stream), "d2h d2") &&
         cuda_ok(cudaMemcpyAsync(cnt_host.data(), cnt_dev, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, c->stream), "d2h cnt") &&
         cuda_ok(cudaStreamSynchronize(c->stream), "knn sync");
    if (!ok) rc = LIO_E_CUDA;
  }
  cudaFree(q_dev);
  cudaFree(nn_dev);
  cudaFree(d2_dev);
  cudaFree(cnt_dev);
  if (rc) return rc;
  for (size_t i = 0; i < (size_t)n * k; ++i) {
    std::memcpy(&ids[i], &nn_host[i].w, 4);
    if (pts) {
      pts[3 * i] = nn_host[i].x;
      pts[3 * i + 1] = nn_host[i].y;
      pts[3 * i + 2] = nn_host[i].z;
    }
  }
  if (cnt) std::memcpy(cnt, cnt_host.data(), sizeof(int) * (size_t)n);
  return LIO_OK;
}

// ---- scan -------------------------------------------------------------------------------------------
int lio_scan_upload(lio_ctx* c, const float* body_xyz, int n) {
  int rc = check_ctx(c);
  if (rc) return rc;
  if (n < 0 || (uint32_t)n > c->cfg.max_scan || (n > 0 && !body_xyz)) {
    set_error("scan of %d points exceeds max_scan %u (or null pointer)", n, c->cfg.max_scan);
    return LIO_E_ARG;
  }
  c->scan_n = n;
  c->nn_valid = false;
  c->order_valid = false;
  if (n == 0) return LIO_OK;
  if (c->pf_slot >= 0 && c->pf_ptr == body_xyz && c->pf_n == n) {  // this scan was prefetched: no copy on the main stream
    const int k = c->pf_slot;
    c->pf_slot = -1;
    c->pf_ptr = nullptr;
    LIO_CUDA(cudaStreamWaitEvent(c->stream, c->pf_copied[k], 0));
    launch_expand_xyz(c, c->pf_dev[k], n, c->body);
    LIO_CUDA(cudaEventRecord(c->pf_consumed[k], c->stream));
    c->pf_used[k] = true;
    LIO_CUDA(cudaGetLastError());
    return LIO_OK;
  }
  return upload_xyz(c, body_xyz, n, c->body);
}

/* Starts the host->device copy of the NEXT scan on a separate copy stream (double-buffered staging), so it overlaps the
 * kernels of the scan being processed; lio_scan_upload with the same pointer and count then only waits for that copy.
 * The host buffer must stay unchanged (and, to overlap at all, pinned) until that lio_scan_upload. */
int lio_scan_prefetch(lio_ctx* c, const float* body_xyz, int n) {
  int rc = check_ctx(c);
  if (rc) return rc;
  if (n <= 0 || (uint32_t)n > c->cfg.max_scan || !body_xyz) {
    set_error("lio_scan_prefetch: %d points (max_scan %u) or null pointer", n, c->cfg.max_scan);
    return LIO_E_ARG;
  }
  if (!c->copy_stream) {
    LIO_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    for (int k = 0; k < 2; ++k) {
      LIO_CUDA(cudaEventCreateWithFlags(&c->pf_copied[k], cudaEventDisableTiming));
      LIO_CUDA(cudaEventCreateWithFlags(&c->pf_consumed[k], cudaEventDisableTiming));
    }
  }
  const int k = c->pf_next;
  const size_t bytes = sizeof(float) * 3 * (size_t)n;
  if (bytes > c->pf_cap[k]) {
    if (c->pf_dev[k]) cudaFree(c->pf_dev[k]);  // synchronises: nothing still reads the old buffer
    c->pf_dev[k] = nullptr;
    c->pf_cap[k] = 0;
    c->pf_used[k] = false;
    LIO_CUDA(cudaMalloc(&c->pf_dev[k], sizeof(float) * 3 * (size_t)c->cfg.max_scan));
    c->pf_cap[k] = sizeof(float) * 3 * (size_t)c->cfg.max_scan;
  }
  if (c->pf_used[k]) LIO_CUDA(cudaStreamWaitEvent(c->copy_stream, c->pf_consumed[k], 0));  // the slot's last reader is done
  LIO_CUDA(cudaMemcpyAsync(c->pf_dev[k], body_xyz, bytes, cudaMemcpyHostToDevice, c->copy_stream));
  LIO_CUDA(cudaEventRecord(c->pf_copied[k], c->copy_stream));
  c->pf_slot = k;
  c->pf_ptr = body_xyz;
  c->pf_n = n;
  c->pf_next = k ^ 1;
  return LIO_OK;
}

int lio_scan_set_device(lio_ctx* c, const void* body_xyzw_dev, int n) {
  int rc = check_ctx(c);
  if (rc) return rc;
  if (n < 0 || (uint32_t)n > c->cfg.max_scan || (n > 0 && !body_xyzw_dev)) {
    set_error("scan of %d points exceeds max_scan %u (or null pointer)", n, c->cfg.max_scan);
    return LIO_E_ARG;
  }
  c->scan_n = n;
  c->nn_valid = false;
  c->order_valid = false;
  if (n == 0) return LIO_OK;
  LIO_CUDA(cudaMemcpyAsync(c->body, body_xyzw_dev, sizeof(float4) * (size_t)n, cudaMemcpyDeviceToDevice, c->stream));
  return LIO_OK;
}

int lio_scan_undistort(lio_ctx* c, const float* t_ms, const double* imu_pose, int n_pose, const double* end_state) {
  int rc = check_ctx(c);
  if (rc) return rc;
  if (!t_ms || !imu_pose || !end_state) {
    set_error("lio_scan_undistort: null argument");
    return LIO_E_ARG;
  }
  if (c->scan_n <= 0) {
    set_error("no scan uploaded");
    return LIO_E_STATE;
  }
  c->nn_valid = false;
  c->order_valid = false;
  return scan_undistort(c, t_ms, imu_pose, n_pose, end_state);
}

int lio_scan_get_body(lio_ctx* c, float* body_xyz) {
  int rc = check_ctx(c);
  if (rc) return rc;
  const int n = c->scan_n;
  if (n <= 0 || !body_xyz) return LIO_OK;
  std::vector<float4> b((size_t)n);
  LIO_CUDA(cudaMemcpyAsync(b.data(), c->body, sizeof(float4) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
  LIO_CUDA(cudaStreamSynchronize(c->stream));
  for (int i = 0; i < n; ++i) {
    body_xyz[3 * i] = b[i].x;
    body_xyz[3 * i + 1] = b[i].y;
    body_xyz[3 * i + 2] = b[i].z;
  }
  return LIO_OK;
}

static int h_share_common(lio_ctx* c, const double* state, int redo_knn, bool want_rows) {
  if (!state) {
    set_error("null state");
    return LIO_E_ARG;
  }
  if (c->scan_n <= 0) {
    set_error("no scan uploaded");
    return LIO_E_STATE;
  }
  if (!redo_knn && !c->nn_valid) {
    set_error("redo_knn=0 but no cached neighbours for this scan");
    return LIO_E_STATE;
  }
  int rc = LIO_OK;
  if (redo_knn) rc = scan_knn(c, state);
  if (rc == LIO_OK) rc = scan_plane_reduce(c, state, want_rows, redo_knn != 0);
  if (rc) return rc;
  // The reduction's last block writes the result block to mapped host memory and bumps `seq` last: poll it instead of
  // paying a stream-synchronise wake-up per IEKF iteration; cudaStreamQuery surfaces a failed launch.
  {
    const volatile unsigned long long* seq = &c->out_host->seq;
    for (unsigned spin = 1; *seq != c->out_seq; ++spin) {
      if ((spin & 0x3FFu) == 0) {
        const cudaError_t q = cudaStreamQuery(c->stream);
        if (q == cudaSuccess) break;  // stream drained: the check below decides
        if (q != cudaErrorNotReady) LIO_CUDA(q);
      }
#if defined(__x86_64__)
      __builtin_ia32_pause();
#endif
    }
    if (want_rows) LIO_CUDA(cudaStreamSynchronize(c->stream));
  }
  if (c->out_host->seq != c->out_seq) {
    set_error("measurement kernel did not complete (seq %llu, expected %llu)", c->out_host->seq, c->out_seq);
    return LIO_E_CUDA;
  }
  return LIO_OK;
}

int lio_h_share(lio_ctx* c, const double* state, int redo_knn, double* HTH, double* HTz, int* n_eff, double* res_sum) {
  int rc = check_ctx(c);
  if (rc) return rc;
  rc = h_share_common(c, state, redo_knn, false);
  if (rc) return rc;
  const HShareOut* o = c->out_host;
  if (HTH) std::memcpy(HTH, o->HTH, sizeof(o->HTH));
  if (HTz) std::memcpy(HTz, o->HTz, sizeof(o->HTz));
  if (n_eff) *n_eff = (int)o->n_eff;
  if (res_sum) *res_sum = o->res_sum;
  return LIO_OK;
}

int lio_h_share_rows(lio_ctx* c, const double* state, int redo_knn, uint8_t* sel, double* rows, double* z, int* n_eff) {
  int rc = check_ctx(c);
  if (rc) return rc;
  rc = h_share_common(c, state, redo_knn, true);
  if (rc) return rc;
  const int n = c->scan_n;
  if (n_eff) *n_eff = (int)c->out_host->n_eff;
  if (sel) LIO_CUDA(cudaMemcpyAsync(sel, c->sel, (size_t)n, cudaMemcpyDeviceToHost, c->stream));
  std::vector<double> tmp;
  if (rows || z) {
    tmp.resize((size_t)n * 13);
    LIO_CUDA(cudaMemcpyAsync(tmp.data(), c->rows, sizeof(double) * 13 * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
  }
  LIO_CUDA(cudaStreamSynchronize(c->stream));
  for (int i = 0; i < n && (rows || z); ++i) {
    if (rows) std::memcpy(rows + (size_t)i * 12, tmp.data() + (size_t)i * 13, sizeof(double) * 12);
    if (z) z[i] = tmp[(size_t)i * 13 + 12];
  }
  return LIO_OK;
}

int lio_scan_get(lio_ctx* c, float* world, uint32_t* nn_ids, float* nn_xyz, int* nn_cnt, uint8_t* sel, float* plane) {
  int rc = check_ctx(c);
  if (rc) return rc;
  const int n = c->scan_n;
  if (n <= 0) return LIO_OK;
  std::vector<float4> w, nn, pl;
  std::vector<int> cnt((size_t)n);
  if (world) {
    w.resize(n);
    LIO_CUDA(cudaMemcpyAsync(w.data(), c->world, sizeof(float4) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
  }
  if (nn_ids || nn_xyz) {
    nn.resize((size_t)n * kK);
    LIO_CUDA(cudaMemcpyAsync(nn.data(), c->nn, sizeof(float4) * kK * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
  }
  LIO_CUDA(cudaMemcpyAsync(cnt.data(), c->nn_cnt, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
  if (sel) LIO_CUDA(cudaMemcpyAsync(sel, c->sel, (size_t)n, cudaMemcpyDeviceToHost, c->stream));
  if (plane) {
    pl.resize(n);
    LIO_CUDA(cudaMemcpyAsync(pl.data(), c->plane, sizeof(float4) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
  }
  LIO_CUDA(cudaStreamSynchronize(c->stream));
  for (int i = 0; i < n; ++i) {
    if (world) {
      world[3 * i] = w[i].x;
      world[3 * i + 1] = w[i].y;
      world[3 * i + 2] = w[i].z;
    }
    if (nn_cnt) nn_cnt[i] = cnt[i];
    if (plane) std::memcpy(plane + 4 * (size_t)i, &pl[i], 16);
    for (int k = 0; k < kK && (nn_ids || nn_xyz); ++k) {
      const size_t o = (size_t)i * kK + k;
      const bool ok = k < cnt[i];
      if (nn_ids) {
        uint32_t id = 0xFFFFFFFFu;
        if (ok) std::memcpy(&id, &nn[o].w, 4);
        nn_ids[o] = id;
      }
      if (nn_xyz) {
        nn_xyz[3 * o] = ok ? nn[o].x : 0.f;
        nn_xyz[3 * o + 1] = ok ? nn[o].y : 0.f;
        nn_xyz[3 * o + 2] = ok ? nn[o].z : 0.f;
      }
    }
  }
  return LIO_OK;
}

static int iekf_common(lio_ctx* c, double* state, double* P, double R, int max_iter, const double* limit, double* iter_states,
                       int* n_eff_trace, int* knn_trace, int* iters_done, bool with_map, int flg_ekf_inited) {
  int rc = check_ctx(c);
  if (rc) return rc;
  if (!state || !P || !limit || max_iter < 1 || !(R > 0)) {
    set_error("lio_iekf_update: bad arguments");
    return LIO_E_ARG;
  }
  if (c->scan_n <= 0) {
    set_error("no scan uploaded");
    return LIO_E_STATE;
  }
  rc = map_check_health(c);  // what the previous scan's inserts left in the mapped status word (capacity error, ids running out)
  if (rc) return rc;
  // default: host loop over lio_h_share (one mapped-memory poll per iteration, ~12 us; measured 0.53 ms/scan on B200).
  // LIO_IEKF_MODE=device enqueues the whole loop on the device (lio_iekf.cuh: no host round trip, but the 23x23 FP64
  // algebra then runs in one thread block, ~60 us per iteration -> 0.79 ms/scan): kept as a tested alternative.
  if (c->iekf_mode == 1 && c->knn_mode == 1 && max_iter + 1 <= kIekfMaxCalls)
    return iekf_update_device(c, state, P, R, max_iter, limit, iter_states, n_eff_trace, knn_trace, iters_done, with_map, flg_ekf_inited);
  rc = iekf_update_host(c, state, P, R, max_iter, limit, iter_states, n_eff_trace, knn_trace, iters_done);
  if (rc || !with_map) return rc;
  return scan_classify_insert(c, state, flg_ekf_inited);
}

int lio_iekf_update(lio_ctx* c, double* state, double* P, double R, int max_iter, const double* limit, double* iter_states,
                    int* n_eff_trace, int* knn_trace, int* iters_done) {
  return iekf_common(c, state, P, R, max_iter, limit, iter_states, n_eff_trace, knn_trace, iters_done, false, 0);
}

int lio_iekf_update_map(lio_ctx* c, double* state, double* P, double R, int max_iter, const double* limit, int flg_ekf_inited,
                        double* iter_states, int* n_eff_trace, int* knn_trace, int* iters_done) {
  return iekf_common(c, state, P, R, max_iter, limit, iter_states, n_eff_trace, knn_trace, iters_done, true, flg_ekf_inited);
}

int lio_map_incremental(lio_ctx* c, const double* state, int flg_ekf_inited, int* n_add, int* n_nodown) {
  int rc = check_ctx(c);
  if (rc) return rc;
  if (!state) {
    set_error("null state");
    return LIO_E_ARG;
  }
  if (c->scan_n <= 0) {
    set_error("no scan uploaded");
    return LIO_E_STATE;
  }
  if (!c->nn_valid) {
    // no search has run for this scan: Nearest_Points[i] is empty for every point (reference: `!Nearest_Points[i].empty()` is
    // false -> PointToAdd).  nn_cnt = -1 says exactly that; 0 would mean "searched, nothing within max_sqdist", for which
    // map_incremental looks up the far nearest neighbour.
    LIO_CUDA(cudaMemsetAsync(c->nn_cnt, 0xFF, sizeof(int) * (size_t)c->scan_n, c->stream));
  }
  rc = scan_classify_insert(c, state, flg_ekf_inited);
  if (rc) return rc;
  if (n_add || n_nodown) {
    MapCounters ctr;
    rc = map_read_counters(c, &ctr);
    if (rc) return rc;
    if (n_add) *n_add = (int)ctr.list_add;
    if (n_nodown) *n_nodown = (int)ctr.list_nodown;
  }
  return LIO_OK;
}

int lio_fov_segment(lio_ctx* c, const double* pos_lid, float* boxes_out, int* n_boxes) {
  int rc = check_ctx(c);
  if (rc) return rc;
  if (!pos_lid) {
    set_error("null position");
    return LIO_E_ARG;
  }
  if (n_boxes) *n_boxes = 0;
  const float kMov = 1.5f;  // MOV_THRESHOLD
  const double cube = c->cfg.cube_side_length;
  const float det = c->cfg.det_range;
  if (!c->lm_init) {
    for (int i = 0; i < 3; ++i) {
      c->lm_min[i] = (float)(pos_lid[i] - cube / 2.0);
      c->lm_max[i] = (float)(pos_lid[i] + cube / 2.0);
    }
    c->lm_init = true;
    return LIO_OK;
  }
  float d[3][2];
  bool need_move = false;
  for (int i = 0; i < 3; ++i) {
    d[i][0] = (float)std::fabs(pos_lid[i] - c->lm_min[i]);
    d[i][1] = (float)std::fabs(pos_lid[i] - c->lm_max[i]);
    if (d[i][0] <= kMov * det || d[i][1] <= kMov * det) need_move = true;
  }
  if (!need_move) return LIO_OK;
  const float mov = (float)std::max((cube - 2.0 * kMov * det) * 0.5 * 0.9, (double)(det * (kMov - 1)));
  float boxes[18];
  float nmin[3], nmax[3];
  for (int i = 0; i < 3; ++i) {
    nmin[i] = c->lm_min[i];
    nmax[i] = c->lm_max[i];
  }
  int nb = 0;
  for (int i = 0; i < 3; ++i) {
    float b[6] = {c->lm_min[0], c->lm_min[1], c->lm_min[2], c->lm_max[0], c->lm_max[1], c->lm_max[2]};
    if (d[i][0] <= kMov * det) {
      nmax[i] -= mov;
      nmin[i] -= mov;
      b[i] = c->lm_max[i] - mov;
      std::memcpy(boxes + 6 * nb++, b, sizeof(b));
    } else if (d[i][1] <= kMov * det) {
      nmax[i] += mov;
      nmin[i] += mov;
      b[3 + i] = c->lm_min[i] + mov;
      std::memcpy(boxes + 6 * nb++, b, sizeof(b));
    }
  }
  for (int i = 0; i < 3; ++i) {
    c->lm_min[i] = nmin[i];
    c->lm_max[i] = nmax[i];
  }
  if (boxes_out && nb) std::memcpy(boxes_out, boxes, sizeof(float) * 6 * nb);
  if (n_boxes) *n_boxes = nb;
  if (nb) return map_delete_boxes(c, boxes, nb);
  return LIO_OK;
}

int lio_timing_enable(lio_ctx* c, int on) {
  if (!c) return LIO_E_ARG;
  c->timing = on != 0;
  return LIO_OK;
}

int lio_timing_get(lio_ctx* c, const char* name, double* mean_ms, int* launches) {
  if (!c || !name) return LIO_E_ARG;
  cudaSetDevice(c->cfg.device);
  timer_collect(c);
  int id = -1;
  if (!std::strcmp(name, "knn")) id = T_KNN;
  else if (!std::strcmp(name, "plane_reduce")) id = T_PLANE;
  else if (!std::strcmp(name, "insert")) id = T_INSERT;
  else if (!std::strcmp(name, "classify")) id = T_CLASSIFY;
  else if (!std::strcmp(name, "sort")) id = T_SORT;
  if (id < 0) {
    set_error("unknown timer '%s'", name);
    return LIO_E_ARG;
  }
  if (mean_ms) *mean_ms = c->t_n[id] ? c->t_sum[id] / c->t_n[id] : 0.0;
  if (launches) *launches = c->t_n[id];
  return LIO_OK;
}

int lio_timing_reset(lio_ctx* c) {
  if (!c) return LIO_E_ARG;
  cudaSetDevice(c->cfg.device);
  timer_collect(c);
  for (int i = 0; i < T_COUNT; ++i) {
    c->t_sum[i] = 0;
    c->t_n[i] = 0;
  }
  return LIO_OK;
}

int lio_knn_stats(lio_ctx* c, uint64_t* fallback_queries) {
  int rc = check_ctx(c);
  if (rc) return rc;
  uint32_t v[4] = {0, 0, 0, 0};
  LIO_CUDA(cudaMemcpyAsync(v, c->tile_stats, sizeof(v), cudaMemcpyDeviceToHost, c->stream));
  LIO_CUDA(cudaStreamSynchronize(c->stream));
  if (fallback_queries) *fallback_queries = v[0];
  return LIO_OK;
}

int lio_knn_stats_detail(lio_ctx* c, uint64_t* out4) {
  int rc = check_ctx(c);
  if (rc) return rc;
  uint32_t v[4] = {0, 0, 0, 0};
  LIO_CUDA(cudaMemcpyAsync(v, c->tile_stats, sizeof(v), cudaMemcpyDeviceToHost, c->stream));
  LIO_CUDA(cudaStreamSynchronize(c->stream));
  if (out4)
    for (int i = 0; i < 4; ++i) out4[i] = v[i];
  return LIO_OK;
}

int lio_tile_profile(lio_ctx* c, uint32_t* out8, int cap_tiles, int* n_tiles) {
  int rc = check_ctx(c);
  if (rc) return rc;
  if (!c->tile_prof) {
    set_error("lio_tile_profile: create the context with LIO_TILE_PROF=1");
    return LIO_E_STATE;
  }
  uint32_t tasks = 0;
  rc = tile_tasks_of_last_search(c, &tasks);
  if (rc) return rc;
  const int tiles = (int)std::min(tasks, c->tile_prof_cap);
  if (n_tiles) *n_tiles = tiles;
  const int m = std::min(tiles, cap_tiles);
  if (out8 && m > 0) {
    LIO_CUDA(cudaMemcpyAsync(out8, c->tile_prof, sizeof(uint4) * 2 * (size_t)m, cudaMemcpyDeviceToHost, c->stream));
    LIO_CUDA(cudaStreamSynchronize(c->stream));
  }
  return LIO_OK;
}

int64_t lio_launch_count(lio_ctx* c) { return c ? c->launches : 0; }

}  // extern "C"

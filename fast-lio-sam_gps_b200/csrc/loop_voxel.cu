// Submap accumulation + VoxelGrid downsample — kernels K6 of SURVEY.md §2.2.
// Reference: LoopClosure::setSrcAndDstCloud (fast_lio_sam/src/loop_closure.cpp:42-67), transformPcd / voxelizePcd
// (fast_lio_sam/include/utilities.hpp:132-171); pcl::VoxelGrid::applyFilter and pcl::transformPointCloud semantics per
// SURVEY.md §9.5 (PCL 1.10 is not in the reference mount).
//
// All clouds ("jobs") of a batch are processed together:
//   k_submap    : every raw point = pose_corrected * keyframe point (f64 math, f32 store), written at its place in the
//                 concatenated cloud of its job; per-job bounding box and finite-point count by warp-aggregated atomics;
//   k_grid      : per job, PCL's min/max box indices, the division multipliers and its index-overflow guard;
//   k_keys      : 64-bit sort key (job << 32 | linear voxel index), value = raw index;
//   radix sort  : one stable device-wide sort for the whole batch (cub::DeviceRadixSort — the only library call on this path);
//   k_heads/scan: voxel heads -> output slots; k_centroids: one thread per voxel sums its points in ORIGINAL cloud order
//                 in f32 (the order the oracle defines as canonical) and divides by the count.
#include <algorithm>
#include <cstring>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include "loop_ctx.h"
#include "loop_device.cuh"

namespace loopb {

namespace {

__global__ void k_box_init(BoxAcc* acc, int n_jobs) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_jobs) return;
  for (int a = 0; a < 3; ++a) {
    acc[j].lo[a] = 0xFFFFFFFFu;
    acc[j].hi[a] = 0u;
  }
  acc[j].n_finite = 0;
}

__device__ __forceinline__ int find_seg(const SegDev* segs, int n_segs, uint32_t i) {
  int lo = 0, hi = n_segs - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (segs[mid].dst_off <= i) lo = mid; else hi = mid - 1;
  }
  return lo;
}

__device__ __forceinline__ int find_job(const JobDev* jobs, int n_jobs, uint32_t i) {
  int lo = 0, hi = n_jobs - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (jobs[mid].raw_off <= i) lo = mid; else hi = mid - 1;
  }
  return lo;
}

// raw[i] = T_seg * kf_pts[...]; with segs == nullptr the raw cloud is already in place (host-supplied)
__global__ void __launch_bounds__(256) k_submap(const SegDev* __restrict__ segs, int n_segs, const JobDev* __restrict__ jobs, int n_jobs,
                                                const float4* __restrict__ kf_pts, float4* __restrict__ raw, uint32_t n_raw,
                                                BoxAcc* __restrict__ acc) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  int job = -1;
  float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
  bool finite = false;
  if (i < n_raw) {
    if (segs) {
      const int s = find_seg(segs, n_segs, i);
      const SegDev& sg = segs[s];
      job = (int)sg.job;
      const float4 p = __ldg(kf_pts + sg.src_off + (i - sg.dst_off));
      const double x = p.x, y = p.y, z = p.z;
      const double* T = sg.T;
      o.x = (float)(((T[0] * x + T[1] * y) + T[2] * z) + T[3]);
      o.y = (float)(((T[4] * x + T[5] * y) + T[6] * z) + T[7]);
      o.z = (float)(((T[8] * x + T[9] * y) + T[10] * z) + T[11]);
      o.w = p.w;
      raw[i] = o;
    } else {
      job = find_job(jobs, n_jobs, i);
      o = raw[i];
    }
    finite = isfinite(o.x) && isfinite(o.y) && isfinite(o.z);
  }
  block_box_update(acc, job, finite, o.x, o.y, o.z);
}

// PCL VoxelGrid::applyFilter prologue, one thread for all jobs (n_jobs is small); also the sorted offsets
// idx_bits_out: bits needed by the largest linear voxel index of any job (the sort key is job << idx_bits | index, so the
// radix passes cover only bits that can differ)
__global__ void k_grid(const BoxAcc* __restrict__ acc, GridDev* __restrict__ grids, const JobDev* __restrict__ jobs, int n_jobs, float leaf,
                       uint32_t* __restrict__ idx_bits_out) {
  if (blockIdx.x || threadIdx.x) return;
  const float inv = 1.0f / leaf;
  uint32_t run = 0;
  unsigned long long max_cells = 2;
  for (int j = 0; j < n_jobs; ++j) {
    GridDev g;
    g.n_finite = acc[j].n_finite;
    g.sorted_off = run;
    run += g.n_finite;
    g.overflow = 0;
    g.vox_off = g.vox_n = 0;
    g.mul1 = g.mul2 = 0;
    long long d[3] = {1, 1, 1};
    for (int a = 0; a < 3; ++a) {
      g.lo[a] = g.n_finite ? ord2f(acc[j].lo[a]) : 0.f;
      g.hi[a] = g.n_finite ? ord2f(acc[j].hi[a]) : 0.f;
      g.minb[a] = (int)floorf(g.lo[a] * inv);
      const int maxb = (int)floorf(g.hi[a] * inv);
      d[a] = (long long)maxb - g.minb[a] + 1;
    }
    if (d[0] * d[1] * d[2] > 2147483647LL) {
      g.overflow = 1;
      max_cells = max(max_cells, (unsigned long long)jobs[j].raw_n);
    } else {
      g.mul1 = (int)d[0];
      g.mul2 = (int)(d[0] * d[1]);
      max_cells = max(max_cells, (unsigned long long)(d[0] * d[1] * d[2]));
    }
    grids[j] = g;
  }
  uint32_t bits = 1;
  while (bits < 32 && (1ull << bits) < max_cells) ++bits;
  *idx_bits_out = bits;
}

template <typename KeyT>
__global__ void __launch_bounds__(256) k_keys(const float4* __restrict__ raw, uint32_t n_raw, const JobDev* __restrict__ jobs, int n_jobs,
                                              const GridDev* __restrict__ grids, float leaf, uint32_t sentinel, int shift,
                                              KeyT* __restrict__ keys, uint32_t* __restrict__ vals) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_raw) return;
  const float4 p = raw[i];
  const int j = find_job(jobs, n_jobs, i);
  KeyT key;
  if (!(isfinite(p.x) && isfinite(p.y) && isfinite(p.z))) {
    key = (KeyT)sentinel << shift;
  } else {
    const GridDev& g = grids[j];
    uint32_t lin;
    if (g.overflow) {
      lin = i - jobs[j].raw_off;  // every point its own voxel, original order: the cloud passes through unfiltered
    } else {
      const float inv = 1.0f / leaf;
      const int i0 = (int)(floorf(p.x * inv) - (float)g.minb[0]);
      const int i1 = (int)(floorf(p.y * inv) - (float)g.minb[1]);
      const int i2 = (int)(floorf(p.z * inv) - (float)g.minb[2]);
      lin = (uint32_t)(i0 + i1 * g.mul1 + i2 * g.mul2);
    }
    key = ((KeyT)(uint32_t)j << shift) | (KeyT)lin;
  }
  keys[i] = key;
  vals[i] = i;
}

template <typename KeyT>
__global__ void __launch_bounds__(256) k_heads(const KeyT* __restrict__ keys, uint32_t n, uint32_t sentinel, int shift,
                                               uint32_t* __restrict__ flags) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const KeyT k = keys[i];
  const bool live = (uint32_t)(k >> shift) != sentinel;
  flags[i] = (live && (i == 0 || keys[i - 1] != k)) ? 1u : 0u;
}

// sorted[i] = raw[vals[i]]: the random 16-byte gathers of the sorted order, one per thread with nothing depending on
// them, so the centroid pass below reads its voxel's points as a contiguous run (it used to gather inside its serial
// per-voxel loop: 3.1 ms of a 256-pair batch)
__global__ void __launch_bounds__(256) k_gather_sorted(const uint32_t* __restrict__ vals, uint32_t n, const float4* __restrict__ raw,
                                                       float4* __restrict__ sorted) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) sorted[i] = __ldg(raw + vals[i]);
}

// head_pos[v] = sorted position of the first point of output voxel v
__global__ void __launch_bounds__(256) k_headpos(const uint32_t* __restrict__ flags, const uint32_t* __restrict__ slot, uint32_t n,
                                                 uint32_t* __restrict__ head_pos) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && flags[i]) head_pos[slot[i]] = i;
}

// one thread per OUTPUT voxel (every lane has work; the per-element form left 7 of 8 lanes idle at 8.5 points per voxel).
// A voxel's points are a contiguous run of the sorted order and MUST be added one after the other in that order (the f32
// sums have to round like the sequential reference's), so a run is a serial chain.  Its end is known up front (the next
// voxel's head), which takes the key comparison - a dependent load per point - out of the loop; runs longer than kLongRun
// (the voxels next to the sensor collect thousands of returns from the 25 keyframes of a submap: one such run kept the
// launch alive for 1.3 ms whatever the batch size) are handed to k_centroids_long.
constexpr uint32_t kLongRun = 128;
template <typename KeyT>
__device__ __forceinline__ uint32_t run_end(const KeyT* __restrict__ keys, const uint32_t* __restrict__ head_pos, uint32_t v, uint32_t n_vox,
                                            uint32_t i, uint32_t n) {
  if (v + 1 < n_vox) return head_pos[v + 1];
  const KeyT k = keys[i];  // last voxel: its run ends where the keys change (sentinel keys of non-finite points sort behind it)
  uint32_t lo = i, hi = n;
  while (lo + 1 < hi) {
    const uint32_t mid = lo + (hi - lo) / 2;
    if (keys[mid] == k) lo = mid; else hi = mid;
  }
  return hi;
}

template <typename KeyT>
__global__ void __launch_bounds__(256) k_centroids(const KeyT* __restrict__ keys, const float4* __restrict__ sorted,
                                                   const uint32_t* __restrict__ flags, const uint32_t* __restrict__ slot,
                                                   const uint32_t* __restrict__ head_pos, uint32_t n, float4* __restrict__ vox,
                                                   uint32_t* __restrict__ long_list) {
  const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t n_vox = slot[n - 1] + flags[n - 1];
  if (v >= n_vox) return;
  const uint32_t i = head_pos[v];
  const uint32_t e = run_end(keys, head_pos, v, n_vox, i, n);
  if (e - i > kLongRun) {
    const uint32_t pos = atomicAdd(long_list, 1u);
    long_list[1 + 2 * pos] = v;
    long_list[2 + 2 * pos] = e;
    return;
  }
  float sx = 0.f, sy = 0.f, sz = 0.f, si = 0.f;
  uint32_t j = i;
  for (; j + 4 <= e; j += 4) {  // loads first, then the additions in order
    const float4 p0 = sorted[j], p1 = sorted[j + 1], p2 = sorted[j + 2], p3 = sorted[j + 3];
    sx = sx + p0.x; sy = sy + p0.y; sz = sz + p0.z; si = si + p0.w;
    sx = sx + p1.x; sy = sy + p1.y; sz = sz + p1.z; si = si + p1.w;
    sx = sx + p2.x; sy = sy + p2.y; sz = sz + p2.z; si = si + p2.w;
    sx = sx + p3.x; sy = sy + p3.y; sz = sz + p3.z; si = si + p3.w;
  }
  for (; j < e; ++j) {
    const float4 p = sorted[j];
    sx = sx + p.x; sy = sy + p.y; sz = sz + p.z; si = si + p.w;
  }
  const float cnt = (float)(e - i);
  vox[v] = make_float4(sx / cnt, sy / cnt, sz / cnt, si / cnt);
}

// one warp per long run: 32 points per coalesced load (the next chunk is in flight while this one is added), the additions
// still strictly in run order (every lane carries the same sums; lane 0 writes)
__global__ void __launch_bounds__(256) k_centroids_long(const float4* __restrict__ sorted, const uint32_t* __restrict__ head_pos,
                                                        const uint32_t* __restrict__ long_list, float4* __restrict__ vox) {
  const int lane = threadIdx.x & 31;
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
  const uint32_t n_long = long_list[0];
  for (uint32_t r = warp; r < n_long; r += n_warps) {
    const uint32_t v = long_list[1 + 2 * r], e = long_list[2 + 2 * r];
    const uint32_t i = head_pos[v];
    float sx = 0.f, sy = 0.f, sz = 0.f, si = 0.f;
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 p = (i + lane < e) ? sorted[i + lane] : zero;
    for (uint32_t j0 = i; j0 < e; j0 += 32) {
      const uint32_t jn = j0 + 32 + lane;
      const float4 nxt = (jn < e) ? sorted[jn] : zero;
      const int cnt = (int)min(32u, e - j0);
      for (int t = 0; t < cnt; ++t) {
        sx = sx + __shfl_sync(0xFFFFFFFFu, p.x, t);
        sy = sy + __shfl_sync(0xFFFFFFFFu, p.y, t);
        sz = sz + __shfl_sync(0xFFFFFFFFu, p.z, t);
        si = si + __shfl_sync(0xFFFFFFFFu, p.w, t);
      }
      p = nxt;
    }
    const float cnt = (float)(e - i);
    if (lane == 0) vox[v] = make_float4(sx / cnt, sy / cnt, sz / cnt, si / cnt);
  }
}

// output ranges of the jobs in the voxelised buffer
__global__ void k_job_ranges(GridDev* __restrict__ grids, int n_jobs, const uint32_t* __restrict__ flags, const uint32_t* __restrict__ slot,
                             uint32_t n_raw) {
  if (blockIdx.x || threadIdx.x) return;
  uint32_t total_finite = 0;
  for (int j = 0; j < n_jobs; ++j) total_finite += grids[j].n_finite;
  const uint32_t total_vox = total_finite ? slot[total_finite - 1] + flags[total_finite - 1] : 0u;
  for (int j = 0; j < n_jobs; ++j) {
    GridDev& g = grids[j];
    const uint32_t b = g.n_finite ? slot[g.sorted_off] : 0u;
    const uint32_t e_pos = g.sorted_off + g.n_finite;
    const uint32_t e = (g.n_finite == 0) ? b : (e_pos < total_finite ? slot[e_pos] : total_vox);
    g.vox_off = g.n_finite ? b : (j ? grids[j - 1].vox_off + grids[j - 1].vox_n : 0u);
    g.vox_n = e - b;
  }
}

int grid_of(size_t n, int block) { return (int)std::max<size_t>(1, (n + block - 1) / block); }

}  // namespace

// ---------------------------------------------------------------------------------------------------------------------
// shared tail: raw cloud(s) are in c->raw, c->jobs describes them; produces c->vox + c->grids_host
static int voxelize_raw(loop_ctx* c, const std::vector<JobDev>& jd, const SegDev* segs_dev, int n_segs, float leaf) {
  const int n_jobs = (int)jd.size();
  const uint32_t n_raw = n_jobs ? jd.back().raw_off + jd.back().raw_n : 0;
  c->n_jobs = n_jobs;
  c->grids_host.assign(n_jobs, GridDev{});
  if (n_jobs == 0) return LOOP_OK;
  cudaStream_t s = c->stream;
  int rc;
  if ((rc = c->jobs.ensure(n_jobs))) return rc;
  if ((rc = c->grids.ensure(n_jobs))) return rc;
  LOOP_CUDA(cudaMemcpyAsync(c->jobs.p, jd.data(), sizeof(JobDev) * n_jobs, cudaMemcpyHostToDevice, s));
  // BoxAcc cells live at the head of the cub scratch buffer for the duration of the box pass
  DevBuf<unsigned char>& tmp = c->cub_tmp;
  if ((rc = tmp.ensure(sizeof(BoxAcc) * ((size_t)n_jobs + 1)))) return rc;  // + one cell for k_grid's scalar when there are no points
  BoxAcc* acc = reinterpret_cast<BoxAcc*>(tmp.p);
  k_box_init<<<grid_of(n_jobs, 128), 128, 0, s>>>(acc, n_jobs);
  c->launches += 1;
  if (n_raw) {
    if ((rc = c->keys_a.ensure(n_raw)) || (rc = c->keys_b.ensure(n_raw)) || (rc = c->vals_a.ensure(n_raw)) ||
        (rc = c->vals_b.ensure(n_raw)) || (rc = c->head_scan.ensure(n_raw)) || (rc = c->vox.ensure(n_raw)))
      return rc;
    tbegin(c, LT_SUBMAP);
    k_submap<<<grid_of(n_raw, 256), 256, 0, s>>>(segs_dev, n_segs, c->jobs.p, n_jobs, c->kf_pts.p, c->raw.p, n_raw, acc);
    tend(c);
    c->launches += 1;
  }
  tbegin(c, LT_VOXEL);
  // head_scan[0] carries the voxel-index width to the host (the buffer is not in use before the scan below)
  k_grid<<<1, 32, 0, s>>>(acc, c->grids.p, c->jobs.p, n_jobs, leaf, n_raw ? c->head_scan.p : reinterpret_cast<uint32_t*>(acc + n_jobs));
  c->launches += 1;
  if (n_raw) {
    int bits = 1;
    while ((1u << bits) - 1u < (uint32_t)n_jobs) ++bits;  // sentinel = 2^bits - 1 >= n_jobs
    const uint32_t sentinel = (1u << bits) - 1u;
    uint32_t shift = 32;
    LOOP_CUDA(cudaMemcpyAsync(&shift, c->head_scan.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    LOOP_CUDA(cudaStreamSynchronize(s));  // also orders k_grid (last reader of the BoxAcc cells) before a scratch reallocation
    const int end_bit = (int)shift + bits;
    const bool narrow = end_bit <= 32;  // job and voxel index fit one 32-bit key: 8 B instead of 12 B per element and pass
    uint32_t* ka32 = reinterpret_cast<uint32_t*>(c->keys_a.p);
    uint32_t* kb32 = reinterpret_cast<uint32_t*>(c->keys_b.p);
    size_t need = 0, need2 = 0;
    if (narrow) cub::DeviceRadixSort::SortPairs(nullptr, need, ka32, kb32, c->vals_a.p, c->vals_b.p, (int)n_raw, 0, end_bit, s);
    else cub::DeviceRadixSort::SortPairs(nullptr, need, c->keys_a.p, c->keys_b.p, c->vals_a.p, c->vals_b.p, (int)n_raw, 0, end_bit, s);
    cub::DeviceScan::ExclusiveSum(nullptr, need2, c->vals_a.p, c->head_scan.p, (int)n_raw, s);
    need = std::max(need, need2);
    if ((rc = tmp.ensure(need))) return rc;
    size_t bytes = tmp.cap;
    if (narrow) {
      k_keys<uint32_t><<<grid_of(n_raw, 256), 256, 0, s>>>(c->raw.p, n_raw, c->jobs.p, n_jobs, c->grids.p, leaf, sentinel, (int)shift, ka32, c->vals_a.p);
      LOOP_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, bytes, ka32, kb32, c->vals_a.p, c->vals_b.p, (int)n_raw, 0, end_bit, s));
      k_heads<uint32_t><<<grid_of(n_raw, 256), 256, 0, s>>>(kb32, n_raw, sentinel, (int)shift, c->vals_a.p);
    } else {
      k_keys<unsigned long long><<<grid_of(n_raw, 256), 256, 0, s>>>(c->raw.p, n_raw, c->jobs.p, n_jobs, c->grids.p, leaf, sentinel, (int)shift,
                                                                      c->keys_a.p, c->vals_a.p);
      LOOP_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, bytes, c->keys_a.p, c->keys_b.p, c->vals_a.p, c->vals_b.p, (int)n_raw, 0, end_bit, s));
      k_heads<unsigned long long><<<grid_of(n_raw, 256), 256, 0, s>>>(c->keys_b.p, n_raw, sentinel, (int)shift, c->vals_a.p);
    }
    bytes = tmp.cap;
    LOOP_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, bytes, c->vals_a.p, c->head_scan.p, (int)n_raw, s));
    if ((rc = c->cur.ensure(n_raw))) return rc;  // the ICP working buffer is idle here: scratch for the points in sorted order
    k_gather_sorted<<<grid_of(n_raw, 256), 256, 0, s>>>(c->vals_b.p, n_raw, c->raw.p, c->cur.p);
    k_headpos<<<grid_of(n_raw, 256), 256, 0, s>>>(c->vals_a.p, c->head_scan.p, n_raw, c->vals_b.p);  // vals_b was consumed by the gather
    // the voxel count is only known on the device: launch for the upper bound n_raw / 1 and let the surplus threads exit
    // list of the long runs: counter + (voxel, end) pairs in the sort's input key buffer, which is free by now (2 * n_raw / kLongRun + 1 words)
    uint32_t* long_list = reinterpret_cast<uint32_t*>(c->keys_a.p);
    LOOP_CUDA(cudaMemsetAsync(long_list, 0, sizeof(uint32_t), s));
    if (narrow) k_centroids<uint32_t><<<grid_of(n_raw, 256), 256, 0, s>>>(kb32, c->cur.p, c->vals_a.p, c->head_scan.p, c->vals_b.p, n_raw, c->vox.p, long_list);
    else k_centroids<unsigned long long><<<grid_of(n_raw, 256), 256, 0, s>>>(c->keys_b.p, c->cur.p, c->vals_a.p, c->head_scan.p, c->vals_b.p, n_raw, c->vox.p, long_list);
    k_centroids_long<<<c->n_sms * 4, 256, 0, s>>>(c->cur.p, c->vals_b.p, long_list, c->vox.p);
    k_job_ranges<<<1, 32, 0, s>>>(c->grids.p, n_jobs, c->vals_a.p, c->head_scan.p, n_raw);
    c->launches += 11;  // keys/heads/gather/headpos/centroids (2)/ranges + the library sort and scan passes (counted as one each)
  }
  tend(c);
  LOOP_CUDA(cudaGetLastError());
  LOOP_CUDA(cudaMemcpyAsync(c->grids_host.data(), c->grids.p, sizeof(GridDev) * n_jobs, cudaMemcpyDeviceToHost, s));
  LOOP_CUDA(cudaStreamSynchronize(s));
  return LOOP_OK;
}

namespace {
// pts[i] = T * pts[i] in place, the arithmetic of k_submap (f64 matrix applied to f32 points, transformPcd / pcl::transformPointCloud
// with a Matrix4d): the keyframe constructor's re-expression of the published world-frame cloud in the LiDAR frame
struct Mat34 {
  double T[12];
};
__global__ void __launch_bounds__(256) k_transform_inplace(float4* __restrict__ pts, uint32_t n, Mat34 m) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 p = pts[i];
  const double x = p.x, y = p.y, z = p.z;
  const double* T = m.T;
  float4 o;
  o.x = (float)(((T[0] * x + T[1] * y) + T[2] * z) + T[3]);
  o.y = (float)(((T[4] * x + T[5] * y) + T[6] * z) + T[7]);
  o.z = (float)(((T[8] * x + T[9] * y) + T[10] * z) + T[11]);
  o.w = p.w;
  pts[i] = o;
}
}  // namespace

int transform_store_inplace(loop_ctx* c, size_t off, int n, const double* T16) {
  if (n <= 0) return LOOP_OK;
  Mat34 m;
  std::memcpy(m.T, T16, sizeof(double) * 12);
  k_transform_inplace<<<grid_of((size_t)n, 256), 256, 0, c->stream>>>(c->kf_pts.p + off, (uint32_t)n, m);
  c->launches += 1;
  LOOP_CUDA(cudaGetLastError());
  return LOOP_OK;
}

int voxelize_jobs(loop_ctx* c, const std::vector<JobHost>& jobs_host, float leaf) {
  std::vector<JobDev> jd;
  std::vector<SegDev> sd;
  uint64_t off = 0;
  for (size_t j = 0; j < jobs_host.size(); ++j) {
    JobDev d;
    d.raw_off = (uint32_t)off;
    d.seg_off = (uint32_t)sd.size();
    for (int k : jobs_host[j].kf) {
      if (c->kf_n[k] == 0) continue;
      SegDev s;
      s.src_off = c->kf_off[k];
      s.n = c->kf_n[k];
      s.dst_off = (uint32_t)off;
      s.job = (uint32_t)j;
      std::memcpy(s.T, &c->kf_pose[16 * (size_t)k], sizeof(double) * 12);
      sd.push_back(s);
      off += s.n;
    }
    d.raw_n = (uint32_t)(off - d.raw_off);
    d.seg_n = (uint32_t)sd.size() - d.seg_off;
    jd.push_back(d);
  }
  if (off >= 0x7FFFFFFFull) {
    lio::set_error("batch chunk holds %llu raw points (limit 2^31-1): use smaller chunks", (unsigned long long)off);
    return LOOP_E_CAPACITY;
  }
  int rc;
  if ((rc = c->raw.ensure((size_t)off))) return rc;
  if (!sd.empty()) {
    if ((rc = c->segs.ensure(sd.size()))) return rc;
    LOOP_CUDA(cudaMemcpyAsync(c->segs.p, sd.data(), sizeof(SegDev) * sd.size(), cudaMemcpyHostToDevice, c->stream));
  }
  // sd must outlive the async copy: voxelize_raw ends with a stream synchronise
  return voxelize_raw(c, jd, sd.empty() ? nullptr : c->segs.p, (int)sd.size(), leaf);
}

int voxelize_host_cloud(loop_ctx* c, const float* xyzi, int n, float leaf) {
  std::vector<JobDev> jd(1);
  jd[0] = JobDev{0u, (uint32_t)n, 0u, 0u};
  int rc;
  if ((rc = c->raw.ensure((size_t)std::max(n, 1)))) return rc;
  if (n) LOOP_CUDA(cudaMemcpyAsync(c->raw.p, xyzi, sizeof(float4) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
  return voxelize_raw(c, jd, nullptr, 0, leaf);
}

int transform_host_cloud(loop_ctx* c, const float* xyzi, int n, const double* T16, float* out) {
  if (n == 0) return LOOP_OK;
  int rc;
  // stage the input behind the keyframe store's live range, transform into raw
  if ((rc = c->kf_pts.ensure(c->kf_used + (size_t)n, c->stream, true))) return rc;
  if ((rc = c->raw.ensure((size_t)n)) || (rc = c->segs.ensure(1)) || (rc = c->jobs.ensure(1)) || (rc = c->cub_tmp.ensure(sizeof(BoxAcc))))
    return rc;
  LOOP_CUDA(cudaMemcpyAsync(c->kf_pts.p + c->kf_used, xyzi, sizeof(float4) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
  SegDev s;
  s.src_off = (uint32_t)c->kf_used;
  s.n = (uint32_t)n;
  s.dst_off = 0;
  s.job = 0;
  std::memcpy(s.T, T16, sizeof(double) * 12);
  JobDev j{0u, (uint32_t)n, 0u, 1u};
  LOOP_CUDA(cudaMemcpyAsync(c->segs.p, &s, sizeof(s), cudaMemcpyHostToDevice, c->stream));
  LOOP_CUDA(cudaMemcpyAsync(c->jobs.p, &j, sizeof(j), cudaMemcpyHostToDevice, c->stream));
  BoxAcc* acc = reinterpret_cast<BoxAcc*>(c->cub_tmp.p);
  k_box_init<<<1, 32, 0, c->stream>>>(acc, 1);
  k_submap<<<grid_of(n, 256), 256, 0, c->stream>>>(c->segs.p, 1, c->jobs.p, 1, c->kf_pts.p, c->raw.p, (uint32_t)n, acc);
  c->launches += 2;
  LOOP_CUDA(cudaGetLastError());
  LOOP_CUDA(cudaMemcpyAsync(out, c->raw.p, sizeof(float4) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
  LOOP_CUDA(cudaStreamSynchronize(c->stream));
  return LOOP_OK;
}

// explicit clouds become the "voxelised" jobs of the batch, untouched (loop_icp_align)
int load_clouds(loop_ctx* c, const float* const* clouds, const int* ns, int n_clouds) {
  size_t total = 0;
  for (int i = 0; i < n_clouds; ++i) total += (size_t)ns[i];
  int rc;
  if ((rc = c->vox.ensure(std::max<size_t>(total, 1)))) return rc;
  c->n_jobs = n_clouds;
  c->grids_host.assign(n_clouds, GridDev{});
  size_t off = 0;
  for (int i = 0; i < n_clouds; ++i) {
    GridDev& g = c->grids_host[i];
    g.vox_off = (uint32_t)off;
    g.vox_n = (uint32_t)ns[i];
    if (ns[i]) LOOP_CUDA(cudaMemcpyAsync(c->vox.p + off, clouds[i], sizeof(float4) * (size_t)ns[i], cudaMemcpyHostToDevice, c->stream));
    off += (size_t)ns[i];
  }
  LOOP_CUDA(cudaStreamSynchronize(c->stream));
  return LOOP_OK;
}

}  // namespace loopb

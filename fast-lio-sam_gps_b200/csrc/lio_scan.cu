// Measurement-model kernels: K1 body->world, K3 kNN (lio_knn.cuh), K4 plane/residual/Jacobian
// (lio_plane.cuh), K5 FP64 reduction to H^T H / H^T z, and the device side of map_incremental.
// Reference: h_share_model / map_incremental in third_party/FAST_LIO/src/laserMapping.cpp
// (absent submodule; behaviour per SURVEY.md §9.1, §9.3; consumer per §9.6).
#include <algorithm>
#include <cstring>
#include "lio_ctx.h"
#include "lio_knn.cuh"
#include "lio_plane.cuh"

namespace lio {

namespace {

constexpr int kKnnBlock = 128;
constexpr int kRedBlock = 256;
constexpr int kRedWarps = kRedBlock / 32;
constexpr int kMaxVals = 92;  // 12 cols: 78 + 12 + 2

// ---------------------------------------------------------------- K1 + K3
template <int G>
__global__ void __launch_bounds__(kKnnBlock) k_scan_knn(MapView m, const float4* __restrict__ body, int n, StateArg st, float max_d2,
                                                        float4* __restrict__ nn, float* __restrict__ nn_d2,
                                                        int* __restrict__ nn_cnt) {
  const Group<G> g;
  const int groups_per_block = kKnnBlock / G;
  const int gid = blockIdx.x * groups_per_block + (threadIdx.x / G);
  const int gstride = gridDim.x * groups_per_block;
  for (int i = gid; i < n; i += gstride) {
    const float4 b = __ldg(body + i);
    float qx, qy, qz;
    body_to_world(st, b.x, b.y, b.z, qx, qy, qz);
    unsigned long long key;
    uint32_t ref;
    int cnt;
    knn_search<G>(m, g, qx, qy, qz, max_d2, key, ref, cnt);
    if (g.gl < kK) {
      float4 p = make_float4(0.f, 0.f, 0.f, __uint_as_float(kEmptyId));
      float d2 = __int_as_float(0x7f800000);
      if (g.gl < cnt) {
        p = __ldg(m.slots + ref);
        d2 = __uint_as_float((uint32_t)(key >> 32));
      }
      nn[(size_t)i * kK + g.gl] = p;
      if (nn_d2) nn_d2[(size_t)i * kK + g.gl] = d2;
    }
    if (g.gl == 0) nn_cnt[i] = cnt;
  }
}

// plain batched Nearest_Search on explicit query points (lio_map_knn)
template <int G>
__global__ void __launch_bounds__(kKnnBlock) k_knn_batch(MapView m, const float4* __restrict__ q, int n, int k, float max_d2,
                                                         float4* __restrict__ nn, float* __restrict__ nn_d2,
                                                         int* __restrict__ nn_cnt) {
  const Group<G> g;
  const int groups_per_block = kKnnBlock / G;
  const int gid = blockIdx.x * groups_per_block + (threadIdx.x / G);
  const int gstride = gridDim.x * groups_per_block;
  for (int i = gid; i < n; i += gstride) {
    const float4 b = __ldg(q + i);
    unsigned long long key;
    uint32_t ref;
    int cnt;
    knn_search<G>(m, g, b.x, b.y, b.z, max_d2, key, ref, cnt);
    cnt = min(cnt, k);
    if (g.gl < k) {
      float4 p = make_float4(0.f, 0.f, 0.f, __uint_as_float(kEmptyId));
      float d2 = __int_as_float(0x7f800000);
      if (g.gl < cnt) {
        p = __ldg(m.slots + ref);
        d2 = __uint_as_float((uint32_t)(key >> 32));
      }
      nn[(size_t)i * k + g.gl] = p;
      nn_d2[(size_t)i * k + g.gl] = d2;
    }
    if (g.gl == 0) nn_cnt[i] = cnt;
  }
}

// ---------------------------------------------------------------- K4 + K5
// One thread per scan point: plane fit, validity, residual, Jacobian row; then per-warp pair sums in
// FP64 over rows staged in shared memory, block tree, and a deterministic last-block pass.
template <int NCOL>
__global__ void __launch_bounds__(kRedBlock) k_plane_reduce(const float4* __restrict__ body, int n, StateArg st,
                                                            const float4* __restrict__ nn, const int* __restrict__ nn_cnt,
                                                            float4* __restrict__ world, uint8_t* __restrict__ sel,
                                                            float4* __restrict__ plane, double* __restrict__ rows_out,
                                                            double* __restrict__ partials, MapCounters* ctr, HShareOut* out,
                                                            unsigned long long seq) {
  constexpr int NC = NCOL + 3;                          // row | z | 1 | |pd2|
  constexpr int NV = NCOL * (NCOL + 1) / 2 + NCOL + 2;  // pair sums produced
  constexpr int VPL = (NV + 31) / 32;                   // values per lane
  __shared__ double s_rows[kRedWarps][32][NC];
  __shared__ double s_blk[kRedWarps][NV];
  __shared__ uint8_t s_pa[NV], s_pb[NV];
  __shared__ bool s_last;

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x < NV) {
    int v = threadIdx.x, a = 0, b = 0;
    if (v < NCOL * (NCOL + 1) / 2) {
      int rem = v;
      a = 0;
      while (rem >= NCOL - a) {
        rem -= NCOL - a;
        ++a;
      }
      b = a + rem;
    } else if (v < NCOL * (NCOL + 1) / 2 + NCOL) {
      a = v - NCOL * (NCOL + 1) / 2;
      b = NCOL;
    } else if (v == NV - 2) {
      a = b = NCOL + 1;
    } else {
      a = NCOL + 2;
      b = NCOL + 1;
    }
    s_pa[v] = (uint8_t)a;
    s_pb[v] = (uint8_t)b;
  }
  __syncthreads();

  double acc[VPL];
#pragma unroll
  for (int j = 0; j < VPL; ++j) acc[j] = 0.0;

  const double* s = st.s;
  const int tiles = (n + 31) / 32;
  for (int tile = blockIdx.x * kRedWarps + warp; tile < tiles; tile += gridDim.x * kRedWarps) {
    const int i = tile * 32 + lane;
    bool valid = false;
    double row[NC];
#pragma unroll
    for (int j = 0; j < NC; ++j) row[j] = 0.0;
    if (i < n) {
      const float4 b = __ldg(body + i);
      float wx, wy, wz;
      body_to_world(st, b.x, b.y, b.z, wx, wy, wz);
      world[i] = make_float4(wx, wy, wz, 0.f);
      float4 pl = make_float4(0.f, 0.f, 0.f, 0.f);
      if (nn_cnt[i] >= kK) {
        float P[kK][3];
#pragma unroll
        for (int k = 0; k < kK; ++k) {
          const float4 q = __ldg(nn + (size_t)i * kK + k);
          P[k][0] = q.x;
          P[k][1] = q.y;
          P[k][2] = q.z;
        }
        float pabcd[4];
        if (esti_plane(pabcd, P, 0.1f)) {
          float pd2 = pabcd[0] * wx + pabcd[1] * wy;
          pd2 = pd2 + pabcd[2] * wz;
          pd2 = pd2 + pabcd[3];
          const double bx = b.x, by = b.y, bz = b.z;
          const double range = sqrt((bx * bx + by * by) + bz * bz);
          const float sw = (float)(1.0 - 0.9 * fabs((double)pd2) / sqrt(range));
          if ((double)sw > 0.9) {
            valid = true;
            pl = make_float4(pabcd[0], pabcd[1], pabcd[2], pd2);
            // Jacobian row (f64): [n, A, B, C] with A = [R_LI p + t_LI]x * (R^T n)
            const double nx = pabcd[0], ny = pabcd[1], nz = pabcd[2];
            double pt[3], Cv[3];
#pragma unroll
            for (int r = 0; r < 3; ++r)
              pt[r] = ((s[12 + 3 * r] * bx + s[12 + 3 * r + 1] * by) + s[12 + 3 * r + 2] * bz) + s[21 + r];
#pragma unroll
            for (int r = 0; r < 3; ++r) Cv[r] = (s[3 + r] * nx + s[3 + 3 + r] * ny) + s[3 + 6 + r] * nz;  // rot^T n
            row[0] = nx;
            row[1] = ny;
            row[2] = nz;
            row[3] = pt[1] * Cv[2] - pt[2] * Cv[1];
            row[4] = pt[2] * Cv[0] - pt[0] * Cv[2];
            row[5] = pt[0] * Cv[1] - pt[1] * Cv[0];
            if (NCOL == 12) {
              double D[3];
#pragma unroll
              for (int r = 0; r < 3; ++r) D[r] = (s[12 + r] * Cv[0] + s[12 + 3 + r] * Cv[1]) + s[12 + 6 + r] * Cv[2];  // offR^T C
              row[6] = by * D[2] - bz * D[1];
              row[7] = bz * D[0] - bx * D[2];
              row[8] = bx * D[1] - by * D[0];
              row[9] = Cv[0];
              row[10] = Cv[1];
              row[11] = Cv[2];
            }
            row[NCOL] = -(double)pd2;
            row[NCOL + 1] = 1.0;
            row[NCOL + 2] = fabs((double)pd2);
          }
        }
      }
      sel[i] = valid ? 1 : 0;
      plane[i] = pl;
      if (rows_out) {
#pragma unroll
        for (int j = 0; j < 12; ++j) rows_out[(size_t)i * 13 + j] = (j < NCOL) ? row[j] : 0.0;
        rows_out[(size_t)i * 13 + 12] = row[NCOL];
      }
    }
    const unsigned vmask = __ballot_sync(0xFFFFFFFFu, valid);
    if (vmask) {
      if (valid) {
#pragma unroll
        for (int j = 0; j < NC; ++j) s_rows[warp][lane][j] = row[j];
      }
      __syncwarp();
#pragma unroll
      for (int j = 0; j < VPL; ++j) {
        const int v = lane + 32 * j;
        if (v < NV) {
          const int a = s_pa[v], b = s_pb[v];
          double sum = acc[j];
          unsigned mm = vmask;
          while (mm) {
            const int r = __ffs(mm) - 1;
            mm &= mm - 1;
            sum = fma(s_rows[warp][r][a], s_rows[warp][r][b], sum);
          }
          acc[j] = sum;
        }
      }
      __syncwarp();
    }
  }
  // block tree: warps -> block partial (fixed order)
#pragma unroll
  for (int j = 0; j < VPL; ++j) {
    const int v = lane + 32 * j;
    if (v < NV) s_blk[warp][v] = acc[j];
  }
  __syncthreads();
  if (threadIdx.x < NV) {
    double sum = 0.0;
#pragma unroll
    for (int w = 0; w < kRedWarps; ++w) sum += s_blk[w][threadIdx.x];
    partials[(size_t)blockIdx.x * kMaxVals + threadIdx.x] = sum;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t t = atomicAdd(&ctr->ticket, 1u);
    s_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // last block: sum the block partials in block order, two halves in parallel
  __shared__ double s_half[2][NV];
  const int v = threadIdx.x & 127, half = threadIdx.x >> 7;
  if (v < NV) {
    const int nb = gridDim.x, mid = (nb + 1) / 2;
    const int lo = half ? mid : 0, hi = half ? nb : mid;
    double sum = 0.0;
    for (int b = lo; b < hi; ++b) sum += __ldcg(partials + (size_t)b * kMaxVals + v);
    s_half[half][v] = sum;
  }
  __syncthreads();
  if (threadIdx.x < NV) {
    const double val = s_half[0][threadIdx.x] + s_half[1][threadIdx.x];
    const int a = s_pa[threadIdx.x], b = s_pb[threadIdx.x];
    if (a < NCOL && b < NCOL) {
      out->HTH[a * 12 + b] = val;
      out->HTH[b * 12 + a] = val;
    } else if (b == NCOL) {
      out->HTz[a] = val;
    } else if (a == NCOL + 1) {
      out->n_eff = val;
    } else {
      out->res_sum = val;
    }
  }
  if (NCOL == 6) {
    // columns 6..11 are identically zero when extrinsic_est_en is off
    for (int e = threadIdx.x; e < 144; e += blockDim.x) {
      const int a = e / 12, b = e % 12;
      if (a >= 6 || b >= 6) out->HTH[e] = 0.0;
    }
    if (threadIdx.x < 6) out->HTz[6 + threadIdx.x] = 0.0;
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    ctr->ticket = 0;
    *reinterpret_cast<volatile unsigned long long*>(&out->seq) = seq;
  }
}

// ---------------------------------------------------------------- map_incremental (device side)
// class per point: 0 skip, 1 PointToAdd, 2 PointNoNeedDownsample; block counts for the ordered compaction
__global__ void __launch_bounds__(256) k_classify(const float4* __restrict__ body, int n, StateArg st, const float4* __restrict__ nn,
                                                  const int* __restrict__ nn_cnt, int flg_ekf_inited, float fs, MapCounters* ctr,
                                                  float4* __restrict__ world, uint8_t* __restrict__ cls,
                                                  uint32_t* __restrict__ blk_cnt) {
  __shared__ uint32_t s_cnt[2];
  if (threadIdx.x < 2) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  int c = 0;
  if (i < n) {
    const float4 b = __ldg(body + i);
    float wx, wy, wz;
    body_to_world(st, b.x, b.y, b.z, wx, wy, wz);
    world[i] = make_float4(wx, wy, wz, 0.f);
    const int cnt = nn_cnt[i];
    c = 1;
    if (cnt > 0 && flg_ekf_inited) {
      const float mx = floorf(wx / fs) * fs + 0.5f * fs, my = floorf(wy / fs) * fs + 0.5f * fs,
                  mz = floorf(wz / fs) * fs + 0.5f * fs;
      const float dist = sqdist3(wx, wy, wz, mx, my, mz);
      const float4 n0 = __ldg(nn + (size_t)i * kK);
      if (fabsf(n0.x - mx) > 0.5f * fs && fabsf(n0.y - my) > 0.5f * fs && fabsf(n0.z - mz) > 0.5f * fs) {
        c = 2;
      } else if (ctr->live >= kK) {
        for (int k = 0; k < cnt; ++k) {
          const float4 q = __ldg(nn + (size_t)i * kK + k);
          if (sqdist3(q.x, q.y, q.z, mx, my, mz) < dist) {
            c = 0;
            break;
          }
        }
      }
    }
    cls[i] = (uint8_t)c;
  }
  const unsigned m1 = __ballot_sync(0xFFFFFFFFu, c == 1), m2 = __ballot_sync(0xFFFFFFFFu, c == 2);
  if ((threadIdx.x & 31) == 0) {
    if (m1) atomicAdd(&s_cnt[0], __popc(m1));
    if (m2) atomicAdd(&s_cnt[1], __popc(m2));
  }
  __syncthreads();
  if (threadIdx.x < 2) blk_cnt[2 * blockIdx.x + threadIdx.x] = s_cnt[threadIdx.x];
}

// exclusive scan of the per-block counts (single block, chunked with a running carry)
__global__ void __launch_bounds__(1024) k_scan_blocks(uint32_t* __restrict__ blk_cnt, int nblocks, MapCounters* ctr) {
  __shared__ uint32_t s_w[2][32];
  __shared__ uint32_t s_carry[2];
  if (threadIdx.x < 2) s_carry[threadIdx.x] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int base = 0; base < nblocks; base += 1024) {
    const int b = base + threadIdx.x;
    uint32_t v[2] = {0, 0};
    if (b < nblocks) {
      v[0] = blk_cnt[2 * b];
      v[1] = blk_cnt[2 * b + 1];
    }
    uint32_t inc[2];
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      uint32_t x = v[t];
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, off);
        if (lane >= off) x += y;
      }
      inc[t] = x;
      if (lane == 31) s_w[t][warp] = x;
    }
    __syncthreads();
    if (warp == 0) {
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        uint32_t x = s_w[t][lane];
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
          const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, off);
          if (lane >= off) x += y;
        }
        s_w[t][lane] = x;
      }
    }
    __syncthreads();
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      const uint32_t before = s_carry[t] + (warp ? s_w[t][warp - 1] : 0) + inc[t] - v[t];
      if (b < nblocks) blk_cnt[2 * b + t] = before;
    }
    __syncthreads();
    if (threadIdx.x < 2) s_carry[threadIdx.x] += s_w[threadIdx.x][31];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    ctr->list_add = s_carry[0];
    ctr->list_nodown = s_carry[1];
  }
}

// ordered scatter of the world points into PointToAdd / PointNoNeedDownsample
__global__ void __launch_bounds__(256) k_scatter_lists(const float4* __restrict__ world, const uint8_t* __restrict__ cls, int n,
                                                       const uint32_t* __restrict__ blk_off, float4* __restrict__ list_add,
                                                       float4* __restrict__ list_nodown) {
  __shared__ uint32_t s_w[2][8];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = (i < n) ? cls[i] : 0;
  const unsigned m1 = __ballot_sync(0xFFFFFFFFu, c == 1), m2 = __ballot_sync(0xFFFFFFFFu, c == 2);
  if (lane == 0) {
    s_w[0][warp] = __popc(m1);
    s_w[1][warp] = __popc(m2);
  }
  __syncthreads();
  if (c == 0) return;
  const int t = c - 1;
  uint32_t off = blk_off[2 * blockIdx.x + t];
  for (int w = 0; w < warp; ++w) off += s_w[t][w];
  off += __popc((t ? m2 : m1) & ((1u << lane) - 1u));
  (t ? list_nodown : list_add)[off] = world[i];
}

__global__ void k_expand_xyz(const float* __restrict__ xyz, int n, float4* __restrict__ out) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    out[i] = make_float4(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], 0.f);
}

StateArg make_state(const double* s) {
  StateArg a;
  std::memcpy(a.s, s, sizeof(a.s));
  return a;
}

int knn_grid(const lio_ctx* c, int n, int lanes) {
  const int groups_per_block = kKnnBlock / lanes;
  const int want = (n + groups_per_block - 1) / groups_per_block;
  return std::max(1, std::min(want, c->n_sms * 16));
}

}  // namespace

void launch_expand_xyz(lio_ctx* c, const float* xyz_dev, int n, float4* out) {
  if (n <= 0) return;
  const int grid = std::max(1, std::min((n + 255) / 256, c->n_sms * 8));
  k_expand_xyz<<<grid, 256, 0, c->stream>>>(xyz_dev, n, out);
  c->launches += 1;
}

int scan_alloc(lio_ctx* c) {
  const size_t n = c->cfg.max_scan;
  LIO_CUDA(cudaMalloc(&c->body, sizeof(float4) * n));
  LIO_CUDA(cudaMalloc(&c->world, sizeof(float4) * n));
  LIO_CUDA(cudaMalloc(&c->nn, sizeof(float4) * kK * n));
  LIO_CUDA(cudaMalloc(&c->nn_d2, sizeof(float) * kK * n));
  LIO_CUDA(cudaMalloc(&c->nn_cnt, sizeof(int) * n));
  LIO_CUDA(cudaMalloc(&c->sel, n));
  LIO_CUDA(cudaMalloc(&c->plane, sizeof(float4) * n));
  LIO_CUDA(cudaMalloc(&c->cls, n));
  LIO_CUDA(cudaMalloc(&c->blk_cnt, sizeof(uint32_t) * 2 * ((n + 255) / 256 + 1)));
  LIO_CUDA(cudaMalloc(&c->list_add, sizeof(float4) * n));
  LIO_CUDA(cudaMalloc(&c->list_nodown, sizeof(float4) * n));
  c->partial_blocks = c->n_sms * 2;
  LIO_CUDA(cudaMalloc(&c->partials, sizeof(double) * kMaxVals * (size_t)c->partial_blocks));
  LIO_CUDA(cudaHostAlloc(&c->out_host, sizeof(HShareOut), cudaHostAllocMapped));
  std::memset(c->out_host, 0, sizeof(HShareOut));
  LIO_CUDA(cudaHostGetDevicePointer(&c->out_dev, c->out_host, 0));
  return LIO_OK;
}

void scan_free(lio_ctx* c) {
  cudaFree(c->body);
  cudaFree(c->world);
  cudaFree(c->nn);
  cudaFree(c->nn_d2);
  cudaFree(c->nn_cnt);
  cudaFree(c->sel);
  cudaFree(c->plane);
  cudaFree(c->rows);
  cudaFree(c->cls);
  cudaFree(c->blk_cnt);
  cudaFree(c->list_add);
  cudaFree(c->list_nodown);
  cudaFree(c->partials);
  if (c->out_host) cudaFreeHost(c->out_host);
}

int scan_knn(lio_ctx* c, const double* state_host) {
  const int n = c->scan_n;
  timer_begin(c, T_KNN);
  k_scan_knn<8><<<knn_grid(c, n, 8), kKnnBlock, 0, c->stream>>>(c->map, c->body, n, make_state(state_host), c->cfg.max_sqdist,
                                                                 c->nn, c->nn_d2, c->nn_cnt);
  c->launches += 1;
  timer_end(c);
  LIO_CUDA(cudaGetLastError());
  c->nn_valid = true;
  return LIO_OK;
}

int knn_batch(lio_ctx* c, const float4* q_dev, int n, int k, float max_d2, float4* out_nn, float* out_d2, int* out_cnt) {
  timer_begin(c, T_KNN);
  k_knn_batch<8><<<knn_grid(c, n, 8), kKnnBlock, 0, c->stream>>>(c->map, q_dev, n, k, max_d2, out_nn, out_d2, out_cnt);
  c->launches += 1;
  timer_end(c);
  LIO_CUDA(cudaGetLastError());
  return LIO_OK;
}

// launches K4+K5; the result lands in c->out_host once the stream reaches the end of the kernel
int scan_plane_reduce(lio_ctx* c, const double* state_host, bool want_rows) {
  const int n = c->scan_n;
  if (want_rows && !c->rows) LIO_CUDA(cudaMalloc(&c->rows, sizeof(double) * 13 * (size_t)c->cfg.max_scan));
  const int tiles = (n + 31) / 32;
  const int grid = std::max(1, std::min((tiles + kRedWarps - 1) / kRedWarps, c->partial_blocks));
  const unsigned long long seq = ++c->out_seq;
  timer_begin(c, T_PLANE);
  if (c->cfg.extrinsic_est_en)
    k_plane_reduce<12><<<grid, kRedBlock, 0, c->stream>>>(c->body, n, make_state(state_host), c->nn, c->nn_cnt, c->world, c->sel,
                                                          c->plane, want_rows ? c->rows : nullptr, c->partials, c->map.ctr,
                                                          c->out_dev, seq);
  else
    k_plane_reduce<6><<<grid, kRedBlock, 0, c->stream>>>(c->body, n, make_state(state_host), c->nn, c->nn_cnt, c->world, c->sel,
                                                         c->plane, want_rows ? c->rows : nullptr, c->partials, c->map.ctr,
                                                         c->out_dev, seq);
  c->launches += 1;
  timer_end(c);
  LIO_CUDA(cudaGetLastError());
  return LIO_OK;
}

int scan_classify_insert(lio_ctx* c, const double* state_host, int flg_ekf_inited) {
  const int n = c->scan_n;
  if (n <= 0) return LIO_OK;
  const int blocks = (n + 255) / 256;
  timer_begin(c, T_CLASSIFY);
  k_classify<<<blocks, 256, 0, c->stream>>>(c->body, n, make_state(state_host), c->nn, c->nn_cnt, flg_ekf_inited,
                                            c->cfg.filter_size_map, c->map.ctr, c->world, c->cls, c->blk_cnt);
  k_scan_blocks<<<1, 1024, 0, c->stream>>>(c->blk_cnt, blocks, c->map.ctr);
  k_scatter_lists<<<blocks, 256, 0, c->stream>>>(c->world, c->cls, n, c->blk_cnt, c->list_add, c->list_nodown);
  c->launches += 3;
  timer_end(c);
  LIO_CUDA(cudaGetLastError());
  int rc = map_add_device(c, c->list_add, (uint32_t)n, &c->map.ctr->list_add, true);
  if (rc) return rc;
  return map_add_device(c, c->list_nodown, (uint32_t)n, &c->map.ctr->list_nodown, false);
}

}  // namespace lio

// Device helpers shared by the loop-closure kernels (loop_voxel.cu, loop_icp.cu).
// Built with --fmad=false: every f32/f64 expression is evaluated in the operation order of oracle/orc_icp.cpp.
#pragma once
#include <cstdint>
#include <cstring>
#include <cuda_runtime.h>

namespace loopb {

// order-preserving float <-> uint mapping for atomicMin / atomicMax on bounding boxes
__device__ __forceinline__ uint32_t f2ord(float f) {
  const uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t u) {
  const uint32_t b = (u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u;
  return __uint_as_float(b);
}

// per-cloud reduction cells: 6 ordered-uint box bounds + finite count
struct BoxAcc {
  uint32_t lo[3], hi[3], n_finite, pad;
};

// exact, order-independent accumulation: 64-bit fixed point (the oracle's fixc / fixp)
constexpr double kFixCoord = 16777216.0;  // 2^24
constexpr double kFixProd = 1048576.0;    // 2^20
__device__ __forceinline__ long long fixc(float v) { return __double2ll_rn((double)v * kFixCoord); }
__device__ __forceinline__ long long fixp(double v) { return __double2ll_rn(v * kFixProd); }

__device__ __forceinline__ long long warp_sum_ll(long long v) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, off);
  return v;
}

// block-wide sums of NV int64 values; the result is valid in thread 0..NV-1 of warp 0 (value t in thread t)
template <int NV, int WARPS>
__device__ __forceinline__ long long block_sum_ll(long long (&v)[NV], long long (*smem)[NV]) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    const long long s = warp_sum_ll(v[k]);
    if (lane == 0) smem[warp][k] = s;
  }
  __syncthreads();
  long long out = 0;
  if (threadIdx.x < NV) {
#pragma unroll
    for (int w = 0; w < WARPS; ++w) out += smem[w][threadIdx.x];
  }
  return out;
}

__device__ __forceinline__ void atomic_add_ll(long long* p, long long v) {
  atomicAdd(reinterpret_cast<unsigned long long*>(p), (unsigned long long)v);
}

__device__ __forceinline__ float sqd3(float qx, float qy, float qz, float bx, float by, float bz) {
  const float dx = qx - bx, dy = qy - by, dz = qz - bz;
  const float s = dx * dx + dy * dy;
  return s + dz * dz;
}

// Bounding-box / finite-count accumulation of one point per thread into acc[job] (job < 0: thread has no point).
// Must be called by every thread of a block of <= 256 threads (contains a barrier).
__device__ __forceinline__ void block_box_update(BoxAcc* __restrict__ acc, int job, bool finite, float ox, float oy, float oz) {
  const float4 o = make_float4(ox, oy, oz, 0.f);
  // box update: warp shuffle reduction, then one thread merges the block's eight warps so a block costs one set of
  // atomics per job it touches (per-warp atomics on ~7 addresses per job were serialising in L2: 2.6 M warps x 7)
  __shared__ uint32_t s_lo[8][3], s_hi[8][3], s_nf[8];
  __shared__ int s_job[8];
  const unsigned full = 0xFFFFFFFFu;
  const int warp = threadIdx.x >> 5;
  const int job0 = __shfl_sync(full, job, 0);
  const bool uniform = __all_sync(full, job == job0 || job < 0) && job0 >= 0;
  if (uniform) {
    uint32_t lo[3] = {finite ? f2ord(o.x) : 0xFFFFFFFFu, finite ? f2ord(o.y) : 0xFFFFFFFFu, finite ? f2ord(o.z) : 0xFFFFFFFFu};
    uint32_t hi[3] = {finite ? f2ord(o.x) : 0u, finite ? f2ord(o.y) : 0u, finite ? f2ord(o.z) : 0u};
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      lo[a] = __reduce_min_sync(full, lo[a]);
      hi[a] = __reduce_max_sync(full, hi[a]);
    }
    const unsigned nf = __popc(__ballot_sync(full, finite));
    if ((threadIdx.x & 31) == 0) {
      s_job[warp] = nf ? job0 : -1;
      s_nf[warp] = nf;
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        s_lo[warp][a] = lo[a];
        s_hi[warp][a] = hi[a];
      }
    }
  } else {
    if ((threadIdx.x & 31) == 0) s_job[warp] = -1;
    if (finite && job >= 0) {
      atomicMin(&acc[job].lo[0], f2ord(o.x));
      atomicMin(&acc[job].lo[1], f2ord(o.y));
      atomicMin(&acc[job].lo[2], f2ord(o.z));
      atomicMax(&acc[job].hi[0], f2ord(o.x));
      atomicMax(&acc[job].hi[1], f2ord(o.y));
      atomicMax(&acc[job].hi[2], f2ord(o.z));
      atomicAdd(&acc[job].n_finite, 1u);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int cur = -1;
    uint32_t lo[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, hi[3] = {0u, 0u, 0u}, nf = 0;
    for (int w = 0; w <= 8; ++w) {
      const int j = (w < 8 && w < (int)(blockDim.x >> 5)) ? s_job[w] : -2;
      if (j != cur && cur >= 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
          atomicMin(&acc[cur].lo[a], lo[a]);
          atomicMax(&acc[cur].hi[a], hi[a]);
          lo[a] = 0xFFFFFFFFu;
          hi[a] = 0u;
        }
        atomicAdd(&acc[cur].n_finite, nf);
        nf = 0;
      }
      if (j >= 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
          lo[a] = min(lo[a], s_lo[w][a]);
          hi[a] = max(hi[a], s_hi[w][a]);
        }
        nf += s_nf[w];
      }
      if (j != -1) cur = j;  // a warp without finite points (-1) does not end a run
    }
  }
}

}  // namespace loopb

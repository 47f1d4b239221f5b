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

}  // namespace loopb

// Measurement hook: achievable random-gather bandwidth of this GPU ("a micro-kernel doing random
// 16-byte float4 loads from a 320 MB table") — the ceiling the HBM-bound kNN regime (config 5) is compared with.
#include <algorithm>
#include "lio_ctx.h"

namespace lio {
namespace {

__device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16;
  x *= 0x7feb352du;
  x ^= x >> 15;
  x *= 0x846ca68bu;
  x ^= x >> 16;
  return x;
}

// every thread performs `per_thread` independent random float4 loads (4 in flight), result folded into a checksum
__global__ void __launch_bounds__(256) k_gather_probe(const float4* __restrict__ table, uint32_t mask, int per_thread, float* __restrict__ sink) {
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
  float acc = 0.f;
  uint32_t s = mix32(tid * 2654435761u + 12345u);
  for (int i = 0; i < per_thread; i += 4) {
    const uint32_t a = mix32(s), b = mix32(s + 1), c = mix32(s + 2), d = mix32(s + 3);
    s += 4;
    const float4 pa = __ldg(table + (a & mask)), pb = __ldg(table + (b & mask)), pc = __ldg(table + (c & mask)), pd = __ldg(table + (d & mask));
    acc += (pa.x + pb.y) + (pc.z + pd.w);
  }
  if (acc == 123.456f) sink[0] = acc;  // keeps the loads alive
}

__global__ void k_fill(float4* t, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    t[i] = make_float4((float)i, 1.f, 2.f, 3.f);
}

}  // namespace
}  // namespace lio

extern "C" int lio_gather_probe(lio_ctx* c, uint64_t table_bytes, int reps, double* gbps_16B, double* gbps_sector) {
  using namespace lio;
  if (!c || table_bytes < (1u << 20) || reps < 1) {
    set_error("lio_gather_probe: bad arguments");
    return LIO_E_ARG;
  }
  uint64_t n = 1;
  while (n * 2 * sizeof(float4) <= table_bytes) n *= 2;  // power of two entries
  float4* table = nullptr;
  float* sink = nullptr;
  LIO_CUDA(cudaMalloc(&table, n * sizeof(float4)));
  LIO_CUDA(cudaMalloc(&sink, sizeof(float)));
  k_fill<<<c->n_sms * 8, 256, 0, c->stream>>>(table, n);
  const int per_thread = 64, grid = c->n_sms * 32;
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  k_gather_probe<<<grid, 256, 0, c->stream>>>(table, (uint32_t)(n - 1), per_thread, sink);  // warm-up
  cudaEventRecord(a, c->stream);
  for (int r = 0; r < reps; ++r) k_gather_probe<<<grid, 256, 0, c->stream>>>(table, (uint32_t)(n - 1), per_thread, sink);
  cudaEventRecord(b, c->stream);
  cudaError_t e = cudaStreamSynchronize(c->stream);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, a, b);
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  cudaFree(table);
  cudaFree(sink);
  c->launches += reps + 2;
  if (!cuda_ok(e, "gather probe")) return LIO_E_CUDA;
  const double loads = (double)grid * 256.0 * per_thread * reps;
  if (gbps_16B) *gbps_16B = loads * 16.0 / (ms * 1e-3) / 1e9;
  if (gbps_sector) *gbps_sector = loads * 32.0 / (ms * 1e-3) / 1e9;  // a 16-byte request drags a 32-byte sector
  return LIO_OK;
}

// lio_hth_reduce_static: the FP64 H^T H / H^T z reduction of the measurement model for callers that may not
// allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  Same sum as the tail of
// k_plane_reduce (lio_scan.cu), which keeps its per-block partial sums in a cudaMalloc'ed buffer of the context;
// here the partial sums live in a file-scope __device__ array, so the call allocates nothing.
//
// No counterpart in the reference.
//
// The array exists once per GPU that loads the library and stays until the process exits.  Every call on a
// device shares it and finds the last call's contents in it: the ticket word is cleared on the caller's stream
// before each launch (and again by the block that finishes the sum), the partial sums are fully rewritten for
// the blocks of the call before anything reads them.  Two calls that overlap on one GPU (different streams,
// or different host threads) race on it.
#include <mutex>
#include "lio_ctx.h"

namespace lio {
namespace {

constexpr int kSrThreads = 128;         // rows staged per pass
constexpr int kSrRowsPerBlock = 1024;   // rows one block sums
constexpr int kSrCols = 12;             // h_x columns
constexpr int kSrVals = 90;             // 78 upper-triangle products + 12 of H^T z
constexpr int kSrStride = 96;           // doubles per block in the scratch
constexpr int kSrMaxBlocks = LIO_STATIC_REDUCE_MAX_ROWS / kSrRowsPerBlock;
constexpr size_t kSrHeadBytes = 256;    // ticket word, alone in its 256 bytes
constexpr size_t kSrScratchBytes = kSrHeadBytes + sizeof(double) * kSrStride * (size_t)kSrMaxBlocks;
constexpr int kSrMaxDevices = 64;

static_assert(LIO_STATIC_REDUCE_MAX_ROWS % kSrRowsPerBlock == 0, "whole blocks");
static_assert(kSrScratchBytes <= (1u << 20), "the static scratch is meant to stay modest");

// 768.25 KiB on every GPU that loads the library.  In HOST code this name is only a handle for
// cudaGetSymbolAddress, never a device address.
static __device__ __align__(256) unsigned char g_static_scratch[kSrScratchBytes];

std::mutex g_sr_mu;
void* g_sr_ptr[kSrMaxDevices] = {};

// Device address of the array on the CURRENT device, looked up once per device.
int static_scratch_of_current_device(unsigned char** out) {
  int dev = -1;
  LIO_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kSrMaxDevices) {
    set_error("lio_hth_reduce_static: device ordinal %d out of range", dev);
    return LIO_E_ARG;
  }
  std::lock_guard<std::mutex> lock(g_sr_mu);
  if (!g_sr_ptr[dev]) {
    void* p = nullptr;
    LIO_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
    if (!p) {
      set_error("lio_hth_reduce_static: cudaGetSymbolAddress returned a null address on device %d", dev);
      return LIO_E_CUDA;
    }
    g_sr_ptr[dev] = p;
  }
  *out = (unsigned char*)g_sr_ptr[dev];
  return LIO_OK;
}

// Block b sums rows [1024 b, 1024 (b+1)): 128 rows at a time are staged in shared memory (coalesced), then
// thread v < 90 adds its product over the staged rows in row order.  The block that draws the last ticket adds
// the per-block sums in block order, so the result does not depend on scheduling.
__global__ void __launch_bounds__(kSrThreads) k_hth_reduce_static(const double* __restrict__ rows, const double* __restrict__ z, int n,
                                                                  unsigned char* scratch, double* __restrict__ HTH,
                                                                  double* __restrict__ HTz) {
  __shared__ double sh[kSrThreads * kSrCols];
  __shared__ double shz[kSrThreads];
  __shared__ int is_last;
  unsigned int* ticket = (unsigned int*)scratch;
  double* partials = (double*)(scratch + kSrHeadBytes);
  const int v = threadIdx.x;
  int vi = 0, vj = 0;  // v < 78: the v-th (i <= j) pair, row by row; 78 <= v < 90: column v - 78 against z
  if (v < 78) {
    int left = v;
    while (left >= kSrCols - vi) {
      left -= kSrCols - vi;
      ++vi;
    }
    vj = vi + left;
  } else if (v < kSrVals) {
    vi = v - 78;
  }
  const bool with_z = v >= 78;
  double sum = 0.0;
  const long long base = (long long)blockIdx.x * kSrRowsPerBlock;
  for (int s = 0; s < kSrRowsPerBlock; s += kSrThreads) {
    const long long r0 = base + s;
    if (r0 >= n) break;  // same for the whole block
    const int live = (int)min((long long)kSrThreads, (long long)n - r0);
    for (int e = threadIdx.x; e < kSrThreads * kSrCols; e += kSrThreads)
      sh[e] = (e < live * kSrCols) ? rows[r0 * kSrCols + e] : 0.0;
    shz[threadIdx.x] = (threadIdx.x < live) ? z[r0 + threadIdx.x] : 0.0;
    __syncthreads();
    if (v < kSrVals) {
      for (int r = 0; r < kSrThreads; ++r) {
        const double a = sh[r * kSrCols + vi];
        const double b = with_z ? shz[r] : sh[r * kSrCols + vj];
        sum += a * b;
      }
    }
    __syncthreads();
  }
  if (v < kSrVals) partials[(size_t)blockIdx.x * kSrStride + v] = sum;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1) ? 1 : 0;
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  if (v < kSrVals) {
    double total = 0.0;
    for (unsigned int b = 0; b < gridDim.x; ++b) total += __ldcg(partials + (size_t)b * kSrStride + v);
    if (with_z) {
      HTz[vi] = total;
    } else {
      HTH[vi * kSrCols + vj] = total;
      HTH[vj * kSrCols + vi] = total;
    }
  }
  if (threadIdx.x == 0) *ticket = 0u;  // leave the word cleared for the next call
}

}  // namespace
}  // namespace lio

extern "C" int lio_hth_reduce_static(const double* rows_dev, const double* z_dev, int n, double* HTH_dev, double* HTz_dev,
                                     void* stream) {
  using namespace lio;
  if (n < 0) {
    set_error("lio_hth_reduce_static: n = %d is negative", n);
    return LIO_E_ARG;
  }
  if (n > LIO_STATIC_REDUCE_MAX_ROWS) {
    set_error("lio_hth_reduce_static: n = %d rows exceeds LIO_STATIC_REDUCE_MAX_ROWS = %d (the static scratch holds one partial "
              "sum per %d rows); split the rows and add the results, or use lio_h_share",
              n, LIO_STATIC_REDUCE_MAX_ROWS, kSrRowsPerBlock);
    return LIO_E_ARG;
  }
  if (!HTH_dev || !HTz_dev || (n > 0 && (!rows_dev || !z_dev))) {
    set_error("lio_hth_reduce_static: null pointer");
    return LIO_E_ARG;
  }
  unsigned char* scratch = nullptr;
  const int rc = static_scratch_of_current_device(&scratch);
  if (rc != LIO_OK) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  const int grid = n > 0 ? (n + kSrRowsPerBlock - 1) / kSrRowsPerBlock : 1;
  LIO_CUDA(cudaMemsetAsync(scratch, 0, kSrHeadBytes, s));  // the ticket word must start at zero
  k_hth_reduce_static<<<grid, kSrThreads, 0, s>>>(rows_dev, z_dev, n, scratch, HTH_dev, HTz_dev);
  LIO_CUDA(cudaGetLastError());
  return LIO_OK;
}

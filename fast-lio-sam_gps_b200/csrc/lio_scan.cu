// Measurement-model kernels: K1 body->world, K3 kNN (lio_knn.cuh), K4 plane/residual/Jacobian
// (lio_plane.cuh), K5 FP64 reduction to H^T H / H^T z, and the device side of map_incremental.
// Reference: h_share_model / map_incremental in third_party/FAST_LIO/src/laserMapping.cpp
// (absent submodule; behaviour per SURVEY.md §9.1, §9.3; consumer per §9.6).
#include <algorithm>
#include <cstddef>
#include <cstdlib>
#include <cstring>
#include "lio_ctx.h"
#include "lio_knn.cuh"
#include "lio_knn_tile.cuh"
#include "lio_plane.cuh"
#include "lio_iekf.cuh"

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
                                                        int* __restrict__ nn_cnt, double* __restrict__ knn_state) {
  if (knn_state && blockIdx.x == 0 && threadIdx.x < 36) knn_state[threadIdx.x] = st.s[threadIdx.x];
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

// ---------------------------------------------------------------- K3, tile form (lio_knn_tile.cuh)
// Scan ordering: the queries are put in curve order (tile_key: 24-bit Hilbert index of the query's map cell) by a two-level
// counting sort written for this size (131 k keys that fall in a few thousand distinct cells) - three short, fully parallel
// kernels instead of the library's radix passes (round 1: cub::DeviceRadixSort, 56 us per scan = 14 % of the step; its
// chained-scan passes run on 17 blocks at this size):
//   k_sort_count   key per query; level-1 histogram over the TOP 12 key bits (a 16-cell cube), level-2 histogram over the LOW
//                  12 bits inside the cube's slot (slots are handed to cubes on first touch); warp-aggregated atomics;
//   k_sort_scan    block 0: exclusive scan of the 4096 level-1 bins -> start[]; block 1+s: exclusive scan of slot s;
//   k_sort_place   order[start[cube] + offset[slot][low]++] = query (the same warp aggregation);
//   k_sort_clear   re-zeroes what this sort touched (off the critical path of the search that follows? no: 2 us, kept simple).
// The order inside one cell is whatever the atomics produce: it only decides which queries share a warp, every result is
// written to the query's own row, so nothing downstream depends on it.
struct KeysArgs {
  double s[36];
  int n;
  int pad;
};
constexpr int kSortBins = 4096;  // 12 bits per level
constexpr uint32_t kSortLowMask = kSortBins - 1;
constexpr uint32_t kSlotPending = 0xFFFFFFFFu;
struct SortScratch {
  uint32_t hist1[kSortBins];      // level 1: queries per cube
  uint32_t start[kSortBins + 1];  // exclusive scan of hist1
  uint32_t slot_of[kSortBins];    // cube -> slot + 1 (0: none yet)
  uint32_t n_slots;
  uint32_t pad[2];  // [0]: ticket of k_sort_clear; sized so that the counters behind the header are 16-byte aligned
  // followed by hist2[kSortBins][kSortBins] (slot-major), allocated behind this header
};
static_assert(sizeof(SortScratch) % 16 == 0, "level-2 counters must start 16-byte aligned (k_sort_clear stores uint4)");
__device__ __forceinline__ uint32_t* sort_hist2(SortScratch* sc) { return reinterpret_cast<uint32_t*>(sc + 1); }

// a != nullptr: graph form, (state, n) come from a small device block refreshed by a memcpy node, so one captured graph
// serves every scan of up to n_grid points.
template <bool XFORM>
__global__ void __launch_bounds__(256) k_sort_count(MapView m, const float4* __restrict__ pts, const KeysArgs* __restrict__ a, StateArg st_val, int n_val,
                                                    uint32_t* __restrict__ keys, SortScratch* __restrict__ sc) {
  __shared__ StateArg s_st;
  if (XFORM && threadIdx.x < 36) s_st.s[threadIdx.x] = a ? a->s[threadIdx.x] : st_val.s[threadIdx.x];
  __syncthreads();
  const int n = a ? a->n : n_val;
  const int lane = threadIdx.x & 31;
  uint32_t* hist2 = sort_hist2(sc);
  for (int i0 = (blockIdx.x * blockDim.x + threadIdx.x) - lane; i0 < n; i0 += gridDim.x * blockDim.x) {
    const int i = i0 + lane;
    const unsigned act = __ballot_sync(0xFFFFFFFFu, i < n);
    if (i >= n) continue;
    const float4 b = __ldg(pts + i);
    float qx = b.x, qy = b.y, qz = b.z;
    if (XFORM) body_to_world(s_st, b.x, b.y, b.z, qx, qy, qz);
    const uint32_t key = tile_key(m, qx, qy, qz);
    keys[i] = key;
    const uint32_t cube = key >> 12;
    // one atomic per distinct cube / cell of the warp (consecutive raw points mostly share their cell)
    const unsigned pc = __match_any_sync(act, cube);
    const int lc = __ffs(pc) - 1;
    uint32_t slot = 0;
    if (lane == lc) {
      atomicAdd(&sc->hist1[cube], (uint32_t)__popc(pc));
      slot = *reinterpret_cast<volatile uint32_t*>(&sc->slot_of[cube]);
      if (slot == 0) {
        const uint32_t old = atomicCAS(&sc->slot_of[cube], 0u, kSlotPending);
        if (old == 0) {
          slot = atomicAdd(&sc->n_slots, 1u) + 1u;
          atomicExch(&sc->slot_of[cube], slot);
        } else {
          slot = old;
        }
      }
      while (slot == kSlotPending) slot = *reinterpret_cast<volatile uint32_t*>(&sc->slot_of[cube]);  // another warp is assigning it
    }
    slot = __shfl_sync(pc, slot, lc);
    const unsigned pk = __match_any_sync(act, key);
    if (lane == __ffs(pk) - 1) atomicAdd(&hist2[(size_t)(slot - 1) * kSortBins + (key & kSortLowMask)], (uint32_t)__popc(pk));
  }
}

// exclusive scan of 4096 counters by one block of 512 threads (8 consecutive counters per thread)
constexpr int kSortScanBlock = 512;
__device__ __forceinline__ uint32_t block_scan_4096(uint32_t* __restrict__ cnt, uint32_t* s_scan) {
  constexpr int PER = kSortBins / kSortScanBlock;
  uint32_t v[PER], sum = 0;
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    v[j] = cnt[threadIdx.x * PER + j];  // (generic pointer: global counters of an earlier launch, or the level-1 copy in shared memory)
    sum += v[j];
  }
  // warp scan, then scan of the 16 warp totals
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t inc = sum;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, off);
    if (lane >= off) inc += t;
  }
  if (lane == 31) s_scan[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = lane < kSortScanBlock / 32 ? s_scan[lane] : 0u;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, w, off);
      if (lane >= off) w += t;
    }
    s_scan[lane] = w;  // inclusive warp totals
  }
  __syncthreads();
  uint32_t run = (warp ? s_scan[warp - 1] : 0u) + inc - sum;
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    cnt[threadIdx.x * PER + j] = run;
    run += v[j];
  }
  return s_scan[kSortScanBlock / 32 - 1];  // total
}

__global__ void __launch_bounds__(kSortScanBlock) k_sort_scan(SortScratch* __restrict__ sc) {
  __shared__ uint32_t s_scan[32];
  if (blockIdx.x == 0) {  // level 1 -> start[] (hist1 itself is left zeroed for the next sort)
    __shared__ uint32_t s_tmp[kSortBins];
    for (int e = threadIdx.x; e < kSortBins; e += kSortScanBlock) s_tmp[e] = __ldcg(&sc->hist1[e]);
    __syncthreads();
    const uint32_t total = block_scan_4096(s_tmp, s_scan);
    __syncthreads();
    for (int e = threadIdx.x; e < kSortBins; e += kSortScanBlock) {
      sc->start[e] = s_tmp[e];
      sc->hist1[e] = 0;
    }
    if (threadIdx.x == 0) sc->start[kSortBins] = total;
    return;
  }
  const uint32_t n_slots = sc->n_slots;
  for (uint32_t slot = blockIdx.x - 1; slot < n_slots; slot += gridDim.x - 1) {
    block_scan_4096(sort_hist2(sc) + (size_t)slot * kSortBins, s_scan);
    __syncthreads();  // s_scan is reused by the next slot
  }
}

__global__ void __launch_bounds__(256) k_sort_place(const uint32_t* __restrict__ keys, const KeysArgs* __restrict__ a, int n_val,
                                                    SortScratch* __restrict__ sc, uint32_t* __restrict__ order,
                                                    const float4* __restrict__ pts, float4* __restrict__ sorted) {
  const int n = a ? a->n : n_val;
  const int lane = threadIdx.x & 31;
  uint32_t* hist2 = sort_hist2(sc);
  for (int i0 = (blockIdx.x * blockDim.x + threadIdx.x) - lane; i0 < n; i0 += gridDim.x * blockDim.x) {
    const int i = i0 + lane;
    const unsigned act = __ballot_sync(0xFFFFFFFFu, i < n);
    if (i >= n) continue;
    const uint32_t key = keys[i];
    const uint32_t cube = key >> 12;
    const unsigned pk = __match_any_sync(act, key);
    const int leader = __ffs(pk) - 1;
    uint32_t base = 0;
    if (lane == leader) {
      const uint32_t slot = sc->slot_of[cube] - 1u;
      base = sc->start[cube] + atomicAdd(&hist2[(size_t)slot * kSortBins + (key & kSortLowMask)], (uint32_t)__popc(pk));
    }
    base = __shfl_sync(pk, base, leader);
    const uint32_t pos = base + __popc(pk & ((1u << lane) - 1u));
    order[pos] = (uint32_t)i;
    sorted[pos] = __ldg(pts + i);  // the search reads its 32 queries with one coalesced load instead of a gather through order[]
  }
}

// zero what the sort touched: slot_of[] and the level-2 counters of the slots in use (hist1 was zeroed by k_sort_scan)
__global__ void __launch_bounds__(256) k_sort_clear(SortScratch* __restrict__ sc) {
  const uint32_t n_slots = sc->n_slots;
  uint4* h = reinterpret_cast<uint4*>(sort_hist2(sc));
  const size_t total = (size_t)n_slots * (kSortBins / 4);
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) h[e] = make_uint4(0u, 0u, 0u, 0u);
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < kSortBins; e += gridDim.x * blockDim.x) sc->slot_of[e] = 0;
  __syncthreads();
  // n_slots is reset by the LAST block (every block has read it by then)
  __shared__ bool s_last;
  if (threadIdx.x == 0) s_last = atomicAdd(&sc->pad[0], 1u) == gridDim.x - 1;
  __syncthreads();
  if (s_last && threadIdx.x == 0) {
    sc->n_slots = 0;
    sc->pad[0] = 0;
  }
}

constexpr int kTileWarps = 4;
#ifndef LIO_TILE_BLOCKS_PER_SM
#define LIO_TILE_BLOCKS_PER_SM 8
#endif
constexpr int kTileBlocksPerSm = LIO_TILE_BLOCKS_PER_SM;  // resident blocks of k_knn_tile per SM (register budget 65536 / (128 * this))

// DEV kernels belong to the device-resident IEKF loop (lio_iekf.cuh): the state comes from ctl->x instead of the launch
// parameters.  Returns the state the block should use (block-wide; contains a barrier when DEV).
template <bool DEV>
__device__ __forceinline__ const StateArg& block_state(const StateArg& st_param, const IekfCtl* ctl, StateArg& s_st) {
  if (DEV) {
    if (threadIdx.x < 36) s_st.s[threadIdx.x] = ctl->x[threadIdx.x];
    __syncthreads();
    return s_st;
  }
  return st_param;
}

// Persistent warps draw tasks from the search's work queue (lio_knn_tile.cuh): a tile of 32 curve-ordered queries, or a
// (tile, lane mask) sub-tile another warp pushed.  `sorted` holds the query points in curve order (written by k_sort_place), so
// a task reads its 32 queries with one coalesced load; results are scattered back to the queries' own rows through `order`.
// q = this search's counter set (zeroed by the previous search's k_knn_left).
template <bool XFORM, bool DEV>
__global__ void __launch_bounds__(kTileWarps * 32, kTileBlocksPerSm) k_knn_tile(MapView m, const float4* __restrict__ sorted, const uint32_t* __restrict__ order,
                                                               int n, StateArg st_param, int k, float max_d2, float4* __restrict__ nn,
                                                               float* __restrict__ nn_d2, int* __restrict__ nn_cnt, TileQueue* __restrict__ q,
                                                               unsigned long long* __restrict__ queue, uint32_t qcap,
                                                               uint32_t* __restrict__ left, const IekfCtl* __restrict__ ctl,
                                                               uint32_t* __restrict__ stats, uint4* __restrict__ prof_out, uint32_t prof_cap,
                                                               double* __restrict__ knn_state) {
  __shared__ TileSmem smem[kTileWarps];
  __shared__ StateArg s_st;
  if (DEV && (ctl->finished || !ctl->converge)) return;  // cached neighbours are reused (ekfom_data.converge == false)
  const StateArg& st = block_state<DEV>(st_param, ctl, s_st);
  if (XFORM && knn_state && blockIdx.x == 0 && threadIdx.x < 36) knn_state[threadIdx.x] = st.s[threadIdx.x];  // for k_far_nn (map_incremental)
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned lt = (1u << lane) - 1u;
  TileSmem& sm = smem[warp];
  const uint32_t tiles = (uint32_t)((n + 31) / 32);
  const float4 pad = make_float4(0.f, 0.f, 0.f, __uint_as_float(kEmptyId));
  const float inf = __int_as_float(0x7f800000);
  uint32_t ticket, tile;
  unsigned mask;
  while (tile_next_task(q, queue, qcap, tiles, lane, ticket, tile, mask)) {
    const int t = (int)tile * 32 + lane;
    const bool valid = t < n && ((mask >> lane) & 1u);
    uint32_t i = 0u;
    float qx = 0.f, qy = 0.f, qz = 0.f;
    if (valid) {
      i = __ldg(order + t);
      const float4 b = __ldg(sorted + t);
      qx = b.x, qy = b.y, qz = b.z;
      if (XFORM) body_to_world(st, b.x, b.y, b.z, qx, qy, qz);
    }
    TileProf prof{0, 0, 0, 0};
    const long long t0 = prof_out ? clock64() : 0;
    unsigned long long g0 = 0;
    if (prof_out) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g0));
    const int cx = vox_of(qx, m.ds) >> m.cell_shift, cy = vox_of(qy, m.ds) >> m.cell_shift, cz = vox_of(qz, m.ds) >> m.cell_shift;
    // the sub-tile of the first query stays here; the others are pushed right away, before any memory is touched
    unsigned pending = __ballot_sync(0xFFFFFFFFu, valid);
    unsigned sub = 0, to_left = 0;
    int n_push = 0;
    unsigned my_push = 0;
    {
      unsigned r = pending;
      while (r) {
        const int a = __ffs(r) - 1;
        const int ax0 = __shfl_sync(0xFFFFFFFFu, cx, a), ay0 = __shfl_sync(0xFFFFFFFFu, cy, a), az0 = __shfl_sync(0xFFFFFFFFu, cz, a);
        const bool near = ((r >> lane) & 1u) && (c_tile_block ? ((cx >> 1) == (ax0 >> 1) && (cy >> 1) == (ay0 >> 1) && (cz >> 1) == (az0 >> 1))
                                                              : (abs(cx - ax0) <= kTileSpan && abs(cy - ay0) <= kTileSpan && abs(cz - az0) <= kTileSpan));
        const unsigned s2 = __ballot_sync(0xFFFFFFFFu, near);
        r &= ~s2;
        if (!sub) {
          sub = s2;
        } else {
          if (lane == n_push) my_push = s2;
          ++n_push;
        }
      }
    }
    // the other sub-tiles are offered on the queue before this warp touches the map: free warps start on them right away
    uint32_t push_base = 0;
    bool offered = false;
    if (ticket < tiles) {
      if (n_push) {
        if (lane == 0) push_base = atomicAdd(&q->tail, (uint32_t)n_push);
        push_base = __shfl_sync(0xFFFFFFFFu, push_base, 0);
        offered = c_tile_offer && push_base + (uint32_t)n_push <= qcap;  // (always, with qcap = 2 n)
        if (offered && lane < n_push) atomicExch(queue + push_base + lane, ((unsigned long long)tile << 32) | my_push);
        __syncwarp();
      }
      if (lane == 0) atomicAdd(&q->split_done, 1u);
    }
    auto emit = [&](const uint32_t (&rs)[kK], int cnt) {
      cnt = min(cnt, k);
#pragma unroll
      for (int j = 0; j < kK; ++j)
        if (j < k) {
          const bool have = j < cnt;
          const float4 c = have ? sm.pts[rs[j]] : pad;
          nn[(size_t)i * k + j] = c;
          if (nn_d2) nn_d2[(size_t)i * k + j] = have ? sqdist3(qx, qy, qz, c.x, c.y, c.z) : inf;
        }
      nn_cnt[i] = cnt;
    };
    // Sets of lanes (coherent sub-tiles) are served one after the other through ONE call site of the search: the own
    // sub-tile first, then (stage 1) whatever of the offers nobody took, taken back with a compare-and-swap.  When the
    // candidates of a set overflow the staging area (a dense corner: the box of a sub-tile can reach 5 x 5 x 5 cells) its
    // lanes are served cell by cell - the queries of one cell stage 27 cells; if even that overflows they go to the per-query kernel.
    unsigned set = sub, back = 0u, owners = 0u;
    if (ticket >= tiles) set = pending;  // (an offered mask is coherent by construction: sub == pending; kept for robustness)
    int stage = 0;
    bool by_cell = false;
    while (true) {
      if (!set) {
        if (stage == 0) {
          stage = 1;
          if (ticket < tiles && n_push) {
            back = offered ? 0u : my_push;
            if (offered && lane < n_push) {
              const unsigned long long val = ((unsigned long long)tile << 32) | my_push;
              if (atomicCAS(queue + push_base + lane, val, 0ull) == val) back = my_push;
            }
            owners = __ballot_sync(0xFFFFFFFFu, lane < n_push && back != 0u);
          }
        }
        if (!owners) break;
        const int c = __ffs(owners) - 1;
        owners &= owners - 1;
        set = __shfl_sync(0xFFFFFFFFu, back, c);
        by_cell = false;
      }
      unsigned part = set;
      if (by_cell) {  // the lanes of the set that share the cell of its first lane
        const int a = __ffs(set) - 1;
        const int ax0 = __shfl_sync(0xFFFFFFFFu, cx, a), ay0 = __shfl_sync(0xFFFFFFFFu, cy, a), az0 = __shfl_sync(0xFFFFFFFFu, cz, a);
        part = __ballot_sync(0xFFFFFFFFu, ((set >> lane) & 1u) && cx == ax0 && cy == ay0 && cz == az0);
      }
      const unsigned failed = tile_search_sub(m, sm, lane, qx, qy, qz, cx, cy, cz, part, max_d2, emit, prof_out ? &prof : nullptr);
      set &= ~(part & ~failed);
      if (failed) {
        if (by_cell || !c_tile_halve) {
          to_left |= failed;
          set &= ~failed;
        } else {
          by_cell = true;
        }
      }
    }
    if (to_left) {
      if (lane == 0 && stats) atomicAdd(stats + 1, (uint32_t)__popc(to_left));
      uint32_t base = 0;
      if (lane == 0) base = atomicAdd(&q->left, (uint32_t)__popc(to_left));
      base = __shfl_sync(0xFFFFFFFFu, base, 0);
      if ((to_left >> lane) & 1u) left[base + __popc(to_left & lt)] = i;
    }
    if (lane == 0 && stats && ticket >= tiles) atomicAdd(stats + 2, (uint32_t)__popc(pending));  // queries served through a pushed task
    if (prof_out && lane == 0 && ticket < prof_cap) {
      unsigned long long g1;
      uint32_t smid;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g1));
      asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
      prof_out[2 * ticket] = make_uint4((uint32_t)(clock64() - t0), (uint32_t)prof.passes | ((uint32_t)__popc(pending & ~to_left) << 8) | ((uint32_t)n_push << 16) |
                                        ((ticket >= tiles ? 1u : 0u) << 24), (uint32_t)prof.staged, (uint32_t)prof.cells);
      prof_out[2 * ticket + 1] = make_uint4((uint32_t)g0, (uint32_t)g1, smid, (uint32_t)prof.stage_cycles);
    }
  }
}

// 8-lane-group form of the leftover search (four queries per warp): used for generic query batches (lio_map_knn), whose
// deferred queries are mostly shallow (one ring) - there four queries per warp beat one; the scan path uses the warp form.
template <bool XFORM, bool DEV>
__global__ void __launch_bounds__(kKnnBlock) k_knn_left8(MapView m, const float4* __restrict__ pts, const uint32_t* __restrict__ left,
                                                        StateArg st_param, int k, float max_d2, float4* __restrict__ nn,
                                                        float* __restrict__ nn_d2, int* __restrict__ nn_cnt, TileQueue* __restrict__ ctr,
                                                        TileQueue* __restrict__ ctr_next, uint32_t* __restrict__ stats,
                                                        const IekfCtl* __restrict__ ctl) {
  __shared__ StateArg s_st;
  const Group<8> g;
  const int groups_per_block = kKnnBlock / 8;
  const int gid = blockIdx.x * groups_per_block + (threadIdx.x / 8);
  const int gstride = gridDim.x * groups_per_block;
  const bool skip = DEV && (ctl->finished || !ctl->converge);  // the tile kernel did not run: ctr is still zero
  const int n_left = skip ? 0 : (int)ctr->left;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    *ctr_next = TileQueue{0u, 0u, 0u, 0u};  // the next search uses the other counter set
    if (stats && !skip) atomicAdd(stats, (uint32_t)n_left);
  }
  if (skip) return;
  const StateArg& st = block_state<DEV>(st_param, ctl, s_st);
  const float4 pad = make_float4(0.f, 0.f, 0.f, __uint_as_float(kEmptyId));
  const float inf = __int_as_float(0x7f800000);
  for (int j = gid; j < n_left; j += gstride) {
    const uint32_t i = __ldg(left + j);
    const float4 b = __ldg(pts + i);
    float qx = b.x, qy = b.y, qz = b.z;
    if (XFORM) body_to_world(st, b.x, b.y, b.z, qx, qy, qz);
    unsigned long long key;
    uint32_t ref;
    int cnt;
    knn_search<8>(m, g, qx, qy, qz, max_d2, key, ref, cnt);
    cnt = min(cnt, k);
    if (g.gl < k) {
      const bool have = g.gl < cnt;
      nn[(size_t)i * k + g.gl] = have ? __ldg(m.slots + ref) : pad;
      if (nn_d2) nn_d2[(size_t)i * k + g.gl] = have ? __uint_as_float((uint32_t)(key >> 32)) : inf;
    }
    if (g.gl == 0) nn_cnt[i] = cnt;
  }
}

// Per-query search (lio_knn.cuh, one warp per query: knn_search_warp) of the queries the tile kernel deferred.
constexpr int kLeftGridMult = 4;  // blocks per SM of k_knn_left
template <bool XFORM, bool DEV>
__global__ void __launch_bounds__(kKnnBlock, 4) k_knn_left(MapView m, const float4* __restrict__ pts, const uint32_t* __restrict__ left,
                                                        StateArg st_param, int k, float max_d2, float4* __restrict__ nn,
                                                        float* __restrict__ nn_d2, int* __restrict__ nn_cnt, TileQueue* __restrict__ ctr,
                                                        TileQueue* __restrict__ ctr_next, uint32_t* __restrict__ stats,
                                                        const IekfCtl* __restrict__ ctl) {
  __shared__ StateArg s_st;
  __shared__ WarpList s_wl[kKnnBlock / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int warps_per_block = kKnnBlock / 32;
  const int wid = blockIdx.x * warps_per_block + warp;
  const int wstride = gridDim.x * warps_per_block;
  const bool skip = DEV && (ctl->finished || !ctl->converge);  // the tile kernel did not run: ctr is still zero
  const int n_left = skip ? 0 : (int)ctr->left;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    *ctr_next = TileQueue{0u, 0u, 0u, 0u};  // the next search uses the other counter set
    if (stats && !skip) atomicAdd(stats, (uint32_t)n_left);
  }
  if (skip) return;
  const StateArg& st = block_state<DEV>(st_param, ctl, s_st);
  const float4 pad = make_float4(0.f, 0.f, 0.f, __uint_as_float(kEmptyId));
  const float inf = __int_as_float(0x7f800000);
  for (int j = wid; j < n_left; j += wstride) {
    const uint32_t i = __ldg(left + j);
    const float4 b = __ldg(pts + i);
    float qx = b.x, qy = b.y, qz = b.z;
    if (XFORM) body_to_world(st, b.x, b.y, b.z, qx, qy, qz);
    unsigned long long key;
    uint32_t ref;
    int cnt;
    knn_search_warp(m, s_wl[warp], qx, qy, qz, max_d2, key, ref, cnt);
    cnt = min(cnt, k);
    if (lane < k) {
      const bool have = lane < cnt;
      nn[(size_t)i * k + lane] = have ? __ldg(m.slots + ref) : pad;
      if (nn_d2) nn_d2[(size_t)i * k + lane] = have ? __uint_as_float((uint32_t)(key >> 32)) : inf;
    }
    if (lane == 0) nn_cnt[i] = cnt;
  }
}

// ---------------------------------------------------------------- K4 + K5
// One thread per scan point: plane fit, validity, residual, Jacobian row; then per-warp pair sums in
// FP64 over rows staged in shared memory, block tree, and a deterministic last-block pass.
template <int NCOL, bool DEV>
__global__ void __launch_bounds__(kRedBlock, (NCOL == 6) ? 4 : 2) k_plane_reduce(const float4* __restrict__ body, int n, StateArg st_param,
                                                            const float4* __restrict__ nn, const int* __restrict__ nn_cnt,
                                                            float4* __restrict__ world, uint8_t* __restrict__ sel,
                                                            float4* __restrict__ plane, double* __restrict__ rows_out,
                                                            double* __restrict__ partials, MapCounters* ctr, HShareOut* out,
                                                            unsigned long long seq, IekfCtl* ctl, IekfIO* io, int iter_i,
                                                            float4* __restrict__ pfit, int refit_param) {
  constexpr int NC = NCOL + 3;                          // row | z | 1 | |pd2|
  constexpr int NV = NCOL * (NCOL + 1) / 2 + NCOL + 2;  // pair sums produced
  constexpr int VPL = (NV + 31) / 32;                   // values per lane
  // the staged rows are dead by the time the last block runs the IEKF algebra (DEV): both live in one buffer
  constexpr size_t kRowsBytes = sizeof(double) * kRedWarps * 32 * NC;
  constexpr size_t kRawBytes = (DEV && sizeof(IekfSmem) > kRowsBytes) ? sizeof(IekfSmem) : kRowsBytes;
  __shared__ __align__(16) unsigned char s_raw[kRawBytes];
  double (*s_rows)[32][NC] = reinterpret_cast<double (*)[32][NC]>(s_raw);
  __shared__ double s_blk[kRedWarps][NV];
  __shared__ uint8_t s_pa[NV], s_pb[NV];
  __shared__ bool s_last;
  __shared__ StateArg s_st;
  if (DEV && ctl->finished) return;
  // esti_plane depends only on the cached neighbours: it is redone when the search was (ekfom_data.converge), otherwise the
  // plane of the previous call is reused (bit-identical to refitting the same five points, as the reference does)
  const bool refit = DEV ? (ctl->converge != 0) : (refit_param != 0);
  const StateArg& st = block_state<DEV>(st_param, ctl, s_st);

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
      float pabcd[4];
      bool fit_ok = false;
      // point_selected_surf[i] persists between the calls of one scan: the search re-derives it
      // (five neighbours, the fifth within max_sqdist); a call that reuses the cached neighbours only looks at the
      // points the previous call left selected, so a point rejected by esti_plane or by the weight test stays out
      // until the next search.
      const bool cand = refit ? (nn_cnt[i] >= kK) : (sel[i] != 0);
      if (refit) {
        if (cand) {
          float P[kK][3];
#pragma unroll
          for (int k = 0; k < kK; ++k) {
            const float4 q = __ldg(nn + (size_t)i * kK + k);
            P[k][0] = q.x;
            P[k][1] = q.y;
            P[k][2] = q.z;
          }
          fit_ok = esti_plane(pabcd, P, 0.1f);
        }
        pfit[i] = fit_ok ? make_float4(pabcd[0], pabcd[1], pabcd[2], pabcd[3]) : make_float4(__int_as_float(0x7fc00000), 0.f, 0.f, 0.f);
      } else if (cand) {
        const float4 f = pfit[i];
        fit_ok = f.x == f.x;  // NaN marks "no valid plane"
        pabcd[0] = f.x, pabcd[1] = f.y, pabcd[2] = f.z, pabcd[3] = f.w;
      }
      {
        if (fit_ok) {
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
      // every lane stages its row (all zeros when the point is not selected: fma(0, 0, sum) == sum exactly), so the pair sums
      // run over a fixed 32 rows in ascending order - no bit scan, no data-dependent trip count
#pragma unroll
      for (int j = 0; j < NC; ++j) s_rows[warp][lane][j] = row[j];
      __syncwarp();
#pragma unroll
      for (int j = 0; j < VPL; ++j) {
        const int v = lane + 32 * j;
        if (v < NV) {
          const int a = s_pa[v], b = s_pb[v];
          double sum = acc[j];
#pragma unroll 8
          for (int r = 0; r < 32; ++r) sum = fma(s_rows[warp][r][a], s_rows[warp][r][b], sum);
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
  // last block: sum the block partials in a fixed order (8 interleaved chains per value, then a fixed 8-way combine),
  // assemble the result block in shared memory and ship it to the mapped host block with 16-byte stores
  __shared__ double s_part[kRedWarps][NV];
  __shared__ __align__(16) double s_out[160];
  for (int e = threadIdx.x; e < 160; e += blockDim.x) s_out[e] = 0.0;
  // (the loads of a batch are issued together; the additions keep their fixed order, absent blocks add +0.0)
  for (int v = lane; v < NV; v += 32) {
    const int nb = gridDim.x;
    double sum = 0.0;
    constexpr int U = 16;
    for (int b0 = warp; b0 < nb; b0 += kRedWarps * U) {
      double t[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int b = b0 + u * kRedWarps;
        t[u] = (b < nb) ? __ldcg(partials + (size_t)b * kMaxVals + v) : 0.0;
      }
#pragma unroll
      for (int u = 0; u < U; ++u) sum += t[u];
    }
    s_part[warp][v] = sum;
  }
  __syncthreads();
  if (threadIdx.x < NV) {
    double val = 0.0;
#pragma unroll
    for (int w = 0; w < kRedWarps; ++w) val += s_part[w][threadIdx.x];
    const int a = s_pa[threadIdx.x], b = s_pb[threadIdx.x];
    if (a < NCOL && b < NCOL) {
      s_out[a * 12 + b] = val;
      s_out[b * 12 + a] = val;
    } else if (b == NCOL) {
      s_out[144 + a] = val;
    } else if (a == NCOL + 1) {
      s_out[157] = val;  // n_eff
    } else {
      s_out[156] = val;  // res_sum
    }
  }
  __syncthreads();
  static_assert(offsetof(HShareOut, HTz) == 144 * 8 && offsetof(HShareOut, res_sum) == 156 * 8 && offsetof(HShareOut, n_eff) == 157 * 8 &&
                    offsetof(HShareOut, seq) == 158 * 8,
                "result block layout");
  if (DEV) {  // device-resident loop: the 23x23 update runs right here, nothing goes to the host before the scan is done
    if (threadIdx.x == 0) ctr->ticket = 0;
    iekf_step_block<kRedBlock>(ctl, io, s_out, iter_i, seq, *reinterpret_cast<IekfSmem*>(s_raw));
    return;
  }
  if (threadIdx.x < 79) reinterpret_cast<double2*>(out)[threadIdx.x] = reinterpret_cast<const double2*>(s_out)[threadIdx.x];
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    ctr->ticket = 0;
    *reinterpret_cast<volatile unsigned long long*>(&out->seq) = seq;
  }
}

// ---------------------------------------------------------------- far nearest neighbour (map_incremental's points_near[0])
// For the scan points whose search found nothing within max_sqdist: their unbounded nearest map point, searched from the
// world position they had at the LAST search (knn_state), as upstream's cached Nearest_Points[i][0] was.  far[i].w = id bits,
// 0xFFFFFFFF: none.  A warp takes 32 consecutive points and serves the (few) flagged ones one after the other.
__global__ void __launch_bounds__(128) k_far_nn(MapView m, const float4* __restrict__ body, int n, const double* __restrict__ knn_state,
                                                const int* __restrict__ nn_cnt, float4* __restrict__ far, const int* __restrict__ n_ptr) {
  if (n_ptr) n = *n_ptr;
  __shared__ StateArg s_st;
  if (threadIdx.x < 36) s_st.s[threadIdx.x] = knn_state[threadIdx.x];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool want = i < n && nn_cnt[i] == 0;
  float qx = 0.f, qy = 0.f, qz = 0.f;
  if (want) {
    const float4 b = __ldg(body + i);
    body_to_world(s_st, b.x, b.y, b.z, qx, qy, qz);
  }
  unsigned todo = __ballot_sync(0xFFFFFFFFu, want);
  float4 mine = make_float4(0.f, 0.f, 0.f, __uint_as_float(kEmptyId));
  while (todo) {
    const int src = __ffs(todo) - 1;
    todo &= todo - 1;
    const float x = __shfl_sync(0xFFFFFFFFu, qx, src), y = __shfl_sync(0xFFFFFFFFu, qy, src), z = __shfl_sync(0xFFFFFFFFu, qz, src);
    unsigned long long key;
    uint32_t ref;
    far_search_warp(m, x, y, z, key, ref);
    if (lane == src && key != ~0ull) mine = __ldg(m.slots + ref);
  }
  if (i < n) far[i] = mine;
}

// ---------------------------------------------------------------- map_incremental (device side)
// class per point: 0 skip, 1 PointToAdd, 2 PointNoNeedDownsample; block counts for the ordered compaction
template <bool DEV>
__global__ void __launch_bounds__(256) k_classify(const float4* __restrict__ body, int n, StateArg st_param, const float4* __restrict__ nn,
                                                  const int* __restrict__ nn_cnt, int flg_ekf_inited, float fs, MapCounters* ctr,
                                                  float4* __restrict__ world, uint8_t* __restrict__ cls,
                                                  uint32_t* __restrict__ blk_cnt, const IekfCtl* __restrict__ ctl,
                                                  const int* __restrict__ n_ptr, const float4* __restrict__ far) {
  if (n_ptr) n = *n_ptr;  // graph form: the grid covers the padded size, the true size sits in device memory
  __shared__ uint32_t s_cnt[2];
  __shared__ StateArg s_st;
  if (threadIdx.x < 2) s_cnt[threadIdx.x] = 0;
  const StateArg& st = block_state<DEV>(st_param, ctl, s_st);
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
    // points_near[0]: the nearest of the cached neighbours, or - when the search found none within max_sqdist - the far
    // nearest map point upstream's unbounded search would have cached (k_far_nn); none at all: PointToAdd
    float4 n0 = make_float4(0.f, 0.f, 0.f, __uint_as_float(kEmptyId));
    if (cnt > 0) n0 = __ldg(nn + (size_t)i * kK);
    else if (far) n0 = __ldg(far + i);
    if (__float_as_uint(n0.w) != kEmptyId && flg_ekf_inited) {
      const float mx = floorf(wx / fs) * fs + 0.5f * fs, my = floorf(wy / fs) * fs + 0.5f * fs,
                  mz = floorf(wz / fs) * fs + 0.5f * fs;
      const float dist = sqdist3(wx, wy, wz, mx, my, mz);
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
                                                       float4* __restrict__ list_nodown, const int* __restrict__ n_ptr) {
  if (n_ptr) n = *n_ptr;
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

// ---------------------------------------------------------------- UndistortPcl (the step in front of the path)
// Backward motion compensation of ImuProcess::UndistortPcl (IMU_Processing.hpp, absent submodule): every point is
// re-expressed in the LiDAR frame at the scan end from the IMU pose that precedes its own timestamp.  The reference
// walks the time-sorted cloud backwards with two pointers; here each thread finds its pose by bisection (same segment:
// the last pose whose offset_time is strictly below the point's time; points at or before pose[0] stay untouched).
constexpr int kMaxImuPoses = 512;
__global__ void __launch_bounds__(256) k_undistort(float4* __restrict__ body, const float* __restrict__ t_ms, int n,
                                                   const double* __restrict__ pose, int n_pose, StateArg end) {
  using namespace esekf;
  __shared__ double s_off[kMaxImuPoses];
  for (int k = threadIdx.x; k < n_pose; k += blockDim.x) s_off[k] = pose[22 * k];
  __syncthreads();
  const State e = load(end.s);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const double t = (double)t_ms[i] / double(1000);
    if (!(t > s_off[0])) continue;
    int lo = 0, hi = n_pose - 2;  // largest head index with offset < t
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (s_off[mid] < t) lo = mid;
      else hi = mid - 1;
    }
    const double* head = pose + 22 * lo;
    const double* tail = head + 22;
    const double dt = t - head[0];
    Mat3 R_imu;
    for (int k = 0; k < 9; ++k) R_imu.m[k] = head[13 + k];
    const Vec3 gyr{{tail[4], tail[5], tail[6]}};
    Mat3 Ex = eye3();
    const double wn = norm(gyr);
    if (wn > 1e-7) {
      const Mat3 K = skew(Vec3{{gyr[0] / wn, gyr[1] / wn, gyr[2] / wn}});
      const double ang = wn * dt;
      Ex = add(add(eye3(), scale(K, sin(ang))), scale(mul(K, K), 1.0 - cos(ang)));
    }
    const Mat3 R_i = mul(R_imu, Ex);
    const float4 b = body[i];
    const Vec3 P_i{{(double)b.x, (double)b.y, (double)b.z}};
    Vec3 T_ei;
    for (int k = 0; k < 3; ++k) T_ei[k] = ((head[10 + k] + head[7 + k] * dt) + ((tail[1 + k] * 0.5) * dt) * dt) - e.pos[k];
    Vec3 p = mul(e.offR, P_i);
    for (int k = 0; k < 3; ++k) p[k] += e.offT[k];
    p = mul(R_i, p);
    for (int k = 0; k < 3; ++k) p[k] += T_ei[k];
    p = mul(transpose(e.rot), p);
    for (int k = 0; k < 3; ++k) p[k] -= e.offT[k];
    p = mul(transpose(e.offR), p);
    body[i] = make_float4((float)p[0], (float)p[1], (float)p[2], b.w);
  }
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
  LIO_CUDA(cudaMalloc(&c->pfit, sizeof(float4) * n));
  LIO_CUDA(cudaMalloc(&c->far_nn, sizeof(float4) * n));
  LIO_CUDA(cudaMalloc(&c->knn_state_dev, sizeof(double) * 36));
  LIO_CUDA(cudaMemsetAsync(c->knn_state_dev, 0, sizeof(double) * 36, c->stream));
  LIO_CUDA(cudaMalloc(&c->cls, n));
  LIO_CUDA(cudaMalloc(&c->blk_cnt, sizeof(uint32_t) * 2 * ((n + 255) / 256 + 1)));
  LIO_CUDA(cudaMalloc(&c->list_add, sizeof(float4) * n));
  LIO_CUDA(cudaMalloc(&c->list_nodown, sizeof(float4) * n));
  LIO_CUDA(cudaMalloc(&c->tile_stats, sizeof(uint32_t) * 12));  // [0..3] stats, [4..11] the two TileQueue counter sets
  LIO_CUDA(cudaMemsetAsync(c->tile_stats, 0, sizeof(uint32_t) * 12, c->stream));
  {
    const char* e = getenv("LIO_KNN_MODE");
    c->knn_mode = (e && std::strcmp(e, "group") == 0) ? 0 : 1;
    {  // tuning knobs of the tile search (process-wide constants: written on every context creation)
      e = getenv("LIO_TILE_MAXSUB");
      const int sub = e ? std::max(1, std::min(32, atoi(e))) : kTileMaxSub;
      e = getenv("LIO_TILE_MAXE");
      const int me = e ? std::max(1, std::min(8, atoi(e))) : kTileMaxE;
      LIO_CUDA(cudaMemcpyToSymbolAsync(c_tile_max_sub, &sub, sizeof(int), 0, cudaMemcpyHostToDevice, c->stream));
      LIO_CUDA(cudaMemcpyToSymbolAsync(c_tile_max_e, &me, sizeof(int), 0, cudaMemcpyHostToDevice, c->stream));
      e = getenv("LIO_TILE_ORDER");
      const int hil = (e && (e[0] == 'm' || e[0] == '0')) ? 0 : 1;
      LIO_CUDA(cudaMemcpyToSymbolAsync(c_tile_hilbert, &hil, sizeof(int), 0, cudaMemcpyHostToDevice, c->stream));
      e = getenv("LIO_TILE_OFFER");
      const int offer = (e && e[0] == '0') ? 0 : 1;
      LIO_CUDA(cudaMemcpyToSymbolAsync(c_tile_offer, &offer, sizeof(int), 0, cudaMemcpyHostToDevice, c->stream));
      e = getenv("LIO_TILE_GROUP");
      const int blk = (e && e[0] == 'b') ? 1 : 0;
      LIO_CUDA(cudaMemcpyToSymbolAsync(c_tile_block, &blk, sizeof(int), 0, cudaMemcpyHostToDevice, c->stream));
      e = getenv("LIO_TILE_HALVE");
      const int halve = (e && e[0] == '1') ? 1 : 0;
      LIO_CUDA(cudaMemcpyToSymbolAsync(c_tile_halve, &halve, sizeof(int), 0, cudaMemcpyHostToDevice, c->stream));
      e = getenv("LIO_TILE_SPLIT");
      const int split = e ? std::max(0, std::min(8, atoi(e))) : 8;
      LIO_CUDA(cudaMemcpyToSymbolAsync(c_tile_split, &split, sizeof(int), 0, cudaMemcpyHostToDevice, c->stream));
      LIO_CUDA(cudaStreamSynchronize(c->stream));
    }
    e = getenv("LIO_TILE_PROF");  // per-tile cycle / pass counters of the last scan search (lio_tile_profile)
    if (e && std::strcmp(e, "1") == 0) {
      c->tile_prof_cap = 2 * (uint32_t)((n + 31) / 32 + 1);  // tickets recorded: the tiles and as many pushed tasks
      LIO_CUDA(cudaMalloc(&c->tile_prof, sizeof(uint4) * 2 * (size_t)c->tile_prof_cap));
      LIO_CUDA(cudaMemsetAsync(c->tile_prof, 0, sizeof(uint4) * 2 * (size_t)c->tile_prof_cap, c->stream));
    }
    e = getenv("LIO_BATCH_LEFT");
    c->batch_left_warp = (e && std::strcmp(e, "warp") == 0) ? 1 : 0;
    e = getenv("LIO_RESORT");
    c->resort_each_search = !(e && e[0] == '0');
    e = getenv("LIO_SORT_GRAPH");
    c->sort_graph_ok = !(e && std::strcmp(e, "0") == 0);
    e = getenv("LIO_MAP_GRAPH");
    c->map_graph_ok = !(e && std::strcmp(e, "0") == 0);
    e = getenv("LIO_IEKF_MODE");
    c->iekf_mode = (e && std::strcmp(e, "device") == 0) ? 1 : 0;
  }
  if (int rc = tile_ensure(c, (uint32_t)n)) return rc;
  c->partial_blocks = c->n_sms * 4;
  LIO_CUDA(cudaMalloc(&c->partials, sizeof(double) * kMaxVals * (size_t)c->partial_blocks));
  LIO_CUDA(cudaHostAlloc(&c->out_host, sizeof(HShareOut), cudaHostAllocMapped));
  std::memset(c->out_host, 0, sizeof(HShareOut));
  LIO_CUDA(cudaHostGetDevicePointer(&c->out_dev, c->out_host, 0));
  LIO_CUDA(cudaMalloc(&c->iekf_ctl, sizeof(IekfCtl)));
  LIO_CUDA(cudaMemsetAsync(c->iekf_ctl, 0, sizeof(IekfCtl), c->stream));
  LIO_CUDA(cudaHostAlloc(&c->io_host, sizeof(IekfIO), cudaHostAllocMapped));
  std::memset(c->io_host, 0, sizeof(IekfIO));
  LIO_CUDA(cudaHostGetDevicePointer(&c->io_dev, c->io_host, 0));
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
  cudaFree(c->pfit);
  cudaFree(c->far_nn);
  cudaFree(c->knn_state_dev);
  cudaFree(c->rows);
  cudaFree(c->cls);
  cudaFree(c->blk_cnt);
  cudaFree(c->list_add);
  cudaFree(c->list_nodown);
  cudaFree(c->partials);
  cudaFree(c->tile_stats);
  cudaFree(c->sort_scratch);
  cudaFree(c->tile_prof);
  for (int j = 0; j < 2; ++j) {
    cudaFree(c->tile_keys[j]);
    cudaFree(c->tile_vals[j]);
  }
  cudaFree(c->tile_sorted);
  cudaFree(c->tile_queue);
  if (c->out_host) cudaFreeHost(c->out_host);
  for (auto& g : c->sort_graphs) cudaGraphExecDestroy((cudaGraphExec_t)g.second);
  c->sort_graphs.clear();
  for (auto& g : c->map_graphs) cudaGraphExecDestroy((cudaGraphExec_t)g.second);
  c->map_graphs.clear();
  if (c->map_args_host) cudaFreeHost(c->map_args_host);
  cudaFree(c->map_args_dev);
  if (c->keys_args_host) cudaFreeHost(c->keys_args_host);
  cudaFree(c->keys_args_dev);
  cudaFree(c->iekf_ctl);
  cudaFree(c->scan_t);
  cudaFree(c->imu_pose);
  if (c->io_host) cudaFreeHost(c->io_host);
}

int scan_undistort(lio_ctx* c, const float* t_ms_host, const double* pose_host, int n_pose, const double* end_state) {
  const int n = c->scan_n;
  if (n_pose < 2 || n_pose > kMaxImuPoses) {
    set_error("lio_scan_undistort: need 2..%d IMU poses (got %d)", kMaxImuPoses, n_pose);
    return LIO_E_ARG;
  }
  if (!c->scan_t) LIO_CUDA(cudaMalloc(&c->scan_t, sizeof(float) * (size_t)c->cfg.max_scan));
  if (!c->imu_pose) LIO_CUDA(cudaMalloc(&c->imu_pose, sizeof(double) * 22 * kMaxImuPoses));
  LIO_CUDA(cudaMemcpyAsync(c->scan_t, t_ms_host, sizeof(float) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
  LIO_CUDA(cudaMemcpyAsync(c->imu_pose, pose_host, sizeof(double) * 22 * (size_t)n_pose, cudaMemcpyHostToDevice, c->stream));
  const int grid = std::max(1, std::min((n + 255) / 256, c->n_sms * 8));
  k_undistort<<<grid, 256, 0, c->stream>>>(c->body, c->scan_t, n, c->imu_pose, n_pose, make_state(end_state));
  c->launches += 1;
  LIO_CUDA(cudaGetLastError());
  LIO_CUDA(cudaStreamSynchronize(c->stream));  // the host buffers may be reused on return
  return LIO_OK;
}

constexpr size_t kSortScratchBytes = sizeof(SortScratch) + sizeof(uint32_t) * (size_t)kSortBins * kSortBins;  // 64 MB of level-2 counters
int tile_ensure(lio_ctx* c, uint32_t n) {
  if (!c->sort_scratch) {
    LIO_CUDA(cudaMalloc(&c->sort_scratch, kSortScratchBytes));
    LIO_CUDA(cudaMemsetAsync(c->sort_scratch, 0, kSortScratchBytes, c->stream));
  }
  if (n <= c->tile_cap) return LIO_OK;
  for (auto& g : c->sort_graphs) cudaGraphExecDestroy((cudaGraphExec_t)g.second);  // they captured the old buffers
  c->sort_graphs.clear();
  for (int j = 0; j < 2; ++j) {
    cudaFree(c->tile_keys[j]);
    cudaFree(c->tile_vals[j]);
    c->tile_keys[j] = c->tile_vals[j] = nullptr;
  }
  cudaFree(c->tile_sorted);
  cudaFree(c->tile_queue);
  c->tile_sorted = nullptr;
  c->tile_queue = nullptr;
  c->tile_cap = 0;
  const uint32_t cap = n + n / 8 + 1024;
  LIO_CUDA(cudaMalloc(&c->tile_keys[0], sizeof(uint32_t) * cap));  // keys
  for (int j = 0; j < 2; ++j) LIO_CUDA(cudaMalloc(&c->tile_vals[j], sizeof(uint32_t) * cap));  // [0] leftover list, [1] order
  LIO_CUDA(cudaMalloc(&c->tile_sorted, sizeof(float4) * cap));                                  // the query points in curve order
  LIO_CUDA(cudaMalloc(&c->tile_queue, sizeof(unsigned long long) * 2 * (size_t)cap));           // pushed (tile, lane mask) tasks
  LIO_CUDA(cudaMemsetAsync(c->tile_queue, 0, sizeof(unsigned long long) * 2 * (size_t)cap, c->stream));  // self-cleaning afterwards
  c->tile_cap = cap;
  c->order_valid = false;
  return LIO_OK;
}

// the sort kernels on c->stream: keys -> tile_keys[0], order -> tile_vals[1].  `args` non-null: graph form (state and n read
// from the device block).
static void sort_launch(lio_ctx* c, const float4* pts, int n_grid, const KeysArgs* args, const StateArg& st, int n, bool xform) {
  SortScratch* sc = (SortScratch*)c->sort_scratch;
  const int grid = std::max(1, std::min((n_grid + 255) / 256, c->n_sms * 8));
  if (xform)
    k_sort_count<true><<<grid, 256, 0, c->stream>>>(c->map, pts, args, st, n, c->tile_keys[0], sc);
  else
    k_sort_count<false><<<grid, 256, 0, c->stream>>>(c->map, pts, args, st, n, c->tile_keys[0], sc);
  k_sort_scan<<<1 + c->n_sms * 3, kSortScanBlock, 0, c->stream>>>(sc);  // block 0: level 1; the others stride over the slots in use
  k_sort_place<<<grid, 256, 0, c->stream>>>(c->tile_keys[0], args, n, sc, c->tile_vals[1], pts, c->tile_sorted);
  k_sort_clear<<<c->n_sms * 2, 256, 0, c->stream>>>(sc);
  c->launches += 4;
}

// tile order (curve order) of n query points (body points at `st` when xform) -> c->tile_vals[1]
static int tile_order(lio_ctx* c, const float4* pts, int n, const StateArg& st, bool xform) {
  int rc = tile_ensure(c, (uint32_t)n);
  if (rc) return rc;
  sort_launch(c, pts, n, nullptr, st, n, xform);
  LIO_CUDA(cudaGetLastError());
  return LIO_OK;
}

// Scan ordering through a captured graph (one launch for the argument copy + the three kernels).  Graphs are cached per padded
// size (multiples of 8192); the true size and the state sit in a small device block refreshed by the graph's memcpy node.
static int tile_order_graph(lio_ctx* c, int n, const double* state_host) {
  int rc = tile_ensure(c, (uint32_t)n);
  if (rc) return rc;
  const int n_pad = std::min<int>((int)c->tile_cap, ((n + 8191) / 8192) * 8192);
  if (!c->keys_args_host) {
    LIO_CUDA(cudaHostAlloc(&c->keys_args_host, sizeof(KeysArgs), cudaHostAllocDefault));
    LIO_CUDA(cudaMalloc(&c->keys_args_dev, sizeof(KeysArgs)));
  }
  cudaGraphExec_t exec = nullptr;
  for (auto& g : c->sort_graphs)
    if (g.first == n_pad) exec = (cudaGraphExec_t)g.second;
  if (!exec) {
    cudaGraph_t graph = nullptr;
    LIO_CUDA(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeRelaxed));
    cudaMemcpyAsync(c->keys_args_dev, c->keys_args_host, sizeof(KeysArgs), cudaMemcpyHostToDevice, c->stream);
    const int64_t l0 = c->launches;
    StateArg unused{};
    sort_launch(c, c->body, n_pad, (const KeysArgs*)c->keys_args_dev, unused, 0, true);
    c->launches = l0;
    const cudaError_t ce = cudaStreamEndCapture(c->stream, &graph);
    if (ce != cudaSuccess || !graph) {
      if (graph) cudaGraphDestroy(graph);
      cudaGetLastError();
      c->sort_graph_ok = false;  // fall back to plain launches for the rest of this context's life
      return LIO_E_STATE;
    }
    const cudaError_t ie = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ie != cudaSuccess || !exec) {
      cudaGetLastError();
      c->sort_graph_ok = false;
      return LIO_E_STATE;
    }
    c->sort_graphs.emplace_back(n_pad, (void*)exec);
  }
  KeysArgs* a = (KeysArgs*)c->keys_args_host;
  std::memcpy(a->s, state_host, sizeof(a->s));
  a->n = n;
  LIO_CUDA(cudaGraphLaunch(exec, c->stream));
  c->launches += 4;
  return LIO_OK;
}

// a quarter more warps than tiles (up to 8 blocks per SM): the spare ones start on pushed sub-tiles right away
static int tile_grid(const lio_ctx* c, int n) {
  const int tiles = (n + 31) / 32;
  const int blocks = (tiles + kTileWarps - 1) / kTileWarps;
  return std::max(1, std::min(blocks + blocks / 4 + 1, c->n_sms * kTileBlocksPerSm));
}

// tasks (tiles + pushed sub-tiles) of the last search; its counter set stays intact until the next search has run
int tile_tasks_of_last_search(lio_ctx* c, uint32_t* n_tasks) {
  TileQueue q;
  const TileQueue* sets = reinterpret_cast<const TileQueue*>(c->tile_stats + 4);
  LIO_CUDA(cudaMemcpyAsync(&q, sets + ((c->tile_flip + 1) & 1), sizeof(TileQueue), cudaMemcpyDeviceToHost, c->stream));
  LIO_CUDA(cudaStreamSynchronize(c->stream));
  *n_tasks = (uint32_t)((c->scan_n + 31) / 32) + q.tail;
  return LIO_OK;
}

int scan_knn(lio_ctx* c, const double* state_host, bool dev_loop) {
  const int n = c->scan_n;
  const StateArg st = make_state(state_host);
  // The curve order is recomputed when the state has moved enough since the last ordering for queries to change cells
  // (> 2 cm or > 1 mrad: 5 cm at 50 m); between the searches of a scan that is the size of the IEKF correction.
  bool moved = true;
  if (c->order_valid && c->resort_each_search) {
    double dp = 0.0, dr = 0.0;
    for (int k = 0; k < 3; ++k) dp = std::max(dp, std::fabs(state_host[k] - c->sort_state[k]));
    for (int k = 3; k < 21; ++k) dr = std::max(dr, std::fabs(state_host[k] - c->sort_state[k]));  // rot, offset_R (entries), offset_T
    moved = dp > 0.02 || dr > 0.001;
  }
  if (c->knn_mode == 1 && (!c->order_valid || (c->resort_each_search && moved))) {  // LIO_RESORT=0: once per scan
    timer_begin(c, T_SORT);
    int rc = c->sort_graph_ok ? tile_order_graph(c, n, state_host) : LIO_E_STATE;
    if (rc == LIO_E_STATE && !c->sort_graph_ok) rc = tile_order(c, c->body, n, st, true);
    timer_end(c);
    if (rc) return rc;
    c->order_valid = true;
    std::memcpy(c->sort_state, state_host, sizeof(c->sort_state));
  }
  timer_begin(c, T_KNN);
  if (c->knn_mode == 1) {
    TileQueue* sets = reinterpret_cast<TileQueue*>(c->tile_stats + 4);
    TileQueue* ctr = sets + (c->tile_flip & 1);
    TileQueue* ctr_next = sets + ((c->tile_flip + 1) & 1);
    c->tile_flip ^= 1;
    const uint32_t qcap = 2 * c->tile_cap;
    // nn_d2 is not written on the scan path: nothing downstream reads it (the plane fit and map_incremental use the points)
    if (dev_loop) {
      k_knn_tile<true, true><<<tile_grid(c, n), kTileWarps * 32, 0, c->stream>>>(c->map, c->tile_sorted, c->tile_vals[1], n, st, kK, c->cfg.max_sqdist,
                                                                                  c->nn, nullptr, c->nn_cnt, ctr, c->tile_queue, qcap, c->tile_vals[0],
                                                                                  c->iekf_ctl, c->tile_stats, c->tile_prof, c->tile_prof_cap, c->knn_state_dev);
      k_knn_left<true, true><<<c->n_sms * kLeftGridMult, kKnnBlock, 0, c->stream>>>(c->map, c->body, c->tile_vals[0], st, kK, c->cfg.max_sqdist, c->nn,
                                                                         nullptr, c->nn_cnt, ctr, ctr_next, c->tile_stats, c->iekf_ctl);
    } else {
      k_knn_tile<true, false><<<tile_grid(c, n), kTileWarps * 32, 0, c->stream>>>(c->map, c->tile_sorted, c->tile_vals[1], n, st, kK, c->cfg.max_sqdist,
                                                                                   c->nn, nullptr, c->nn_cnt, ctr, c->tile_queue, qcap, c->tile_vals[0],
                                                                                   nullptr, c->tile_stats, c->tile_prof, c->tile_prof_cap, c->knn_state_dev);
      k_knn_left<true, false><<<c->n_sms * kLeftGridMult, kKnnBlock, 0, c->stream>>>(c->map, c->body, c->tile_vals[0], st, kK, c->cfg.max_sqdist, c->nn,
                                                                          nullptr, c->nn_cnt, ctr, ctr_next, c->tile_stats, nullptr);
    }
    c->launches += 1;
  } else {
    k_scan_knn<8><<<knn_grid(c, n, 8), kKnnBlock, 0, c->stream>>>(c->map, c->body, n, st, c->cfg.max_sqdist, c->nn, c->nn_d2, c->nn_cnt, c->knn_state_dev);
  }
  c->launches += 1;
  timer_end(c);
  LIO_CUDA(cudaGetLastError());
  c->nn_valid = true;
  return LIO_OK;
}

int knn_batch(lio_ctx* c, const float4* q_dev, int n, int k, float max_d2, float4* out_nn, float* out_d2, int* out_cnt) {
  if (c->knn_mode == 1) {
    StateArg st{};
    timer_begin(c, T_SORT);
    int rc = tile_order(c, q_dev, n, st, false);
    timer_end(c);
    if (rc) return rc;
    c->order_valid = false;  // the scan's order lived in the same buffers
  }
  timer_begin(c, T_KNN);
  if (c->knn_mode == 1) {
    StateArg st{};
    TileQueue* sets = reinterpret_cast<TileQueue*>(c->tile_stats + 4);
    TileQueue* ctr = sets + (c->tile_flip & 1);
    TileQueue* ctr_next = sets + ((c->tile_flip + 1) & 1);
    c->tile_flip ^= 1;
    k_knn_tile<false, false><<<tile_grid(c, n), kTileWarps * 32, 0, c->stream>>>(c->map, c->tile_sorted, c->tile_vals[1], n, st, k, max_d2, out_nn,
                                                                                  out_d2, out_cnt, ctr, c->tile_queue, 2 * c->tile_cap, c->tile_vals[0],
                                                                                  nullptr, c->tile_stats, nullptr, 0u, nullptr);
    if (c->batch_left_warp)
      k_knn_left<false, false><<<c->n_sms * kLeftGridMult, kKnnBlock, 0, c->stream>>>(c->map, q_dev, c->tile_vals[0], st, k, max_d2, out_nn, out_d2, out_cnt,
                                                                         ctr, ctr_next, c->tile_stats, nullptr);
    else
      k_knn_left8<false, false><<<c->n_sms * 4, kKnnBlock, 0, c->stream>>>(c->map, q_dev, c->tile_vals[0], st, k, max_d2, out_nn, out_d2, out_cnt,
                                                                         ctr, ctr_next, c->tile_stats, nullptr);
    c->launches += 1;
  } else {
    k_knn_batch<8><<<knn_grid(c, n, 8), kKnnBlock, 0, c->stream>>>(c->map, q_dev, n, k, max_d2, out_nn, out_d2, out_cnt);
  }
  c->launches += 1;
  timer_end(c);
  LIO_CUDA(cudaGetLastError());
  return LIO_OK;
}

// launches K4+K5; the result lands in c->out_host once the stream reaches the end of the kernel
int scan_plane_reduce(lio_ctx* c, const double* state_host, bool want_rows, bool refit, bool dev_loop, int iter_i,
                      unsigned long long io_seq) {
  const int n = c->scan_n;
  if (want_rows && !c->rows) LIO_CUDA(cudaMalloc(&c->rows, sizeof(double) * 13 * (size_t)c->cfg.max_scan));
  const int tiles = (n + 31) / 32;
  const int grid = std::max(1, std::min((tiles + kRedWarps - 1) / kRedWarps, c->partial_blocks));
  timer_begin(c, T_PLANE);
  if (dev_loop) {
    StateArg unused{};
    if (c->cfg.extrinsic_est_en)
      k_plane_reduce<12, true><<<grid, kRedBlock, 0, c->stream>>>(c->body, n, unused, c->nn, c->nn_cnt, c->world, c->sel, c->plane, nullptr,
                                                                  c->partials, c->map.ctr, nullptr, io_seq, c->iekf_ctl, c->io_dev, iter_i, c->pfit, 1);
    else
      k_plane_reduce<6, true><<<grid, kRedBlock, 0, c->stream>>>(c->body, n, unused, c->nn, c->nn_cnt, c->world, c->sel, c->plane, nullptr,
                                                                 c->partials, c->map.ctr, nullptr, io_seq, c->iekf_ctl, c->io_dev, iter_i, c->pfit, 1);
  } else {
    const unsigned long long seq = ++c->out_seq;
    if (c->cfg.extrinsic_est_en)
      k_plane_reduce<12, false><<<grid, kRedBlock, 0, c->stream>>>(c->body, n, make_state(state_host), c->nn, c->nn_cnt, c->world, c->sel,
                                                                   c->plane, want_rows ? c->rows : nullptr, c->partials, c->map.ctr,
                                                                   c->out_dev, seq, nullptr, nullptr, 0, c->pfit, refit ? 1 : 0);
    else
      k_plane_reduce<6, false><<<grid, kRedBlock, 0, c->stream>>>(c->body, n, make_state(state_host), c->nn, c->nn_cnt, c->world, c->sel,
                                                                  c->plane, want_rows ? c->rows : nullptr, c->partials, c->map.ctr,
                                                                  c->out_dev, seq, nullptr, nullptr, 0, c->pfit, refit ? 1 : 0);
  }
  c->launches += 1;
  timer_end(c);
  LIO_CUDA(cudaGetLastError());
  return LIO_OK;
}

// map_incremental as one captured graph (13 small launches otherwise: the GPU waits for the host between them).  Cached per
// (padded size, flg_ekf_inited); state and true size come from a device block refreshed by the graph's memcpy node.
static int classify_insert_graph(lio_ctx* c, const double* state_host, int flg_ekf_inited) {
  const int n = c->scan_n;
  const int n_pad = std::min<int>((int)c->cfg.max_scan, ((n + 8191) / 8192) * 8192);
  int rc = map_ensure_batch(c, (uint32_t)n_pad);  // allocations must not happen inside the capture
  if (rc) return rc;
  if (!c->map_args_host) {
    LIO_CUDA(cudaHostAlloc(&c->map_args_host, sizeof(KeysArgs), cudaHostAllocDefault));
    LIO_CUDA(cudaMalloc(&c->map_args_dev, sizeof(IekfCtl)));  // read through the IekfCtl view (x[36] first), n right after
    LIO_CUDA(cudaMemsetAsync(c->map_args_dev, 0, sizeof(IekfCtl), c->stream));
  }
  const int key = n_pad * 2 + (flg_ekf_inited ? 1 : 0);
  cudaGraphExec_t exec = nullptr;
  for (auto& g : c->map_graphs)
    if (g.first == key) exec = (cudaGraphExec_t)g.second;
  if (!exec) {
    const int blocks = (n_pad + 255) / 256;
    const IekfCtl* st_view = (const IekfCtl*)c->map_args_dev;  // block_state<true> reads x[0..35] only
    const int* n_ptr = (const int*)((const char*)c->map_args_dev + offsetof(KeysArgs, n));
    cudaGraph_t graph = nullptr;
    const bool was_timing = c->timing;
    c->timing = false;  // no event records inside the capture
    LIO_CUDA(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeRelaxed));
    cudaMemcpyAsync(c->map_args_dev, c->map_args_host, sizeof(KeysArgs), cudaMemcpyHostToDevice, c->stream);
    StateArg unused{};
    k_far_nn<<<(n_pad + 127) / 128, 128, 0, c->stream>>>(c->map, c->body, n_pad, c->knn_state_dev, c->nn_cnt, c->far_nn, n_ptr);
    k_classify<true><<<blocks, 256, 0, c->stream>>>(c->body, n_pad, unused, c->nn, c->nn_cnt, flg_ekf_inited, c->cfg.filter_size_map,
                                                    c->map.ctr, c->world, c->cls, c->blk_cnt, st_view, n_ptr, c->far_nn);
    k_scan_blocks<<<1, 1024, 0, c->stream>>>(c->blk_cnt, blocks, c->map.ctr);
    k_scatter_lists<<<blocks, 256, 0, c->stream>>>(c->world, c->cls, n_pad, c->blk_cnt, c->list_add, c->list_nodown, n_ptr);
    const int64_t l0 = c->launches;
    int r1 = map_add_device(c, c->list_add, (uint32_t)n_pad, &c->map.ctr->list_add, true);
    int r2 = map_add_device(c, c->list_nodown, (uint32_t)n_pad, &c->map.ctr->list_nodown, false);
    c->map_graph_launches = 4 + (int)(c->launches - l0);
    c->launches = l0;
    const cudaError_t ce = cudaStreamEndCapture(c->stream, &graph);
    c->timing = was_timing;
    cudaError_t ie = cudaErrorUnknown;
    if (r1 == LIO_OK && r2 == LIO_OK && ce == cudaSuccess && graph) ie = cudaGraphInstantiate(&exec, graph, 0);
    if (graph) cudaGraphDestroy(graph);
    if (ie != cudaSuccess || !exec) {
      cudaGetLastError();
      c->map_graph_ok = false;
      return LIO_E_STATE;
    }
    c->map_graphs.emplace_back(key, (void*)exec);
  }
  KeysArgs* a = (KeysArgs*)c->map_args_host;
  std::memcpy(a->s, state_host, sizeof(a->s));
  a->n = n;
  timer_begin(c, T_INSERT);  // graph form: the whole map update is one timed region
  LIO_CUDA(cudaGraphLaunch(exec, c->stream));
  timer_end(c);
  c->launches += c->map_graph_launches;
  return LIO_OK;
}

int scan_classify_insert(lio_ctx* c, const double* state_host, int flg_ekf_inited, bool dev_loop) {
  const int n = c->scan_n;
  if (n <= 0) return LIO_OK;
  if (!dev_loop && c->map_graph_ok) {
    const int rc = classify_insert_graph(c, state_host, flg_ekf_inited);
    if (rc != LIO_E_STATE || c->map_graph_ok) return rc;  // LIO_E_STATE with the flag cleared: capture failed, plain launches
  }
  const int blocks = (n + 255) / 256;
  timer_begin(c, T_CLASSIFY);
  k_far_nn<<<(n + 127) / 128, 128, 0, c->stream>>>(c->map, c->body, n, c->knn_state_dev, c->nn_cnt, c->far_nn, nullptr);
  if (dev_loop) {
    StateArg unused{};
    k_classify<true><<<blocks, 256, 0, c->stream>>>(c->body, n, unused, c->nn, c->nn_cnt, flg_ekf_inited, c->cfg.filter_size_map, c->map.ctr,
                                                    c->world, c->cls, c->blk_cnt, c->iekf_ctl, nullptr, c->far_nn);
  } else {
    k_classify<false><<<blocks, 256, 0, c->stream>>>(c->body, n, make_state(state_host), c->nn, c->nn_cnt, flg_ekf_inited,
                                                     c->cfg.filter_size_map, c->map.ctr, c->world, c->cls, c->blk_cnt, nullptr, nullptr, c->far_nn);
  }
  k_scan_blocks<<<1, 1024, 0, c->stream>>>(c->blk_cnt, blocks, c->map.ctr);
  k_scatter_lists<<<blocks, 256, 0, c->stream>>>(c->world, c->cls, n, c->blk_cnt, c->list_add, c->list_nodown, nullptr);
  c->launches += 4;
  timer_end(c);
  LIO_CUDA(cudaGetLastError());
  int rc = map_add_device(c, c->list_add, (uint32_t)n, &c->map.ctr->list_add, true);
  if (rc) return rc;
  return map_add_device(c, c->list_nodown, (uint32_t)n, &c->map.ctr->list_nodown, false);
}

// ---------------------------------------------------------------- device-resident IEKF loop (lio_iekf.cuh)
namespace {
__global__ void __launch_bounds__(256) k_iekf_begin(IekfCtl* ctl, const IekfIO* io) {
  const int tid = threadIdx.x;
  for (int e = tid; e < 36; e += 256) {
    const double v = io->state_in[e];
    ctl->x[e] = v;
    ctl->x_prop[e] = v;
  }
  for (int e = tid; e < kNN; e += 256) ctl->P_prop[e] = io->P_in[e];
  if (tid < kN) ctl->limit[tid] = io->limit[tid];
  if (tid == 0) {
    ctl->R = io->R;
    ctl->max_iter = io->max_iter;
    ctl->converge = 1;
    ctl->t = 0;
    ctl->done = 0;
    ctl->finished = 0;
  }
}
}  // namespace

// esekfom::update_iterated_dyn_share_modified (+ map_incremental when with_map) enqueued as one chain of kernels: every
// model call reads the state the previous call's reduction left in ctl->x; the host waits once, on the mapped result block.
int iekf_update_device(lio_ctx* c, double* state, double* P_io, double R, int max_iter, const double* limit, double* iter_states,
                       int* n_eff_trace, int* knn_trace, int* iters_done, bool with_map, int flg_ekf_inited) {
  IekfIO* io = c->io_host;
  std::memcpy(io->state_in, state, sizeof(io->state_in));
  std::memcpy(io->P_in, P_io, sizeof(io->P_in));
  std::memcpy(io->limit, limit, sizeof(io->limit));
  io->R = R;
  io->max_iter = max_iter;
  const unsigned long long seq = ++c->io_seq;
  k_iekf_begin<<<1, 256, 0, c->stream>>>(c->iekf_ctl, c->io_dev);
  c->launches += 1;
  for (int i = -1; i < max_iter; ++i) {
    int rc = scan_knn(c, state, true);  // by-value state: only the once-per-scan tile keys use it (the propagated state)
    if (rc == LIO_OK) rc = scan_plane_reduce(c, nullptr, false, true, true, i, seq);
    if (rc) return rc;
  }
  if (with_map) {
    int rc = scan_classify_insert(c, nullptr, flg_ekf_inited, true);
    if (rc) return rc;
  }
  const volatile unsigned long long* pseq = &io->seq;
  for (unsigned spin = 1; *pseq != seq; ++spin) {
    if ((spin & 0x3FFu) == 0) {
      const cudaError_t q = cudaStreamQuery(c->stream);
      if (q == cudaSuccess) break;
      if (q != cudaErrorNotReady) LIO_CUDA(q);
    }
#if defined(__x86_64__)
    __builtin_ia32_pause();
#endif
  }
  if (*pseq != seq) {
    set_error("device IEKF loop did not complete (seq %llu, expected %llu)", (unsigned long long)*pseq, seq);
    return LIO_E_CUDA;
  }
  const int done = io->iters_done;
  if (iters_done) *iters_done = done;
  if (iter_states) std::memcpy(iter_states, io->iter_states, sizeof(double) * 36 * (size_t)done);
  if (n_eff_trace) std::memcpy(n_eff_trace, io->n_eff, sizeof(int) * (size_t)done);
  if (knn_trace) std::memcpy(knn_trace, io->knn, sizeof(int) * (size_t)done);
  if (io->rc) {
    set_error("lio_iekf_update: covariance / information matrix is singular");
    return io->rc;
  }
  std::memcpy(state, io->state_out, sizeof(io->state_out));
  std::memcpy(P_io, io->P_out, sizeof(io->P_out));
  return LIO_OK;
}

}  // namespace lio

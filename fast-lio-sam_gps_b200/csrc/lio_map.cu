// GPU-resident map: voxel-hashed point store with ikd-Tree's observable semantics
// (Build, Add_Points with/without downsample, Delete_Point_Boxes, size, flatten) — kernel K2 of SURVEY.md §2.2.
// Reference behaviour: third_party/FAST_LIO/include/ikd-Tree/ikd_Tree.cpp (absent submodule).
//
// Downsample-on-insert is order dependent in the reference (points are processed one by one).  The
// kernels below reproduce the RESULT of that sequential process without a sort:
//   scan   : every batch point finds the best existing point of its filter_size_map voxel
//            (min (dist-to-centre, id)) and takes part in a per-voxel atomicMin arbitration among the
//            batch (min dist, ties -> later point, exactly what "new point wins ties" yields sequentially);
//   commit : the batch winner of a voxel replaces the existing points when its distance is <= theirs,
//            otherwise the voxel is collapsed to its best existing point.
#include <algorithm>
#include <cstring>
#include "lio_ctx.h"

namespace lio {

namespace {

__device__ __forceinline__ uint32_t batch_n(uint32_t n_max, const uint32_t* n_dev) {
  if (!n_dev) return n_max;
  uint32_t n = *n_dev;
  return n < n_max ? n : n_max;
}

// The batch size is only known on the device (map_incremental's lists), so the grid is sized for the largest batch and
// a small batch would land in the first few blocks.  Consecutive points go to different blocks instead: every SM
// gets its share of the per-point latency chains (probe, bucket reads, atomics).
__device__ __forceinline__ uint32_t spread_index() { return threadIdx.x * gridDim.x + blockIdx.x; }

// ---- pass 1: make sure the cell of every batch point exists -----------------------------------
__global__ void k_cells_ensure(MapView m, const float4* __restrict__ pts, uint32_t n_max, const uint32_t* n_dev,
                               uint32_t* __restrict__ ent) {
  const uint32_t n = batch_n(n_max, n_dev);
  for (uint32_t i = spread_index(); i < n; i += gridDim.x * blockDim.x) {
    const float4 p = pts[i];
    const int cx = vox_of(p.x, m.ds) >> m.cell_shift, cy = vox_of(p.y, m.ds) >> m.cell_shift,
              cz = vox_of(p.z, m.ds) >> m.cell_shift;
    const unsigned long long key = pack3(cx, cy, cz);
    uint32_t s = hash64(key) & m.mask;
    uint32_t found = kNone;
    bool pool_fail = false, table_fail = false;
    for (uint32_t probes = 0; probes <= m.mask;) {
      unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(&m.table[s].key);
      if (cur == kEmptyKey) cur = atomicCAS(&m.table[s].key, kEmptyKey, kPendingKey);
      if (cur == kPendingKey) continue;  // another thread is creating a cell here: its key (or "empty" again) appears shortly
      if (cur == kEmptyKey) {
        // this thread owns the entry until it publishes the key.  Occupied entries never exceed cell_limit (capacity / 2):
        // a read-only probe (table_find and the search kernels' own probe loops) always meets an empty entry and terminates.
        if (atomicAdd(&m.ctr->n_cells, 1u) >= m.cell_limit) {
          atomicSub(&m.ctr->n_cells, 1u);
          atomicExch(&m.table[s].key, kEmptyKey);
          table_fail = true;
          break;
        }
        uint32_t b = atomicAdd(&m.ctr->pool_top, 1u);
        if (b >= m.pool_cap) {
          // pool exhausted: the cell points at the null bucket (index pool_cap: always empty, never written), readers see an
          // empty cell, writers refuse (claim_slot_g), the sticky error reaches the host with the next result
          atomicOr(&m.ctr->error, kErrPoolFull);
          b = m.pool_cap;
          pool_fail = true;
        }
        m.table[s].head = b;
        atomicOr(&m.occ[occ_word_of(m, cx >> 2, cy >> 2, cz >> 2)], 1ull << occ_bit_of(cx, cy, cz));
        __threadfence();
        atomicExch(&m.table[s].key, key);  // publish
        found = pool_fail ? kNone : s;
        break;
      }
      if (cur == key) {
        found = s;
        break;
      }
      s = (s + 1) & m.mask;
      ++probes;
    }
    if (found == kNone && !pool_fail) atomicOr(&m.ctr->error, kErrHashFull);
    (void)table_fail;
    ent[i] = found;
  }
}

__device__ __forceinline__ void voxel_centre(int kx, int ky, int kz, float ds, float& mx, float& my, float& mz) {
  const float minx = (float)kx * ds, miny = (float)ky * ds, minz = (float)kz * ds;
  const float maxx = minx + ds, maxy = miny + ds, maxz = minz + ds;
  mx = minx + (maxx - minx) / 2.0f;
  my = miny + (maxy - miny) / 2.0f;
  mz = minz + (maxz - minz) / 2.0f;
}

// ---- 8-lane group forms: lane j of a group looks at slot j of a bucket, so a bucket is ONE coalesced 128-byte access
// and the per-point chains of dependent slot reads of the thread-per-point kernels above become a ballot -----------
struct Grp {
  int lane, gl;
  unsigned mask;
  __device__ Grp() {
    lane = threadIdx.x & 31;
    gl = lane & 7;
    mask = 0xFFu << (lane & 24);
  }
  __device__ __forceinline__ unsigned ballot(bool p) const { return (__ballot_sync(mask, p) >> (lane & 24)) & 0xFFu; }
  template <typename T>
  __device__ __forceinline__ T bcast(T v, int src) const { return __shfl_sync(mask, v, src, 8); }
};
// group index spread over the blocks like spread_index(): consecutive batch points go to different blocks
__device__ __forceinline__ uint32_t spread_group() { return (threadIdx.x >> 3) * gridDim.x + blockIdx.x; }
__device__ __forceinline__ uint32_t groups_total() { return gridDim.x * (blockDim.x >> 3); }

// claim a free slot in the chain of entry `ei` for point (x,y,z,id); every lane of the group returns the outcome
__device__ bool claim_slot_g(const MapView& m, const Grp& g, uint32_t ei, float x, float y, float z, uint32_t id) {
  HashEntry* e = m.table + ei;
  const uint32_t head = *reinterpret_cast<volatile uint32_t*>(&e->head);
  if (head >= m.pool_cap) return false;  // cell created while the pool was exhausted (null bucket): error already flagged
  uint32_t b = head, spare = kNone;
  while (true) {
    float4* s = m.slots + (size_t)b * kSlots;
    uint32_t* w = reinterpret_cast<uint32_t*>(&s[g.gl].w);
    unsigned empties = g.ballot(*reinterpret_cast<volatile uint32_t*>(w) == kEmptyId);
    while (empties) {
      const int j = __ffs(empties) - 1;
      int won = 0;
      if (g.gl == j) won = atomicCAS(w, kEmptyId, id) == kEmptyId;
      won = g.bcast(won, j);
      if (won) {
        if (g.gl == j) {
          s[j].x = x;
          s[j].y = y;
          s[j].z = z;
        }
        return true;
      }
      empties &= empties - 1;
    }
    // bucket full: follow the chain, linking a new bucket at its tail when needed (lane 0 does the bookkeeping)
    uint32_t nb = kNone;
    int fail = 0;
    if (g.gl == 0) {
      uint32_t* link = (b == head) ? &e->ovf : &m.next[b];
      nb = *reinterpret_cast<volatile uint32_t*>(link);
      if (nb == kNone) {
        if (spare == kNone) {
          spare = atomicAdd(&m.ctr->pool_top, 1u);
          if (spare >= m.pool_cap) {
            atomicOr(&m.ctr->error, kErrPoolFull);
            fail = 1;
          }
        }
        if (!fail) {
          const uint32_t prev = atomicCAS(link, kNone, spare);
          if (prev == kNone) {
            nb = spare;
            spare = kNone;
          } else {
            nb = prev;
          }
        }
      }
    }
    fail = g.bcast(fail, 0);
    if (fail) return false;
    b = g.bcast(nb, 0);
  }
}

// IDS_IN_W: the batch carries its ids in .w (map_compact re-inserting the live points under their old ids)
template <bool IDS_IN_W>
__global__ void __launch_bounds__(256) k_insert_plain_g(MapView m, const float4* __restrict__ pts, uint32_t n_max, const uint32_t* n_dev,
                                                        const uint32_t* __restrict__ ent) {
  const Grp g;
  const uint32_t n = batch_n(n_max, n_dev);
  const uint32_t base = m.ctr->next_id;
  for (uint32_t i = spread_group(); i < n; i += groups_total()) {
    const uint32_t ei = ent[i];
    if (ei == kNone) continue;
    const float4 p = pts[i];
    const uint32_t id = IDS_IN_W ? __float_as_uint(p.w) : base + i;
    if (claim_slot_g(m, g, ei, p.x, p.y, p.z, id) && g.gl == 0) atomicAdd(&m.ctr->live, 1);
  }
}

__global__ void __launch_bounds__(256) k_ds_scan_g(MapView m, const float4* __restrict__ pts, uint32_t n_max, const uint32_t* n_dev,
                                                   const uint32_t* __restrict__ ent, unsigned long long* vox_keys,
                                                   unsigned long long* vox_best, uint32_t vox_mask, uint32_t* __restrict__ tslot,
                                                   unsigned long long* __restrict__ ebest, uint32_t* __restrict__ ecnt) {
  const Grp g;
  const uint32_t n = batch_n(n_max, n_dev);
  for (uint32_t i = spread_group(); i < n; i += groups_total()) {
    const uint32_t ei = ent[i];
    if (ei == kNone) {
      if (g.gl == 0) tslot[i] = kNone;
      continue;
    }
    const float4 p = pts[i];
    const int kx = vox_of(p.x, m.ds), ky = vox_of(p.y, m.ds), kz = vox_of(p.z, m.ds);
    float mx, my, mz;
    voxel_centre(kx, ky, kz, m.ds, mx, my, mz);
    unsigned long long best = ~0ull;
    uint32_t cnt = 0;
    const HashEntry e = m.table[ei];
    for (uint32_t b = e.head; b != kNone; b = chain_next(m, e.head, e.ovf, b)) {
      const float4 q = m.slots[(size_t)b * kSlots + g.gl];
      const uint32_t id = __float_as_uint(q.w);
      const bool hit = id != kEmptyId && vox_of(q.x, m.ds) == kx && vox_of(q.y, m.ds) == ky && vox_of(q.z, m.ds) == kz;
      cnt += __popc(g.ballot(hit));
      if (hit) {
        const unsigned long long k = key64(sqdist3(q.x, q.y, q.z, mx, my, mz), id);
        if (k < best) best = k;
      }
    }
#pragma unroll
    for (int off = 4; off; off >>= 1) {
      const unsigned long long o = __shfl_xor_sync(g.mask, best, off, 8);
      if (o < best) best = o;
    }
    if (g.gl == 0) {
      ebest[i] = best;
      ecnt[i] = cnt;
      const unsigned long long vkey = pack3(kx, ky, kz);
      uint32_t s = hash64(vkey) & vox_mask;
      while (true) {
        unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(&vox_keys[s]);
        if (cur == kEmptyKey) cur = atomicCAS(&vox_keys[s], kEmptyKey, vkey);
        if (cur == kEmptyKey || cur == vkey) break;
        s = (s + 1) & vox_mask;
      }
      const float d = sqdist3(p.x, p.y, p.z, mx, my, mz);
      atomicMin(&vox_best[s], ((unsigned long long)__float_as_uint(d) << 32) | (0xFFFFFFFFu - i));
      tslot[i] = s;
    }
  }
}

// group form of erase_voxel
__device__ void erase_voxel_g(const MapView& m, const Grp& g, const HashEntry& e, int kx, int ky, int kz, uint32_t base_id, uint32_t keep) {
  for (uint32_t b = e.head; b != kNone; b = chain_next(m, e.head, e.ovf, b)) {
    float4* s = m.slots + (size_t)b * kSlots + g.gl;
    uint32_t* w = reinterpret_cast<uint32_t*>(&s->w);
    const uint32_t id = *reinterpret_cast<volatile uint32_t*>(w);
    if (id == kEmptyId || id >= base_id || id == keep) continue;
    const float x = *reinterpret_cast<volatile float*>(&s->x), y = *reinterpret_cast<volatile float*>(&s->y),
                z = *reinterpret_cast<volatile float*>(&s->z);
    if (vox_of(x, m.ds) != kx || vox_of(y, m.ds) != ky || vox_of(z, m.ds) != kz) continue;
    atomicCAS(w, id, kEmptyId);
  }
}

__global__ void __launch_bounds__(256) k_ds_commit_g(MapView m, const float4* __restrict__ pts, uint32_t n_max, const uint32_t* n_dev,
                                                     const uint32_t* __restrict__ ent, const unsigned long long* __restrict__ vox_best,
                                                     const uint32_t* __restrict__ tslot, const unsigned long long* __restrict__ ebest,
                                                     const uint32_t* __restrict__ ecnt) {
  const Grp g;
  const uint32_t n = batch_n(n_max, n_dev);
  const uint32_t base = m.ctr->next_id;
  for (uint32_t i = spread_group(); i < n; i += groups_total()) {
    const uint32_t ts = tslot[i];
    if (ts == kNone) continue;
    const float4 p = pts[i];
    const int kx = vox_of(p.x, m.ds), ky = vox_of(p.y, m.ds), kz = vox_of(p.z, m.ds);
    float mx, my, mz;
    voxel_centre(kx, ky, kz, m.ds, mx, my, mz);
    const uint32_t dbits = __float_as_uint(sqdist3(p.x, p.y, p.z, mx, my, mz));
    if (vox_best[ts] != (((unsigned long long)dbits << 32) | (0xFFFFFFFFu - i))) continue;  // not the voxel's winner
    const uint32_t ei = ent[i];
    const HashEntry e = m.table[ei];
    const unsigned long long eb = ebest[i];
    const uint32_t ec = ecnt[i];
    if (ec == 0 || dbits <= (uint32_t)(eb >> 32)) {
      if (ec) erase_voxel_g(m, g, e, kx, ky, kz, base, kEmptyId);
      __syncwarp(g.mask);
      if (claim_slot_g(m, g, ei, p.x, p.y, p.z, base + i) && g.gl == 0) {
        atomicAdd(&m.ctr->live, 1 - (int)ec);
        atomicAdd(&m.ctr->batch_alive, 1u);
      }
    } else if (ec > 1) {
      erase_voxel_g(m, g, e, kx, ky, kz, base, (uint32_t)eb);
      if (g.gl == 0) atomicAdd(&m.ctr->live, -(int)(ec - 1));
    }
  }
}

// last kernel of every add: advances the id counter and mirrors the sticky error word (plus the "ids running out" warning)
// into mapped host memory, where the next API call sees it without a device read
__global__ void k_advance_ids(MapCounters* ctr, uint32_t n_max, const uint32_t* n_dev, uint32_t* __restrict__ err_host) {
  ctr->next_id += batch_n(n_max, n_dev);
  const uint32_t w = ctr->error | (ctr->next_id > kIdRenumberAt ? kWarnIdsHigh : 0u);
  if (w) *err_host = w;
}
__global__ void k_set_next_id(MapCounters* ctr, uint32_t v) { ctr->next_id = v; }
__global__ void k_reset_batch(MapCounters* ctr) { ctr->batch_alive = 0; }

// ---- Delete_Point_Boxes --------------------------------------------------------------------------
__global__ void k_delete_boxes(MapView m, const float* __restrict__ boxes, int nb) {
  const uint32_t top = min(m.ctr->pool_top, m.pool_cap);
  const size_t total = (size_t)top * kSlots;
  uint32_t removed = 0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const float4 q = m.slots[i];
    if (__float_as_uint(q.w) == kEmptyId) continue;
    bool hit = false;
    for (int b = 0; b < nb && !hit; ++b) {
      const float* bx = boxes + 6 * b;
      hit = q.x >= bx[0] && q.x < bx[3] && q.y >= bx[1] && q.y < bx[4] && q.z >= bx[2] && q.z < bx[5];
    }
    if (hit) {
      reinterpret_cast<uint32_t*>(&m.slots[i].w)[0] = kEmptyId;
      ++removed;
    }
  }
  if (removed) {
    atomicAdd(&m.ctr->deleted, removed);
    atomicAdd(&m.ctr->live, -(int)removed);
  }
}
__global__ void k_reset_deleted(MapCounters* ctr) { ctr->deleted = 0; }

// ---- flatten ---------------------------------------------------------------------------------------
__global__ void k_export(MapView m, float4* __restrict__ out, uint32_t cap, uint32_t* cursor) {
  const uint32_t top = min(m.ctr->pool_top, m.pool_cap);
  const size_t total = (size_t)top * kSlots;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const float4 q = m.slots[i];
    if (__float_as_uint(q.w) == kEmptyId) continue;
    const uint32_t pos = atomicAdd(cursor, 1u);
    if (pos < cap) out[pos] = q;
  }
}

uint32_t next_pow2(uint64_t v) {
  uint64_t p = 1;
  while (p < v) p <<= 1;
  return (uint32_t)std::min<uint64_t>(p, 1ull << 31);
}

int grid_for(const lio_ctx* c, uint32_t n, int block) {
  const int want = (int)((n + block - 1) / block);
  return std::max(1, std::min(want, c->n_sms * 16));
}

}  // namespace

int map_alloc(lio_ctx* c) {
  const lio_config& cfg = c->cfg;
  MapView& m = c->map;
  m.ds = cfg.filter_size_map;
  int mult = cfg.cell_mult > 0 ? cfg.cell_mult : 2;
  int shift = 0;
  while ((1 << shift) < mult) ++shift;
  if ((1 << shift) != mult) {
    set_error("cell_mult must be a power of two (got %d)", mult);
    return LIO_E_ARG;
  }
  m.cell_shift = shift;
  m.cell_size = m.ds * (float)mult;
  const uint64_t cells = cfg.max_cells ? cfg.max_cells : cfg.max_points;
  c->table_cap = next_pow2(std::max<uint64_t>(1024, cells * 2));
  m.mask = c->table_cap - 1;
  // one head bucket per cell plus overflow room for cells holding more than kSlots points
  m.pool_cap = (uint32_t)std::min<uint64_t>(cells + cfg.max_points / kSlots + 1024, 0x7FFFFFFFull);
  LIO_CUDA(cudaMalloc(&m.table, sizeof(HashEntry) * (size_t)c->table_cap));
  m.cell_limit = c->table_cap / 2;
  // + the null bucket (index pool_cap): what a cell created under pool exhaustion points at; always empty
  LIO_CUDA(cudaMalloc(&m.slots, sizeof(float4) * kSlots * ((size_t)m.pool_cap + 1)));
  LIO_CUDA(cudaMalloc(&m.next, sizeof(uint32_t) * ((size_t)m.pool_cap + 1)));
  // occupancy words: >= one per 2 cells of capacity (a 4x4x4 block usually holds several cells, so most words stay private)
  const uint64_t occ_words = next_pow2(std::max<uint64_t>(1u << 14, cells / 2));
  m.occ_mask = (uint32_t)(occ_words - 1);
  LIO_CUDA(cudaMalloc(&m.occ, sizeof(unsigned long long) * (size_t)occ_words));
  LIO_CUDA(cudaMalloc(&m.ctr, sizeof(MapCounters)));
  LIO_CUDA(cudaMalloc(&c->boxes_dev, sizeof(float) * 6 * 64));
  LIO_CUDA(cudaHostAlloc(&c->map_err_host, 64, cudaHostAllocMapped));
  std::memset(c->map_err_host, 0, 64);
  LIO_CUDA(cudaHostGetDevicePointer(&c->map_err_dev, c->map_err_host, 0));
  {
    const char* e = getenv("LIO_MAP_COMPACT");
    c->auto_compact = !(e && std::strcmp(e, "0") == 0);
  }
  LIO_CUDA(cudaMemsetAsync(m.ctr, 0, sizeof(MapCounters), c->stream));
  return map_clear(c);
}

void map_free(lio_ctx* c) {
  MapView& m = c->map;
  cudaFree(m.table);
  cudaFree(m.slots);
  cudaFree(m.next);
  cudaFree(m.occ);
  cudaFree(m.ctr);
  cudaFree(c->boxes_dev);
  if (c->map_err_host) cudaFreeHost(c->map_err_host);
  c->map_err_host = c->map_err_dev = nullptr;
  cudaFree(c->batch_pts);
  cudaFree(c->batch_ent);
  cudaFree(c->batch_tslot);
  cudaFree(c->batch_ebest);
  cudaFree(c->batch_ecnt);
  cudaFree(c->vox_keys);
  cudaFree(c->vox_best);
  m = MapView{};
}

// empties the store; the id counter keeps running (the oracle's Build does the same)
int map_clear(lio_ctx* c) {
  MapView& m = c->map;
  LIO_CUDA(cudaMemsetAsync(m.table, 0xFF, sizeof(HashEntry) * (size_t)c->table_cap, c->stream));
  LIO_CUDA(cudaMemsetAsync(m.slots, 0xFF, sizeof(float4) * kSlots * ((size_t)m.pool_cap + 1), c->stream));
  LIO_CUDA(cudaMemsetAsync(m.next, 0xFF, sizeof(uint32_t) * ((size_t)m.pool_cap + 1), c->stream));
  LIO_CUDA(cudaMemsetAsync(m.occ, 0, sizeof(unsigned long long) * ((size_t)m.occ_mask + 1), c->stream));
  // keep next_id and the sticky error word, zero the rest
  LIO_CUDA(cudaMemsetAsync(&m.ctr->pool_top, 0, sizeof(uint32_t) * 3, c->stream));
  return LIO_OK;
}

int map_ensure_batch(lio_ctx* c, uint32_t n) {
  if (n <= c->batch_cap) return LIO_OK;
  for (auto& g : c->map_graphs) cudaGraphExecDestroy((cudaGraphExec_t)g.second);  // they captured the old scratch buffers
  c->map_graphs.clear();
  cudaFree(c->batch_pts);
  cudaFree(c->batch_ent);
  cudaFree(c->batch_tslot);
  cudaFree(c->batch_ebest);
  cudaFree(c->batch_ecnt);
  cudaFree(c->vox_keys);
  cudaFree(c->vox_best);
  c->batch_cap = 0;
  const uint32_t cap = std::max<uint32_t>(n, 1024);
  LIO_CUDA(cudaMalloc(&c->batch_pts, sizeof(float4) * (size_t)cap));
  LIO_CUDA(cudaMalloc(&c->batch_ent, sizeof(uint32_t) * (size_t)cap));
  LIO_CUDA(cudaMalloc(&c->batch_tslot, sizeof(uint32_t) * (size_t)cap));
  LIO_CUDA(cudaMalloc(&c->batch_ebest, sizeof(unsigned long long) * (size_t)cap));
  LIO_CUDA(cudaMalloc(&c->batch_ecnt, sizeof(uint32_t) * (size_t)cap));
  c->vox_cap = next_pow2((uint64_t)cap * 2);
  LIO_CUDA(cudaMalloc(&c->vox_keys, sizeof(unsigned long long) * (size_t)c->vox_cap));
  LIO_CUDA(cudaMalloc(&c->vox_best, sizeof(unsigned long long) * (size_t)c->vox_cap));
  c->batch_cap = cap;
  return LIO_OK;
}

// Add_Points on a device-resident batch.  n_dev (optional) holds the actual count (<= n_max).
int map_add_device(lio_ctx* c, const float4* pts, uint32_t n_max, const uint32_t* n_dev, bool downsample) {
  if (n_max == 0) return LIO_OK;
  int rc = map_ensure_batch(c, n_max);
  if (rc) return rc;
  const int block = 256, grid = grid_for(c, n_max, block);
  const int ggrid = grid_for(c, (uint32_t)std::min<uint64_t>((uint64_t)n_max * 8, 0x7FFFFFFFull), block);  // 8 lanes per point
  MapView& m = c->map;
  timer_begin(c, T_INSERT);
  k_cells_ensure<<<grid, block, 0, c->stream>>>(m, pts, n_max, n_dev, c->batch_ent);
  if (!downsample) {
    k_insert_plain_g<false><<<ggrid, block, 0, c->stream>>>(m, pts, n_max, n_dev, c->batch_ent);
    c->launches += 2;
  } else {
    // the arbitration table only needs to cover this batch
    const uint32_t vcap = std::min<uint32_t>(c->vox_cap, next_pow2((uint64_t)std::max<uint32_t>(n_max, 512) * 2));
    cudaMemsetAsync(c->vox_keys, 0xFF, sizeof(unsigned long long) * (size_t)vcap, c->stream);
    cudaMemsetAsync(c->vox_best, 0xFF, sizeof(unsigned long long) * (size_t)vcap, c->stream);
    k_reset_batch<<<1, 1, 0, c->stream>>>(m.ctr);
    k_ds_scan_g<<<ggrid, block, 0, c->stream>>>(m, pts, n_max, n_dev, c->batch_ent, c->vox_keys, c->vox_best, vcap - 1,
                                                c->batch_tslot, c->batch_ebest, c->batch_ecnt);
    k_ds_commit_g<<<ggrid, block, 0, c->stream>>>(m, pts, n_max, n_dev, c->batch_ent, c->vox_best, c->batch_tslot,
                                                  c->batch_ebest, c->batch_ecnt);
    c->launches += 4;
  }
  k_advance_ids<<<1, 1, 0, c->stream>>>(m.ctr, n_max, n_dev, c->map_err_dev);
  c->launches += 1;
  timer_end(c);
  LIO_CUDA(cudaGetLastError());
  return LIO_OK;
}

int map_delete_boxes(lio_ctx* c, const float* boxes_host, int nb) {
  if (nb <= 0) return LIO_OK;
  if (nb > 64) {
    set_error("at most 64 boxes per call (got %d)", nb);
    return LIO_E_ARG;
  }
  LIO_CUDA(cudaMemcpyAsync(c->boxes_dev, boxes_host, sizeof(float) * 6 * nb, cudaMemcpyHostToDevice, c->stream));
  k_reset_deleted<<<1, 1, 0, c->stream>>>(c->map.ctr);
  k_delete_boxes<<<c->n_sms * 8, 256, 0, c->stream>>>(c->map, c->boxes_dev, nb);
  c->launches += 2;
  LIO_CUDA(cudaGetLastError());
  // the host copy of `boxes_host` may go away once we return; the counters decide whether the store is worth compacting
  MapCounters ctr;
  LIO_CUDA(cudaMemcpyAsync(&ctr, c->map.ctr, sizeof(MapCounters), cudaMemcpyDeviceToHost, c->stream));
  LIO_CUDA(cudaStreamSynchronize(c->stream));
  c->last_deleted = ctr.deleted;
  c->dead_since_compact += ctr.deleted;
  // a sliding local map deletes as much as it inserts: without reclaiming, pool_top / n_cells only grow.  Compact when one of
  // the pools is three quarters used and a good part of what it holds is dead.
  const bool crowded = (uint64_t)ctr.pool_top * 4 > (uint64_t)c->map.pool_cap * 3 || (uint64_t)ctr.n_cells * 4 > (uint64_t)c->map.cell_limit * 3;
  if (c->auto_compact && crowded && c->dead_since_compact * 4 > (uint64_t)std::max(ctr.live, 1)) return map_compact(c, false);
  return LIO_OK;
}

// Rebuilds the store from its live points: dead cells, emptied buckets and overflow chains left behind by
// Delete_Point_Boxes / downsample replacements are reclaimed (table, pool and occupancy words start over).  Ids are kept, so
// every search and flatten() returns what it returned before; `renumber` replaces them by their rank (same order, same
// tie-breaks) and restarts the insertion counter at the live count - what keeps the 32-bit id from wrapping on long drives.
int map_compact(lio_ctx* c, bool renumber) {
  std::vector<float4> pts;
  int rc = map_export(c, pts);  // sorted by id
  if (rc) return rc;
  rc = map_clear(c);
  if (rc) return rc;
  const uint32_t n = (uint32_t)pts.size();
  if (renumber)
    for (uint32_t i = 0; i < n; ++i) std::memcpy(&pts[i].w, &i, 4);
  if (n) {
    rc = map_ensure_batch(c, n);
    if (rc) return rc;
    LIO_CUDA(cudaMemcpyAsync(c->batch_pts, pts.data(), sizeof(float4) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    const int block = 256, grid = grid_for(c, n, block);
    const int ggrid = grid_for(c, (uint32_t)std::min<uint64_t>((uint64_t)n * 8, 0x7FFFFFFFull), block);
    k_cells_ensure<<<grid, block, 0, c->stream>>>(c->map, c->batch_pts, n, nullptr, c->batch_ent);
    k_insert_plain_g<true><<<ggrid, block, 0, c->stream>>>(c->map, c->batch_pts, n, nullptr, c->batch_ent);
    c->launches += 2;
  }
  if (renumber) {
    k_set_next_id<<<1, 1, 0, c->stream>>>(c->map.ctr, n);
    c->launches += 1;
  }
  LIO_CUDA(cudaGetLastError());
  LIO_CUDA(cudaStreamSynchronize(c->stream));  // `pts` goes away
  if (renumber) *c->map_err_host &= ~kWarnIdsHigh;
  c->dead_since_compact = 0;
  // (cached neighbours of the scan in flight keep their coordinates, which is all map_incremental reads; after a renumbering
  // the ids lio_scan_get reports for that one scan are the old ones)
  return LIO_OK;
}

int map_read_counters(lio_ctx* c, MapCounters* out) {
  LIO_CUDA(cudaMemcpyAsync(out, c->map.ctr, sizeof(MapCounters), cudaMemcpyDeviceToHost, c->stream));
  LIO_CUDA(cudaStreamSynchronize(c->stream));
  if (out->error) {
    set_error("map capacity exhausted (%s%s): raise lio_config.max_points / max_cells",
              (out->error & kErrHashFull) ? "hash table " : "", (out->error & kErrPoolFull) ? "bucket pool" : "");
    return LIO_E_CAPACITY;
  }
  return LIO_OK;
}

// Called at the top of the hot-path entry points: the mapped word carries what the last insert left behind.
int map_check_health(lio_ctx* c) {
  const uint32_t w = *reinterpret_cast<volatile uint32_t*>(c->map_err_host);
  if (w & (kErrHashFull | kErrPoolFull)) {
    set_error("map capacity exhausted (%s%s): points were dropped; raise lio_config.max_points / max_cells",
              (w & kErrHashFull) ? "hash table " : "", (w & kErrPoolFull) ? "bucket pool" : "");
    return LIO_E_CAPACITY;
  }
  if (w & kWarnIdsHigh) return map_compact(c, true);
  return LIO_OK;
}

int map_export(lio_ctx* c, std::vector<float4>& out) {
  MapCounters ctr;
  int rc = map_read_counters(c, &ctr);
  if (rc) return rc;
  const uint32_t live = ctr.live > 0 ? (uint32_t)ctr.live : 0;
  out.resize(live);
  if (!live) return LIO_OK;
  float4* dev = nullptr;
  uint32_t* cursor = nullptr;
  LIO_CUDA(cudaMalloc(&dev, sizeof(float4) * (size_t)live));
  LIO_CUDA(cudaMalloc(&cursor, sizeof(uint32_t)));
  cudaMemsetAsync(cursor, 0, sizeof(uint32_t), c->stream);
  k_export<<<c->n_sms * 8, 256, 0, c->stream>>>(c->map, dev, live, cursor);
  c->launches += 1;
  cudaMemcpyAsync(out.data(), dev, sizeof(float4) * (size_t)live, cudaMemcpyDeviceToHost, c->stream);
  cudaError_t e = cudaStreamSynchronize(c->stream);
  cudaFree(dev);
  cudaFree(cursor);
  if (!cuda_ok(e, "map_export")) return LIO_E_CUDA;
  std::sort(out.begin(), out.end(), [](const float4& a, const float4& b) {
    uint32_t ia, ib;
    std::memcpy(&ia, &a.w, 4);
    std::memcpy(&ib, &b.w, 4);
    return ia < ib;
  });
  return LIO_OK;
}

}  // namespace lio

// Device-side data layout and inline helpers shared by the map, kNN and measurement kernels.
//
// The map replaces ikd-Tree (third_party/FAST_LIO/include/ikd-Tree — absent submodule)
// with a voxel-hashed point store:
//   * hash table (open addressing, linear probing) keyed by the packed coordinates of a CELL of
//     cell_mult x filter_size_map metres; an entry names the head bucket of the cell and, when the
//     cell overflowed once, its second bucket (deeper chains go through next[]);
//   * bucket pool: a bucket is 8 float4 slots = ONE 128-byte line, so an 8-lane group reads a whole
//     bucket with a single coalesced load; slot.w carries the point id, 0xFFFFFFFF = empty slot.
// Built with --fmad=false: every f32 expression below is evaluated in the same order as oracle/orc_*.{hpp,cpp}.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace lio {

constexpr uint32_t kEmptyId = 0xFFFFFFFFu;
constexpr uint32_t kNone = 0xFFFFFFFFu;
constexpr unsigned long long kEmptyKey = ~0ull;
constexpr unsigned long long kPendingKey = ~0ull - 1ull;  // entry being created (insert kernels only; real keys use 63 bits)
constexpr int kSlots = 8;  // points per bucket
constexpr int kK = 5;      // NUM_MATCH_POINTS

constexpr uint32_t kErrHashFull = 1u;
constexpr uint32_t kErrPoolFull = 2u;
constexpr uint32_t kWarnIdsHigh = 4u;              // host-visible only: the 32-bit insertion counter is three quarters used
constexpr uint32_t kIdRenumberAt = 0xC0000000u;  // -> the next API call compacts the store and renumbers the ids by rank

struct __align__(16) HashEntry {
  unsigned long long key;
  uint32_t head;
  uint32_t ovf;
};

// device-resident counters (one cache line)
struct MapCounters {
  uint32_t next_id;      // running insertion counter
  uint32_t pool_top;     // buckets handed out
  uint32_t n_cells;      // occupied hash entries
  int32_t live;          // live points
  uint32_t error;        // kErr* bits, sticky
  uint32_t batch_alive;  // points of the last add batch alive afterwards
  uint32_t deleted;      // points removed by the last delete_boxes
  uint32_t list_add;     // map_incremental: |PointToAdd|
  uint32_t list_nodown;  // map_incremental: |PointNoNeedDownsample|
  uint32_t ticket;       // last-block ticket of the reduction kernel
  uint32_t pad[6];
};

struct MapView {
  HashEntry* table;
  uint32_t mask;  // capacity - 1
  uint32_t cell_limit;  // occupied entries never exceed this (capacity / 2): read-only probes always terminate
  float4* slots;  // pool_cap * kSlots
  uint32_t* next; // pool_cap
  uint32_t pool_cap;  // buckets; bucket index pool_cap is the always-empty null bucket
  MapCounters* ctr;
  float ds;        // filter_size_map
  int cell_shift;  // log2(cell_mult)
  float cell_size; // ds * cell_mult
  // coarse occupancy: one 64-bit word per 4x4x4 block of cells, addressed by a hash of the block's coordinates.  Bit
  // (x&3) | (y&3)<<2 | (z&3)<<4 is set when the cell was ever created.  Blocks that collide share a word, so a set bit
  // means "may exist" (the hash table is still probed), a clear bit means "does not exist" - never the other way round.
  unsigned long long* occ;
  uint32_t occ_mask;  // words - 1
};

#ifdef __CUDACC__
__device__ __forceinline__ float sqdist3(float ax, float ay, float az, float bx, float by, float bz) {
  float dx = ax - bx, dy = ay - by, dz = az - bz;
  float xx = dx * dx, yy = dy * dy, zz = dz * dz;
  float s = xx + yy;
  return s + zz;
}

__device__ __forceinline__ int vox_of(float v, float ds) { return (int)floorf(v / ds); }

__device__ __forceinline__ unsigned long long pack3(int x, int y, int z) {
  const unsigned long long m = 0x1FFFFFull;
  return (((unsigned long long)(x + (1 << 20)) & m) << 42) | (((unsigned long long)(y + (1 << 20)) & m) << 21) |
         ((unsigned long long)(z + (1 << 20)) & m);
}

__device__ __forceinline__ uint32_t hash64(unsigned long long k) {
  k ^= k >> 33;
  k *= 0xff51afd7ed558ccdull;
  k ^= k >> 33;
  k *= 0xc4ceb9fe1a85ec53ull;
  k ^= k >> 33;
  return (uint32_t)k;
}

__device__ __forceinline__ uint32_t occ_word_of(const MapView& m, int bx, int by, int bz) {
  return hash64(pack3(bx, by, bz) ^ 0x9E3779B97F4A7C15ull) & m.occ_mask;
}
__device__ __forceinline__ int occ_bit_of(int cx, int cy, int cz) { return (cx & 3) | ((cy & 3) << 2) | ((cz & 3) << 4); }

// read-only lookup; returns the entry (key == kEmptyKey when the cell does not exist)
__device__ __forceinline__ HashEntry table_find(const MapView& m, unsigned long long key) {
  uint32_t s = hash64(key) & m.mask;
  while (true) {
    const uint4 raw = __ldg(reinterpret_cast<const uint4*>(m.table + s));
    HashEntry e;
    e.key = ((unsigned long long)raw.y << 32) | raw.x;
    e.head = raw.z;
    e.ovf = raw.w;
    if (e.key == key || e.key == kEmptyKey) return e;
    s = (s + 1) & m.mask;
  }
}

// next bucket of a chain: head -> entry.ovf -> next[ovf] -> ...
__device__ __forceinline__ uint32_t chain_next(const MapView& m, uint32_t head, uint32_t ovf, uint32_t cur) {
  if (cur == head) return ovf;
  return m.next[cur];
}

__device__ __forceinline__ unsigned long long key64(float d2, uint32_t id) {
  return ((unsigned long long)__float_as_uint(d2) << 32) | id;
}

#endif  // __CUDACC__

}  // namespace lio

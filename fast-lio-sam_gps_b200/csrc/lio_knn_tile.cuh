// Tile kNN: one warp serves 32 spatially sorted queries and stages the candidate voxels they share in shared memory.
// Replaces KD_TREE::Nearest_Search for a whole scan (ikd-Tree, absent submodule) with the same result
// as lio_knn.cuh's per-query search: the exact k<=5 nearest map points within max_d2, ascending (squared distance, id).
//
//   order   : queries are visited in space-filling-curve order (Hilbert by default) of their map cell (k_tile_keys + one radix sort per scan), so the 32
//             queries of a warp usually fall in one or two cells;
//   stage   : the warp probes every cell of the tile's bounding box grown by E cells (lane-parallel hash probes), then
//             reads the hit buckets four at a time (one 128-byte line per 8 lanes, float4 per lane) and compacts the
//             live points into shared memory;
//   scan    : every lane runs over the staged points (broadcast LDS.128) and keeps a sorted private top-5 of
//             (d2 bits, id) keys;
//   exact   : a lane is done when its 5th distance (max_d2 if it found fewer) is below its distance to the boundary
//             of the staged box; otherwise the tile is re-staged with E+1 (E = 1..3) for the lanes still pending;
//   leftover: queries of incoherent tiles (beyond the first kTileMaxSub sub-tiles) and of sub-tiles whose box or
//             candidate count overflows the staging area are appended to a list that the per-query search (one warp per query on the scan path) of
//             lio_knn.cuh (k_knn_left) finishes, spread over the whole GPU.
#pragma once
#include "lio_knn.cuh"

namespace lio {

constexpr int kTileCap = 256;       // staged points per warp (4 KB)
constexpr int kTileMaxCells = 1024;  // cells probed per staging pass
constexpr int kTileSpan = 1;        // a sub-tile spans at most 2*kTileSpan+1 cells per axis
constexpr int kTileMaxWork = 128;   // buckets per staging pass
constexpr int kTileMaxE = 3;
constexpr int kTileMaxSub = 2;      // sub-tiles served per tile; the remaining queries are deferred to the per-query kernel
constexpr int kStageUnroll = 4;     // bucket loads in flight per lane while staging
// run-time tuning knobs (defaults = the constants above; LIO_TILE_MAXSUB / LIO_TILE_MAXE at context creation)
__device__ __constant__ int c_tile_max_sub = kTileMaxSub;
__device__ __constant__ int c_tile_max_e = kTileMaxE;
__device__ __constant__ int c_tile_hilbert = 1;  // LIO_TILE_ORDER=morton|hilbert

struct __align__(16) TileSmem {
  float4 pts[kTileCap];
  uint32_t wl_b[kTileMaxWork];  // bucket work list
  uint32_t wl_n[kTileMaxWork];  // successor of each listed bucket (kNone: end of chain)
};

// Per-lane sorted top-5 of the tile scan: 32-bit distance keys with the staged index they came from.  Ties on the exact
// distance (rare) are ordered by point id, read back from the staged points, so the order equals TopList's (d2 bits, id).
struct Top32 {
  uint32_t b0, b1, b2, b3, b4;
  uint32_t r0, r1, r2, r3, r4;
  __device__ __forceinline__ void clear() {
    b0 = b1 = b2 = b3 = b4 = 0xFFFFFFFFu;
    r0 = r1 = r2 = r3 = r4 = kNone;
  }
  __device__ __forceinline__ int count() const {
    return (b0 != 0xFFFFFFFFu) + (b1 != 0xFFFFFFFFu) + (b2 != 0xFFFFFFFFu) + (b3 != 0xFFFFFFFFu) + (b4 != 0xFFFFFFFFu);
  }
  // lexicographic (bits, id) "new < entry t"
  __device__ __forceinline__ static bool less(uint32_t bits, uint32_t id, uint32_t bt, uint32_t rt, const float4* pts) {
    if (bits != bt) return bits < bt;
    return id < __float_as_uint(pts[rt].w);  // bt != sentinel here because bits is a finite distance
  }
  __device__ __forceinline__ void insert(uint32_t bits, uint32_t id, uint32_t j, const float4* pts) {
    if (bits > b4) return;
    bool c0, c1, c2, c3;
    if (bits == b0 || bits == b1 || bits == b2 || bits == b3 || bits == b4) {  // exact tie: order by id
      if (!less(bits, id, b4, r4, pts)) return;
      c0 = less(bits, id, b0, r0, pts), c1 = less(bits, id, b1, r1, pts), c2 = less(bits, id, b2, r2, pts), c3 = less(bits, id, b3, r3, pts);
    } else {
      c0 = bits < b0, c1 = bits < b1, c2 = bits < b2, c3 = bits < b3;
    }
    b4 = c3 ? b3 : bits;  r4 = c3 ? r3 : j;
    b3 = c2 ? b2 : (c3 ? bits : b3);  r3 = c2 ? r2 : (c3 ? j : r3);
    b2 = c1 ? b1 : (c2 ? bits : b2);  r2 = c1 ? r1 : (c2 ? j : r2);
    b1 = c0 ? b0 : (c1 ? bits : b1);  r1 = c0 ? r0 : (c1 ? j : r1);
    b0 = c0 ? bits : b0;              r0 = c0 ? j : r0;
  }
};

// 8 bits per axis Morton code of a cell (wraps every 256 cells: locality is all that matters, and 24 key bits are three
// radix passes instead of four)
constexpr int kTileKeyBits = 24;
__device__ __forceinline__ uint32_t spread8(uint32_t x) {
  x &= 0xffu;
  x = (x | (x << 8)) & 0x0000f00fu;
  x = (x | (x << 4)) & 0x000c30c3u;
  x = (x | (x << 2)) & 0x00249249u;
  return x;
}
// 24-bit Hilbert index of an 8-bit-per-axis cell (Skilling's axes-to-transpose form).  Consecutive index ranges of the
// Hilbert curve are spatially connected, so the 32 queries of a tile stay in neighbouring cells where a Morton range
// jumps across the cube at every octant boundary (tiles the search then has to split or defer).
__device__ __forceinline__ uint32_t hilbert24(uint32_t x, uint32_t y, uint32_t z) {
  uint32_t X[3] = {x & 0xffu, y & 0xffu, z & 0xffu};
  const uint32_t M = 1u << 7;
#pragma unroll
  for (uint32_t Q = M; Q > 1; Q >>= 1) {
    const uint32_t P = Q - 1;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      if (X[i] & Q) {
        X[0] ^= P;
      } else {
        const uint32_t t = (X[0] ^ X[i]) & P;
        X[0] ^= t;
        X[i] ^= t;
      }
    }
  }
  X[1] ^= X[0];
  X[2] ^= X[1];
  uint32_t t = 0;
#pragma unroll
  for (uint32_t Q = M; Q > 1; Q >>= 1)
    if (X[2] & Q) t ^= Q - 1;
  X[0] ^= t; X[1] ^= t; X[2] ^= t;
  return (spread8(X[0]) << 2) | (spread8(X[1]) << 1) | spread8(X[2]);
}
__device__ __forceinline__ uint32_t tile_key(const MapView& m, float x, float y, float z) {
  const int cx = vox_of(x, m.ds) >> m.cell_shift, cy = vox_of(y, m.ds) >> m.cell_shift, cz = vox_of(z, m.ds) >> m.cell_shift;
  if (c_tile_hilbert) return hilbert24((uint32_t)cx, (uint32_t)cy, (uint32_t)cz);
  return spread8((uint32_t)cx) | (spread8((uint32_t)cy) << 1) | (spread8((uint32_t)cz) << 2);
}

// Stages every live point of cells [lo, hi] (inclusive cell coordinates).  Returns the number of staged points, or -1
// when the work list or the staging area overflowed.  Warp-uniform.
__device__ __forceinline__ int tile_stage(const MapView& m, TileSmem& sm, int lane, int lox, int loy, int loz, int dx, int dy, int dz) {
  const unsigned lt = (1u << lane) - 1u;
  const int total = dx * dy * dz;
  int nb = 0;
  for (int c0 = 0; c0 < total; c0 += 32) {
    const int idx = c0 + lane;
    bool hit = false;
    HashEntry e;
    e.head = e.ovf = kNone;
    if (idx < total) {
      const int iz = idx / (dx * dy), r = idx - iz * dx * dy, iy = r / dx, ix = r - iy * dx;
      const unsigned long long key = pack3(lox + ix, loy + iy, loz + iz);
      e = table_find(m, key);
      hit = e.key == key;
    }
    const unsigned hb = __ballot_sync(0xFFFFFFFFu, hit);
    const int pos = nb + __popc(hb & lt);
    if (hit && pos < kTileMaxWork) {
      sm.wl_b[pos] = e.head;
      sm.wl_n[pos] = e.ovf;
    }
    nb += __popc(hb);
  }
  if (nb > kTileMaxWork) return -1;
  __syncwarp();
  int cnt = 0, begin = 0, end = nb;
  while (begin < end) {
    int new_end = end;
    for (int b0 = begin; b0 < end; b0 += 4 * kStageUnroll) {
      // issue every load of this round before any is consumed
      float4 p[kStageUnroll];
      uint32_t succ[kStageUnroll];
#pragma unroll
      for (int u = 0; u < kStageUnroll; ++u) {
        const int wi = b0 + 4 * u + (lane >> 3);
        p[u] = make_float4(0.f, 0.f, 0.f, __uint_as_float(kEmptyId));
        succ[u] = kNone;
        if (wi < end) {
          p[u] = __ldg(m.slots + (size_t)sm.wl_b[wi] * kSlots + (lane & 7));
          if ((lane & 7) == 0) succ[u] = sm.wl_n[wi];
        }
      }
#pragma unroll
      for (int u = 0; u < kStageUnroll; ++u) {
        const bool live = __float_as_uint(p[u].w) != kEmptyId;
        const unsigned lb = __ballot_sync(0xFFFFFFFFu, live);
        const int pos = cnt + __popc(lb & lt);
        if (live && pos < kTileCap) sm.pts[pos] = p[u];
        cnt += __popc(lb);
        const bool f = succ[u] != kNone;
        const unsigned fb = __ballot_sync(0xFFFFFFFFu, f);
        if (fb) {
          const int fpos = new_end + __popc(fb & lt);
          if (f && fpos < kTileMaxWork) {
            sm.wl_b[fpos] = succ[u];
            sm.wl_n[fpos] = __ldg(m.next + succ[u]);
          }
          new_end += __popc(fb);
        }
      }
    }
    if (new_end > kTileMaxWork || cnt > kTileCap) return -1;
    __syncwarp();
    begin = end;
    end = new_end;
  }
  __syncwarp();
  return cnt;
}

// Searches one tile.  Per lane: (qx,qy,qz) and `valid`.  On return, lanes whose bit is set in the result mask still
// need the per-query fallback; every other valid lane has had `emit(L, cnt)` called while its staged points were live.
// A tile whose queries are spread out (a jump in the curve order, or a sparse region) is served as up to kTileMaxSub sub-tiles: the lanes
// within kTileSpan cells of the first pending lane go first; what is left is returned as not served (serialising many
// one-query sub-tiles in one warp is a chain of dependent memory latencies: the per-query kernel spreads them out).
struct TileProf {  // warp-uniform counters of one tile (debug / profiling hook: lio_tile_profile)
  int passes, staged, cells;
};
template <typename Emit>
__device__ __forceinline__ unsigned tile_search(const MapView& m, TileSmem& sm, int lane, float qx, float qy, float qz, bool valid,
                                                float max_d2, Emit&& emit, unsigned* overflow_mask = nullptr, TileProf* prof = nullptr) {
  const float h = m.cell_size;
  const int cx = vox_of(qx, m.ds) >> m.cell_shift, cy = vox_of(qy, m.ds) >> m.cell_shift, cz = vox_of(qz, m.ds) >> m.cell_shift;
  const int big = 0x3fffffff;
  unsigned pending = __ballot_sync(0xFFFFFFFFu, valid);
  unsigned failed = 0;
  for (int round = 0; round < c_tile_max_sub && pending; ++round) {
    // members of this sub-tile
    const int a = __ffs(pending) - 1;
    const int ax0 = __shfl_sync(0xFFFFFFFFu, cx, a), ay0 = __shfl_sync(0xFFFFFFFFu, cy, a), az0 = __shfl_sync(0xFFFFFFFFu, cz, a);
    const bool near = ((pending >> lane) & 1u) && abs(cx - ax0) <= kTileSpan && abs(cy - ay0) <= kTileSpan && abs(cz - az0) <= kTileSpan;
    unsigned sub = __ballot_sync(0xFFFFFFFFu, near);
    pending &= ~sub;
    int E = 1;
    while (sub) {
      const bool mine = (sub >> lane) & 1u;
      const int lox = __reduce_min_sync(0xFFFFFFFFu, mine ? cx : big) - E, hix = __reduce_max_sync(0xFFFFFFFFu, mine ? cx : -big) + E;
      const int loy = __reduce_min_sync(0xFFFFFFFFu, mine ? cy : big) - E, hiy = __reduce_max_sync(0xFFFFFFFFu, mine ? cy : -big) + E;
      const int loz = __reduce_min_sync(0xFFFFFFFFu, mine ? cz : big) - E, hiz = __reduce_max_sync(0xFFFFFFFFu, mine ? cz : -big) + E;
      const int dx = hix - lox + 1, dy = hiy - loy + 1, dz = hiz - loz + 1;
      int n_st = -1;
      if (E <= c_tile_max_e && dx * dy * dz <= kTileMaxCells) n_st = tile_stage(m, sm, lane, lox, loy, loz, dx, dy, dz);
      if (prof) {
        prof->passes += 1;
        prof->staged += max(n_st, 0);
        prof->cells += dx * dy * dz;
      }
      if (n_st < 0) {  // too many cells / candidates: these queries take the per-query path
        failed |= sub;
        break;
      }
      bool done = false;
      int need = 0;
      if (mine) {
        Top32 L;
        L.clear();
        const uint32_t max_bits = __float_as_uint(max_d2);
        int j = 0;
        for (; j + 4 <= n_st; j += 4) {
          const float4 c0 = sm.pts[j], c1 = sm.pts[j + 1], c2 = sm.pts[j + 2], c3 = sm.pts[j + 3];
          const uint32_t d0 = __float_as_uint(sqdist3(qx, qy, qz, c0.x, c0.y, c0.z)), d1 = __float_as_uint(sqdist3(qx, qy, qz, c1.x, c1.y, c1.z));
          const uint32_t d2 = __float_as_uint(sqdist3(qx, qy, qz, c2.x, c2.y, c2.z)), d3 = __float_as_uint(sqdist3(qx, qy, qz, c3.x, c3.y, c3.z));
          if (d0 <= max_bits) L.insert(d0, __float_as_uint(c0.w), j, sm.pts);
          if (d1 <= max_bits) L.insert(d1, __float_as_uint(c1.w), j + 1, sm.pts);
          if (d2 <= max_bits) L.insert(d2, __float_as_uint(c2.w), j + 2, sm.pts);
          if (d3 <= max_bits) L.insert(d3, __float_as_uint(c3.w), j + 3, sm.pts);
        }
        for (; j < n_st; ++j) {
          const float4 c0 = sm.pts[j];
          const uint32_t d0 = __float_as_uint(sqdist3(qx, qy, qz, c0.x, c0.y, c0.z));
          if (d0 <= max_bits) L.insert(d0, __float_as_uint(c0.w), j, sm.pts);
        }
        const int cnt = L.count();
        // everything closer than `reach` to the query lies inside the staged box
        const float rx = fminf(qx - (float)lox * h, (float)(hix + 1) * h - qx);
        const float ry = fminf(qy - (float)loy * h, (float)(hiy + 1) * h - qy);
        const float rz = fminf(qz - (float)loz * h, (float)(hiz + 1) * h - qz);
        const float reach = fmaxf(fminf(fminf(rx, ry), rz), 0.f) * 0.9999f;
        const float tau = (cnt == kK) ? __uint_as_float(L.b4) : max_d2;
        done = tau < reach * reach;
        if (done) emit(L, cnt);
        else need = (int)(sqrtf(tau) / (h * 0.9999f)) + 1;  // growth that is certain to cover sqrt(tau)
      }
      sub &= ~__ballot_sync(0xFFFFFFFFu, done);
      const int want = __reduce_max_sync(0xFFFFFFFFu, need);
      E = max(E + 1, min(want, c_tile_max_e));
      __syncwarp();  // staged points are overwritten by the next pass
    }
  }
  if (overflow_mask) *overflow_mask = failed;  // lanes deferred because the staging area / growth limit was exceeded
  return failed | pending;
}

}  // namespace lio

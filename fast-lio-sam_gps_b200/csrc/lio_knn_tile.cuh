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

constexpr int kTileCap = 256;       // staged points per warp (4 KB); the fast scan keeps the staged index in 8 key bits
static_assert(kTileCap <= 256, "Top6T carries the staged index in the low 8 bits of its keys");
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
__device__ __constant__ int c_tile_offer = 1;    // LIO_TILE_OFFER=0: a tile serves all its sub-tiles itself (A/B of the work queue)
// LIO_TILE_GROUP=block: a sub-tile = the lanes in the same aligned 2x2x2 block of cells (box <= 4^3 cells, never overflows,
// but half again as many sub-tiles: 86 vs 62 us per search on B200); default "span": the lanes within kTileSpan cells of the
// sub-tile's first lane (box up to 5^3 cells)
__device__ __constant__ int c_tile_block = 0;
// LIO_TILE_HALVE=1: a sub-tile whose candidates overflow the staging area is served cell by cell inside the tile kernel
// (27 cells staged per pass); default 0: its queries go to the per-query kernel (60 vs 62 us per search: the dense corners
// that overflow are the slowest tasks of the launch, and serving them in the same warp lengthens its tail)
__device__ __constant__ int c_tile_halve = 0;
// LIO_TILE_SPLIT=n (0..8, default 8): a staging pass with at most n lanes still pending (offered sub-tile tasks, second passes
// with a grown box) scans the staged points with FOUR lanes per query - lane 4g+p takes every 4th candidate for the g-th
// pending lane - and merges the four top-6 lists by two exchange rounds, instead of leaving 24+ lanes idle for the whole scan.
// Keys are unique (they carry the staged index), so the merged list is the list one lane would have built: same result.
__device__ __constant__ int c_tile_split = 8;

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

// Fast form of the tile scan: six sorted 32-bit keys = (d2 bits with the low 8 bits replaced by the staged index), kept by a
// branch-free min/max insertion network - every candidate costs the same 11 integer min/max whether it enters the list or not,
// where the exact list above pays a divergent ~25-instruction insert as soon as ONE lane of the warp has a hit (which, with 32
// queries per warp, is on almost every candidate).  Truncation drops the low 8 of the 23 mantissa bits of d2: when the six
// smallest truncated distances are pairwise different (and differ from the truncated search radius) the five smallest keys
// are exactly the five nearest points in exact (d2, id) order; otherwise the lane redoes its scan with the exact list.
struct Top6T {
  uint32_t k0, k1, k2, k3, k4, k5;
  __device__ __forceinline__ void clear() { k0 = k1 = k2 = k3 = k4 = k5 = 0xFFFFFFFFu; }
  __device__ __forceinline__ void insert(uint32_t x) {
    uint32_t t;
    t = min(k0, x); x = max(k0, x); k0 = t;
    t = min(k1, x); x = max(k1, x); k1 = t;
    t = min(k2, x); x = max(k2, x); k2 = t;
    t = min(k3, x); x = max(k3, x); k3 = t;
    t = min(k4, x); x = max(k4, x); k4 = t;
    k5 = min(k5, x);
  }
  // number of neighbours within max_d2 (0..5), or -1 when truncation leaves the answer ambiguous
  __device__ __forceinline__ int resolve(uint32_t max_bits) const {
    const uint32_t M = 0xFFFFFF00u, mt = max_bits & M;
    const uint32_t t0 = k0 & M, t1 = k1 & M, t2 = k2 & M, t3 = k3 & M, t4 = k4 & M, t5 = k5 & M;
    // an entry counts when its truncated distance is below the truncated radius; equal to it: cannot tell
    const int cnt = (t0 < mt) + (t1 < mt) + (t2 < mt) + (t3 < mt) + (t4 < mt);
    bool amb = (t0 == mt) | (t1 == mt) | (t2 == mt) | (t3 == mt) | (t4 == mt);
    // two of the counted entries (or the last counted and the next one) with the same truncated distance: order / membership unknown
    amb |= (t0 == t1) & (t0 <= mt);
    amb |= (t1 == t2) & (t1 <= mt);
    amb |= (t2 == t3) & (t2 <= mt);
    amb |= (t3 == t4) & (t3 <= mt);
    amb |= (t4 == t5) & (t4 <= mt);
    return amb ? -1 : cnt;
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
// when the work list or the staging area overflowed.  Warp-uniform.  The hash probes of two rounds of 32 cells are in
// flight together (a sub-tile's box is 27..64 cells: one dependent round trip instead of two).
__device__ __forceinline__ int tile_stage(const MapView& m, TileSmem& sm, int lane, int lox, int loy, int loz, int dx, int dy, int dz) {
  const unsigned lt = (1u << lane) - 1u;
  const int total = dx * dy * dz;
  const int dxy = dx * dy;
  int nb = 0;
  for (int c0 = 0; c0 < total; c0 += 64) {
    unsigned long long key[2];
    uint32_t slot[2];
    uint4 raw[2];
    bool want[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int idx = c0 + 32 * u + lane;
      want[u] = idx < total;
      const int iz = idx / dxy, r = idx - iz * dxy, iy = r / dx, ix = r - iy * dx;
      key[u] = pack3(lox + ix, loy + iy, loz + iz);
      slot[u] = hash64(key[u]) & m.mask;
      raw[u] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, kNone, kNone);
      if (want[u]) raw[u] = __ldg(reinterpret_cast<const uint4*>(m.table + slot[u]));
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      unsigned long long ek = ((unsigned long long)raw[u].y << 32) | raw[u].x;
      while (want[u] && ek != key[u] && ek != kEmptyKey) {  // linear probing past a collision
        slot[u] = (slot[u] + 1) & m.mask;
        raw[u] = __ldg(reinterpret_cast<const uint4*>(m.table + slot[u]));
        ek = ((unsigned long long)raw[u].y << 32) | raw[u].x;
      }
      const bool hit = want[u] && ek == key[u];
      const unsigned hb = __ballot_sync(0xFFFFFFFFu, hit);
      const int pos = nb + __popc(hb & lt);
      if (hit && pos < kTileMaxWork) {
        sm.wl_b[pos] = raw[u].z;
        sm.wl_n[pos] = raw[u].w;
      }
      nb += __popc(hb);
    }
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

struct TileProf {  // warp-uniform counters of one task (debug / profiling hook: lio_tile_profile)
  int passes, staged, cells;
  long long stage_cycles;  // SM cycles inside tile_stage (hash probes + bucket reads + compaction)
};

// The lanes of `sub` (a coherent sub-tile: every member within kTileSpan cells of its first lane) are searched together:
// stage the box of their cells grown by E, scan, and grow E for the lanes whose answer could still change.  `emit(L, cnt)`
// is called for every finished lane while its staged points are live.  Returns the lanes that could not be served because
// the staging area overflowed (the caller splits them or hands them to the per-query search).
template <typename Emit>
__device__ __forceinline__ unsigned tile_search_sub(const MapView& m, TileSmem& sm, int lane, float qx, float qy, float qz, int cx, int cy,
                                                    int cz, unsigned sub, float max_d2, Emit&& emit, TileProf* prof = nullptr) {
  const float h = m.cell_size;
  const int big = 0x3fffffff;
  int E = 1;
  while (sub) {
    const bool mine = (sub >> lane) & 1u;
    const int lox = __reduce_min_sync(0xFFFFFFFFu, mine ? cx : big) - E, hix = __reduce_max_sync(0xFFFFFFFFu, mine ? cx : -big) + E;
    const int loy = __reduce_min_sync(0xFFFFFFFFu, mine ? cy : big) - E, hiy = __reduce_max_sync(0xFFFFFFFFu, mine ? cy : -big) + E;
    const int loz = __reduce_min_sync(0xFFFFFFFFu, mine ? cz : big) - E, hiz = __reduce_max_sync(0xFFFFFFFFu, mine ? cz : -big) + E;
    const int dx = hix - lox + 1, dy = hiy - loy + 1, dz = hiz - loz + 1;
    int n_st = -1;
    const long long ts0 = prof ? clock64() : 0;
    if (E <= c_tile_max_e && dx * dy * dz <= kTileMaxCells) n_st = tile_stage(m, sm, lane, lox, loy, loz, dx, dy, dz);
    if (prof) {
      prof->stage_cycles += clock64() - ts0;
      prof->passes += 1;
      prof->staged += max(n_st, 0);
      prof->cells += dx * dy * dz;
    }
    if (n_st < 0) return sub;  // too many cells / candidates
    bool done = false;
    int need = 0;
    Top6T F;
    F.clear();
    const int act = __popc(sub);
    if (act <= c_tile_split) {  // warp-uniform: four lanes per pending query
      const int grp = lane >> 2, part = lane & 3;
      const bool gact = grp < act;
      unsigned rest = sub;
      for (int i = 0; i < grp && i < 7; ++i) rest &= rest - 1u;
      const int src = gact ? (__ffs(rest) - 1) : lane;  // the grp-th pending lane
      const float sx = __shfl_sync(0xFFFFFFFFu, qx, src), sy = __shfl_sync(0xFFFFFFFFu, qy, src), sz = __shfl_sync(0xFFFFFFFFu, qz, src);
      if (gact) {
        for (int j = part; j < n_st; j += 4) {
          const float4 c0 = sm.pts[j];
          F.insert((__float_as_uint(sqdist3(sx, sy, sz, c0.x, c0.y, c0.z)) & 0xFFFFFF00u) | (uint32_t)j);
        }
      }
#pragma unroll
      for (int mm = 1; mm <= 2; mm <<= 1) {
        const uint32_t o0 = __shfl_xor_sync(0xFFFFFFFFu, F.k0, mm), o1 = __shfl_xor_sync(0xFFFFFFFFu, F.k1, mm);
        const uint32_t o2 = __shfl_xor_sync(0xFFFFFFFFu, F.k2, mm), o3 = __shfl_xor_sync(0xFFFFFFFFu, F.k3, mm);
        const uint32_t o4 = __shfl_xor_sync(0xFFFFFFFFu, F.k4, mm), o5 = __shfl_xor_sync(0xFFFFFFFFu, F.k5, mm);
        F.insert(o0), F.insert(o1), F.insert(o2), F.insert(o3), F.insert(o4), F.insert(o5);
      }
      const int from = (__popc(sub & ((1u << lane) - 1u)) << 2) & 31;  // a pending lane reads the list of its group
      F.k0 = __shfl_sync(0xFFFFFFFFu, F.k0, from), F.k1 = __shfl_sync(0xFFFFFFFFu, F.k1, from), F.k2 = __shfl_sync(0xFFFFFFFFu, F.k2, from);
      F.k3 = __shfl_sync(0xFFFFFFFFu, F.k3, from), F.k4 = __shfl_sync(0xFFFFFFFFu, F.k4, from), F.k5 = __shfl_sync(0xFFFFFFFFu, F.k5, from);
    } else if (mine) {
      int j = 0;
      for (; j + 4 <= n_st; j += 4) {
        const float4 c0 = sm.pts[j], c1 = sm.pts[j + 1], c2 = sm.pts[j + 2], c3 = sm.pts[j + 3];
        const uint32_t d0 = __float_as_uint(sqdist3(qx, qy, qz, c0.x, c0.y, c0.z)), d1 = __float_as_uint(sqdist3(qx, qy, qz, c1.x, c1.y, c1.z));
        const uint32_t d2 = __float_as_uint(sqdist3(qx, qy, qz, c2.x, c2.y, c2.z)), d3 = __float_as_uint(sqdist3(qx, qy, qz, c3.x, c3.y, c3.z));
        F.insert((d0 & 0xFFFFFF00u) | (uint32_t)j);
        F.insert((d1 & 0xFFFFFF00u) | (uint32_t)(j + 1));
        F.insert((d2 & 0xFFFFFF00u) | (uint32_t)(j + 2));
        F.insert((d3 & 0xFFFFFF00u) | (uint32_t)(j + 3));
      }
      for (; j < n_st; ++j) {
        const float4 c0 = sm.pts[j];
        F.insert((__float_as_uint(sqdist3(qx, qy, qz, c0.x, c0.y, c0.z)) & 0xFFFFFF00u) | (uint32_t)j);
      }
    }
    if (mine) {
      const uint32_t max_bits = __float_as_uint(max_d2);
      uint32_t r0, r1, r2, r3, r4;  // staged indices of the answer, ascending (d2, id)
      int cnt = F.resolve(max_bits);
      r0 = F.k0 & 0xFFu, r1 = F.k1 & 0xFFu, r2 = F.k2 & 0xFFu, r3 = F.k3 & 0xFFu, r4 = F.k4 & 0xFFu;
      if (cnt < 0) {  // near-tie at the truncated precision (rare): exact list
        Top32 L;
        L.clear();
        for (int j = 0; j < n_st; ++j) {
          const float4 c0 = sm.pts[j];
          const uint32_t d0 = __float_as_uint(sqdist3(qx, qy, qz, c0.x, c0.y, c0.z));
          if (d0 <= max_bits) L.insert(d0, __float_as_uint(c0.w), j, sm.pts);
        }
        cnt = L.count();
        r0 = L.r0, r1 = L.r1, r2 = L.r2, r3 = L.r3, r4 = L.r4;
      }
      // everything closer than `reach` to the query lies inside the staged box
      const float rx = fminf(qx - (float)lox * h, (float)(hix + 1) * h - qx);
      const float ry = fminf(qy - (float)loy * h, (float)(hiy + 1) * h - qy);
      const float rz = fminf(qz - (float)loz * h, (float)(hiz + 1) * h - qz);
      const float reach = fmaxf(fminf(fminf(rx, ry), rz), 0.f) * 0.9999f;
      float tau = max_d2;
      if (cnt == kK) {
        const float4 c4 = sm.pts[r4];
        tau = sqdist3(qx, qy, qz, c4.x, c4.y, c4.z);
      }
      done = tau < reach * reach;
      if (done) {
        const uint32_t rs[kK] = {r0, r1, r2, r3, r4};
        emit(rs, cnt);
      } else {
        need = (int)(sqrtf(tau) / (h * 0.9999f)) + 1;  // growth that is certain to cover sqrt(tau)
      }
    }
    sub &= ~__ballot_sync(0xFFFFFFFFu, done);
    const int want = __reduce_max_sync(0xFFFFFFFFu, need);
    E = max(E + 1, min(want, c_tile_max_e));
    __syncwarp();  // staged points are overwritten by the next pass
  }
  return 0u;
}

// ---- work queue of one search (device memory; two sets are used alternately so that a set is re-zeroed a launch ahead) ----
// Tickets [0, tiles) are the tiles themselves (32 consecutive queries of the curve order).  A tile whose queries do not fit one
// sub-tile (a jump of the curve, or sparse far-range returns) keeps the sub-tile of its first query and OFFERS the others as
// (tile, lane mask) tasks before it touches the map; offered task j is taken (atomic exchange with 0) by the warp that draws
// ticket tiles + j - warps whose own tile was light, and the spare warps of the launch.  When the owner has finished its own
// sub-tile it takes back (compare-and-swap) whatever nobody took and serves it itself, with the queries it still holds in
// registers.  So an offer is never lost whatever a taker decides, no ordering between the offer and the counters is needed
// (no fence), and a warp that finds its slot empty may simply leave once every tile has made its offers (`split_done` ==
// tiles, a couple of microseconds into the launch).  Round 1 served these queries in a second, one-warp-per-query kernel:
// 27 us behind a 50 us launch that was two thirds idle.  Slots return to 0 by construction: the queue needs no clearing.
struct TileQueue {
  uint32_t head;        // next ticket
  uint32_t tail;        // tasks offered
  uint32_t split_done;  // tiles that have made their offers (or had none)
  uint32_t left;        // queries handed to the per-query kernel (single queries whose candidates overflow the staging area)
};

// Warp-uniform: draws the next task.  Returns false when there is none left for this warp.
__device__ __forceinline__ bool tile_next_task(TileQueue* q, unsigned long long* __restrict__ queue, uint32_t qcap, uint32_t tiles, int lane,
                                               uint32_t& ticket, uint32_t& tile, unsigned& mask) {
  uint32_t tk = 0;
  if (lane == 0) tk = atomicAdd(&q->head, 1u);
  tk = __shfl_sync(0xFFFFFFFFu, tk, 0);
  ticket = tk;
  if (tk < tiles) {
    tile = tk;
    mask = 0xFFFFFFFFu;
    return true;
  }
  const uint32_t j = tk - tiles;
  if (j >= qcap) return false;
  unsigned long long v = 0;
  if (lane == 0) {
    volatile unsigned long long* slot = queue + j;
    while (true) {
      if (*slot) {
        v = atomicExch(queue + j, 0ull);  // 0: the owner took it back in between
        break;
      }
      if (*reinterpret_cast<volatile uint32_t*>(&q->split_done) == tiles) {  // every offer there will be has been made
        v = atomicExch(queue + j, 0ull);
        break;
      }
      __nanosleep(400);
    }
  }
  v = __shfl_sync(0xFFFFFFFFu, v, 0);
  if (!v) return false;
  tile = (uint32_t)(v >> 32);
  mask = (unsigned)v;
  return true;
}

}  // namespace lio

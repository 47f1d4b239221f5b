// Exact k<=5 nearest-neighbour search on the voxel-hashed map — kernel K3 of SURVEY.md §2.2.
// Replaces KD_TREE::Nearest_Search (ikd-Tree, absent submodule): same result set,
// canonical order ascending (squared distance, id).
//
// A group of G lanes (G = 8: one bucket = one 128-byte line per load) serves one query:
//   ring 1 : the 27 cells around the query are probed by the lanes in parallel (all first-probe loads
//            are issued before any is consumed), then every hit bucket is read with one coalesced
//            float4-per-lane load; each lane keeps a sorted private top-5 of the slots it saw;
//   merge  : 5 rounds of a butterfly min over the lanes' list heads yield the group's top-5 in lanes 0..4;
//   rings 2..R are visited only while the 5th distance (or max_d2 when fewer than 5 were found) still
//            exceeds the distance from the query to the boundary of the block already searched, with
//            cells pruned by their box distance.  This is what makes the search exact.
#pragma once
#include "lio_device.cuh"

namespace lio {

struct TopList {  // ascending (d2 bits, id) keys with the slot they came from
  unsigned long long k0, k1, k2, k3, k4;
  uint32_t r0, r1, r2, r3, r4;
  __device__ __forceinline__ void clear() {
    k0 = k1 = k2 = k3 = k4 = ~0ull;
    r0 = r1 = r2 = r3 = r4 = kNone;
  }
  __device__ __forceinline__ void insert(unsigned long long k, uint32_t r) {
    if (k >= k4) return;
    const bool c0 = k < k0, c1 = k < k1, c2 = k < k2, c3 = k < k3;
    k4 = c3 ? k3 : k;  r4 = c3 ? r3 : r;
    k3 = c2 ? k2 : (c3 ? k : k3);  r3 = c2 ? r2 : (c3 ? r : r3);
    k2 = c1 ? k1 : (c2 ? k : k2);  r2 = c1 ? r1 : (c2 ? r : r2);
    k1 = c0 ? k0 : (c1 ? k : k1);  r1 = c0 ? r0 : (c1 ? r : r1);
    k0 = c0 ? k : k0;              r0 = c0 ? r : r0;
  }
  __device__ __forceinline__ void pop() {
    k0 = k1; k1 = k2; k2 = k3; k3 = k4; k4 = ~0ull;
    r0 = r1; r1 = r2; r2 = r3; r3 = r4; r4 = kNone;
  }
};

template <int G>
struct Group {
  static_assert(G == 8 || G == 16 || G == 32, "lanes per query");
  static constexpr int kSub = G / kSlots;  // buckets read per load instruction
  unsigned mask;                           // lanes of this group within the warp
  int gl;                                  // lane within the group
  int base;                                // first warp lane of the group
  __device__ __forceinline__ Group() {
    const int lane = threadIdx.x & 31;
    gl = lane & (G - 1);
    base = lane - gl;
    mask = (G == 32) ? 0xFFFFFFFFu : (((1u << G) - 1u) << base);
  }
  __device__ __forceinline__ unsigned ballot(bool p) const { return (__ballot_sync(mask, p) >> base) & ((G == 32) ? 0xFFFFFFFFu : ((1u << G) - 1u)); }
  template <typename T>
  __device__ __forceinline__ T shfl(T v, int src) const { return __shfl_sync(mask, v, base + src); }
};

// distance^2 from q to the axis-aligned box of cell (cx,cy,cz)
__device__ __forceinline__ float cell_box_d2(float qx, float qy, float qz, int cx, int cy, int cz, float h) {
  const float lx = (float)cx * h, ly = (float)cy * h, lz = (float)cz * h;
  float dx = 0.f, dy = 0.f, dz = 0.f;
  if (qx < lx) dx = lx - qx; else if (qx > lx + h) dx = qx - (lx + h);
  if (qy < ly) dy = ly - qy; else if (qy > ly + h) dy = qy - (ly + h);
  if (qz < lz) dz = lz - qz; else if (qz > lz + h) dz = qz - (lz + h);
  return dx * dx + dy * dy + dz * dz;
}

// read the chain of one cell; every lane of the group looks at slot (gl % 8) of bucket sub = gl / 8
template <int G>
__device__ __forceinline__ void scan_chain(const MapView& m, const Group<G>& g, uint32_t head, uint32_t ovf, float qx, float qy,
                                           float qz, uint32_t max_bits, TopList& L) {
  // kSub > 1: callers hand different chains to different sub-groups, so this routine only walks one chain
  uint32_t b = head;
  while (b != kNone) {
    const uint32_t ref = b * kSlots + (g.gl & (kSlots - 1));
    const float4 p = __ldg(m.slots + ref);
    const uint32_t id = __float_as_uint(p.w);
    const float d2 = sqdist3(qx, qy, qz, p.x, p.y, p.z);
    const uint32_t bits = __float_as_uint(d2);
    if (id != kEmptyId && d2 <= __uint_as_float(max_bits)) L.insert(((unsigned long long)bits << 32) | id, ref);
    b = (b == head) ? ovf : __ldg(m.next + b);
  }
}

// Search result, valid in lanes gl < cnt of the group: key (d2 bits << 32 | id), ref (slot index).
template <int G>
__device__ __forceinline__ void merge_top(const Group<G>& g, TopList& L, unsigned long long& okey, uint32_t& oref, int& cnt) {
  okey = ~0ull;
  oref = kNone;
  cnt = 0;
#pragma unroll
  for (int j = 0; j < kK; ++j) {
    unsigned long long k = L.k0;
    uint32_t r = L.r0;
#pragma unroll
    for (int off = G / 2; off >= 1; off >>= 1) {
      const unsigned long long ok = __shfl_xor_sync(g.mask, k, off);
      const uint32_t orf = __shfl_xor_sync(g.mask, r, off);
      if (ok < k) {
        k = ok;
        r = orf;
      }
    }
    if (k != ~0ull) ++cnt;
    if (L.k0 == k && k != ~0ull) L.pop();  // ids are unique: exactly one lane owns the minimum
    if (g.gl == j) {
      okey = k;
      oref = r;
    }
  }
}

// Scans the chains of up to 32 hit cells: bit b of `hits` = the cell probed by lane (b % G) in slot (b / G) of raw[].
// Bucket loads are issued four at a time before any is consumed, so a query waits for ~hits/4 memory round trips
// instead of one per cell.  G == 8 only (one bucket per load instruction).
template <int G, int NR>
__device__ __forceinline__ void scan_hits(const MapView& m, const Group<G>& g, unsigned hits, const uint4 (&raw)[NR], float qx, float qy,
                                          float qz, uint32_t max_bits, TopList& L) {
  static_assert(G == 8 && NR <= 4, "scan_hits: one bucket per 8-lane load");
  while (hits) {
    uint32_t hd[4], ov[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      hd[u] = ov[u] = kNone;
      if (hits) {
        const int b = __ffs(hits) - 1;
        hits &= hits - 1;
        const int r = b / G, src = b % G;
        uint32_t z = raw[0].z, w = raw[0].w;
#pragma unroll
        for (int t = 1; t < NR; ++t) {
          z = (r == t) ? raw[t].z : z;
          w = (r == t) ? raw[t].w : w;
        }
        hd[u] = g.shfl(z, src);
        ov[u] = g.shfl(w, src);
      }
    }
    float4 p[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (hd[u] != kNone) p[u] = __ldg(m.slots + hd[u] * kSlots + g.gl);
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (hd[u] != kNone) {
        const uint32_t id = __float_as_uint(p[u].w);
        const float d2 = sqdist3(qx, qy, qz, p[u].x, p[u].y, p[u].z);
        if (id != kEmptyId && d2 <= __uint_as_float(max_bits))
          L.insert(((unsigned long long)__float_as_uint(d2) << 32) | id, hd[u] * kSlots + g.gl);
        if (ov[u] != kNone) scan_chain<G>(m, g, ov[u], __ldg(m.next + ov[u]), qx, qy, qz, max_bits, L);  // overflow buckets: rare
      }
  }
}

template <int G>
__device__ void knn_search(const MapView& m, const Group<G>& g, float qx, float qy, float qz, float max_d2,
                           unsigned long long& okey, uint32_t& oref, int& cnt) {
  const float h = m.cell_size;
  const int cx = vox_of(qx, m.ds) >> m.cell_shift, cy = vox_of(qy, m.ds) >> m.cell_shift, cz = vox_of(qz, m.ds) >> m.cell_shift;
  const uint32_t max_bits = __float_as_uint(max_d2);
  TopList L;
  L.clear();

  // ---- ring 1: 27 cells, probes issued up front -------------------------------------------------
  constexpr int NR = (27 + G - 1) / G;
  unsigned long long keys[NR];
  uint32_t slot[NR];
  uint4 raw[NR];
#pragma unroll
  for (int r = 0; r < NR; ++r) {
    const int idx = g.gl + r * G;
    const int dz = idx / 9, dy = (idx / 3) % 3, dx = idx % 3;
    keys[r] = pack3(cx + dx - 1, cy + dy - 1, cz + dz - 1);
    slot[r] = hash64(keys[r]) & m.mask;
    if (idx < 27) raw[r] = __ldg(reinterpret_cast<const uint4*>(m.table + slot[r]));
    else raw[r] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, kNone, kNone);
  }
  unsigned all_hits = 0;
#pragma unroll
  for (int r = 0; r < NR; ++r) {
    unsigned long long ek = ((unsigned long long)raw[r].y << 32) | raw[r].x;
    while (ek != keys[r] && ek != kEmptyKey) {  // linear probing past a collision
      slot[r] = (slot[r] + 1) & m.mask;
      raw[r] = __ldg(reinterpret_cast<const uint4*>(m.table + slot[r]));
      ek = ((unsigned long long)raw[r].y << 32) | raw[r].x;
    }
    const bool hit = (ek == keys[r]) && (g.gl + r * G < 27);
    unsigned hits = g.ballot(hit);
    if (G == 8) {
      all_hits |= hits << (r * G);
      continue;
    }
    while (hits) {
      // hand one chain to each sub-group of 8 lanes
      uint32_t head = kNone, ovf = kNone;
      unsigned rest = hits;
#pragma unroll
      for (int s = 0; s < Group<G>::kSub; ++s) {
        const int src = rest ? (__ffs(rest) - 1) : 0;
        const uint32_t hh = g.shfl(raw[r].z, src), oo = g.shfl(raw[r].w, src);
        if (rest && (g.gl / kSlots) == s) {
          head = hh;
          ovf = oo;
        }
        if (rest) rest &= rest - 1;
      }
      hits = rest;
      scan_chain<G>(m, g, head, ovf, qx, qy, qz, max_bits, L);
    }
  }
  if constexpr (G == 8) scan_hits<G, NR>(m, g, all_hits, raw, qx, qy, qz, max_bits, L);
  merge_top<G>(g, L, okey, oref, cnt);

  // ---- rings 2.. : only while the answer could still change ------------------------------------------
  const float fx = qx - (float)cx * h, fy = qy - (float)cy * h, fz = qz - (float)cz * h;  // offset inside the centre cell
  const float lo = fminf(fminf(fx, fy), fz), hi = fminf(fminf(h - fx, h - fy), h - fz);
  const float edge = fmaxf(fminf(lo, hi), 0.f);  // distance to the nearest face of the centre cell
  const int max_ring = (int)ceilf(sqrtf(max_d2) / h) + 1;
  for (int ring = 2; ring <= max_ring; ++ring) {
    // everything within `reach` of the query has been seen
    const float reach = ((float)(ring - 1) * h + edge) * 0.9999f;
    const unsigned long long k5 = g.shfl(okey, kK - 1);
    const float tau = (cnt == kK) ? __uint_as_float((uint32_t)(k5 >> 32)) : max_d2;
    if (tau < reach * reach) break;
    // re-seed the private lists with the current answer
    L.clear();
    if (g.gl < cnt) L.insert(okey, oref);
    const int side = 2 * ring + 1, total = side * side * side;
    if constexpr (G == 8) {
      // four probe rounds (32 cells) in flight per step
      for (int idx0 = 0; idx0 < total; idx0 += 4 * G) {
        unsigned long long key4[4];
        uint32_t slot4[4];
        uint4 raw4[4];
        bool want4[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int idx = idx0 + u * G + g.gl;
          const int dz = idx / (side * side) - ring, dy = (idx / side) % side - ring, dx = idx % side - ring;
          bool want = idx < total && max(max(abs(dx), abs(dy)), abs(dz)) == ring;
          if (want) want = cell_box_d2(qx, qy, qz, cx + dx, cy + dy, cz + dz, h) * 0.9999f <= tau;
          want4[u] = want;
          key4[u] = pack3(cx + dx, cy + dy, cz + dz);
          slot4[u] = hash64(key4[u]) & m.mask;
          raw4[u] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, kNone, kNone);
          if (want) raw4[u] = __ldg(reinterpret_cast<const uint4*>(m.table + slot4[u]));
        }
        unsigned hits = 0;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          unsigned long long ek = ((unsigned long long)raw4[u].y << 32) | raw4[u].x;
          while (want4[u] && ek != key4[u] && ek != kEmptyKey) {
            slot4[u] = (slot4[u] + 1) & m.mask;
            raw4[u] = __ldg(reinterpret_cast<const uint4*>(m.table + slot4[u]));
            ek = ((unsigned long long)raw4[u].y << 32) | raw4[u].x;
          }
          hits |= g.ballot(want4[u] && ek == key4[u]) << (u * G);
        }
        scan_hits<G, 4>(m, g, hits, raw4, qx, qy, qz, max_bits, L);
      }
    } else
    for (int idx0 = 0; idx0 < total; idx0 += G) {
      const int idx = idx0 + g.gl;
      const int dz = idx / (side * side) - ring, dy = (idx / side) % side - ring, dx = idx % side - ring;
      bool want = idx < total && max(max(abs(dx), abs(dy)), abs(dz)) == ring;
      if (want) want = cell_box_d2(qx, qy, qz, cx + dx, cy + dy, cz + dz, h) * 0.9999f <= tau;
      HashEntry e;
      e.key = kEmptyKey;
      e.head = e.ovf = kNone;
      unsigned long long key = 0;
      if (want) {
        key = pack3(cx + dx, cy + dy, cz + dz);
        e = table_find(m, key);
      }
      unsigned hits = g.ballot(want && e.key == key);
      while (hits) {
        uint32_t head = kNone, ovf = kNone;
        unsigned rest = hits;
#pragma unroll
        for (int s = 0; s < Group<G>::kSub; ++s) {
          const int src = rest ? (__ffs(rest) - 1) : 0;
          const uint32_t hh = g.shfl(e.head, src), oo = g.shfl(e.ovf, src);
          if (rest && (g.gl / kSlots) == s) {
            head = hh;
            ovf = oo;
          }
          if (rest) rest &= rest - 1;
        }
        hits = rest;
        scan_chain<G>(m, g, head, ovf, qx, qy, qz, max_bits, L);
      }
    }
    merge_top<G>(g, L, okey, oref, cnt);
  }
}

// ------------------------------------------------------------------------------------------------ warp per query
// One query per warp, 128 cells probed per step (four rounds of lane-parallel hash probes in flight), the hit buckets
// read 16 at a time (one 128-byte line per 8 lanes, four loads in flight per lane).  Serves the queries the tile kernel
// defers: those are few (a few % of a scan) but deep — a query with fewer than five neighbours inside max_d2 has to see
// every cell within sqrt(max_d2), 343 of them at 1 m cells — so the sequential chain per query is what bounds the
// launch, and a warp cuts it 4x against the 8-lane group.  Same exact result as knn_search<G>.
struct WarpList {
  uint32_t b[128];      // hit buckets of the current step
  uint32_t n[128];      // their overflow bucket (kNone: none)
  uint32_t cells[256];  // rings 2-3: offsets (dx+8 | dy+8 << 8 | dz+8 << 16) of the cells worth probing (occupancy words)
};

// RING_T > 0: the ring is a compile-time constant (cell index -> offset is then multiply-shift instead of three integer
// divisions, which were a third of the instructions of a deep query); RING_T == 0: run-time ring.  ROUNDS <= 4 probe
// rounds of 32 cells per step.
// LIST: the cells come from wl.cells[0, ring_rt) (already filtered by shell, box distance and occupancy) instead of the cube.
template <int RING_T, int ROUNDS, bool LIST = false>
__device__ __forceinline__ void warp_scan_step(const MapView& m, WarpList& wl, int lane, int cx, int cy, int cz, int ring_rt, int idx0,
                                               float qx, float qy, float qz, float tau, uint32_t max_bits, TopList& L) {
  const float h = m.cell_size;
  const int ring = RING_T ? RING_T : ring_rt;
  const int side = 2 * ring + 1, total = LIST ? ring_rt : side * side * side;
  const unsigned lt = (1u << lane) - 1u;
  unsigned long long key4[ROUNDS];
  uint32_t slot4[ROUNDS];
  uint4 raw4[ROUNDS];
  bool want4[ROUNDS];
#pragma unroll
  for (int u = 0; u < ROUNDS; ++u) {
    const int idx = idx0 + u * 32 + lane;
    int dx, dy, dz;
    bool want;
    if (LIST) {
      want = idx < total;
      const uint32_t pk = want ? wl.cells[idx] : 0x080808u;
      dx = (int)(pk & 0xffu) - 8, dy = (int)((pk >> 8) & 0xffu) - 8, dz = (int)((pk >> 16) & 0xffu) - 8;
    } else {
      dz = idx / (side * side) - ring, dy = (idx / side) % side - ring, dx = idx % side - ring;
      want = idx < total && (ring == 1 || max(max(abs(dx), abs(dy)), abs(dz)) == ring);
      if (want && ring > 1) want = cell_box_d2(qx, qy, qz, cx + dx, cy + dy, cz + dz, h) * 0.9999f <= tau;
    }
    want4[u] = want;
    key4[u] = pack3(cx + dx, cy + dy, cz + dz);
    slot4[u] = hash64(key4[u]) & m.mask;
    raw4[u] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, kNone, kNone);
    if (want) raw4[u] = __ldg(reinterpret_cast<const uint4*>(m.table + slot4[u]));
  }
  int nb = 0;
#pragma unroll
  for (int u = 0; u < ROUNDS; ++u) {
    unsigned long long ek = ((unsigned long long)raw4[u].y << 32) | raw4[u].x;
    while (want4[u] && ek != key4[u] && ek != kEmptyKey) {  // linear probing past a collision
      slot4[u] = (slot4[u] + 1) & m.mask;
      raw4[u] = __ldg(reinterpret_cast<const uint4*>(m.table + slot4[u]));
      ek = ((unsigned long long)raw4[u].y << 32) | raw4[u].x;
    }
    const bool hit = want4[u] && ek == key4[u];
    const unsigned hb = __ballot_sync(0xFFFFFFFFu, hit);
    if (hit) {
      const int pos = nb + __popc(hb & lt);
      wl.b[pos] = raw4[u].z;
      wl.n[pos] = raw4[u].w;
    }
    nb += __popc(hb);
  }
  __syncwarp();
  const int sub = lane >> 3, sl = lane & 7;
  for (int b0 = 0; b0 < nb; b0 += 16) {
    float4 p[4];
    uint32_t bk[4], ov[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int wi = b0 + 4 * u + sub;
      bk[u] = ov[u] = kNone;
      if (wi < nb) {
        bk[u] = wl.b[wi];
        ov[u] = wl.n[wi];
        p[u] = __ldg(m.slots + (size_t)bk[u] * kSlots + sl);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (bk[u] != kNone) {
        const uint32_t id = __float_as_uint(p[u].w);
        const float d2 = sqdist3(qx, qy, qz, p[u].x, p[u].y, p[u].z);
        if (id != kEmptyId && d2 <= __uint_as_float(max_bits)) L.insert(key64(d2, id), bk[u] * kSlots + sl);
      }
      uint32_t b = ov[u];  // overflow buckets of the cell: rare, walked by the 8-lane sub-group that read its head
      while (__any_sync(0xFFFFFFFFu, b != kNone)) {
        if (b != kNone) {
          const float4 q = __ldg(m.slots + (size_t)b * kSlots + sl);
          const uint32_t id = __float_as_uint(q.w);
          const float d2 = sqdist3(qx, qy, qz, q.x, q.y, q.z);
          if (id != kEmptyId && d2 <= __uint_as_float(max_bits)) L.insert(key64(d2, id), b * kSlots + sl);
          b = __ldg(m.next + b);
        }
      }
    }
  }
  __syncwarp();  // the list is rewritten by the next step
}

// 64-bit mask of the cells of block (bx,by,bz) (4x4x4 cells) that lie within Chebyshev distance r of cell (cx,cy,cz)
__device__ __forceinline__ unsigned long long occ_cube_mask(int bx, int by, int bz, int cx, int cy, int cz, int r) {
  unsigned xm = 0, ym = 0, zm = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    xm |= (abs(bx * 4 + k - cx) <= r) ? (1u << k) : 0u;
    ym |= (abs(by * 4 + k - cy) <= r) ? (1u << k) : 0u;
    zm |= (abs(bz * 4 + k - cz) <= r) ? (1u << k) : 0u;
  }
  const unsigned long long mx = (unsigned long long)xm * 0x1111111111111111ull;
  unsigned long long my = 0, mz = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    my |= ((ym >> k) & 1u) ? (0x000F000F000F000Full << (4 * k)) : 0ull;
    mz |= ((zm >> k) & 1u) ? (0xFFFFull << (16 * k)) : 0ull;
  }
  return mx & my & mz;
}

// Rings 2 and 3 through the occupancy words: lane l < 27 holds the word of one 4x4x4 block around the query (`occ`, block
// coordinates bx,by,bz).  The cells of shell `ring` that were ever created and whose box is within tau go to wl.cells;
// returns their number.  In the sparse far-range regions where the deferred queries live this list has a handful of
// entries where the cube enumeration probed 98 / 218 mostly empty cells.
__device__ __forceinline__ int occ_ring_list(const MapView& m, WarpList& wl, int lane, unsigned long long occ, int bx, int by, int bz, int cx,
                                             int cy, int cz, int ring, float qx, float qy, float qz, float tau) {
  const unsigned lt = (1u << lane) - 1u;
  unsigned long long bits = occ ? (occ & occ_cube_mask(bx, by, bz, cx, cy, cz, ring) & ~occ_cube_mask(bx, by, bz, cx, cy, cz, ring - 1)) : 0ull;
  int n = 0;
  while (__any_sync(0xFFFFFFFFu, bits != 0ull)) {
    bool want = false;
    uint32_t pk = 0;
    if (bits) {
      const int b = __ffsll((long long)bits) - 1;
      bits &= bits - 1;
      const int dx = bx * 4 + (b & 3) - cx, dy = by * 4 + ((b >> 2) & 3) - cy, dz = bz * 4 + (b >> 4) - cz;
      want = cell_box_d2(qx, qy, qz, cx + dx, cy + dy, cz + dz, m.cell_size) * 0.9999f <= tau;
      pk = (uint32_t)(dx + 8) | ((uint32_t)(dy + 8) << 8) | ((uint32_t)(dz + 8) << 16);
    }
    const unsigned wb = __ballot_sync(0xFFFFFFFFu, want);
    if (want) wl.cells[n + __popc(wb & lt)] = pk;
    n += __popc(wb);
  }
  __syncwarp();
  return n;
}

// Result valid in lanes < cnt of the warp: key (d2 bits << 32 | id), ref (slot index).
__device__ void knn_search_warp(const MapView& m, WarpList& wl, float qx, float qy, float qz, float max_d2, unsigned long long& okey,
                                uint32_t& oref, int& cnt) {
  const Group<32> g;
  const int lane = g.gl;
  const float h = m.cell_size;
  const int cx = vox_of(qx, m.ds) >> m.cell_shift, cy = vox_of(qy, m.ds) >> m.cell_shift, cz = vox_of(qz, m.ds) >> m.cell_shift;
  const uint32_t max_bits = __float_as_uint(max_d2);
  TopList L;
  L.clear();
  warp_scan_step<1, 1>(m, wl, lane, cx, cy, cz, 1, 0, qx, qy, qz, max_d2, max_bits, L);
  merge_top<32>(g, L, okey, oref, cnt);
  const float fx = qx - (float)cx * h, fy = qy - (float)cy * h, fz = qz - (float)cz * h;  // offset inside the centre cell
  const float lo = fminf(fminf(fx, fy), fz), hi = fminf(fminf(h - fx, h - fy), h - fz);
  const float edge = fmaxf(fminf(lo, hi), 0.f);  // distance to the nearest face of the centre cell
  const int max_ring = (int)ceilf(sqrtf(max_d2) / h) + 1;
  // occupancy words of the (at most 3x3x3) blocks that cover rings 2 and 3, fetched on first use
  unsigned long long occ = 0;
  int obx = 0, oby = 0, obz = 0;
  bool occ_loaded = false;
  for (int ring = 2; ring <= max_ring; ++ring) {
    const float reach = ((float)(ring - 1) * h + edge) * 0.9999f;  // everything within `reach` of the query has been seen
    const unsigned long long k5 = g.shfl(okey, kK - 1);
    const float tau = (cnt == kK) ? __uint_as_float((uint32_t)(k5 >> 32)) : max_d2;
    if (tau < reach * reach) break;
    L.clear();
    if (lane < cnt) L.insert(okey, oref);  // re-seed the private lists with the current answer
    if (ring <= 3) {
      if (!occ_loaded) {
        const int bx0 = (cx - 3) >> 2, by0 = (cy - 3) >> 2, bz0 = (cz - 3) >> 2;
        const int nx = ((cx + 3) >> 2) - bx0 + 1, ny = ((cy + 3) >> 2) - by0 + 1, nz = ((cz + 3) >> 2) - bz0 + 1;  // 2 or 3 each
        if (lane < nx * ny * nz) {
          obx = bx0 + lane % nx, oby = by0 + (lane / nx) % ny, obz = bz0 + lane / (nx * ny);
          occ = __ldg(m.occ + occ_word_of(m, obx, oby, obz));
        }
        occ_loaded = true;
      }
      const int n_list = occ_ring_list(m, wl, lane, occ, obx, oby, obz, cx, cy, cz, ring, qx, qy, qz, tau);
#pragma unroll 1
      for (int idx0 = 0; idx0 < n_list; idx0 += 128) warp_scan_step<0, 4, true>(m, wl, lane, cx, cy, cz, n_list, idx0, qx, qy, qz, tau, max_bits, L);
    } else {
      const int side = 2 * ring + 1, total = side * side * side;
#pragma unroll 1
      for (int idx0 = 0; idx0 < total; idx0 += 128) warp_scan_step<0, 4>(m, wl, lane, cx, cy, cz, ring, idx0, qx, qy, qz, tau, max_bits, L);
    }
    merge_top<32>(g, L, okey, oref, cnt);
  }
}

// ------------------------------------------------------------------------------------------------ unbounded 1-NN
// Nearest live map point of q with NO distance bound (upstream's Nearest_Search is unbounded: a scan point whose 5-NN search
// found nothing within max_d2 still has its far nearest neighbour at the head of Nearest_Points[i], and map_incremental
// reads it).  One warp per query; rare path (points in front of the mapped area).  Shells of 4x4x4-cell blocks around the
// query are walked through the occupancy words - a lane per block, the set bits name the cells worth probing - until the
// best distance is inside the searched cube; when kFarRings shells (32 m at 1 m cells) hold nothing, every bucket of the
// pool is scanned.  Result: key (d2 bits << 32 | id, ~0: the map is empty) and slot, the same in every lane.
constexpr int kFarRings = 8;
__device__ __forceinline__ void far_scan_cell(const MapView& m, int cx, int cy, int cz, float qx, float qy, float qz, unsigned long long& best,
                                              uint32_t& bref) {
  const unsigned long long key = pack3(cx, cy, cz);
  const HashEntry e = table_find(m, key);
  if (e.key != key) return;
  for (uint32_t b = e.head; b != kNone; b = chain_next(m, e.head, e.ovf, b)) {
    if (b >= m.pool_cap) break;  // null bucket of a cell created under pool exhaustion
#pragma unroll
    for (int j = 0; j < kSlots; ++j) {
      const float4 p = __ldg(m.slots + (size_t)b * kSlots + j);
      const uint32_t id = __float_as_uint(p.w);
      if (id == kEmptyId) continue;
      const unsigned long long k = key64(sqdist3(qx, qy, qz, p.x, p.y, p.z), id);
      if (k < best) {
        best = k;
        bref = b * kSlots + j;
      }
    }
  }
}

__device__ void far_search_warp(const MapView& m, float qx, float qy, float qz, unsigned long long& okey, uint32_t& oref) {
  const int lane = threadIdx.x & 31;
  const float h = m.cell_size, hb = 4.f * m.cell_size;
  const int cx = vox_of(qx, m.ds) >> m.cell_shift, cy = vox_of(qy, m.ds) >> m.cell_shift, cz = vox_of(qz, m.ds) >> m.cell_shift;
  const int bx = cx >> 2, by = cy >> 2, bz = cz >> 2;
  // distance from q to the nearest face of its own block
  const float fx = qx - (float)bx * hb, fy = qy - (float)by * hb, fz = qz - (float)bz * hb;
  const float edge = fmaxf(fminf(fminf(fminf(fx, fy), fz), fminf(fminf(hb - fx, hb - fy), hb - fz)), 0.f);
  unsigned long long best = ~0ull;
  uint32_t bref = kNone;
  bool found_in_rings = false;
  for (int ring = 0; ring <= kFarRings; ++ring) {
    if (ring >= 1 && best != ~0ull) {  // everything within `reach` of q lies in the cube of rings < ring
      const float reach = ((float)(ring - 1) * hb + edge) * 0.9999f;
      if (__uint_as_float((uint32_t)(best >> 32)) < reach * reach) {
        found_in_rings = true;
        break;
      }
    }
    const int side = 2 * ring + 1, total = side * side * side;
    for (int idx = lane; idx < total; idx += 32) {
      const int dz = idx / (side * side) - ring, dy = (idx / side) % side - ring, dx = idx % side - ring;
      if (max(max(abs(dx), abs(dy)), abs(dz)) != ring) continue;
      const int ox = bx + dx, oy = by + dy, oz = bz + dz;
      unsigned long long word = __ldg(m.occ + occ_word_of(m, ox, oy, oz));
      while (word) {
        const int b = __ffsll((long long)word) - 1;
        word &= word - 1;
        const int ccx = ox * 4 + (b & 3), ccy = oy * 4 + ((b >> 2) & 3), ccz = oz * 4 + (b >> 4);
        const float bd = cell_box_d2(qx, qy, qz, ccx, ccy, ccz, h) * 0.9999f;
        if (best != ~0ull && bd > __uint_as_float((uint32_t)(best >> 32))) continue;
        far_scan_cell(m, ccx, ccy, ccz, qx, qy, qz, best, bref);
      }
    }
    // the warp's best so far, in every lane (tightens the pruning of the next shell)
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
      const unsigned long long ok = __shfl_xor_sync(0xFFFFFFFFu, best, off);
      const uint32_t orf = __shfl_xor_sync(0xFFFFFFFFu, bref, off);
      if (ok < best) {
        best = ok;
        bref = orf;
      }
    }
    if (ring == kFarRings && best != ~0ull) {
      const float reach = ((float)ring * hb + edge) * 0.9999f;
      found_in_rings = __uint_as_float((uint32_t)(best >> 32)) < reach * reach;
    }
  }
  if (!found_in_rings) {  // nothing (provably nearest) within the shells: every bucket handed out so far
    const uint32_t top = min(m.ctr->pool_top, m.pool_cap);
    const size_t total = (size_t)top * kSlots;
    for (size_t i = lane; i < total; i += 32) {
      const float4 p = __ldg(m.slots + i);
      const uint32_t id = __float_as_uint(p.w);
      if (id == kEmptyId) continue;
      const unsigned long long k = key64(sqdist3(qx, qy, qz, p.x, p.y, p.z), id);
      if (k < best) {
        best = k;
        bref = (uint32_t)i;
      }
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
      const unsigned long long ok = __shfl_xor_sync(0xFFFFFFFFu, best, off);
      const uint32_t orf = __shfl_xor_sync(0xFFFFFFFFu, bref, off);
      if (ok < best) {
        best = ok;
        bref = orf;
      }
    }
  }
  okey = best;
  oref = bref;
}

}  // namespace lio

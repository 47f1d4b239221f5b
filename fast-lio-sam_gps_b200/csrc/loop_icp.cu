// Batched point-to-point ICP verification — kernels K7 (exact 1-NN), K8 (fused reductions + closed-form solve) and
// K9 (all pairs of a batch advanced together) of SURVEY.md §2.2.
// Reference: LoopClosure::icpAlignment (fast_lio_sam/src/loop_closure.cpp:69-92, parameters :7-11); the PCL 1.10 pieces it
// calls (IterativeClosestPoint::computeTransformation, CorrespondenceEstimation over KdTreeFLANN, TransformationEstimationSVD
// = Eigen::umeyama without scaling, DefaultConvergenceCriteria, getFitnessScore) follow SURVEY.md §9.4; operation order is
// that of oracle/orc_icp.cpp so that transforms, iteration counts and scores compare bit-for-bit.
//
// Target side: one implicit balanced tree per target cloud over its points in Hilbert order; a leaf is 8 float4 = one 128-byte
// line, a node holds the boxes of its two children (64 bytes).  Search is exact (box pruning with a relative slack, ties
// -> smaller original index) and unbounded, as pcl's kd-tree search is; max_corr_dist is applied afterwards.
// Reductions are exact 64-bit fixed-point sums (order independent), so the parallel result equals the serial oracle's.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <cub/device/device_radix_sort.cuh>
#include "loop_ctx.h"
#include "loop_device.cuh"

namespace loopb {

namespace {

constexpr int kBlock = 256;
constexpr int kWarps = kBlock / 32;
constexpr uint32_t kPadIdx = 0x7FFFFFFFu;

int grid_of(size_t n, int block) { return (int)std::max<size_t>(1, (n + block - 1) / block); }

// compact index space of all tree points: tree t owns [cmp_off[t], cmp_off[t] + n)
__device__ __forceinline__ int find_tree(const uint32_t* __restrict__ cmp_off, int n_trees, uint32_t i) {
  int lo = 0, hi = n_trees - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (cmp_off[mid] <= i) lo = mid; else hi = mid - 1;
  }
  return lo;
}

__global__ void k_tree_box_init(BoxAcc* acc, int n) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  for (int a = 0; a < 3; ++a) {
    acc[j].lo[a] = 0xFFFFFFFFu;
    acc[j].hi[a] = 0u;
  }
  acc[j].n_finite = 0;
}

__global__ void __launch_bounds__(kBlock) k_tree_box(const TreeDev* __restrict__ trees, const uint32_t* __restrict__ cmp_off, int n_trees,
                                                     uint32_t n_total, const float4* __restrict__ vox, BoxAcc* __restrict__ acc) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  int t = -1;
  float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
  bool finite = false;
  if (i < n_total) {
    t = find_tree(cmp_off, n_trees, i);
    p = __ldg(vox + trees[t].src_off + (i - cmp_off[t]));
    finite = isfinite(p.x) && isfinite(p.y) && isfinite(p.z);
  }
  block_box_update(acc, t, finite, p.x, p.y, p.z);  // one set of atomics per block (per-thread atomics took 14 ms of a batch)
}

__device__ __forceinline__ uint32_t spread10(uint32_t v) {
  v &= 0x3FFu;
  v = (v | (v << 16)) & 0x030000FFu;
  v = (v | (v << 8)) & 0x0300F00Fu;
  v = (v | (v << 4)) & 0x030C30C3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}

// 30-bit Hilbert index of a 10-bit-per-axis cell (Skilling's axes-to-transpose form).  Index ranges of a Hilbert curve are
// spatially connected, so the halves of an index-balanced tree get tighter, less overlapping boxes than with Morton
// order, whose ranges jump across the cube at every octant boundary.
__device__ __forceinline__ uint32_t hilbert30(uint32_t x, uint32_t y, uint32_t z) {
  uint32_t X[3] = {x, y, z};
  const uint32_t M = 1u << 9;
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
  for (uint32_t Q = M; Q > 1; Q >>= 1)
    if (X[2] & Q) t ^= Q - 1;
  X[0] ^= t; X[1] ^= t; X[2] ^= t;
  return (spread10(X[0]) << 2) | (spread10(X[1]) << 1) | spread10(X[2]);
}

// sort key (tree << 32 | 30-bit space-filling-curve index inside the tree's box); only the ORDER matters, not exactness
__global__ void __launch_bounds__(kBlock) k_tree_keys(const TreeDev* __restrict__ trees, const uint32_t* __restrict__ cmp_off, int n_trees,
                                                      uint32_t n_total, const float4* __restrict__ vox, const BoxAcc* __restrict__ acc,
                                                      unsigned long long* __restrict__ keys, uint32_t* __restrict__ vals, int hilbert) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_total) return;
  const int t = find_tree(cmp_off, n_trees, i);
  const uint32_t l = i - cmp_off[t];
  const float4 p = __ldg(vox + trees[t].src_off + l);
  uint32_t code = 0x3FFFFFFFu;  // non-finite points sort last inside their tree and are never matched
  if (isfinite(p.x) && isfinite(p.y) && isfinite(p.z)) {
    const float lo[3] = {ord2f(acc[t].lo[0]), ord2f(acc[t].lo[1]), ord2f(acc[t].lo[2])};
    const float hi[3] = {ord2f(acc[t].hi[0]), ord2f(acc[t].hi[1]), ord2f(acc[t].hi[2])};
    const float ext = fmaxf(fmaxf(hi[0] - lo[0], hi[1] - lo[1]), fmaxf(hi[2] - lo[2], 1e-6f));
    const float sc = 1023.0f / ext;
    const uint32_t qx = (uint32_t)fminf(fmaxf((p.x - lo[0]) * sc, 0.f), 1023.f);
    const uint32_t qy = (uint32_t)fminf(fmaxf((p.y - lo[1]) * sc, 0.f), 1023.f);
    const uint32_t qz = (uint32_t)fminf(fmaxf((p.z - lo[2]) * sc, 0.f), 1023.f);
    code = hilbert ? hilbert30(qx, qy, qz) : (spread10(qx) | (spread10(qy) << 1) | (spread10(qz) << 2));
    if (code == 0x3FFFFFFFu) code = 0x3FFFFFFEu;  // the all-ones code is reserved for non-finite points
  }
  keys[i] = ((unsigned long long)(uint32_t)t << 32) | code;
  vals[i] = l;
}

// tree_pts[pts_off + r] = (xyz of the r-th point in curve order, original index); leaves padded with +inf points
__global__ void __launch_bounds__(kBlock) k_tree_gather(const TreeDev* __restrict__ trees, int n_trees, const uint32_t* __restrict__ cmp_off,
                                                        const uint32_t* __restrict__ pad_off, uint32_t n_pad_total,
                                                        const uint32_t* __restrict__ vals_sorted, const float4* __restrict__ vox,
                                                        float4* __restrict__ tree_pts, const unsigned long long* __restrict__ keys_sorted,
                                                        uint32_t* __restrict__ leaf_code) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_pad_total) return;
  const int t = find_tree(pad_off, n_trees, i);
  const TreeDev& tr = trees[t];
  const uint32_t r = i - pad_off[t];
  if ((r & (kLeafPts - 1)) == 0) leaf_code[i / kLeafPts] = (uint32_t)keys_sorted[cmp_off[t] + r];  // curve code of the leaf's first point
  float4 o = make_float4(INFINITY, INFINITY, INFINITY, __uint_as_float(kPadIdx));
  if (r < tr.n) {
    const uint32_t l = vals_sorted[cmp_off[t] + r];
    const float4 p = __ldg(vox + tr.src_off + l);
    if (isfinite(p.x) && isfinite(p.y) && isfinite(p.z)) o = make_float4(p.x, p.y, p.z, __uint_as_float(l));
  }
  tree_pts[i] = o;
}

__device__ __forceinline__ void box_of_leaf(const float4* __restrict__ pts, uint32_t leaf, uint32_t nleaf, float4& lo, float4& hi) {
  lo = make_float4(INFINITY, INFINITY, INFINITY, 0.f);
  hi = make_float4(-INFINITY, -INFINITY, -INFINITY, 0.f);
  if (leaf >= nleaf) return;
#pragma unroll
  for (int k = 0; k < kLeafPts; ++k) {
    const float4 p = pts[(size_t)leaf * kLeafPts + k];
    if (__float_as_uint(p.w) == kPadIdx) continue;
    lo.x = fminf(lo.x, p.x); lo.y = fminf(lo.y, p.y); lo.z = fminf(lo.z, p.z);
    hi.x = fmaxf(hi.x, p.x); hi.y = fmaxf(hi.y, p.y); hi.z = fmaxf(hi.z, p.z);
  }
}

// one block per tree: bottom level from the leaves, then level by level up to the root
__global__ void __launch_bounds__(1024) k_tree_nodes(const TreeDev* __restrict__ trees, const float4* __restrict__ tree_pts,
                                                     Node2* __restrict__ nodes) {
  const TreeDev tr = trees[blockIdx.x];
  Node2* nd = nodes + tr.node_off;
  const float4* pts = tree_pts + tr.pts_off;
  const uint32_t half = tr.nleaf_pad >> 1;
  for (uint32_t i = half + threadIdx.x; i < tr.nleaf_pad; i += blockDim.x) {
    Node2 n;
    box_of_leaf(pts, 2 * i - tr.nleaf_pad, tr.nleaf, n.lo0, n.hi0);
    box_of_leaf(pts, 2 * i + 1 - tr.nleaf_pad, tr.nleaf, n.lo1, n.hi1);
    nd[i] = n;
  }
  for (uint32_t lvl = half >> 1; lvl >= 1; lvl >>= 1) {
    __syncthreads();
    for (uint32_t i = lvl + threadIdx.x; i < 2 * lvl; i += blockDim.x) {
      const Node2 a = nd[2 * i], b = nd[2 * i + 1];
      Node2 n;
      n.lo0 = make_float4(fminf(a.lo0.x, a.lo1.x), fminf(a.lo0.y, a.lo1.y), fminf(a.lo0.z, a.lo1.z), 0.f);
      n.hi0 = make_float4(fmaxf(a.hi0.x, a.hi1.x), fmaxf(a.hi0.y, a.hi1.y), fmaxf(a.hi0.z, a.hi1.z), 0.f);
      n.lo1 = make_float4(fminf(b.lo0.x, b.lo1.x), fminf(b.lo0.y, b.lo1.y), fminf(b.lo0.z, b.lo1.z), 0.f);
      n.hi1 = make_float4(fmaxf(b.hi0.x, b.hi1.x), fmaxf(b.hi0.y, b.hi1.y), fmaxf(b.hi0.z, b.hi1.z), 0.f);
      nd[i] = n;
    }
    if (lvl == 1) break;
  }
}

// ------------------------------------------------------------------------------------------------ radix tree (LBVH)
// The balanced tree above splits an index range in the middle, wherever that falls in space: sibling boxes overlap and a
// seeded query still visits ~50 nodes.  Here the internal nodes are those of the binary radix tree over the leaves' curve
// keys (Karras 2012): a node covers the leaves that share a key prefix, i.e. the points of one cell of the curve's
// hierarchy, and its children split that cell at the highest differing bit - disjoint sub-cells, tight boxes.  Keys are made
// unique by appending the leaf index.  Child references live in the unused .w of the two low corners: bit 31 = leaf.
constexpr uint32_t kLeafFlag = 0x80000000u;
__device__ __forceinline__ int lbvh_delta(const uint32_t* __restrict__ code, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  const unsigned long long a = ((unsigned long long)code[i] << 32) | (uint32_t)i, b = ((unsigned long long)code[j] << 32) | (uint32_t)j;
  return __clzll((long long)(a ^ b));
}

// one thread per internal node of every tree (tree t has nleaf - 1 of them, numbered from 0 = root)
__global__ void __launch_bounds__(kBlock) k_lbvh_build(const TreeDev* __restrict__ trees, const uint32_t* __restrict__ pad_off, int n_trees,
                                                       uint32_t n_leaf_total, const uint32_t* __restrict__ leaf_code, Node2* __restrict__ nodes,
                                                       uint32_t* __restrict__ node_flag) {
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n_leaf_total) return;
  const int t = find_tree(pad_off, n_trees, g * kLeafPts);
  const TreeDev tr = trees[t];
  const uint32_t leaf0 = pad_off[t] / kLeafPts;  // first leaf of the tree in the global leaf arrays
  const int n = (int)tr.nleaf, i = (int)(g - leaf0);
  node_flag[tr.node_off + i] = 0u;
  if (i >= n - 1) return;
  const uint32_t* code = leaf_code + leaf0;
  // direction of the range, its other end (exponential + binary search), then the split position
  const int d = (lbvh_delta(code, n, i, i + 1) - lbvh_delta(code, n, i, i - 1)) >= 0 ? 1 : -1;
  const int dmin = lbvh_delta(code, n, i, i - d);
  int lmax = 2;
  while (lbvh_delta(code, n, i, i + lmax * d) > dmin) lmax <<= 1;
  int l = 0;
  for (int step = lmax >> 1; step >= 1; step >>= 1)
    if (lbvh_delta(code, n, i, i + (l + step) * d) > dmin) l += step;
  const int j = i + l * d;
  const int dnode = lbvh_delta(code, n, i, j);
  int sp = 0;
  for (int step = (l + 1) >> 1, len = l;; step = (step + 1) >> 1) {  // ceil halving down to 1
    if (sp + step <= len && lbvh_delta(code, n, i, i + (sp + step) * d) > dnode) sp += step;
    if (step == 1) break;
  }
  const int gamma = i + sp * d + min(d, 0);
  const int lo = min(i, j), hi = max(i, j);
  const uint32_t c0 = (lo == gamma) ? (kLeafFlag | (uint32_t)gamma) : (uint32_t)gamma;
  const uint32_t c1 = (hi == gamma + 1) ? (kLeafFlag | (uint32_t)(gamma + 1)) : (uint32_t)(gamma + 1);
  Node2& nd = nodes[tr.node_off + i];
  nd.lo0.w = __uint_as_float(c0);
  nd.lo1.w = __uint_as_float(c1);
}

// Child boxes of every internal node, one block per tree, bottom-up in rounds: a node is filled in once both its children are
// available (a leaf: always; an internal child: filled in an earlier round), so the number of rounds is the depth of the tree.
// Barriers of the block order the rounds - the one-thread-per-leaf form with a device-wide fence and an atomic ticket at
// every level (the usual LBVH refit) took 1.76 ms for the 131 trees of a batch, this one a tenth of that.
__global__ void __launch_bounds__(1024) k_lbvh_boxes(const TreeDev* __restrict__ trees, const float4* __restrict__ tree_pts, Node2* __restrict__ nodes,
                                                     uint32_t* __restrict__ node_flag) {
  const TreeDev tr = trees[blockIdx.x];
  if (tr.nleaf == 0) return;
  Node2* nd = nodes + tr.node_off;
  const float4* pts = tree_pts + tr.pts_off;
  if (tr.nleaf == 1) {  // a single leaf: node 0 holds it as child 0, child 1 is empty
    if (threadIdx.x == 0) {
      float4 lo, hi;
      box_of_leaf(pts, 0, 1, lo, hi);
      nd[0].lo0 = make_float4(lo.x, lo.y, lo.z, __uint_as_float(kLeafFlag));
      nd[0].hi0 = hi;
      nd[0].lo1 = make_float4(INFINITY, INFINITY, INFINITY, __uint_as_float(kLeafFlag | 0x7FFFFFFFu));
      nd[0].hi1 = make_float4(-INFINITY, -INFINITY, -INFINITY, 0.f);
    }
    return;
  }
  volatile uint32_t* flag = node_flag + tr.node_off;  // 0: open, 2: filled in this round, 1: available (zeroed by k_lbvh_build)
  const uint32_t n_int = tr.nleaf - 1;
  auto child_box = [&](uint32_t c, float4& lo, float4& hi) {
    if (c & kLeafFlag) {
      box_of_leaf(pts, c & ~kLeafFlag, tr.nleaf, lo, hi);
    } else {
      const Node2 k = nd[c];
      lo = make_float4(fminf(k.lo0.x, k.lo1.x), fminf(k.lo0.y, k.lo1.y), fminf(k.lo0.z, k.lo1.z), 0.f);
      hi = make_float4(fmaxf(k.hi0.x, k.hi1.x), fmaxf(k.hi0.y, k.hi1.y), fmaxf(k.hi0.z, k.hi1.z), 0.f);
    }
  };
  for (int round = 0; round < 80; ++round) {
    for (uint32_t i = threadIdx.x; i < n_int; i += blockDim.x) {
      if (flag[i]) continue;
      const uint32_t c0 = __float_as_uint(nd[i].lo0.w), c1 = __float_as_uint(nd[i].lo1.w);
      const bool r0 = (c0 & kLeafFlag) || flag[c0] == 1u, r1 = (c1 & kLeafFlag) || flag[c1] == 1u;
      if (!(r0 && r1)) continue;
      float4 lo0, hi0, lo1, hi1;
      child_box(c0, lo0, hi0);
      child_box(c1, lo1, hi1);
      nd[i].lo0 = make_float4(lo0.x, lo0.y, lo0.z, __uint_as_float(c0));
      nd[i].hi0 = hi0;
      nd[i].lo1 = make_float4(lo1.x, lo1.y, lo1.z, __uint_as_float(c1));
      nd[i].hi1 = hi1;
      flag[i] = 2u;
    }
    __syncthreads();
    int open = 0;
    for (uint32_t i = threadIdx.x; i < n_int; i += blockDim.x) {
      const uint32_t f = flag[i];
      if (f == 2u) flag[i] = 1u;
      open |= (f == 0u);
    }
    if (!__syncthreads_or(open)) break;
  }
}

// ------------------------------------------------------------------------------------------------ target grid
// A VoxelGrid output holds about one point per leaf-sized cell, and from the second ICP iteration on a query's previous
// correspondence lies within millimetres of the new one: every target point that could beat it sits in the one to eight
// cells the ball around the query touches.  Those are looked up directly (hash of the cell coordinates, points of a cell
// chained) instead of walking the tree: ~5 dependent loads against ~50 node visits + 8 leaf lines.  The result is the same
// minimum of (distance, index) over the whole target: every point with d2 <= the seed's distance lies in a probed cell.
__device__ __forceinline__ unsigned long long gkey(int x, int y, int z) {
  const unsigned long long m = 0x1FFFFFull;
  return (((unsigned long long)(x + (1 << 20)) & m) << 42) | (((unsigned long long)(y + (1 << 20)) & m) << 21) | ((unsigned long long)(z + (1 << 20)) & m);
}
__device__ __forceinline__ uint32_t ghash(unsigned long long k) {
  k ^= k >> 33;
  k *= 0xff51afd7ed558ccdull;
  k ^= k >> 33;
  k *= 0xc4ceb9fe1a85ec53ull;
  k ^= k >> 33;
  return (uint32_t)k;
}
__device__ __forceinline__ int gcell_of(float v, float inv) { return (int)floorf(v * inv); }

__global__ void __launch_bounds__(kBlock) k_tree_grid(const TreeDev* __restrict__ trees, const uint32_t* __restrict__ cmp_off, int n_trees,
                                                      uint32_t n_total, const float4* __restrict__ vox, GCell* __restrict__ gtab,
                                                      uint32_t* __restrict__ gnext, float inv) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_total) return;
  const int t = find_tree(cmp_off, n_trees, i);
  const TreeDev tr = trees[t];
  if (!tr.grid_mask) return;
  const uint32_t l = i - cmp_off[t];
  const float4 p = __ldg(vox + tr.src_off + l);
  gnext[i] = 0xFFFFFFFFu;
  if (!(isfinite(p.x) && isfinite(p.y) && isfinite(p.z))) return;
  const unsigned long long key = gkey(gcell_of(p.x, inv), gcell_of(p.y, inv), gcell_of(p.z, inv));
  GCell* tab = gtab + tr.grid_off;
  uint32_t s = ghash(key) & tr.grid_mask;
  while (true) {
    unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(&tab[s].key);
    if (cur == ~0ull) cur = atomicCAS(&tab[s].key, ~0ull, key);
    if (cur == ~0ull || cur == key) break;
    s = (s + 1) & tr.grid_mask;  // (the table is at least twice the point count: an empty entry always exists)
  }
  gnext[i] = atomicExch(&tab[s].head, l);
}

__device__ __constant__ int c_grid_max_cells = 8;  // LOOP_GRID_MAXC
// Seeded exact 1-NN through the grid.  Returns false when the ball touches more than c_grid_max_cells cells (the caller walks the tree).
__device__ __forceinline__ bool grid_probe(const TreeDev& tr, const GCell* __restrict__ gtab, const uint32_t* __restrict__ gnext,
                                           const float4* __restrict__ target, float inv, float qx, float qy, float qz, float& bd,
                                           uint32_t& bi, float4& bp) {
  const float r = sqrtf(bd) * 1.0001f + 1e-6f;
  const int x0 = gcell_of(qx - r, inv), x1 = gcell_of(qx + r, inv);
  const int y0 = gcell_of(qy - r, inv), y1 = gcell_of(qy + r, inv);
  const int z0 = gcell_of(qz - r, inv), z1 = gcell_of(qz + r, inv);
  const long long nc = (long long)(x1 - x0 + 1) * (y1 - y0 + 1) * (z1 - z0 + 1);
  if (nc > c_grid_max_cells || nc <= 0) return false;
  const GCell* tab = gtab + tr.grid_off;
  const uint32_t* nx = gnext + tr.cmp_off;
  for (int z = z0; z <= z1; ++z)
    for (int y = y0; y <= y1; ++y)
      for (int x = x0; x <= x1; ++x) {
        const unsigned long long key = gkey(x, y, z);
        uint32_t s = ghash(key) & tr.grid_mask;
        uint32_t head = 0xFFFFFFFFu;
        while (true) {
          const uint4 e = __ldg(reinterpret_cast<const uint4*>(tab + s));
          const unsigned long long ek = ((unsigned long long)e.y << 32) | e.x;
          if (ek == key) {
            head = e.z;
            break;
          }
          if (ek == ~0ull) break;
          s = (s + 1) & tr.grid_mask;
        }
        for (uint32_t l = head; l != 0xFFFFFFFFu; l = __ldg(nx + l)) {
          const float4 p = __ldg(target + l);
          const float d2 = sqd3(qx, qy, qz, p.x, p.y, p.z);
          if (d2 < bd || (d2 == bd && l < bi)) {
            bd = d2;
            bi = l;
            bp = p;
          }
        }
      }
  return true;
}

// ------------------------------------------------------------------------------------------------ exact 1-NN
__device__ __forceinline__ float box_d2(float qx, float qy, float qz, const float4& lo, const float4& hi) {
  float dx = 0.f, dy = 0.f, dz = 0.f;
  if (qx < lo.x) dx = lo.x - qx; else if (qx > hi.x) dx = qx - hi.x;
  if (qy < lo.y) dy = lo.y - qy; else if (qy > hi.y) dy = qy - hi.y;
  if (qz < lo.z) dz = lo.z - qz; else if (qz > hi.z) dz = qz - hi.z;
  return dx * dx + dy * dy + dz * dz;  // an empty box (lo = +inf) yields +inf
}

__device__ __forceinline__ void scan_leaf(const float4* __restrict__ pts, uint32_t leaf, float qx, float qy, float qz, float& bd,
                                          uint32_t& bi, float4& bp) {
  const float4* lp = pts + (size_t)leaf * kLeafPts;
#pragma unroll
  for (int k = 0; k < kLeafPts; ++k) {
    const float4 p = __ldg(lp + k);
    const uint32_t idx = __float_as_uint(p.w);
    const float d2 = sqd3(qx, qy, qz, p.x, p.y, p.z);
    if (idx != kPadIdx && (d2 < bd || (d2 == bd && idx < bi))) {
      bd = d2;
      bi = idx;
      bp = p;
    }
  }
}

// a subtree is visited unless it is empty (box distance +inf) or provably farther than the best so far; the box bound is
// rounded differently from point distances, hence the relative slack (more visits, never fewer: exactness is kept)
__device__ __forceinline__ bool visit(float d, float bd) { return d < INFINITY && d * 0.9999f <= bd; }

__device__ void nn_search_lbvh(const TreeDev& tr, const Node2* __restrict__ nodes, const float4* __restrict__ tree_pts, float qx, float qy,
                               float qz, float& bd, uint32_t& bi, float4& bp);

// Exact nearest neighbour of q in tree `tr`; bi = original index in the target cloud (kPadIdx: none), bp = its point.
// `seed` (original index of a target point, or -1) starts the search from that point's distance: the correspondence of the
// previous ICP iteration moves by millimetres, so its distance prunes nearly every subtree the cold search would still
// enter.  The result is the same minimum of (distance, index) over the whole target either way.
__device__ void nn_search(const TreeDev& tr, const Node2* __restrict__ nodes, const float4* __restrict__ tree_pts, float qx, float qy,
                          float qz, float& bd, uint32_t& bi, float4& bp, int seed = -1, const float4* __restrict__ target = nullptr,
                          const GCell* __restrict__ gtab = nullptr, const uint32_t* __restrict__ gnext = nullptr, float ginv = 0.f) {
  bd = INFINITY;
  bi = kPadIdx;
  bp = make_float4(0.f, 0.f, 0.f, 0.f);
  if (tr.n == 0 || !(isfinite(qx) && isfinite(qy) && isfinite(qz))) return;
  if (seed >= 0) {
    const float4 s = __ldg(target + seed);
    bd = sqd3(qx, qy, qz, s.x, s.y, s.z);
    bi = (uint32_t)seed;
    bp = s;
    if (gtab && tr.grid_mask && grid_probe(tr, gtab, gnext, target, ginv, qx, qy, qz, bd, bi, bp)) return;
  }
  if (tr.lbvh) {
    nn_search_lbvh(tr, nodes, tree_pts, qx, qy, qz, bd, bi, bp);
    return;
  }
  const Node2* nd = nodes + tr.node_off;
  const float4* pts = tree_pts + tr.pts_off;
  const uint32_t first_leaf_parent = tr.nleaf_pad >> 1;
  uint32_t st_node[24];
  float st_d[24];
  int sp = 0;
  uint32_t node = 1;
  while (true) {
    const Node2 n = nd[node];
    float d0 = box_d2(qx, qy, qz, n.lo0, n.hi0), d1 = box_d2(qx, qy, qz, n.lo1, n.hi1);
    uint32_t c0 = 2 * node, c1 = c0 + 1;
    if (d1 < d0) {  // nearer child first
      const float td = d0; d0 = d1; d1 = td;
      const uint32_t tc = c0; c0 = c1; c1 = tc;
    }
    uint32_t next = 0;
    if (node >= first_leaf_parent) {  // children are leaves
      if (visit(d0, bd)) scan_leaf(pts, c0 - tr.nleaf_pad, qx, qy, qz, bd, bi, bp);
      if (visit(d1, bd)) scan_leaf(pts, c1 - tr.nleaf_pad, qx, qy, qz, bd, bi, bp);
    } else {
      const bool v0 = visit(d0, bd), v1 = visit(d1, bd);
      if (v0) {
        next = c0;
        if (v1) {
          st_node[sp] = c1;
          st_d[sp] = d1;
          ++sp;
        }
      } else if (v1) {
        next = c1;
      }
    }
    while (!next && sp > 0) {
      --sp;
      if (visit(st_d[sp], bd)) next = st_node[sp];
    }
    if (!next) break;
    node = next;
  }
}

// the same exact search over the radix tree: children are named explicitly (bit 31: a leaf)
__device__ void nn_search_lbvh(const TreeDev& tr, const Node2* __restrict__ nodes, const float4* __restrict__ tree_pts, float qx, float qy,
                               float qz, float& bd, uint32_t& bi, float4& bp) {
  const Node2* nd = nodes + tr.node_off;
  const float4* pts = tree_pts + tr.pts_off;
  uint32_t st_node[64];  // a radix tree over 62-bit keys is at most 62 levels deep
  float st_d[64];
  int sp = 0;
  uint32_t node = 0;
  while (true) {
    const Node2 n = nd[node];
    float d0 = box_d2(qx, qy, qz, n.lo0, n.hi0), d1 = box_d2(qx, qy, qz, n.lo1, n.hi1);
    uint32_t c0 = __float_as_uint(n.lo0.w), c1 = __float_as_uint(n.lo1.w);
    if (d1 < d0) {  // nearer child first
      const float td = d0; d0 = d1; d1 = td;
      const uint32_t tc = c0; c0 = c1; c1 = tc;
    }
    uint32_t next = 0xFFFFFFFFu;
    if (visit(d0, bd)) {
      if (c0 & kLeafFlag) scan_leaf(pts, c0 & ~kLeafFlag, qx, qy, qz, bd, bi, bp);
      else next = c0;
    }
    if (visit(d1, bd)) {  // (bd may just have shrunk)
      if (c1 & kLeafFlag) scan_leaf(pts, c1 & ~kLeafFlag, qx, qy, qz, bd, bi, bp);
      else if (next == 0xFFFFFFFFu) next = c1;
      else if (sp < 64) {
        st_node[sp] = c1;
        st_d[sp] = d1;
        ++sp;
      }
    }
    while (next == 0xFFFFFFFFu && sp > 0) {
      --sp;
      if (visit(st_d[sp], bd)) next = st_node[sp];
    }
    if (next == 0xFFFFFFFFu) break;
    node = next;
  }
}

// ------------------------------------------------------------------------------------------------ ICP iteration
__device__ __forceinline__ void tf_f(const float* T, float x, float y, float z, float& ox, float& oy, float& oz) {
  ox = ((T[0] * x + T[1] * y) + T[2] * z) + T[3];
  oy = ((T[4] * x + T[5] * y) + T[6] * z) + T[7];
  oz = ((T[8] * x + T[9] * y) + T[10] * z) + T[11];
}
// pcl::transformPointCloud<PointT, float> SSE association used by getFitnessScore: p0 + (p1 + (p2 + c3))
__device__ __forceinline__ void tf_f_sse(const float* T, float x, float y, float z, float& ox, float& oy, float& oz) {
  ox = T[0] * x + (T[1] * y + (T[2] * z + T[3]));
  oy = T[4] * x + (T[5] * y + (T[6] * z + T[7]));
  oz = T[8] * x + (T[9] * y + (T[10] * z + T[11]));
}

// correspondence search + first-moment sums.  mode 0: ICP iteration, mode 1: fitness score pass.
// phase 0: every point (seeded with its previous correspondence from the second iteration on).  The FIRST iteration has no
// seeds and a cold tree walk costs twice a seeded one, so it runs in two launches: phase 1 searches every 8th point of the
// curve order cold (warp 0 of each block takes the block's 32 samples), phase 2 searches the others seeded with the match
// of their sample - a point a few decimetres from their own answer - and adds everybody's sums.  Seeds only ever supply
// an initial candidate, so the result is the exact minimum of (distance, index) in every case.
template <int MODE>
__global__ void __launch_bounds__(kBlock) k_icp_nn(PairDev* __restrict__ pairs, const uint32_t* __restrict__ blk2pair,
                                                   const float4* __restrict__ vox, float4* __restrict__ cur, int* __restrict__ match,
                                                   float* __restrict__ md2, const TreeDev* __restrict__ trees,
                                                   const Node2* __restrict__ nodes, const float4* __restrict__ tree_pts, double max_d2,
                                                   const GCell* __restrict__ gtab, const uint32_t* __restrict__ gnext, float ginv, int phase) {
  __shared__ long long s_red[kWarps][8];
  __shared__ float s_T[16];
  const uint32_t p = blk2pair[2 * blockIdx.x], first = blk2pair[2 * blockIdx.x + 1];
  PairDev& pr = pairs[p];
  if (MODE == 0 && !pr.active) return;
  const TreeDev& tr = trees[pr.dst_job];
  const float4* target = vox + tr.src_off;
  if (MODE == 0 && phase == 1) {  // sample pass of the first iteration: no sums, 32 cold searches per block
    if (threadIdx.x >= 32) return;
    const uint32_t l = first + 8u * threadIdx.x;
    if (l >= pr.src_pad_n) return;
    const float4 s = __ldg(tree_pts + pr.src_pts_off + l);
    if (__float_as_uint(s.w) == kPadIdx) return;
    const size_t ci = (size_t)pr.cur_off + l;
    float bd;
    uint32_t bi;
    float4 bp;
    nn_search(tr, nodes, tree_pts, s.x, s.y, s.z, bd, bi, bp);
    match[ci] = (bi != kPadIdx && !((double)bd > max_d2)) ? (int)bi : -1;
    md2[ci] = bd;
    return;
  }
  if (threadIdx.x < 16) s_T[threadIdx.x] = (MODE == 0) ? pr.Tr[threadIdx.x] : pr.Tfinal[threadIdx.x];
  __syncthreads();
  const uint32_t l = first + threadIdx.x;
  long long v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  // the source points are visited in Hilbert order (the warp's 32 queries walk the same part of the target tree: fewer
  // divergent traversal steps, node lines shared through L1); every sum below is an exact fixed-point sum, so the
  // visiting order does not change a bit of the result
  const float4 s = (l < pr.src_pad_n) ? __ldg(tree_pts + pr.src_pts_off + l) : make_float4(0.f, 0.f, 0.f, __uint_as_float(kPadIdx));
  const size_t ci = (size_t)pr.cur_off + l;
  if (__float_as_uint(s.w) == kPadIdx) {
    if (MODE == 0 && l < pr.src_pad_n && pr.iters == 0) match[ci] = -1;
  } else {
    float x, y, z;
    if (MODE == 0) {
      if (pr.iters == 0) {
        x = s.x; y = s.y; z = s.z;
      } else {
        const float4 c = cur[ci];
        tf_f(s_T, c.x, c.y, c.z, x, y, z);
      }
      cur[ci] = make_float4(x, y, z, 0.f);
    } else {
      tf_f_sse(s_T, s.x, s.y, s.z, x, y, z);
    }
    float bd;
    uint32_t bi;
    float4 bp;
    if (MODE == 0 && phase == 2 && (l & 7u) == 0u) {  // searched by the sample pass
      const int m = match[ci];
      bd = md2[ci];
      bi = m >= 0 ? (uint32_t)m : kPadIdx;
      bp = m >= 0 ? __ldg(target + m) : make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
      int seed = -1;
      if (pr.iters > 0) seed = match[ci];  // written by the previous iteration of this pair
      else if (phase == 2) seed = match[(size_t)pr.cur_off + (l & ~7u)];
      nn_search(tr, nodes, tree_pts, x, y, z, bd, bi, bp, seed, target, gtab, gnext, ginv);
    }
    if (MODE == 0) {
      const bool ok = bi != kPadIdx && !((double)bd > max_d2);
      match[ci] = ok ? (int)bi : -1;
      md2[ci] = bd;
      if (ok) {
        v[0] = fixc(x); v[1] = fixc(y); v[2] = fixc(z);
        v[3] = fixc(bp.x); v[4] = fixc(bp.y); v[5] = fixc(bp.z);
        v[6] = fixp((double)bd);
        v[7] = 1;
      }
    } else if (bi != kPadIdx) {
      v[6] = fixp((double)bd);
      v[7] = 1;
    }
  }
  const long long tot = block_sum_ll<8, kWarps>(v, s_red);
  if (threadIdx.x < 8 && tot != 0) {
    long long* dst;
    if (MODE == 0) dst = (threadIdx.x < 3) ? &pr.ss[threadIdx.x] : (threadIdx.x < 6) ? &pr.sd[threadIdx.x - 3] : (threadIdx.x == 6) ? &pr.mse_fix : &pr.cnt;
    else dst = (threadIdx.x == 6) ? &pr.score_fix : &pr.score_n;
    if (MODE == 0 || threadIdx.x >= 6) atomic_add_ll(dst, tot);
  }
}

// second-moment sums about the f32 means (Eigen::umeyama: src_demean / dst_demean, sigma = dst_demean * src_demean^T / n)
__global__ void __launch_bounds__(kBlock) k_icp_cov(PairDev* __restrict__ pairs, const uint32_t* __restrict__ blk2pair,
                                                    const float4* __restrict__ vox, const float4* __restrict__ cur,
                                                    const int* __restrict__ match, const TreeDev* __restrict__ trees, int min_corr) {
  __shared__ long long s_red[kWarps][9];
  const uint32_t p = blk2pair[2 * blockIdx.x], first = blk2pair[2 * blockIdx.x + 1];
  PairDev& pr = pairs[p];
  if (!pr.active) return;
  const long long n = pr.cnt;
  if (n < min_corr) return;
  float ms[3], md[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    ms[a] = (float)(((double)pr.ss[a] / kFixCoord) / (double)n);
    md[a] = (float)(((double)pr.sd[a] / kFixCoord) / (double)n);
  }
  const uint32_t l = first + threadIdx.x;
  long long v[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  if (l < pr.src_pad_n) {
    const size_t ci = (size_t)pr.cur_off + l;
    const int m = match[ci];
    if (m >= 0) {
      const float4 c = cur[ci];
      const float4 d = __ldg(vox + trees[pr.dst_job].src_off + m);
      const float a[3] = {c.x - ms[0], c.y - ms[1], c.z - ms[2]};
      const float b[3] = {d.x - md[0], d.y - md[1], d.z - md[2]};
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int cc = 0; cc < 3; ++cc) v[r * 3 + cc] = fixp((double)b[r] * (double)a[cc]);
    }
  }
  const long long tot = block_sum_ll<9, kWarps>(v, s_red);
  if (threadIdx.x < 9 && tot != 0) atomic_add_ll(&pr.cov[threadIdx.x], tot);
}

// ------------------------------------------------------------------------------------------------ closed-form solve
// 3x3 two-sided Jacobi SVD in f32 with Eigen::JacobiSVD's structure; same operation order as oracle/orc_icp.cpp svd3().
struct Rot {
  float c, s;
};
__device__ __forceinline__ void rot_apply(float& x, float& y, Rot j) {
  const float xn = j.c * x + j.s * y;
  const float yn = j.c * y - j.s * x;
  x = xn;
  y = yn;
}
__device__ __forceinline__ void rot_left(float* W, int p, int q, Rot j) {
  for (int k = 0; k < 3; ++k) rot_apply(W[p * 3 + k], W[q * 3 + k], j);
}
__device__ __forceinline__ void rot_right(float* W, int p, int q, Rot j) {
  const Rot t = {j.c, -j.s};
  for (int k = 0; k < 3; ++k) rot_apply(W[k * 3 + p], W[k * 3 + q], t);
}
__device__ __forceinline__ void make_jacobi(float x, float y, float z, Rot& j) {
  const float deno = 2.0f * fabsf(y);
  if (deno < FLT_MIN) {
    j.c = 1.f;
    j.s = 0.f;
    return;
  }
  const float tau = (x - z) / deno;
  const float w = sqrtf(tau * tau + 1.0f);
  const float t = tau > 0.f ? 1.0f / (tau + w) : 1.0f / (tau - w);
  const float sign_t = t > 0.f ? 1.0f : -1.0f;
  const float n = 1.0f / sqrtf(t * t + 1.0f);
  const float ysign = y > 0.f ? 1.0f : -1.0f;
  j.s = -sign_t * ysign * fabsf(t) * n;
  j.c = n;
}

__device__ void svd3(const float* M, float* U, float* S, float* V) {
  float W[9];
  float scale = 0.f;
  for (int i = 0; i < 9; ++i) scale = fmaxf(scale, fabsf(M[i]));
  if (!(scale > 0.f)) scale = 1.0f;
  for (int i = 0; i < 9; ++i) {
    W[i] = M[i] / scale;
    U[i] = V[i] = (i % 4 == 0) ? 1.f : 0.f;
  }
  const float precision = 2.0f * FLT_EPSILON;
  float maxdiag = fmaxf(fabsf(W[0]), fmaxf(fabsf(W[4]), fabsf(W[8])));
  bool finished = false;
  int sweeps = 0;
  while (!finished && sweeps < 64) {
    finished = true;
    ++sweeps;
    for (int p = 1; p < 3; ++p)
      for (int q = 0; q < p; ++q) {
        const float thr = fmaxf(FLT_MIN, precision * maxdiag);
        if (fabsf(W[p * 3 + q]) > thr || fabsf(W[q * 3 + p]) > thr) {
          finished = false;
          float m00 = W[p * 3 + p], m01 = W[p * 3 + q], m10 = W[q * 3 + p], m11 = W[q * 3 + q];
          const float t = m00 + m11, d = m10 - m01;
          Rot r1;
          if (fabsf(d) < FLT_MIN) {
            r1.s = 0.f;
            r1.c = 1.f;
          } else {
            const float u = t / d;
            const float tmp = sqrtf(1.0f + u * u);
            r1.s = 1.0f / tmp;
            r1.c = u / tmp;
          }
          rot_apply(m00, m10, r1);
          rot_apply(m01, m11, r1);
          Rot jr;
          make_jacobi(m00, m01, m11, jr);
          const Rot jrt = {jr.c, -jr.s};
          const Rot jl = {r1.c * jrt.c - r1.s * jrt.s, r1.c * jrt.s + r1.s * jrt.c};
          rot_left(W, p, q, jl);
          {
            const Rot jlt = {jl.c, -jl.s};
            rot_right(U, p, q, jlt);
          }
          rot_right(W, p, q, jr);
          rot_right(V, p, q, jr);
          maxdiag = fmaxf(maxdiag, fmaxf(fabsf(W[p * 3 + p]), fabsf(W[q * 3 + q])));
        }
      }
  }
  for (int i = 0; i < 3; ++i) {
    const float a = W[i * 3 + i];
    S[i] = fabsf(a);
    if (a < 0.f)
      for (int k = 0; k < 3; ++k) U[k * 3 + i] = -U[k * 3 + i];
    S[i] = S[i] * scale;
  }
  for (int i = 0; i < 3; ++i) {
    int b = i;
    for (int k = i + 1; k < 3; ++k)
      if (S[k] > S[b]) b = k;
    if (b != i) {
      float t = S[i]; S[i] = S[b]; S[b] = t;
      for (int k = 0; k < 3; ++k) {
        t = U[k * 3 + i]; U[k * 3 + i] = U[k * 3 + b]; U[k * 3 + b] = t;
        t = V[k * 3 + i]; V[k * 3 + i] = V[k * 3 + b]; V[k * 3 + b] = t;
      }
    }
  }
}

__device__ __forceinline__ float det3(const float* A) {
  const float a = A[0] * (A[4] * A[8] - A[5] * A[7]);
  const float b = A[1] * (A[3] * A[8] - A[5] * A[6]);
  const float c = A[2] * (A[3] * A[7] - A[4] * A[6]);
  return (a - b) + c;
}

// one thread per pair: umeyama from the sums, compose, DefaultConvergenceCriteria, reset the accumulators
__global__ void k_icp_solve(PairDev* __restrict__ pairs, int n_pairs, IcpParams prm, int* __restrict__ n_active) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_pairs) return;
  PairDev& pr = pairs[p];
  if (!pr.active) return;
  const long long n = pr.cnt;
  bool done = false;
  if (n < prm.min_corr) {
    pr.state = 5;
    pr.converged = 0;
    done = true;
  } else {
    float ms[3], md[3], sigma[9];
    for (int a = 0; a < 3; ++a) {
      ms[a] = (float)(((double)pr.ss[a] / kFixCoord) / (double)n);
      md[a] = (float)(((double)pr.sd[a] / kFixCoord) / (double)n);
    }
    for (int k = 0; k < 9; ++k) sigma[k] = (float)(((double)pr.cov[k] / kFixProd) / (double)n);
    float U[9], S[3], V[9], R[9], Tr[16];
    svd3(sigma, U, S, V);
    const float sgn = (det3(U) * det3(V) < 0.f) ? -1.0f : 1.0f;
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) {
        const float v = U[r * 3 + 0] * V[c * 3 + 0] + U[r * 3 + 1] * V[c * 3 + 1];
        R[r * 3 + c] = v + (U[r * 3 + 2] * sgn) * V[c * 3 + 2];
      }
    for (int r = 0; r < 3; ++r) {
      const float rs = (R[r * 3 + 0] * ms[0] + R[r * 3 + 1] * ms[1]) + R[r * 3 + 2] * ms[2];
      Tr[r * 4 + 0] = R[r * 3 + 0];
      Tr[r * 4 + 1] = R[r * 3 + 1];
      Tr[r * 4 + 2] = R[r * 3 + 2];
      Tr[r * 4 + 3] = md[r] - rs;
    }
    Tr[12] = Tr[13] = Tr[14] = 0.f;
    Tr[15] = 1.f;
    // final_transformation_ = transformation_ * final_transformation_
    float Tf[16];
    for (int r = 0; r < 4; ++r)
      for (int c = 0; c < 4; ++c) {
        float s = Tr[r * 4 + 0] * pr.Tfinal[0 * 4 + c] + Tr[r * 4 + 1] * pr.Tfinal[1 * 4 + c];
        s = s + Tr[r * 4 + 2] * pr.Tfinal[2 * 4 + c];
        Tf[r * 4 + c] = s + Tr[r * 4 + 3] * pr.Tfinal[3 * 4 + c];
      }
    for (int k = 0; k < 16; ++k) {
      pr.Tr[k] = Tr[k];
      pr.Tfinal[k] = Tf[k];
    }
    pr.iters += 1;
    // DefaultConvergenceCriteria::hasConverged
    if (pr.iters >= prm.max_iter) {
      pr.state = 1;
      pr.converged = 1;
      done = true;
    } else {
      bool is_similar = false;
      const double cos_angle = 0.5 * (((double)Tr[0] + (double)Tr[5] + (double)Tr[10]) - 1.0);
      const double tr2 = ((double)Tr[3] * (double)Tr[3] + (double)Tr[7] * (double)Tr[7]) + (double)Tr[11] * (double)Tr[11];
      if (cos_angle >= prm.rot_thr && tr2 <= prm.trans_eps) {
        if (pr.similar >= prm.max_similar) {
          pr.state = 2;
          pr.converged = 1;
          done = true;
        }
        is_similar = true;
      }
      const double mse = ((double)pr.mse_fix / kFixProd) / (double)n;
      if (!done && fabs(mse - pr.prev_mse) < prm.mse_abs) {
        if (pr.similar >= prm.max_similar) {
          pr.state = 3;
          pr.converged = 1;
          done = true;
        }
        is_similar = true;
      }
      if (!done && fabs(mse - pr.prev_mse) / pr.prev_mse < prm.fit_eps) {
        if (pr.similar >= prm.max_similar) {
          pr.state = 4;
          pr.converged = 1;
          done = true;
        }
        is_similar = true;
      }
      if (!done) {
        pr.similar = is_similar ? pr.similar + 1 : 0;
        pr.prev_mse = mse;
      }
    }
  }
  for (int a = 0; a < 3; ++a) pr.ss[a] = pr.sd[a] = 0;
  for (int k = 0; k < 9; ++k) pr.cov[k] = 0;
  pr.mse_fix = 0;
  pr.cnt = 0;
  if (done) {
    pr.active = 0;
    atomicSub(n_active, 1);
  }
}

__global__ void k_icp_finish(const PairDev* __restrict__ pairs, int n_pairs, const TreeDev* __restrict__ trees, double score_thr,
                             loop_result* __restrict__ out) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_pairs) return;
  const PairDev& pr = pairs[p];
  loop_result r;
  for (int k = 0; k < 16; ++k) r.T[k] = pr.Tfinal[k];
  r.score = pr.score_n > 0 ? ((double)pr.score_fix / kFixProd) / (double)pr.score_n : DBL_MAX;
  r.converged = pr.converged;
  r.valid = (pr.converged && r.score < score_thr) ? 1 : 0;
  r.iterations = pr.iters;
  r.state = pr.state;
  r.n_src = (int)pr.src_n;
  r.n_dst = (int)trees[pr.dst_job].n;
  out[p] = r;
}

__global__ void __launch_bounds__(kBlock) k_aligned(const float4* __restrict__ src, int n, const float* __restrict__ Tfinal,
                                                    float4* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float T[16];
  for (int k = 0; k < 16; ++k) T[k] = Tfinal[k];
  const float4 s = src[i];
  float x, y, z;
  tf_f(T, s.x, s.y, s.z, x, y, z);
  out[i] = make_float4(x, y, z, s.w);
}

}  // namespace

static uint32_t pow2_ceil(uint32_t v) {
  uint32_t p = 1;
  while (p < v) p <<= 1;
  return p;
}

// builds one search tree per distinct target cloud of the batch
static int build_trees(loop_ctx* c, const std::vector<int>& tree_jobs, int n_targets) {
  const int nt = (int)tree_jobs.size();
  cudaStream_t s = c->stream;
  c->trees_host.assign(nt, TreeDev{});
  std::vector<uint32_t> offs(2 * (size_t)nt + 2, 0);  // [0, nt): compact offsets, [nt, 2nt): padded offsets
  uint32_t cmp = 0, pad = 0, node = 0, gtot = 0, pad_targets = 0;  // pad_targets: padded points of the clouds that are searched
  for (int t = 0; t < nt; ++t) {
    const GridDev& g = c->grids_host[tree_jobs[t]];
    TreeDev& tr = c->trees_host[t];
    tr.src_off = g.vox_off;
    tr.n = g.vox_n;
    tr.nleaf = (tr.n + kLeafPts - 1) / kLeafPts;
    tr.nleaf_pad = std::max<uint32_t>(2u, pow2_ceil(tr.nleaf));
    tr.pts_off = pad;
    tr.node_off = node;
    tr.cmp_off = cmp;
    tr.lbvh = c->tree_kind ? 1u : 0u;
    tr.grid_off = gtot;
    tr.grid_mask = 0;
    if (c->use_grid && t < n_targets && tr.n > 0) {  // grids for the clouds that are searched (the sources are only ordered)
      const uint32_t gcap = std::max<uint32_t>(16u, pow2_ceil(2u * tr.n));
      tr.grid_mask = gcap - 1;
      gtot += gcap;
    }
    offs[t] = cmp;
    offs[nt + t] = pad;
    cmp += tr.n;
    pad += tr.nleaf * kLeafPts;
    node += tr.nleaf_pad;
    if (t == n_targets - 1) pad_targets = pad;
  }
  int rc;
  if ((rc = c->trees.ensure(std::max(nt, 1))) || (rc = c->blk2job.ensure(offs.size())) || (rc = c->tree_pts.ensure(std::max<uint32_t>(pad, 1))) ||
      (rc = c->nodes.ensure(std::max<uint32_t>(node, 1))) || (rc = c->gtab.ensure(std::max<uint32_t>(gtot, 1))) ||
      (rc = c->gnext.ensure(std::max<uint32_t>(cmp, 1))) || (rc = c->leaf_code.ensure(pad / kLeafPts + 1)) ||
      (rc = c->node_flag.ensure(std::max<uint32_t>(node, 1))))
    return rc;
  if (nt == 0) return LOOP_OK;
  LOOP_CUDA(cudaMemcpyAsync(c->trees.p, c->trees_host.data(), sizeof(TreeDev) * nt, cudaMemcpyHostToDevice, s));
  LOOP_CUDA(cudaMemcpyAsync(c->blk2job.p, offs.data(), sizeof(uint32_t) * offs.size(), cudaMemcpyHostToDevice, s));
  const uint32_t* cmp_off = c->blk2job.p;
  const uint32_t* pad_off = c->blk2job.p + nt;
  tbegin(c, LT_TREE);
  if (cmp) {
    if ((rc = c->keys_a.ensure(cmp)) || (rc = c->keys_b.ensure(cmp)) || (rc = c->vals_a.ensure(cmp)) || (rc = c->vals_b.ensure(cmp))) return rc;
    int bits = 1;
    while ((1u << bits) < (uint32_t)nt) ++bits;
    size_t need = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, need, c->keys_a.p, c->keys_b.p, c->vals_a.p, c->vals_b.p, (int)cmp, 0, 32 + bits, s);
    need = std::max(need, sizeof(BoxAcc) * (size_t)nt);
    if ((rc = c->cub_tmp.ensure(need))) return rc;
    BoxAcc* acc = reinterpret_cast<BoxAcc*>(c->cub_tmp.p);
    if (gtot) {
      LOOP_CUDA(cudaMemsetAsync(c->gtab.p, 0xFF, sizeof(GCell) * (size_t)gtot, s));
      k_tree_grid<<<grid_of(cmp, kBlock), kBlock, 0, s>>>(c->trees.p, cmp_off, nt, cmp, c->vox.p, c->gtab.p, c->gnext.p, 1.0f / c->grid_cell);
      c->launches += 1;
    }
    k_tree_box_init<<<grid_of(nt, 128), 128, 0, s>>>(acc, nt);
    k_tree_box<<<grid_of(cmp, kBlock), kBlock, 0, s>>>(c->trees.p, cmp_off, nt, cmp, c->vox.p, acc);
    k_tree_keys<<<grid_of(cmp, kBlock), kBlock, 0, s>>>(c->trees.p, cmp_off, nt, cmp, c->vox.p, acc, c->keys_a.p, c->vals_a.p, c->tree_order);
    size_t bytes = c->cub_tmp.cap;
    LOOP_CUDA(cub::DeviceRadixSort::SortPairs(c->cub_tmp.p, bytes, c->keys_a.p, c->keys_b.p, c->vals_a.p, c->vals_b.p, (int)cmp, 0, 32 + bits, s));
    c->launches += 4;
  }
  if (pad) {
    k_tree_gather<<<grid_of(pad, kBlock), kBlock, 0, s>>>(c->trees.p, nt, cmp_off, pad_off, pad, c->vals_b.p, c->vox.p, c->tree_pts.p, c->keys_b.p,
                                                          c->leaf_code.p);
    c->launches += 1;
  }
  // nodes only for the clouds that are searched: the targets come first in the list, the source clouds behind them went
  // through the ordering pass for their curve order alone
  if (c->tree_kind && pad_targets) {
    const uint32_t n_leaf_total = pad_targets / kLeafPts;
    k_lbvh_build<<<grid_of(n_leaf_total, kBlock), kBlock, 0, s>>>(c->trees.p, pad_off, n_targets, n_leaf_total, c->leaf_code.p, c->nodes.p, c->node_flag.p);
    k_lbvh_boxes<<<n_targets, 1024, 0, s>>>(c->trees.p, c->tree_pts.p, c->nodes.p, c->node_flag.p);
    c->launches += 2;
  } else if (n_targets > 0) {
    k_tree_nodes<<<n_targets, 1024, 0, s>>>(c->trees.p, c->tree_pts.p, c->nodes.p);
    c->launches += 1;
  }
  tend(c);
  LOOP_CUDA(cudaGetLastError());
  // offs / trees_host feed async copies: keep them alive until the copies are done
  LOOP_CUDA(cudaStreamSynchronize(s));
  return LOOP_OK;
}

int run_icp(loop_ctx* c, const std::vector<PairHost>& ph, double score_thr, loop_result* out) {
  const int np = (int)ph.size();
  c->n_pairs = np;
  c->st_iters = c->st_src = c->st_queries = 0;
  if (np == 0) return LOOP_OK;
  cudaStream_t s = c->stream;
  // distinct target clouds -> trees
  std::vector<int> tree_of_job(c->n_jobs, -1), tree_jobs;
  for (const PairHost& p : ph)
    if (tree_of_job[p.dst_job] < 0) {
      tree_of_job[p.dst_job] = (int)tree_jobs.size();
      tree_jobs.push_back(p.dst_job);
    }
  const int n_targets = (int)tree_jobs.size();
  if (const char* e = getenv("LOOP_GRID_MAXC")) {
    const int v = std::max(1, atoi(e));
    cudaMemcpyToSymbolAsync(c_grid_max_cells, &v, sizeof(int), 0, cudaMemcpyHostToDevice, s);
  }
  for (const PairHost& p : ph)  // the source clouds go through the same ordering pass: their points are visited in Hilbert order
    if (tree_of_job[p.src_job] < 0) {
      tree_of_job[p.src_job] = (int)tree_jobs.size();
      tree_jobs.push_back(p.src_job);
    }
  int rc = build_trees(c, tree_jobs, n_targets);
  if (rc) return rc;

  c->pairs_host.assign(np, PairDev{});
  std::vector<uint32_t> b2p;
  uint64_t cur_total = 0;
  for (int p = 0; p < np; ++p) {
    PairDev& pr = c->pairs_host[p];
    const GridDev& g = c->grids_host[ph[p].src_job];
    pr.src_off = g.vox_off;
    pr.src_n = g.vox_n;
    pr.dst_job = (uint32_t)tree_of_job[ph[p].dst_job];
    const TreeDev& stree = c->trees_host[tree_of_job[ph[p].src_job]];
    pr.src_pts_off = stree.pts_off;
    pr.src_pad_n = stree.nleaf * kLeafPts;
    pr.cur_off = (uint32_t)cur_total;
    for (int k = 0; k < 16; ++k) pr.Tr[k] = pr.Tfinal[k] = (k % 5 == 0) ? 1.f : 0.f;
    pr.prev_mse = DBL_MAX;
    pr.active = 1;
    for (uint32_t b = 0; b < pr.src_pad_n; b += kBlock) {
      b2p.push_back((uint32_t)p);
      b2p.push_back(b);
    }
    cur_total += pr.src_pad_n;
  }
  if (cur_total >= 0xFFFFFFFFull) {
    lio::set_error("batch chunk holds %llu source points (limit 2^32-1)", (unsigned long long)cur_total);
    return LOOP_E_CAPACITY;
  }
  const int n_blocks = (int)(b2p.size() / 2);
  if ((rc = c->pairs.ensure(np)) || (rc = c->blk2pair.ensure(std::max<size_t>(b2p.size(), 2))) || (rc = c->cur.ensure(std::max<uint64_t>(cur_total, 1))) ||
      (rc = c->match.ensure(std::max<uint64_t>(cur_total, 1))) || (rc = c->md2.ensure(std::max<uint64_t>(cur_total, 1))))
    return rc;
  LOOP_CUDA(cudaMemcpyAsync(c->pairs.p, c->pairs_host.data(), sizeof(PairDev) * np, cudaMemcpyHostToDevice, s));
  if (n_blocks) LOOP_CUDA(cudaMemcpyAsync(c->blk2pair.p, b2p.data(), sizeof(uint32_t) * b2p.size(), cudaMemcpyHostToDevice, s));
  *c->active_host = np;
  LOOP_CUDA(cudaMemcpyAsync(c->active_dev, c->active_host, sizeof(int), cudaMemcpyHostToDevice, s));
  LOOP_CUDA(cudaStreamSynchronize(s));  // b2p / pairs_host are host temporaries

  const IcpParams& prm = c->prm;
  for (int it = 0; it < prm.max_iter; ++it) {
    if (n_blocks) {
      tbegin(c, LT_NN);
      const float ginv = 1.0f / c->grid_cell;
      const GCell* gt = c->use_grid ? c->gtab.p : nullptr;
      if (it == 0 && c->sample_first) {
        k_icp_nn<0><<<n_blocks, kBlock, 0, s>>>(c->pairs.p, c->blk2pair.p, c->vox.p, c->cur.p, c->match.p, c->md2.p, c->trees.p, c->nodes.p,
                                               c->tree_pts.p, prm.max_d2, gt, c->gnext.p, ginv, 1);
        k_icp_nn<0><<<n_blocks, kBlock, 0, s>>>(c->pairs.p, c->blk2pair.p, c->vox.p, c->cur.p, c->match.p, c->md2.p, c->trees.p, c->nodes.p,
                                               c->tree_pts.p, prm.max_d2, gt, c->gnext.p, ginv, 2);
        c->launches += 1;
      } else {
        k_icp_nn<0><<<n_blocks, kBlock, 0, s>>>(c->pairs.p, c->blk2pair.p, c->vox.p, c->cur.p, c->match.p, c->md2.p, c->trees.p, c->nodes.p,
                                               c->tree_pts.p, prm.max_d2, gt, c->gnext.p, ginv, 0);
      }
      tend(c);
      tbegin(c, LT_COV);
      k_icp_cov<<<n_blocks, kBlock, 0, s>>>(c->pairs.p, c->blk2pair.p, c->vox.p, c->cur.p, c->match.p, c->trees.p, prm.min_corr);
      tend(c);
      c->launches += 2;
    }
    tbegin(c, LT_SOLVE);
    k_icp_solve<<<grid_of(np, 64), 64, 0, s>>>(c->pairs.p, np, prm, c->active_dev);
    tend(c);
    c->launches += 1;
    if ((it & 3) == 3 || it == prm.max_iter - 1) {
      LOOP_CUDA(cudaMemcpyAsync(c->active_host, c->active_dev, sizeof(int), cudaMemcpyDeviceToHost, s));
      LOOP_CUDA(cudaStreamSynchronize(s));
      if (*c->active_host <= 0) break;
    }
  }
  if (n_blocks) {
    tbegin(c, LT_SCORE);
    k_icp_nn<1><<<n_blocks, kBlock, 0, s>>>(c->pairs.p, c->blk2pair.p, c->vox.p, c->cur.p, c->match.p, c->md2.p, c->trees.p, c->nodes.p,
                                           c->tree_pts.p, prm.max_d2, c->use_grid ? c->gtab.p : nullptr, c->gnext.p, 1.0f / c->grid_cell, 0);
    tend(c);
    c->launches += 1;
  }
  // results: staged in the (now free) md2 head? no - md2 is part of the debug getters; use a dedicated small buffer in keys_a
  if ((rc = c->keys_a.ensure((sizeof(loop_result) * (size_t)np + 7) / 8))) return rc;
  loop_result* res_dev = reinterpret_cast<loop_result*>(c->keys_a.p);
  k_icp_finish<<<grid_of(np, 64), 64, 0, s>>>(c->pairs.p, np, c->trees.p, score_thr, res_dev);
  c->launches += 1;
  LOOP_CUDA(cudaGetLastError());
  LOOP_CUDA(cudaMemcpyAsync(out, res_dev, sizeof(loop_result) * (size_t)np, cudaMemcpyDeviceToHost, s));
  LOOP_CUDA(cudaMemcpyAsync(c->pairs_host.data(), c->pairs.p, sizeof(PairDev) * np, cudaMemcpyDeviceToHost, s));
  LOOP_CUDA(cudaStreamSynchronize(s));
  for (int p = 0; p < np; ++p) {
    c->st_iters += out[p].iterations;
    c->st_src += out[p].n_src;
    c->st_queries += (int64_t)out[p].n_src * (out[p].iterations + (out[p].state == 5 ? 1 : 0) + 1);
  }
  return LOOP_OK;
}

int fetch_aligned(loop_ctx* c, int pair, float* out, int cap, int* n_out) {
  const PairDev& pr = c->pairs_host[pair];
  if (n_out) *n_out = (int)pr.src_n;
  const int n = std::min<int>(cap, (int)pr.src_n);
  if (n <= 0 || !out) return LOOP_OK;
  int rc;
  if ((rc = c->raw.ensure(pr.src_n))) return rc;
  k_aligned<<<grid_of(pr.src_n, kBlock), kBlock, 0, c->stream>>>(c->vox.p + pr.src_off, (int)pr.src_n, c->pairs.p[pair].Tfinal, c->raw.p);
  c->launches += 1;
  LOOP_CUDA(cudaGetLastError());
  LOOP_CUDA(cudaMemcpyAsync(out, c->raw.p, sizeof(float4) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
  LOOP_CUDA(cudaStreamSynchronize(c->stream));
  return LOOP_OK;
}

int fetch_corr(loop_ctx* c, int pair, int* out) {
  const PairDev& pr = c->pairs_host[pair];
  if (!pr.src_n) return LOOP_OK;
  // match[] is indexed by Hilbert position: put it back in the source cloud's own order (points that are not in the
  // ordering, i.e. non-finite ones, have no correspondence)
  std::vector<int> m(pr.src_pad_n);
  std::vector<float4> pts(pr.src_pad_n);
  LOOP_CUDA(cudaMemcpyAsync(m.data(), c->match.p + pr.cur_off, sizeof(int) * m.size(), cudaMemcpyDeviceToHost, c->stream));
  LOOP_CUDA(cudaMemcpyAsync(pts.data(), c->tree_pts.p + pr.src_pts_off, sizeof(float4) * pts.size(), cudaMemcpyDeviceToHost, c->stream));
  LOOP_CUDA(cudaStreamSynchronize(c->stream));
  for (uint32_t i = 0; i < pr.src_n; ++i) out[i] = -1;
  for (uint32_t l = 0; l < pr.src_pad_n; ++l) {
    uint32_t orig;
    std::memcpy(&orig, &pts[l].w, sizeof(orig));
    if (orig != kPadIdx && orig < pr.src_n) out[orig] = m[l];
  }
  return LOOP_OK;
}

}  // namespace loopb

// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).  PARITY UNPINNED.
//
// CPU restatement of the OBSERVABLE semantics of ikd-Tree as FAST-LIO uses it
// (reference: third_party/FAST_LIO/include/ikd-Tree/ikd_Tree.{h,cpp} — submodule
// missing from /root/reference; behaviour from SURVEY.md §9.3
// [RECALL] and the published ikd-Tree paper):
//   Build, Nearest_Search(k, max_dist), Add_Points(downsample on/off),
//   Delete_Point_Boxes, size / validnum, flatten.
// Structure: a k-d tree with one point per node, incremental leaf insertion, lazy
// deletion and scapegoat-style partial rebuilds (synchronous; the reference's
// background rebuild thread has no observable effect on query results).
//
// Canonical choices made here because the reference leaves them to traversal order:
//   * every point handed to Build/Add_Points gets id = running insertion counter,
//     in list order, whether or not downsampling later drops it;
//   * kNN ties at equal squared distance are broken by smaller id;
//   * voxel membership for downsample-on-insert is floorf(p/ds) equality per axis.
#pragma once
#include <cstdint>
#include <cmath>
#include <vector>
#include <algorithm>

namespace orc {

struct OPt { float x, y, z; uint32_t id; };

struct KnnHit { float d2; uint32_t id; float x, y, z; };

inline float sqdist3(float ax, float ay, float az, float bx, float by, float bz) {
  float dx = ax - bx, dy = ay - by, dz = az - bz;
  float xx = dx * dx, yy = dy * dy, zz = dz * dz;
  float s = xx + yy;
  return s + zz;
}

class KdMap {
 public:
  explicit KdMap(float ds) : ds_(ds) {}

  float ds() const { return ds_; }
  uint32_t next_id() const { return next_id_; }
  int size() const { return root_ < 0 ? 0 : nodes_[root_].size - nodes_[root_].invalid; }

  void build(const float* xyz, int n) {
    nodes_.clear(); free_.clear(); root_ = -1;
    std::vector<OPt> pts(n);
    for (int i = 0; i < n; ++i) pts[i] = {xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], next_id_++};
    nodes_.reserve(size_t(n) * 2);
    root_ = build_rec(pts.data(), 0, n);
  }

  // Add_Points.  Returns the number of Add_by_point calls the sequential reference
  // algorithm performs (its return value); *net_new = points of THIS batch alive afterwards.
  int add_points(const float* xyz, int n, bool downsample, int* net_new) {
    int add_calls = 0;
    uint32_t base = next_id_;
    for (int i = 0; i < n; ++i) {
      OPt p = {xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], next_id_++};
      if (!downsample) { insert(p); ++add_calls; continue; }
      int kx = (int)std::floor(p.x / ds_), ky = (int)std::floor(p.y / ds_), kz = (int)std::floor(p.z / ds_);
      float minx = kx * ds_, miny = ky * ds_, minz = kz * ds_;
      float maxx = minx + ds_, maxy = miny + ds_, maxz = minz + ds_;
      float mx = minx + (maxx - minx) / 2.0f, my = miny + (maxy - miny) / 2.0f, mz = minz + (maxz - minz) / 2.0f;
      std::vector<int> in_vox;
      collect_voxel(root_, kx, ky, kz, in_vox);
      // existing points iterate in ascending id (canonical); strict '<' so the new point wins ties
      std::sort(in_vox.begin(), in_vox.end(), [&](int a, int b) { return nodes_[a].p.id < nodes_[b].p.id; });
      float min_dist = sqdist3(p.x, p.y, p.z, mx, my, mz);
      OPt best = p;
      for (int ni : in_vox) {
        const OPt& e = nodes_[ni].p;
        float d = sqdist3(e.x, e.y, e.z, mx, my, mz);
        if (d < min_dist) { min_dist = d; best = e; }
      }
      if (in_vox.size() > 1 || best.id == p.id) {
        for (int ni : in_vox) mark_deleted(ni);
        insert(best);
        ++add_calls;
      }
    }
    if (net_new) {
      int alive = 0;
      count_alive_ge(root_, base, alive);
      *net_new = alive;
    }
    maybe_rebuild_all();
    return add_calls;
  }

  // Delete_Point_Boxes: half-open boxes [min, max) per axis.  Returns points deleted.
  int delete_boxes(const float* boxes, int nb) {
    int cnt = 0;
    for (int b = 0; b < nb; ++b) cnt += delete_box_rec(root_, boxes + 6 * b);
    maybe_rebuild_all();
    return cnt;
  }

  // exact k-NN among live points with d2 <= max_d2; ascending (d2, id). returns count.
  int knn(const float q[3], int k, float max_d2, KnnHit* out) const {
    int cnt = 0;
    if (root_ >= 0) knn_rec(root_, q, k, max_d2, out, cnt);
    return cnt;
  }

  void flatten(std::vector<OPt>& out) const {
    out.clear();
    if (root_ >= 0) flatten_rec(root_, out);
    std::sort(out.begin(), out.end(), [](const OPt& a, const OPt& b) { return a.id < b.id; });
  }

 private:
  struct Node {
    OPt p;
    int l = -1, r = -1;
    int size = 1, invalid = 0;
    float bmin[3], bmax[3];
    uint8_t axis = 0;
    bool deleted = false;
  };

  float ds_;
  uint32_t next_id_ = 0;
  std::vector<Node> nodes_;
  std::vector<int> free_;
  int root_ = -1;

  static float coord(const OPt& p, int a) { return a == 0 ? p.x : (a == 1 ? p.y : p.z); }

  int alloc(const OPt& p, uint8_t axis) {
    int i;
    if (!free_.empty()) { i = free_.back(); free_.pop_back(); nodes_[i] = Node(); }
    else { i = (int)nodes_.size(); nodes_.emplace_back(); }
    Node& n = nodes_[i];
    n.p = p; n.axis = axis;
    n.bmin[0] = n.bmax[0] = p.x; n.bmin[1] = n.bmax[1] = p.y; n.bmin[2] = n.bmax[2] = p.z;
    return i;
  }

  void pull(int i) {
    Node& n = nodes_[i];
    n.size = 1; n.invalid = n.deleted ? 1 : 0;
    n.bmin[0] = n.bmax[0] = n.p.x; n.bmin[1] = n.bmax[1] = n.p.y; n.bmin[2] = n.bmax[2] = n.p.z;
    for (int c : {n.l, n.r}) {
      if (c < 0) continue;
      const Node& m = nodes_[c];
      n.size += m.size; n.invalid += m.invalid;
      for (int a = 0; a < 3; ++a) { n.bmin[a] = std::min(n.bmin[a], m.bmin[a]); n.bmax[a] = std::max(n.bmax[a], m.bmax[a]); }
    }
  }

  int build_rec(OPt* pts, int lo, int hi) {
    if (lo >= hi) return -1;
    float mn[3] = {1e30f, 1e30f, 1e30f}, mx[3] = {-1e30f, -1e30f, -1e30f};
    for (int i = lo; i < hi; ++i)
      for (int a = 0; a < 3; ++a) { float c = coord(pts[i], a); mn[a] = std::min(mn[a], c); mx[a] = std::max(mx[a], c); }
    int axis = 0;
    if (mx[1] - mn[1] > mx[axis] - mn[axis]) axis = 1;
    if (mx[2] - mn[2] > mx[axis] - mn[axis]) axis = 2;
    int mid = (lo + hi) / 2;
    std::nth_element(pts + lo, pts + mid, pts + hi, [axis](const OPt& a, const OPt& b) {
      float ca = coord(a, axis), cb = coord(b, axis);
      return ca < cb || (ca == cb && a.id < b.id);
    });
    int i = alloc(pts[mid], (uint8_t)axis);
    int l = build_rec(pts, lo, mid);
    int r = build_rec(pts, mid + 1, hi);
    nodes_[i].l = l; nodes_[i].r = r;
    pull(i);
    return i;
  }

  void collect_live(int i, std::vector<OPt>& out) {
    if (i < 0) return;
    collect_live(nodes_[i].l, out);
    if (!nodes_[i].deleted) out.push_back(nodes_[i].p);
    collect_live(nodes_[i].r, out);
    free_.push_back(i);
  }

  int rebuild(int i) {
    std::vector<OPt> pts;
    collect_live(i, pts);
    return build_rec(pts.data(), 0, (int)pts.size());
  }

  bool unbalanced(int i) const {
    const Node& n = nodes_[i];
    if (n.size <= 10) return false;
    if ((float)n.invalid / (float)n.size > 0.5f) return true;
    int ls = n.l >= 0 ? nodes_[n.l].size : 0;
    float bal = (float)ls / (float)(n.size - 1);
    return bal > 0.6f || bal < 0.4f;
  }

  void insert(const OPt& p) {
    if (root_ < 0) { root_ = alloc(p, 0); return; }
    std::vector<int> path;
    int cur = root_;
    while (true) {
      path.push_back(cur);
      Node& n = nodes_[cur];
      bool left = coord(p, n.axis) < coord(n.p, n.axis);
      int nxt = left ? n.l : n.r;
      if (nxt < 0) {
        int leaf = alloc(p, (uint8_t)((nodes_[cur].axis + 1) % 3));
        if (left) nodes_[cur].l = leaf; else nodes_[cur].r = leaf;
        break;
      }
      cur = nxt;
    }
    for (int j = (int)path.size() - 1; j >= 0; --j) pull(path[j]);
    // scapegoat: rebuild the highest unbalanced subtree on the path
    for (size_t j = 0; j < path.size(); ++j) {
      if (!unbalanced(path[j])) continue;
      int nb = rebuild(path[j]);
      if (j == 0) root_ = nb;
      else {
        Node& par = nodes_[path[j - 1]];
        if (par.l == path[j]) par.l = nb; else par.r = nb;
        for (int t = (int)j - 1; t >= 0; --t) pull(path[t]);
      }
      break;
    }
  }

  void maybe_rebuild_all() {
    if (root_ >= 0 && nodes_[root_].size > 10 && (float)nodes_[root_].invalid / (float)nodes_[root_].size > 0.5f)
      root_ = rebuild(root_);
  }

  // lazy delete of one node; fixes the counters on the root path by a top-down search
  void mark_deleted(int target) {
    std::vector<int> path;
    int cur = root_;
    const OPt& p = nodes_[target].p;
    find_path(cur, target, p, path);
    nodes_[target].deleted = true;
    for (int j = (int)path.size() - 1; j >= 0; --j) pull(path[j]);
  }

  bool find_path(int cur, int target, const OPt& p, std::vector<int>& path) {
    if (cur < 0) return false;
    const Node& n = nodes_[cur];
    for (int a = 0; a < 3; ++a)
      if (coord(p, a) < n.bmin[a] || coord(p, a) > n.bmax[a]) return false;
    path.push_back(cur);
    if (cur == target) return true;
    if (find_path(n.l, target, p, path)) return true;
    if (find_path(n.r, target, p, path)) return true;
    path.pop_back();
    return false;
  }

  void collect_voxel(int i, int kx, int ky, int kz, std::vector<int>& out) const {
    if (i < 0) return;
    const Node& n = nodes_[i];
    // conservative prune with a one-voxel margin (membership is decided by floor() below)
    float lo[3] = {(kx - 1) * ds_, (ky - 1) * ds_, (kz - 1) * ds_};
    float hi[3] = {(kx + 2) * ds_, (ky + 2) * ds_, (kz + 2) * ds_};
    for (int a = 0; a < 3; ++a)
      if (n.bmax[a] < lo[a] || n.bmin[a] > hi[a]) return;
    if (!n.deleted && (int)std::floor(n.p.x / ds_) == kx && (int)std::floor(n.p.y / ds_) == ky &&
        (int)std::floor(n.p.z / ds_) == kz)
      out.push_back(i);
    collect_voxel(n.l, kx, ky, kz, out);
    collect_voxel(n.r, kx, ky, kz, out);
  }

  int delete_box_rec(int i, const float* b) {
    if (i < 0) return 0;
    Node& n = nodes_[i];
    for (int a = 0; a < 3; ++a)
      if (n.bmax[a] < b[a] || n.bmin[a] >= b[3 + a]) return 0;
    int cnt = 0;
    if (!n.deleted && n.p.x >= b[0] && n.p.x < b[3] && n.p.y >= b[1] && n.p.y < b[4] && n.p.z >= b[2] && n.p.z < b[5]) {
      n.deleted = true;
      cnt = 1;
    }
    cnt += delete_box_rec(n.l, b);
    cnt += delete_box_rec(nodes_[i].r, b);
    if (cnt) pull(i);
    return cnt;
  }

  void count_alive_ge(int i, uint32_t base, int& alive) const {
    if (i < 0) return;
    const Node& n = nodes_[i];
    if (!n.deleted && n.p.id >= base) ++alive;
    count_alive_ge(n.l, base, alive);
    count_alive_ge(n.r, base, alive);
  }

  void flatten_rec(int i, std::vector<OPt>& out) const {
    const Node& n = nodes_[i];
    if (!n.deleted) out.push_back(n.p);
    if (n.l >= 0) flatten_rec(n.l, out);
    if (n.r >= 0) flatten_rec(n.r, out);
  }

  static bool hit_less(float d2a, uint32_t ida, float d2b, uint32_t idb) {
    return d2a < d2b || (d2a == d2b && ida < idb);
  }

  float box_d2(const Node& n, const float q[3]) const {
    float s = 0.f;
    for (int a = 0; a < 3; ++a) {
      float d = 0.f;
      if (q[a] < n.bmin[a]) d = n.bmin[a] - q[a];
      else if (q[a] > n.bmax[a]) d = q[a] - n.bmax[a];
      s += d * d;
    }
    return s;
  }

  void knn_rec(int i, const float q[3], int k, float max_d2, KnnHit* out, int& cnt) const {
    const Node& n = nodes_[i];
    // the box bound is computed with a different rounding than point distances, so
    // prune only with a small relative slack; exactness is preserved (slack => more visits)
    float bd = box_d2(n, q) * 0.9999f;
    if (bd > max_d2) return;
    if (cnt == k && bd > out[k - 1].d2) return;
    if (!n.deleted) {
      float d2 = sqdist3(q[0], q[1], q[2], n.p.x, n.p.y, n.p.z);
      if (d2 <= max_d2 && (cnt < k || hit_less(d2, n.p.id, out[k - 1].d2, out[k - 1].id))) {
        int j = cnt < k ? cnt++ : k - 1;
        while (j > 0 && hit_less(d2, n.p.id, out[j - 1].d2, out[j - 1].id)) { out[j] = out[j - 1]; --j; }
        out[j] = {d2, n.p.id, n.p.x, n.p.y, n.p.z};
      }
    }
    float qa = q[n.axis], pa = coord(n.p, n.axis);
    int first = qa < pa ? n.l : n.r, second = qa < pa ? n.r : n.l;
    if (first >= 0) knn_rec(first, q, k, max_d2, out, cnt);
    if (second >= 0) knn_rec(second, q, k, max_d2, out, cnt);
  }
};

}  // namespace orc

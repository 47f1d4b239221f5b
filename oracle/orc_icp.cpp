// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).  PARITY UNPINNED.
//
// CPU restatement of the loop-closure verification path of fast_lio_sam:
//   /root/reference/fast_lio_sam/src/loop_closure.cpp:18-40   fetchClosestKeyframeIdx
//   /root/reference/fast_lio_sam/src/loop_closure.cpp:42-67   setSrcAndDstCloud
//   /root/reference/fast_lio_sam/src/loop_closure.cpp:69-92   icpAlignment (+ ICP params :7-11)
//   /root/reference/fast_lio_sam/include/utilities.hpp:132-142 transformPcd
//   /root/reference/fast_lio_sam/include/utilities.hpp:158-171 voxelizePcd
// The arithmetic those functions delegate to lives in PCL 1.10 + Eigen
// (fast_lio_sam/CMakeLists.txt:30-32), neither of which is in the container; the PCL
// pieces (VoxelGrid::applyFilter, transformPointCloud, IterativeClosestPoint::
// computeTransformation, CorrespondenceEstimation, TransformationEstimationSVD/umeyama,
// DefaultConvergenceCriteria, getFitnessScore) are restated from SURVEY.md §9.4-9.5 [RECALL].
//
// Canonical choices where PCL's result depends on unspecified order:
//   * VoxelGrid: points of one voxel are accumulated in ORIGINAL cloud order (PCL uses an
//     unstable std::sort, so its own order is implementation-defined);
//   * 1-NN ties at equal squared distance -> smaller target index;
//   * the umeyama sums, the MSE and the fitness score are accumulated EXACTLY in 64-bit
//     fixed point (order-independent), then rounded once; PCL accumulates in f32/f64 in
//     cloud order.  This is what lets a parallel GPU reduction match this oracle bit-for-bit.
#include <cstdint>
#include <cstdio>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <vector>
#include <algorithm>
#include <numeric>
#include <omp.h>

namespace {

struct P4 { float x, y, z, i; };  // pcl::PointXYZI payload (utilities.hpp:29)

// ---------------------------------------------------------------- transforms
// pcl::transformPointCloud<PointT,double> (transformPcd passes an Eigen::Matrix4d): f64 math, f32 store
inline P4 tf_d(const double* T, const P4& p) {
  double x = p.x, y = p.y, z = p.z;
  P4 o;
  o.x = (float)(((T[0] * x + T[1] * y) + T[2] * z) + T[3]);
  o.y = (float)(((T[4] * x + T[5] * y) + T[6] * z) + T[7]);
  o.z = (float)(((T[8] * x + T[9] * y) + T[10] * z) + T[11]);
  o.i = p.i;
  return o;
}
// ICP's internal transformCloud: Eigen Matrix4f * Vector4f(x,y,z,1)
inline void tf_f(const float* T, float x, float y, float z, float* o) {
  o[0] = ((T[0] * x + T[1] * y) + T[2] * z) + T[3];
  o[1] = ((T[4] * x + T[5] * y) + T[6] * z) + T[7];
  o[2] = ((T[8] * x + T[9] * y) + T[10] * z) + T[11];
}
// pcl::transformPointCloud<PointT,float> SSE path (getFitnessScore): p0 + (p1 + (p2 + c3))
inline void tf_f_sse(const float* T, float x, float y, float z, float* o) {
  o[0] = T[0] * x + (T[1] * y + (T[2] * z + T[3]));
  o[1] = T[4] * x + (T[5] * y + (T[6] * z + T[7]));
  o[2] = T[8] * x + (T[9] * y + (T[10] * z + T[11]));
}

// ---------------------------------------------------------------- VoxelGrid (PCL 1.10 applyFilter)
// returns 0 ok, 1 = index overflow (PCL warns and returns the input unfiltered)
int voxel_grid(const std::vector<P4>& in, float leaf, std::vector<P4>& out) {
  out.clear();
  if (in.empty()) return 0;
  const float inv = 1.0f / leaf;
  float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  for (const P4& p : in) {
    if (!std::isfinite(p.x) || !std::isfinite(p.y) || !std::isfinite(p.z)) continue;
    mn[0] = std::min(mn[0], p.x); mn[1] = std::min(mn[1], p.y); mn[2] = std::min(mn[2], p.z);
    mx[0] = std::max(mx[0], p.x); mx[1] = std::max(mx[1], p.y); mx[2] = std::max(mx[2], p.z);
  }
  int64_t d[3];
  int minb[3], maxb[3];
  for (int a = 0; a < 3; ++a) {
    minb[a] = (int)std::floor(mn[a] * inv);
    maxb[a] = (int)std::floor(mx[a] * inv);
    d[a] = (int64_t)maxb[a] - minb[a] + 1;
  }
  if (d[0] * d[1] * d[2] > (int64_t)INT32_MAX) { out = in; return 1; }
  const int mul1 = (int)d[0], mul2 = (int)(d[0] * d[1]);
  std::vector<std::pair<int, int>> iv;
  iv.reserve(in.size());
  for (int k = 0; k < (int)in.size(); ++k) {
    const P4& p = in[k];
    if (!std::isfinite(p.x) || !std::isfinite(p.y) || !std::isfinite(p.z)) continue;
    int i0 = (int)(std::floor(p.x * inv) - (float)minb[0]);
    int i1 = (int)(std::floor(p.y * inv) - (float)minb[1]);
    int i2 = (int)(std::floor(p.z * inv) - (float)minb[2]);
    iv.emplace_back(i0 + i1 * mul1 + i2 * mul2, k);
  }
  std::stable_sort(iv.begin(), iv.end(), [](const std::pair<int, int>& a, const std::pair<int, int>& b) { return a.first < b.first; });
  size_t s = 0;
  while (s < iv.size()) {
    size_t e = s + 1;
    while (e < iv.size() && iv[e].first == iv[s].first) ++e;
    float sx = 0.f, sy = 0.f, sz = 0.f, si = 0.f;
    for (size_t j = s; j < e; ++j) {
      const P4& p = in[iv[j].second];
      sx = sx + p.x; sy = sy + p.y; sz = sz + p.z; si = si + p.i;
    }
    float cnt = (float)(e - s);
    out.push_back({sx / cnt, sy / cnt, sz / cnt, si / cnt});
    s = e;
  }
  return 0;
}

// ---------------------------------------------------------------- static kd-tree (stands in for pcl::KdTreeFLANN)
inline float sqd(const float* a, float bx, float by, float bz) {
  float dx = a[0] - bx, dy = a[1] - by, dz = a[2] - bz;
  float s = dx * dx + dy * dy;
  return s + dz * dz;
}

class KdStatic {
 public:
  void build(const std::vector<P4>& pts) {
    pts_ = &pts;
    idx_.resize(pts.size());
    std::iota(idx_.begin(), idx_.end(), 0);
    nodes_.clear();
    nodes_.reserve(pts.size() / 4 + 16);
    if (!pts.empty()) build_rec(0, (int)pts.size());
  }
  // exact 1-NN, ties -> smaller index
  void nn(const float* q, int& best_i, float& best_d2) const {
    best_i = -1; best_d2 = INFINITY;
    if (!nodes_.empty()) nn_rec(0, q, best_i, best_d2);
  }

 private:
  struct Node { int lo, hi, l, r; float mn[3], mx[3]; int axis; float split; };
  const std::vector<P4>* pts_ = nullptr;
  std::vector<int> idx_;
  std::vector<Node> nodes_;
  static constexpr int kLeaf = 12;
  float c(int i, int a) const { const P4& p = (*pts_)[i]; return a == 0 ? p.x : (a == 1 ? p.y : p.z); }
  int build_rec(int lo, int hi) {
    int id = (int)nodes_.size();
    nodes_.emplace_back();
    Node n;
    n.lo = lo; n.hi = hi; n.l = n.r = -1;
    for (int a = 0; a < 3; ++a) { n.mn[a] = FLT_MAX; n.mx[a] = -FLT_MAX; }
    for (int j = lo; j < hi; ++j)
      for (int a = 0; a < 3; ++a) { float v = c(idx_[j], a); n.mn[a] = std::min(n.mn[a], v); n.mx[a] = std::max(n.mx[a], v); }
    n.axis = 0;
    if (n.mx[1] - n.mn[1] > n.mx[n.axis] - n.mn[n.axis]) n.axis = 1;
    if (n.mx[2] - n.mn[2] > n.mx[n.axis] - n.mn[n.axis]) n.axis = 2;
    if (hi - lo > kLeaf) {
      int mid = (lo + hi) / 2, ax = n.axis;
      std::nth_element(idx_.begin() + lo, idx_.begin() + mid, idx_.begin() + hi, [&](int a, int b) {
        float ca = c(a, ax), cb = c(b, ax);
        return ca < cb || (ca == cb && a < b);
      });
      n.split = c(idx_[mid], ax);
      nodes_[id] = n;
      int l = build_rec(lo, mid);
      int r = build_rec(mid, hi);
      nodes_[id].l = l; nodes_[id].r = r;
    } else {
      nodes_[id] = n;
    }
    return id;
  }
  void nn_rec(int ni, const float* q, int& bi, float& bd) const {
    const Node& n = nodes_[ni];
    float s = 0.f;
    for (int a = 0; a < 3; ++a) {
      float d = 0.f;
      if (q[a] < n.mn[a]) d = n.mn[a] - q[a]; else if (q[a] > n.mx[a]) d = q[a] - n.mx[a];
      s += d * d;
    }
    if (s * 0.9999f > bd) return;
    if (n.l < 0) {
      for (int j = n.lo; j < n.hi; ++j) {
        int i = idx_[j];
        const P4& p = (*pts_)[i];
        float d2 = sqd(q, p.x, p.y, p.z);
        if (d2 < bd || (d2 == bd && i < bi)) { bd = d2; bi = i; }
      }
      return;
    }
    bool left_first = q[n.axis] < n.split;
    nn_rec(left_first ? n.l : n.r, q, bi, bd);
    nn_rec(left_first ? n.r : n.l, q, bi, bd);
  }
};

// ---------------------------------------------------------------- 3x3 Jacobi SVD, f32 (Eigen::JacobiSVD structure)
struct Rot { float c, s; };

// rows/cols helper: x' = c x + s y ; y' = -s x + c y
inline void rot_apply(float& x, float& y, Rot j) {
  float xn = j.c * x + j.s * y;
  float yn = j.c * y - j.s * x;
  x = xn; y = yn;
}
inline void left(float* W, int p, int q, Rot j) { for (int k = 0; k < 3; ++k) rot_apply(W[p * 3 + k], W[q * 3 + k], j); }
inline void right(float* W, int p, int q, Rot j) {  // applyOnTheRight uses the transposed rotation
  Rot t = {j.c, -j.s};
  for (int k = 0; k < 3; ++k) rot_apply(W[k * 3 + p], W[k * 3 + q], t);
}

inline bool make_jacobi(float x, float y, float z, Rot& j) {
  float deno = 2.0f * std::fabs(y);
  if (deno < FLT_MIN) { j.c = 1.f; j.s = 0.f; return false; }
  float tau = (x - z) / deno;
  float w = std::sqrt(tau * tau + 1.0f);
  float t = tau > 0.f ? 1.0f / (tau + w) : 1.0f / (tau - w);
  float sign_t = t > 0.f ? 1.0f : -1.0f;
  float n = 1.0f / std::sqrt(t * t + 1.0f);
  float ysign = y > 0.f ? 1.0f : -1.0f;
  j.s = -sign_t * ysign * std::fabs(t) * n;
  j.c = n;
  return true;
}

void svd3(const float* M, float* U, float* S, float* V) {
  float W[9];
  float scale = 0.f;
  for (int i = 0; i < 9; ++i) scale = std::max(scale, std::fabs(M[i]));
  if (!(scale > 0.f)) scale = 1.0f;
  for (int i = 0; i < 9; ++i) { W[i] = M[i] / scale; U[i] = V[i] = (i % 4 == 0) ? 1.f : 0.f; }
  const float precision = 2.0f * FLT_EPSILON;
  float maxdiag = std::max(std::fabs(W[0]), std::max(std::fabs(W[4]), std::fabs(W[8])));
  bool finished = false;
  int sweeps = 0;
  while (!finished && sweeps < 64) {
    finished = true; ++sweeps;
    for (int p = 1; p < 3; ++p)
      for (int q = 0; q < p; ++q) {
        float thr = std::max(FLT_MIN, precision * maxdiag);
        if (std::fabs(W[p * 3 + q]) > thr || std::fabs(W[q * 3 + p]) > thr) {
          finished = false;
          // real_2x2_jacobi_svd on [[W(p,p) W(p,q)],[W(q,p) W(q,q)]]
          float m00 = W[p * 3 + p], m01 = W[p * 3 + q], m10 = W[q * 3 + p], m11 = W[q * 3 + q];
          float t = m00 + m11, d = m10 - m01;
          Rot r1;
          if (std::fabs(d) < FLT_MIN) { r1.s = 0.f; r1.c = 1.f; }
          else { float u = t / d; float tmp = std::sqrt(1.0f + u * u); r1.s = 1.0f / tmp; r1.c = u / tmp; }
          // m.applyOnTheLeft(0,1,r1)
          rot_apply(m00, m10, r1);
          rot_apply(m01, m11, r1);
          Rot jr;
          make_jacobi(m00, m01, m11, jr);
          // j_left = r1 * jr^T
          Rot jrt = {jr.c, -jr.s};
          Rot jl = {r1.c * jrt.c - r1.s * jrt.s, r1.c * jrt.s + r1.s * jrt.c};
          left(W, p, q, jl);
          { Rot jlt = {jl.c, -jl.s}; right(U, p, q, jlt); }
          right(W, p, q, jr);
          right(V, p, q, jr);
          maxdiag = std::max(maxdiag, std::max(std::fabs(W[p * 3 + p]), std::fabs(W[q * 3 + q])));
        }
      }
  }
  for (int i = 0; i < 3; ++i) {
    float a = W[i * 3 + i];
    S[i] = std::fabs(a);
    if (a < 0.f) for (int k = 0; k < 3; ++k) U[k * 3 + i] = -U[k * 3 + i];
    S[i] = S[i] * scale;
  }
  for (int i = 0; i < 3; ++i) {
    int b = i;
    for (int k = i + 1; k < 3; ++k) if (S[k] > S[b]) b = k;
    if (b != i) {
      std::swap(S[i], S[b]);
      for (int k = 0; k < 3; ++k) { std::swap(U[k * 3 + i], U[k * 3 + b]); std::swap(V[k * 3 + i], V[k * 3 + b]); }
    }
  }
}

inline float det3(const float* A) {
  float a = A[0] * (A[4] * A[8] - A[5] * A[7]);
  float b = A[1] * (A[3] * A[8] - A[5] * A[6]);
  float c = A[2] * (A[3] * A[7] - A[4] * A[6]);
  return (a - b) + c;
}

// fixed-point helpers (exact, order-independent accumulation)
constexpr double kFixCoord = 16777216.0;  // 2^24
constexpr double kFixProd = 1048576.0;    // 2^20
inline int64_t fixc(float v) { return (int64_t)std::llrint((double)v * kFixCoord); }
inline int64_t fixp(double v) { return (int64_t)std::llrint(v * kFixProd); }

// Eigen::umeyama(src, dst, with_scaling=false) on matched pairs -> 4x4 row-major f32
void umeyama(const std::vector<float>& src, const std::vector<float>& dst, float* T) {
  const size_t n = src.size() / 3;
  int64_t ss[3] = {0, 0, 0}, sd[3] = {0, 0, 0};
  for (size_t i = 0; i < n; ++i)
    for (int a = 0; a < 3; ++a) { ss[a] += fixc(src[3 * i + a]); sd[a] += fixc(dst[3 * i + a]); }
  float ms[3], md[3];
  for (int a = 0; a < 3; ++a) {
    ms[a] = (float)(((double)ss[a] / kFixCoord) / (double)n);
    md[a] = (float)(((double)sd[a] / kFixCoord) / (double)n);
  }
  int64_t cov[9] = {0};
  for (size_t i = 0; i < n; ++i) {
    float a[3], b[3];
    for (int k = 0; k < 3; ++k) { a[k] = src[3 * i + k] - ms[k]; b[k] = dst[3 * i + k] - md[k]; }
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) cov[r * 3 + c] += fixp((double)b[r] * (double)a[c]);
  }
  float sigma[9];
  for (int k = 0; k < 9; ++k) sigma[k] = (float)(((double)cov[k] / kFixProd) / (double)n);
  float U[9], S[3], V[9];
  svd3(sigma, U, S, V);
  float sgn = (det3(U) * det3(V) < 0.f) ? -1.0f : 1.0f;
  float R[9];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) {
      float v = U[r * 3 + 0] * V[c * 3 + 0] + U[r * 3 + 1] * V[c * 3 + 1];
      R[r * 3 + c] = v + (U[r * 3 + 2] * sgn) * V[c * 3 + 2];
    }
  for (int r = 0; r < 3; ++r) {
    float rs = (R[r * 3 + 0] * ms[0] + R[r * 3 + 1] * ms[1]) + R[r * 3 + 2] * ms[2];
    T[r * 4 + 0] = R[r * 3 + 0]; T[r * 4 + 1] = R[r * 3 + 1]; T[r * 4 + 2] = R[r * 3 + 2];
    T[r * 4 + 3] = md[r] - rs;
  }
  T[12] = T[13] = T[14] = 0.f; T[15] = 1.f;
}

inline void mat4_mul(const float* A, const float* B, float* C) {
  float t[16];
  for (int r = 0; r < 4; ++r)
    for (int c = 0; c < 4; ++c) {
      float s = A[r * 4 + 0] * B[0 * 4 + c] + A[r * 4 + 1] * B[1 * 4 + c];
      s = s + A[r * 4 + 2] * B[2 * 4 + c];
      t[r * 4 + c] = s + A[r * 4 + 3] * B[3 * 4 + c];
    }
  std::memcpy(C, t, sizeof(t));
}

struct IcpParams {
  double max_corr_dist;      // loop_closure.cpp:7
  double trans_eps;          // :8  setTransformationEpsilon(0.01)
  double fit_eps;            // :9  setEuclideanFitnessEpsilon(0.01) -> DefaultConvergenceCriteria relative MSE
  int max_iter;              // :10
  double rot_eps;            // transformation_rotation_epsilon_ (0 -> 1 - trans_eps)
  double mse_abs;            // DefaultConvergenceCriteria::mse_threshold_absolute_ (1e-12)
  int max_similar;           // max_iterations_similar_transforms_ (0)
  int min_corr;              // Registration::min_number_correspondences_ (3)
};

struct IcpResult {
  float T[16];
  double score;
  int converged;
  int iterations;
  int state;  // 0 not converged, 1 iterations, 2 transform, 3 abs mse, 4 rel mse, 5 no correspondences
};

void icp_align(const std::vector<P4>& src, const std::vector<P4>& dst, const IcpParams& prm, int nthreads, IcpResult& res,
               std::vector<P4>* aligned, std::vector<int>* last_corr) {
  KdStatic tree;
  tree.build(dst);
  const int S = (int)src.size();
  std::vector<float> cur(size_t(S) * 3);
  for (int i = 0; i < S; ++i) { cur[3 * i] = src[i].x; cur[3 * i + 1] = src[i].y; cur[3 * i + 2] = src[i].z; }
  float Tfinal[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
  float Tr[16];
  std::memcpy(Tr, Tfinal, sizeof(Tr));
  const double max_d2 = prm.max_corr_dist * prm.max_corr_dist;
  const double rot_thr = prm.rot_eps > 0 ? prm.rot_eps : 1.0 - prm.trans_eps;
  double prev_mse = DBL_MAX;
  int similar = 0, iters = 0, state = 0;
  bool converged = false;
  std::vector<int> match(S);
  std::vector<float> md2(S);
  std::vector<float> cs, cd;
  while (true) {
#pragma omp parallel for num_threads(nthreads) schedule(static)
    for (int i = 0; i < S; ++i) {
      int bi; float bd;
      tree.nn(&cur[3 * i], bi, bd);
      if (bi < 0 || (double)bd > max_d2) bi = -1;
      match[i] = bi; md2[i] = bd;
    }
    cs.clear(); cd.clear();
    int64_t mse_fix = 0;
    size_t cnt = 0;
    for (int i = 0; i < S; ++i) {
      if (match[i] < 0) continue;
      const P4& d = dst[match[i]];
      cs.insert(cs.end(), {cur[3 * i], cur[3 * i + 1], cur[3 * i + 2]});
      cd.insert(cd.end(), {d.x, d.y, d.z});
      mse_fix += fixp((double)md2[i]);
      ++cnt;
    }
    if ((int)cnt < prm.min_corr) { state = 5; converged = false; break; }
    umeyama(cs, cd, Tr);
    for (int i = 0; i < S; ++i) {
      float o[3];
      tf_f(Tr, cur[3 * i], cur[3 * i + 1], cur[3 * i + 2], o);
      cur[3 * i] = o[0]; cur[3 * i + 1] = o[1]; cur[3 * i + 2] = o[2];
    }
    mat4_mul(Tr, Tfinal, Tfinal);
    ++iters;
    // DefaultConvergenceCriteria::hasConverged
    if (iters >= prm.max_iter) { state = 1; converged = true; break; }
    bool is_similar = false;
    double cos_angle = 0.5 * (((double)Tr[0] + (double)Tr[5] + (double)Tr[10]) - 1.0);
    double tr2 = ((double)Tr[3] * (double)Tr[3] + (double)Tr[7] * (double)Tr[7]) + (double)Tr[11] * (double)Tr[11];
    if (cos_angle >= rot_thr && tr2 <= prm.trans_eps) {
      if (similar >= prm.max_similar) { state = 2; converged = true; break; }
      is_similar = true;
    }
    double mse = ((double)mse_fix / kFixProd) / (double)cnt;
    if (std::fabs(mse - prev_mse) < prm.mse_abs) {
      if (similar >= prm.max_similar) { state = 3; converged = true; break; }
      is_similar = true;
    }
    if (std::fabs(mse - prev_mse) / prev_mse < prm.fit_eps) {
      if (similar >= prm.max_similar) { state = 4; converged = true; break; }
      is_similar = true;
    }
    if (is_similar) ++similar; else similar = 0;
    prev_mse = mse;
  }
  std::memcpy(res.T, Tfinal, sizeof(Tfinal));
  res.converged = converged ? 1 : 0;
  res.iterations = iters;
  res.state = state;
  if (last_corr) *last_corr = match;
  // aligned_ = transformCloud(*input_, final_transformation_)
  if (aligned) {
    aligned->resize(S);
    for (int i = 0; i < S; ++i) {
      float o[3];
      tf_f(Tfinal, src[i].x, src[i].y, src[i].z, o);
      (*aligned)[i] = {o[0], o[1], o[2], src[i].i};
    }
  }
  // getFitnessScore(): transformPointCloud(input, final) then mean squared 1-NN distance
  int64_t fs = 0;
  int nr = 0;
  std::vector<float> fd(S);
#pragma omp parallel for num_threads(nthreads) schedule(static)
  for (int i = 0; i < S; ++i) {
    float o[3];
    tf_f_sse(Tfinal, src[i].x, src[i].y, src[i].z, o);
    int bi; float bd;
    tree.nn(o, bi, bd);
    fd[i] = bi >= 0 ? bd : -1.0f;
  }
  for (int i = 0; i < S; ++i)
    if (fd[i] >= 0.f) { fs += fixp((double)fd[i]); ++nr; }
  res.score = nr > 0 ? ((double)fs / kFixProd) / (double)nr : DBL_MAX;
}

struct Keyframe {
  std::vector<P4> pcd;  // LiDAR frame (pose_pcd.hpp:39)
  double pose[16];      // pose_corrected_eig_, row-major
  double stamp;
};

struct LoopOracle {
  std::vector<Keyframe> kfs;
  IcpParams prm;
  int nthreads = 1;
};

}  // namespace

extern "C" {

void* orc_loop_create(double max_corr_dist, double trans_eps, double fit_eps, int max_iter, int nthreads) {
  LoopOracle* L = new LoopOracle();
  L->prm = {max_corr_dist, trans_eps, fit_eps, max_iter, 0.0, 1e-12, 0, 3};
  L->nthreads = nthreads > 0 ? nthreads : omp_get_max_threads();
  return L;
}
void orc_loop_destroy(void* h) { delete (LoopOracle*)h; }

int orc_loop_add_keyframe(void* h, const float* xyzi, int n, const double* pose16, double stamp) {
  LoopOracle* L = (LoopOracle*)h;
  Keyframe k;
  k.pcd.resize(n);
  std::memcpy(k.pcd.data(), xyzi, size_t(n) * 16);
  std::memcpy(k.pose, pose16, sizeof(k.pose));
  k.stamp = stamp;
  L->kfs.push_back(std::move(k));
  return (int)L->kfs.size() - 1;
}

// pose_corrected_eig_ = ... after a pose-graph update
void orc_loop_set_pose(void* h, int idx, const double* pose16) {
  LoopOracle* L = (LoopOracle*)h;
  std::memcpy(L->kfs[idx].pose, pose16, sizeof(L->kfs[idx].pose));
}

// loop_closure.cpp:18-40
int orc_loop_fetch_closest(void* h, int query_idx, double radius, double tdiff) {
  LoopOracle* L = (LoopOracle*)h;
  const Keyframe& q = L->kfs[query_idx];
  double shortest = radius * 3.0;
  int best = -1;
  for (size_t i = 0; i + 1 < L->kfs.size(); ++i) {
    const Keyframe& k = L->kfs[i];
    double dx = k.pose[3] - q.pose[3], dy = k.pose[7] - q.pose[7], dz = k.pose[11] - q.pose[11];
    double d = std::sqrt(dx * dx + dy * dy + dz * dz);
    if (radius > d && tdiff < (q.stamp - k.stamp)) {
      if (d < shortest) { shortest = d; best = (int)i; }
    }
  }
  return best;
}

// pose_eig_.inverse() of the PosePcd constructor (/root/reference/fast_lio_sam/include/pose_pcd.hpp:38): Eigen's fixed-size
// 4x4 inverse is a cofactor expansion [RECALL: Eigen/src/LU/InverseImpl.h, not in the mount]; restated here as adjugate /
// determinant in f64, laid out by 2x2 minors.  Returns 0 for a singular matrix.
int orc_pose_inverse(const double* m, double* o) {
  // minors of the lower and the upper two rows
  const double s0 = m[0] * m[5] - m[4] * m[1], s1 = m[0] * m[6] - m[4] * m[2], s2 = m[0] * m[7] - m[4] * m[3];
  const double s3 = m[1] * m[6] - m[5] * m[2], s4 = m[1] * m[7] - m[5] * m[3], s5 = m[2] * m[7] - m[6] * m[3];
  const double c5 = m[10] * m[15] - m[14] * m[11], c4 = m[9] * m[15] - m[13] * m[11], c3 = m[9] * m[14] - m[13] * m[10];
  const double c2 = m[8] * m[15] - m[12] * m[11], c1 = m[8] * m[14] - m[12] * m[10], c0 = m[8] * m[13] - m[12] * m[9];
  const double det = s0 * c5 - s1 * c4 + s2 * c3 + s3 * c2 - s4 * c1 + s5 * c0;
  if (det == 0.0 || !std::isfinite(det)) return 0;
  const double r = 1.0 / det;
  o[0] = (m[5] * c5 - m[6] * c4 + m[7] * c3) * r;
  o[1] = (-m[1] * c5 + m[2] * c4 - m[3] * c3) * r;
  o[2] = (m[13] * s5 - m[14] * s4 + m[15] * s3) * r;
  o[3] = (-m[9] * s5 + m[10] * s4 - m[11] * s3) * r;
  o[4] = (-m[4] * c5 + m[6] * c2 - m[7] * c1) * r;
  o[5] = (m[0] * c5 - m[2] * c2 + m[3] * c1) * r;
  o[6] = (-m[12] * s5 + m[14] * s2 - m[15] * s1) * r;
  o[7] = (m[8] * s5 - m[10] * s2 + m[11] * s1) * r;
  o[8] = (m[4] * c4 - m[5] * c2 + m[7] * c0) * r;
  o[9] = (-m[0] * c4 + m[1] * c2 - m[3] * c0) * r;
  o[10] = (m[12] * s4 - m[13] * s2 + m[15] * s0) * r;
  o[11] = (-m[8] * s4 + m[9] * s2 - m[11] * s0) * r;
  o[12] = (-m[4] * c3 + m[5] * c1 - m[6] * c0) * r;
  o[13] = (m[0] * c3 - m[1] * c1 + m[2] * c0) * r;
  o[14] = (-m[12] * s3 + m[13] * s1 - m[14] * s0) * r;
  o[15] = (m[8] * s3 - m[9] * s1 + m[10] * s0) * r;
  return 1;
}

void orc_transform_pcd(const float* in, int n, const double* T, float* out) {
  for (int i = 0; i < n; ++i) {
    P4 p = {in[4 * i], in[4 * i + 1], in[4 * i + 2], in[4 * i + 3]};
    P4 o = tf_d(T, p);
    out[4 * i] = o.x; out[4 * i + 1] = o.y; out[4 * i + 2] = o.z; out[4 * i + 3] = o.i;
  }
}

// utilities.hpp:158-171; returns number of output points (out must hold n)
int orc_voxelize(const float* in, int n, float leaf, float* out) {
  std::vector<P4> v(n), o;
  std::memcpy(v.data(), in, size_t(n) * 16);
  voxel_grid(v, leaf, o);
  std::memcpy(out, o.data(), o.size() * 16);
  return (int)o.size();
}

static void build_submap(LoopOracle* L, int center, int range, float leaf, std::vector<P4>& out) {
  std::vector<P4> acc;
  const int K = (int)L->kfs.size();
  for (int i = center - range; i < center + range + 1; ++i) {
    if (i >= 0 && i < K - 1) {  // loop_closure.cpp:54 / :61 — the newest keyframe is excluded
      const Keyframe& k = L->kfs[i];
      for (const P4& p : k.pcd) acc.push_back(tf_d(k.pose, p));
    }
  }
  voxel_grid(acc, leaf, out);
}

// setSrcAndDstCloud; buffers sized by the caller (cap points each); returns sizes
void orc_loop_set_src_dst(void* h, int src_idx, int dst_idx, int range, float leaf, float* src_out, int* n_src, float* dst_out,
                          int* n_dst, int cap) {
  LoopOracle* L = (LoopOracle*)h;
  std::vector<P4> s, d;
  build_submap(L, src_idx, range, leaf, s);
  build_submap(L, dst_idx, range, leaf, d);
  *n_src = (int)s.size(); *n_dst = (int)d.size();
  if (src_out) std::memcpy(src_out, s.data(), std::min<size_t>(s.size(), cap) * 16);
  if (dst_out) std::memcpy(dst_out, d.data(), std::min<size_t>(d.size(), cap) * 16);
}

// icpAlignment on explicit clouds.  out: T[16] f32 row-major, score, converged, iterations, state
void orc_icp_align(void* h, const float* src, int ns, const float* dst, int nd, float* T, double* score, int* converged,
                   int* iterations, int* state, float* aligned_out, int* corr_out) {
  LoopOracle* L = (LoopOracle*)h;
  std::vector<P4> s(ns), d(nd), al;
  std::memcpy(s.data(), src, size_t(ns) * 16);
  std::memcpy(d.data(), dst, size_t(nd) * 16);
  IcpResult r;
  std::vector<int> corr;
  icp_align(s, d, L->prm, L->nthreads, r, aligned_out ? &al : nullptr, corr_out ? &corr : nullptr);
  std::memcpy(T, r.T, sizeof(r.T));
  *score = r.score; *converged = r.converged; *iterations = r.iterations; *state = r.state;
  if (aligned_out) std::memcpy(aligned_out, al.data(), al.size() * 16);
  if (corr_out) std::memcpy(corr_out, corr.data(), corr.size() * 4);
}

// performLoopClosure(query, keyframes, closest) = setSrcAndDstCloud + icpAlignment + score gate
void orc_loop_verify(void* h, int src_idx, int dst_idx, int range, float leaf, double score_thr, float* T, double* score,
                     int* converged, int* valid, int* iterations, int* n_src, int* n_dst) {
  LoopOracle* L = (LoopOracle*)h;
  std::vector<P4> s, d;
  build_submap(L, src_idx, range, leaf, s);
  build_submap(L, dst_idx, range, leaf, d);
  IcpResult r;
  icp_align(s, d, L->prm, L->nthreads, r, nullptr, nullptr);
  std::memcpy(T, r.T, sizeof(r.T));
  *score = r.score; *converged = r.converged; *iterations = r.iterations;
  *valid = (r.converged && r.score < score_thr) ? 1 : 0;  // loop_closure.cpp:85
  *n_src = (int)s.size(); *n_dst = (int)d.size();
}

void orc_svd3(const float* M, float* U, float* S, float* V) { svd3(M, U, S, V); }

}  // extern "C"

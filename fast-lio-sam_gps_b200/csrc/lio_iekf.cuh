// Device-resident half of esekfom::update_iterated_dyn_share_modified: the 23x23 FP64 algebra of one IEKF iteration,
// run by the LAST block of the measurement reduction (k_plane_reduce<.., true>) right after it has assembled
// H^T H / H^T z, so a whole scan (up to max_iter+1 model calls + map_incremental) is enqueued without a host round trip.
// Reference: third_party/FAST_LIO/include/IKFoM_toolkit/esekfom/esekfom.hpp (absent submodule).
//
// The arithmetic is that of esekf_host.cpp, operation for operation (same headers: esekf_math.hpp / esekf_state.hpp;
// --fmad=false): the serial manifold pieces run on one thread, the dense pieces are spread over the block with every
// output element computed by one thread in the host's summation order, and the Gauss-Jordan inverse keeps the host's
// pivot choice (first row of maximal magnitude) and per-element update expression.
#pragma once
#include "esekf_state.hpp"
#include "lio_device.cuh"

namespace lio {

constexpr int kIekfMaxCalls = 16;  // max_iter + 1 model calls the device loop can trace
constexpr int kN = esekf::kDof;
constexpr int kNN = kN * kN;

// device-side loop state (global memory)
struct IekfCtl {
  double x[36];        // current state: every kernel of the loop reads it from here
  double x_prop[36];   // propagated state (input)
  double P_prop[kNN];  // propagated covariance (input)
  double limit[kN];
  double R;
  int max_iter;
  int converge;  // ekfom_data.converge of the NEXT model call (1: redo kNN)
  int t;         // converged-iteration counter
  int done;      // model calls made
  int finished;  // 1: the remaining enqueued calls of this scan are skipped
  int pad;
};

// mapped pinned block: the host writes the inputs, the device the outputs and, last, `seq`
struct IekfIO {
  double state_in[36];
  double P_in[kNN];
  double limit[kN];
  double R;
  int max_iter;
  int pad0;
  double state_out[36];
  double P_out[kNN];
  double iter_states[kIekfMaxCalls * 36];
  int n_eff[kIekfMaxCalls];
  int knn[kIekfMaxCalls];
  int iters_done;
  int rc;  // LIO_OK or LIO_E_ARG (singular matrix)
  unsigned long long seq;
};

#ifdef __CUDACC__

struct IekfSmem {
  double P[kNN], A[kNN], B[kNN], Kx[kNN], L[kNN];
  double dx[kN], dxn[kN], dxu[kN], Kh[kN];
  double J[3][9];  // block Jacobians: rot, offset_R (3x3), grav (2x2)
  double xs[36];
  int mode;  // 0 valid round, 1 invalid round
  int fin;   // 0 go on, 1 finish with covariance update, 2 finish without (every remaining round invalid)
};

// rows [idx, idx+D) of M (23 columns) <- J * rows, column c only (esekf::left_block, one column per thread)
__device__ __forceinline__ void blk_left_col(double* M, int idx, int D, const double* J, int c) {
  double t[3];
  for (int r = 0; r < D; ++r) {
    t[r] = 0;
    for (int k = 0; k < D; ++k) t[r] += J[r * D + k] * M[(idx + k) * kN + c];
  }
  for (int r = 0; r < D; ++r) M[(idx + r) * kN + c] = t[r];
}
// columns [idx, idx+D) of M <- cols * J^T, row r only (esekf::right_block, one row per thread)
__device__ __forceinline__ void blk_right_row(double* M, int idx, int D, const double* J, int r) {
  double t[3];
  for (int c = 0; c < D; ++c) {
    t[c] = 0;
    for (int k = 0; k < D; ++k) t[c] += M[r * kN + idx + k] * J[c * D + k];
  }
  for (int c = 0; c < D; ++c) M[r * kN + idx + c] = t[c];
}

// Inv <- A^-1 (A is destroyed).  Block-wide Gauss-Jordan with partial pivoting; returns false (uniformly) when singular.
template <int NT>
__device__ bool gj_invert(double* A, double* Inv) {
  constexpr int kElems = kN * 2 * kN;
  constexpr int kPer = (kElems + NT - 1) / NT;
  const int tid = threadIdx.x, lane = tid & 31;
  for (int e = tid; e < kNN; e += NT) Inv[e] = (e / kN == e % kN) ? 1.0 : 0.0;
  __syncthreads();
  for (int c = 0; c < kN; ++c) {
    // pivot: first row r >= c of maximal |A[r][c]| (every warp finds it for itself)
    int pr = c + lane;
    double pv = (pr < kN) ? fabs(A[pr * kN + c]) : -1.0;
#pragma unroll
    for (int off = 16; off; off >>= 1) {
      const double ov = __shfl_xor_sync(0xFFFFFFFFu, pv, off);
      const int orow = __shfl_xor_sync(0xFFFFFFFFu, pr, off);
      if (ov > pv || (ov == pv && orow < pr)) {
        pv = ov;
        pr = orow;
      }
    }
    if (pv == 0.0) return false;
    const int piv = pr;
    const double d = 1.0 / A[piv * kN + c];
    double nv[kPer];
#pragma unroll
    for (int j = 0; j < kPer; ++j) {
      const int e = tid + j * NT;
      nv[j] = 0.0;
      if (e < kElems) {
        const int r = e / (2 * kN), k = e - r * 2 * kN;
        const double* M = (k < kN) ? A : Inv;
        const int kk = (k < kN) ? k : k - kN;
        const double pc = M[piv * kN + kk] * d;  // row c after the swap and the scaling
        if (r == c) {
          nv[j] = pc;
        } else {
          const int sr = (r == piv) ? c : r;  // the row that sits at r after the swap
          const double f = A[sr * kN + c];
          const double old = M[sr * kN + kk];
          nv[j] = (f == 0.0) ? old : old - f * pc;
        }
      }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < kPer; ++j) {
      const int e = tid + j * NT;
      if (e < kElems) {
        const int r = e / (2 * kN), k = e - r * 2 * kN;
        if (k < kN) A[r * kN + k] = nv[j];
        else Inv[r * kN + (k - kN)] = nv[j];
      }
    }
    __syncthreads();
  }
  return true;
}

// Ships the outputs of the scan to the mapped host block; `seq` goes last.
template <int NT>
__device__ void iekf_publish(IekfCtl* ctl, IekfIO* io, const double* x36, const double* P_out /*shared or global*/, int rc,
                             unsigned long long seq) {
  const int tid = threadIdx.x;
  for (int e = tid; e < 36; e += NT) io->state_out[e] = x36[e];
  for (int e = tid; e < kNN; e += NT) io->P_out[e] = P_out[e];
  if (tid == 0) {
    io->iters_done = ctl->done;
    io->rc = rc;
    ctl->finished = 1;
  }
  __threadfence_system();
  __syncthreads();
  if (tid == 0) *reinterpret_cast<volatile unsigned long long*>(&io->seq) = seq;
}

// One iteration of the IEKF update, executed by a whole block of NT threads.  hth: HTH[144] | HTz[12] | res_sum | n_eff
// (shared memory).  iter_i runs -1 .. max_iter-1 as in esekfom.  sm may alias any shared memory the caller no longer needs.
template <int NT>
__device__ __noinline__ void iekf_step_block(IekfCtl* ctl, IekfIO* io, const double* hth, int iter_i, unsigned long long seq, IekfSmem& sm) {
  using namespace esekf;
  const int tid = threadIdx.x;
  const int max_iter = ctl->max_iter;
  const int done0 = ctl->done;
  const int n_eff = (int)hth[157];
  __syncthreads();  // everyone has read ctl->done before thread 0 bumps it
  if (tid == 0) {
    io->n_eff[done0] = n_eff;
    io->knn[done0] = ctl->converge;
  }
  if (n_eff < 1) {  // ekfom_data.valid == false: the round is skipped; state, converge flag and t stay
    for (int e = tid; e < 36; e += NT) io->iter_states[36 * done0 + e] = ctl->x[e];
    __syncthreads();
    if (tid == 0) ctl->done = done0 + 1;
    __syncthreads();
    if (iter_i == max_iter - 1) iekf_publish<NT>(ctl, io, ctl->x, ctl->P_prop, 0, seq);  // covariance stays as propagated
    return;
  }

  // ---- serial manifold part (thread 0) while the rest of the block copies P_prop
  for (int e = tid; e < kNN; e += NT) sm.P[e] = ctl->P_prop[e];
  if (tid == 0) {
    const State x = load(ctl->x), x_prop = load(ctl->x_prop);
    double dx[kDof];
    state_boxminus(x, x_prop, dx);
    for (int k = 0; k < kDof; ++k) sm.dx[k] = sm.dxn[k] = dx[k];
    for (int b = 0; b < 2; ++b) {
      const int idx = 3 + 3 * b;
      const Mat3 J = transpose(a_matrix(Vec3{{dx[idx], dx[idx + 1], dx[idx + 2]}}));
      const Vec3 seg = mul(J, Vec3{{dx[idx], dx[idx + 1], dx[idx + 2]}});
      for (int k = 0; k < 3; ++k) sm.dxn[idx + k] = seg[k];
      for (int k = 0; k < 9; ++k) sm.J[b][k] = J.m[k];
    }
    double J2[4];
    s2_jacobian(x.grav, x_prop.grav, dx + 21, J2);
    const double a = dx[21], b2 = dx[22];
    sm.dxn[21] = J2[0] * a + J2[1] * b2;
    sm.dxn[22] = J2[2] * a + J2[3] * b2;
    for (int k = 0; k < 4; ++k) sm.J[2][k] = J2[k];
  }
  __syncthreads();
  // P <- J P J^T block by block, in the host's order (left then right, blocks rot, offset_R, grav)
  for (int b = 0; b < 3; ++b) {
    const int idx = (b == 2) ? 21 : 3 + 3 * b, D = (b == 2) ? 2 : 3;
    if (tid < kN) blk_left_col(sm.P, idx, D, sm.J[b], tid);
    __syncthreads();
    if (tid < kN) blk_right_row(sm.P, idx, D, sm.J[b], tid);
    __syncthreads();
  }
  // information form: B = (P/R)^-1 ; B[:12,:12] += HTH ; A = B^-1
  {
    const double rinv = 1.0 / ctl->R;
    for (int e = tid; e < kNN; e += NT) sm.A[e] = sm.P[e] * rinv;
  }
  __syncthreads();
  bool ok = gj_invert<NT>(sm.A, sm.B);
  if (ok) {
    if (tid < 144) sm.B[(tid / 12) * kN + (tid % 12)] += hth[tid];
    __syncthreads();
    ok = gj_invert<NT>(sm.B, sm.A);
  }
  if (!ok) {  // singular covariance / information matrix: the host loop returns LIO_E_ARG here
    __syncthreads();
    iekf_publish<NT>(ctl, io, ctl->x, ctl->P_prop, -2, seq);
    return;
  }
  // K_h = Pi[:, :12] HTz ; K_x[:, :12] = Pi[:, :12] HTH ; the other columns of K_x are zero
  for (int e = tid; e < kNN; e += NT) {
    const int r = e / kN, cc = e - r * kN;
    double q = 0.0;
    if (cc < 12) {
      q = 0;
      for (int k = 0; k < 12; ++k) q += sm.A[r * kN + k] * hth[k * 12 + cc];
    }
    sm.Kx[e] = q;
  }
  if (tid < kN) {
    double s = 0;
    for (int k = 0; k < 12; ++k) s += sm.A[tid * kN + k] * hth[144 + k];
    sm.Kh[tid] = s;
  }
  __syncthreads();
  if (tid < kN) {
    double s = sm.Kh[tid];
    for (int k = 0; k < kN; ++k) s += (sm.Kx[tid * kN + k] - (tid == k ? 1.0 : 0.0)) * sm.dxn[k];
    sm.dxu[tid] = s;
  }
  __syncthreads();
  if (tid == 0) {
    State x = load(ctl->x);
    const State x_prop = load(ctl->x_prop);
    double dxu[kDof];
    for (int k = 0; k < kDof; ++k) dxu[k] = sm.dxu[k];
    state_boxplus(x, dxu);
    store(x, sm.xs);
    bool converge = true;
    for (int k = 0; k < kDof; ++k)
      if (fabs(dxu[k]) > ctl->limit[k]) {
        converge = false;
        break;
      }
    int t = ctl->t;
    if (converge) ++t;
    if (!t && iter_i == max_iter - 2) converge = true;
    ctl->t = t;
    ctl->converge = converge ? 1 : 0;
    ctl->done = done0 + 1;
    const int fin = (t > 1 || iter_i == max_iter - 1) ? 1 : 0;
    sm.fin = fin;
    if (fin) {
      for (int b = 0; b < 2; ++b) {
        const int idx = 3 + 3 * b;
        const Mat3 J = transpose(a_matrix(Vec3{{dxu[idx], dxu[idx + 1], dxu[idx + 2]}}));
        for (int k = 0; k < 9; ++k) sm.J[b][k] = J.m[k];
      }
      double J2[4];
      s2_jacobian(x.grav, x_prop.grav, dxu + 21, J2);  // Mx at x_propagated, as esekf_host.cpp
      for (int k = 0; k < 4; ++k) sm.J[2][k] = J2[k];
    }
  }
  __syncthreads();
  for (int e = tid; e < 36; e += NT) {
    const double v = sm.xs[e];
    ctl->x[e] = v;
    io->iter_states[36 * done0 + e] = v;
  }
  if (!sm.fin) {
    __threadfence();
    return;
  }
  // covariance update: P_out = L - K_x[:, :12] * P[:12, :] with the manifold re-projections
  for (int e = tid; e < kNN; e += NT) sm.L[e] = sm.P[e];
  __syncthreads();
  const int warp = tid >> 5, lane = tid & 31;
  for (int b = 0; b < 3; ++b) {
    const int idx = (b == 2) ? 21 : 3 + 3 * b, D = (b == 2) ? 2 : 3;
    // independent matrices in parallel: left(L) | left(K_x) | right(P); then right(L)
    if (warp == 0 && lane < kN) blk_left_col(sm.L, idx, D, sm.J[b], lane);
    if (warp == 1 && lane < kN) blk_left_col(sm.Kx, idx, D, sm.J[b], lane);
    if (warp == 2 && lane < kN) blk_right_row(sm.P, idx, D, sm.J[b], lane);
    __syncthreads();
    if (warp == 0 && lane < kN) blk_right_row(sm.L, idx, D, sm.J[b], lane);
    __syncthreads();
  }
  for (int e = tid; e < kNN; e += NT) {
    const int r = e / kN, cc = e - r * kN;
    double s = 0;
    for (int k = 0; k < 12; ++k) s += sm.Kx[r * kN + k] * sm.P[k * kN + cc];
    sm.A[e] = sm.L[e] - s;
  }
  __syncthreads();
  iekf_publish<NT>(ctl, io, sm.xs, sm.A, 0, seq);
}

#endif  // __CUDACC__

}  // namespace lio

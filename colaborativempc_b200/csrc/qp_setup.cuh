// qp_setup.cuh - per-QP problem build on chip: planes + weights, LPV linearisation, prediction rows, H and g.
//
// Reference functions replaced (paths relative to planner/lib/plan_lib/):
//   planes/compute_plane.py:41-68 (compute_hyperplane, keep_sign=True), utilities/misc.py:10-26 (compute_weights),
//   distributedPlanner/LPV_Planner.py:477-591 (_EstimateABC), utilities/misc.py:78-126 (curvature, get_ey),
//   LPV_Planner.py:382-427 (_buildMatCost) and :429-475 (_buildMatEqConst, eliminated here by condensation).
#pragma once
#include "qp_engine.cuh"

// planes (N,3,nn) + distance weights folded into the linear cost of x,y (LPV_Planner.py:416-422).
// Returns nonzero if ego and a neighbour coincide (compute_plane.py:52 divides by zero there).
template <class CT>
__device__ __forceinline__ int setup_planes(const CT& c, const DevParams& P, const double* nbr, const double* pose,
                                            double* planes_out) {
  int bad = 0;
  for (int h = threadIdx.x; h < c.N; h += QP_THREADS) {
    const double ex = pose[2 * h], ey = pose[2 * h + 1];
    const double fx = pose[2 * (h + 1)], fy = pose[2 * (h + 1) + 1];
    double sx = 0.0, sy = 0.0;
    for (int n = 0; n < c.nn; ++n) {
      double ax = 0.0, ay = 0.0, bb = 0.0;
      if (n < c.cnt) {
        const double gx = nbr[(h * c.nn + n) * 2], gy = nbr[(h * c.nn + n) * 2 + 1];
        ax = gx - ex; ay = gy - ey;
        const double nrm = sqrt(ax * ax + ay * ay);
        if (!(nrm > 0.0)) bad = 1;
        ax /= nrm; ay /= nrm;
        bb = -0.5 * (ax * (ex + gx) + ay * (ey + gy));
        const double dx = fx - nbr[((h + 1) * c.nn + n) * 2], dy = fy - nbr[((h + 1) * c.nn + n) * 2 + 1];
        const double w = (2.0 * P.min_dist - sqrt(dx * dx + dy * dy)) / (double)c.cnt;  // rows 1..N (misc.py:15)
        sx = fma(w, ax, sx); sy = fma(w, ay, sy);
      }
      c.ax[h * c.nn + n] = ax; c.ay[h * c.nn + n] = ay;
      c.hh[h * c.nn + n] = -0.5 * P.min_dist - bb;
      if (planes_out) {
        planes_out[(h * 3 + 0) * c.nn + n] = ax;
        planes_out[(h * 3 + 1) * c.nn + n] = ay;
        planes_out[(h * 3 + 2) * c.nn + n] = bb;
      }
    }
    c.pxy[2 * h] = P.wq * sx; c.pxy[2 * h + 1] = P.wq * sy;
  }
  return bad;
}

// A_k = I + dt*Ac, B_k = dt*Bc along the previous prediction (LPV_Planner.py:491-589).  Only the first three
// columns of Ac are non-zero, so A_k is stored as its 9x3 column block.  Returns nonzero on a failed map lookup.
template <class CT>
__device__ __forceinline__ int setup_linearise(const CT& c, const DevParams& P, const double* xprev, int rows,
                                               const double* uprev) {
  int bad = 0;
  for (int k = threadIdx.x; k < max(c.N, rows); k += QP_THREADS) {
    if (k >= c.N) {  // get_ey is evaluated on every row of Last_xPredicted (LPV_Planner.py:489)
      if (map_segment(P, xprev[k * NSTATE + 6]) < 0) bad = 1;
      continue;
    }
    const int re = k < rows ? k : rows - 1;
    const int sege = map_segment(P, xprev[re * NSTATE + 6]);
    const double* x = xprev + k * NSTATE;
    const double vx = x[0], vy = x[1], ey = x[3], epsi = x[4], theta = x[5];
    const int seg = map_segment(P, x[6]);
    if (seg < 0 || sege < 0) { bad = 1; continue; }
    c.eyb[k] = P.seg_hw[sege];  // get_ey's own sm=1.0; the controller's sm is never applied (misc.py:105)
    const double cur = P.seg_curv[seg];
    const double delta = uprev[2 * k];
    double sd, cd, se, ce, st, ct;
    sincos(delta, &sd, &cd); sincos(epsi, &se, &ce); sincos(theta, &st, &ct);
    double A12 = 0, A13 = 0, A22 = 0, A23 = 0, A32 = 0, A33 = 0, B11 = 0;
    if (!(vx < 0.2)) {
      A12 = (sd * P.Cf) / (P.m * vx);
      A13 = (sd * P.Cf * P.lf) / (P.m * vx) + vy;
      A22 = -(P.Cr + P.Cf * cd) / (P.m * vx);
      A23 = -(P.lf * P.Cf * cd - P.lr * P.Cr) / (P.m * vx) - vx;
      A32 = -(P.lf * P.Cf * cd - P.lr * P.Cr) / (P.I * vx);
      A33 = -(P.lf * P.lf * P.Cf * cd + P.lr * P.lr * P.Cr) / (P.I * vx);
      B11 = -(sd * P.Cf) / P.m;
    }
    const double den = 1.0 - ey * cur;
    const double iden = 1.0 / den;
    const double Ac[27] = {-P.mu, A12, A13,
                           0.0, A22, A23,
                           0.0, A32, A33,
                           se, ce, 0.0,
                           iden * (-ce * cur), iden * (se * cur), 1.0,
                           0.0, 0.0, 1.0,
                           ce / den, -se / den, 0.0,
                           ct, -st, 0.0,
                           st, ct, 0.0};
    double* A = c.Acol + 27 * k;
#pragma unroll
    for (int r = 0; r < 9; ++r)
#pragma unroll
      for (int j = 0; j < 3; ++j) A[r * 3 + j] = (r == j ? 1.0 : 0.0) + P.dt * Ac[r * 3 + j];
    double* B = c.Bcol + 4 * k;
    B[0] = P.dt * B11;
    B[1] = P.dt * ((cd * P.Cf) / P.m);
    B[2] = P.dt * ((P.lf * P.Cf * cd) / P.I);
    B[3] = P.dt;
  }
  return bad;
}

__device__ __forceinline__ void lpv_step(const double* A, double* v) {  // v <- A_k v with A_k = [Acol | e_3..e_8]
  const double v0 = v[0], v1 = v[1], v2 = v[2];
#pragma unroll
  for (int r = 0; r < 9; ++r) {
    double s = (r >= 3) ? v[r] : 0.0;
    s = fma(A[r * 3 + 0], v0, s); s = fma(A[r * 3 + 1], v1, s); s = fma(A[r * 3 + 2], v2, s);
    v[r] = s;
  }
}

// Prediction rows: x_t = cr_t + sum_j Gr_t[:, j] u_j for the kept state rows (one thread per input column,
// plus one thread for the free response from x0).
__device__ __forceinline__ void setup_prediction(const Ctx& c, const DevParams& P, const double* x0) {
  for (int j = threadIdx.x; j <= c.nu; j += QP_THREADS) {
    double v[9];
    if (j == c.nu) {
#pragma unroll
      for (int r = 0; r < 9; ++r) v[r] = x0[r];
      for (int t = 0; t < c.N; ++t) {
        lpv_step(c.Acol + 27 * t, v);
        for (int r = 0; r < c.NR; ++r) c.cr[t * c.NR + r] = v[P.rowsel[r]];
      }
    } else {
      const int k = j >> 1;
      for (int t = 0; t < k; ++t)
        for (int r = 0; r < c.NR; ++r) GR(c, t, r, j) = 0.0;
#pragma unroll
      for (int r = 0; r < 9; ++r) v[r] = 0.0;
      const double* B = c.Bcol + 4 * k;
      if (j & 1) v[0] = B[3]; else { v[0] = B[0]; v[1] = B[1]; v[2] = B[2]; }
      for (int t = k; t < c.N; ++t) {
        if (t > k) lpv_step(c.Acol + 27 * t, v);
        for (int r = 0; r < c.NR; ++r) GR(c, t, r, j) = v[P.rowsel[r]];
      }
    }
  }
}

// H = 2 (Gamma' Qbar Gamma + Rbar + D' dRbar D),  g = 2 Gamma' (Qbar c + p) - 2 D' dRbar d0   (condensed cost).
// `ql` is an N x NR scratch (uses c.Zb).
__device__ __forceinline__ void setup_cost(const Ctx& c, const DevParams& P, const double* uold) {
  double* ql = c.Zb;
  for (int t = threadIdx.x; t < c.N; t += QP_THREADS) {
    double* q = ql + t * c.NR;
    for (int r = 0; r < c.NR; ++r) q[r] = 0.0;
    const double* ct = c.cr + t * c.NR;
    for (int e = 0; e < P.nq; ++e) {
      const int a = P.qa[e], b = P.qb[e];
      q[a] = fma(P.qv[e], ct[b], q[a]);
      if (a != b) q[b] = fma(P.qv[e], ct[a], q[b]);
    }
    q[0] += -P.vx_ref * P.Q00;      // LPV_Planner.py:412
    q[2] += c.pxy[2 * t];           // :421
    q[3] += c.pxy[2 * t + 1];       // :422
  }
  __syncthreads();
  for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) {
    double s = 0.0;
    for (int t = j >> 1; t < c.N; ++t)
      for (int r = 0; r < c.NR; ++r) s = fma(GR(c, t, r, j), ql[t * c.NR + r], s);
    s *= 2.0;
    if (j < 2) s -= 2.0 * (P.dR[j * 2 + 0] * uold[0] + P.dR[j * 2 + 1] * uold[1]);
    c.g[j] = s;
  }
  const int ntri = c.nu * (c.nu + 1) / 2;
  for (int e = threadIdx.x; e < ntri; e += QP_THREADS) {
    int i = (int)((sqrt(8.0 * e + 1.0) - 1.0) * 0.5);
    while ((i + 1) * (i + 2) / 2 <= e) ++i;
    while (i * (i + 1) / 2 > e) --i;
    const int j = e - i * (i + 1) / 2;
    double s = 0.0;
    for (int t = i >> 1; t < c.N; ++t)
      for (int q = 0; q < P.nq; ++q) {
        const int a = P.qa[q], b = P.qb[q];
        double term = GR(c, t, a, i) * GR(c, t, b, j);
        if (a != b) term = fma(GR(c, t, b, i), GR(c, t, a, j), term);
        s = fma(P.qv[q], term, s);
      }
    s *= 2.0;
    const int ki = i >> 1, kj = j >> 1, ai = i & 1, aj = j & 1;
    if (ki == kj) {
      s += 2.0 * (P.R[ai * 2 + aj] + P.dR[ai * 2 + aj]);
      if (ki < c.N - 1) s += 2.0 * P.dR[ai * 2 + aj];
    } else if (ki == kj + 1) {
      s -= 2.0 * P.dR[ai * 2 + aj];
    }
    c.H[j * c.ldk + i] = s;
    c.H[i * c.ldk + j] = s;
  }
  __syncthreads();
}

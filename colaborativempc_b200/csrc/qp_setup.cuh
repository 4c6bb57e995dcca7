// qp_setup.cuh - per-QP problem build on chip: planes + weights, LPV linearisation, prediction rows, H and g.
//
// Reference functions replaced (paths relative to planner/lib/plan_lib/):
//   planes/compute_plane.py:41-68 (compute_hyperplane, keep_sign=True), utilities/misc.py:10-26 (compute_weights),
//   distributedPlanner/LPV_Planner.py:477-591 (_EstimateABC), utilities/misc.py:78-126 (curvature, get_ey),
//   (the condensed engine's prediction rows and cost, LPV_Planner.py:382-475, are in qp2_kernel.cuh).
#pragma once
#include "qp_common.cuh"

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

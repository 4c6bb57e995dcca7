// lpv_sim.cuh - batched LPV model roll-out.
//
// Replaces planner/lib/plan_lib/distributedPlanner/LPV_val.py:77-94 (LPV_Model.sim) with its own linearisation
// LPV_val.py:96-195 (_EstimateABC of that file: the planner's formulas WITHOUT the "vx < 0.2" low-velocity branch):
//   A_k, B_k along xPred (rows 0..N-1) and u;  x_{k+1} = A_k x_k + B_k u_k from x_0 = xPred[0];
//   result = [x_1 .. x_N, ones(9)]  (np.delete(xPred, 0, 0) + a row of ones).
// get_ey is evaluated on every row of xPred by the reference (its value is unused here, a failed lookup raises).
//
// One warp per scenario: lane k linearises stage k (rows of xPred are contiguous, so a warp reads them coalesced) into
// shared memory, then lanes 0..8 carry one state each through the N sequential steps (the three non-trivial columns of
// A_k are broadcast by shuffle) and write each predicted row as one 72-byte segment.  HBM traffic = compulsory:
// (N+1)*72 + N*16 bytes read, (N+1)*72 written per scenario.
#pragma once
#include "qp_common.cuh"

#define SIM_WARPS 4

__global__ void __launch_bounds__(32 * SIM_WARPS) lpv_sim_kernel(const __grid_constant__ DevParams P, int B,
                                                                   const double* __restrict__ xpred, const double* __restrict__ u,
                                                                   double* __restrict__ out, int* __restrict__ status) {
  extern __shared__ __align__(16) double sim_sm[];  // per warp: N x (27 + 4)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * SIM_WARPS + warp;
  if (b >= B) return;
  const int N = P.N;
  double* AB = sim_sm + (size_t)warp * N * 31;
  const double* xp = xpred + (size_t)b * (N + 1) * NSTATE;
  const double* ub = u + (size_t)b * N * 2;
  double* o = out + (size_t)b * (N + 1) * NSTATE;
  int bad = 0;
  for (int k = lane; k <= N; k += 32) {
    const double* x = xp + k * NSTATE;
    const int seg = map_segment(P, x[6]);
    if (seg < 0) { bad = 1; continue; }
    if (k == N) continue;  // row N: only the get_ey lookup
    const double vx = x[0], vy = x[1], ey = x[3], epsi = x[4], theta = x[5];
    const double cur = P.seg_curv[seg];
    const double delta = ub[2 * k];
    double sd, cd, se, ce, st, ct;
    sincos(delta, &sd, &cd); sincos(epsi, &se, &ce); sincos(theta, &st, &ct);
    const double A12 = (sd * P.Cf) / (P.m * vx);
    const double A13 = (sd * P.Cf * P.lf) / (P.m * vx) + vy;
    const double A22 = -(P.Cr + P.Cf * cd) / (P.m * vx);
    const double A23 = -(P.lf * P.Cf * cd - P.lr * P.Cr) / (P.m * vx) - vx;
    const double A32 = -(P.lf * P.Cf * cd - P.lr * P.Cr) / (P.I * vx);
    const double A33 = -(P.lf * P.lf * P.Cf * cd + P.lr * P.lr * P.Cr) / (P.I * vx);
    const double B11 = -(sd * P.Cf) / P.m;
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
    double* A = AB + 31 * k;
#pragma unroll
    for (int r = 0; r < 9; ++r)
#pragma unroll
      for (int j = 0; j < 3; ++j) A[r * 3 + j] = (r == j ? 1.0 : 0.0) + P.dt * Ac[r * 3 + j];
    A[27] = P.dt * B11;
    A[28] = P.dt * ((cd * P.Cf) / P.m);
    A[29] = P.dt * ((P.lf * P.Cf * cd) / P.I);
    A[30] = P.dt;
  }
  bad = __any_sync(0xffffffffu, bad);
  __syncwarp();
  if (bad) {  // the reference raises (misc.py:97: int() of an empty index)
    const double qnan = __longlong_as_double(0x7ff8000000000000ULL);
    for (int j = lane; j < (N + 1) * NSTATE; j += 32) o[j] = qnan;
    if (lane == 0 && status) status[b] = DMPC_MAP_LOOKUP_FAILED;
    return;
  }
  // lanes 0..8: state r of the roll-out
  const int r = lane < NSTATE ? lane : 0;
  double v = xp[r];
#pragma unroll 1
  for (int k = 0; k < N; ++k) {
    const double* A = AB + 31 * k;
    const double v0 = __shfl_sync(0xffffffffu, v, 0), v1 = __shfl_sync(0xffffffffu, v, 1), v2 = __shfl_sync(0xffffffffu, v, 2);
    double s = (r >= 3) ? v : 0.0;
    s = fma(A[r * 3 + 0], v0, s); s = fma(A[r * 3 + 1], v1, s); s = fma(A[r * 3 + 2], v2, s);
    const double dl = ub[2 * k], ac = ub[2 * k + 1];
    if (r == 0) s = fma(A[27], dl, fma(A[30], ac, s));
    else if (r == 1) s = fma(A[28], dl, s);
    else if (r == 2) s = fma(A[29], dl, s);
    v = s;
    if (lane < NSTATE) o[k * NSTATE + lane] = v;  // row k of the result = x_{k+1}
  }
  if (lane < NSTATE) o[N * NSTATE + lane] = 1.0;
  if (lane == 0 && status) status[b] = 0;
}

// lpv_sim.cuh - batched LPV model roll-out.
//
// Replaces planner/lib/plan_lib/distributedPlanner/LPV_val.py:77-94 (LPV_Model.sim) with its own linearisation
// LPV_val.py:96-195 (_EstimateABC of that file: the planner's formulas WITHOUT the "vx < 0.2" low-velocity branch):
//   A_k, B_k along xPred (rows 0..N-1) and u;  x_{k+1} = A_k x_k + B_k u_k from x_0 = xPred[0];
//   result = [x_1 .. x_N, ones(9)]  (np.delete(xPred, 0, 0) + a row of ones).
// get_ey is evaluated on every row of xPred by the reference (its value is unused here, a failed lookup raises).
//
// One warp per THREE models: a lane linearises one (model, stage) pair (rows of xPred are contiguous, so a warp reads
// them coalesced) into shared memory - the 9x3 column block of A_k and the input term B_k u_k, 30 doubles - then lanes
// 9g..9g+8 carry the nine states of model g through the N sequential steps (the three non-trivial columns of A_k are
// broadcast by shuffle inside the group) and write each predicted row as one 72-byte segment.  HBM traffic =
// compulsory: (N+1)*72 + N*16 bytes read, (N+1)*72 + 4 written per model.
#pragma once
#include "qp_common.cuh"

#define SIM_WARPS 2
#define SIM_GROUP 3  // models per warp

__global__ void __launch_bounds__(32 * SIM_WARPS) lpv_sim_kernel(const __grid_constant__ DevParams P, int B,
                                                                   const double* __restrict__ xpred, const double* __restrict__ u,
                                                                   double* __restrict__ out, int* __restrict__ status) {
  extern __shared__ __align__(16) double sim_sm[];  // per warp: SIM_GROUP x N x 30
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b0 = (blockIdx.x * SIM_WARPS + warp) * SIM_GROUP;
  if (b0 >= B) return;
  const int N = P.N;
  const int nm = min(SIM_GROUP, B - b0);  // models of this warp
  double* AB = sim_sm + (size_t)warp * SIM_GROUP * N * 30;
  unsigned badmask = 0;  // bit g: a map lookup of model g failed
  for (int t = lane; t < nm * (N + 1); t += 32) {
    const int g = t / (N + 1), k = t - g * (N + 1);
    const double* x = xpred + ((size_t)(b0 + g) * (N + 1) + k) * NSTATE;
    const int seg = map_segment(P, x[6]);
    if (seg < 0) { badmask |= 1u << g; continue; }
    if (k == N) continue;  // row N: only the get_ey lookup
    const double* uk = u + ((size_t)(b0 + g) * N + k) * 2;
    const double vx = x[0], vy = x[1], ey = x[3], epsi = x[4], theta = x[5];
    const double cur = P.seg_curv[seg];
    const double delta = uk[0], acc = uk[1];
    double sd, cd, se, ce, st, ct;
    sincos(delta, &sd, &cd); sincos(epsi, &se, &ce); sincos(theta, &st, &ct);
    const double imvx = 1.0 / (P.m * vx), iIvx = 1.0 / (P.I * vx);
    const double A12 = (sd * P.Cf) * imvx;
    const double A13 = (sd * P.Cf * P.lf) * imvx + vy;
    const double A22 = -(P.Cr + P.Cf * cd) * imvx;
    const double A23 = -(P.lf * P.Cf * cd - P.lr * P.Cr) * imvx - vx;
    const double A32 = -(P.lf * P.Cf * cd - P.lr * P.Cr) * iIvx;
    const double A33 = -(P.lf * P.lf * P.Cf * cd + P.lr * P.lr * P.Cr) * iIvx;
    const double B11 = -(sd * P.Cf) / P.m;
    const double iden = 1.0 / (1.0 - ey * cur);
    const double Ac[27] = {-P.mu, A12, A13,
                           0.0, A22, A23,
                           0.0, A32, A33,
                           se, ce, 0.0,
                           iden * (-ce * cur), iden * (se * cur), 1.0,
                           0.0, 0.0, 1.0,
                           ce * iden, -se * iden, 0.0,
                           ct, -st, 0.0,
                           st, ct, 0.0};
    double* A = AB + ((size_t)g * N + k) * 30;
#pragma unroll
    for (int r = 0; r < 9; ++r)
#pragma unroll
      for (int j = 0; j < 3; ++j) A[r * 3 + j] = (r == j ? 1.0 : 0.0) + P.dt * Ac[r * 3 + j];
    // B_k u_k: only the first three states have an input term
    A[27] = fma(P.dt * B11, delta, P.dt * acc);
    A[28] = P.dt * ((cd * P.Cf) / P.m) * delta;
    A[29] = P.dt * ((P.lf * P.Cf * cd) / P.I) * delta;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) badmask |= __shfl_xor_sync(0xffffffffu, badmask, o);
  __syncwarp();
  // lanes 9g .. 9g+8: state r of model g
  const int g = lane / NSTATE, r = lane - g * NSTATE;
  const bool live = g < nm;
  const bool bad = live && ((badmask >> g) & 1u);
  const int gb = live ? g : 0;
  const double* xp = xpred + (size_t)(b0 + gb) * (N + 1) * NSTATE;
  double* o = out + (size_t)(b0 + gb) * (N + 1) * NSTATE;
  const double qnan = __longlong_as_double(0x7ff8000000000000ULL);
  double v = xp[r];
  const int base = g * NSTATE;  // first lane of the group (lanes 27..31 shuffle among themselves, unused)
#pragma unroll 2
  for (int k = 0; k < N; ++k) {
    const double* A = AB + ((size_t)gb * N + k) * 30;
    const double v0 = __shfl_sync(0xffffffffu, v, base), v1 = __shfl_sync(0xffffffffu, v, base + 1),
                 v2 = __shfl_sync(0xffffffffu, v, base + 2);
    double s = (r >= 3) ? v : A[27 + r];
    s = fma(A[r * 3 + 0], v0, s); s = fma(A[r * 3 + 1], v1, s); s = fma(A[r * 3 + 2], v2, s);
    v = s;
    if (live) o[k * NSTATE + r] = bad ? qnan : v;  // row k of the result = x_{k+1}; the reference raises on a failed lookup
  }
  if (live) {
    o[N * NSTATE + r] = bad ? qnan : 1.0;
    if (r == 0 && status) status[b0 + g] = bad ? DMPC_MAP_LOOKUP_FAILED : 0;
  }
}

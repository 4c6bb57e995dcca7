// qp_common.cuh - device-side parameter block, batch pointers and small helpers for the LPV-MPC QP kernels.
//
// Replaces (reference paths relative to planner/lib/plan_lib/):
//   distributedPlanner/LPV_Planner.py:138-157 (planes, weights, linearisation, QP data) and the OSQP call :222-249.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/dmpc_b200.h"

#define QP_THREADS 128
#define QP_WARPS (QP_THREADS / 32)
#define NSTATE 9

struct DevParams {
  int N, nn, nseg, nrows;   // nrows: state rows of the prediction matrices kept on chip
  int rowsel[NSTATE];       // state index of each kept row; the first four are always vx(0), e_y(3), x(7), y(8)
  int nq;                   // nonzeros of the upper triangle of Q
  int qa[45], qb[45];       // positions (in rowsel order) of Q's nonzeros
  double qv[45];
  double dt, wq, Q00;
  double Qs[9], R[4], dR[4];
  double lf, lr, m, I, Cf, Cr, mu;
  double vx_ref, min_dist, max_vel, min_vel, max_rs, max_ls, max_ac, max_dc;
  double seg_s0[DMPC_MAX_SEG], seg_len[DMPC_MAX_SEG], seg_curv[DMPC_MAX_SEG], seg_hw[DMPC_MAX_SEG];
  // solver options
  double delta;             // polish regularisation (multiplier-refinement penalty 1/delta)
  double cert_tol;          // relative KKT certificate tolerance
  double ipm_mu0;           // first interior-point complementarity target
  int refine;               // multiplier refinement steps per polish round
  int pdas_first;           // active-set rounds tried before the interior-point phase
  int pdas_after;           // active-set rounds after each interior-point phase
  int ipm_max;              // interior-point iteration cap per phase
  int gr_global, hk_global; // 1: prediction rows / Hessian+factor live in the global workspace (long horizons)
};

struct BatchPtrs {
  int B, xprev_rows;
  const double *x0, *xprev, *uprev, *uold, *nbr_xy, *pose;
  const int* nbr_count;
  double *z, *planes, *y_ineq, *y_eq, *obj;
  int *status, *iters, *pivots;
  double* gws;              // global workspace, gws_stride doubles per QP
  long long gws_stride;
  // launch order of the fast engine: CTA i solves QP order[i] (nullptr = identity) and records what the solve cost in
  // cost[]; order_by_cost_kernel turns that into the next call's order, most expensive first, so the few QPs that
  // need the interior-point phase do not start at the end of the grid and become the tail of the launch
  const int* order;
  int* cost;
  // split launch: order[0 .. *nhard) are the QPs that needed the interior-point phase last time; part 1 (side stream,
  // one CTA per SM) solves exactly those, part 2 (main launch) skips them; part 0: one launch does everything
  const int* nhard;
  int part;
};

// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Block-wide reductions over QP_THREADS threads; `red` is a QP_WARPS-entry shared scratch. All threads get the result.
__device__ __forceinline__ double block_max(double v, double* red) {
  v = warp_max(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double r = red[0];
#pragma unroll
  for (int w = 1; w < QP_WARPS; ++w) r = fmax(r, red[w]);
  return r;
}
__device__ __forceinline__ double block_min(double v, double* red) { return -block_max(-v, red); }
__device__ __forceinline__ double block_sum(double v, double* red) {
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double r = red[0];
#pragma unroll
  for (int w = 1; w < QP_WARPS; ++w) r += red[w];
  return r;
}
__device__ __forceinline__ int block_or(int v, int* redi) {
  v = __any_sync(0xffffffffu, v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) redi[threadIdx.x >> 5] = v;
  __syncthreads();
  int r = 0;
#pragma unroll
  for (int w = 0; w < QP_WARPS; ++w) r |= redi[w];
  return r;
}

// Track lookup with the semantics of utilities/misc.py:78-101 / :105-126: wrap while s > L, clamp s < 0,
// segment i with s0[i] <= s < s0[i]+len[i]; -1 where the reference's int(np.where(...)) raises.
__device__ __forceinline__ int map_segment(const DevParams& P, double s) {
  const double L = P.seg_s0[P.nseg - 1] + P.seg_len[P.nseg - 1];
  if (!(s == s) || !(L > 0.0) || fabs(s) > 1e12) return -1;
  while (s > L) s -= L;
  if (s < 0.0) s = 0.0;
  int found = -1, cnt = 0;
  for (int i = 0; i < P.nseg; ++i)
    if (s >= P.seg_s0[i] && s < P.seg_s0[i] + P.seg_len[i]) { found = i; ++cnt; }
  return cnt == 1 ? found : -1;
}

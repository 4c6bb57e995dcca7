// qp_types.h - plain parameter / batch-pointer structs shared by every engine (no device code: also compiled by
// the host emulation of the stage-wise engine in tests/hostsim, which is test infrastructure only).
#pragma once
#include <stdint.h>

#include "../../include/dmpc_b200.h"

#define NSTATE 9

struct DevParams {
  int N, nn, nseg, nrows;   // nrows: state rows of the prediction matrices kept on chip
  int rowsel[NSTATE];       // state index of each kept row; the first four are always vx(0), e_y(3), x(7), y(8)
  int nq;                   // nonzeros of the upper triangle of Q
  int qa[45], qb[45];       // positions (in rowsel order) of Q's nonzeros
  double qv[45];
  double dt, wq, Q00;
  double Qf[81];            // Q as OSQP sees it (upper triangle mirrored), state order: the stage-wise engine's cost
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
  int seg_tiled;            // 1: the track segments tile [s0[0], end) exactly (s0[i+1] == s0[i] + len[i], len >= 0): one compare per segment finds the only match
};

struct BatchPtrs {
  int B, xprev_rows;
  int bulk_in;              // stage-wise engine: 1 = the per-QP inputs are staged in shared memory by bulk async copies (TMA)
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
  // stage-wise engine only.  only_status: solve just the QPs whose status[] equals this value (0 = all): the re-solve of
  // what the condensed engine could not factorise.  The nx_* / hold pointers (all or none) make the epilogue do the
  // caller's bookkeeping of planner/scripts/LPV_HP_N_main.py:106,113-117 itself (dmpc_shift_io).
  int only_status;
  // warm start: warm[b][row] (bytes, in/out) = the working set the previous call certified for QP b; warm_shift = 1 stores
  // it moved one stage towards the present (the next closed-loop step sees stage k+1 as its stage k)
  unsigned char* warm;
  int warm_shift;
  double *nx_x0, *nx_xprev, *nx_uprev, *nx_uold, *nx_xy;
  int* hold;
  double* flags;            // [B] 1 * ended + 2 * failed (dmpc_shift_io.flags)
  const int* stop;          // per-scenario stop (dmpc_shift_io.stop), scenario of QP b = b / stop_div
  int stop_div, laps;
  double end_s;
};


// qp_params.h - host-side conversion of the C-ABI parameter block (include/dmpc_b200.h, the constructor arguments of
// PlannerLPV, LPV_Planner.py:17-105) into the device parameter block.  Plain C++ (no device code).
#pragma once
#include "qp_types.h"

// symmetric matrix as OSQP sees it: the upper triangle of the user's array (LPV_Planner.py:156 hands csr(M), and
// OSQP reads triu(P)).
static inline double sym_up(const double* M, int n, int i, int j) { return i <= j ? M[i * n + j] : M[j * n + i]; }

static inline void fill_dev_params(const dmpc_params* p, const dmpc_options* o, DevParams& P) {
  P.N = p->N; P.nn = p->nn; P.nseg = p->nseg;
  P.dt = p->dt; P.wq = p->wq; P.Q00 = p->Q[0];
  for (int i = 0; i < 9; ++i) for (int j = 0; j < 9; ++j) P.Qf[i * 9 + j] = sym_up(p->Q, 9, i, j);
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) P.Qs[i * 3 + j] = sym_up(p->Qs, 3, i, j);
  for (int i = 0; i < 2; ++i) for (int j = 0; j < 2; ++j) { P.R[i * 2 + j] = sym_up(p->R, 2, i, j); P.dR[i * 2 + j] = sym_up(p->dR, 2, i, j); }
  P.lf = p->lf; P.lr = p->lr; P.m = p->m; P.I = p->I; P.Cf = p->Cf; P.Cr = p->Cr; P.mu = p->mu;
  P.vx_ref = p->vx_ref; P.min_dist = p->min_dist; P.max_vel = p->max_vel; P.min_vel = p->min_vel;
  P.max_rs = p->max_rs; P.max_ls = p->max_ls; P.max_ac = p->max_ac; P.max_dc = p->max_dc;
  for (int i = 0; i < DMPC_MAX_SEG; ++i) { P.seg_s0[i] = p->seg_s0[i]; P.seg_len[i] = p->seg_len[i]; P.seg_curv[i] = p->seg_curv[i]; P.seg_hw[i] = p->seg_hw[i]; }
  // state rows the condensed engine keeps on chip: the four constrained outputs first, then every other state the cost touches
  int pos[9];
  for (int a = 0; a < 9; ++a) pos[a] = -1;
  const int base[4] = {0, 3, 7, 8};
  P.nrows = 0;
  for (int i = 0; i < 4; ++i) { pos[base[i]] = P.nrows; P.rowsel[P.nrows++] = base[i]; }
  for (int a = 0; a < 9; ++a) {
    bool used = false;
    for (int b = 0; b < 9; ++b) used |= (sym_up(p->Q, 9, a, b) != 0.0);
    if (used && pos[a] < 0) { pos[a] = P.nrows; P.rowsel[P.nrows++] = a; }
  }
  P.nq = 0;
  for (int a = 0; a < 9; ++a) for (int b = a; b < 9; ++b) {
    const double v = sym_up(p->Q, 9, a, b);
    if (v != 0.0) { P.qa[P.nq] = pos[a]; P.qb[P.nq] = pos[b]; P.qv[P.nq] = v; ++P.nq; }
  }
  P.delta = o->polish_delta; P.cert_tol = o->cert_tol; P.ipm_mu0 = o->ipm_mu0;
  P.refine = o->polish_refine; P.pdas_first = o->pdas_first; P.pdas_after = o->pdas_after;
  P.ipm_max = o->ipm_max_iter;
  P.gr_global = 0; P.hk_global = 0;
  P.seg_tiled = p->nseg >= 1;
  for (int i = 0; i < p->nseg; ++i) {
    if (!(p->seg_len[i] >= 0.0)) P.seg_tiled = 0;  // zero-length segments (the reference's tracks start and end with one) never match and are skipped
    if (i + 1 < p->nseg && p->seg_s0[i + 1] != p->seg_s0[i] + p->seg_len[i]) P.seg_tiled = 0;
  }
}

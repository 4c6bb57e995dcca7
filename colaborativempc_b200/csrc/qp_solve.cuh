// qp_solve.cuh - exact QP solve on the condensed problem: primal-dual active-set polish rounds, with a Mehrotra
// predictor-corrector interior-point phase as the globalisation when the active-set rounds cycle.
//
// Replaces the OSQP call of distributedPlanner/LPV_Planner.py:222-249 (setup + solve + polish).  The returned
// point is the KKT-certified optimum (what OSQP converges to as eps -> 0), see DESIGN.md "Why not ADMM".
#pragma once
#include "qp_setup.cuh"

// res_j = (A v)_j - b_j for all rows, from inputs c.u and slacks c.sig.   Ends with a barrier.
__device__ __forceinline__ void eval_residuals(const Ctx& c, const DevParams& P) {
  stage_outputs(c, c.u, c.o, true);
  __syncthreads();
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS)
    c.res[j] = row_valid(c, j) ? row_val(c, j, c.o, c.sig, c.u) - row_h(c, P, j) : -1.0;
  __syncthreads();
}

// One polish round on the active set c.act: equality-constrained QP solved by the proximal method of multipliers
// with penalty 1/delta (OSQP's regularised KKT + iterative refinement, restated on the condensed system).
// On exit: c.u, c.sig, c.lam (multipliers, zero on inactive rows), c.res.
__device__ __forceinline__ void polish_round(const Ctx& c, const DevParams& P) {
  const double Dv = 1.0 / P.delta;
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) {
    c.D[j] = (c.act[j] && row_valid(c, j)) ? Dv : 0.0;
    c.lam[j] = 0.0;
  }
  for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) c.u[j] = 0.0;
  for (int j = threadIdx.x; j < 3 * c.N; j += QP_THREADS) c.sig[j] = 0.0;
  __syncthreads();
  stage_weights(c, P);
  __syncthreads();
  build_K(c);
  cholesky(c);
  eval_residuals(c, P);
  for (int k = 0; k <= P.refine; ++k) {
    // gradient of the augmented Lagrangian: w = lam + D*res
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) c.om[j] = c.lam[j] + c.D[j] * c.res[j];
    __syncthreads();
    rows_to_stage(c, c.om, c.wo, c.rsig);  // rsig <- A_s' w
    apply_H(c, c.u, c.tv);
    __syncthreads();
    // ru = -(H u + g + A_u' w)
    apply_Gt(c, c.wo, c.om + c.Rs, -1.0, c.tv, -1.0, c.rhs);
    for (int t = threadIdx.x; t < c.N; t += QP_THREADS) {
      const double s0 = c.sig[3 * t], s1 = c.sig[3 * t + 1], s2 = c.sig[3 * t + 2];
      c.rsig[3 * t + 0] = -(2.0 * (P.Qs[0] * s0 + P.Qs[1] * s1 + P.Qs[2] * s2) + c.rsig[3 * t + 0]);
      c.rsig[3 * t + 1] = -(2.0 * (P.Qs[3] * s0 + P.Qs[4] * s1 + P.Qs[5] * s2) + c.rsig[3 * t + 1]);
      c.rsig[3 * t + 2] = -(2.0 * (P.Qs[6] * s0 + P.Qs[7] * s1 + P.Qs[8] * s2) + c.rsig[3 * t + 2]);
    }
    __syncthreads();
    for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) c.rhs[j] -= c.g[j];
    __syncthreads();
    solve_direction(c, c.rhs, c.rsig);
    for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) c.u[j] += c.du[j];
    for (int j = threadIdx.x; j < 3 * c.N; j += QP_THREADS) c.sig[j] += c.dsig[j];
    __syncthreads();
    eval_residuals(c, P);
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) c.lam[j] += c.D[j] * c.res[j];
    __syncthreads();
  }
}

// KKT certificate + primal-dual active-set update.  Returns 1 when (u, sig, lam) is optimal for the full QP.
__device__ __forceinline__ int certify_update(const Ctx& c, const DevParams& P) {
  double lm = 0.0;
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) lm = fmax(lm, fabs(c.lam[j]));
  lm = 1.0 + block_max(lm, c.red);
  int changed = 0;
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) {
    if (!row_valid(c, j)) continue;
    const double h = row_h(c, P, j);
    if (!(fabs(c.res[j]) < 1e300) || !(fabs(c.lam[j]) < 1e300)) { changed = 1; continue; }  // NaN / inf never certifies
    if (!c.act[j]) {
      if (c.res[j] > P.cert_tol * (1.0 + fabs(h))) { c.act[j] = 1; changed = 1; }
    } else if (c.lam[j] < -P.cert_tol * lm) {
      c.act[j] = 0; changed = 1;
    }
  }
  changed = block_or(changed, c.redi);
  return !changed;
}

__device__ __forceinline__ unsigned long long act_hash(const Ctx& c) {
  unsigned long long h = 1469598103934665603ull;
  for (int j = 0; j < c.RT; ++j) { h ^= (unsigned long long)(c.act[j] + 1); h *= 1099511628211ull; }
  return h;
}

// Up to `rounds` polish rounds; stops early on a certificate (returns 1) or when an active set repeats (0).
__device__ __forceinline__ int pdas(const Ctx& c, const DevParams& P, int rounds, int* used) {
  __shared__ int s_cycle;
  int nh = 0;
  for (int r = 0; r < rounds; ++r) {
    polish_round(c, P);
    ++*used;
    if (certify_update(c, P)) return 1;
    __syncthreads();
    if (threadIdx.x == 0) {
      const unsigned long long h = act_hash(c);
      int cyc = 0;
      for (int i = 0; i < nh && i < 16; ++i) cyc |= (c.hist[i] == h);
      if (nh < 16) c.hist[nh] = h;
      s_cycle = cyc;
    }
    ++nh;
    __syncthreads();
    if (s_cycle) return 0;
  }
  return 0;
}

// dv for the current (D, rdu, rds) and complementarity right-hand side held in c.om (= rc/t - D*rp),
// then dt = -rp - A dv (c.dtt) ; caller finishes dlam.
__device__ __forceinline__ void ipm_direction(const Ctx& c, const DevParams& P) {
  rows_to_stage(c, c.om, c.wo, c.rsig);
  __syncthreads();
  apply_Gt(c, c.wo, c.om + c.Rs, 1.0, c.rdu, -1.0, c.rhs);  // -rdu + A_u' w
  for (int j = threadIdx.x; j < 3 * c.N; j += QP_THREADS) c.rsig[j] -= c.rds[j];
  __syncthreads();
  solve_direction(c, c.rhs, c.rsig);
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS)
    c.dtt[j] = row_valid(c, j) ? row_val(c, j, c.dob, c.dsig, c.du) : 0.0;  // A dv
  __syncthreads();
}

__device__ __forceinline__ double step_to_boundary(const Ctx& c, const double* x, const double* dx) {
  double a = 1.0;
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS)
    if (row_valid(c, j) && dx[j] < 0.0) a = fmin(a, -x[j] / dx[j]);
  return block_min(a, c.red);
}

// Mehrotra predictor-corrector on the condensed QP down to complementarity mu_tol.
// Returns 0 converged, 1 diverged (only a QP with infeasible hard rows does that), 2 iteration cap, 3 NaN (breakdown).
__device__ __forceinline__ int ipm(const Ctx& c, const DevParams& P, double mu_tol, int* iters) {
  const int m = c.N * (4 + c.cnt) + 4 * c.N;
  // ---- starting point: least squares with unit row weights, then shift into the cone
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) c.D[j] = row_valid(c, j) ? 1.0 : 0.0;
  for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) c.u[j] = 0.0;
  for (int j = threadIdx.x; j < 3 * c.N; j += QP_THREADS) c.sig[j] = 0.0;
  __syncthreads();
  stage_weights(c, P);
  __syncthreads();
  build_K(c);
  cholesky(c);
  eval_residuals(c, P);
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) c.om[j] = row_valid(c, j) ? c.res[j] : 0.0;
  __syncthreads();
  rows_to_stage(c, c.om, c.wo, c.rsig);
  __syncthreads();
  apply_Gt(c, c.wo, c.om + c.Rs, -1.0, c.g, -1.0, c.rhs);
  for (int j = threadIdx.x; j < 3 * c.N; j += QP_THREADS) c.rsig[j] = -c.rsig[j];
  __syncthreads();
  solve_direction(c, c.rhs, c.rsig);
  for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) c.u[j] = c.du[j];
  for (int j = threadIdx.x; j < 3 * c.N; j += QP_THREADS) c.sig[j] = c.dsig[j];
  __syncthreads();
  eval_residuals(c, P);
  {
    double mt = 1e300, ml = 1e300;
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS)
      if (row_valid(c, j)) { mt = fmin(mt, -c.res[j]); ml = fmin(ml, c.res[j]); }
    mt = block_min(mt, c.red);
    ml = block_min(ml, c.red);
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) {
      if (!row_valid(c, j)) { c.tt[j] = 1.0; c.lam[j] = 0.0; continue; }
      double t = -c.res[j], l = c.res[j];
      if (mt < 0.0) t += 1.0 - mt;
      if (ml < 0.0) l += 1.0 - ml;
      c.tt[j] = fmax(t, 1e-8); c.lam[j] = fmax(l, 1e-8);
    }
    __syncthreads();
  }
  double v_g = 0.0;
  for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) v_g = fmax(v_g, fabs(c.g[j]));
  const double gmax = 1.0 + block_max(v_g, c.red);
  for (int it = 0; it < P.ipm_max; ++it) {
    // ---- residuals
    eval_residuals(c, P);
    rows_to_stage(c, c.lam, c.wo, c.rds);
    apply_H(c, c.u, c.tv);
    __syncthreads();
    apply_Gt(c, c.wo, c.lam + c.Rs, 1.0, c.tv, 1.0, c.rdu);
    for (int t = threadIdx.x; t < c.N; t += QP_THREADS) {
      const double s0 = c.sig[3 * t], s1 = c.sig[3 * t + 1], s2 = c.sig[3 * t + 2];
      c.rds[3 * t + 0] += 2.0 * (P.Qs[0] * s0 + P.Qs[1] * s1 + P.Qs[2] * s2);
      c.rds[3 * t + 1] += 2.0 * (P.Qs[3] * s0 + P.Qs[4] * s1 + P.Qs[5] * s2);
      c.rds[3 * t + 2] += 2.0 * (P.Qs[6] * s0 + P.Qs[7] * s1 + P.Qs[8] * s2);
    }
    __syncthreads();
    double v_mu = 0.0, v_rp = 0.0, v_rd = 0.0, v_obj = 0.0, v_lm = 0.0;
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS)
      if (row_valid(c, j)) { v_mu = fma(c.lam[j], c.tt[j], v_mu); v_rp = fmax(v_rp, fabs(c.res[j] + c.tt[j])); v_lm = fmax(v_lm, c.lam[j]); }
    for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) {
      c.rdu[j] += c.g[j];
      v_rd = fmax(v_rd, fabs(c.rdu[j]));
      v_obj += c.u[j] * (0.5 * c.tv[j] + c.g[j]);
    }
    for (int j = threadIdx.x; j < 3 * c.N; j += QP_THREADS) v_rd = fmax(v_rd, fabs(c.rds[j]));
    const double mu = block_sum(v_mu, c.red) / m;
    const double rp = block_max(v_rp, c.red), rd = block_max(v_rd, c.red);
    const double obj = fabs(block_sum(v_obj, c.red));
    const double lmax = block_max(v_lm, c.red);  // the dual residual carries the rounding of A' lam: scale with it
    *iters = it;
    if (!isfinite(mu) || !isfinite(rd)) return 3;  // NaN/inf: the factorisation broke down (ill-conditioned condensed H), not a verdict
    if (!(mu < 1e40) || !(rd < 1e40)) return 1;
    if (mu <= mu_tol * (1.0 + obj) && rp <= 10.0 * mu_tol && rd <= 10.0 * mu_tol * (gmax + lmax)) return 0;
    // ---- factorise for D = lam / t
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) c.D[j] = row_valid(c, j) ? c.lam[j] / c.tt[j] : 0.0;
    __syncthreads();
    stage_weights(c, P);
    __syncthreads();
    build_K(c);
    cholesky(c);
    // ---- predictor: rc = lam*t  ->  w = rc/t - D*rp = lam - D*(res + t)
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS)
      c.om[j] = row_valid(c, j) ? c.lam[j] - c.D[j] * (c.res[j] + c.tt[j]) : 0.0;
    __syncthreads();
    ipm_direction(c, P);
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) {
      if (!row_valid(c, j)) { c.dtt[j] = 0.0; c.dlam[j] = 0.0; continue; }
      const double dt = -(c.res[j] + c.tt[j]) - c.dtt[j];
      c.dtt[j] = dt;
      c.dlam[j] = -(c.lam[j] * c.tt[j] + c.lam[j] * dt) / c.tt[j];
    }
    __syncthreads();
    double ap = step_to_boundary(c, c.tt, c.dtt), ad = step_to_boundary(c, c.lam, c.dlam);
    double v_aff = 0.0;
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS)
      if (row_valid(c, j)) v_aff = fma(c.lam[j] + ad * c.dlam[j], c.tt[j] + ap * c.dtt[j], v_aff);
    const double mu_aff = block_sum(v_aff, c.red) / m;
    const double sr = mu_aff / mu, smu = sr * sr * sr * mu;
    // ---- corrector: rc = lam*t + dt_aff*dlam_aff - sigma*mu
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) {
      if (!row_valid(c, j)) { c.om[j] = 0.0; continue; }
      const double rc = c.lam[j] * c.tt[j] + c.dtt[j] * c.dlam[j] - smu;
      c.dlam[j] = rc;  // keep rc
      c.om[j] = rc / c.tt[j] - c.D[j] * (c.res[j] + c.tt[j]);
    }
    __syncthreads();
    ipm_direction(c, P);
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) {
      if (!row_valid(c, j)) { c.dtt[j] = 0.0; c.dlam[j] = 0.0; continue; }
      const double dt = -(c.res[j] + c.tt[j]) - c.dtt[j];
      const double rc = c.dlam[j];
      c.dtt[j] = dt;
      c.dlam[j] = -(rc + c.lam[j] * dt) / c.tt[j];
    }
    __syncthreads();
    ap = fmin(1.0, 0.99 * step_to_boundary(c, c.tt, c.dtt));
    ad = fmin(1.0, 0.99 * step_to_boundary(c, c.lam, c.dlam));
    for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) c.u[j] = fma(ap, c.du[j], c.u[j]);
    for (int j = threadIdx.x; j < 3 * c.N; j += QP_THREADS) c.sig[j] = fma(ap, c.dsig[j], c.sig[j]);
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) {
      c.tt[j] = fma(ap, c.dtt[j], c.tt[j]);
      c.lam[j] = fma(ad, c.dlam[j], c.lam[j]);
    }
    __syncthreads();
  }
  return 2;
}

// Full solve.  Returns dmpc_status; *n_ipm / *n_rounds report the work done.
__device__ __forceinline__ int solve_qp(const Ctx& c, const DevParams& P, int* n_ipm, int* n_rounds) {
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) c.act[j] = 0;
  __syncthreads();
  *n_ipm = 0; *n_rounds = 0;
  if (pdas(c, P, P.pdas_first, n_rounds)) return DMPC_SOLVED;
  double mu_tol = P.ipm_mu0, mu_ok = 0.0;
  while (mu_tol > 1e-13) {
    int its = 0;
    const int rc = ipm(c, P, mu_tol, &its);
    *n_ipm += its;
    if (rc == 1 || rc == 3) {  // divergence of the FIRST phase = infeasible hard rows; later (tighter) phases can only lose accuracy
      if (mu_ok == 0.0) return rc == 3 ? DMPC_UNSOLVED : DMPC_PRIMAL_INFEASIBLE;  // NaN is a numerical failure, not a certificate
      break;
    }
    if (rc == 0) mu_ok = mu_tol;
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) c.act[j] = row_valid(c, j) && (c.lam[j] > c.tt[j]);
    __syncthreads();
    if (pdas(c, P, P.pdas_after, n_rounds)) return DMPC_SOLVED;
    mu_tol *= 1e-2;
  }
  if (mu_ok > 0.0) {  // no active set certified (degenerate rows): hand back the tightest interior-point iterate
    int its = 0;
    ipm(c, P, mu_ok, &its);
    *n_ipm += its;
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) if (!row_valid(c, j)) c.lam[j] = 0.0;
    __syncthreads();
    return DMPC_SOLVED_INACCURATE;
  }
  return DMPC_MAX_ITER_REACHED;
}

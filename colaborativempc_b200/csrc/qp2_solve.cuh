// qp2_solve.cuh - exact QP solve on the fast engine (qp2_engine.cuh): primal-dual active-set polish rounds with a
// Mehrotra predictor-corrector interior-point phase as globalisation.  Same algorithm as qp_solve.cuh.
//
// Replaces the OSQP call of distributedPlanner/LPV_Planner.py:222-249 (setup + solve + polish).  The returned point is
// the KKT-certified optimum (what OSQP converges to as eps -> 0), see DESIGN.md.
#pragma once
#include "qp2_engine.cuh"

// res_j = (A v)_j - b_j for all rows, from inputs c.u and slacks c.sig.   Ends with a barrier.
__device__ __noinline__ void eval_residuals2() {
  CTX2_LOCAL(c);
  stage_outputs2(c, c.u, c.dob, true);
  __syncthreads();
#pragma unroll 1
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS)
    c.res[j] = row_valid2(c, j) ? row_val2(c, j, c.dob, c.sig, c.u) - c.bv[j] : -1.0;
  __syncthreads();
}

// One polish round on the active set c.act: equality-constrained QP solved by the proximal method of multipliers
// with penalty 1/delta (OSQP's regularised KKT + iterative refinement, restated on the condensed system).
// On exit: c.u, c.sig, c.lam (multipliers, zero on inactive rows), c.res.
__device__ __forceinline__ void polish_round2(const Ctx2& c) {
  const double Dv = 1.0 / c.hd->delta;
  const double* Qs = c.hd->Qs;
  const int refine = c.hd->refine;
#pragma unroll 1
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) {
    c.D[j] = (c.act[j] && row_valid2(c, j)) ? Dv : 0.0;
    c.lam[j] = 0.0;
  }
#pragma unroll 1
  for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) c.u[j] = 0.0;
#pragma unroll 1
  for (int j = threadIdx.x; j < 3 * c.N; j += QP_THREADS) c.sig[j] = 0.0;
  __syncthreads();
  factorise2();
  eval_residuals2();
#pragma unroll 1
  for (int k = 0; k <= refine; ++k) {
    // gradient of the augmented Lagrangian: w = lam + D*res
#pragma unroll 1
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) c.om[j] = c.lam[j] + c.D[j] * c.res[j];
    __syncthreads();
    rows_to_stage2(c, c.om, c.wo, c.rsig);  // rsig <- A_s' w
    apply_H2(c, c.u, c.tv);
    __syncthreads();
    // ru = -(H u + g + A_u' w)
    apply_Gt2(c, c.wo, c.om + c.Rs, -1.0, c.tv, -1.0, c.rhs, true);
#pragma unroll 1
    for (int t = threadIdx.x; t < c.N; t += QP_THREADS) {
      const double s0 = c.sig[3 * t], s1 = c.sig[3 * t + 1], s2 = c.sig[3 * t + 2];
      c.rsig[3 * t + 0] = -(2.0 * (Qs[0] * s0 + Qs[1] * s1 + Qs[2] * s2) + c.rsig[3 * t + 0]);
      c.rsig[3 * t + 1] = -(2.0 * (Qs[3] * s0 + Qs[4] * s1 + Qs[5] * s2) + c.rsig[3 * t + 1]);
      c.rsig[3 * t + 2] = -(2.0 * (Qs[6] * s0 + Qs[7] * s1 + Qs[8] * s2) + c.rsig[3 * t + 2]);
    }
    __syncthreads();
#pragma unroll 1
    for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) c.rhs[j] -= c.g[j];
    __syncthreads();
    solve_direction2(1);
#pragma unroll 1
    for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) c.u[j] += c.du[j];
#pragma unroll 1
    for (int j = threadIdx.x; j < 3 * c.N; j += QP_THREADS) c.sig[j] += c.dsig[j];
    __syncthreads();
    eval_residuals2();
    double v = 0.0;
#pragma unroll 1
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) {
      c.lam[j] += c.D[j] * c.res[j];
      if (c.D[j] > 0.0) v = fmax(v, fabs(c.res[j]) / (1.0 + fabs(c.bv[j])));
    }
    // stationarity holds for (u, lam) after every pass; the passes only drive the active rows onto their bounds
    if (block_max(v, c.red) <= 0.1 * c.hd->cert_tol) break;
  }
  __syncthreads();
}

// KKT certificate + primal-dual active-set update.  Returns 1 when (u, sig, lam) is optimal for the full QP.
__device__ __forceinline__ int certify_update2(const Ctx2& c) {
  const double cert_tol = c.hd->cert_tol;
  double lm = 0.0;
#pragma unroll 1
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) lm = fmax(lm, fabs(c.lam[j]));
  lm = 1.0 + block_max(lm, c.red);
  int changed = 0;
#pragma unroll 1
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) {
    if (!row_valid2(c, j)) continue;
    const double h = c.bv[j];
    if (!(fabs(c.res[j]) < 1e300) || !(fabs(c.lam[j]) < 1e300)) { changed = 1; continue; }  // NaN / inf never certifies
    if (!c.act[j]) {
      if (c.res[j] > cert_tol * (1.0 + fabs(h))) { c.act[j] = 1; changed = 1; }
    } else if (c.lam[j] < -cert_tol * lm) {
      c.act[j] = 0; changed = 1;
    }
  }
  changed = block_or(changed, c.redi);
  return !changed;
}

// order-independent 64-bit hash of the active set (sum of per-row mixes), all threads get the result
__device__ __forceinline__ unsigned long long act_hash2(const Ctx2& c) {
  unsigned long long h = 0;
#pragma unroll 1
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS)
    if (c.act[j]) {
      unsigned long long x = (unsigned long long)(j + 1) * 0x9E3779B97F4A7C15ull;
      x ^= x >> 29; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 32;
      h += x;
    }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) h += __shfl_xor_sync(0xffffffffu, h, o);
  unsigned long long* red = (unsigned long long*)c.red;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = h;
  __syncthreads();
  unsigned long long r = 0;
#pragma unroll
  for (int w = 0; w < QP_WARPS; ++w) r += red[w];
  return r;
}

// Up to `rounds` polish rounds; stops early on a certificate (returns 1) or when an active set repeats (0).
__device__ __noinline__ int pdas2(int rounds) {
  CTX2_LOCAL(c);
  int nh = 0;
#pragma unroll 1
  for (int r = 0; r < rounds; ++r) {
    polish_round2(c);
    if (threadIdx.x == 0) ++((Hdr2*)smem_fast)->n_rounds;
    if (certify_update2(c)) return 1;
    const unsigned long long h = act_hash2(c);
    int cyc = 0;
    for (int i = 0; i < nh && i < 16; ++i) cyc |= (c.hist[i] == h);
    __syncthreads();
    if (threadIdx.x == 0 && nh < 16) c.hist[nh] = h;
    ++nh;
    __syncthreads();
    if (cyc) return 0;
  }
  return 0;
}

// dv for the current (D, rdu, rds) and complementarity right-hand side held in c.om (= rc/t - D*rp),
// then A dv (c.dtt); caller finishes dt, dlam.
__device__ __noinline__ void ipm_direction2() {
  CTX2_LOCAL(c);
  rows_to_stage2(c, c.om, c.wo, c.rsig);
  __syncthreads();
  apply_Gt2(c, c.wo, c.om + c.Rs, 1.0, c.rdu, -1.0, c.rhs, false);  // -rdu + A_u' w
#pragma unroll 1
  for (int j = threadIdx.x; j < 3 * c.N; j += QP_THREADS) c.rsig[j] -= c.rds[j];
  __syncthreads();
  solve_direction2(0);
#pragma unroll 1
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS)
    c.dtt[j] = row_valid2(c, j) ? row_val2(c, j, c.dob, c.dsig, c.du) : 0.0;  // A dv
  __syncthreads();
}

__device__ __forceinline__ double step_to_boundary2(const Ctx2& c, const double* x, const double* dx) {
  double a = 1.0;
#pragma unroll 1
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS)
    if (row_valid2(c, j) && dx[j] < 0.0) a = fmin(a, -x[j] / dx[j]);
  return block_min(a, c.red);
}

// Interior-point iterate (u, sigma, t, lambda) <-> c.sv.  Ends with a barrier.
__device__ __forceinline__ void ipm_keep2(const Ctx2& c, int save) {
  double* p = c.sv;
  double* const v[4] = {c.u, c.sig, c.tt, c.lam};
  const int len[4] = {c.nu, 3 * c.N, c.RT, c.RT};
#pragma unroll
  for (int q = 0; q < 4; ++q) {
#pragma unroll 1
    for (int j = threadIdx.x; j < len[q]; j += QP_THREADS) {
      if (save) p[j] = v[q][j];
      else v[q][j] = p[j];
    }
    p += len[q];
  }
  __syncthreads();
}

// Mehrotra predictor-corrector on the condensed QP down to complementarity mu_tol.  warm: continue from the iterate
// kept in c.sv (the previous, looser phase) instead of building a starting point.
// Returns 0 converged, 1 diverged (only a QP with infeasible hard rows does that), 2 iteration cap, 3 NaN (breakdown).
__device__ __noinline__ int ipm2(double mu_tol, int warm) {
  CTX2_LOCAL(c);
  const double* Qs = c.hd->Qs;
  const int ipm_max = c.hd->ipm_max;
  const int m = c.N * (4 + c.cnt) + 4 * c.N;
  if (warm) ipm_keep2(c, 0);
  else {
  // ---- starting point: least squares with unit row weights, then shift into the cone
#pragma unroll 1
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) c.D[j] = row_valid2(c, j) ? 1.0 : 0.0;
#pragma unroll 1
  for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) c.u[j] = 0.0;
#pragma unroll 1
  for (int j = threadIdx.x; j < 3 * c.N; j += QP_THREADS) c.sig[j] = 0.0;
  __syncthreads();
  factorise2();
  eval_residuals2();
#pragma unroll 1
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) c.om[j] = row_valid2(c, j) ? c.res[j] : 0.0;
  __syncthreads();
  rows_to_stage2(c, c.om, c.wo, c.rsig);
  __syncthreads();
  apply_Gt2(c, c.wo, c.om + c.Rs, -1.0, c.g, -1.0, c.rhs, false);
#pragma unroll 1
  for (int j = threadIdx.x; j < 3 * c.N; j += QP_THREADS) c.rsig[j] = -c.rsig[j];
  __syncthreads();
  solve_direction2(0);
#pragma unroll 1
  for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) c.u[j] = c.du[j];
#pragma unroll 1
  for (int j = threadIdx.x; j < 3 * c.N; j += QP_THREADS) c.sig[j] = c.dsig[j];
  __syncthreads();
  eval_residuals2();
  {
    // Multipliers of violated soft rows end near 2 Qs sigma: with the reference's slack weights (1e7) a start at O(1)
    // spends ten iterations with tiny dual steps growing them.  Start the duals at 1e-3 max(Qs) instead (measured on the
    // working-set-overflow QPs of the bench batch: 18 -> 9.5 iterations on average, 24 -> 12 at worst).
    const double lam_scale = fmax(1.0, 1e-3 * fmax(Qs[0], fmax(Qs[4], Qs[8])));
    double mt = 1e300, ml = 1e300;
#pragma unroll 1
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS)
      if (row_valid2(c, j)) { mt = fmin(mt, -c.res[j]); ml = fmin(ml, c.res[j]); }
    mt = block_min(mt, c.red);
    ml = block_min(ml, c.red);
#pragma unroll 1
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) {
      if (!row_valid2(c, j)) { c.tt[j] = 1.0; c.lam[j] = 0.0; continue; }
      double t = -c.res[j], l = c.res[j];
      if (mt < 0.0) t += 1.0 - mt;
      if (ml < 0.0) l += 1.0 - ml;
      c.tt[j] = fmax(t, 1e-8); c.lam[j] = lam_scale * fmax(l, 1e-8);
    }
    __syncthreads();
  }
  }
  double v_g = 0.0;
#pragma unroll 1
  for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) v_g = fmax(v_g, fabs(c.g[j]));
  const double gmax = 1.0 + block_max(v_g, c.red);
#pragma unroll 1
  for (int it = 0; it < ipm_max; ++it) {
    // ---- residuals
    eval_residuals2();
    rows_to_stage2(c, c.lam, c.wo, c.rds);
    apply_H2(c, c.u, c.tv);
    __syncthreads();
    apply_Gt2(c, c.wo, c.lam + c.Rs, 1.0, c.tv, 1.0, c.rdu, false);
#pragma unroll 1
    for (int t = threadIdx.x; t < c.N; t += QP_THREADS) {
      const double s0 = c.sig[3 * t], s1 = c.sig[3 * t + 1], s2 = c.sig[3 * t + 2];
      c.rds[3 * t + 0] += 2.0 * (Qs[0] * s0 + Qs[1] * s1 + Qs[2] * s2);
      c.rds[3 * t + 1] += 2.0 * (Qs[3] * s0 + Qs[4] * s1 + Qs[5] * s2);
      c.rds[3 * t + 2] += 2.0 * (Qs[6] * s0 + Qs[7] * s1 + Qs[8] * s2);
    }
    __syncthreads();
    double v_mu = 0.0, v_rp = 0.0, v_rd = 0.0, v_obj = 0.0, v_lm = 0.0;
#pragma unroll 1
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS)
      if (row_valid2(c, j)) { v_mu = fma(c.lam[j], c.tt[j], v_mu); v_rp = fmax(v_rp, fabs(c.res[j] + c.tt[j])); v_lm = fmax(v_lm, c.lam[j]); }
#pragma unroll 1
    for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) {
      c.rdu[j] += c.g[j];
      v_rd = fmax(v_rd, fabs(c.rdu[j]));
      v_obj += c.u[j] * (0.5 * c.tv[j] + c.g[j]);
    }
#pragma unroll 1
    for (int j = threadIdx.x; j < 3 * c.N; j += QP_THREADS) v_rd = fmax(v_rd, fabs(c.rds[j]));
    const double mu = block_sum(v_mu, c.red) / m;
    const double rp = block_max(v_rp, c.red), rd = block_max(v_rd, c.red);
    const double obj = fabs(block_sum(v_obj, c.red));
    const double lmax = block_max(v_lm, c.red);  // the dual residual carries the rounding of A' lam: scale with it
    if (threadIdx.x == 0) ++((Hdr2*)smem_fast)->n_ipm;
    if (!isfinite(mu) || !isfinite(rd)) return 3;  // NaN/inf: the factorisation broke down (ill-conditioned condensed H), not a verdict
    if (!(mu < 1e40) || !(rd < 1e40)) return 1;
    if (mu <= mu_tol * (1.0 + obj) && rp <= 10.0 * mu_tol && rd <= 10.0 * mu_tol * (gmax + lmax)) return 0;
    // ---- factorise for D = lam / t
#pragma unroll 1
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) c.D[j] = row_valid2(c, j) ? c.lam[j] / c.tt[j] : 0.0;
    __syncthreads();
    factorise2();
    // ---- predictor: rc = lam*t  ->  w = rc/t - D*rp = lam - D*(res + t)
#pragma unroll 1
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS)
      c.om[j] = row_valid2(c, j) ? c.lam[j] - c.D[j] * (c.res[j] + c.tt[j]) : 0.0;
    __syncthreads();
    ipm_direction2();
#pragma unroll 1
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) {
      if (!row_valid2(c, j)) { c.dtt[j] = 0.0; c.dlam[j] = 0.0; continue; }
      const double dt = -(c.res[j] + c.tt[j]) - c.dtt[j];
      c.dtt[j] = dt;
      c.dlam[j] = -(c.lam[j] * c.tt[j] + c.lam[j] * dt) / c.tt[j];
    }
    __syncthreads();
    double ap = step_to_boundary2(c, c.tt, c.dtt), ad = step_to_boundary2(c, c.lam, c.dlam);
    double v_aff = 0.0;
#pragma unroll 1
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS)
      if (row_valid2(c, j)) v_aff = fma(c.lam[j] + ad * c.dlam[j], c.tt[j] + ap * c.dtt[j], v_aff);
    const double mu_aff = block_sum(v_aff, c.red) / m;
    const double sr = mu_aff / mu, smu = sr * sr * sr * mu;
    // ---- corrector: rc = lam*t + dt_aff*dlam_aff - sigma*mu
#pragma unroll 1
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) {
      if (!row_valid2(c, j)) { c.om[j] = 0.0; continue; }
      const double rc = c.lam[j] * c.tt[j] + c.dtt[j] * c.dlam[j] - smu;
      c.dlam[j] = rc;  // keep rc
      c.om[j] = rc / c.tt[j] - c.D[j] * (c.res[j] + c.tt[j]);
    }
    __syncthreads();
    ipm_direction2();
#pragma unroll 1
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) {
      if (!row_valid2(c, j)) { c.dtt[j] = 0.0; c.dlam[j] = 0.0; continue; }
      const double dt = -(c.res[j] + c.tt[j]) - c.dtt[j];
      const double rc = c.dlam[j];
      c.dtt[j] = dt;
      c.dlam[j] = -(rc + c.lam[j] * dt) / c.tt[j];
    }
    __syncthreads();
    ap = fmin(1.0, 0.99 * step_to_boundary2(c, c.tt, c.dtt));
    ad = fmin(1.0, 0.99 * step_to_boundary2(c, c.lam, c.dlam));
#pragma unroll 1
    for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) c.u[j] = fma(ap, c.du[j], c.u[j]);
#pragma unroll 1
    for (int j = threadIdx.x; j < 3 * c.N; j += QP_THREADS) c.sig[j] = fma(ap, c.dsig[j], c.sig[j]);
#pragma unroll 1
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) {
      c.tt[j] = fma(ap, c.dtt[j], c.tt[j]);
      c.lam[j] = fma(ad, c.dlam[j], c.lam[j]);
    }
    __syncthreads();
  }
  return 2;
}


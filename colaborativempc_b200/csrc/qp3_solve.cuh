// qp3_solve.cuh - exact QP solve on the stage-wise engine (qp3_engine.cuh).
//
// Replaces the OSQP call of distributedPlanner/LPV_Planner.py:222-249 (setup + solve + polish).  The returned point is
// the KKT-certified optimum of the reference's QP (what OSQP converges to as eps -> 0):
//   1. primal-dual active-set rounds: the rows of the guess get the penalty 1/delta (OSQP's polish, LPV_Planner.py:233
//      polish=True, restated stage-wise), the multipliers are refined by the method of multipliers, the guess is updated
//      from the signs (violated rows in, negative multipliers out) until the KKT certificate holds;
//   2. a Mehrotra predictor-corrector interior-point phase when the rounds cycle or run out, whose complementarity
//      pattern seeds the next rounds; the phase continues from its own iterate at a tighter target if they fail.
// Every linear solve is a Riccati sweep; nothing here depends on the horizon except loop lengths.
#pragma once
#include "qp3_engine.cuh"

// loose thresholds that decide "the guess is clearly wrong, do not spend refinement passes on it"
#define Q3_LOOSE 1e-6

// One pass of the method of multipliers, ONE LANE PER STAGE and straight-line per row (no row decoding): the residuals of
// the stage's rows at the current iterate (as eval_residuals3), lam += D res, and the pass statistics: v = how far the
// active rows are from their bounds (relative), lm = max |lam|, dl = max |D res| (the multiplier update), bad = non-finite.
// All lanes return the reduced values.  Ends with a barrier.
Q3D int pass_update3(const Ctx3& c, double& v_out, double& lm_out, double& dl_out, double& vin_out, double& lmin_out) {
  const double* lim = c.hd->lim;
  const double hm = 0.5 * c.hd->mind;
  double v = 0.0, lm = 0.0, dl = 0.0, vin = 0.0, lmin = 0.0;
  int bad = 0;
  auto row = [&](int j, double res, double babs) {
    const double D = c.D[j];
    const double dr = D * res, l = c.lam[j] + dr;
    c.res[j] = res; c.lam[j] = l;
    dl = fmax(dl, fabs(dr)); lm = fmax(lm, fabs(l));
    if (!(fabs(res) < 1e300) || !(fabs(l) < 1e300)) bad = 1;
    // what judging the guess needs, gathered in the same pass: the worst violation of a row outside the guess (relative
    // to 1 + |b|; only violated rows pay the reciprocal) and the most negative multiplier inside it
    if (D > 0.0) { v = fmax(v, fabs(res) * q3_rcp(1.0 + babs)); lmin = fmin(lmin, l); }
    else if (res > 0.0) vin = fmax(vin, res * q3_rcp(1.0 + babs));
  };
  Q3_FOR(k, c.N) {
    const double* xk = c.x + 9 * k;
    const double* sk = c.sig + 3 * k;
    const int j0 = k * c.rst, jb = c.Rs + 4 * k;
    const double eb = c.eyb[k];
    row(j0, lim[0] - xk[0], fabs(lim[0]));
    row(j0 + 1, (xk[0] - lim[1]) + sk[0], fabs(lim[1]));
    row(j0 + 2, (xk[3] - eb) + sk[1], fabs(eb));
    row(j0 + 3, (-xk[3] - eb) + sk[1], fabs(eb));
#pragma unroll 1
    for (int i = 0; i < c.nn; ++i) {
      const int q = k * c.nn + i;
      if (i < c.cnt) {
        const double ax = c.ax[q], ay = c.ay[q], mx = c.mx[q], my = c.my[q];
        row(j0 + 4 + i, fma(ax, xk[7] - mx, ay * (xk[8] - my)) + (hm - sk[2]), fabs(fma(ax, mx, ay * my) - hm));
      } else {
        c.res[j0 + 4 + i] = -1.0;
      }
    }
    const double u0 = c.u[2 * k], u1 = c.u[2 * k + 1];
    row(jb, u0 - lim[2], fabs(lim[2])); row(jb + 1, -u0 - lim[3], fabs(lim[3]));
    row(jb + 2, u1 - lim[4], fabs(lim[4])); row(jb + 3, -u1 - lim[5], fabs(lim[5]));
  }
  bad = q3_wor(bad);
  v_out = q3_wmax(v); lm_out = q3_wmax(lm); dl_out = q3_wmax(dl); vin_out = q3_wmax(vin); lmin_out = q3_wmin(lmin);
  Q3_SYNC();
  return bad;
}

// Judging / updating the guess, one lane per stage: a row outside the guess that is violated beyond tol (relative to
// 1 + |b|) or a row inside it whose multiplier is below -tol lm.  update = 0: returns whether any row would move;
// update = 1: moves them (c.act) and ends with a barrier.
Q3D int judge_rows3(const Ctx3& c, double tol, double lm, int update) {
  const double* lim = c.hd->lim;
  const double hm = 0.5 * c.hd->mind;
  int changed = 0;
  auto row = [&](int j, double babs) {
    const int a = c.act[j];
    const bool mv = a ? (c.lam[j] < -tol * lm) : (c.res[j] > tol * (1.0 + babs));
    if (mv) { changed = 1; if (update) c.act[j] = (unsigned char)(a ? 0 : 1); }
  };
  Q3_FOR(k, c.N) {
    const int j0 = k * c.rst, jb = c.Rs + 4 * k;
    const double eb = fabs(c.eyb[k]);
    row(j0, fabs(lim[0])); row(j0 + 1, fabs(lim[1])); row(j0 + 2, eb); row(j0 + 3, eb);
#pragma unroll 1
    for (int i = 0; i < c.cnt; ++i) {
      const int q = k * c.nn + i;
      row(j0 + 4 + i, fabs(fma(c.ax[q], c.mx[q], c.ay[q] * c.my[q]) - hm));
    }
    row(jb, fabs(lim[2])); row(jb + 1, fabs(lim[3])); row(jb + 2, fabs(lim[4])); row(jb + 3, fabs(lim[5]));
  }
  changed = q3_wor(changed);
  if (update) Q3_SYNC();
  return changed;
}

// One active-set round on c.act.  first: the iterate is the bounded starting point, not yet dynamics-consistent.
// Returns 1 certified, 0 guess updated (c.act), 2 numerical failure (non-finite).
// On exit c.x, c.u, c.sig, c.lam, c.res belong to the guess that was solved.
Q3N int polish_round3(int first) {
  Q3_CTX(c);
  Hdr3* hd = c.hd;
  // Penalty of the rows in the guess: 10 / delta.  The multiplier iteration contracts by (curvature seen by the row) / D
  // per pass, and that curvature is up to 2 Qs = 2e7 for a soft row (after the slack elimination its weight is
  // D 2Qs / (2Qs + D) -> 2Qs) AND for a hard row whose state is coupled to active soft rows through the dynamics.
#ifndef Q3_SOFT
#define Q3_SOFT 10.0
#endif
  const double Dv = Q3_SOFT / hd->delta, Ds = Dv, cert_tol = hd->cert_tol;
  Q3_FOR(j, c.RT) {
    const bool soft = j < c.Rs && (c.rinfo[j] & 127) != 0;
    c.D[j] = (c.act[j] && row_ok3(c, j)) ? (soft ? Ds : Dv) : 0.0;
    c.lam[j] = 0.0;
  }
  Q3_SYNC();
  eval_residuals3(c);
  factor_solve3(G3_ALM, first);
  primal_step3(c, 1.0);
  int pass = 0;
  double v_prev = 1e300;
#pragma unroll 1
  for (;;) {
    // residuals, multiplier update, then: how far are the active rows from their bounds, is the guess still plausible
    double v, lm, dl, vin, lmin;
    if (pass_update3(c, v, lm, dl, vin, lmin)) return 2;
    lm = 1.0 + lm;
#ifdef QP3_TRACE
    { double dl = 0; for (int j = 0; j < c.RT; ++j) dl = fmax(dl, fabs(c.D[j] * c.res[j])); int na = 0; for (int j = 0; j < c.RT; ++j) na += c.act[j];
      printf("  pass %d v %.3e dlam %.3e lm %.3e nact %d\n", pass, v, dl, lm, na); }
#endif
    // converged: active rows on their bounds AND the multiplier iteration has stopped moving (the last update is what
    // the multipliers are still off by, times the contraction of a pass)
    // pass budget: polish_refine passes, more (up to 3x) while each pass still brings the active rows 8x closer
    // a pass that does not even halve a violation of 1e-3 and more: the rows of the guess cannot all be met (an
    // inconsistent guess keeps v where it is while the multipliers grow without bound) - do not refine it
    const bool stuck = pass >= 1 && v > 1e-3 && !(v < 0.5 * v_prev);
    const bool last = stuck || pass >= 3 * hd->refine || (pass >= hd->refine && !(v < 0.125 * v_prev));
    v_prev = v;
    const bool tight = pass >= 1 && v <= 0.1 * cert_tol && (dl <= 1e-7 * lm || last);
    // before that, only a CLEARLY wrong guess ends the round early: a pass leaves the active rows lam / D off their
    // bounds (v), and every other row is off by about as much, so "violated" has to mean well beyond v
    const double tol = tight ? cert_tol : fmax(Q3_LOOSE, 8.0 * v);
    // ... and while the active rows are still far from their bounds (multipliers of 1e7 and more leave them 1e-2 off
    // after the first pass) neither the other rows' residuals nor the multipliers' signs mean anything yet: refine first
    const bool judge = tight || v <= 1e-4 || last;
    const int changed = judge && (vin > tol || lmin < -tol * lm);  // = judge_rows3(c, tol, lm, 0), from the pass's own statistics
    if (tight && !changed) return 1;
    if (changed || last) {
      // update the guess with the tight thresholds (a row the loose test let through still moves)
      judge_rows3(c, cert_tol, lm, 1);
      return 0;
    }
    affine_solve3(G3_ALM, 0.0);
    primal_step3(c, 1.0);
    ++pass;
  }
}

// order-independent 64-bit hash of the active set
Q3D unsigned long long act_hash3(const Ctx3& c) {
  unsigned long long h = 0;
  Q3_FOR(j, c.RT)
    if (c.act[j]) {
      unsigned long long x = (unsigned long long)(j + 1) * 0x9E3779B97F4A7C15ull;
      x ^= x >> 29; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 32;
      h += x;
    }
  return q3_wsum64(h);
}

// Up to `rounds` rounds; 1 certified, 0 not (cycle or round cap), 2 numerical failure.
// bail > 0: give up after any round from `bail_from` on that leaves more than `bail` rows in the guess (a QP whose
// guesses hold that many rows ends in the interior point anyway: do not spend the rounds; a warm first guess may
// legitimately be that large, so for it the test starts after the second round).
Q3N int pdas3(int rounds, int first, int bail, int bail_from) {
  Q3_CTX(c);
  int nh = 0;
#pragma unroll 1
  for (int r = 0; r < rounds; ++r) {
    const int st = polish_round3(first && r == 0);
    if (Q3_LANE == 0) ++c.hd->n_rounds;
#ifdef QP3_TRACE2
    { int na = 0; for (int j = 0; j < c.RT; ++j) na += c.act[j]; printf("R %d %d %d\n", r, st, na); }
#endif
    if (st) return st;
    if (bail > 0 && r >= bail_from) {
      double na = 0.0;
      Q3_FOR(j, c.RT) na += c.act[j] ? 1.0 : 0.0;
      if (q3_wsum(na) > (double)bail) return 0;
    }
    const unsigned long long h = act_hash3(c);
    int cyc = 0;
    for (int i = 0; i < nh && i < 8; ++i) cyc |= (c.hd->hist[i] == h);
    Q3_SYNC();
    if (Q3_LANE == 0) c.hd->hist[nh & 7] = h;
    ++nh;
    Q3_SYNC();
    if (cyc) return 0;
  }
  return 0;
}

// interior-point iterate (x, u, sigma, lambda, t) <-> the QP's global scratch slice.  Ends with a barrier.
Q3D void ipm_park3(const Ctx3& c, double* park, int save) {
  double* const v[5] = {c.x, c.u, c.sig, c.lam, c.tt};
  const int len[5] = {9 * c.N, 2 * c.N, 3 * c.N, c.RT, c.RT};
  double* p = park;
  for (int q = 0; q < 5; ++q) {
    Q3_FOR(j, len[q]) {
      if (save) p[j] = v[q][j];
      else v[q][j] = p[j];
    }
    p += len[q];
  }
  Q3_SYNC();
}

// Top of an interior-point iteration, ONE LANE PER STAGE (straight-line per row, no row decoding): the residuals at the
// iterate (as eval_residuals3), the per-lane partial sums of the complementarity lam't, the primal residual |res + t| and the
// largest multiplier, and the weights D = lam / t of the next factorisation.  Unused neighbour rows are inert by
// construction (lam = 0, t = 1, res = -1, steps 0), so the flat loops of the iteration need no validity test.  Ends with
// a barrier.
Q3D void ipm_rows_eval3(const Ctx3& c, double& v_mu, double& v_rp, double& v_lm) {
  const double* lim = c.hd->lim;
  const double hm = 0.5 * c.hd->mind;
  auto row = [&](int j, double res) {
    const double t = c.tt[j], l = c.lam[j];
    c.res[j] = res;
    v_mu = fma(l, t, v_mu); v_rp = fmax(v_rp, fabs(res + t)); v_lm = fmax(v_lm, l);
    c.D[j] = l * q3_rcp(t);
  };
  Q3_FOR(k, c.N) {
    const double* xk = c.x + 9 * k;
    const double* sk = c.sig + 3 * k;
    const int j0 = k * c.rst, jb = c.Rs + 4 * k;
    const double eb = c.eyb[k];
    row(j0, lim[0] - xk[0]);
    row(j0 + 1, (xk[0] - lim[1]) + sk[0]);
    row(j0 + 2, (xk[3] - eb) + sk[1]);
    row(j0 + 3, (-xk[3] - eb) + sk[1]);
#pragma unroll 1
    for (int i = 0; i < c.nn; ++i) {
      const int q = k * c.nn + i;
      if (i < c.cnt) row(j0 + 4 + i, fma(c.ax[q], xk[7] - c.mx[q], c.ay[q] * (xk[8] - c.my[q])) + (hm - sk[2]));
      else { c.res[j0 + 4 + i] = -1.0; c.D[j0 + 4 + i] = 0.0; }
    }
    const double u0 = c.u[2 * k], u1 = c.u[2 * k + 1];
    row(jb, u0 - lim[2]); row(jb + 1, -u0 - lim[3]); row(jb + 2, u1 - lim[4]); row(jb + 3, -u1 - lim[5]);
  }
  Q3_SYNC();
}

// Slack and multiplier steps from the primal direction, one lane per stage:  dt = -(res + t) - (A d)_j;  predictor
// dlam = -(lam + D dt);  corrector dlam = -(lam t + dt_aff dlam_aff - smu + lam dt) / t  (c.dtt / c.dlam hold the affine
// step on entry).  Returns the largest step in (0, 1] that keeps t and lam non-negative (all lanes).  Ends with a barrier.
Q3D double ipm_row_steps3(const Ctx3& c, const int corrector, const double smu) {
  double a = 1.0;
  auto row = [&](int j, double av) {
    const double t = c.tt[j], l = c.lam[j];
    const double dt = -(c.res[j] + t) - av;
    double dl;
    if (corrector) dl = -((fma(c.dtt[j], c.dlam[j], l * t) - smu) + l * dt) * q3_rcp(t);
    else dl = -(l + c.D[j] * dt);
    c.dtt[j] = dt; c.dlam[j] = dl;
    if (fma(a, dt, t) < 0.0) a = -t * q3_rcp(dt);   // dt < 0 and -t / dt < a
    if (fma(a, dl, l) < 0.0) a = -l * q3_rcp(dl);
  };
  Q3_FOR(k, c.N) {
    const double* dk = c.dx + 9 * k;
    const double* sk = c.dsig + 3 * k;
    const int j0 = k * c.rst, jb = c.Rs + 4 * k;
    row(j0, -dk[0]); row(j0 + 1, dk[0] + sk[0]); row(j0 + 2, dk[3] + sk[1]); row(j0 + 3, -dk[3] + sk[1]);
#pragma unroll 1
    for (int i = 0; i < c.nn; ++i) {
      const int q = k * c.nn + i;
      if (i < c.cnt) row(j0 + 4 + i, c.ax[q] * dk[7] + c.ay[q] * dk[8] - sk[2]);
      else { c.dtt[j0 + 4 + i] = 0.0; c.dlam[j0 + 4 + i] = 0.0; }
    }
    const double d0 = c.du[2 * k], d1 = c.du[2 * k + 1];
    row(jb, d0); row(jb + 1, -d0); row(jb + 2, d1); row(jb + 3, -d1);
  }
  a = q3_wmin(a);
  Q3_SYNC();
  return a;
}

// Mehrotra predictor-corrector down to complementarity mu_tol.  warm: continue from the parked iterate.
// One step length for primal and dual (the Hessian couples them): every linear residual - dual, primal, dynamics -
// then contracts by exactly (1 - alpha) per iteration, which is how the dual residual is tracked (its costates are
// implicit in the Riccati sweep).  Returns 0 converged, 1 diverged, 2 iteration cap, 3 non-finite.
Q3N int ipm3(double mu_tol, int warm, double* park) {
  Q3_CTX(c);
  Hdr3* hd = c.hd;
  const int ipm_max = hd->ipm_max;
  const int m = c.N * (4 + c.cnt) + 4 * c.N;
  const double inv_rows = 1.0 / (double)m;
  double lin = 1.0;  // product of (1 - alpha): scale of the linear residuals relative to the phase's start
  if (warm) {
    ipm_park3(c, park, 0);
  } else {
    // ---- starting point: least squares with unit row weights from a bounded point, then shift into the cone
    Q3_FOR(j, c.RT) { c.D[j] = row_ok3(c, j) ? 1.0 : 0.0; c.lam[j] = 0.0; }
    Q3_FOR(j, 9 * c.N) c.x[j] = hd->x0[j % 9];
    Q3_FOR(j, 2 * c.N) c.u[j] = 0.0;
    Q3_FOR(j, 3 * c.N) c.sig[j] = 0.0;
    Q3_SYNC();
    eval_residuals3(c);
    factor_solve3(G3_ALM, 1);
    primal_step3(c, 1.0);
    eval_residuals3(c);
    const double qsm = 0.5 * fmax(hd->Qs2[0], fmax(hd->Qs2[4], hd->Qs2[8]));
    // multipliers of violated soft rows end near 2 Qs sigma: start the duals at 1e-4 max(Qs) rather than O(1)
    const double lam_scale = fmax(1.0, 1e-4 * qsm);
    double mt = 1e300, ml = 1e300;
    Q3_FOR(j, c.RT)
      if (row_ok3(c, j)) { mt = fmin(mt, -c.res[j]); ml = fmin(ml, c.res[j]); }
    mt = q3_wmin(mt); ml = q3_wmin(ml);
    Q3_FOR(j, c.RT) {
      if (!row_ok3(c, j)) { c.tt[j] = 1.0; c.lam[j] = 0.0; continue; }
      double t = -c.res[j], l = c.res[j];
      if (mt < 0.0) t += 1.0 - mt;
      if (ml < 0.0) l += 1.0 - ml;
      c.tt[j] = fmax(t, 1e-8); c.lam[j] = lam_scale * fmax(l, 1e-8);
    }
    Q3_SYNC();
  }
  // scale of the dual residual at the start: the least-squares point is stationary for lam = res, so what is left is
  // C'(lam - res); bounded by the multiplier scale (a warm start continues a converged phase: same bound)
  double v_l = 0.0;
  Q3_FOR(j, c.RT) if (row_ok3(c, j)) v_l = fmax(v_l, c.lam[j]);
  const double rd0 = (double)(4 + c.cnt) * (1.0 + q3_wmax(v_l));
  double gmax = 1.0 + fabs(hd->q0);
#pragma unroll 1
  for (int it = 0; it < ipm_max; ++it) {
    double v_mu = 0.0, v_rp = 0.0, v_lm = 0.0;
    ipm_rows_eval3(c, v_mu, v_rp, v_lm);
    double v_obj = 0.0;
    Q3_FOR(k, c.N) {  // scale of the objective (the exact value is computed by the epilogue)
      const double* xk = c.x + 9 * k;
#pragma unroll 1
      for (int q = 0; q < hd->nq; ++q) { const int e = hd->nzl[q], i = hd->tri[e] >> 4; if (i == (hd->tri[e] & 15)) v_obj += xk[i] * (0.5 * hd->Q2u[e] * xk[i]); }
      v_obj += hd->q0 * xk[0];
    }
    const double mu = q3_wsum(v_mu) * inv_rows;
    const double rp = q3_wmax(v_rp), lmax = q3_wmax(v_lm);
    const double obj = fabs(q3_wsum(v_obj));
    const double rd = lin * rd0;
    if (Q3_LANE == 0) ++hd->n_ipm;
    if (!isfinite(mu) || !isfinite(rp)) return 3;
    if (!(mu < 1e40) || !(rp < 1e40)) return 1;
    if (mu <= mu_tol * (1.0 + obj) && rp <= 10.0 * mu_tol && rd <= 10.0 * mu_tol * (gmax + lmax)) return 0;
    // ---- predictor (factorisation for D = lam / t, set by ipm_rows_eval3)
    factor_solve3(G3_AFF, 0);
    const double aa = ipm_row_steps3(c, 0, 0.0);
    double v_aff = 0.0;
    Q3_FOR(j, c.RT) v_aff = fma(c.lam[j] + aa * c.dlam[j], c.tt[j] + aa * c.dtt[j], v_aff);
    const double mu_aff = q3_wsum(v_aff) * inv_rows;
    const double sr = mu_aff * q3_rcp(mu), smu = sr * sr * sr * mu;
    // ---- corrector: same factor, full right-hand side with the second-order term (smu - dt_aff dlam_aff) / t
    affine_solve3(G3_COR, smu);
    const double al = fmin(1.0, 0.99 * ipm_row_steps3(c, 1, smu));
#ifdef QP3_TRACE3
    printf("  ipm it %d mu %.3e rp %.3e lin %.2e aa %.3f sig %.3f al %.4f lmax %.2e\n", it, mu, rp, lin, aa, sr*sr*sr, al, lmax);
#endif
    primal_step3(c, al);
    Q3_FOR(j, c.RT) {
      c.tt[j] = fma(al, c.dtt[j], c.tt[j]);
      c.lam[j] = fma(al, c.dlam[j], c.lam[j]);
    }
    Q3_SYNC();
    lin *= 1.0 - al;
    if (Q3_LANE == 0) hd->ipm_al = al;
    Q3_SYNC();
  }
  return 2;
}

// Working-set guess from the interior-point iterate.  Tapia indicators (the ratios of successive iterates: lam+/lam -> 1
// and t+/t -> 0 on an active row, the other way round on an inactive one) are scale-free, unlike lam > t: a row that is
// nearly active (t ~ 1e-6) next to multipliers of 1e7 is told apart decades of mu earlier.  Ends with a barrier.
Q3D void ipm_guess3(const Ctx3& c) {
  const double al = c.hd->ipm_al;
  Q3_FOR(j, c.RT) {
    int a = 0;
    if (row_ok3(c, j)) {
      if (al > 0.0) {
        const double l0 = c.lam[j] - al * c.dlam[j], t0 = c.tt[j] - al * c.dtt[j];
        a = (l0 > 0.0 && t0 > 0.0) ? (c.lam[j] * t0 > c.tt[j] * l0) : (c.lam[j] > c.tt[j]);
      } else {
        a = c.lam[j] > c.tt[j];
      }
    }
    c.act[j] = (unsigned char)a;
  }
  Q3_SYNC();
}

// Full solve.  Returns dmpc_status; the header counts the work done.  `park`: the QP's global scratch slice.
Q3D int solve_qp3(const Ctx3& c, double* park, const unsigned char* warm, const int hard_hint) {
  Hdr3* hd = c.hd;
  // first guess: the working set the previous solve of this QP certified (consecutive MPC steps are alike: usually it IS
  // the answer and one round certifies it), else the empty set; the rounds correct a wrong guess either way.  The
  // iterate starts from a bounded (not dynamics-consistent) point; the first solve removes the dynamics defects.
  int any_guess = 0;
  Q3_FOR(j, c.RT) {
    const int a = warm ? (warm[j] != 0 && row_ok3(c, j)) : 0;
    c.act[j] = (unsigned char)a;
    any_guess |= a;
  }
  any_guess = q3_wor(any_guess);
  Q3_FOR(j, 9 * c.N) c.x[j] = hd->x0[j % 9];
  Q3_FOR(j, 2 * c.N) c.u[j] = 0.0;
  Q3_FOR(j, 3 * c.N) c.sig[j] = 0.0;
  Q3_SYNC();
  {
    // budget: 3 pdas_first rounds (as the condensed engine's free rounds); a QP that ended in the interior point last
    // time gets one round on its previous working set and then goes there directly (95 % of them end there again)
        const int st = pdas3(hard_hint ? 1 : 3 * hd->pdas_first, 1, c.N + c.N / 2, any_guess ? 1 : 0);
    if (st == 1) return DMPC_SOLVED;
  }
  double mu_tol = hd->ipm_mu0, mu_ok = 0.0;
#pragma unroll 1
  while (mu_tol > 1e-13) {
    const int rc = ipm3(mu_tol, mu_ok > 0.0, park);
    if (rc == 1 || rc == 3) {  // divergence of the FIRST phase = infeasible hard rows; later phases only lose accuracy
      if (mu_ok == 0.0) return rc == 3 ? DMPC_UNSOLVED : DMPC_PRIMAL_INFEASIBLE;
      break;
    }
    if (rc == 0) { mu_ok = mu_tol; ipm_park3(c, park, 1); }
    ipm_guess3(c);
    if (pdas3(hd->pdas_after, 0, 0, 0) == 1) return DMPC_SOLVED;
    if (rc == 2) break;  // iteration cap: a tighter target from the same start would hit it again
    mu_tol *= 1e-2;
  }
  if (mu_ok > 0.0) {  // no active set certified (degenerate rows): hand back the tightest interior-point iterate
    ipm_park3(c, park, 0);
    Q3_FOR(j, c.RT) if (!row_ok3(c, j)) c.lam[j] = 0.0;
    Q3_SYNC();
    return DMPC_SOLVED_INACCURATE;
  }
  return DMPC_MAX_ITER_REACHED;
}

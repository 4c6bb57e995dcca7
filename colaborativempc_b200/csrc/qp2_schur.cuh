// qp2_schur.cuh - Schur-complement active-set rounds on ONE factorisation of the condensed Hessian.
//
// The working sets of this QP family are small (bench workload: median 2 rows, 95 % <= 11, almost all input-box rows),
// so instead of re-building and re-factorising K = H + A'DA for every active-set guess (qp2_solve.cuh), H is factorised
// once (L_H^-1 kept packed in c.Hi) and a round on the working set W solves the exact equality-constrained KKT system
//       H u + g + A_W' lam = 0,   2 Qs sig + E_W' lam = 0,   A_W u + E_W sig = b_W
// through its Schur complement     S lam = A_W u0 - b_W,   S = Y'Y + E_W (2Qs)^-1 E_W',   Y = L_H^-1 A_W'   (|W| x |W|),
// u = u0 - L_H^-T (Y lam),  sig = -(2Qs)^-1 E_W' lam.  No penalty, hence no multiplier refinement loop; the certificate
// and the primal-dual working-set update are the ones of the penalty rounds (certify_update2).  A working set larger than
// c.WS rows, or a numerically singular S (dependent hard rows), hands over to the penalty rounds (pdas2).
// Same role as OSQP's polish (LPV_Planner.py:233 polish=True), restated on the condensed problem.
#pragma once
#include "qp2_solve.cuh"

// slack slot / sign of stage row type r (row_val2): r=1 -> +sig0, r=2,3 -> +sig1, collision -> -sig2, r=0 -> none
__device__ __forceinline__ int slack_slot(int r) { return r == 0 ? -1 : (r == 1 ? 0 : (r < 4 ? 1 : 2)); }

// In-place Cholesky of the packed w x w matrix S by warp 0 (lane i owns row i).  Returns 1 if a pivot is not safely
// positive relative to its original diagonal.
__device__ __forceinline__ int small_chol(double* S, int w) {
  const int lane = threadIdx.x & 31;
  const double d_own = lane < w ? S[TRI(lane) + lane] : 1.0;
  int sing = 0;
#pragma unroll 1
  for (int k = 0; k < w; ++k) {
    __syncwarp();
    const double d = S[TRI(k) + k];
    const double dk0 = __shfl_sync(0xffffffffu, d_own, k);
    if (!(d > 1e-11 * dk0) || !(d > 1e-300)) { sing = 1; break; }
    const double ip = rsqrt(d);
    const bool mine = lane > k && lane < w;
    const double lik = mine ? S[TRI(lane) + k] * ip : 0.0;
    __syncwarp();
    if (mine) S[TRI(lane) + k] = lik;
    if (lane == k) S[TRI(k) + k] = d * ip;
    __syncwarp();
    if (mine) {
      double* row = S + TRI(lane);
#pragma unroll 1
      for (int j = k + 1; j <= lane; ++j) row[j] = fma(-lik, S[TRI(j) + k], row[j]);
    }
  }
  __syncwarp();
  return sing;
}

// Solves L L' x = b with the factor left by small_chol; warp 0, lane i passes b_i and receives x_i.
__device__ __forceinline__ double small_solve(const double* S, int w, double b) {
  const int lane = threadIdx.x & 31;
  double v = b;
#pragma unroll 1
  for (int k = 0; k < w; ++k) {
    const double xk = __shfl_sync(0xffffffffu, v, k) / S[TRI(k) + k];
    if (lane == k) v = xk;
    else if (lane > k && lane < w) v = fma(-S[TRI(lane) + k], xk, v);
  }
#pragma unroll 1
  for (int k = w - 1; k >= 0; --k) {
    const double xk = __shfl_sync(0xffffffffu, v, k) / S[TRI(k) + k];
    if (lane == k) v = xk;
    else if (lane < k) v = fma(-S[TRI(k) + lane], xk, v);
  }
  return v;
}

// x = S^-1 b from the inverse factor left in S by chol_inverse_blk (S^-1 = Linv' Linv); warp 0, lane i passes b_i
// and receives x_i (w <= 32).
__device__ __forceinline__ double inv_solve(const double* S, int w, double b) {
  const int lane = threadIdx.x & 31;
  const double* row = S + TRI(lane < w ? lane : 0);
  double y = 0.0;
#pragma unroll 1
  for (int j = 0; j < w; ++j) {
    const double bj = __shfl_sync(0xffffffffu, b, j);
    if (j <= lane && lane < w) y = fma(row[j], bj, y);
  }
  double x = 0.0;
#pragma unroll 1
  for (int i = 0; i < w; ++i) {
    const double yi = __shfl_sync(0xffffffffu, y, i);
    if (i >= lane && lane < w) x = fma(S[TRI(i) + lane], yi, x);
  }
  return x;
}

// Block-wide arg-max of (v, idx) (ties -> smaller idx); all threads get the winner.  Scratch: c.red / c.asl.
__device__ __forceinline__ int block_argmax(const Ctx2& c, double v, int idx, double* vmax) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double v2 = __shfl_xor_sync(0xffffffffu, v, o);
    const int i2 = __shfl_xor_sync(0xffffffffu, idx, o);
    if (v2 > v || (v2 == v && i2 < idx)) { v = v2; idx = i2; }
  }
  __syncthreads();
  if ((threadIdx.x & 31) == 0) { c.red[threadIdx.x >> 5] = v; c.asl[threadIdx.x >> 5] = idx; }
  __syncthreads();
  double bv = c.red[0];
  int bi = c.asl[0];
#pragma unroll
  for (int w = 1; w < QP_WARPS; ++w) {
    const double v2 = c.red[w];
    const int i2 = c.asl[w];
    if (v2 > bv || (v2 == bv && i2 < bi)) { bv = v2; bi = i2; }
  }
  *vmax = bv;
  return bi;
}

// Working-set update of the Schur rounds.  Returns 1 when (u, sig, lam) passes the KKT certificate, 2 when the set
// cannot grow any further (overflow), 0 otherwise (c.act updated).
//   normal mode: drop every row with a negative multiplier, add the violated rows - all of them if they fit in the WS
//                slots, else the most violated ones;
//   safe mode (after a repeated working set): one change per round - drop the most negative multiplier, else add the
//                most violated row.
__device__ __forceinline__ int schur_update(const Ctx2& c, int safe, int first) {
  const double cert_tol = c.hd->cert_tol;
  double lm = 0.0;
#pragma unroll 1
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) lm = fmax(lm, fabs(c.lam[j]));
  lm = 1.0 + block_max(lm, c.red);
  double* score = c.om;  // > 0: violated inactive row (relative violation); < 0: active row with a negative multiplier
  double cnt = 0.0;
#pragma unroll 1
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) {
    double sc = 0.0;
    if (row_valid2(c, j)) {
      const double hs = 1.0 + fabs(c.bv[j]);
      if (!(fabs(c.res[j]) < 1e300) || !(fabs(c.lam[j]) < 1e300)) cnt += 68719476736.0;  // NaN / inf never certifies
      else if (!c.act[j]) { if (c.res[j] > cert_tol * hs) { sc = c.res[j] / hs; cnt += 1.0; } }
      else if (c.lam[j] < -cert_tol * lm) { sc = c.lam[j]; cnt += 4096.0; }
      else cnt += 16777216.0;
    }
    score[j] = sc;
  }
  cnt = block_sum(cnt, c.red);
  const long long total = (long long)cnt;  // four 12-bit counters in one exact FP64 sum (RT < 4096)
  const int nviol = (int)(total & 4095), nneg = (int)((total >> 12) & 4095), nkeep = (int)((total >> 24) & 4095);
  if ((total >> 36) != 0) return 2;  // non-finite iterate: leave the rounds
  if (nviol == 0 && nneg == 0) return 1;
  // far more rows violated at the unconstrained optimum than the working set can hold: the rounds would end in overflow
  // (measured: of the QPs with more than 1.5 WS violated rows at u0, 95 % do) - do not spend them
  if (first && 2 * nviol > 3 * c.WS) return 2;
  if (!safe) {
    const int room = c.WS - nkeep;
    if (nviol <= room) {
#pragma unroll 1
      for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) { if (score[j] > 0.0) c.act[j] = 1; else if (score[j] < 0.0) c.act[j] = 0; }
    } else {
#pragma unroll 1
      for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) if (score[j] < 0.0) c.act[j] = 0;
      if (room <= 0 && nneg == 0) { __syncthreads(); return 2; }
#pragma unroll 1
      for (int k = 0; k < room; ++k) {
        double v = 0.0, vm;
        int bi = 0x7fffffff;
#pragma unroll 1
        for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) if (score[j] > v) { v = score[j]; bi = j; }
        bi = block_argmax(c, v, bi, &vm);
        if (threadIdx.x == 0) { c.act[bi] = 1; score[bi] = 0.0; }
        __syncthreads();
      }
    }
  } else {
    double v = 0.0, vm;
    int bi = 0x7fffffff;
    if (nneg > 0) {
#pragma unroll 1
      for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) if (-score[j] > v) { v = -score[j]; bi = j; }
      bi = block_argmax(c, v, bi, &vm);
      if (threadIdx.x == 0) c.act[bi] = 0;
    } else {
      if (nkeep >= c.WS) return 2;
#pragma unroll 1
      for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) if (score[j] > v) { v = score[j]; bi = j; }
      bi = block_argmax(c, v, bi, &vm);
      if (threadIdx.x == 0) c.act[bi] = 1;
    }
  }
  __syncthreads();
  return 0;
}

// L_H^-1 -> c.Hi, unconstrained optimum u0 = -H^-1 g -> c.rdu.  Ends with a barrier.
__device__ __noinline__ void schur_prepare2() {
  CTX2_LOCAL(c);
#pragma unroll 1
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) c.D[j] = 0.0;
  __syncthreads();
  factorise2();  // D = 0: K = H, then K <- L_H^-1
#pragma unroll 1
  for (int e = threadIdx.x; e < c.ntri; e += QP_THREADS) c.Hi[e] = c.K[e];
#pragma unroll 1
  for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) c.rdu[j] = -c.g[j];
  __syncthreads();
  chol_solve2(c, c.rdu);
}

// Up to `rounds` Schur-complement rounds from the working set c.act (all-zero at the first call).
// Returns 1 certified (c.u, c.sig, c.lam, c.res final), 0 round cap, 2 working set cannot grow within WS rows,
// 3 S numerically singular (dependent hard rows; c.act holds the current guess for the penalty rounds).
__device__ __noinline__ int schur_pdas2(int rounds, int safe_rounds) {
  CTX2_LOCAL(c);
  const double* Qsi = c.hd->Qsi;
  const double cert_tol = c.hd->cert_tol;
  const int nu = c.nu, ldy = c.ldy, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double* u0 = c.rdu;
  double* res0 = c.tt;
  int* nw = c.redi + 2;
  int* sing = c.redi + 3;
#pragma unroll 1
  for (int j = threadIdx.x; j < nu; j += QP_THREADS) c.u[j] = u0[j];
#pragma unroll 1
  for (int j = threadIdx.x; j < 3 * c.N; j += QP_THREADS) c.sig[j] = 0.0;
#pragma unroll 1
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) c.lam[j] = 0.0;
  __syncthreads();
  eval_residuals2();
#pragma unroll 1
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) res0[j] = c.res[j];  // every row at (u0, sig = 0)
  __syncthreads();
  int nh = 0, safe = 0;
#pragma unroll 1
  for (int r = 0; r < rounds + safe_rounds; ++r) {
    {
      const int st = schur_update(c, safe, r == 0);
      if (st) return st;
    }
    if (!safe) {
      const unsigned long long h = act_hash2(c);
      int cyc = 0;
      for (int i = 0; i < nh && i < 16; ++i) cyc |= (c.hist[i] == h);
      __syncthreads();
      if (threadIdx.x == 0) c.hist[nh & 15] = h;
      ++nh;
      if (cyc || r + 1 >= rounds) safe = 1;  // repeated working set (or too many rounds): one change per round from here
    }
    if (warp == 0) {  // working-set list, rows ascending (ballot compaction)
      int w = 0;
#pragma unroll 1
      for (int j0 = 0; j0 < c.RT; j0 += 32) {
        const int j = j0 + lane;
        const unsigned m = __ballot_sync(0xffffffffu, j < c.RT && c.act[j]);
        if ((m >> lane) & 1u) {
          const int pos = w + __popc(m & ((1u << lane) - 1u));
          if (pos < c.WS) c.wl[pos] = j;
        }
        w += __popc(m);
      }
      if (lane == 0) { *nw = w; ++((Hdr2*)smem_fast)->n_rounds; }
    }
    __syncthreads();
    const int w = *nw;
    if (w > c.WS) return 2;  // (cannot happen: schur_update keeps the set within WS)
    // ---- Y = L_H^-1 A_W'   (entry (rr, i) = row rr of L_H^-1 times row wl[i] of A)
#pragma unroll 1
    for (int e = threadIdx.x; e < w * nu; e += QP_THREADS) {
      const int i = e / nu, rr = e - i * nu, j = c.wl[i];
      const int ri = c.rinfo[j], rt = ri & 15, t = ri >> 4;
      const double* hr = c.Hi + TRI(rr);
      double s;
      if (j >= c.Rs) {
        const int col = 2 * t + (rt >> 1);
        s = rr >= col ? ((rt & 1) ? -hr[col] : hr[col]) : 0.0;
      } else {
        const int n = min(rr + 1, 2 * (t + 1));
        double s0 = 0.0, s1 = 0.0;
        if (rt < 4) {
          const double* p = &GB(c, t, rt >> 1, 0);
          int q = 0;
#pragma unroll 1
          for (; q + 1 < n; q += 2) { s0 = fma(hr[q], p[q], s0); s1 = fma(hr[q + 1], p[q + 1], s1); }
          if (q < n) s0 = fma(hr[q], p[q], s0);
          s = (rt == 0 || rt == 3) ? -(s0 + s1) : (s0 + s1);
        } else {
          const int ic = j - t * c.rst - 4;
          const double* p = &GB(c, t, 2, 0);
          const double* pq = &GB(c, t, 3, 0);
#pragma unroll 1
          for (int q = 0; q < n; ++q) { s0 = fma(hr[q], p[q], s0); s1 = fma(hr[q], pq[q], s1); }
          s = c.ax[t * c.nn + ic] * s0 + c.ay[t * c.nn + ic] * s1;
        }
      }
      c.Y[rr * ldy + i] = s;
    }
    __syncthreads();
    // ---- S = Y'Y + E_W (2Qs)^-1 E_W'  (packed lower), right-hand side = residual of the rows at u0
#pragma unroll 1
    for (int p = threadIdx.x; p < ((w * (w + 1)) >> 1); p += QP_THREADS) {
      const int ij = c.tri[p], i = ij >> 8, k = ij & 255;
      double s0 = 0.0, s1 = 0.0;
      int q = 0;
#pragma unroll 1
      for (; q + 1 < nu; q += 2) {
        s0 = fma(c.Y[q * ldy + i], c.Y[q * ldy + k], s0);
        s1 = fma(c.Y[(q + 1) * ldy + i], c.Y[(q + 1) * ldy + k], s1);
      }
      double s = s0 + s1;
      const int ji = c.wl[i], jk = c.wl[k];
      if (ji < c.Rs && jk < c.Rs) {
        const int ra = c.rinfo[ji], rb = c.rinfo[jk];
        if ((ra >> 4) == (rb >> 4)) {
          const int sa = slack_slot(ra & 15), sb = slack_slot(rb & 15);
          if (sa >= 0 && sb >= 0) s += ((sa == 2) != (sb == 2) ? -1.0 : 1.0) * Qsi[sa * 3 + sb];
        }
      }
      c.S[p] = s;
    }
    if (threadIdx.x < w) {
      c.rw[threadIdx.x] = res0[c.wl[threadIdx.x]];
      c.rs[threadIdx.x] = c.S[TRI(threadIdx.x) + threadIdx.x];
    }
    if (threadIdx.x == 0) *sing = 0;
    __syncthreads();
    // small working sets: one warp, right-looking in place (S <- L); larger ones: the blocked routine (S <- L^-1)
    const bool blocked = w > 12;
    if (blocked) {
      chol_inverse_sh((int)(c.S - c.Gb), w, (int)(c.dlam - c.Gb), (int)(c.rs - c.Gb));
    } else if (warp == 0) {
      const int sg = small_chol(c.S, w);
      if (lane == 0) *sing = sg;
    }
    if (warp == 0) {
      __syncwarp();
      if (!*(volatile int*)sing) {
        const double b = lane < w ? c.rw[lane] : 0.0;
        const double x = blocked ? inv_solve(c.S, w, b) : small_solve(c.S, w, b);
        if (lane < w) c.lw[lane] = x;
      }
    }
    __syncthreads();
    if (*sing) return 3;
    // ---- back-substitute; up to two corrections of lam_W from the residual of the working-set rows
#pragma unroll 1
    for (int pass = 0; pass < 3; ++pass) {
#pragma unroll 1
      for (int q = threadIdx.x; q < nu; q += QP_THREADS) {
        double s = 0.0;
        for (int i = 0; i < w; ++i) s = fma(c.Y[q * ldy + i], c.lw[i], s);
        c.yt[q] = s;
      }
#pragma unroll 1
      for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) c.lam[j] = 0.0;
      __syncthreads();
#pragma unroll 1
      for (int j = threadIdx.x; j < nu; j += QP_THREADS) {
        double s0 = 0.0, s1 = 0.0;
        int i = j;
#pragma unroll 1
        for (; i + 1 < nu; i += 2) { s0 = fma(c.Hi[TRI(i) + j], c.yt[i], s0); s1 = fma(c.Hi[TRI(i + 1) + j], c.yt[i + 1], s1); }
        if (i < nu) s0 = fma(c.Hi[TRI(i) + j], c.yt[i], s0);
        c.u[j] = u0[j] - (s0 + s1);
      }
#pragma unroll 1
      for (int t = threadIdx.x; t < c.N; t += QP_THREADS) {
        double a0 = 0.0, a1 = 0.0, a2 = 0.0;
        for (int i = 0; i < w; ++i) {
          const int j = c.wl[i];
          if (j >= c.Rs) continue;
          const int ri = c.rinfo[j];
          if ((ri >> 4) != t) continue;
          const int sl = slack_slot(ri & 15);
          if (sl == 0) a0 += c.lw[i]; else if (sl == 1) a1 += c.lw[i]; else if (sl == 2) a2 -= c.lw[i];
        }
        c.sig[3 * t + 0] = -(Qsi[0] * a0 + Qsi[1] * a1 + Qsi[2] * a2);
        c.sig[3 * t + 1] = -(Qsi[3] * a0 + Qsi[4] * a1 + Qsi[5] * a2);
        c.sig[3 * t + 2] = -(Qsi[6] * a0 + Qsi[7] * a1 + Qsi[8] * a2);
      }
      if (threadIdx.x < w) c.lam[c.wl[threadIdx.x]] = c.lw[threadIdx.x];
      __syncthreads();
      eval_residuals2();
      double v = 0.0;
      if (threadIdx.x < w) { const int j = c.wl[threadIdx.x]; v = fabs(c.res[j]) / (1.0 + fabs(c.bv[j])); }
      v = block_max(v, c.red);
      if (v <= 0.1 * cert_tol || pass == 2) break;
      if (warp == 0) {
        const double b = lane < w ? c.res[c.wl[lane]] : 0.0;
        const double x = blocked ? inv_solve(c.S, w, b) : small_solve(c.S, w, b);
        if (lane < w) c.lw[lane] += x;
      }
      __syncthreads();
    }
  }
  return 0;
}

// Full solve.  Returns dmpc_status; the header counts the work done (n_rounds, n_ipm).
__device__ __forceinline__ int solve_qp2(const Ctx2& c) {
#pragma unroll 1
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) c.act[j] = 0;
  __syncthreads();
  schur_prepare2();
  // 3 pdas_first rounds of free updates, then at most 2 pdas_first one-change rounds: past that an interior-point solve
  // (about 15 factorisations) is the cheaper way out
  const int rs = schur_pdas2(3 * c.hd->pdas_first, 2 * c.hd->pdas_first);
  if (rs == 1) return DMPC_SOLVED;
  // singular S: the penalty rounds tolerate dependent rows.  Overflow / round cap: large or cycling working sets, for
  // which the penalty rounds almost never certify either (measured) - straight to the interior-point phase.
  if (rs == 3 && pdas2(c.hd->pdas_first)) return DMPC_SOLVED;
  // Interior-point ladder: each phase CONTINUES from the iterate the previous one converged to (kept in c.sv: the
  // active-set rounds in between overwrite u, sigma, lambda), so a tighter target costs a few iterations, not a restart.
  double mu_tol = c.hd->ipm_mu0, mu_ok = 0.0;
#pragma unroll 1
  while (mu_tol > 1e-13) {
    const int rc = ipm2(mu_tol, mu_ok > 0.0);
    if (rc == 1 || rc == 3) {  // divergence of the FIRST phase = infeasible hard rows; later (tighter) phases can only lose accuracy
      if (mu_ok == 0.0) return rc == 3 ? DMPC_UNSOLVED : DMPC_PRIMAL_INFEASIBLE;  // NaN is a numerical failure, not a certificate
      break;
    }
    if (rc == 0) { mu_ok = mu_tol; ipm_keep2(c, 1); }
#pragma unroll 1
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) c.act[j] = row_valid2(c, j) && (c.lam[j] > c.tt[j]);
    __syncthreads();
    if (pdas2(c.hd->pdas_after)) return DMPC_SOLVED;
    if (rc == 2) break;  // iteration cap: a tighter target from the same start would hit it again
    mu_tol *= 1e-2;
  }
  if (mu_ok > 0.0) {  // no active set certified (degenerate rows): hand back the tightest interior-point iterate
    ipm_keep2(c, 0);
#pragma unroll 1
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) if (!row_valid2(c, j)) c.lam[j] = 0.0;
    __syncthreads();
    return DMPC_SOLVED_INACCURATE;
  }
  return DMPC_MAX_ITER_REACHED;
}

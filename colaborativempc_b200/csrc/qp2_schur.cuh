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
// Returns 1 certified (c.u, c.sig, c.lam, c.res final), 0 working set repeated / round cap, 2 working set too large or
// S singular (c.act holds the current guess for the penalty rounds).
__device__ __noinline__ int schur_pdas2(int rounds) {
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
  int nh = 0;
#pragma unroll 1
  for (int r = 0; r < rounds; ++r) {
    if (certify_update2(c)) return 1;
    {
      const unsigned long long h = act_hash2(c);
      int cyc = 0;
      for (int i = 0; i < nh && i < 16; ++i) cyc |= (c.hist[i] == h);
      __syncthreads();
      if (threadIdx.x == 0 && nh < 16) c.hist[nh] = h;
      ++nh;
      if (cyc) { __syncthreads(); return 0; }
    }
    if (threadIdx.x == 0) {  // working-set list
      int w = 0;
      for (int j = 0; j < c.RT; ++j)
        if (c.act[j]) { if (w < c.WS) c.wl[w] = j; ++w; }
      *nw = w;
      ++((Hdr2*)smem_fast)->n_rounds;
    }
    __syncthreads();
    const int w = *nw;
    if (w > c.WS) return 2;
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
    if (threadIdx.x < w) c.rw[threadIdx.x] = res0[c.wl[threadIdx.x]];
    __syncthreads();
    if (warp == 0) {
      const int sg = small_chol(c.S, w);
      if (!sg) {
        const double x = small_solve(c.S, w, lane < w ? c.rw[lane] : 0.0);
        if (lane < w) c.lw[lane] = x;
      }
      if (lane == 0) *sing = sg;
    }
    __syncthreads();
    if (*sing) return 2;
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
        const double x = small_solve(c.S, w, lane < w ? c.res[c.wl[lane]] : 0.0);
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
  {
    const int rs = schur_pdas2(3 * c.hd->pdas_first);
    if (rs == 1) return DMPC_SOLVED;
    if (rs == 2 && pdas2(c.hd->pdas_first)) return DMPC_SOLVED;
  }
  double mu_tol = c.hd->ipm_mu0, mu_ok = 0.0;
#pragma unroll 1
  while (mu_tol > 1e-13) {
    const int rc = ipm2(mu_tol);
    if (rc == 1) {  // divergence of the FIRST phase = infeasible hard rows; later (tighter) phases can only lose accuracy
      if (mu_ok == 0.0) return DMPC_PRIMAL_INFEASIBLE;
      break;
    }
    if (rc == 0) mu_ok = mu_tol;
#pragma unroll 1
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) c.act[j] = row_valid2(c, j) && (c.lam[j] > c.tt[j]);
    __syncthreads();
    if (pdas2(c.hd->pdas_after)) return DMPC_SOLVED;
    mu_tol *= 1e-2;
  }
  if (mu_ok > 0.0) {  // no active set certified (degenerate rows): hand back the tightest interior-point iterate
    ipm2(mu_ok);
#pragma unroll 1
    for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) if (!row_valid2(c, j)) c.lam[j] = 0.0;
    __syncthreads();
    return DMPC_SOLVED_INACCURATE;
  }
  return DMPC_MAX_ITER_REACHED;
}

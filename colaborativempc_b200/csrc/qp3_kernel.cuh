// qp3_kernel.cuh - the per-agent LPV-MPC solve on the stage-wise engine: one WARP per QP (a one-warp CTA), the O(N)
// workspace in shared memory, everything between the inputs of PlannerLPV.solve and its outputs
// (distributedPlanner/LPV_Planner.py:115-182) on chip - and, when asked, the caller's bookkeeping of
// planner/scripts/LPV_HP_N_main.py:102-117 (shift, uOld, packed (x,y) for the exchange, stop on a failed solve) in the
// same epilogue, so that the next DMPC sweep starts from buffers this kernel wrote.
#pragma once
#include "qp3_solve.cuh"

#define QP3_THREADS 32

Q3D void solve_one_qp3(const DevParams& P, const BatchPtrs& io, const int b) {
  const int N = P.N, nn = P.nn;
  const int n = 16 * N + 12;
  {
    Hdr3* hd = (Hdr3*)Q3_SMEM;
    if (Q3_LANE == 0) {
      int cnt = io.nbr_count ? io.nbr_count[b] : nn;
      hd->N = N; hd->nn = nn; hd->cnt = cnt < 0 ? 0 : (cnt > nn ? nn : cnt); hd->rows = io.xprev_rows;
      hd->refine = P.refine; hd->pdas_first = P.pdas_first; hd->pdas_after = P.pdas_after; hd->ipm_max = P.ipm_max;
      hd->n_rounds = 0; hd->n_ipm = 0; hd->n_fact = 0; hd->n_solve = 0;
      hd->delta = P.delta; hd->cert_tol = P.cert_tol; hd->ipm_mu0 = P.ipm_mu0; hd->ipm_al = 0.0;
      hd->q0 = 2.0 * (-P.vx_ref * P.Q00);
      hd->lim[0] = P.min_vel; hd->lim[1] = P.max_vel; hd->lim[2] = P.max_rs; hd->lim[3] = P.max_ls;
      hd->lim[4] = P.max_ac; hd->lim[5] = P.max_dc; hd->mind = P.min_dist;
      for (int i = 0; i < 4; ++i) { hd->R2[i] = 2.0 * P.R[i]; hd->dR2[i] = 2.0 * P.dR[i]; }
      for (int i = 0; i < 9; ++i) hd->Qs2[i] = 2.0 * P.Qs[i];
      hd->uold[0] = io.uold[(size_t)b * 2]; hd->uold[1] = io.uold[(size_t)b * 2 + 1];
      int nq = 0, e = 0;
      for (int i = 0; i < 9; ++i) for (int j = i; j < 9; ++j, ++e) {
        hd->Q2u[e] = 2.0 * P.Qf[i * 9 + j];
        if (P.Qf[i * 9 + j] != 0.0) hd->nzl[nq++] = (unsigned char)e;
      }
      hd->nq = nq;
    }
    Q3_FOR(i, 9) hd->x0[i] = io.x0[(size_t)b * NSTATE + i];
    Q3_SYNC();
  }
  Q3_CTX(c);
  Hdr3* hd = c.hd;
  const double* xprev = io.xprev + (size_t)b * io.xprev_rows * NSTATE;
  const double* uprev = io.uprev + (size_t)b * N * 2;
  const double* nbr = io.nbr_xy + (size_t)b * (N + 1) * nn * 2;
  const double* pose = io.pose + (size_t)b * (N + 1) * 2;
#ifndef QP3_HOST
  // The QP's ~2.9 KB of stacked problem data staged in shared memory by bulk async copies (TMA, cp.async.bulk +
  // mbarrier): four copies issued by one lane land while the tables are built, instead of ~90 dependent per-lane
  // loads from HBM (or, with page-locked host buffers, over PCIe).  The staging area is workspace that is not live
  // before the solve (lam .. res).  A copy covers the 16-byte aligned interior of its array; the odd head / tail
  // double (Last_xPredicted rows are 72 bytes) is loaded plainly, so nothing outside the QP's own slice is read.
  if (io.bulk_in && io.xprev_rows <= N + 1) {
    const double* src[4] = {xprev, uprev, nbr, pose};
    const int len[4] = {io.xprev_rows * NSTATE, 2 * N, 2 * (N + 1) * nn, 2 * (N + 1)};
    double* dstv[4];
    double* stg = c.lam + (((size_t)c.lam >> 3) & 1);  // 16-byte aligned
    const unsigned bar = (unsigned)__cvta_generic_to_shared(&hd->mbar);
    if (Q3_LANE == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    unsigned total = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int head = (int)(((size_t)src[q] >> 3) & 1);        // doubles before the first 16-byte boundary
      dstv[q] = stg + head;                                      // same alignment class as the source
      stg += (len[q] + head + 1) & ~1;
      const int body = (len[q] - head) & ~1;                    // doubles of the aligned interior
      if (body > 0) total += (unsigned)body * 8u;
    }
    if (Q3_LANE == 0) {
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(total) : "memory");
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int head = (int)(((size_t)src[q] >> 3) & 1);
        const int body = (len[q] - head) & ~1;
        if (body > 0)
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                           (unsigned)__cvta_generic_to_shared(dstv[q] + head)),
                       "l"(src[q] + head), "r"((unsigned)body * 8u), "r"(bar)
                       : "memory");
      }
    }
    // head / tail doubles: lanes 0-3 one array each
    if (Q3_LANE < 4) {
      const int q = Q3_LANE;
      const double* s_ = q == 0 ? src[0] : (q == 1 ? src[1] : (q == 2 ? src[2] : src[3]));
      double* d_ = q == 0 ? dstv[0] : (q == 1 ? dstv[1] : (q == 2 ? dstv[2] : dstv[3]));
      const int n_ = q == 0 ? len[0] : (q == 1 ? len[1] : (q == 2 ? len[2] : len[3]));
      const int head = (int)(((size_t)s_ >> 3) & 1);
      const int body = (n_ - head) & ~1;
      if (head && n_ > 0) d_[0] = s_[0];
      if (head + body < n_) d_[n_ - 1] = s_[n_ - 1];
    }
    build_tables3(c);  // overlaps the copies
    unsigned done = 0;
    while (!done)
      asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(bar) : "memory");
    __syncwarp();
    xprev = dstv[0]; uprev = dstv[1]; nbr = dstv[2]; pose = dstv[3];
  } else
#endif
  build_tables3(c);

  // ---------------- problem build ----------------
  int bad_plane = setup_planes3(c, P, nbr, pose, io.planes ? io.planes + (size_t)b * N * 3 * nn : nullptr);
  int bad_map = setup_linearise3(c, P, xprev, io.xprev_rows, uprev);
  bad_plane = q3_wor(bad_plane);
  bad_map = q3_wor(bad_map);
  Q3_SYNC();
  int status;
  if (bad_plane || bad_map) {  // the reference raises here; report and leave NaNs
#ifdef QP3_HOST
    const double qnan = NAN;
#else
    const double qnan = __longlong_as_double(0x7ff8000000000000ULL);
#endif
    if (io.z) Q3_FOR(j, n) io.z[(size_t)b * n + j] = qnan;
    if (io.y_ineq) Q3_FOR(j, c.RT) io.y_ineq[(size_t)b * c.RT + j] = qnan;
    if (io.y_eq) Q3_FOR(j, 14 * N + 12) io.y_eq[(size_t)b * (14 * N + 12) + j] = qnan;
    status = bad_map ? DMPC_MAP_LOOKUP_FAILED : DMPC_DEGENERATE_PLANE;
    if (Q3_LANE == 0) {
      if (io.obj) io.obj[b] = qnan;
      if (io.iters) io.iters[b] = 0;
      if (io.pivots) io.pivots[b] = 0;
      if (io.cost) io.cost[b] = 0;
    }
  } else {
    // ---------------- solve ----------------
    double* park = io.gws + (size_t)b * io.gws_stride;
    // what the previous call of this batch size spent on this QP index (valid exactly when the launch order is)
    const int hard_hint = io.order && io.cost && ((io.cost[b] >> 21) & 1);
    status = solve_qp3(c, park, io.warm ? io.warm + (size_t)b * c.RT : nullptr, hard_hint);
    if (io.warm) {  // the certified working set is the next call's first guess (nothing to offer if not certified)
      unsigned char* w = io.warm + (size_t)b * c.RT;
      const int ok = status == DMPC_SOLVED;
      Q3_FOR(j, c.RT) {
        int src = j;
        if (io.warm_shift) {  // stage k of the next step is stage k+1 of this one; the last stage repeats
          const int per = j < c.Rs ? c.rst : 4, base = j < c.Rs ? 0 : c.Rs;
          const int kk = (j - base) / per;
          if (kk < N - 1) src = j + per;
        }
        w[j] = ok ? c.act[src] : 0;
      }
    }

    // ---------------- outputs (LPV_Planner.py:162-180 layout of res.x) ----------------
    if (io.z) {
      double* z = io.z + (size_t)b * n;
      Q3_FOR(w, 12 * (N + 1)) {
        const int t = w / 12, r = w - 12 * t;
        double v;
        if (t == 0) v = r < 9 ? hd->x0[r] : 0.0;
        else v = r < 9 ? c.x[9 * (t - 1) + r] : c.sig[3 * (t - 1) + (r - 9)];
        z[w] = v;
      }
      Q3_FOR(j, 2 * N) {
        z[12 * (N + 1) + j] = c.u[j];
        z[12 * (N + 1) + 2 * N + j] = c.u[j] - (j < 2 ? hd->uold[j] : c.u[j - 2]);
      }
    }
    if (io.y_ineq) Q3_FOR(j, c.RT) io.y_ineq[(size_t)b * c.RT + j] = c.lam[j];
    if (io.obj) {  // objective 1/2 z'Pz + q'z (LPV_Planner.py:382-427)
      double v = 0.0;
      Q3_FOR(t, N + 1) {
        const double* x = t == 0 ? hd->x0 : c.x + 9 * (t - 1);
        double s = 0.0;
#pragma unroll 1
        for (int q = 0; q < hd->nq; ++q) {
          const int e = hd->nzl[q], i = hd->tri[e] >> 4, j = hd->tri[e] & 15;
          s = fma((i == j ? 0.5 : 1.0) * hd->Q2u[e] * x[i], x[j], s);
        }
        s = fma(hd->q0, x[0], s);
        if (t >= 1) {
          const int k = t - 1;
          s += 2.0 * (c.pxy[2 * k] * x[7] + c.pxy[2 * k + 1] * x[8]);
          const double* sg = c.sig + 3 * k;
          for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) s = fma(0.5 * hd->Qs2[i * 3 + j] * sg[i], sg[j], s);
          const double u0 = c.u[2 * k], u1 = c.u[2 * k + 1];
          const double d0 = u0 - (k == 0 ? hd->uold[0] : c.u[2 * k - 2]), d1 = u1 - (k == 0 ? hd->uold[1] : c.u[2 * k - 1]);
          s += 0.5 * (hd->R2[0] * u0 * u0 + (hd->R2[1] + hd->R2[2]) * u0 * u1 + hd->R2[3] * u1 * u1);
          s += 0.5 * (hd->dR2[0] * d0 * d0 + (hd->dR2[1] + hd->dR2[2]) * d0 * d1 + hd->dR2[3] * d1 * d1);
        }
        v += s;
      }
      v = q3_wsum(v);
      if (Q3_LANE == 0) io.obj[b] = v;
    }
    // multipliers of the dynamics rows (rows of G, LPV_Planner.py:429-475) = costates of the Riccati sweep at the
    // solution: one affine backward pass on the Lagrangian gradient with the factor of the last solve
    if (io.y_eq) {
      double* ye = io.y_eq + (size_t)b * (14 * N + 12);
      assemble_gradient3(c, G3_LAG, 0.0);
      ric_affine_backward3(c, ye, status == DMPC_SOLVED);
      Q3_FOR(t, N + 1) { ye[12 * t + 9] = 0.0; ye[12 * t + 10] = 0.0; ye[12 * t + 11] = 0.0; }
      Q3_FOR(k, N) {
        const double d0 = c.u[2 * k] - (k == 0 ? hd->uold[0] : c.u[2 * k - 2]);
        const double d1 = c.u[2 * k + 1] - (k == 0 ? hd->uold[1] : c.u[2 * k - 1]);
        const double m0 = hd->dR2[0] * d0 + hd->dR2[1] * d1, m1 = hd->dR2[2] * d0 + hd->dR2[3] * d1;
        ye[12 * (N + 1) + 2 * k] = k == 0 ? m0 : -m0;
        ye[12 * (N + 1) + 2 * k + 1] = k == 0 ? m1 : -m1;
      }
    }
    if (Q3_LANE == 0) {
      if (io.iters) io.iters[b] = hd->n_ipm;
      if (io.pivots) io.pivots[b] = hd->n_rounds;
      if (io.cost) io.cost[b] = (3 * hd->n_fact + hd->n_solve) | (1 << 20) | (hd->n_ipm > 0 ? 1 << 21 : 0);  // bit 20: stage-wise cost scale, bit 21: interior point used
    }
  }
  if (Q3_LANE == 0 && io.status) io.status[b] = status;

  // ---------------- fused caller bookkeeping (LPV_HP_N_main.py:102-117, LPV_Planner.py:179-180) ----------------
  // feasible (LPV_Planner.py:246-248: solved, solved inaccurate, max-iter): x0 <- xPred[1], x_old <- xPred[1:],
  // u_old <- uPred (not shifted), uOld <- uPred[0], (x,y) of the whole prediction packed for the exchange.
  // Not feasible: the reference stops the experiment (LPV_HP_N_main.py:102-104).  Here the agent's buffers are left as
  // they are - its last valid plan stays in the exchange, nothing non-finite leaves this QP - and hold[b] counts the
  // failed solves, which the caller turns into the per-scenario stop.
  if (io.nx_x0) {
    const bool feas = status == DMPC_SOLVED || status == DMPC_SOLVED_INACCURATE || status == DMPC_MAX_ITER_REACHED;
    if (feas) {
      Q3_FOR(i, 9) io.nx_x0[(size_t)b * 9 + i] = c.x[i];
      Q3_FOR(i, 9 * N) io.nx_xprev[(size_t)b * 9 * N + i] = c.x[i];
      Q3_FOR(i, 2 * N) io.nx_uprev[(size_t)b * 2 * N + i] = c.u[i];
      Q3_FOR(i, 2) io.nx_uold[(size_t)b * 2 + i] = c.u[i];
      Q3_FOR(i, 2 * (N + 1)) {
        const int t = i >> 1, q = i & 1;
        io.nx_xy[(size_t)b * 2 * (N + 1) + i] = t == 0 ? hd->x0[7 + q] : c.x[9 * (t - 1) + 7 + q];
      }
      if (Q3_LANE == 0 && io.hold) io.hold[b] = 0;
    } else if (Q3_LANE == 0 && io.hold) {
      io.hold[b] += 1;
    }
    if (io.flags && Q3_LANE == 0) {  // utilities/misc.py:28-48 on row 0 of this prediction (= the state solved from)
      const double s = hd->x0[6], L = io.end_s;
      const bool near = fabs(s - L) <= 0.15 + 1e-5 * fabs(L) || s > L;   // np.isclose(s, L, atol=0.15) or s > L
      const bool ended = near && floor(s / L) == (double)io.laps;
      io.flags[b] = (ended ? 1.0 : 0.0) + (feas ? 0.0 : 2.0);
    }
  }
}

#ifndef QP3_HOST
__global__ void __launch_bounds__(QP3_THREADS) lpv_qp_stage_kernel(const __grid_constant__ DevParams P, const BatchPtrs io) {
  if ((int)blockIdx.x >= io.B) return;
  const int b = io.order ? io.order[blockIdx.x] : blockIdx.x;  // most expensive QPs of the previous call first
  if (io.only_status && io.status[b] != io.only_status) return;
  if (io.stop && io.stop[b / io.stop_div]) return;  // the scenario's experiment is over: nothing of it changes any more
  solve_one_qp3(P, io, b);
}
#endif

// qp_kernel.cuh - the per-agent LPV-MPC solve kernel: one CTA per QP, everything between the inputs of
// PlannerLPV.solve and its outputs (distributedPlanner/LPV_Planner.py:115-182) on chip.
#pragma once
#include "qp_solve.cuh"

// Carves the per-CTA workspace.  `sm` = dynamic shared memory, `gw` = this QP's slice of the global workspace.
// Returns the shared-memory bytes used; *gdoubles (if not null) receives the global doubles used.
__host__ __device__ inline size_t carve_workspace(Ctx& c, char* sm, double* gw, int N, int nn, int NR, int gr_global,
                                                  int hk_global, long long* gdoubles) {
  c.N = N; c.nn = nn; c.nu = 2 * N; c.NR = NR;
  c.ldg = c.nu + 1; c.ldk = c.nu + 1; c.rst = 4 + nn; c.Rs = N * c.rst; c.RT = c.Rs + 4 * N;
  double* p = (double*)sm;
  double* q = gw;
  auto take = [&](size_t n) { double* r = p; p += n; return r; };
  auto gtake = [&](size_t n) { double* r = q; q += n; return r; };
  const size_t szGr = (size_t)N * NR * c.ldg, szZb = (size_t)N * 4 * c.ldg, szK = (size_t)c.nu * c.ldk;
  c.Gr = gr_global ? gtake(szGr) : take(szGr);
  c.Zb = gr_global ? gtake(szZb > (size_t)(N + 1) * 9 ? szZb : (size_t)(N + 1) * 9) : take(szZb > (size_t)(N + 1) * 9 ? szZb : (size_t)(N + 1) * 9);
  c.H = hk_global ? gtake(szK) : take(szK);
  c.K = hk_global ? gtake(szK) : take(szK);
  c.Acol = take(27 * N); c.Bcol = take(4 * N); c.eyb = take(N); c.pxy = take(2 * N);
  c.ax = take((size_t)N * nn); c.ay = take((size_t)N * nn); c.hh = take((size_t)N * nn);
  c.cr = take((size_t)N * NR); c.g = take(c.nu);
  c.u = take(c.nu); c.rhs = take(c.nu); c.du = take(c.nu); c.tv = take(c.nu); c.rdu = take(c.nu);
  c.lam = take(c.RT); c.tt = take(c.RT); c.D = take(c.RT); c.res = take(c.RT); c.om = take(c.RT);
  c.dlam = take(c.RT); c.dtt = take(c.RT);
  c.sig = take(3 * N); c.dsig = take(3 * N); c.rsig = take(3 * N); c.rds = take(3 * N);
  c.Weff = take(10 * N); c.Si = take(6 * N); c.bh = take(4 * N); c.ro = take(4 * N); c.o = take(4 * N);
  c.dob = take(4 * N); c.wo = take(4 * N);
  c.red = take(8);
  c.hist = (unsigned long long*)take(16);
  c.redi = (int*)take(4);
  c.act = (unsigned char*)take((c.RT + 7) / 8 + 1);
  if (gdoubles) *gdoubles = (long long)(q - gw);
  return (size_t)((char*)p - sm);
}

// x_{t+1} = A_t x_t + B_t u_t from x0 with the solved inputs: the full predicted trajectory (N+1, 9) into xs.
__device__ __forceinline__ void rollout(const Ctx& c, const double* x0, double* xs) {
  if (threadIdx.x == 0) {
    double v[9];
#pragma unroll
    for (int r = 0; r < 9; ++r) { v[r] = x0[r]; xs[r] = v[r]; }
    for (int t = 0; t < c.N; ++t) {
      lpv_step(c.Acol + 27 * t, v);
      const double* B = c.Bcol + 4 * t;
      const double d = c.u[2 * t], a = c.u[2 * t + 1];
      v[0] = fma(B[0], d, fma(B[3], a, v[0]));
      v[1] = fma(B[1], d, v[1]);
      v[2] = fma(B[2], d, v[2]);
#pragma unroll
      for (int r = 0; r < 9; ++r) xs[9 * (t + 1) + r] = v[r];
    }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(QP_THREADS) lpv_qp_solve_kernel(const __grid_constant__ DevParams P, const BatchPtrs io) {
  extern __shared__ __align__(16) char smem_raw[];
  const int b = blockIdx.x;
  if (b >= io.B) return;
  Ctx c;
  carve_workspace(c, smem_raw, io.gws ? io.gws + (long long)b * io.gws_stride : nullptr, P.N, P.nn, P.nrows,
                  P.gr_global, P.hk_global, nullptr);
  const int N = P.N, nn = P.nn;
  int cnt = io.nbr_count ? io.nbr_count[b] : nn;
  c.cnt = cnt < 0 ? 0 : (cnt > nn ? nn : cnt);
  const double* x0 = io.x0 + (size_t)b * NSTATE;
  const double* xprev = io.xprev + (size_t)b * io.xprev_rows * NSTATE;
  const double* uprev = io.uprev + (size_t)b * N * 2;
  const double* uold = io.uold + (size_t)b * 2;
  const double* nbr = io.nbr_xy + (size_t)b * (N + 1) * nn * 2;
  const double* pose = io.pose + (size_t)b * (N + 1) * 2;
  const int n = 16 * N + 12;

  // ---------------- problem build ----------------
  int bad_plane = setup_planes(c, P, nbr, pose, io.planes ? io.planes + (size_t)b * N * 3 * nn : nullptr);
  int bad_map = setup_linearise(c, P, xprev, io.xprev_rows, uprev);
  bad_plane = block_or(bad_plane, c.redi);
  bad_map = block_or(bad_map, c.redi);
  if (bad_plane || bad_map) {  // the reference raises here; report and leave NaNs
    const double qnan = __longlong_as_double(0x7ff8000000000000ULL);
    if (io.z) for (int j = threadIdx.x; j < n; j += QP_THREADS) io.z[(size_t)b * n + j] = qnan;
    if (io.y_ineq) for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) io.y_ineq[(size_t)b * c.RT + j] = qnan;
    if (io.y_eq) for (int j = threadIdx.x; j < 14 * N + 12; j += QP_THREADS) io.y_eq[(size_t)b * (14 * N + 12) + j] = qnan;
    if (threadIdx.x == 0) {
      if (io.status) io.status[b] = bad_map ? DMPC_MAP_LOOKUP_FAILED : DMPC_DEGENERATE_PLANE;
      if (io.obj) io.obj[b] = qnan;
      if (io.iters) io.iters[b] = 0;
      if (io.pivots) io.pivots[b] = 0;
    }
    return;
  }
  setup_prediction(c, P, x0);
  __syncthreads();
  setup_cost(c, P, uold);

  // ---------------- solve ----------------
  int n_ipm = 0, n_rounds = 0;
  const int status = solve_qp(c, P, &n_ipm, &n_rounds);

  // ---------------- outputs (LPV_Planner.py:162-180 layout of res.x) ----------------
  double* xs = c.Zb;  // (N+1) x 9 scratch
  rollout(c, x0, xs);
  if (io.z) {
    double* z = io.z + (size_t)b * n;
    for (int w = threadIdx.x; w < 12 * (N + 1); w += QP_THREADS) {
      const int t = w / 12, r = w % 12;
      z[w] = r < 9 ? xs[9 * t + r] : (t == 0 ? 0.0 : c.sig[3 * (t - 1) + (r - 9)]);
    }
    for (int j = threadIdx.x; j < 2 * N; j += QP_THREADS) {
      z[12 * (N + 1) + j] = c.u[j];
      z[12 * (N + 1) + 2 * N + j] = c.u[j] - (j < 2 ? uold[j] : c.u[j - 2]);
    }
  }
  if (io.y_ineq) for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) io.y_ineq[(size_t)b * c.RT + j] = c.lam[j];
  // objective 1/2 z'Pz + q'z (LPV_Planner.py:382-427)
  {
    double v = 0.0;
    for (int t = threadIdx.x; t <= N; t += QP_THREADS) {
      const double* x = xs + 9 * t;
      double s = 0.0;
      for (int e = 0; e < P.nq; ++e) {
        const double xa = x[P.rowsel[P.qa[e]]], xb = x[P.rowsel[P.qb[e]]];
        s = fma(P.qv[e] * (P.qa[e] == P.qb[e] ? 1.0 : 2.0), xa * xb, s);
      }
      s += 2.0 * (-P.vx_ref * P.Q00) * x[0];
      if (t >= 1) {
        s += 2.0 * (c.pxy[2 * (t - 1)] * x[7] + c.pxy[2 * (t - 1) + 1] * x[8]);
        const double* sg = c.sig + 3 * (t - 1);
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) s = fma(P.Qs[i * 3 + j] * sg[i], sg[j], s);
        const int k = t - 1;
        const double u0 = c.u[2 * k], u1 = c.u[2 * k + 1];
        const double d0 = u0 - (k == 0 ? uold[0] : c.u[2 * k - 2]), d1 = u1 - (k == 0 ? uold[1] : c.u[2 * k - 1]);
        s += P.R[0] * u0 * u0 + (P.R[1] + P.R[2]) * u0 * u1 + P.R[3] * u1 * u1;
        s += P.dR[0] * d0 * d0 + (P.dR[1] + P.dR[2]) * d0 * d1 + P.dR[3] * d1 * d1;
      }
      v += s;
    }
    v = block_sum(v, c.red);
    if (threadIdx.x == 0 && io.obj) io.obj[b] = v;
  }
  // multipliers of the eliminated equality rows (rows of G, LPV_Planner.py:429-475) by the adjoint recursion
  if (io.y_eq) {
    double* ye = io.y_eq + (size_t)b * (14 * N + 12);
    rows_to_stage(c, c.lam, c.wo, nullptr);
    __syncthreads();
    if (threadIdx.x == 0) {
      double nu[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
      for (int t = N; t >= 0; --t) {
        double nn_[9];
        if (t < N) {  // A_t' nu_{t+1}
          const double* A = c.Acol + 27 * t;
          for (int j = 0; j < 9; ++j) {
            double s = j >= 3 ? nu[j] : 0.0;
            if (j < 3) for (int r = 0; r < 9; ++r) s = fma(A[r * 3 + j], nu[r], s);
            nn_[j] = s;
          }
        } else for (int j = 0; j < 9; ++j) nn_[j] = 0.0;
        const double* x = xs + 9 * t;
        double gq[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        for (int e = 0; e < P.nq; ++e) {
          const int a = P.rowsel[P.qa[e]], bq = P.rowsel[P.qb[e]];
          gq[a] = fma(2.0 * P.qv[e], x[bq], gq[a]);
          if (a != bq) gq[bq] = fma(2.0 * P.qv[e], x[a], gq[bq]);
        }
        gq[0] += 2.0 * (-P.vx_ref * P.Q00);
        if (t >= 1) {
          gq[7] += 2.0 * c.pxy[2 * (t - 1)]; gq[8] += 2.0 * c.pxy[2 * (t - 1) + 1];
          const double* w4 = c.wo + 4 * (t - 1);
          gq[0] += w4[0]; gq[3] += w4[1]; gq[7] += w4[2]; gq[8] += w4[3];
        }
        for (int j = 0; j < 9; ++j) { nu[j] = nn_[j] - gq[j]; ye[12 * t + j] = nu[j]; }
        ye[12 * t + 9] = 0.0; ye[12 * t + 10] = 0.0; ye[12 * t + 11] = 0.0;
      }
      for (int k = 0; k < N; ++k) {
        const double d0 = c.u[2 * k] - (k == 0 ? uold[0] : c.u[2 * k - 2]);
        const double d1 = c.u[2 * k + 1] - (k == 0 ? uold[1] : c.u[2 * k - 1]);
        const double m0 = 2.0 * (P.dR[0] * d0 + P.dR[1] * d1), m1 = 2.0 * (P.dR[2] * d0 + P.dR[3] * d1);
        ye[12 * (N + 1) + 2 * k] = k == 0 ? m0 : -m0;
        ye[12 * (N + 1) + 2 * k + 1] = k == 0 ? m1 : -m1;
      }
    }
  }
  if (threadIdx.x == 0) {
    if (io.status) io.status[b] = status;
    if (io.iters) io.iters[b] = n_ipm;
    if (io.pivots) io.pivots[b] = n_rounds;
  }
}

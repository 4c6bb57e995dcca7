// qp2_kernel.cuh - the per-agent LPV-MPC solve kernel on the fast engine: one CTA (4 warps) per QP, the whole
// workspace packed in shared memory, everything between the inputs of PlannerLPV.solve and its outputs
// (distributedPlanner/LPV_Planner.py:115-182) on chip.
#pragma once
#include "qp_setup.cuh"
#include "qp2_schur.cuh"

__device__ __forceinline__ double pick9(const double* v, int idx) {  // v[idx] without spilling v to local memory
  double r = v[0];
#pragma unroll
  for (int a = 1; a < 9; ++a) r = (idx == a) ? v[a] : r;
  return r;
}

// Prediction rows: x_t = cr_t + sum_j Gr_t[:, j] u_j for the kept state rows (one thread per input column,
// plus one thread for the free response from x0).  Rows 0-3 (vx, e_y, x, y) go to the packed Gb, cost-only rows to Gx.
__device__ __forceinline__ void setup_prediction2(const Ctx2& c, const DevParams& P, const double* x0) {
#pragma unroll 1
  for (int j = threadIdx.x; j <= c.nu; j += QP_THREADS) {
    double v[9];
    if (j == c.nu) {
#pragma unroll
      for (int r = 0; r < 9; ++r) v[r] = x0[r];
      for (int t = 0; t < c.N; ++t) {
        lpv_step(c.Acol + 27 * t, v);
        double* ct = c.cr + t * c.NR;
        ct[0] = v[0]; ct[1] = v[3]; ct[2] = v[7]; ct[3] = v[8];
        for (int r = 4; r < c.NR; ++r) ct[r] = pick9(v, P.rowsel[r]);
      }
    } else {
      const int k = j >> 1;
#pragma unroll
      for (int r = 0; r < 9; ++r) v[r] = 0.0;
      const double* B = c.Bcol + 4 * k;
      if (j & 1) v[0] = B[3]; else { v[0] = B[0]; v[1] = B[1]; v[2] = B[2]; }
      for (int t = k; t < c.N; ++t) {
        if (t > k) lpv_step(c.Acol + 27 * t, v);
        GB(c, t, 0, j) = v[0]; GB(c, t, 1, j) = v[3]; GB(c, t, 2, j) = v[7]; GB(c, t, 3, j) = v[8];
        for (int r = 4; r < c.NR; ++r) GX(c, t, r, j) = pick9(v, P.rowsel[r]);
      }
    }
  }
}

__device__ __forceinline__ double gr2(const Ctx2& c, int t, int r, int j) { return r < 4 ? GB(c, t, r, j) : GX(c, t, r, j); }

// H = 2 (Gamma' Qbar Gamma + Rbar + D' dRbar D),  g = 2 Gamma' (Qbar c + p) - 2 D' dRbar d0   (condensed cost).
// `ql` is an N x NR scratch (c.om .. c.dlam).  Ends with a barrier.
__device__ __forceinline__ void setup_cost2(const Ctx2& c, const DevParams& P, const double* uold) {
  double* ql = c.om;
#pragma unroll 1
  for (int t = threadIdx.x; t < c.N; t += QP_THREADS) {
    double* q = ql + t * c.NR;
    for (int r = 0; r < c.NR; ++r) q[r] = 0.0;
    const double* ct = c.cr + t * c.NR;
    for (int e = 0; e < P.nq; ++e) {
      const int a = P.qa[e], b = P.qb[e];
      q[a] = fma(P.qv[e], ct[b], q[a]);
      if (a != b) q[b] = fma(P.qv[e], ct[a], q[b]);
    }
    q[0] += -P.vx_ref * P.Q00;      // LPV_Planner.py:412
    q[2] += c.pxy[2 * t];           // :421
    q[3] += c.pxy[2 * t + 1];       // :422
  }
  __syncthreads();
#pragma unroll 1
  for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) {
    double s = 0.0;
    for (int t = j >> 1; t < c.N; ++t)
      for (int r = 0; r < c.NR; ++r) s = fma(gr2(c, t, r, j), ql[t * c.NR + r], s);
    s *= 2.0;
    if (j < 2) s -= 2.0 * (P.dR[j * 2 + 0] * uold[0] + P.dR[j * 2 + 1] * uold[1]);
    c.g[j] = s;
  }
#pragma unroll 1
  for (int e = threadIdx.x; e < c.ntri; e += QP_THREADS) {
    const int ij = c.tri[e], i = ij >> 8, j = ij & 255;
    double s = 0.0;
    for (int q = 0; q < P.nq; ++q) {
      const int a = P.qa[q], b = P.qb[q];
      double sq = 0.0;
      if (a == b) {
        for (int t = i >> 1; t < c.N; ++t) sq = fma(gr2(c, t, a, i), gr2(c, t, a, j), sq);
      } else {
        for (int t = i >> 1; t < c.N; ++t) sq += gr2(c, t, a, i) * gr2(c, t, b, j) + gr2(c, t, b, i) * gr2(c, t, a, j);
      }
      s = fma(P.qv[q], sq, s);
    }
    s *= 2.0;
    const int ki = i >> 1, kj = j >> 1, ai = i & 1, aj = j & 1;
    if (ki == kj) {
      s += 2.0 * (P.R[ai * 2 + aj] + P.dR[ai * 2 + aj]);
      if (ki < c.N - 1) s += 2.0 * P.dR[ai * 2 + aj];
    } else if (ki == kj + 1) {
      s -= 2.0 * P.dR[ai * 2 + aj];
    }
    c.H[e] = s;
  }
  __syncthreads();
}

// x_{t+1} = A_t x_t + B_t u_t from x0 with the solved inputs: the full predicted trajectory (N+1, 9) into xs.
__device__ __forceinline__ void rollout2(const Ctx2& c, const double* x0, double* xs) {
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

__global__ void __launch_bounds__(QP_THREADS, 4) lpv_qp_fast_kernel(const __grid_constant__ DevParams P, const BatchPtrs io) {
  if (io.part == 1) asm volatile("griddepcontrol.launch_dependents;");  // split launch: the bulk may be scheduled now
  if (blockIdx.x >= io.B) return;
  if (io.part) {
    const int nh = *io.nhard;  // <= number of SMs = part 1's grid (order_by_cost_kernel clamps it)
    if (io.part == 1 ? (int)blockIdx.x >= nh : (int)blockIdx.x < nh) {
      // part 2 shares no data with part 1, but whatever follows in the stream needs both: one idle CTA of part 2 stays
      // until part 1 has completed and flushed, so that "part 2 complete" implies "part 1 complete"
      if (io.part == 2 && blockIdx.x == 0) asm volatile("griddepcontrol.wait;" ::: "memory");
      return;
    }
  }
  const int b = io.order ? io.order[blockIdx.x] : blockIdx.x;
  const int N = P.N, nn = P.nn;
  if (threadIdx.x == 0) {
    Hdr2* hd = (Hdr2*)smem_fast;
    int cnt = io.nbr_count ? io.nbr_count[b] : nn;
    hd->N = N; hd->nn = nn; hd->NR = P.nrows; hd->cnt = cnt < 0 ? 0 : (cnt > nn ? nn : cnt);
    hd->refine = P.refine; hd->pdas_first = P.pdas_first; hd->pdas_after = P.pdas_after; hd->ipm_max = P.ipm_max;
    hd->n_rounds = 0; hd->n_ipm = 0;
    { Ctx2 c0; carve_workspace2(c0, nullptr, N, nn, P.nrows); hd->WS = c0.WS; }
    hd->delta = P.delta; hd->cert_tol = P.cert_tol; hd->ipm_mu0 = P.ipm_mu0;
    for (int i = 0; i < 9; ++i) hd->Qs[i] = P.Qs[i];
    {  // (2 Qs)^-1 by cofactors (Qs symmetric positive definite, checked in dmpc_create)
      const double a = 2 * P.Qs[0], b = 2 * P.Qs[1], cc = 2 * P.Qs[2], d = 2 * P.Qs[4], e = 2 * P.Qs[5], f = 2 * P.Qs[8];
      const double c00 = d * f - e * e, c01 = cc * e - b * f, c02 = b * e - cc * d;
      const double c11 = a * f - cc * cc, c12 = b * cc - a * e, c22 = a * d - b * b;
      const double idet = 1.0 / (a * c00 + b * c01 + cc * c02);
      hd->Qsi[0] = c00 * idet; hd->Qsi[1] = c01 * idet; hd->Qsi[2] = c02 * idet;
      hd->Qsi[3] = c01 * idet; hd->Qsi[4] = c11 * idet; hd->Qsi[5] = c12 * idet;
      hd->Qsi[6] = c02 * idet; hd->Qsi[7] = c12 * idet; hd->Qsi[8] = c22 * idet;
    }
  }
  __syncthreads();
  CTX2_LOCAL(c);
  const double* x0 = io.x0 + (size_t)b * NSTATE;
  const double* xprev = io.xprev + (size_t)b * io.xprev_rows * NSTATE;
  const double* uprev = io.uprev + (size_t)b * N * 2;
  const double* uold = io.uold + (size_t)b * 2;
  const double* nbr = io.nbr_xy + (size_t)b * (N + 1) * nn * 2;
  const double* pose = io.pose + (size_t)b * (N + 1) * 2;
  const int n = 16 * N + 12;

  // ---------------- problem build ----------------
  build_tables(c);
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
      if (io.cost) io.cost[b] = 0;
    }
    return;
  }
#pragma unroll 1
  for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) c.bv[j] = row_valid2(c, j) ? row_h2_build(c, P, j) : 0.0;
  setup_prediction2(c, P, x0);
  __syncthreads();
  setup_cost2(c, P, uold);

  // ---------------- solve ----------------
  const int status = solve_qp2(c);

  // ---------------- outputs (LPV_Planner.py:162-180 layout of res.x) ----------------
  setup_linearise(c, P, xprev, io.xprev_rows, uprev);  // A_k, B_k again: their block was reused by the solver
  __syncthreads();
  double* xs = c.K;  // (N+1) x 9 scratch
  rollout2(c, x0, xs);
  if (io.z) {
    double* z = io.z + (size_t)b * n;
#pragma unroll 1
    for (int w = threadIdx.x; w < 12 * (N + 1); w += QP_THREADS) {
      const int t = w / 12, r = w % 12;
      z[w] = r < 9 ? xs[9 * t + r] : (t == 0 ? 0.0 : c.sig[3 * (t - 1) + (r - 9)]);
    }
#pragma unroll 1
    for (int j = threadIdx.x; j < 2 * N; j += QP_THREADS) {
      z[12 * (N + 1) + j] = c.u[j];
      z[12 * (N + 1) + 2 * N + j] = c.u[j] - (j < 2 ? uold[j] : c.u[j - 2]);
    }
  }
  if (io.y_ineq) for (int j = threadIdx.x; j < c.RT; j += QP_THREADS) io.y_ineq[(size_t)b * c.RT + j] = c.lam[j];
  // objective 1/2 z'Pz + q'z (LPV_Planner.py:382-427)
  {
    double v = 0.0;
#pragma unroll 1
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
    rows_to_stage2(c, c.lam, c.wo, nullptr);
    __syncthreads();
    if (threadIdx.x == 0) {
      double nu[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
      for (int t = N; t >= 0; --t) {
        double nn_[9];
        if (t < N) {  // A_t' nu_{t+1}
          const double* A = c.Acol + 27 * t;
#pragma unroll
          for (int j = 0; j < 9; ++j) {
            double s = j >= 3 ? nu[j] : 0.0;
            if (j < 3) {
#pragma unroll
              for (int r = 0; r < 9; ++r) s = fma(A[r * 3 + j], nu[r], s);
            }
            nn_[j] = s;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 9; ++j) nn_[j] = 0.0;
        }
        const double* x = xs + 9 * t;
        double* gq = c.yt;  // 9-entry scratch in shared memory (dynamic indexing)
        for (int j = 0; j < 9; ++j) gq[j] = 0.0;
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
#pragma unroll
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
    if (io.iters) io.iters[b] = c.hd->n_ipm;
    if (io.pivots) io.pivots[b] = c.hd->n_rounds;
    if (io.cost) io.cost[b] = 24 * c.hd->n_ipm + c.hd->n_rounds;  // an interior-point iteration ~ 24 Schur rounds
  }
}

// ---------------------------------------------------------------------------------------------------
// order[] for the next call from this call's cost[]: a counting sort into 8 cost classes, most expensive class first,
// QP index ascending inside a class (deterministic).  One CTA; each thread owns a contiguous chunk of the batch.
// Only the launch order depends on it - every QP is solved from scratch on every call whatever the order.
#define ORDER_THREADS 1024
#define ORDER_CLASSES 8
__device__ __forceinline__ int cost_class(int cst) {
  if (cst & (1 << 20)) {  // stage-wise engine: 3 factorisations + solves; one active-set round ~ 5, an interior-point solve ~ 80
    cst &= (1 << 20) - 1;  // (bit 21: interior point used)
    return cst >= 110 ? 7 : cst >= 80 ? 6 : cst >= 50 ? 5 : cst >= 30 ? 4 : cst >= 20 ? 3 : cst >= 12 ? 2 : cst >= 7 ? 1 : 0;
  }
  return cst >= 24 * 16 ? 7 : cst >= 24 ? 6 : cst >= 12 ? 5 : cst >= 6 ? 4 : cst >= 4 ? 3 : cst >= 3 ? 2 : cst >= 2 ? 1 : 0;
}

__global__ void __launch_bounds__(ORDER_THREADS, 1) order_by_cost_kernel(const int* __restrict__ cost, int* __restrict__ order,
                                                                         int* __restrict__ nhard, int B, int nsm) {
  // One pass: every thread counts its chunk into the eight classes, ONE block-wide scan of the eight counters (two
  // barriers instead of four per class), then the scatter.  Same order as a class-by-class counting sort.
  __shared__ int wsum[ORDER_THREADS / 32][ORDER_CLASSES];  // inclusive per-warp totals, then their scan over the warps
  const int chunk = (B + ORDER_THREADS - 1) / ORDER_THREADS;
  const int lo = min(B, (int)threadIdx.x * chunk), hi = min(B, lo + chunk);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int mine[ORDER_CLASSES], incl[ORDER_CLASSES];
#pragma unroll
  for (int c = 0; c < ORDER_CLASSES; ++c) mine[c] = 0;
  for (int i = lo; i < hi; ++i) {
    const int cl = cost_class(cost[i]);
#pragma unroll
    for (int c = 0; c < ORDER_CLASSES; ++c) mine[c] += (cl == c);
  }
#pragma unroll
  for (int c = 0; c < ORDER_CLASSES; ++c) {
    int v = mine[c];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int w = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += w;
    }
    incl[c] = v;
    if (lane == 31) wsum[warp][c] = v;
  }
  __syncthreads();
  if (warp < ORDER_CLASSES) {  // warp c scans class c over the 32 warps
    int v = wsum[lane][warp];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int w = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += w;
    }
    wsum[lane][warp] = v;
  }
  __syncthreads();
  // most expensive class first: where class c starts = the totals of the classes above it
  int pos[ORDER_CLASSES];
  int base = 0;
#pragma unroll
  for (int c = ORDER_CLASSES - 1; c >= 0; --c) {
    pos[c] = base + incl[c] - mine[c] + (warp ? wsum[warp - 1][c] : 0);
    base += wsum[ORDER_THREADS / 32 - 1][c];
    if (c == 6 && threadIdx.x == 0) *nhard = min(base, nsm);  // classes 7 and 6: at least one interior-point iteration
  }
  for (int i = lo; i < hi; ++i) {
    const int cl = cost_class(cost[i]);
    int p = 0;
#pragma unroll
    for (int c = 0; c < ORDER_CLASSES; ++c)
      if (cl == c) { p = pos[c]; pos[c] = p + 1; }
    order[p] = i;
  }
}

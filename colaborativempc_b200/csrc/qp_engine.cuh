// qp_engine.cuh - one-CTA-per-QP engine: weighted normal equations of the condensed LPV-MPC QP in shared memory.
//
// Condensed problem (decision v = (u in R^{2N}, sigma in R^{3N})), obtained from the QP that
// PlannerLPV.solve assembles (distributedPlanner/LPV_Planner.py:279-475) by eliminating the dynamics and the
// du-definition equalities:
//     min 1/2 u'Hu + g'u + sum_t sigma_t' Qs sigma_t     s.t.  rows of F (LPV_Planner.py:279-380) <= b.
// Every linear solve the active-set polish and the interior-point phase need has the form
//     (blkdiag(H, 2Qs) + A' D A) dv = r,  D = diag(row weights) >= 0,
// and is reduced on chip to a dense 2N x 2N SPD system  K du = rhs  by eliminating the 3 slacks of each stage
// (3x3 blocks) and by writing the stage rows through the four predicted outputs (vx, e_y, x, y) of that stage:
//     K = H + sum_t Gb_t' Weff_t Gb_t + diag(box weights),   Gb_t = rows {vx,e_y,x,y} of the prediction matrix.
#pragma once
#include "qp_common.cuh"

struct Ctx {
  int N, nn, cnt, nu, NR, ldg, ldk, rst, Rs, RT;  // rst: rows per stage (4+nn); Rs = N*rst; RT = Rs + 4N
  // linearisation / problem data
  double *Acol, *Bcol, *eyb, *pxy, *ax, *ay, *hh;
  double *Gr, *cr, *Zb, *H, *K, *g;
  // iterates and work vectors
  double *u, *rhs, *du, *tv, *rdu;
  double *lam, *tt, *D, *res, *om, *dlam, *dtt;
  unsigned char* act;
  double *sig, *dsig, *rsig, *rds;
  double *Weff, *Si, *bh, *ro, *o, *dob, *wo;
  double* red;
  int* redi;
  unsigned long long* hist;
};

#define GR(c, t, r, j) (c).Gr[((t) * (c).NR + (r)) * (c).ldg + (j)]
#define ZB(c, t, r, j) (c).Zb[((t) * 4 + (r)) * (c).ldg + (j)]

__device__ __forceinline__ bool row_valid(const Ctx& c, int j) {
  if (j >= c.Rs) return j < c.RT;
  return (j % c.rst) < 4 + c.cnt;
}

// right-hand side b of row j (LPV_Planner.py:304-369 after the row re-ordering :350-362)
__device__ __forceinline__ double row_h(const Ctx& c, const DevParams& P, int j) {
  if (j >= c.Rs) {
    const int r = (j - c.Rs) & 3;
    return r == 0 ? P.max_rs : (r == 1 ? P.max_ls : (r == 2 ? P.max_ac : P.max_dc));
  }
  const int t = j / c.rst, r = j % c.rst;
  if (r == 0) return -P.min_vel;
  if (r == 1) return P.max_vel;
  if (r < 4) return c.eyb[t];
  return c.hh[t * c.nn + (r - 4)];
}

// value of row j of A v for stage outputs `o` (N x 4), slacks `sg` (N x 3) and inputs `uv`
__device__ __forceinline__ double row_val(const Ctx& c, int j, const double* o, const double* sg, const double* uv) {
  if (j >= c.Rs) {
    const int k = (j - c.Rs) >> 2, r = (j - c.Rs) & 3;
    const double v = uv[2 * k + (r >> 1)];
    return (r & 1) ? -v : v;
  }
  const int t = j / c.rst, r = j % c.rst;
  const double* ot = o + 4 * t;
  const double* st = sg + 3 * t;
  if (r == 0) return -ot[0];
  if (r == 1) return ot[0] + st[0];
  if (r == 2) return ot[1] + st[1];
  if (r == 3) return -ot[1] + st[1];
  const int i = r - 4;
  return c.ax[t * c.nn + i] * ot[2] + c.ay[t * c.nn + i] * ot[3] - st[2];
}

// o[t][q] = Gb_t[q,:] . uv (+ offset): the four constrained outputs of every stage.   (syncs: caller)
__device__ __forceinline__ void stage_outputs(const Ctx& c, const double* uv, double* o, bool with_offset) {
  for (int w = threadIdx.x; w < 4 * c.N; w += QP_THREADS) {
    const int t = w >> 2, q = w & 3;
    const double* g = &GR(c, t, q, 0);
    double s = with_offset ? c.cr[t * c.NR + q] : 0.0;
    const int jn = min(c.nu, 2 * (t + 1));
    for (int j = 0; j < jn; ++j) s = fma(g[j], uv[j], s);
    o[w] = s;
  }
}

// A' w split into the stage-output part wo (N x 4), the slack part gs (N x 3); box part is added by apply_Gt.
__device__ __forceinline__ void rows_to_stage(const Ctx& c, const double* w, double* wo, double* gs) {
  for (int t = threadIdx.x; t < c.N; t += QP_THREADS) {
    const double* wr = w + t * c.rst;
    double sx = 0.0, sy = 0.0, sw = 0.0;
    for (int i = 0; i < c.cnt; ++i) {
      sx = fma(c.ax[t * c.nn + i], wr[4 + i], sx);
      sy = fma(c.ay[t * c.nn + i], wr[4 + i], sy);
      sw += wr[4 + i];
    }
    wo[4 * t + 0] = wr[1] - wr[0];
    wo[4 * t + 1] = wr[2] - wr[3];
    wo[4 * t + 2] = sx;
    wo[4 * t + 3] = sy;
    if (gs) { gs[3 * t + 0] = wr[1]; gs[3 * t + 1] = wr[2] + wr[3]; gs[3 * t + 2] = -sw; }
  }
}

// out[j] = base[j]*bs + sum_t Gb_t[:,j]' wo_t * ws + box(wbox)*ws
__device__ __forceinline__ void apply_Gt(const Ctx& c, const double* wo, const double* wbox, double ws,
                                         const double* base, double bs, double* out) {
  for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) {
    double s = 0.0;
    for (int t = j >> 1; t < c.N; ++t) {
      const double* w4 = wo + 4 * t;
      s = fma(GR(c, t, 0, j), w4[0], s);
      s = fma(GR(c, t, 1, j), w4[1], s);
      s = fma(GR(c, t, 2, j), w4[2], s);
      s = fma(GR(c, t, 3, j), w4[3], s);
    }
    if (wbox) {
      const double* wb = wbox + 4 * (j >> 1) + 2 * (j & 1);
      s += wb[0] - wb[1];
    }
    out[j] = (base ? bs * base[j] : 0.0) + ws * s;
  }
}

// out = H uv (H symmetric, full storage)
__device__ __forceinline__ void apply_H(const Ctx& c, const double* uv, double* out) {
  for (int i = threadIdx.x; i < c.nu; i += QP_THREADS) {
    double s = 0.0;
    for (int j = 0; j < c.nu; ++j) s = fma(c.H[j * c.ldk + i], uv[j], s);
    out[i] = s;
  }
}

// Per-stage slack elimination for row weights D: Si = (2Qs + A_s'DA_s)^-1 (3x3), bh, Weff (4x4 symmetric).
__device__ __forceinline__ void stage_weights(const Ctx& c, const DevParams& P) {
  for (int t = threadIdx.x; t < c.N; t += QP_THREADS) {
    const double* d = c.D + t * c.rst;
    double sxx = 0, sxy = 0, syy = 0, sx = 0, sy = 0, sd = 0;
    for (int i = 0; i < c.cnt; ++i) {
      const double di = d[4 + i], x = c.ax[t * c.nn + i], y = c.ay[t * c.nn + i];
      sxx = fma(di * x, x, sxx); sxy = fma(di * x, y, sxy); syy = fma(di * y, y, syy);
      sx = fma(di, x, sx); sy = fma(di, y, sy); sd += di;
    }
    const double b0 = d[1], b1 = d[2] - d[3], b2 = -sx, b3 = -sy;
    // S = 2 Qs + diag(d1, d2+d3, sum d4)
    const double s00 = 2 * P.Qs[0] + d[1], s01 = 2 * P.Qs[1], s02 = 2 * P.Qs[2];
    const double s11 = 2 * P.Qs[4] + d[2] + d[3], s12 = 2 * P.Qs[5], s22 = 2 * P.Qs[8] + sd;
    const double c00 = s11 * s22 - s12 * s12, c01 = s02 * s12 - s01 * s22, c02 = s01 * s12 - s02 * s11;
    const double c11 = s00 * s22 - s02 * s02, c12 = s01 * s02 - s00 * s12, c22 = s00 * s11 - s01 * s01;
    const double idet = 1.0 / (s00 * c00 + s01 * c01 + s02 * c02);
    double* si = c.Si + 6 * t;  // 00 01 02 11 12 22
    si[0] = c00 * idet; si[1] = c01 * idet; si[2] = c02 * idet; si[3] = c11 * idet; si[4] = c12 * idet; si[5] = c22 * idet;
    double* bh = c.bh + 4 * t;
    bh[0] = b0; bh[1] = b1; bh[2] = b2; bh[3] = b3;
    // Weff = W - Bh Si Bh'; rows 0,1,2,3 of Bh use slack columns 0,1,2,2
    double* w = c.Weff + 10 * t;  // 00 01 02 03 11 12 13 22 23 33
    w[0] = d[0] + d[1] - b0 * b0 * si[0];
    w[1] = -b0 * b1 * si[1];
    w[2] = -b0 * b2 * si[2];
    w[3] = -b0 * b3 * si[2];
    w[4] = d[2] + d[3] - b1 * b1 * si[3];
    w[5] = -b1 * b2 * si[4];
    w[6] = -b1 * b3 * si[4];
    w[7] = sxx - b2 * b2 * si[5];
    w[8] = sxy - b2 * b3 * si[5];
    w[9] = syy - b3 * b3 * si[5];
  }
}

// ro_t = -Bh Si rsig_t (stage-output space), used to fold the slack right-hand side into the u system
__device__ __forceinline__ void slack_rhs_to_outputs(const Ctx& c, const double* rsig, double* ro) {
  for (int t = threadIdx.x; t < c.N; t += QP_THREADS) {
    const double* si = c.Si + 6 * t;
    const double* bh = c.bh + 4 * t;
    const double* r = rsig + 3 * t;
    const double v0 = si[0] * r[0] + si[1] * r[1] + si[2] * r[2];
    const double v1 = si[1] * r[0] + si[3] * r[1] + si[4] * r[2];
    const double v2 = si[2] * r[0] + si[4] * r[1] + si[5] * r[2];
    ro[4 * t + 0] = -bh[0] * v0; ro[4 * t + 1] = -bh[1] * v1; ro[4 * t + 2] = -bh[2] * v2; ro[4 * t + 3] = -bh[3] * v2;
  }
}

// dsig_t = Si (rsig_t - Bh' dob_t)
__device__ __forceinline__ void slack_backsolve(const Ctx& c, const double* rsig, const double* dob, double* dsig) {
  for (int t = threadIdx.x; t < c.N; t += QP_THREADS) {
    const double* si = c.Si + 6 * t;
    const double* bh = c.bh + 4 * t;
    const double* d = dob + 4 * t;
    const double r0 = rsig[3 * t + 0] - bh[0] * d[0];
    const double r1 = rsig[3 * t + 1] - bh[1] * d[1];
    const double r2 = rsig[3 * t + 2] - bh[2] * d[2] - bh[3] * d[3];
    dsig[3 * t + 0] = si[0] * r0 + si[1] * r1 + si[2] * r2;
    dsig[3 * t + 1] = si[1] * r0 + si[3] * r1 + si[4] * r2;
    dsig[3 * t + 2] = si[2] * r0 + si[4] * r1 + si[5] * r2;
  }
}

// K = H + sum_t Gb_t' Weff_t Gb_t + diag(box weights); lower triangle (i >= j) written, column-major K[j*ldk+i].
__device__ __forceinline__ void build_K(const Ctx& c) {
  // Z_t = Weff_t Gb_t
  for (int w = threadIdx.x; w < c.N * c.ldg; w += QP_THREADS) {
    const int t = w / c.ldg, j = w % c.ldg;
    if (j >= c.nu) continue;
    const double* we = c.Weff + 10 * t;
    const double g0 = GR(c, t, 0, j), g1 = GR(c, t, 1, j), g2 = GR(c, t, 2, j), g3 = GR(c, t, 3, j);
    ZB(c, t, 0, j) = we[0] * g0 + we[1] * g1 + we[2] * g2 + we[3] * g3;
    ZB(c, t, 1, j) = we[1] * g0 + we[4] * g1 + we[5] * g2 + we[6] * g3;
    ZB(c, t, 2, j) = we[2] * g0 + we[5] * g1 + we[7] * g2 + we[8] * g3;
    ZB(c, t, 3, j) = we[3] * g0 + we[6] * g1 + we[8] * g2 + we[9] * g3;
  }
  __syncthreads();
  const int ntri = c.nu * (c.nu + 1) / 2;
  for (int e = threadIdx.x; e < ntri; e += QP_THREADS) {
    // e -> (i >= j): row-major enumeration of the lower triangle
    int i = (int)((sqrt(8.0 * e + 1.0) - 1.0) * 0.5);
    while ((i + 1) * (i + 2) / 2 <= e) ++i;
    while (i * (i + 1) / 2 > e) --i;
    const int j = e - i * (i + 1) / 2;
    double s = c.H[j * c.ldk + i];
    for (int t = i >> 1; t < c.N; ++t) {
      s = fma(GR(c, t, 0, i), ZB(c, t, 0, j), s);
      s = fma(GR(c, t, 1, i), ZB(c, t, 1, j), s);
      s = fma(GR(c, t, 2, i), ZB(c, t, 2, j), s);
      s = fma(GR(c, t, 3, i), ZB(c, t, 3, j), s);
    }
    if (i == j) {
      const double* db = c.D + c.Rs + 4 * (i >> 1) + 2 * (i & 1);
      s += db[0] + db[1];
    }
    c.K[j * c.ldk + i] = s;
  }
  __syncthreads();
}

// In-place Cholesky of the lower triangle of K (column-major). Non-positive pivots are floored (returns count).
__device__ __forceinline__ int cholesky(const Ctx& c) {
  int bad = 0;
  const int n = c.nu;
  for (int k = 0; k < n; ++k) {
    double* col = c.K + k * c.ldk;
    __syncthreads();
    double d = col[k];
    if (!(d > 1e-300)) { d = 1e-30; bad = 1; }
    const double piv = sqrt(d), ipiv = 1.0 / piv;
    __syncthreads();
    for (int i = k + threadIdx.x; i < n; i += QP_THREADS) col[i] = (i == k) ? piv : col[i] * ipiv;
    __syncthreads();
    const int rem = n - k - 1;
    // trailing update: K[i][j] -= L[i][k] L[j][k] for k < j <= i
    for (int e = threadIdx.x; e < rem * rem; e += QP_THREADS) {
      const int jj = e / rem, ii = e % rem;
      if (ii < jj) continue;
      const int i = k + 1 + ii, j = k + 1 + jj;
      c.K[j * c.ldk + i] = fma(-col[i], col[j], c.K[j * c.ldk + i]);
    }
  }
  __syncthreads();
  return bad;
}

// Solve L L' x = rhs in place (x in rhs) by warp 0; other warps wait at the barrier.
__device__ __forceinline__ void chol_solve(const Ctx& c, double* x) {
  const int n = c.nu;
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    for (int k = 0; k < n; ++k) {
      const double* col = c.K + k * c.ldk;
      if (lane == 0) x[k] = x[k] / col[k];
      __syncwarp();
      const double xk = x[k];
      for (int i = k + 1 + lane; i < n; i += 32) x[i] = fma(-col[i], xk, x[i]);
      __syncwarp();
    }
    for (int k = n - 1; k >= 0; --k) {
      const double* col = c.K + k * c.ldk;
      double s = 0.0;
      for (int i = k + 1 + lane; i < n; i += 32) s = fma(col[i], x[i], s);
      s = warp_sum(s);
      if (lane == 0) x[k] = (x[k] - s) / col[k];
      __syncwarp();
    }
  }
  __syncthreads();
}

// Direction for (ru, rsig): du, dsig, dob (= Gb du).  K must be factorised for the current D.
__device__ __forceinline__ void solve_direction(const Ctx& c, const double* ru, const double* rsig) {
  slack_rhs_to_outputs(c, rsig, c.ro);
  __syncthreads();
  apply_Gt(c, c.ro, nullptr, 1.0, ru, 1.0, c.du);
  __syncthreads();
  chol_solve(c, c.du);
  stage_outputs(c, c.du, c.dob, false);
  __syncthreads();
  slack_backsolve(c, rsig, c.dob, c.dsig);
  __syncthreads();
}

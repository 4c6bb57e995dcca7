// qp3_engine.cuh - STAGE-WISE engine: the LPV-MPC QP of PlannerLPV.solve solved in its UNCONDENSED form, the way the
// reference hands it to OSQP (distributedPlanner/LPV_Planner.py:222-239: states, slacks, inputs and input increments are
// all variables, the dynamics are equality rows), with every linear solve done by a Riccati recursion over the stages.
//
// Why: condensing (qp2_*.cuh) multiplies the A_k along the horizon; for crawling vehicles the forward-Euler lateral
// dynamics of _EstimateABC (LPV_Planner.py:505-517) are unstable (|1 - dt (Cf+Cr)/(m vx)| up to 7), the products reach
// 1e11+ and the condensed Hessian cannot be factorised in FP64.  The stage-wise factorisation never forms a product of
// two A_k: it carries the cost-to-go  V_k(xi) = 1/2 xi' P_k xi + p_k' xi  backwards (closed loop, bounded) and the state
// forwards under the optimal feedback.  Memory and work are O(N): ~190 doubles per stage, so N = 125 (the reference's
// shipped horizon, config_files/config_LPV.py:24) still fits in one SM's shared memory.
//
// One WARP per QP.  The recursion is sequential over the stages; inside a stage the 11 x 11 cost-to-go update is spread
// over the 32 lanes (three phases, __syncwarp between them).  Everything that has no stage-to-stage dependence (rows,
// slack elimination, gradients, step lengths) runs with one lane per stage or per row.
//
// Variables, per stage k = 0 .. N-1:  u_k (2),  x_{k+1} (9),  sigma_{k+1} (3).  Augmented state xi_k = (x_k, v_k = u_{k-1})
// carries the input-increment cost (u_k - u_{k-1})' dR (u_k - u_{k-1}) of LPV_Planner.py:404-409 without du variables.
// Rows (LPV_Planner.py:304-369), stage k on (x_{k+1}, sigma_{k+1}):  r0 -vx <= -min_vel,  r1 vx + s0 <= max_vel,
// r2 e_y + s1 <= ey,  r3 -e_y + s1 <= ey,  r4+i  a_i . (x,y) - s2 <= -min_dist/2 - b_i;   box rows on u_k (4).
//
// The same source compiles for the host (QP3_HOST: one "lane", no barriers) so that the algorithm can be checked in the
// CPU-only test tier (tests/hostsim, test infrastructure); the product only ever runs the sm_100a kernel.
#pragma once
#include <math.h>

#include "qp_types.h"

#ifdef QP3_HOST
struct Q3_V2 { double x, y; };
#define Q3HD inline
#define Q3D static inline
#define Q3N static
#define Q3_NL 1
#define Q3_LANE 0
#define Q3_SYNC() ((void)0)
static inline double q3_wsum(double v) { return v; }
static inline double q3_wmax(double v) { return v; }
static inline double q3_wmin(double v) { return v; }
static inline int q3_wor(int v) { return v; }
static inline unsigned long long q3_wsum64(unsigned long long v) { return v; }
static inline double q3_rcp(double d) { return 1.0 / d; }
static thread_local char* q3_host_smem = nullptr;
#define Q3_SMEM q3_host_smem
#else
#include <cuda_runtime.h>
typedef double2 Q3_V2;
#define Q3HD __host__ __device__ inline
#define Q3D __device__ __forceinline__
#define Q3N __device__ __noinline__
#define Q3_NL 32
#define Q3_LANE ((int)(threadIdx.x & 31))
#define Q3_SYNC() __syncwarp()
__device__ __forceinline__ double q3_wsum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double q3_wmax(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double q3_wmin(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ int q3_wor(int v) { return __any_sync(0xffffffffu, v); }
__device__ __forceinline__ unsigned long long q3_wsum64(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// 1 / d without the slow-path branch of the division: the hardware's double-precision reciprocal seed (MUFU.RCP64H,
// about 20 bits, full exponent range) and three Newton steps, the last in residual form: relative error about one ulp.
// Zero, infinite and non-finite d give NaN (the callers' finiteness checks see them).
__device__ __forceinline__ double q3_rcp(double d) {
  double x;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(d));
  x = x * fma(-d, x, 2.0);
  x = x * fma(-d, x, 2.0);
  return fma(x, fma(-d, x, 1.0), x);
}
extern __shared__ __align__(16) char smem_stage[];
#define Q3_SMEM smem_stage
#endif

// lane-strided item loop: on the device the 32 lanes share the items, on the host one "lane" walks them all
// (not unrolled: the kernel is 20 k instructions, its warps sit in different loops, and instruction fetch - 12 % of the
// stall samples with these loops unrolled four times - costs more than the unrolling gains)
#define Q3_FOR(i, n) _Pragma("unroll 1") for (int i = Q3_LANE; i < (n); i += Q3_NL)

// Header at the start of the workspace: sizes, options, the cost / limit constants and the small Riccati scratch.
struct Hdr3 {
  int N, nn, cnt, rows;
  int refine, pdas_first, pdas_after, ipm_max;
  int n_rounds, n_ipm, n_fact, n_solve;
  double delta, cert_tol, ipm_mu0;
  double ipm_al;      // length of the interior point's last step (0: none taken yet), for the Tapia indicators
  double Qs2[9];      // 2 Qs
  double R2[4], dR2[4];
  double q0;          // 2 * (-vx_ref Q[0,0])   (LPV_Planner.py:412)
  double lim[6];      // min_vel max_vel max_rs max_ls max_ac max_dc
  double mind;        // min_dist
  double x0[9], uold[2];
  int nq;                   // nonzeros of the upper triangle of 2 Q (P = 2 blockdiag(...), LPV_Planner.py:399-410)
  double Q2u[45];           // 2 Q at the 45 upper-triangle pairs (tri order)
  // cost-to-go and the per-stage temporaries of the Riccati passes (layout: see ric_factor_backward3)
  alignas(16) double Pf[99];  // [Pxx (9 x 9); Pvx (2 x 9)], row stride 9 (bank-conflict free for one row per lane pair)
                            // (device: Pf .. dbuf, 248 doubles, double as the 2 x 88 exchange buffer of ric_factor_backward3)
  double Pvv[4], px[2][9], pv[2][2], pe[9];
  double Y[44], Yb[22];     // Pf [Acol | d] (11 x 4),  Pf [b0 | b1] (11 x 2)
  double TA[9], Apx[3], Fc[18];
  double dbuf[2][9];        // dynamics defect of the stage being processed / the next one
  unsigned long long hist[8];
  unsigned long long mbar;  // mbarrier of the bulk input copies (device)
  unsigned char tri[48];  // (i << 4) | j of the 45 upper-triangle pairs of a 9 x 9 matrix
  unsigned char wix[48];  // index into the stage's Weff of pair e (outputs 0, 3, 7, 8), 255 = none
  unsigned char nzl[48];  // the nq pairs e with 2 Q != 0
};

struct Ctx3 {
  int N, nn, cnt, rst, Rs, RT;
  Hdr3* hd;
  double *Acol, *Bcol, *ax, *ay, *mx, *my, *eyb, *pxy;   // problem data (mx, my: midpoint ego-neighbour, on the plane)
  double *x, *u, *sig, *lam, *tt;                         // iterate: x[k] = x_{k+1}
  double *dx, *du, *dsig, *dlam, *dtt;                    // direction
  double *gx, *gu, *gs;                                   // gradient handed to a solve: SAME storage as dx, du, dsig (the
                                                          // forward sweep overwrites stage k's gradient with stage k's step)
  double *D, *res;                                        // row weights, row residuals
  double *F, *Wi, *zk, *Si, *bh, *Weff;                   // Riccati factor per stage, slack elimination per stage
  unsigned short* rinfo;                                  // row j -> (k << 7) | r
  unsigned char* act;
};

#define Q3_HDR_BYTES ((sizeof(Hdr3) + 15) & ~(size_t)15)
#define Q3_AS 28  // doubles per stage of Acol (27 used): keeps every stage's block 16-byte aligned for 128-bit loads

// doubles per QP after the header
Q3HD size_t carve_workspace3(Ctx3& c, char* sm, int N, int nn) {
  c.N = N; c.nn = nn; c.rst = 4 + nn; c.Rs = N * c.rst; c.RT = c.Rs + 4 * N;
  c.hd = (Hdr3*)sm;
  double* p = (double*)(sm + Q3_HDR_BYTES);
  auto take = [&](size_t n) { double* r = p; p += n; return r; };
  const size_t n = (size_t)N, RT = (size_t)c.RT;
  c.Acol = take(Q3_AS * n); c.Bcol = take(4 * n);
  c.F = take(18 * n); c.Wi = take(4 * n); c.zk = take(2 * n);  // even offsets, even strides: 128-bit loads per stage
  c.ax = take(n * nn); c.ay = take(n * nn); c.mx = take(n * nn); c.my = take(n * nn);
  c.eyb = take(n); c.pxy = take(2 * n);
  c.x = take(9 * n); c.u = take(2 * n); c.sig = take(3 * n); c.lam = take(RT); c.tt = take(RT);
  c.dx = take(9 * n); c.du = take(2 * n); c.dsig = take(3 * n); c.dlam = take(RT); c.dtt = take(RT);
  c.gx = c.dx; c.gu = c.du; c.gs = c.dsig;
  c.D = take(RT); c.res = take(RT);
  c.Si = take(6 * n); c.bh = take(4 * n); c.Weff = take(10 * n);
  c.rinfo = (unsigned short*)take((RT + 3) / 4);
  c.act = (unsigned char*)take((RT + 7) / 8);
  return (size_t)((char*)p - sm);
}

// doubles the IPM iterate occupies when parked in the global scratch slice of a QP (x, u, sigma, lambda, t)
Q3HD size_t stage_park_doubles(int N, int nn) { return (size_t)14 * N + 2 * (size_t)N * (8 + nn); }

#define Q3_CTX(c)                                              \
  Ctx3 c;                                                      \
  {                                                            \
    const Hdr3* h_ = (const Hdr3*)Q3_SMEM;                     \
    carve_workspace3(c, Q3_SMEM, h_->N, h_->nn);               \
    c.cnt = h_->cnt;                                           \
  }

// ---------------------------------------------------------------------------------------------------
// rows
Q3D bool row_ok3(const Ctx3& c, int j) { return j >= c.Rs || (c.rinfo[j] & 127) < 4 + c.cnt; }

Q3D double row_b3(const Ctx3& c, int j) {
  const int ri = c.rinfo[j], r = ri & 127, k = ri >> 7;
  const double* lim = c.hd->lim;
  if (j >= c.Rs) return lim[2 + r];
  if (r == 0) return -lim[0];
  if (r == 1) return lim[1];
  if (r < 4) return c.eyb[k];
  const int i = k * c.nn + (r - 4);
  return fma(c.ax[i], c.mx[i], c.ay[i] * c.my[i]) - 0.5 * c.hd->mind;  // -min_dist/2 - b, b = -a.m (compute_plane.py:54)
}

// (A v)_j for states xs (x_{k+1} at 9k), slacks sg, inputs uv
Q3D double row_val3(const Ctx3& c, int j, const double* xs, const double* sg, const double* uv) {
  const int ri = c.rinfo[j], r = ri & 127, k = ri >> 7;
  if (j >= c.Rs) {
    const double v = uv[2 * k + (r >> 1)];
    return (r & 1) ? -v : v;
  }
  const double* xk = xs + 9 * k;
  const double* sk = sg + 3 * k;
  if (r == 0) return -xk[0];
  if (r == 1) return xk[0] + sk[0];
  if (r == 2) return xk[3] + sk[1];
  if (r == 3) return -xk[3] + sk[1];
  return c.ax[k * c.nn + (r - 4)] * xk[7] + c.ay[k * c.nn + (r - 4)] * xk[8] - sk[2];
}

Q3D void build_tables3(const Ctx3& c) {
  Q3_FOR(j, c.RT) {
    int k, r;
    if (j < c.Rs) { k = j / c.rst; r = j - k * c.rst; } else { k = (j - c.Rs) >> 2; r = (j - c.Rs) & 3; }
    c.rinfo[j] = (unsigned short)((k << 7) | r);
  }
  Q3_FOR(e, 45) {
    int i = 0, rem = e;
    while (rem >= 9 - i) { rem -= 9 - i; ++i; }
    const int j = i + rem;
    c.hd->tri[e] = (unsigned char)((i << 4) | j);
    const int a = i == 0 ? 0 : (i == 3 ? 1 : (i == 7 ? 2 : (i == 8 ? 3 : -1)));
    const int b = j == 0 ? 0 : (j == 3 ? 1 : (j == 7 ? 2 : (j == 8 ? 3 : -1)));
    c.hd->wix[e] = (a >= 0 && b >= 0) ? (unsigned char)(a * 4 - ((a * (a - 1)) >> 1) + (b - a)) : (unsigned char)255;
  }
}

// (A z)_j - b_j at the current iterate, evaluated so that the terms that cancel on an active row are subtracted first
// (state minus its bound is exact by Sterbenz when the row is tight; collision rows are taken relative to the plane's
// own point, the ego-neighbour midpoint): the rounding of an active row's residual is then far below an ulp of the
// state, which is what the multiplier iteration (penalty x residual) amplifies.
Q3D double row_res3(const Ctx3& c, int j) {
  const int ri = c.rinfo[j], r = ri & 127, k = ri >> 7;
  const double* lim = c.hd->lim;
  if (j >= c.Rs) {
    const double v = c.u[2 * k + (r >> 1)];
    return (r & 1) ? -v - lim[2 + r] : v - lim[2 + r];
  }
  const double* xk = c.x + 9 * k;
  const double* sk = c.sig + 3 * k;
  if (r == 0) return lim[0] - xk[0];
  if (r == 1) return (xk[0] - lim[1]) + sk[0];
  if (r == 2) return (xk[3] - c.eyb[k]) + sk[1];
  if (r == 3) return (-xk[3] - c.eyb[k]) + sk[1];
  const int i = k * c.nn + (r - 4);
  return fma(c.ax[i], xk[7] - c.mx[i], c.ay[i] * (xk[8] - c.my[i])) + (0.5 * c.hd->mind - sk[2]);
}

// res_j at the current iterate, one lane per stage (straight-line code: no per-row decoding); invalid rows (unused
// neighbour slots) read -1.  Ends with a barrier.
Q3D void eval_residuals3(const Ctx3& c) {
  const double* lim = c.hd->lim;
  const double hm = 0.5 * c.hd->mind;
  Q3_FOR(k, c.N) {
    const double* xk = c.x + 9 * k;
    const double* sk = c.sig + 3 * k;
    double* r = c.res + k * c.rst;
    r[0] = lim[0] - xk[0];
    r[1] = (xk[0] - lim[1]) + sk[0];
    r[2] = (xk[3] - c.eyb[k]) + sk[1];
    r[3] = (-xk[3] - c.eyb[k]) + sk[1];
#pragma unroll 1
    for (int i = 0; i < c.nn; ++i) {
      const int q = k * c.nn + i;
      r[4 + i] = i < c.cnt ? fma(c.ax[q], xk[7] - c.mx[q], c.ay[q] * (xk[8] - c.my[q])) + (hm - sk[2]) : -1.0;
    }
    double* rb = c.res + c.Rs + 4 * k;
    const double u0 = c.u[2 * k], u1 = c.u[2 * k + 1];
    rb[0] = u0 - lim[2]; rb[1] = -u0 - lim[3]; rb[2] = u1 - lim[4]; rb[3] = -u1 - lim[5];
  }
  Q3_SYNC();
}

// ---------------------------------------------------------------------------------------------------
// Track lookup with the semantics of utilities/misc.py:78-101 / :105-126: wrap while s > L, clamp s < 0, segment i with
// s0[i] <= s < s0[i]+len[i]; -1 where the reference's int(np.where(...)) raises.  The wrap is one floor-based reduction
// plus a fix-up that keeps the reference's s == k L behaviour (a multiple of L stays at L, where no segment matches).
Q3D int map_segment3(const DevParams& P, double s) {
  const double L = P.seg_s0[P.nseg - 1] + P.seg_len[P.nseg - 1];
  if (!(s == s) || !(L > 0.0) || fabs(s) > 1e6 * L) return -1;
  if (s > L) {
    s -= L * floor(s / L);
    if (s <= 0.0) s += L;   // the reference's loop stops at the first value <= L: exact multiples end at L
    while (s > L) s -= L;
  }
  if (s < 0.0) s = 0.0;
  int found = -1, cnt = 0;
  for (int i = 0; i < P.nseg; ++i)
    if (s >= P.seg_s0[i] && s < P.seg_s0[i] + P.seg_len[i]) { found = i; ++cnt; }
  return cnt == 1 ? found : -1;
}

// planes (N,3,nn) (planes/compute_plane.py:41-68, keep_sign=True) + distance weights (utilities/misc.py:10-18) folded
// into the linear cost of x,y (LPV_Planner.py:416-422).  Nonzero if ego and a neighbour coincide.
Q3D int setup_planes3(const Ctx3& c, const DevParams& P, const double* nbr, const double* pose, double* planes_out) {
  int bad = 0;
  Q3_FOR(h, c.N) {
    const double ex = pose[2 * h], ey = pose[2 * h + 1];
    const double fx = pose[2 * (h + 1)], fy = pose[2 * (h + 1) + 1];
    double sx = 0.0, sy = 0.0;
#pragma unroll 1
    for (int n = 0; n < c.nn; ++n) {
      double ax = 0.0, ay = 0.0, bb = 0.0;
      if (n < c.cnt) {
        const double gx = nbr[(h * c.nn + n) * 2], gy = nbr[(h * c.nn + n) * 2 + 1];
        ax = gx - ex; ay = gy - ey;
        const double nrm = sqrt(ax * ax + ay * ay);
        if (!(nrm > 0.0)) bad = 1;
        ax /= nrm; ay /= nrm;
        bb = -0.5 * (ax * (ex + gx) + ay * (ey + gy));
        const double dx = fx - nbr[((h + 1) * c.nn + n) * 2], dy = fy - nbr[((h + 1) * c.nn + n) * 2 + 1];
        const double w = (2.0 * P.min_dist - sqrt(dx * dx + dy * dy)) / (double)c.cnt;  // rows 1..N (misc.py:15)
        sx = fma(w, ax, sx); sy = fma(w, ay, sy);
      }
      c.ax[h * c.nn + n] = ax; c.ay[h * c.nn + n] = ay;
      c.mx[h * c.nn + n] = n < c.cnt ? 0.5 * (ex + nbr[(h * c.nn + n) * 2]) : 0.0;
      c.my[h * c.nn + n] = n < c.cnt ? 0.5 * (ey + nbr[(h * c.nn + n) * 2 + 1]) : 0.0;
      if (planes_out) c.res[h * c.nn + n] = bb;  // offsets parked for the coalesced write below (res is not live yet)
    }
    c.pxy[2 * h] = P.wq * sx; c.pxy[2 * h + 1] = P.wq * sy;
  }
  if (planes_out) {  // planes (N, 3, nn) written as consecutive doubles: full lines, not 8-byte fragments 48 bytes apart
    Q3_SYNC();
    const int per = 3 * c.nn;
    Q3_FOR(i, c.N * per) {
      const int h = i / per, rem = i - h * per, cc = rem / c.nn, n = rem - cc * c.nn;
      const int q = h * c.nn + n;
      planes_out[i] = cc == 0 ? c.ax[q] : (cc == 1 ? c.ay[q] : c.res[q]);
    }
  }
  return bad;
}

// A_k = I + dt Ac, B_k = dt Bc along the previous prediction (LPV_Planner.py:491-589); A_k stored as its 9 x 3 column
// block (only columns 0-2 of Ac are non-zero).  Nonzero on a failed map lookup.
Q3D int setup_linearise3(const Ctx3& c, const DevParams& P, const double* xprev, int rows, const double* uprev) {
  int bad = 0;
  const int nk = c.N > rows ? c.N : rows;
  Q3_FOR(k, nk) {
    if (k >= c.N) {  // get_ey is evaluated on every row of Last_xPredicted (LPV_Planner.py:489)
      if (map_segment3(P, xprev[k * NSTATE + 6]) < 0) bad = 1;
      continue;
    }
    const int re = k < rows ? k : rows - 1;
    const int sege = map_segment3(P, xprev[re * NSTATE + 6]);
    const double* x = xprev + k * NSTATE;
    const double vx = x[0], vy = x[1], ey = x[3], epsi = x[4], theta = x[5];
    const int seg = map_segment3(P, x[6]);
    if (seg < 0 || sege < 0) { bad = 1; continue; }
    c.eyb[k] = P.seg_hw[sege];  // get_ey's own sm=1.0; the controller's sm is never applied (misc.py:105)
    const double cur = P.seg_curv[seg];
    const double delta = uprev[2 * k];
    double sd, cd, se, ce, st, ct;
    sincos(delta, &sd, &cd); sincos(epsi, &se, &ce); sincos(theta, &st, &ct);
    double A12 = 0, A13 = 0, A22 = 0, A23 = 0, A32 = 0, A33 = 0, B11 = 0;
    if (!(vx < 0.2)) {
      A12 = (sd * P.Cf) / (P.m * vx);
      A13 = (sd * P.Cf * P.lf) / (P.m * vx) + vy;
      A22 = -(P.Cr + P.Cf * cd) / (P.m * vx);
      A23 = -(P.lf * P.Cf * cd - P.lr * P.Cr) / (P.m * vx) - vx;
      A32 = -(P.lf * P.Cf * cd - P.lr * P.Cr) / (P.I * vx);
      A33 = -(P.lf * P.lf * P.Cf * cd + P.lr * P.lr * P.Cr) / (P.I * vx);
      B11 = -(sd * P.Cf) / P.m;
    }
    const double den = 1.0 - ey * cur;
    const double iden = 1.0 / den;
    const double Ac[27] = {-P.mu, A12, A13,
                           0.0, A22, A23,
                           0.0, A32, A33,
                           se, ce, 0.0,
                           iden * (-ce * cur), iden * (se * cur), 1.0,
                           0.0, 0.0, 1.0,
                           ce / den, -se / den, 0.0,
                           ct, -st, 0.0,
                           st, ct, 0.0};
    double* A = c.Acol + Q3_AS * k;
    for (int r = 0; r < 9; ++r)
      for (int j = 0; j < 3; ++j) A[r * 3 + j] = (r == j ? 1.0 : 0.0) + P.dt * Ac[r * 3 + j];
    double* B = c.Bcol + 4 * k;
    B[0] = P.dt * B11;
    B[1] = P.dt * ((cd * P.Cf) / P.m);
    B[2] = P.dt * ((P.lf * P.Cf * cd) / P.I);
    B[3] = P.dt;
  }
  return bad;
}

// y = A_k v for the 9 x 9 A_k = [Acol | e_3 .. e_8], row j
Q3D double arow3(const double* A, const double* v, int j) {
  double s = j >= 3 ? v[j] : 0.0;
  s = fma(A[j * 3 + 0], v[0], s); s = fma(A[j * 3 + 1], v[1], s); s = fma(A[j * 3 + 2], v[2], s);
  return s;
}
// (B_k du)_j
Q3D double brow3(const double* B, double d0, double d1, int j) {
  const double b0 = j < 3 ? B[j] : 0.0, b1 = j == 0 ? B[3] : 0.0;
  return fma(b0, d0, b1 * d1);
}

// ---------------------------------------------------------------------------------------------------
// Per-stage slack elimination for the row weights D: Si = (2Qs + E'DE)^-1 (3x3 symmetric: 00 01 02 11 12 22),
// bh = E'D C restricted to the outputs (vx, e_y, x, y), Weff = C'DC - bh Si bh' (4x4 symmetric: 00 01 02 03 11 12 13 22
// 23 33) = what the stage rows add to the Hessian of (vx, e_y, x, y)_{k+1} once the slacks are minimised out.
Q3D void stage_weights3(const Ctx3& c) {
  const double* Qs2 = c.hd->Qs2;
  Q3_FOR(t, c.N) {
    const double* d = c.D + t * c.rst;
    double sxx = 0, sxy = 0, syy = 0, sx = 0, sy = 0, sd = 0;
#pragma unroll 1
    for (int i = 0; i < c.cnt; ++i) {
      const double di = d[4 + i], x = c.ax[t * c.nn + i], y = c.ay[t * c.nn + i];
      sxx = fma(di * x, x, sxx); sxy = fma(di * x, y, sxy); syy = fma(di * y, y, syy);
      sx = fma(di, x, sx); sy = fma(di, y, sy); sd += di;
    }
    const double b0 = d[1], b1 = d[2] - d[3], b2 = -sx, b3 = -sy;
    const double s00 = Qs2[0] + d[1], s01 = Qs2[1], s02 = Qs2[2];
    const double s11 = Qs2[4] + d[2] + d[3], s12 = Qs2[5], s22 = Qs2[8] + sd;
    const double c00 = s11 * s22 - s12 * s12, c01 = s02 * s12 - s01 * s22, c02 = s01 * s12 - s02 * s11;
    const double c11 = s00 * s22 - s02 * s02, c12 = s01 * s02 - s00 * s12, c22 = s00 * s11 - s01 * s01;
    const double idet = q3_rcp(s00 * c00 + s01 * c01 + s02 * c02);
    double* si = c.Si + 6 * t;
    si[0] = c00 * idet; si[1] = c01 * idet; si[2] = c02 * idet; si[3] = c11 * idet; si[4] = c12 * idet; si[5] = c22 * idet;
    double* bh = c.bh + 4 * t;
    bh[0] = b0; bh[1] = b1; bh[2] = b2; bh[3] = b3;
    double* w = c.Weff + 10 * t;
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

// Hessian entry of the state x_{k+1} at upper-triangle pair e: 2Q + S' Weff_k S with S = rows (0, 3, 7, 8)
Q3D double qt3(const Ctx3& c, int k, int e) {
  const int w = c.hd->wix[e];
  return w != 255 ? c.hd->Q2u[e] + c.Weff[10 * k + w] : c.hd->Q2u[e];
}

// What a solve's gradient is made of: cost gradient + C'w with the per-row w of the mode.
//   G3_ALM   w = lam + D res                                       (method-of-multipliers polish pass)
//   G3_AFF   w = D (res + t)                                       (interior point, predictor; lam cancels)
//   G3_COR   w = D (res + t) + (smu - dtt dlam) / t                (interior point, corrector; dtt/dlam = affine step)
//   G3_LAG   w = lam                                               (Lagrangian gradient: the costate pass)
enum { G3_ALM = 0, G3_AFF = 1, G3_COR = 2, G3_LAG = 3 };

Q3D double row_w3(const Ctx3& c, int j, int mode, double smu) {
  if (mode == G3_ALM) return fma(c.D[j], c.res[j], c.lam[j]);
  if (mode == G3_AFF) return c.D[j] * (c.res[j] + c.tt[j]);
  if (mode == G3_COR) return fma(c.D[j], c.res[j] + c.tt[j], (smu - c.dtt[j] * c.dlam[j]) * q3_rcp(c.tt[j]));
  return c.lam[j];
}

// Gradient of the solve into c.gx / c.gu / c.gs (= the direction's storage), slack part folded into gx through the
// stage's Si / bh (stage_weights3 of the current D must have run).  One lane per stage.  Ends with a barrier.
Q3D void assemble_gradient3(const Ctx3& c, int mode, double smu) {
  const Hdr3* hd = c.hd;
  Q3_FOR(k, c.N) {
    const double* xk1 = c.x + 9 * k;
    const double* sk = c.sig + 3 * k;
    const double* uk = c.u + 2 * k;
    const int j0 = k * c.rst, jb = c.Rs + 4 * k;
    const double w0 = row_w3(c, j0, mode, smu), w1 = row_w3(c, j0 + 1, mode, smu);
    const double w2 = row_w3(c, j0 + 2, mode, smu), w3 = row_w3(c, j0 + 3, mode, smu);
    double wx = 0.0, wy = 0.0, ws = 0.0;
#pragma unroll 1
    for (int i = 0; i < c.cnt; ++i) {
      const double wi = row_w3(c, j0 + 4 + i, mode, smu);
      wx = fma(c.ax[k * c.nn + i], wi, wx); wy = fma(c.ay[k * c.nn + i], wi, wy); ws += wi;
    }
    double* g = c.gx + 9 * k;
    for (int i = 0; i < 9; ++i) g[i] = 0.0;
#pragma unroll 1
    for (int q = 0; q < hd->nq; ++q) {
      const int e = hd->nzl[q], i = hd->tri[e] >> 4, j = hd->tri[e] & 15;
      g[i] = fma(hd->Q2u[e], xk1[j], g[i]);
      if (i != j) g[j] = fma(hd->Q2u[e], xk1[i], g[j]);
    }
    double s0 = hd->Qs2[0] * sk[0] + hd->Qs2[1] * sk[1] + hd->Qs2[2] * sk[2] + w1;
    double s1 = hd->Qs2[3] * sk[0] + hd->Qs2[4] * sk[1] + hd->Qs2[5] * sk[2] + w2 + w3;
    double s2 = hd->Qs2[6] * sk[0] + hd->Qs2[7] * sk[1] + hd->Qs2[8] * sk[2] - ws;
    const double* um = k > 0 ? uk - 2 : hd->uold;
    const double d0 = uk[0] - um[0], d1 = uk[1] - um[1];
    double gu0 = hd->R2[0] * uk[0] + hd->R2[1] * uk[1] + hd->dR2[0] * d0 + hd->dR2[1] * d1;
    double gu1 = hd->R2[2] * uk[0] + hd->R2[3] * uk[1] + hd->dR2[2] * d0 + hd->dR2[3] * d1;
    if (k < c.N - 1) {
      const double e0 = uk[2] - uk[0], e1 = uk[3] - uk[1];
      gu0 -= hd->dR2[0] * e0 + hd->dR2[1] * e1;
      gu1 -= hd->dR2[2] * e0 + hd->dR2[3] * e1;
    }
    gu0 += row_w3(c, jb, mode, smu) - row_w3(c, jb + 1, mode, smu);
    gu1 += row_w3(c, jb + 2, mode, smu) - row_w3(c, jb + 3, mode, smu);
    // slack elimination: the slack gradient reaches the outputs as -bh Si gs
    const double* si = c.Si + 6 * k;
    const double* bh = c.bh + 4 * k;
    const double v0 = si[0] * s0 + si[1] * s1 + si[2] * s2;
    const double v1 = si[1] * s0 + si[3] * s1 + si[4] * s2;
    const double v2 = si[2] * s0 + si[4] * s1 + si[5] * s2;
    g[0] += hd->q0 + (w1 - w0) - bh[0] * v0;
    g[3] += (w2 - w3) - bh[1] * v1;
    g[7] += 2.0 * c.pxy[2 * k] + wx - bh[2] * v2;
    g[8] += 2.0 * c.pxy[2 * k + 1] + wy - bh[3] * v2;
    c.gs[3 * k] = s0; c.gs[3 * k + 1] = s1; c.gs[3 * k + 2] = s2;
    c.gu[2 * k] = gu0; c.gu[2 * k + 1] = gu1;
  }
  Q3_SYNC();
}

// dynamics defect of stage k at the current iterate, component j:  (A_k x_k + B_k u_k - x_{k+1})_j
Q3D double defect3(const Ctx3& c, int k, int j) {
  const double* xk = k > 0 ? c.x + 9 * (k - 1) : c.hd->x0;
  return arow3(c.Acol + Q3_AS * k, xk, j) + brow3(c.Bcol + 4 * k, c.u[2 * k], c.u[2 * k + 1], j) - c.x[9 * k + j];
}

// The dynamics defects of every stage at the current iterate, into the scratch c.dlam (9N <= 2 RT doubles over dlam, dtt:
// free whenever a solve runs with defects - the bounded starting point of the rounds and of the interior point).
// Ends with a barrier.
Q3D void stage_defects3(const Ctx3& c) {
  Q3_FOR(i, 9 * c.N) { const int k = i / 9; c.dlam[i] = defect3(c, k, i - 9 * k); }
  Q3_SYNC();
}

// One work item per lane.  On the device a phase is STRAIGHT-LINE code: every lane runs every item group's body on a
// clamped index and only the stores are predicated (l_ok) - a divergent region per group costs this latency-bound warp a
// reconvergence bubble each and keeps the compiler from interleaving the groups' dependent chains.  On the host (one
// "lane") the same bodies run as loops over the items.
#ifdef QP3_HOST
#define Q3_ITEM(l, n) for (int l = 0; l < (n); ++l)
#define Q3_ITEMS(l, lo, hi) for (int l = (lo); l < (hi); ++l)
#define Q3_PRED(l, n) for (int l = 0; l < (n); ++l) if (const bool l##_ok = true; true)
#define Q3_PREDS(l, lo, hi) for (int l = (lo); l < (hi); ++l) if (const bool l##_ok = true; true)
#define Q3_PRED2(l, n) for (int l = 0; l < (n); ++l) if (const bool l##_ok = true; true) if (const int l##_p = 0; true)
#define Q3_LOCAL_N(n) (n)              // per-lane copies of a table: the host keeps all n entries
#define Q3_LOCAL_E(q) (q)              // entry q of the per-lane copy = item q
#define Q3_LOCAL_AT(arr, l) arr[l]
#else
#define Q3_ITEM(l, n) for (int l = Q3_LANE; l < (n); l += 32)
#define Q3_ITEMS(l, lo, hi) for (int l = Q3_LANE; l >= (lo) && l < (hi); l += 32)
#define Q3_PRED(l, n) if (const bool l##_ok = Q3_LANE < (n); true) if (const int l = l##_ok ? Q3_LANE : (n) - 1; true)
#define Q3_PREDS(l, lo, hi) if (const bool l##_ok = Q3_LANE >= (lo) && Q3_LANE < (hi); true) if (const int l = l##_ok ? Q3_LANE : (lo); true)
#define Q3_PRED2(l, n)                                                            \
  _Pragma("unroll") for (int l##_p = 0; l##_p < 2; ++l##_p)                        \
    if (const bool l##_ok = Q3_LANE + 32 * l##_p < (n); true)                      \
      if (const int l = l##_ok ? Q3_LANE + 32 * l##_p : (n) - 1; true)
#define Q3_LOCAL_N(n) 2                // a lane keeps the entries of its (at most two) items
#define Q3_LOCAL_E(q) (Q3_LANE + 32 * (q) < 45 ? Q3_LANE + 32 * (q) : 44)
#define Q3_LOCAL_AT(arr, l) arr[l##_p]
#endif

// ---------------------------------------------------------------------------------------------------
// Riccati recursion.  Cost-to-go of the augmented state xi = (dx_k, dv_k = du_{k-1}) from stage k on, including the cost
// of x_k itself:  W_k(xi) = 1/2 xi' [Pxx Pxv; Pvx Pvv] xi + (px, pv)' xi.  Stage k, with P+ = P_{k+1}, defect d_k, the
// two columns b0 = (B0, B1, B2, 0..), b1 = (B3, 0..) of B_k and A_k = [Acol | e_3 .. e_8]:
//   G   = B'Pxx+ + Pvx+                       (2 x 9)         Huu = Rt + dR2 + G B + B'Pxv+ + Pvv+
//   F   = G A                                 (2 x 9)         hu  = gu + B'(px+ + Pxx+ d) + pv+ + Pvx+ d
//   du  = -Huu^-1 (hu + F dx - dR2 dv)
//   Pxx = A'Pxx+A - F'Huu^-1 F + Qt_{k-1}     Pvx = dR2 Huu^-1 F     Pvv = dR2 - dR2 Huu^-1 dR2
//   px  = A'(px+ + Pxx+ d) - F'Huu^-1 hu + gx_{k-1}               pv  = dR2 Huu^-1 hu
// Three phases per stage, every lane running the same code on a statically assigned item:
//   1  Y = Pf [Acol | d] (11 x 4) and Yb = Pf [b0 | b1] (11 x 2), Pf = [Pxx+; Pvx+]: lane (r, h) holds row r of Pf in
//      registers and produces two columns of Y and one of Yb;  pe = px+ + Pxx+ d
//   2  TA = Acol' Y[0:9, 0:3] and Acol' pe (twelve 9-term products), F (18 short products); Huu, its inverse and
//      z = Huu^-1 hu are computed by EVERY lane (uniform scalar code: no broadcast, no extra barrier)
//   3  the 45 + 18 + 3 + 9 + 2 entries of P_k, p_k, and the defect of the next stage to be processed
// Kept per stage for later solves with the same weights: F (18), Huu^-1 (3), z (2).
// In: c.gx / c.gu (gradient, slack part folded), c.D (box rows), c.Weff.  defects: d_k from the iterate, else zero.
Q3D void ric_factor_backward3(const Ctx3& c, const int defects) {
  Hdr3* hd = c.hd;
  const int N = c.N;
  double* Pf = hd->Pf;
  double* px = hd->px[0];
  double* pv = hd->pv[0];
  const double* dR2 = hd->dR2;
  // the triangle tables of this lane's entries, out of the stage loop (the loop's stores could alias the byte tables)
  int t_i[Q3_LOCAL_N(45)], t_j[Q3_LOCAL_N(45)], t_w[Q3_LOCAL_N(45)];
  double t_q[Q3_LOCAL_N(45)];
  for (int q = 0; q < Q3_LOCAL_N(45); ++q) {
    const int e = Q3_LOCAL_E(q), ij = hd->tri[e];
    t_i[q] = ij >> 4; t_j[q] = ij & 15; t_w[q] = hd->wix[e] != 255 ? hd->wix[e] : -1; t_q[q] = hd->Q2u[e];
  }
  {  // terminal cost-to-go: the cost of x_N (straight-line like the phases below, tables from the registers above)
    const double* wN = c.Weff + 10 * (N - 1);
    Q3_PRED2(e, 45) {
      const int i = Q3_LOCAL_AT(t_i, e), j = Q3_LOCAL_AT(t_j, e), w = Q3_LOCAL_AT(t_w, e);
      const double v = Q3_LOCAL_AT(t_q, e) + (w >= 0 ? wN[w >= 0 ? w : 0] : 0.0);
      if (e_ok) { Pf[i * 9 + j] = v; Pf[j * 9 + i] = v; }
    }
    Q3_PRED(e, 18) { if (e_ok) Pf[81 + e] = 0.0; }
    Q3_PRED(e, 9) {
      const double g = c.gx[9 * (N - 1) + e], dd = defects ? c.dlam[9 * (N - 1) + e] : 0.0;
      if (e_ok) { px[e] = g; hd->dbuf[(N - 1) & 1][e] = dd; hd->dbuf[N & 1][e] = 0.0; }
    }
    Q3_PREDS(l, 9, 13) { if (l_ok) hd->Pvv[l - 9] = 0.0; }
    Q3_PREDS(l, 13, 15) { if (l_ok) pv[l - 13] = 0.0; }
  }
  Q3_SYNC();
#pragma unroll 1
  for (int k = N - 1; k >= 0; --k) {
    const double* A = c.Acol + Q3_AS * k;
    const double* B = c.Bcol + 4 * k;
    const double* d = hd->dbuf[k & 1];
    const Q3_V2 bA = ((const Q3_V2*)B)[0], bB = ((const Q3_V2*)B)[1];
    const double B0 = bA.x, B1 = bA.y, B2 = bB.x, B3 = bB.y;
    // ---- phase 1
#if !defined(QP3_HOST) && defined(Q3_DMMA)
    // A/B variant (tensor cores): [Y | Yb] = Pf (11 x 9) [Acol | d | b0 | b1] (9 x 6) as six DMMA m8n8k4
    // (two row tiles x three k steps) instead of 18 FMAs on 22 lanes
    {
      const int lane = Q3_LANE, g = lane >> 2, t = lane & 3;
      double c00 = 0.0, c01 = 0.0, c10 = 0.0, c11 = 0.0;
#pragma unroll
      for (int sK = 0; sK < 3; ++sK) {
        const int kk = 4 * sK + t;
        const int kc = kk < 9 ? kk : 8;
        const double bv_a = A[3 * kc + (g < 3 ? g : 0)], bv_d = d[kc], bv_b = B[kc < 3 ? kc : 0];
        double bv = g < 3 ? bv_a : (g == 3 ? bv_d : (g == 4 ? (kk < 3 ? bv_b : 0.0) : ((g == 5 && kk == 0) ? B3 : 0.0)));
        bv = kk < 9 ? bv : 0.0;
        const double a0v = Pf[g * 9 + kc];
        const double a1v = Pf[(g < 3 ? 8 + g : 8) * 9 + kc];
        const double av0 = kk < 9 ? a0v : 0.0, av1 = (kk < 9 && g < 3) ? a1v : 0.0;
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c00), "+d"(c01) : "d"(av0), "d"(bv));
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c10), "+d"(c11) : "d"(av1), "d"(bv));
      }
      // C fragment: rows g (tile 0) and 8 + g (tile 1), columns 2t, 2t + 1: t = 0, 1 -> Y, t = 2 -> Yb
      const double px0 = px[g], px1 = px[8];
      if (t < 2) {
        hd->Y[g * 4 + 2 * t] = c00; hd->Y[g * 4 + 2 * t + 1] = c01;
        if (g < 3) { hd->Y[(8 + g) * 4 + 2 * t] = c10; hd->Y[(8 + g) * 4 + 2 * t + 1] = c11; }
        if (t == 1) { hd->pe[g] = px0 + c01; if (g == 0) hd->pe[8] = px1 + c11; }
      } else if (t == 2) {
        hd->Yb[g * 2] = c00; hd->Yb[g * 2 + 1] = c01;
        if (g < 3) { hd->Yb[(8 + g) * 2] = c10; hd->Yb[(8 + g) * 2 + 1] = c11; }
      }
    }
#else
    Q3_PRED(l, 22) {
      const int r = l >> 1, h = l & 1;
      double p[9];
#pragma unroll
      for (int q = 0; q < 9; ++q) p[q] = Pf[r * 9 + q];
      const double* z0p = A + 2 * h;            // column 2h of Acol, stride 3
      const double* z1p = h ? d : A + 1;        // column 1 of Acol (stride 3) or the defect (stride 1)
      const int st1 = h ? 1 : 3;
      double a0 = 0.0, a1 = 0.0, b0 = 0.0, b1 = 0.0;
#pragma unroll
      for (int q = 0; q < 8; q += 2) {
        a0 = fma(p[q], z0p[3 * q], a0); a1 = fma(p[q + 1], z0p[3 * q + 3], a1);
        b0 = fma(p[q], z1p[st1 * q], b0); b1 = fma(p[q + 1], z1p[st1 * (q + 1)], b1);
      }
      a0 = fma(p[8], z0p[24], a0) + a1;
      b0 = fma(p[8], z1p[st1 * 8], b0) + b1;
      const double yb = h ? B3 * p[0] : fma(B0, p[0], fma(B1, p[1], B2 * p[2]));
      const double pxe = px[r < 9 ? r : 0] + b0;
      if (l_ok) {
        hd->Y[r * 4 + 2 * h] = a0;
        hd->Y[r * 4 + 2 * h + 1] = b0;
        hd->Yb[r * 2 + h] = yb;
        if (h && r < 9) hd->pe[r] = pxe;
      }
    }
#endif
    Q3_SYNC();
    // ---- phase 2: scalar section, identical on every lane
    const double* Yb = hd->Yb;
    const double* db = c.D + c.Rs + 4 * k;
    const double pe0 = hd->pe[0], pe1 = hd->pe[1], pe2 = hd->pe[2];
    const double hu0 = c.gu[2 * k] + pv[0] + hd->Y[39] + fma(B0, pe0, fma(B1, pe1, B2 * pe2));
    const double hu1 = c.gu[2 * k + 1] + pv[1] + hd->Y[43] + B3 * pe0;
    const double H00 = hd->R2[0] + db[0] + db[1] + dR2[0] + fma(B0, Yb[0], fma(B1, Yb[2], B2 * Yb[4])) + 2.0 * Yb[18] + hd->Pvv[0];
    const double H01 = hd->R2[1] + dR2[1] + fma(B0, Yb[1], fma(B1, Yb[3], B2 * Yb[5])) + Yb[19] + Yb[20] + hd->Pvv[1];
    const double H11 = hd->R2[3] + db[2] + db[3] + dR2[3] + B3 * Yb[1] + 2.0 * Yb[21] + hd->Pvv[3];
    const double idet = q3_rcp(H00 * H11 - H01 * H01);
    const double W00 = H11 * idet, W01 = -H01 * idet, W11 = H00 * idet;
    const double z0 = W00 * hu0 + W01 * hu1, z1 = W01 * hu0 + W11 * hu1;
    Q3_PRED(l, 12) {
      const int col = l < 9 ? l / 3 : l - 9;
      const double* v = l < 9 ? hd->Y + (l % 3) : hd->pe;
      const int sv = l < 9 ? 4 : 1;
      double a0 = 0.0, a1 = 0.0;
#pragma unroll
      for (int q = 0; q < 8; q += 2) { a0 = fma(A[3 * q + col], v[sv * q], a0); a1 = fma(A[3 * q + 3 + col], v[sv * (q + 1)], a1); }
      a0 = fma(A[24 + col], v[sv * 8], a0) + a1;
      if (l_ok) { if (l < 9) hd->TA[l] = a0; else hd->Apx[l - 9] = a0; }
    }
    Q3_PREDS(l, 12, 30) {
      const int e = l - 12, m = e >= 9, j = e - 9 * m;
      const int j3 = j < 3 ? j : 0;
      // j < 3: (B'Y)[m][j] + Y[9 + m][j];  else Yb[j][m] + Pvx[m][j]
      const double fl = m ? fma(B3, hd->Y[j3], hd->Y[40 + j3]) : fma(B0, hd->Y[j3], fma(B1, hd->Y[4 + j3], fma(B2, hd->Y[8 + j3], hd->Y[36 + j3])));
      const double fh = Yb[j * 2 + m] + Pf[81 + e];
      const double f = j < 3 ? fl : fh;
      if (l_ok) { hd->Fc[e] = f; c.F[18 * k + e] = f; }
    }
    Q3_PREDS(l, 30, 31) { if (l_ok) { c.Wi[4 * k] = W00; c.Wi[4 * k + 1] = W01; c.Wi[4 * k + 2] = W11; } }
    Q3_PREDS(l, 31, 32) { if (l_ok) { c.zk[2 * k] = z0; c.zk[2 * k + 1] = z1; } }
    Q3_SYNC();
    // ---- phase 3
    if (k > 0) {
      const double* Fc = hd->Fc;
      const double* wk = c.Weff + 10 * (k - 1);
      Q3_PRED2(e, 45) {
        const int i = Q3_LOCAL_AT(t_i, e), j = Q3_LOCAL_AT(t_j, e), w = Q3_LOCAL_AT(t_w, e);
        const double f0i = Fc[i], f1i = Fc[9 + i], f0j = Fc[j], f1j = Fc[9 + j];
        const double T = j < 3 ? hd->TA[i * 3 + j] : (i < 3 ? hd->Y[j * 4 + i] : Pf[i * 9 + j]);
        const double qt = Q3_LOCAL_AT(t_q, e) + (w >= 0 ? wk[w >= 0 ? w : 0] : 0.0);
        const double v = T - (W00 * (f0i * f0j) + W01 * (f0i * f1j + f1i * f0j) + W11 * (f1i * f1j)) + qt;
        if (e_ok) { Pf[i * 9 + j] = v; Pf[j * 9 + i] = v; }
      }
      Q3_PRED(e, 18) {
        const int m = e >= 9, j = e - 9 * m;
        const double fw0 = W00 * Fc[j] + W01 * Fc[9 + j], fw1 = W01 * Fc[j] + W11 * Fc[9 + j];
        if (e_ok) Pf[81 + e] = dR2[m * 2] * fw0 + dR2[m * 2 + 1] * fw1;
      }
      Q3_PRED(j, 9) {
        const double v = (j < 3 ? hd->Apx[j < 3 ? j : 0] : hd->pe[j]) - Fc[j] * z0 - Fc[9 + j] * z1 + c.gx[9 * (k - 1) + j];
        if (j_ok) px[j] = v;
      }
      Q3_PREDS(l, 9, 11) { const int m = l - 9; if (l_ok) pv[m] = dR2[m * 2] * z0 + dR2[m * 2 + 1] * z1; }
      Q3_PREDS(l, 11, 14) {
        const int q = l - 11, m = q == 2, n = q != 0;  // (0,0) (0,1) (1,1)
        const double a0 = dR2[m * 2] * W00 + dR2[m * 2 + 1] * W01, a1 = dR2[m * 2] * W01 + dR2[m * 2 + 1] * W11;
        const double v = dR2[m * 2 + n] - (a0 * dR2[n] + a1 * dR2[2 + n]);
        if (l_ok) { hd->Pvv[m * 2 + n] = v; hd->Pvv[n * 2 + m] = v; }
      }
      Q3_PREDS(l, 14, 23) { const double v = defects ? c.dlam[9 * (k - 1) + (l - 14)] : 0.0; if (l_ok) hd->dbuf[(k - 1) & 1][l - 14] = v; }
    }
    Q3_SYNC();
  }
}

// Backward pass for a new gradient with the factor kept from ric_factor_backward3 (zero defects).  One phase per stage.
// ye != nullptr: the costates are written as the multipliers of the dynamics rows (rows of G, LPV_Planner.py:429-475):
// ye[12 t + j] = -px_t[j], t = N .. 1, and row block 0 from one adjoint step (gq0 = gradient of the cost in x_0).
// fb = 0: no feedback term (z = 0), i.e. the plain adjoint recursion px_k = A_k' px_{k+1} + gx_{k-1}: the costates that make
// the state rows of the stationarity condition hold exactly at ANY point (used when no working set was certified and
// the kept factor does not belong to the returned iterate).
Q3D void ric_affine_backward3(const Ctx3& c, double* ye, const int fb = 1) {
  Hdr3* hd = c.hd;
  const int N = c.N;
  const double* dR2 = hd->dR2;
  Q3_ITEM(e, 9) hd->px[0][e] = c.gx[9 * (N - 1) + e];
  Q3_ITEM(e, 2) hd->pv[0][e] = 0.0;
  Q3_SYNC();
  int cur = 0;
#pragma unroll 1
  for (int k = N - 1; k >= 0; --k) {
    const double* A = c.Acol + Q3_AS * k;
    const double* B = c.Bcol + 4 * k;
    const double* pin = hd->px[cur];
    const double* pvin = hd->pv[cur];
    const double* Fk = c.F + 18 * k;
    // what does not depend on the recursion first (the loads overlap the previous stage's barrier)
    const Q3_V2 bA = ((const Q3_V2*)B)[0], bB = ((const Q3_V2*)B)[1];
    const Q3_V2 wA = ((const Q3_V2*)(c.Wi + 4 * k))[0], wB = ((const Q3_V2*)(c.Wi + 4 * k))[1];
    const double B0 = bA.x, B1 = bA.y, B2 = bB.x, B3 = bB.y;
    const double gu0 = c.gu[2 * k], gu1 = c.gu[2 * k + 1];
    const double W00 = wA.x, W01 = wA.y, W11 = wB.x;
    const double hu0 = gu0 + pvin[0] + fma(B0, pin[0], fma(B1, pin[1], B2 * pin[2]));
    const double hu1 = gu1 + pvin[1] + B3 * pin[0];
    const double z0 = fb ? W00 * hu0 + W01 * hu1 : 0.0, z1 = fb ? W01 * hu0 + W11 * hu1 : 0.0;
    Q3_PRED(j, 9) {
      const int j3 = j < 3 ? j : 0;
      const double f0 = Fk[j], f1 = Fk[9 + j];
      const double gxm = c.gx[9 * (k > 0 ? k - 1 : 0) + j];
      double a0 = 0.0, a1 = 0.0;
#pragma unroll
      for (int q = 0; q < 8; q += 2) { a0 = fma(A[3 * q + j3], pin[q], a0); a1 = fma(A[3 * q + 3 + j3], pin[q + 1], a1); }
      a0 = fma(A[24 + j3], pin[8], a0) + a1;
      const double pj = pin[j];
      const double v = j < 3 ? a0 : pj;
      if (ye && j_ok) ye[12 * (k + 1) + j] = -pj;
      if (k > 0) { if (j_ok) hd->px[cur ^ 1][j] = v - f0 * z0 - f1 * z1 + gxm; }
      else if (ye && j_ok) {
        double gq = j == 0 ? hd->q0 : 0.0;
    #pragma unroll 1
    for (int q = 0; q < hd->nq; ++q) {
          const int e = hd->nzl[q], qi = hd->tri[e] >> 4, qj = hd->tri[e] & 15;
          if (qi == j) gq = fma(hd->Q2u[e], hd->x0[qj], gq);
          else if (qj == j) gq = fma(hd->Q2u[e], hd->x0[qi], gq);
        }
        ye[j] = -(v + gq);
      }
    }
    Q3_PREDS(l, 9, 11) { const int m = l - 9; if (l_ok) hd->pv[cur ^ 1][m] = dR2[m * 2] * z0 + dR2[m * 2 + 1] * z1; }
    Q3_PREDS(l, 11, 12) { if (l_ok) { c.zk[2 * k] = z0; c.zk[2 * k + 1] = z1; } }
    Q3_SYNC();
    cur ^= 1;
  }
}

// Forward pass: c.du, c.dx (over the gradient's storage), then the slack steps (c.dsig over c.gs).  defects: the
// dynamics defect of the iterate is added stage by stage, so that a full step lands on a dynamics-consistent point.
// One phase per stage.  Ends with a barrier.
Q3D void ric_forward3(const Ctx3& c, const int defects) {
  Hdr3* hd = c.hd;
  const double* dR2 = hd->dR2;
  double* fx = hd->Y;        // the running stage's step (ping-pong 2 x 10 (9 used), 2 x 2): over the backward sweeps' scratch
  double* fu = hd->Y + 20;
  Q3_ITEM(e, 20) fx[e] = 0.0;  // dx_0 = 0, dv_0 = 0: stage 0 runs the same code as the others
  Q3_ITEM(e, 4) fu[e] = 0.0;
  Q3_SYNC();
  int cur = 0;
#pragma unroll 1
  for (int k = 0; k < c.N; ++k) {
    const double* A = c.Acol + Q3_AS * k;
    // what does not depend on the recursion first (the loads overlap the previous stage's barrier); 128-bit loads
    double f[18];
    {
      const Q3_V2* F2 = (const Q3_V2*)(c.F + 18 * k);
#pragma unroll
      for (int q = 0; q < 9; ++q) { const Q3_V2 t = F2[q]; f[2 * q] = t.x; f[2 * q + 1] = t.y; }
    }
    const Q3_V2 wA = ((const Q3_V2*)(c.Wi + 4 * k))[0], wB = ((const Q3_V2*)(c.Wi + 4 * k))[1];
    const Q3_V2 zz = ((const Q3_V2*)(c.zk + 2 * k))[0];
    const Q3_V2 bA = ((const Q3_V2*)(c.Bcol + 4 * k))[0], bB = ((const Q3_V2*)(c.Bcol + 4 * k))[1];
    const double W00 = wA.x, W01 = wA.y, W11 = wB.x, zk0 = zz.x, zk1 = zz.y;
    const double B0 = bA.x, B1 = bA.y, B2 = bB.x, B3 = bB.y;
    double xin[10];
    {
      const Q3_V2* X2 = (const Q3_V2*)(fx + 10 * cur);
#pragma unroll
      for (int q = 0; q < 5; ++q) { const Q3_V2 t = X2[q]; xin[2 * q] = t.x; xin[2 * q + 1] = t.y; }
    }
    const Q3_V2 vv = ((const Q3_V2*)(fu + 2 * cur))[0];
    const double vin[2] = {vv.x, vv.y};
    double t0 = 0.0, t1 = 0.0, t2 = 0.0, t3 = 0.0;
#pragma unroll
    for (int l = 0; l < 8; l += 2) {
      t0 = fma(f[l], xin[l], t0); t1 = fma(f[l + 1], xin[l + 1], t1);
      t2 = fma(f[9 + l], xin[l], t2); t3 = fma(f[9 + l + 1], xin[l + 1], t3);
    }
    const double r0 = fma(f[8], xin[8], t0) + t1 - (dR2[0] * vin[0] + dR2[1] * vin[1]);
    const double r1 = fma(f[17], xin[8], t2) + t3 - (dR2[2] * vin[0] + dR2[3] * vin[1]);
    const double du0 = -zk0 - (W00 * r0 + W01 * r1);
    const double du1 = -zk1 - (W01 * r0 + W11 * r1);
    Q3_PRED(j, 9) {
      const double a0 = A[j * 3], a1 = A[j * 3 + 1], a2 = A[j * 3 + 2];
      const double dj = defects ? c.dlam[9 * k + j] : 0.0;
      const double bj0 = j == 0 ? B0 : (j == 1 ? B1 : (j == 2 ? B2 : 0.0)), bj1 = j == 0 ? B3 : 0.0;
      double v = j >= 3 ? fx[10 * cur + j] : 0.0;
      v = fma(a0, xin[0], v); v = fma(a1, xin[1], v); v = fma(a2, xin[2], v);
      v += fma(bj0, du0, bj1 * du1) + dj;
      if (j_ok) { fx[10 * (cur ^ 1) + j] = v; c.dx[9 * k + j] = v; }
    }
    Q3_PREDS(l, 9, 11) {
      const double v = l == 9 ? du0 : du1;
      if (l_ok) { fu[2 * (cur ^ 1) + (l - 9)] = v; c.du[2 * k + (l - 9)] = v; }
    }
    Q3_SYNC();
    cur ^= 1;
  }
  // slack steps: dsig = -Si (bh' do + gs), in place over gs
  Q3_FOR(t, c.N) {
    const double* si = c.Si + 6 * t;
    const double* bh = c.bh + 4 * t;
    const double* d = c.dx + 9 * t;
    const double r0 = -(c.gs[3 * t + 0] + bh[0] * d[0]);
    const double r1 = -(c.gs[3 * t + 1] + bh[1] * d[3]);
    const double r2 = -(c.gs[3 * t + 2] + bh[2] * d[7] + bh[3] * d[8]);
    c.dsig[3 * t + 0] = si[0] * r0 + si[1] * r1 + si[2] * r2;
    c.dsig[3 * t + 1] = si[1] * r0 + si[3] * r1 + si[4] * r2;
    c.dsig[3 * t + 2] = si[2] * r0 + si[4] * r1 + si[5] * r2;
  }
  Q3_SYNC();
}

// One instantiation each of the two solves (they are called from the active-set rounds and from the interior point).
// c.D must be final and barrier-visible.  mode / smu: see row_w3.  defects: the iterate is not dynamics-consistent (the
// bounded starting point): the solve removes the defects, a full step lands on a consistent point.  Every later step is
// a combination of consistent directions, so the iterate stays consistent to rounding and the flag stays off.
Q3N void factor_solve3(int mode, int defects) {
  Q3_CTX(c);
  stage_weights3(c);
  assemble_gradient3(c, mode, 0.0);
  if (defects) stage_defects3(c);
  ric_factor_backward3(c, defects);
  ric_forward3(c, defects);
  if (Q3_LANE == 0) { ++c.hd->n_fact; ++c.hd->n_solve; }
}
Q3N void affine_solve3(int mode, double smu) {
  Q3_CTX(c);
  assemble_gradient3(c, mode, smu);
  ric_affine_backward3(c, nullptr);
  ric_forward3(c, 0);
  if (Q3_LANE == 0) ++c.hd->n_solve;
}

// z <- z + alpha dz for the primal variables; returns max |dz| / (1 + max |z|) (all lanes).  Ends with a barrier.
Q3D double primal_step3(const Ctx3& c, double alpha) {
  double vm = 0.0, zm = 0.0;
  Q3_FOR(j, 9 * c.N) { zm = fmax(zm, fabs(c.x[j])); vm = fmax(vm, fabs(c.dx[j])); c.x[j] = fma(alpha, c.dx[j], c.x[j]); }
  Q3_FOR(j, 2 * c.N) { zm = fmax(zm, fabs(c.u[j])); vm = fmax(vm, fabs(c.du[j])); c.u[j] = fma(alpha, c.du[j], c.u[j]); }
  Q3_FOR(j, 3 * c.N) { c.sig[j] = fma(alpha, c.dsig[j], c.sig[j]); }
  vm = q3_wmax(vm); zm = q3_wmax(zm);
  Q3_SYNC();
  return alpha * vm / (1.0 + zm);
}

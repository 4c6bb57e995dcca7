// qp2_engine.cuh - "fast" one-CTA-per-QP engine (horizons whose whole workspace fits in shared memory).
//
// Same condensed problem and the same linear algebra as qp_engine.cuh (decision v = (u in R^{2N}, sigma in R^{3N}),
// rows of F from LPV_Planner.py:279-380, weighted normal equations reduced to a dense 2N x 2N SPD system), laid out
// for throughput on sm_100a:
//   * every matrix is PACKED: the prediction rows Gb_t (4 x 2(t+1), block lower triangular) and the symmetric H / K
//     as row-major lower triangles (index i(i+1)/2 + j).  Triangular numbers are a permutation mod 32, so a warp that
//     walks a column of a packed triangle touches 32 different banks.  ~50 KB per QP at N = 20 -> 4 CTAs per SM.
//   * factorisation and inversion are fused: one barrier per pivot.  The column is kept unscaled (a_ik = L_ik L_kk),
//     the trailing update uses a_ik a_jk / d_k, and while three warps update the trailing block the fourth writes
//     row k of L^-1 over row k of K (nobody reads that row any more).  Every later solve is two packed triangular
//     mat-vecs on all threads instead of a serial substitution.
//   * K = H + sum_t Gb_t' Weff_t Gb_t is accumulated over the ACTIVE stages only (in an active-set round most stages
//     have no active row and contribute nothing).
#pragma once
#include "qp_common.cuh"

#define TRI(i) ((((i) + 1) * (i)) >> 1)
#define ZCAP 3  // stages per Z chunk (Z aliases the linearisation block, 31N doubles >= 3 * 8N)

// Solver header at the start of dynamic shared memory: what the non-inlined primitives need to rebuild their view of
// the workspace (they take no pointer arguments, so every access stays a shared-memory access) plus the few solver
// options and the slack weights.
struct Hdr2 {
  int N, nn, NR, cnt;
  int refine, pdas_first, pdas_after, ipm_max;
  int n_rounds, n_ipm, pad0, pad1;
  double delta, cert_tol, ipm_mu0;
  double Qs[9];
  double Qsi[9];  // (2 Qs)^-1, for the exact slack elimination of the Schur-complement rounds
};

extern __shared__ __align__(16) char smem_fast[];

struct Ctx2 {
  int N, nn, cnt, nu, NR, rst, Rs, RT, ntri, WS, ldy;
  const Hdr2* hd;
  double *Hi, *Y, *S, *lw, *rw;  // Schur rounds: L_H^-1 (packed), Y = L_H^-1 A_W' (nu x ldy), S (packed), lambda_W, rhs_W
  int* wl;                       // rows of the working set
  double* bv;  // right-hand side b of every row (LPV_Planner.py:304-369)
  // problem data
  double *Gb, *Gx, *H, *K;
  double *Acol, *Bcol, *Z;  // Z aliases Acol/Bcol (re-linearised before the outputs are written)
  double *eyb, *pxy, *ax, *ay, *hh, *cr, *g;
  // iterates / work vectors
  double *u, *rhs, *du, *tv, *rdu, *yt;
  double *lam, *tt, *D, *res, *om, *dlam, *dtt;
  double *sig, *dsig, *rsig, *rds;
  double* sv;  // interior-point iterate kept between the phases of the mu ladder (nu + 3N + 2 RT)
  double *Weff, *Si, *bh, *wo, *dob;  // ro aliases wo, o aliases dob
  double *rs, *lrow;
  double* red;
  unsigned long long* hist;
  int *redi, *asl, *nas;
  unsigned short *tri, *rinfo;
  unsigned char *act, *sflag;
};

// packed prediction rows: stage t holds 4 rows of length 2(t+1); offset of stage t = 4 t (t+1)
#define GB(c, t, q, j) (c).Gb[4 * (t) * ((t) + 1) + (q) * 2 * ((t) + 1) + (j)]
// extra (cost-only) rows r >= 4, same packing per row: row r-4 occupies N(N+1) doubles
#define GX(c, t, r, j) (c).Gx[((r) - 4) * (c).N * ((c).N + 1) + (t) * ((t) + 1) + (j)]

__host__ __device__ inline size_t carve_workspace2(Ctx2& c, char* sm, int N, int nn, int NR) {
  c.N = N; c.nn = nn; c.nu = 2 * N; c.NR = NR;
  c.rst = 4 + nn; c.Rs = N * c.rst; c.RT = c.Rs + 4 * N; c.ntri = N * (2 * N + 1);
  double* p = (double*)sm;
  auto take = [&](size_t n) { double* r = p; p += n; return r; };
  c.Gb = take((size_t)4 * N * (N + 1));
  c.H = take(c.ntri);
  size_t kx = (size_t)(NR > 4 ? NR - 4 : 0) * N * (N + 1);  // cost-only rows live in the K block during set-up
  c.K = take(kx > (size_t)c.ntri ? kx : (size_t)c.ntri);
  c.Gx = c.K;
  c.Acol = take(27 * N); c.Bcol = take(4 * N);
  {  // Schur-round scratch aliases K + the linearisation block: largest working set (<= 32 rows) that fits
    const size_t room = (size_t)((c.Acol + 31 * N) - c.K);
    int ws = 32;
    while (ws > 1 && (size_t)c.nu * (ws + 1) + (size_t)ws * (ws + 1) / 2 > room) --ws;
    c.WS = ws; c.ldy = ws + 1;
    c.Y = c.K; c.S = c.K + (size_t)c.nu * c.ldy;
  }
  c.Z = c.Acol;
  c.eyb = take(N); c.pxy = take(2 * N);
  c.ax = take((size_t)N * nn); c.ay = take((size_t)N * nn); c.hh = take((size_t)N * nn);
  c.cr = take((size_t)N * NR); c.g = take(c.nu);
  c.u = take(c.nu); c.rhs = take(c.nu); c.du = take(c.nu); c.tv = take(c.nu); c.rdu = take(c.nu);
  c.yt = c.tv;  // H u is consumed before any triangular solve starts
  c.lam = take(c.RT); c.tt = take(c.RT); c.D = take(c.RT); c.res = take(c.RT); c.bv = take(c.RT);
  c.om = take(c.RT); c.dlam = take(c.RT); c.dtt = take(c.RT);  // contiguous: set-up scratch ql (N x NR <= 2 RT), Weff
  c.sig = take(3 * N); c.dsig = take(3 * N); c.rsig = take(3 * N); c.rds = take(3 * N);
  c.Weff = c.om;  // 10N <= 2 RT: only alive between stage_weights2 and build_K2, when om/dlam are dead
  c.Si = take(6 * N); c.bh = take(4 * N); c.wo = take(4 * N); c.dob = take(4 * N);
  c.rs = take(c.nu); c.lrow = take(2 * c.nu);
  c.Hi = take(c.ntri);
  {  // the Schur rounds are over (and L_H^-1 dead) once the interior-point phase starts
    const size_t need = (size_t)c.nu + 3 * N + 2 * (size_t)c.RT;
    c.sv = need <= (size_t)c.ntri ? c.Hi : take(need);
  }
  c.lw = take(32); c.rw = take(32); c.wl = (int*)take(16);
  c.red = take(8);
  c.hist = (unsigned long long*)take(16);
  c.redi = (int*)take(4);
  c.asl = (int*)take((N + 2) / 2 + 1);
  c.nas = c.asl + N;
  c.tri = (unsigned short*)take((c.ntri + 3) / 4);
  c.rinfo = (unsigned short*)take((c.RT + 3) / 4);
  c.act = (unsigned char*)take((c.RT + 7) / 8);
  c.sflag = (unsigned char*)take((N + 7) / 8);
  return (size_t)((char*)p - sm);
}

#define FAST_HDR_BYTES ((sizeof(Hdr2) + 15) & ~(size_t)15)
// view of this CTA's workspace, rebuilt from the header (cheap: a handful of integer multiply-adds per pointer used)
#define CTX2_LOCAL(c)                                                            \
  Ctx2 c;                                                                        \
  {                                                                              \
    const Hdr2* hd_ = (const Hdr2*)smem_fast;                                    \
    carve_workspace2(c, smem_fast + FAST_HDR_BYTES, hd_->N, hd_->nn, hd_->NR);   \
    c.cnt = hd_->cnt; c.hd = hd_;                                                \
  }

// index tables: tri[e] = (i << 8) | j for the packed lower triangle, rinfo[row] = (t << 4) | r for stage rows
// (r = 15 marks an unused neighbour slot) and (k << 4) | r for box rows.
__device__ __forceinline__ void build_tables(const Ctx2& c) {
#pragma unroll 1
  for (int i = threadIdx.x; i < c.nu; i += QP_THREADS)
    for (int j = 0; j <= i; ++j) c.tri[TRI(i) + j] = (unsigned short)((i << 8) | j);
#pragma unroll 1
  for (int t = threadIdx.x; t < c.N; t += QP_THREADS) {
    for (int r = 0; r < c.rst; ++r) c.rinfo[t * c.rst + r] = (unsigned short)((t << 4) | (r < 4 + c.cnt ? (r < 4 ? r : 4) : 15));
    for (int r = 0; r < 4; ++r) c.rinfo[c.Rs + 4 * t + r] = (unsigned short)((t << 4) | r);
  }
}

__device__ __forceinline__ bool row_valid2(const Ctx2& c, int j) { return j >= c.Rs || (c.rinfo[j] & 15) != 15; }

// right-hand side b of row j (LPV_Planner.py:304-369 after the row re-ordering :350-362)
__device__ __forceinline__ double row_h2_build(const Ctx2& c, const DevParams& P, int j) {
  const int ri = c.rinfo[j], r = ri & 15, t = ri >> 4;
  if (j >= c.Rs) return r == 0 ? P.max_rs : (r == 1 ? P.max_ls : (r == 2 ? P.max_ac : P.max_dc));
  if (r == 0) return -P.min_vel;
  if (r == 1) return P.max_vel;
  if (r < 4) return c.eyb[t];
  return c.hh[t * c.nn + (j - t * c.rst - 4)];
}

// value of row j of A v for stage outputs `o` (N x 4), slacks `sg` (N x 3) and inputs `uv`
__device__ __forceinline__ double row_val2(const Ctx2& c, int j, const double* o, const double* sg, const double* uv) {
  const int ri = c.rinfo[j], r = ri & 15, t = ri >> 4;
  if (j >= c.Rs) {
    const double v = uv[2 * t + (r >> 1)];
    return (r & 1) ? -v : v;
  }
  const double* ot = o + 4 * t;
  const double* st = sg + 3 * t;
  if (r == 0) return -ot[0];
  if (r == 1) return ot[0] + st[0];
  if (r == 2) return ot[1] + st[1];
  if (r == 3) return -ot[1] + st[1];
  const int i = j - t * c.rst - 4;
  return c.ax[t * c.nn + i] * ot[2] + c.ay[t * c.nn + i] * ot[3] - st[2];
}

// o[t][q] = Gb_t[q,:] . uv (+ offset): the four constrained outputs of every stage.   (syncs: caller)
__device__ __forceinline__ void stage_outputs2(const Ctx2& c, const double* uv, double* o, bool with_offset) {
#pragma unroll 1
  for (int w = threadIdx.x; w < 4 * c.N; w += QP_THREADS) {
    const int t = w >> 2, q = w & 3;
    const double* g = &GB(c, t, q, 0);
    double s0 = with_offset ? c.cr[t * c.NR + q] : 0.0, s1 = 0.0;
    const int jn = 2 * (t + 1);
    for (int j = 0; j < jn; j += 2) { s0 = fma(g[j], uv[j], s0); s1 = fma(g[j + 1], uv[j + 1], s1); }
    o[w] = s0 + s1;
  }
}

// A' w split into the stage-output part wo (N x 4), the slack part gs (N x 3); box part is added by apply_Gt2.
__device__ __forceinline__ void rows_to_stage2(const Ctx2& c, const double* w, double* wo, double* gs) {
#pragma unroll 1
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

// out[j] = base[j]*bs + ws * (sum_t Gb_t[:,j]' wo_t + box(wbox)).  `sparse`: wo is zero outside the active stages.
__device__ __forceinline__ void apply_Gt2(const Ctx2& c, const double* wo, const double* wbox, double ws,
                                          const double* base, double bs, double* out, bool sparse) {
#pragma unroll 1
  for (int j = threadIdx.x; j < c.nu; j += QP_THREADS) {
    double s0 = 0.0, s1 = 0.0;
    if (sparse) {
      const int nas = *c.nas;
      for (int a = 0; a < nas; ++a) {
        const int t = c.asl[a];
        if (2 * (t + 1) <= j) continue;
        const double* w4 = wo + 4 * t;
        s0 = fma(GB(c, t, 0, j), w4[0], s0); s1 = fma(GB(c, t, 1, j), w4[1], s1);
        s0 = fma(GB(c, t, 2, j), w4[2], s0); s1 = fma(GB(c, t, 3, j), w4[3], s1);
      }
    } else {
      for (int t = j >> 1; t < c.N; ++t) {
        const double* w4 = wo + 4 * t;
        s0 = fma(GB(c, t, 0, j), w4[0], s0); s1 = fma(GB(c, t, 1, j), w4[1], s1);
        s0 = fma(GB(c, t, 2, j), w4[2], s0); s1 = fma(GB(c, t, 3, j), w4[3], s1);
      }
    }
    double s = s0 + s1;
    if (wbox) {
      const double* wb = wbox + 4 * (j >> 1) + 2 * (j & 1);
      s += wb[0] - wb[1];
    }
    out[j] = (base ? bs * base[j] : 0.0) + ws * s;
  }
}

// out = H uv (H symmetric, packed lower)
__device__ __forceinline__ void apply_H2(const Ctx2& c, const double* uv, double* out) {
#pragma unroll 1
  for (int i = threadIdx.x; i < c.nu; i += QP_THREADS) {
    double s0 = 0.0, s1 = 0.0;
    const double* hr = c.H + TRI(i);
    int j = 0;
    for (; j + 1 <= i; j += 2) { s0 = fma(hr[j], uv[j], s0); s1 = fma(hr[j + 1], uv[j + 1], s1); }
    for (; j <= i; ++j) s0 = fma(hr[j], uv[j], s0);
    for (j = i + 1; j < c.nu; ++j) s1 = fma(c.H[TRI(j) + i], uv[j], s1);
    out[i] = s0 + s1;
  }
}

// Per-stage slack elimination for row weights D: Si = (2Qs + A_s'DA_s)^-1 (3x3), bh, Weff (4x4 symmetric), and the
// list of stages with a non-zero row weight.  Ends with a barrier.
__device__ __forceinline__ void stage_weights2(const Ctx2& c) {
  const double* Qs = c.hd->Qs;
#pragma unroll 1
  for (int t = threadIdx.x; t < c.N; t += QP_THREADS) {
    const double* d = c.D + t * c.rst;
    double sxx = 0, sxy = 0, syy = 0, sx = 0, sy = 0, sd = 0;
    for (int i = 0; i < c.cnt; ++i) {
      const double di = d[4 + i], x = c.ax[t * c.nn + i], y = c.ay[t * c.nn + i];
      sxx = fma(di * x, x, sxx); sxy = fma(di * x, y, sxy); syy = fma(di * y, y, syy);
      sx = fma(di, x, sx); sy = fma(di, y, sy); sd += di;
    }
    const double b0 = d[1], b1 = d[2] - d[3], b2 = -sx, b3 = -sy;
    const double s00 = 2 * Qs[0] + d[1], s01 = 2 * Qs[1], s02 = 2 * Qs[2];
    const double s11 = 2 * Qs[4] + d[2] + d[3], s12 = 2 * Qs[5], s22 = 2 * Qs[8] + sd;
    const double c00 = s11 * s22 - s12 * s12, c01 = s02 * s12 - s01 * s22, c02 = s01 * s12 - s02 * s11;
    const double c11 = s00 * s22 - s02 * s02, c12 = s01 * s02 - s00 * s12, c22 = s00 * s11 - s01 * s01;
    const double idet = 1.0 / (s00 * c00 + s01 * c01 + s02 * c02);
    double* si = c.Si + 6 * t;  // 00 01 02 11 12 22
    si[0] = c00 * idet; si[1] = c01 * idet; si[2] = c02 * idet; si[3] = c11 * idet; si[4] = c12 * idet; si[5] = c22 * idet;
    double* bh = c.bh + 4 * t;
    bh[0] = b0; bh[1] = b1; bh[2] = b2; bh[3] = b3;
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
    c.sflag[t] = (d[0] + d[1] + d[2] + d[3] + sd) > 0.0;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int n = 0;
    for (int t = 0; t < c.N; ++t)
      if (c.sflag[t]) c.asl[n++] = t;
    *c.nas = n;
  }
  __syncthreads();
}

// ro_t = -Bh Si rsig_t (stage-output space), used to fold the slack right-hand side into the u system
__device__ __forceinline__ void slack_rhs_to_outputs2(const Ctx2& c, const double* rsig, double* ro) {
#pragma unroll 1
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
__device__ __forceinline__ void slack_backsolve2(const Ctx2& c, const double* rsig, const double* dob, double* dsig) {
#pragma unroll 1
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

// K = H + sum_{t active} Gb_t' Weff_t Gb_t + diag(box weights), packed lower.  Ends with a barrier.
__device__ __forceinline__ void build_K2(const Ctx2& c) {
  const int nas = *c.nas;
#pragma unroll 1
  for (int e = threadIdx.x; e < c.ntri; e += QP_THREADS) {
    const int ij = c.tri[e], i = ij >> 8, j = ij & 255;
    double s = c.H[e];
    if (i == j) {
      const double* db = c.D + c.Rs + 4 * (i >> 1) + 2 * (i & 1);
      s += db[0] + db[1];
    }
    c.K[e] = s;
  }
  for (int c0 = 0; c0 < nas; c0 += ZCAP) {
    const int nc = min(ZCAP, nas - c0);
    __syncthreads();
    // Z = Weff_t Gb_t for the stages of this chunk
    for (int sc = 0; sc < nc; ++sc) {
      const int t = c.asl[c0 + sc];
      const double* we = c.Weff + 10 * t;
      double* z = c.Z + (size_t)sc * 4 * c.nu;
#pragma unroll 1
      for (int j = threadIdx.x; j < 2 * (t + 1); j += QP_THREADS) {
        const double g0 = GB(c, t, 0, j), g1 = GB(c, t, 1, j), g2 = GB(c, t, 2, j), g3 = GB(c, t, 3, j);
        z[j] = we[0] * g0 + we[1] * g1 + we[2] * g2 + we[3] * g3;
        z[c.nu + j] = we[1] * g0 + we[4] * g1 + we[5] * g2 + we[6] * g3;
        z[2 * c.nu + j] = we[2] * g0 + we[5] * g1 + we[7] * g2 + we[8] * g3;
        z[3 * c.nu + j] = we[3] * g0 + we[6] * g1 + we[8] * g2 + we[9] * g3;
      }
    }
    __syncthreads();
#pragma unroll 1
    for (int e = threadIdx.x; e < c.ntri; e += QP_THREADS) {
      const int ij = c.tri[e], i = ij >> 8, j = ij & 255;
      double s = c.K[e];
      for (int sc = 0; sc < nc; ++sc) {
        const int t = c.asl[c0 + sc];
        if (2 * (t + 1) <= i) continue;
        const double* z = c.Z + (size_t)sc * 4 * c.nu + j;
        s = fma(GB(c, t, 0, i), z[0], s);
        s = fma(GB(c, t, 1, i), z[c.nu], s);
        s = fma(GB(c, t, 2, i), z[2 * c.nu], s);
        s = fma(GB(c, t, 3, i), z[3 * c.nu], s);
      }
      c.K[e] = s;
    }
  }
  __syncthreads();
}

// In place: K (SPD, packed lower) -> L^-1 (packed lower), K = L L'.  One barrier per pivot.  Non-positive pivots are
// floored.  The trailing triangle of step k is walked through the tri[] table (item e' -> (i', j'), shifted by k+1),
// so every lane of the updating warps has work; the remaining warps write row k of L^-1 over row k of K.
__device__ __forceinline__ void chol_inverse(const Ctx2& c) {
  const int n = c.nu, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    double d = c.K[0];
    if (!(d > 1e-300)) d = 1e-30;
    c.rs[0] = rsqrt(d);
  }
  __syncthreads();
#pragma unroll 1
  for (int k = 0; k < n; ++k) {
    const double rsk = c.rs[k];
    const int nl = k < 14 ? 1 : (k < 27 ? 2 : 3);  // warps on the L^-1 row; the others update the trailing block
    if (warp < QP_WARPS - nl) {
      const double invd = rsk * rsk;
      const int m = n - 1 - k, items = (m * (m + 1)) >> 1, stride = 32 * (QP_WARPS - nl);
#pragma unroll 2
      for (int e = threadIdx.x; e < items; e += stride) {
        const int ij = c.tri[e], i = k + 1 + (ij >> 8), j = k + 1 + (ij & 255);
        const int ti = TRI(i);
        const double v = fma(-c.K[ti + k] * invd, c.K[TRI(j) + k], c.K[ti + j]);
        c.K[ti + j] = v;
        if (e == 0) {  // (i, j) = (k+1, k+1): next pivot
          double d = v;
          if (!(d > 1e-300)) d = 1e-30;
          c.rs[k + 1] = rsqrt(d);
        }
      }
    } else {
      const int lw = warp - (QP_WARPS - nl);  // 0 .. nl-1
      // stage row k+1 of L (scaled) for the next step; all of its entries are final already
      if (lw == 0 && k + 1 < n) {
        double* ln = c.lrow + ((k + 1) & 1) * n;
        const double* row = c.K + TRI(k + 1);
        for (int j = lane; j <= k; j += 32) ln[j] = row[j] * c.rs[j];
      }
      // row k of L^-1:  Linv[k][t] = -(sum_{j=t}^{k-1} L[k][j] Linv[j][t]) / L[k][k]
      const double* lk = c.lrow + (k & 1) * n;
      for (int t = lane + 32 * lw; t <= k; t += 32 * nl) {
        double s0 = 0.0, s1 = 0.0;
        int j = t, tj = TRI(t) + t;
#pragma unroll 1
        for (; j + 1 < k; j += 2) {
          s0 = fma(lk[j], c.K[tj], s0);
          tj += j + 1;
          s1 = fma(lk[j + 1], c.K[tj], s1);
          tj += j + 2;
        }
        if (j < k) s0 = fma(lk[j], c.K[tj], s0);
        c.K[TRI(k) + t] = (t == k) ? rsk : -(s0 + s1) * rsk;
      }
    }
    __syncthreads();
  }
}

// x <- K^-1 x with K^-1 = Linv' Linv (Linv packed lower in c.K).  Ends with a barrier.
__device__ __forceinline__ void chol_solve2(const Ctx2& c, double* x) {
  const int n = c.nu;
#pragma unroll 1
  for (int i = threadIdx.x; i < n; i += QP_THREADS) {
    const double* r = c.K + TRI(i);
    double s0 = 0.0, s1 = 0.0;
    int j = 0;
    for (; j + 1 <= i; j += 2) { s0 = fma(r[j], x[j], s0); s1 = fma(r[j + 1], x[j + 1], s1); }
    for (; j <= i; ++j) s0 = fma(r[j], x[j], s0);
    c.yt[i] = s0 + s1;
  }
  __syncthreads();
#pragma unroll 1
  for (int j = threadIdx.x; j < n; j += QP_THREADS) {
    double s0 = 0.0, s1 = 0.0;
    int i = j;
    for (; i + 1 < n; i += 2) { s0 = fma(c.K[TRI(i) + j], c.yt[i], s0); s1 = fma(c.K[TRI(i + 1) + j], c.yt[i + 1], s1); }
    for (; i < n; ++i) s0 = fma(c.K[TRI(i) + j], c.yt[i], s0);
    x[j] = s0 + s1;
  }
  __syncthreads();
}

// Weights D -> Si/bh/Weff, K = H + A'DA, K <- inverse Cholesky factor.  One instantiation (not inlined).
__device__ __noinline__ void factorise2() {
  CTX2_LOCAL(c);
  stage_weights2(c);
  build_K2(c);
  chol_inverse(c);
}

// Direction for (c.rhs, c.rsig): c.du, c.dsig, c.dob (= Gb du).  K must hold the inverse factor for the current D.
__device__ __noinline__ void solve_direction2(int sparse) {
  CTX2_LOCAL(c);
  slack_rhs_to_outputs2(c, c.rsig, c.wo);
  __syncthreads();
  apply_Gt2(c, c.wo, nullptr, 1.0, c.rhs, 1.0, c.du, sparse);
  __syncthreads();
  chol_solve2(c, c.du);
  stage_outputs2(c, c.du, c.dob, false);
  __syncthreads();
  slack_backsolve2(c, c.rsig, c.dob, c.dsig);
  __syncthreads();
}

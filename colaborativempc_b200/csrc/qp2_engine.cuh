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
  int n_rounds, n_ipm, WS, pad1;  // WS: working-set capacity of the Schur rounds (found once by the first carve)
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

__host__ __device__ inline size_t carve_workspace2(Ctx2& c, char* sm, int N, int nn, int NR, int ws_known = 0) {
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
    int ws = ws_known > 0 ? ws_known : 32;
    while (ws_known <= 0 && ws > 1 && ((size_t)c.nu * (ws + 1) + (size_t)ws * (ws + 1) / 2 > room || 8 * ws > 2 * c.RT)) --ws;  // 8 x ws
    // scratch of the S factorisation lives in dlam/dtt
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
    carve_workspace2(c, smem_fast + FAST_HDR_BYTES, hd_->N, hd_->nn, hd_->NR, hd_->WS);   \
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
    double s0 = with_offset ? c.cr[t * c.NR + q] : 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    const int jn = 2 * (t + 1);
    int j = 0;
    for (; j + 3 < jn; j += 4) {
      s0 = fma(g[j], uv[j], s0); s1 = fma(g[j + 1], uv[j + 1], s1);
      s2 = fma(g[j + 2], uv[j + 2], s2); s3 = fma(g[j + 3], uv[j + 3], s3);
    }
    if (j < jn) { s0 = fma(g[j], uv[j], s0); s1 = fma(g[j + 1], uv[j + 1], s1); }
    o[w] = (s0 + s1) + (s2 + s3);
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
// Two threads per column (even / odd positions of the stage walk), combined by one shuffle.
__device__ __forceinline__ void apply_Gt2(const Ctx2& c, const double* wo, const double* wbox, double ws,
                                          const double* base, double bs, double* out, bool sparse) {
  const int items = 2 * c.nu;
#pragma unroll 1
  for (int e0 = 0; e0 < items; e0 += QP_THREADS) {  // uniform trip count: the shuffle below needs whole warps
    const int e = e0 + threadIdx.x;
    const bool valid = e < items;
    const int j = valid ? e >> 1 : 0, half = e & 1;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    if (valid) {
      if (sparse) {
        const int nas = *c.nas;
        for (int a = half; a < nas; a += 2) {
          const int t = c.asl[a];
          if (2 * (t + 1) <= j) continue;
          const double* w4 = wo + 4 * t;
          s0 = fma(GB(c, t, 0, j), w4[0], s0); s1 = fma(GB(c, t, 1, j), w4[1], s1);
          s2 = fma(GB(c, t, 2, j), w4[2], s2); s3 = fma(GB(c, t, 3, j), w4[3], s3);
        }
      } else {
        for (int t = (j >> 1) + half; t < c.N; t += 2) {
          const double* w4 = wo + 4 * t;
          s0 = fma(GB(c, t, 0, j), w4[0], s0); s1 = fma(GB(c, t, 1, j), w4[1], s1);
          s2 = fma(GB(c, t, 2, j), w4[2], s2); s3 = fma(GB(c, t, 3, j), w4[3], s3);
        }
      }
    }
    double s = (s0 + s1) + (s2 + s3);
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    if (valid && half == 0) {
      if (wbox) {
        const double* wb = wbox + 4 * (j >> 1) + 2 * (j & 1);
        s += wb[0] - wb[1];
      }
      out[j] = (base ? bs * base[j] : 0.0) + ws * s;
    }
  }
}

// out = H uv (H symmetric, packed lower)
__device__ __forceinline__ void apply_H2(const Ctx2& c, const double* uv, double* out) {
#pragma unroll 1
  for (int i = threadIdx.x; i < c.nu; i += QP_THREADS) {
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    const double* hr = c.H + TRI(i);
    int j = 0;
    for (; j + 1 <= i; j += 2) { s0 = fma(hr[j], uv[j], s0); s1 = fma(hr[j + 1], uv[j + 1], s1); }
    for (; j <= i; ++j) s0 = fma(hr[j], uv[j], s0);
    for (j = i + 1; j + 1 < c.nu; j += 2) { s2 = fma(c.H[TRI(j) + i], uv[j], s2); s3 = fma(c.H[TRI(j + 1) + i], uv[j + 1], s3); }
    if (j < c.nu) s2 = fma(c.H[TRI(j) + i], uv[j], s2);
    out[i] = (s0 + s1) + (s2 + s3);
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
// Register-tiled, no staging buffer and no barrier inside: a thread owns 8 rows x 1 column of K (rows 8rb..8rb+7,
// column j <= 8rb+7), walks the active stages that reach its rows (2(t+1) > 8rb), forms z = Weff_t Gb_t[:, j] (4 values)
// and adds Gb_t[:, i]' z for its 8 rows.  Gb rows are read as double2 (all offsets are even).
__device__ __forceinline__ void build_K2(const Ctx2& c) {
  const int nas = *c.nas, nu = c.nu;
  const int nrb = (nu + 7) >> 3;
  int ntile = 0;
  for (int rb = 0; rb < nrb; ++rb) ntile += min(nu, 8 * rb + 8);
#pragma unroll 1
  for (int tile = threadIdx.x; tile < ntile; tile += QP_THREADS) {
    int rb = 0, j = tile;
    while (j >= min(nu, 8 * rb + 8)) { j -= min(nu, 8 * rb + 8); ++rb; }
    const int i0 = 8 * rb;
    double acc[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const int i = i0 + r;
      acc[r] = (i < nu && i >= j) ? c.H[TRI(i) + j] : 0.0;
      if (i == j) {
        const double* db = c.D + c.Rs + 4 * (i >> 1) + 2 * (i & 1);
        acc[r] += db[0] + db[1];
      }
    }
#pragma unroll 1
    for (int a = 0; a < nas; ++a) {
      const int t = c.asl[a], len = 2 * (t + 1);
      if (len <= i0 || len <= j) continue;
      const double* gb = c.Gb + 4 * t * (t + 1);
      const double* we = c.Weff + 10 * t;
      const double g0 = gb[j], g1 = gb[len + j], g2 = gb[2 * len + j], g3 = gb[3 * len + j];
      const double z0 = we[0] * g0 + we[1] * g1 + we[2] * g2 + we[3] * g3;
      const double z1 = we[1] * g0 + we[4] * g1 + we[5] * g2 + we[6] * g3;
      const double z2 = we[2] * g0 + we[5] * g1 + we[7] * g2 + we[8] * g3;
      const double z3 = we[3] * g0 + we[6] * g1 + we[8] * g2 + we[9] * g3;
      const double2* r0 = (const double2*)(gb + i0);
      const double2* r1 = (const double2*)(gb + len + i0);
      const double2* r2 = (const double2*)(gb + 2 * len + i0);
      const double2* r3 = (const double2*)(gb + 3 * len + i0);
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        if (i0 + 2 * p < len) {
          const double2 a0 = r0[p], a1 = r1[p], a2 = r2[p], a3 = r3[p];
          acc[2 * p] = fma(a0.x, z0, fma(a1.x, z1, fma(a2.x, z2, fma(a3.x, z3, acc[2 * p]))));
          acc[2 * p + 1] = fma(a0.y, z0, fma(a1.y, z1, fma(a2.y, z2, fma(a3.y, z3, acc[2 * p + 1]))));
        }
      }
    }
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const int i = i0 + r;
      if (i < nu && i >= j) c.K[TRI(i) + j] = acc[r];
    }
  }
  __syncthreads();
}

// In place: A (SPD, packed lower) -> L^-1 (packed lower), A = L L'.  Blocked, 8 columns per step, with look-ahead:
// two barriers per step and the serial part (the 8 x 8 diagonal block) hidden behind the trailing update.
// Step s (rows/columns b0 = 8s .. b0+bs-1), entered with the diagonal block already replaced by its inverse factor
// Lss^-1 and T = L[blk, <b0] * Linv[<b0, <b0] in the scratch:
//   B    panel L[i, blk] = A[i, blk] * Lss^-T for the rows below;  row block of the inverse Linv[blk, <b0] = -Lss^-1 T.
//   C/A' warp 0 applies the panel to the NEXT diagonal block, factorises it in registers (one row per lane, shuffles)
//        and inverts it (one column per lane); the other warps do the rest of the trailing update
//        A[i][j] -= L[i, blk] . L[j, blk] in 2 x 2 register tiles and then the next step's T.
// Non-positive pivots are floored.  With d0 (the diagonal before the factorisation) a pivot below 1e-11 of its original
// diagonal entry sets *sing (numerically dependent rows); the factorisation still runs to the end.
// A: packed lower n x n (any n >= 1), T: 8 x n scratch, tri: the (i << 8 | j) table of packed-lower indices.
__device__ __forceinline__ void chol_diag_block(double* A, const int b0, const int bs, const double* d0, int* sing) {
  const int lane = threadIdx.x & 31;
  const unsigned full = 0xffffffffu;
  const bool own = lane < bs;
  const double* row = A + TRI(b0 + (own ? lane : 0)) + b0;
  double a[8], rsv[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) a[q] = (own && q <= lane) ? row[q] : (q == lane ? 1.0 : 0.0);
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    double d = __shfl_sync(full, a[k], k);
    if (d0 && k < bs && lane == 0 && !(d > 1e-11 * d0[b0 + k])) *sing = 1;
    if (!(d > 1e-300)) d = 1e-30;
    const double rk = rsqrt(d);
    rsv[k] = rk;
    const double lk = (lane == k) ? d * rk : a[k] * rk;
#pragma unroll
    for (int j = k + 1; j < 8; ++j) {
      const double v = __shfl_sync(full, lk, j);
      a[j] = fma(-lk, v, a[j]);
    }
    a[k] = lk;
  }
  // column `lane` of the inverse: x[i] = -(sum_{m<i} l_im x[m]) / l_ii, x[lane] = 1 / l_ll, zero above
  double x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    double sacc = 0.0;
#pragma unroll
    for (int m = 0; m < i; ++m) {
      const double v = __shfl_sync(full, a[m], i);
      sacc = fma(v, x[m], sacc);
    }
    x[i] = (i == lane) ? rsv[i] : (i > lane ? -sacc * rsv[i] : 0.0);
  }
  if (own) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (i >= lane && i < bs) A[TRI(b0 + i) + b0 + lane] = x[i];
  }
}

__device__ __forceinline__ void chol_inverse_blk(double* A, const int n, double* T, const unsigned short* tri,
                                                 const double* d0, int* sing) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ldt = n;
  if (n <= 0) return;
  if (warp == 0) chol_diag_block(A, 0, min(8, n), d0, sing);
  __syncthreads();
#pragma unroll 1
  for (int b0 = 0; b0 < n; b0 += 8) {
    const int bs = min(8, n - b0);
    // ---------------- B
    {
      const int mrows = n - b0 - bs;
      const int npan = 4 * mrows;  // (row, column pair); the four threads of a row sit in one warp
      const int nit = (npan + b0 + QP_THREADS - 1) / QP_THREADS;
#pragma unroll 1
      for (int it = 0; it < nit; ++it) {
        const int e = threadIdx.x + it * QP_THREADS;
        const bool pan = e < npan;
        const int i = b0 + bs + (e >> 2), cp = (e & 3) * 2;
        double* kr = A + TRI(pan ? i : 0) + b0;
        double kv[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) kv[q] = (pan && q < bs && q <= cp + 1) ? kr[q] : 0.0;
        __syncwarp();  // in place: every thread of the row has read before any of them writes
        if (pan) {
          // L[i][b0+cc] = sum_{m<=cc} A[i][b0+m] Li[cc][m]
          if (cp < bs) {
            const double* li0 = A + TRI(b0 + cp) + b0;
            double o0 = 0.0;
#pragma unroll
            for (int q = 0; q < 8; ++q) if (q <= cp) o0 = fma(kv[q], li0[q], o0);
            kr[cp] = o0;
          }
          if (cp + 1 < bs) {
            const double* li1 = A + TRI(b0 + cp + 1) + b0;
            double o1 = 0.0;
#pragma unroll
            for (int q = 0; q < 8; ++q) if (q <= cp + 1) o1 = fma(kv[q], li1[q], o1);
            kr[cp + 1] = o1;
          }
        } else if (e < npan + b0) {
          const int j = e - npan;  // column of the inverse's row block
          double tv[8];
#pragma unroll
          for (int r = 0; r < 8; ++r) tv[r] = r < bs ? T[r * ldt + j] : 0.0;
#pragma unroll
          for (int r = 0; r < 8; ++r) {
            if (r < bs) {
              const double* li = A + TRI(b0 + r) + b0;
              double o = 0.0;
#pragma unroll
              for (int q = 0; q < 8; ++q) if (q <= r) o = fma(li[q], tv[q], o);
              A[TRI(b0 + r) + j] = -o;
            }
          }
        }
      }
    }
    __syncthreads();
    // ---------------- C / A'
    const int base = b0 + bs, rem = n - base;
    if (rem <= 0) break;
    const int nbs = min(8, rem);                 // next diagonal block
    const int m2 = (rem + 1) >> 1;               // 2 x 2 tiles of the trailing block (last one a row short for odd sizes)
    const int ntl = (m2 * (m2 + 1)) >> 1;
    const int ndiag = rem <= 8 ? ntl : 10;       // tiles inside the next diagonal block (the first 4 tile rows)
    if (warp == 0) {
      // the next diagonal block, element per lane, then factorise / invert it
      const int nel = (nbs * (nbs + 1)) >> 1;
#pragma unroll 1
      for (int e = lane; e < nel; e += 32) {
        const int ij = tri[e], i = base + (ij >> 8), j = base + (ij & 255);
        const double* li = A + TRI(i) + b0;
        const double* lj = A + TRI(j) + b0;
        double sacc = 0.0;
#pragma unroll
        for (int q = 0; q < 8; ++q) if (q < bs) sacc = fma(li[q], lj[q], sacc);
        A[TRI(i) + j] -= sacc;
      }
      __syncwarp();
      chol_diag_block(A, base, nbs, d0, sing);
    } else {
#pragma unroll 1
      for (int e = ndiag + (int)threadIdx.x - 32; e < ntl; e += QP_THREADS - 32) {
        const int ij = tri[e], i = base + 2 * (ij >> 8), j = base + 2 * (ij & 255);
        const bool i1 = i + 1 < n;
        const double* li0 = A + TRI(i) + b0;
        const double* li1 = i1 ? li0 + i + 1 : li0;
        const double* lj0 = A + TRI(j) + b0;
        const double* lj1 = (j + 1 < n) ? lj0 + j + 1 : lj0;
        double s00 = 0.0, s01 = 0.0, s10 = 0.0, s11 = 0.0;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          if (q < bs) {
            const double a0 = li0[q], a1 = li1[q], b0v = lj0[q], b1v = lj1[q];
            s00 = fma(a0, b0v, s00); s01 = fma(a0, b1v, s01);
            s10 = fma(a1, b0v, s10); s11 = fma(a1, b1v, s11);
          }
        }
        double* k0 = A + TRI(i) + j;
        k0[0] -= s00;
        if (i != j) k0[1] -= s01;
        if (i1) {
          double* k1 = k0 + i + 1;
          k1[0] -= s10;
          k1[1] -= s11;
        }
      }
      // next step's T[r][j] = sum_{k=j}^{base-1} L[base+r][k] Linv[k][j]; an item = (column j, half of the block's rows)
#pragma unroll 1
      for (int e = threadIdx.x - 32; e < 2 * base; e += QP_THREADS - 32) {
        const int j = e >> 1, rh = (e & 1) * 4;
        if (rh >= nbs) continue;
        const double* l0 = A + TRI(base + rh);  // rows past the block (nbs < 8) repeat its last row; never stored
        const double* l1 = A + TRI(base + min(rh + 1, nbs - 1));
        const double* l2 = A + TRI(base + min(rh + 2, nbs - 1));
        const double* l3 = A + TRI(base + min(rh + 3, nbs - 1));
        double t0 = 0.0, t1 = 0.0, t2 = 0.0, t3 = 0.0;
        int tk = TRI(j) + j;
#pragma unroll 2
        for (int k = j; k < base; ++k) {
          const double v = A[tk];
          tk += k + 1;
          t0 = fma(l0[k], v, t0);
          t1 = fma(l1[k], v, t1);
          t2 = fma(l2[k], v, t2);
          t3 = fma(l3[k], v, t3);
        }
        if (rh < nbs) T[(rh + 0) * ldt + j] = t0;
        if (rh + 1 < nbs) T[(rh + 1) * ldt + j] = t1;
        if (rh + 2 < nbs) T[(rh + 2) * ldt + j] = t2;
        if (rh + 3 < nbs) T[(rh + 3) * ldt + j] = t3;
      }
    }
    __syncthreads();
  }
}

// The one instantiation of the blocked factorisation (its unrolled register blocks are ~50 KB of code): operands are
// passed as offsets (in doubles) into the workspace so that every access stays a shared-memory access.
// d0_off < 0: no dependence test; else *(c.redi + 3) is set when a pivot collapses.
__device__ __noinline__ void chol_inverse_sh(int a_off, int n, int t_off, int d0_off) {
  CTX2_LOCAL(c);
  double* base = c.Gb;  // first array of the workspace
  chol_inverse_blk(base + a_off, n, base + t_off, c.tri, d0_off >= 0 ? base + d0_off : nullptr, c.redi + 3);
}
__device__ __forceinline__ void chol_inverse(const Ctx2& c) {
  chol_inverse_sh((int)(c.K - c.Gb), c.nu, (int)(c.Z - c.Gb), -1);
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

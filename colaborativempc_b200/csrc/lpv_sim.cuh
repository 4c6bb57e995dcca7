// lpv_sim.cuh - batched LPV model roll-out.
//
// Replaces planner/lib/plan_lib/distributedPlanner/LPV_val.py:77-94 (LPV_Model.sim) with its own linearisation
// LPV_val.py:96-195 (_EstimateABC of that file: the planner's formulas WITHOUT the "vx < 0.2" low-velocity branch):
//   A_k, B_k along xPred (rows 0..N-1) and u;  x_{k+1} = A_k x_k + B_k u_k from x_0 = xPred[0];
//   result = [x_1 .. x_N, ones(9)]  (np.delete(xPred, 0, 0) + a row of ones).
// get_ey is evaluated on every row of xPred by the reference (its value is unused here, a failed lookup raises).
//
// A pure streaming op: (N+1)*72 + N*16 bytes read, (N+1)*72 + 4 written per model, HBM-bound by construction.  One
// 128-thread CTA per TILE of SIM_TILE consecutive models, whose inputs and outputs are single contiguous byte ranges:
//   1  one thread stages the tile's xPred rows and inputs in shared memory with two bulk async copies (TMA,
//      cp.async.bulk + mbarrier) - full-line reads, no per-lane address arithmetic;
//   2  one lane per (model, stage) pair linearises it into 17 doubles (the 3 x 3 block of A_k that feeds back, the input
//      term B_k u_k, the ten non-trivial coefficients of the six integrator rows) - every lane busy on the sincos /
//      division work, conflict-free strides (9, 17);
//   3  one lane per MODEL carries its nine states through the N sequential steps in registers (30 FMAs a step, no
//      shuffles, no barriers) and overwrites the tile in place with the predicted rows;
//   4  one thread writes the tile back with one bulk async store - full-line writes instead of 72-byte row fragments.
// Several CTAs per SM are in different phases at any time, which is what overlaps the copies with the arithmetic.
// A tile whose byte ranges are not 16-byte multiples (odd tail tile, unaligned caller pointers) takes plain loads / stores.
#pragma once
#include "qp_common.cuh"

#ifndef SIM_TILE
#define SIM_TILE 16   // measured on B200 (N = 20): 8 x 64 threads 42 %, 12 x 64 40 %, 16 x 64 33 %, 16 x 128 45 % of the HBM copy peak
#endif
#ifndef SIM_THREADS
#define SIM_THREADS 128
#endif
#define SIM_ABS 17  // doubles per (model, stage)

// shared-memory layout (doubles after the 16-byte barrier slot): tile [T][N+1][9] | u [T][N][2] | AB [T][17 N + 1] | bad [T] ints
__host__ __device__ inline size_t sim_smem_bytes(int N) {
  size_t d = (size_t)SIM_TILE * (N + 1) * 9;
  d = (d + 1) & ~(size_t)1;
  d += (size_t)SIM_TILE * N * 2;
  d += (size_t)SIM_TILE * (SIM_ABS * N + 1);
  return 16 + d * sizeof(double) + SIM_TILE * sizeof(int);
}

// map_segment (qp_common.cuh) for a track whose segments tile [s0[0], end) exactly - checked on the host, DevParams.seg_tiled:
// "exactly one segment holds s" is then "s lies in the track" and its index is the number of later starts not above s.
__device__ __forceinline__ int sim_map_segment(const DevParams& P, double s) {
  if (!P.seg_tiled) return map_segment(P, s);
  const double L = P.seg_s0[P.nseg - 1] + P.seg_len[P.nseg - 1];
  if (!(s == s) || !(L > 0.0) || fabs(s) > 1e6 * L) return -1;
  if (s > L) {
    s -= L * floor(s / L);
    if (s <= 0.0) s += L;
    while (s > L) s -= L;
  }
  if (s < 0.0) s = 0.0;
  if (!(s >= P.seg_s0[0] && s < L)) return -1;
  int f = 0;
#pragma unroll 1
  for (int i = 1; i < P.nseg; ++i) f += (s >= P.seg_s0[i]) ? 1 : 0;
  return f;
}

// inv_m = 1 / m, inv_I = 1 / I.  The per-row reciprocals are branch-free Newton reciprocals (q3_rcp, about one ulp):
// six IEEE divisions per row were 15 % of this kernel's instructions; the roll-out agrees with the reference to 1e-12.
__device__ __forceinline__ void sim_linearise(const DevParams& P, const double* x, double delta, double acc, double cur, double* A,
                                              double inv_m, double inv_I) {
  const double vx = x[0], vy = x[1], ey = x[3], epsi = x[4], theta = x[5];
  double sd, cd, se, ce, st, ct;
  sincos(delta, &sd, &cd); sincos(epsi, &se, &ce); sincos(theta, &st, &ct);
  const double ivx = q3_rcp(vx), imvx = ivx * inv_m, iIvx = ivx * inv_I;
  const double A12 = (sd * P.Cf) * imvx;
  const double A13 = (sd * P.Cf * P.lf) * imvx + vy;
  const double A22 = -(P.Cr + P.Cf * cd) * imvx;
  const double A23 = -(P.lf * P.Cf * cd - P.lr * P.Cr) * imvx - vx;
  const double A32 = -(P.lf * P.Cf * cd - P.lr * P.Cr) * iIvx;
  const double A33 = -(P.lf * P.lf * P.Cf * cd + P.lr * P.lr * P.Cr) * iIvx;
  const double B11 = -(sd * P.Cf) * inv_m;
  const double iden = q3_rcp(1.0 - ey * cur);
  const double dt = P.dt;
  // rows 0-2 of A_k = I + dt Ac, columns 1-2 (column 0 is (1 - dt mu, 0, 0))
  A[0] = dt * A12; A[1] = dt * A13;
  A[2] = 1.0 + dt * A22; A[3] = dt * A23;
  A[4] = dt * A32; A[5] = 1.0 + dt * A33;
  // B_k u_k: only the first three states have an input term
  A[6] = fma(dt * B11, delta, dt * acc);
  A[7] = dt * ((cd * P.Cf) * inv_m) * delta;
  A[8] = dt * ((P.lf * P.Cf * cd) * inv_I) * delta;
  // integrator rows: e_y (vx, vy), e_psi (vx, vy; w has dt), s (vx, vy), x and y (vx, vy)
  A[9] = dt * se; A[10] = dt * ce;
  A[11] = dt * (iden * (-ce * cur)); A[12] = dt * (iden * (se * cur));
  A[13] = dt * (ce * iden); A[14] = dt * (-se * iden);
  A[15] = dt * ct; A[16] = dt * st;
}

__global__ void __launch_bounds__(SIM_THREADS) lpv_sim_kernel(const __grid_constant__ DevParams P, int B,
                                                                const double* __restrict__ xpred, const double* __restrict__ u,
                                                                double* __restrict__ out, int* __restrict__ status) {
  extern __shared__ __align__(16) char sim_raw[];
  const int N = P.N, tid = threadIdx.x;
  const int b0 = blockIdx.x * SIM_TILE;
  if (b0 >= B) return;
  const int nm = min(SIM_TILE, B - b0);  // models of this tile
  const int rowd = (N + 1) * NSTATE;     // doubles of one model's rows
  unsigned long long* mbar = (unsigned long long*)sim_raw;
  double* tile = (double*)(sim_raw + 16);
  double* ut = tile + (((size_t)SIM_TILE * rowd + 1) & ~(size_t)1);
  double* AB = ut + (size_t)SIM_TILE * N * 2;
  int* bad = (int*)(AB + (size_t)SIM_TILE * (SIM_ABS * N + 1));
  const int abm = SIM_ABS * N + 1;       // odd stride between the models' blocks
  const double* gx = xpred + (size_t)b0 * rowd;
  const double* gu = u + (size_t)b0 * N * 2;
  double* go = out + (size_t)b0 * rowd;
  const unsigned xbytes = (unsigned)nm * rowd * 8u, ubytes = (unsigned)nm * N * 16u;
  // bulk copies need 16-byte aligned addresses and sizes (a full tile always has them for aligned arrays)
  const bool bulk = ((((size_t)gx | (size_t)gu | (size_t)go) & 15) == 0) && (xbytes & 15) == 0;
  if (tid < SIM_TILE) bad[tid] = 0;

  // ---- 1: stage the tile
  if (bulk) {
    const unsigned bar = (unsigned)__cvta_generic_to_shared(mbar);
    if (tid == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(xbytes + ubytes) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                       (unsigned)__cvta_generic_to_shared(tile)), "l"(gx), "r"(xbytes), "r"(bar) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                       (unsigned)__cvta_generic_to_shared(ut)), "l"(gu), "r"(ubytes), "r"(bar) : "memory");
    }
    __syncthreads();  // the barrier is initialised before anyone polls it
    unsigned done = 0;
    while (!done)
      asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(bar) : "memory");
  } else {
    for (int i = tid; i < nm * rowd; i += SIM_THREADS) tile[i] = gx[i];
    for (int i = tid; i < nm * N * 2; i += SIM_THREADS) ut[i] = gu[i];
    __syncthreads();
  }

  const double inv_m = 1.0 / P.m, inv_I = 1.0 / P.I;
  // ---- 2: one lane per (model, row): map lookup on every row, linearisation on rows 0 .. N-1
  for (int t = tid; t < nm * (N + 1); t += SIM_THREADS) {
    const int m = t / (N + 1), k = t - m * (N + 1);
    const double* x = tile + (size_t)t * NSTATE;
    const int seg = sim_map_segment(P, x[6]);
    if (seg < 0) { bad[m] = 1; continue; }  // benign race: every writer stores 1
    if (k == N) continue;                   // row N: only the get_ey lookup
    const double2 uk = *(const double2*)(ut + ((size_t)m * N + k) * 2);
    sim_linearise(P, x, uk.x, uk.y, P.seg_curv[seg], AB + (size_t)m * abm + (size_t)k * SIM_ABS, inv_m, inv_I);
  }
  __syncthreads();

  // ---- 3: one lane per model: the N sequential steps in registers, rows written over the tile in place
  if (tid < nm) {
    const int m = tid;
    double* rows = tile + (size_t)m * rowd;
    const double* Am = AB + (size_t)m * abm;
    const double a00 = 1.0 + P.dt * (-P.mu), dt = P.dt;
    const bool isbad = bad[m] != 0;
    const double qnan = __longlong_as_double(0x7ff8000000000000ULL);
    double v[NSTATE];
#pragma unroll
    for (int j = 0; j < NSTATE; ++j) v[j] = rows[j];
#pragma unroll 2
    for (int k = 0; k < N; ++k) {
      const double* A = Am + (size_t)k * SIM_ABS;
      const double v0 = v[0], v1 = v[1], v2 = v[2];
      // same evaluation order per row as A_k x_k + B_k u_k with the zero entries left out
      v[0] = fma(A[1], v2, fma(A[0], v1, fma(a00, v0, A[6])));
      v[1] = fma(A[3], v2, fma(A[2], v1, A[7]));
      v[2] = fma(A[5], v2, fma(A[4], v1, A[8]));
      v[3] = fma(A[10], v1, fma(A[9], v0, v[3]));
      v[4] = fma(dt, v2, fma(A[12], v1, fma(A[11], v0, v[4])));
      v[5] = fma(dt, v2, v[5]);
      v[6] = fma(A[14], v1, fma(A[13], v0, v[6]));
      v[7] = fma(-A[16], v1, fma(A[15], v0, v[7]));
      v[8] = fma(A[15], v1, fma(A[16], v0, v[8]));
#pragma unroll
      for (int j = 0; j < NSTATE; ++j) rows[k * NSTATE + j] = isbad ? qnan : v[j];  // row k of the result = x_{k+1}
    }
#pragma unroll
    for (int j = 0; j < NSTATE; ++j) rows[N * NSTATE + j] = isbad ? qnan : 1.0;  // the reference raises on a failed lookup
    if (status) status[b0 + m] = isbad ? DMPC_MAP_LOOKUP_FAILED : 0;
  }

  // ---- 4: write the tile back
  if (bulk) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the generic-proxy writes above, before the async copy reads them
    __syncthreads();
    if (tid == 0) {
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(go), "r"((unsigned)__cvta_generic_to_shared(tile)), "r"(xbytes) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // shared memory stays valid until the copy has read it
    }
  } else {
    __syncthreads();
    for (int i = tid; i < nm * rowd; i += SIM_THREADS) go[i] = tile[i];
  }
}


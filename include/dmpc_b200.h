/*
 * dmpc_b200.h - C ABI of the B200-native LPV-MPC QP path.
 *
 * The reference (MarcFacerias/ColaborativeMPC-) has no FFI layer: its boundary for this path is the
 * Python object API  PlannerLPV(...).solve(...)  (planner/lib/plan_lib/distributedPlanner/LPV_Planner.py:17,
 * :115-182) called from  agent.one_step  (planner/scripts/LPV_HP_N_main.py:41) and the neighbour exchange of
 * planner/scripts/LPV_HP_N_main.py:117.  The entry points below are what a ctypes binding for that path
 * binds; colaborativempc_b200/planner.py is that binding.
 *
 * Conventions: plain pointers and sizes, no torch types.  All `d_*` pointers are DEVICE pointers owned by the
 * caller, all `h_*` pointers are HOST pointers owned by the caller.  Every function returns 0 on success or a
 * negative dmpc_error; dmpc_last_error() gives text.  A handle is bound to one device and one stream user at
 * a time (not thread-safe, like the reference controller object, which mutates self.*).
 */
#ifndef DMPC_B200_H
#define DMPC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DMPC_NS 9          /* states  [vx vy psidot e_y e_psi theta s x y]  (plan_lib/config/__init__.py:4) */
#define DMPC_NSLACK 3      /* slacks per stage                               (LPV_Planner.py:20-23)          */
#define DMPC_NEXP 12
#define DMPC_NU 2
#define DMPC_MAX_SEG 32    /* rows of Map.PointAndTangent kept per lane      (track_initialization.py:98-300)  */

/* Constructor arguments of PlannerLPV (LPV_Planner.py:17-105) + the map table read by
 * curvature()/get_ey() (utilities/misc.py:78-126). */
typedef struct dmpc_params {
  int32_t N;              /* horizon                                                         */
  int32_t nn;             /* neighbour slots per agent (len(agents_id) of the first call)    */
  int32_t nseg;           /* map segments                                                    */
  int32_t reserved;
  double dt, wq;
  double Q[81];           /* row-major 9x9                                                   */
  double Qs[9];           /* row-major 3x3                                                   */
  double R[4];
  double dR[4];
  double lf, lr, m, I, Cf, Cr, mu;                                     /* model_param        */
  double vx_ref, min_dist, max_vel, min_vel, max_rs, max_ls, max_ac, max_dc; /* sys_lim      */
  double seg_s0[DMPC_MAX_SEG];   /* PointAndTangent[:,3,lane]                                */
  double seg_len[DMPC_MAX_SEG];  /* PointAndTangent[:,4,lane]                                */
  double seg_curv[DMPC_MAX_SEG]; /* PointAndTangent[:,5,lane]                                */
  double seg_hw[DMPC_MAX_SEG];   /* halfWidth[:]                                             */
} dmpc_params;

/* Solver options.  The reference calls OSQP with polish=True, max_iter=500000 (LPV_Planner.py:233); the B200 path
 * returns the KKT-certified optimum of the same QP (OSQP's limit as eps -> 0, i.e. a successful polish), found by
 * active-set polish rounds with an interior-point phase as globalisation. */
typedef struct dmpc_options {
  double polish_delta;    /* regularisation of the polish KKT solve (OSQP's `delta`), 1e-9            */
  double cert_tol;        /* relative KKT certificate tolerance, 1e-9                                  */
  double ipm_mu0;         /* first complementarity target of the interior-point phase, 1e-6            */
  int32_t polish_refine;  /* refinement steps per polish solve (OSQP's polish_refine_iter), 4         */
  int32_t pdas_first;     /* active-set round budget before the interior-point phase, 4 (fast engine: 3x free Schur
                             rounds + 2x one-change rounds; generic engine / singular S: that many penalty rounds)   */
  int32_t pdas_after;     /* active-set polish rounds after each interior-point phase, 2               */
  int32_t ipm_max_iter;   /* interior-point iteration cap per phase, 60                                */
  int32_t engine;         /* dmpc_engine: 0 auto (the stage-wise engine), 2 condensed engine (workspace packed in shared memory;
                             dmpc_create fails if the horizon does not fit; QPs whose condensed Hessian cannot be factorised
                             are re-solved stage-wise in the same call), 3 stage-wise engine (any horizon)           */
  int32_t order_hint;     /* 1 = a call uses what the previous call of the same batch size spent on each QP index (consecutive
                             MPC steps are alike): the launch starts the most expensive QPs first, and a QP that ended in the
                             interior point last time goes there after two active-set rounds instead of twelve (stage-wise
                             engine; the condensed engine also splits its launch).  The certified optimum is the same; the
                             route to it - hence the last bits and the iteration counts - may differ.  Calls on one handle
                             must be stream-ordered with each other.  0 = off                                              */
  int32_t host_direct;    /* dmpc_solve_batch_host: 1 = when every buffer of the call is page-locked host memory
                             (cudaHostAlloc / cudaHostRegister, e.g. torch pin_memory) the kernel reads and writes it
                             in place over PCIe instead of staging copies; pageable buffers always take the staged
                             path.  Results are identical.  0 = always stage                                         */
  int32_t warm_start;     /* stage-wise engine: first working-set guess of a QP = the set the previous call of the same batch
                             size certified for that QP index (handle-owned, starts empty).  1 = as certified, 2 = moved one
                             stage towards the present (consecutive closed-loop steps: dmpc_solve_shift_batch callers),
                             0 = always start from the empty set.  Any guess leads to the same certified optimum (agreement
                             to the certificate tolerance, not bit for bit); a good one saves the active-set rounds.       */
} dmpc_options;

/* Per-QP status values mirror OSQP's status_val so `feasible` is derived as LPV_Planner.py:246-248. */
enum dmpc_status {
  DMPC_SOLVED = 1,
  DMPC_SOLVED_INACCURATE = 2,
  DMPC_MAX_ITER_REACHED = -2,
  DMPC_PRIMAL_INFEASIBLE = -3,
  DMPC_DUAL_INFEASIBLE = -4,
  DMPC_MAP_LOOKUP_FAILED = -20, /* curvature()/get_ey() find no segment (misc.py:97: int() of an empty index raises) */
  DMPC_DEGENERATE_PLANE = -21,  /* ego and neighbour coincide: compute_plane.py:52 divides by zero        */
  DMPC_UNSOLVED = -10           /* numerical breakdown (NaN in the condensed factorisation): no verdict on the QP         */
};

/* DMPC_ENGINE_STAGE: uncondensed QP, Riccati sweeps over the stages, one warp per QP, O(N) shared memory (qp3_*.cuh).
 * DMPC_ENGINE_FAST:  condensed QP, dense 2N x 2N factorisations, one 4-warp CTA per QP (qp2_*.cuh); horizons up to ~40. */
enum dmpc_engine { DMPC_ENGINE_AUTO = 0, DMPC_ENGINE_FAST = 2, DMPC_ENGINE_STAGE = 3 };

enum dmpc_error {
  DMPC_OK = 0,
  DMPC_ERR_ARG = -1,
  DMPC_ERR_CUDA = -2,
  DMPC_ERR_UNSUPPORTED = -3,
  DMPC_ERR_NOMEM = -4
};

/* One batch of B independent agent QPs (B = scenarios x local agents).  Layouts are the reference's numpy
 * layouts with a leading batch axis, C-contiguous FP64. */
typedef struct dmpc_batch_io {
  int32_t B;
  int32_t xprev_rows;        /* N+1 on the first DMPC iteration, N afterwards (LPV_HP_N_main.py:114-115) */
  /* inputs */
  const double* x0;          /* [B,9]                 solve(x0, ...)                                   */
  const double* xprev;       /* [B,xprev_rows,9]      Last_xPredicted                                  */
  const double* uprev;       /* [B,N,2]               uPred                                            */
  const double* uold;        /* [B,2]                 (OldSteering[0], OldAccelera[0]) :153            */
  const double* nbr_xy;      /* [B,N+1,nn,2]          x_agents                                         */
  const double* pose;        /* [B,N+1,2]             pose (ego predicted x,y)                          */
  const int32_t* nbr_count;  /* [B] or NULL (= nn for all); rows beyond the count are left out of the QP */
  /* outputs (any may be NULL) */
  double* z;                 /* [B,16N+12]            Solution = res.x                      :162        */
  double* planes;            /* [B,N,3,nn]            :138                                              */
  double* y_ineq;            /* [B,(8+nn)N]           multipliers of the rows of F (first rows of res.y) */
  double* y_eq;              /* [B,14N+12]            multipliers of the rows of G (adjoint recursion)   */
  double* obj;               /* [B]                   res.info.obj_val                                  */
  int32_t* status;           /* [B]                   res.info.status_val                               */
  int32_t* iters;            /* [B]                   ADMM iterations used                               */
  int32_t* pivots;           /* [B]                   active-set polish rounds used                      */
} dmpc_batch_io;

/* Buffers of the NEXT DMPC iteration, written by the solve's own epilogue (dmpc_solve_shift_batch): the caller's
 * bookkeeping of planner/scripts/LPV_HP_N_main.py:96-117 and LPV_Planner.py:179-180.  A pointer may alias the same
 * input of the call (x0 / xprev with xprev_rows == N / uprev / uold): a QP only ever touches its own rows. */
typedef struct dmpc_shift_io {
  double* x0;        /* [B,9]      <- xPred[1]                              LPV_HP_N_main.py:106 */
  double* xprev;     /* [B,N,9]    <- xPred[1:]                             :114-115             */
  double* uprev;     /* [B,N,2]    <- uPred, not shifted                    :113                 */
  double* uold;      /* [B,2]      <- uPred[0]                              LPV_Planner.py:179-180 */
  double* xy_send;   /* [B,N+1,2]  <- xPred[:,7:9], packed for the exchange LPV_HP_N_main.py:117 */
  int32_t* hold;     /* [B] in/out: consecutive failed solves of this QP.  feasible (LPV_Planner.py:246-248) -> 0 and the
                        buffers above receive this call's plan; otherwise +1 and the buffers keep the last valid plan (the
                        reference breaks its loop here, LPV_HP_N_main.py:102-104)                                         */
  /* optional (stage-wise engine): the reference's two stop conditions, per QP, for the caller's per-scenario stop */
  double* flags;     /* [B] or NULL: 1.0 * ended + 2.0 * failed.  ended = checkEnd (utilities/misc.py:28-48) on the state
                        this solve started from: s within 0.15 m (np.isclose) of end_s or beyond it, on lap `laps`;
                        failed = this solve was not feasible.  Laid out to travel with xy_send in the same all-gather.     */
  const int32_t* stop; /* [ceil(B / stop_div)] or NULL: QPs of a stopped scenario (stop[b / stop_div] != 0) are skipped: none
                        of their buffers, outputs or flags change (the experiment is over, LPV_HP_N_main.py:96,110-111)    */
  double end_s;      /* track length the end rule compares with (Map.TrackLength)                                          */
  int32_t laps;      /* lap number that ends the run (checkEnd's `laps`, 1 in the reference)                               */
  int32_t stop_div;  /* QPs per scenario in this batch (local agents)                                                      */
} dmpc_shift_io;

typedef struct dmpc_handle dmpc_handle;

void dmpc_default_options(dmpc_options* opt);
int dmpc_version(void);

/* device < 0: current device.  max_batch sizes the handle-owned workspaces. */
int dmpc_create(const dmpc_params* params, const dmpc_options* opt, int32_t max_batch, int32_t device,
                dmpc_handle** out);
void dmpc_destroy(dmpc_handle* h);
const char* dmpc_last_error(const dmpc_handle* h);
int dmpc_set_options(dmpc_handle* h, const dmpc_options* opt);

/* Replaces steps 2-9 of PlannerLPV.solve (LPV_Planner.py:138-180) for B QPs with DEVICE buffers.
 * `stream` is a cudaStream_t passed as void* (0 = legacy default stream).  Asynchronous. */
int dmpc_solve_batch(dmpc_handle* h, const dmpc_batch_io* io, void* stream);

/* dmpc_solve_batch + the caller's bookkeeping fused into the solve kernel's epilogue (all six pointers of `next`
 * required, DEVICE buffers): what dmpc_solve_batch followed by dmpc_shift_pack_hold leaves, in one launch. */
int dmpc_solve_shift_batch(dmpc_handle* h, const dmpc_batch_io* io, const dmpc_shift_io* next, void* stream);

/* Same call with HOST buffers (the end-to-end path).  Synchronous.  Pageable buffers: copied to page-locked staging,
 * H2D, solve, D2H.  If every buffer of the call is page-locked (cudaHostAlloc / cudaHostRegister) and
 * dmpc_options.host_direct = 1 (default) the kernel reads and writes them in place over PCIe - no staging copies,
 * same results bit for bit. */
int dmpc_solve_batch_host(dmpc_handle* h, const dmpc_batch_io* io);

/* Caller-side bookkeeping of one DMPC iteration (LPV_HP_N_main.py:106,113-117), on device:
 *   x0 <- xPred[1];  xprev <- xPred[1:] (N rows);  uprev <- uPred;  uold <- uPred[0];
 *   xy_send[b,t,:] = xPred[b,t,7:9]  (packed for the neighbour exchange).
 * z is [B,16N+12]; outputs may alias nothing in z. */
int dmpc_shift_pack(dmpc_handle* h, int32_t B, const double* d_z, double* d_x0, double* d_xprev /*[B,N,9]*/,
                    double* d_uprev, double* d_uold, double* d_xy_send /*[B,N+1,2]*/, void* stream);

/* dmpc_shift_pack with the reference's stop rule (LPV_HP_N_main.py:102-104): d_status [B] of the solve that produced
 * d_z, d_hold [B] in/out.  A QP whose status is feasible (1, 2, -2) is shifted and its hold reset to 0; any other
 * status leaves all five buffers of that QP untouched (last valid plan) and increments its hold.  Nothing non-finite
 * can reach the exchange through this call. */
int dmpc_shift_pack_hold(dmpc_handle* h, int32_t B, const double* d_z, const int32_t* d_status, double* d_x0,
                         double* d_xprev, double* d_uprev, double* d_uold, double* d_xy_send, int32_t* d_hold, void* stream);

/* After the exchange: gathers each agent's neighbour trajectories and its own pose from the all-gathered buffer.
 * xy_all has the layout ncclAllGather leaves behind when every rank sends its xy_send [S, A_local, N+1, 2]:
 *   xy_all [G, S, A_local, N+1, 2],  G = A_total / A_local ranks  (G = 1: plain [S, A_total, N+1, 2]),
 * global agent k of scenario s at rank k / A_local, slot k % A_local.  For local QP b = s * A_local + a_local
 * (global agent a = agent_offset + a_local):
 *   nbr_xy[b,t,j,:] = xy_all[agent nbr_idx[b,j] of scenario s, t, :],   pose[b,t,:] = xy_all[agent a of scenario s, t, :]
 * nbr_idx is [B, nn] int32 of global agent ids (agents_id of the reference, LPV_HP_N_main.py:82-85), -1 = empty slot.
 * Replaces  agents[:, ns[i], :] / agents[:, i, :]  of LPV_HP_N_main.py:100 and the ROS callback LPV_ROS_main.py:71-73. */
int dmpc_gather_neighbours(dmpc_handle* h, int32_t S, int32_t A_total, int32_t A_local, int32_t agent_offset,
                           const double* d_xy_all, const int32_t* d_nbr_idx, double* d_nbr_xy, double* d_pose,
                           void* stream);

/* dmpc_gather_neighbours on a payload that carries the stop flags: every rank sends
 *   [ xy_send (S * A_local * (N+1) * 2 doubles) | flags (S * A_local doubles, dmpc_shift_io.flags) ],
 * so d_all is [G][S * A_local * ((N+1) * 2 + 1)].  Besides the gather, d_stop[s] (int32 [S], in/out, sticky) is set for
 * every scenario with ANY agent on ANY rank ended or failed - the reference's loop condition and break
 * (LPV_HP_N_main.py:96,102-104,110-111; checkEnd returns at the first agent that is at the end).  Every rank computes
 * the same d_stop from the same gathered buffer: no extra collective, no rank can leave the loop alone. */
int dmpc_gather_neighbours_stop(dmpc_handle* h, int32_t S, int32_t A_total, int32_t A_local, int32_t agent_offset,
                                const double* d_all, const int32_t* d_nbr_idx, double* d_nbr_xy, double* d_pose,
                                int32_t* d_stop, void* stream);

/* Batched LPV model roll-out, SURVEY.md 8(f2).  Replaces LPV_Model.sim (LPV_val.py:77-94) with that file's own
 * linearisation (LPV_val.py:96-195: no low-velocity branch) for B independent models, DEVICE buffers:
 *   d_xpred [B,N+1,9] = each model's xPred before the call, d_u [B,N,2];
 *   d_out   [B,N+1,9] = [x_1 .. x_N, ones(9)] (the value sim() returns and stores in xPred); may not alias d_xpred;
 *   d_status [B] (may be NULL): 0, or DMPC_MAP_LOOKUP_FAILED with NaN rows where the reference raises.
 * Uses the handle's N, dt, map and vehicle parameters (sys_lim is not used by sim).  Asynchronous. */
int dmpc_sim_batch(dmpc_handle* h, int32_t B, const double* d_xpred, const double* d_u, double* d_out,
                   int32_t* d_status, void* stream);

/* Dual update of the reference's OCD loop, SURVEY.md 8(f4).  Replaces the triple loop of NL_EU_N_main.py:122-138 with
 * eval_constraintEU (plan_lib/config/NL/config.py:19-23) for S scenarios of A agents, in place, DEVICE buffers:
 *   d_lambdas[s,i,j,k-1] += alpha * (dth - || d_xy[s,i,k,:] - d_xy[s,j,k,:] ||)   for i < j, k = 1..N;
 * entries with i >= j are left as they are (the reference adds 0 there).  d_xy [S,A,N+1,2] is the layout
 * dmpc_shift_pack / the all-gather leave behind, d_lambdas [S,A,A,N].  alpha = get_alpha() = 0.25 in the reference.
 * Bit-exact with the reference's numpy arithmetic.  Asynchronous. */
int dmpc_ocd_lambda_update(dmpc_handle* h, int32_t S, int32_t A, double alpha, double dth, const double* d_xy,
                           double* d_lambdas, void* stream);

/* Hyperplane form of the OCD dual update.  Replaces the triple loop of NL_HP_N_main.py:128-137 with eval_constraintHp
 * (plan_lib/config/NL/config.py:25-29), in place, DEVICE buffers:
 *   d_lambdas[s,i,j,k-1] += alpha * (-a_x * x - a_y * y - b + dth/2),  (a_x,a_y,b) = d_planes[s,i,k-1,:,j],
 *   (x,y) = d_xy[s,j,k,:],  for i < j, k = 1..N;  entries with i >= j are left as they are.
 * d_planes [S,A,N,3,A] stacks the planes each agent's solve returned (planes_raw[i][k-1,:,j], j = absolute agent
 * index).  Bit-exact with the reference's numpy arithmetic.  Asynchronous. */
int dmpc_ocd_lambda_update_hp(dmpc_handle* h, int32_t S, int32_t A, double alpha, double dth, const double* d_xy,
                              const double* d_planes, double* d_lambdas, void* stream);

/* Allocation-free variant of dmpc_solve_batch.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Needs no handle: params / options (NULL = defaults) are passed per call, `device` < 0 = current device,
 * DEVICE buffers in `io`, asynchronous on `stream`.  Stage-wise engine only, nothing carried between calls (no launch
 * order, no warm start): results are those of a handle with warm_start = 0, order_hint = 0, bit for bit.
 * The kernel's global scratch (the parked interior-point iterate of each QP) is a static device array of
 * DMPC_STATIC_SCRATCH_DOUBLES doubles (24 MiB) that every GPU loading the library holds until the process exits:
 *   io->B * (14 N + 2 N (8 + nn), rounded up to even) doubles must fit - 4626 QPs at N = 20, nn = 2, 925 at N = 100, nn = 2 -
 * else DMPC_ERR_NOMEM with the limit spelt out in `err` (caller's buffer of err_len bytes, may be NULL).
 * The call makes no cudaMalloc / cudaFree / allocating library call.  Every call shares the array, which keeps the last
 * call's contents (the kernel writes a QP's slice before reading it: nothing needs clearing); calls that overlap on the
 * GPU race on it - keep them stream-ordered.  The first call on a device resolves the array's address
 * (cudaGetSymbolAddress) and is not thread-safe. */
#define DMPC_STATIC_SCRATCH_DOUBLES (3 * 1024 * 1024)
int dmpc_solve_batch_static(const dmpc_params* params, const dmpc_options* opt, const dmpc_batch_io* io, int32_t device,
                            void* stream, char* err, int32_t err_len);

/* Engine the handle runs on: DMPC_ENGINE_STAGE or DMPC_ENGINE_FAST (what DMPC_ENGINE_AUTO resolved to). */
int dmpc_engine(const dmpc_handle* h);

/* Dynamic shared memory one QP's CTA asks for on the handle's engine (the whole per-QP workspace lives there). */
int64_t dmpc_smem_bytes(const dmpc_handle* h);

/* Number of kernels the handle has launched so far (bench.py's gpu_launches claim). */
int64_t dmpc_launch_count(const dmpc_handle* h);

#ifdef __cplusplus
}
#endif
#endif /* DMPC_B200_H */

// dmpc_api.cu - C ABI (include/dmpc_b200.h) over the sm_100a kernels.  No torch types, no CPU compute path:
// every numeric result comes from lpv_qp_stage_kernel / lpv_qp_fast_kernel / the bookkeeping kernels below.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

#include "qp2_kernel.cuh"
#include "qp3_kernel.cuh"
#include "lpv_sim.cuh"
#include "qp_params.h"

struct dmpc_handle {
  dmpc_params params;
  dmpc_options opt;
  DevParams P;
  int device, max_batch, smem_limit;
  int fast;  // 1: lpv_qp_fast_kernel (condensed, qp2_*.cuh), 0: lpv_qp_stage_kernel (stage-wise, qp3_*.cuh)
  int stage_ok;            // the stage-wise workspace fits in shared memory (always required when !fast)
  size_t smem_bytes;       // fast engine workspace
  size_t stage_smem;       // stage-wise workspace
  long long gws_stride;    // doubles per QP of the stage-wise engine's global scratch (parked interior-point iterate)
  double* d_gws;
  int *d_cost, *d_order;  // fast engine: per-QP cost of the last call and the launch order derived from it
  int order_B;            // batch size d_order is valid for (0: none yet)
  unsigned char* d_warm;  // stage-wise engine: certified working set per QP of the last call ([max_batch][rows] bytes)
  int warm_B;             // batch size d_warm belongs to (0: none yet)
  int nsm;                // SMs of the device
  size_t smem_alone;      // dynamic shared memory request that keeps a CTA alone on its SM
  cudaStream_t host_stream;  // dmpc_solve_batch_host
  int sim_attr;              // lpv_sim_kernel's shared-memory opt-in done (long horizons)
  // staging for the host-buffer entry point
  char* d_stage;
  size_t stage_bytes;
  int64_t launches;
  char err[256];
};

static int fail(dmpc_handle* h, int code, const char* fmt, const char* a = "") {
  if (h) snprintf(h->err, sizeof(h->err), fmt, a);
  return code;
}
#define CUDA_TRY(h, call)                                                         \
  do {                                                                            \
    cudaError_t e_ = (call);                                                      \
    if (e_ != cudaSuccess) return fail((h), DMPC_ERR_CUDA, "CUDA: %s", cudaGetErrorString(e_)); \
  } while (0)

extern "C" int dmpc_version(void) { return 100; }

extern "C" void dmpc_default_options(dmpc_options* o) {
  if (!o) return;
  o->polish_delta = 1e-9; o->cert_tol = 1e-9; o->ipm_mu0 = 1e-6;
  o->polish_refine = 4; o->pdas_first = 4; o->pdas_after = 2; o->ipm_max_iter = 60;
  o->engine = DMPC_ENGINE_AUTO; o->order_hint = 1; o->host_direct = 1; o->warm_start = 1;
}

static void apply_options(dmpc_handle* h) {
  h->P.delta = h->opt.polish_delta; h->P.cert_tol = h->opt.cert_tol; h->P.ipm_mu0 = h->opt.ipm_mu0;
  h->P.refine = h->opt.polish_refine; h->P.pdas_first = h->opt.pdas_first; h->P.pdas_after = h->opt.pdas_after;
  h->P.ipm_max = h->opt.ipm_max_iter;
}

static bool options_ok(const dmpc_options* o) {
  return o->polish_delta > 0 && o->polish_delta < 1e-3 && o->cert_tol > 0 && o->ipm_mu0 > 0 && o->polish_refine >= 0 &&
         o->polish_refine <= 32 && o->pdas_first >= 0 && o->pdas_after >= 1 && o->ipm_max_iter >= 1 && (o->engine == DMPC_ENGINE_AUTO || o->engine == DMPC_ENGINE_FAST || o->engine == DMPC_ENGINE_STAGE) &&
         o->warm_start >= 0 && o->warm_start <= 2;
}

static bool psd(const double* M, int n, bool strict) {  // LDL' without pivoting on the symmetrised matrix
  double a[81];
  double scale = 0;
  for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) { a[i * n + j] = sym_up(M, n, i, j); scale = fmax(scale, fabs(a[i * n + j])); }
  if (!(scale < 1e300)) return false;
  const double tol = 1e-12 * (scale > 0 ? scale : 1.0);
  for (int k = 0; k < n; ++k) {
    const double d = a[k * n + k];
    if (strict ? !(d > tol) : !(d >= -tol)) return false;
    if (d <= tol) {  // zero pivot: the rest of the column must vanish
      for (int i = k + 1; i < n; ++i) if (fabs(a[i * n + k]) > 1e-9 * (scale + 1)) return false;
      continue;
    }
    for (int i = k + 1; i < n; ++i) {
      const double l = a[i * n + k] / d;
      for (int j = k + 1; j < n; ++j) a[i * n + j] -= l * a[k * n + j];
    }
  }
  return true;
}

extern "C" int dmpc_create(const dmpc_params* p, const dmpc_options* opt, int32_t max_batch, int32_t device, dmpc_handle** out) {
  if (!p || !out) return DMPC_ERR_ARG;
  *out = nullptr;
  dmpc_handle* h = new (std::nothrow) dmpc_handle();
  if (!h) return DMPC_ERR_NOMEM;
  memset(h, 0, sizeof(*h));
  *out = h;  // returned even on failure so that dmpc_last_error() can be read; caller destroys it
  if (p->N < 1 || p->N > 512 || p->nn < 0 || p->nn > 255 || p->nseg < 1 || p->nseg > DMPC_MAX_SEG || !(p->dt > 0) || max_batch < 1)
    return fail(h, DMPC_ERR_ARG, "bad N/nn/nseg/dt/max_batch%s");
  if (!psd(p->Q, 9, false) || !psd(p->R, 2, false) || !psd(p->dR, 2, false))
    return fail(h, DMPC_ERR_ARG, "Q, R and dR must be positive semidefinite%s");
  if (!psd(p->Qs, 3, true)) return fail(h, DMPC_ERR_ARG, "Qs must be positive definite%s");
  {
    double rs[4];
    for (int i = 0; i < 4; ++i) rs[i] = p->R[i] + p->dR[i];
    if (!psd(rs, 2, true)) return fail(h, DMPC_ERR_ARG, "R + dR must be positive definite (strictly convex inputs)%s");
  }
  h->params = *p;
  if (opt) h->opt = *opt; else dmpc_default_options(&h->opt);
  if (!options_ok(&h->opt)) return fail(h, DMPC_ERR_ARG, "bad options%s");
  if (device < 0) CUDA_TRY(h, cudaGetDevice(&device));
  h->device = device;
  CUDA_TRY(h, cudaSetDevice(device));
  cudaDeviceProp prop;
  CUDA_TRY(h, cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) return fail(h, DMPC_ERR_UNSUPPORTED, "this library is built for sm_100a (B200) only%s");
  h->max_batch = max_batch;

  DevParams& P = h->P;
  fill_dev_params(p, &h->opt, P);

  int optin = 0;
  CUDA_TRY(h, cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
  h->smem_limit = optin;
  CUDA_TRY(h, cudaDeviceGetAttribute(&h->nsm, cudaDevAttrMultiProcessorCount, device));
  // stage-wise engine: O(N) workspace per QP in shared memory, the parked interior-point iterate in a global slice
  {
    Ctx3 c3;
    const size_t sb3 = carve_workspace3(c3, nullptr, P.N, P.nn);
    h->stage_ok = sb3 + 1024 <= (size_t)optin && P.N <= 511 && P.nn <= 123;  // rinfo packs (stage << 7) | row in 16 bits
    h->stage_smem = sb3;
  }
  bool fast_fits = false;
  size_t sb2 = 0;
  {  // condensed engine: the whole packed workspace in shared memory (qp2_engine.cuh)
    Ctx2 c2;
    sb2 = FAST_HDR_BYTES + carve_workspace2(c2, nullptr, P.N, P.nn, P.nrows);
    fast_fits = sb2 + 1024 <= (size_t)optin && 2 * P.N <= 255 && P.N <= 4095;
  }
  if (h->opt.engine == DMPC_ENGINE_FAST && !fast_fits)
    return fail(h, DMPC_ERR_UNSUPPORTED, "horizon too long for the condensed engine%s");
  h->fast = h->opt.engine == DMPC_ENGINE_FAST;
  if (!h->fast && !h->stage_ok)
    return fail(h, DMPC_ERR_UNSUPPORTED, "horizon/neighbour count too large for the on-chip workspace%s");
  if (h->stage_ok) {
    h->gws_stride = (long long)((stage_park_doubles(P.N, P.nn) + 1) & ~(size_t)1);
    CUDA_TRY(h, cudaMalloc(&h->d_gws, (size_t)h->gws_stride * sizeof(double) * max_batch));
    CUDA_TRY(h, cudaFuncSetAttribute(lpv_qp_stage_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->stage_smem));
    CUDA_TRY(h, cudaMalloc(&h->d_warm, (size_t)max_batch * (size_t)(P.N * (8 + P.nn))));
  }
  CUDA_TRY(h, cudaMalloc(&h->d_cost, sizeof(int) * (size_t)max_batch));
  CUDA_TRY(h, cudaMalloc(&h->d_order, sizeof(int) * ((size_t)max_batch + 1)));  // [max_batch]: expensive-QP count
  if (h->fast) {
    h->smem_bytes = sb2;
    // alone on the SM: a second CTA of this kernel (sb2 + 1 KB system share each) must not fit beside it
    const size_t per_sm = (size_t)optin + 1024;
    h->smem_alone = per_sm > 2 * (sb2 + 1024) ? per_sm - (sb2 + 1024) - 1024 + 16 : sb2;
    if (h->smem_alone > (size_t)optin) h->smem_alone = (size_t)optin;
    if (h->smem_alone < sb2) h->smem_alone = sb2;
    CUDA_TRY(h, cudaFuncSetAttribute(lpv_qp_fast_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem_alone));
  }
  return DMPC_OK;
}

extern "C" void dmpc_destroy(dmpc_handle* h) {
  if (!h) return;
  if (h->d_gws) cudaFree(h->d_gws);
  if (h->d_cost) cudaFree(h->d_cost);
  if (h->d_order) cudaFree(h->d_order);
  if (h->d_warm) cudaFree(h->d_warm);
  if (h->host_stream) cudaStreamDestroy(h->host_stream);
  if (h->d_stage) cudaFree(h->d_stage);
  delete h;
}

extern "C" const char* dmpc_last_error(const dmpc_handle* h) { return h ? h->err : "null handle"; }

extern "C" int dmpc_set_options(dmpc_handle* h, const dmpc_options* opt) {
  if (!h || !opt) return DMPC_ERR_ARG;
  if (!options_ok(opt)) return fail(h, DMPC_ERR_ARG, "bad options%s");
  if (opt->engine != h->opt.engine) return fail(h, DMPC_ERR_ARG, "the engine is chosen at dmpc_create%s");
  h->opt = *opt;
  apply_options(h);
  return DMPC_OK;
}

extern "C" int dmpc_engine(const dmpc_handle* h) { return h ? (h->fast ? DMPC_ENGINE_FAST : DMPC_ENGINE_STAGE) : 0; }

extern "C" int64_t dmpc_smem_bytes(const dmpc_handle* h) { return h ? (int64_t)(h->fast ? h->smem_bytes : h->stage_smem) : 0; }

extern "C" int64_t dmpc_launch_count(const dmpc_handle* h) { return h ? h->launches : 0; }

static int launch_shift_pack(dmpc_handle* h, int32_t B, const double* d_z, const int32_t* d_status, double* d_x0,
                             double* d_xprev, double* d_uprev, double* d_uold, double* d_xy_send, int32_t* d_hold, void* stream);

// One batch: the solve and, with `next`, the caller's bookkeeping.  Stage-wise engine: one launch does both.
// Condensed engine: its (split) launch, then the stage-wise kernel over the QPs it reported DMPC_UNSOLVED (numerical
// breakdown of the condensed factorisation - crawling vehicles), then the shift kernel.
// Input staging of the stage-wise kernel: bulk async copies (TMA) into shared memory unless DMPC_B200_BULK_IN=0
// (development switch for A/B measurements; results are identical either way).
static int bulk_in_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("DMPC_B200_BULK_IN"); v = !(e && e[0] == '0'); }
  return v;
}

static int solve_impl(dmpc_handle* h, const dmpc_batch_io* io, const dmpc_shift_io* next, void* stream) {
  if (!h || !io) return DMPC_ERR_ARG;
  if (io->B == 0) return DMPC_OK;
  if (io->B < 0 || io->B > h->max_batch) return fail(h, DMPC_ERR_ARG, "batch larger than max_batch%s");
  if (io->xprev_rows < h->P.N) return fail(h, DMPC_ERR_ARG, "Last_xPredicted needs at least N rows%s");
  if (!io->x0 || !io->xprev || !io->uprev || !io->uold || !io->pose || (h->P.nn > 0 && !io->nbr_xy))
    return fail(h, DMPC_ERR_ARG, "null input pointer%s");
  if (next && (!next->x0 || !next->xprev || !next->uprev || !next->uold || !next->xy_send || !next->hold))
    return fail(h, DMPC_ERR_ARG, "dmpc_shift_io: all six buffers are required%s");
  if (next && (!io->z || !io->status) && h->fast) return fail(h, DMPC_ERR_ARG, "the condensed engine shifts from z and status%s");
  if (next && (next->flags || next->stop) && h->fast) return fail(h, DMPC_ERR_UNSUPPORTED, "stop flags need the stage-wise engine%s");
  CUDA_TRY(h, cudaSetDevice(h->device));
  cudaStream_t st = (cudaStream_t)stream;
  BatchPtrs bp;
  memset(&bp, 0, sizeof(bp));
  bp.B = io->B; bp.xprev_rows = io->xprev_rows;
  bp.x0 = io->x0; bp.xprev = io->xprev; bp.uprev = io->uprev; bp.uold = io->uold; bp.nbr_xy = io->nbr_xy; bp.pose = io->pose;
  bp.nbr_count = io->nbr_count;
  bp.z = io->z; bp.planes = io->planes; bp.y_ineq = io->y_ineq; bp.y_eq = io->y_eq; bp.obj = io->obj;
  bp.status = io->status; bp.iters = io->iters; bp.pivots = io->pivots;
  bp.gws = h->d_gws; bp.gws_stride = h->gws_stride;
  if (!h->fast) {
    if (next) {
      bp.nx_x0 = next->x0; bp.nx_xprev = next->xprev; bp.nx_uprev = next->uprev; bp.nx_uold = next->uold; bp.nx_xy = next->xy_send; bp.hold = next->hold;
      bp.flags = next->flags; bp.stop = next->stop; bp.stop_div = next->stop_div > 0 ? next->stop_div : 1;
      bp.laps = next->laps; bp.end_s = next->end_s;
      if (next->flags && !(next->end_s > 0.0)) return fail(h, DMPC_ERR_ARG, "dmpc_shift_io.flags needs end_s > 0%s");
    }
    // cost-ordered launch (dmpc_options.order_hint): CTA i solves the QP that was i-th most expensive in the previous
    // call of the same batch size - consecutive MPC steps of a scenario are alike - so the interior-point QPs (8x the mean
    // work) start in the first wave instead of being the tail of the grid.  Scheduling only; results do not change.
    const bool hint = h->opt.order_hint != 0;
    if (hint) { bp.cost = h->d_cost; if (h->order_B == io->B) bp.order = h->d_order; }
    if (h->opt.warm_start) {  // the previous call's certified working sets are this call's first guesses
      if (h->warm_B != io->B) {
        CUDA_TRY(h, cudaMemsetAsync(h->d_warm, 0, (size_t)io->B * (size_t)(h->P.N * (8 + h->P.nn)), st));
        h->warm_B = io->B;
      }
      bp.warm = h->d_warm; bp.warm_shift = h->opt.warm_start == 2;
    }
    bp.bulk_in = bulk_in_enabled();
    lpv_qp_stage_kernel<<<io->B, QP3_THREADS, h->stage_smem, st>>>(h->P, bp);
    h->launches++;
    if (hint) {
      order_by_cost_kernel<<<1, ORDER_THREADS, 0, st>>>(h->d_cost, h->d_order, h->d_order + h->max_batch, io->B, h->nsm);
      h->order_B = io->B;
      h->launches++;
    }
    CUDA_TRY(h, cudaGetLastError());
    return DMPC_OK;
  }
  {
    // cost-ordered launch (dmpc_options.order_hint): this call starts its QPs in decreasing order of what the previous
    // call of the same batch size spent on them - consecutive MPC steps of a scenario are alike, so the few
    // interior-point QPs start first instead of becoming the tail of the grid.  Scheduling only; results do not change.
    const bool hint = h->opt.order_hint != 0;
    if (hint) { bp.cost = h->d_cost; if (h->order_B == io->B) bp.order = h->d_order; }
    if (bp.order && h->smem_alone > h->smem_bytes) {
      // The QPs that needed the interior-point phase last time (order[0 .. nhard), ~1 %) are launched first with a
      // shared-memory request that keeps each of them alone on an SM; the bulk follows in the same stream as a
      // programmatic dependent launch: it may start as soon as every CTA of the first part is resident (they signal at
      // entry), so it fills the other SMs.  Same kernel, same arithmetic per QP.
      bp.nhard = h->d_order + h->max_batch;
      bp.part = 1;
      lpv_qp_fast_kernel<<<io->B < h->nsm ? io->B : h->nsm, QP_THREADS, h->smem_alone, st>>>(h->P, bp);
      bp.part = 2;
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3((unsigned)io->B); cfg.blockDim = dim3(QP_THREADS);
      cfg.dynamicSmemBytes = h->smem_bytes; cfg.stream = st;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      at[0].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = at; cfg.numAttrs = 1;
      CUDA_TRY(h, cudaLaunchKernelEx(&cfg, lpv_qp_fast_kernel, h->P, bp));
      h->launches++;
    } else {
      lpv_qp_fast_kernel<<<io->B, QP_THREADS, h->smem_bytes, st>>>(h->P, bp);
    }
    h->launches++;
    if (hint) {
      order_by_cost_kernel<<<1, ORDER_THREADS, 0, st>>>(h->d_cost, h->d_order, h->d_order + h->max_batch, io->B, h->nsm);
      h->order_B = io->B;
      h->launches++;
    }
    if (h->stage_ok && io->status) {  // what the condensed factorisation could not handle is solved stage-wise
      bp.order = nullptr; bp.cost = nullptr; bp.nhard = nullptr; bp.part = 0;
      bp.only_status = DMPC_UNSOLVED;
      bp.bulk_in = bulk_in_enabled();
      lpv_qp_stage_kernel<<<io->B, QP3_THREADS, h->stage_smem, st>>>(h->P, bp);
      h->launches++;
    }
  }
  CUDA_TRY(h, cudaGetLastError());
  if (next)
    return launch_shift_pack(h, io->B, io->z, io->status, next->x0, next->xprev, next->uprev, next->uold, next->xy_send,
                             next->hold, stream);
  return DMPC_OK;
}

extern "C" int dmpc_solve_batch(dmpc_handle* h, const dmpc_batch_io* io, void* stream) { return solve_impl(h, io, nullptr, stream); }

extern "C" int dmpc_solve_shift_batch(dmpc_handle* h, const dmpc_batch_io* io, const dmpc_shift_io* next, void* stream) {
  if (!next) return h ? fail(h, DMPC_ERR_ARG, "null dmpc_shift_io%s") : DMPC_ERR_ARG;
  return solve_impl(h, io, next, stream);
}

// ---------------------------------------------------------------------------------------------------
// Host-buffer entry point: H2D of the inputs, solve, D2H of the outputs, all inside the call.
// Device view of a page-locked host range (cudaHostAlloc / cudaHostRegister memory is mapped into the device's address
// space under unified addressing); false for pageable memory, device memory, or a range that leaves its allocation.
static bool pinned_view(const void* p, size_t bytes, void** dev) {
  *dev = nullptr;
  if (!p || !bytes) return true;
  cudaPointerAttributes a, e;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess || cudaPointerGetAttributes(&e, (const char*)p + bytes - 1) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  if (a.type != cudaMemoryTypeHost || e.type != cudaMemoryTypeHost || !a.devicePointer || !e.devicePointer) return false;
  if ((const char*)e.devicePointer - (const char*)a.devicePointer != (ptrdiff_t)(bytes - 1)) return false;
  *dev = a.devicePointer;
  return true;
}

extern "C" int dmpc_solve_batch_host(dmpc_handle* h, const dmpc_batch_io* io) {
  if (!h || !io) return DMPC_ERR_ARG;
  if (io->B == 0) return DMPC_OK;
  if (io->B < 0 || io->B > h->max_batch) return fail(h, DMPC_ERR_ARG, "batch larger than max_batch%s");
  const int N = h->P.N, nn = h->P.nn, B = io->B, rows = io->xprev_rows;
  if (rows < N) return fail(h, DMPC_ERR_ARG, "Last_xPredicted needs at least N rows%s");
  const size_t n = 16 * (size_t)N + 12, mi = (8 + (size_t)nn) * N, me = 14 * (size_t)N + 12;
  if (h->opt.host_direct && rows <= N + 1) {
    // Page-locked host buffers: the solve kernel reads its ~3 KB of inputs per QP and writes its results straight
    // through the PCIe mapping, so the transfers of one QP overlap the arithmetic of the others and no copy sits in
    // front of or behind the launch.  Any pageable / unmapped buffer sends the whole call down the staging path below.
    CUDA_TRY(h, cudaSetDevice(h->device));
    dmpc_batch_io dio = *io;
    void* v[15];
    const size_t Bz = (size_t)B;
    bool ok = pinned_view(io->x0, Bz * 9 * 8, &v[0]) && pinned_view(io->xprev, Bz * rows * 9 * 8, &v[1]) &&
              pinned_view(io->uprev, Bz * N * 2 * 8, &v[2]) && pinned_view(io->uold, Bz * 2 * 8, &v[3]) &&
              pinned_view(io->nbr_xy, Bz * (N + 1) * nn * 2 * 8, &v[4]) && pinned_view(io->pose, Bz * (N + 1) * 2 * 8, &v[5]) &&
              pinned_view(io->nbr_count, Bz * 4, &v[6]) && pinned_view(io->z, Bz * n * 8, &v[7]) &&
              pinned_view(io->planes, Bz * N * 3 * nn * 8, &v[8]) && pinned_view(io->y_ineq, Bz * mi * 8, &v[9]) &&
              pinned_view(io->y_eq, Bz * me * 8, &v[10]) && pinned_view(io->obj, Bz * 8, &v[11]) &&
              pinned_view(io->status, Bz * 4, &v[12]) && pinned_view(io->iters, Bz * 4, &v[13]) &&
              pinned_view(io->pivots, Bz * 4, &v[14]);
    if (ok) {
      if (!h->host_stream) CUDA_TRY(h, cudaStreamCreateWithFlags(&h->host_stream, cudaStreamNonBlocking));
      dio.x0 = (const double*)v[0]; dio.xprev = (const double*)v[1]; dio.uprev = (const double*)v[2];
      dio.uold = (const double*)v[3]; dio.nbr_xy = (const double*)v[4]; dio.pose = (const double*)v[5];
      dio.nbr_count = (const int32_t*)v[6]; dio.z = (double*)v[7]; dio.planes = (double*)v[8]; dio.y_ineq = (double*)v[9];
      dio.y_eq = (double*)v[10]; dio.obj = (double*)v[11]; dio.status = (int32_t*)v[12]; dio.iters = (int32_t*)v[13];
      dio.pivots = (int32_t*)v[14];
      const int rc = dmpc_solve_batch(h, &dio, (void*)h->host_stream);
      if (rc) return rc;
      CUDA_TRY(h, cudaStreamSynchronize(h->host_stream));
      return DMPC_OK;
    }
  }
  const size_t Bm = (size_t)h->max_batch;
  // fixed staging layout sized for max_batch (offsets in bytes, 256-aligned)
  size_t off = 0;
  auto seg = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
  const size_t o_x0 = seg(Bm * 9 * 8), o_xp = seg(Bm * (N + 1) * 9 * 8), o_up = seg(Bm * N * 2 * 8), o_uo = seg(Bm * 2 * 8);
  const size_t o_nb = seg(Bm * (N + 1) * (nn > 0 ? nn : 1) * 2 * 8), o_po = seg(Bm * (N + 1) * 2 * 8), o_nc = seg(Bm * 4);
  const size_t o_z = seg(Bm * n * 8), o_pl = seg(Bm * N * 3 * (nn > 0 ? nn : 1) * 8), o_yi = seg(Bm * mi * 8), o_ye = seg(Bm * me * 8);
  const size_t o_ob = seg(Bm * 8), o_st = seg(Bm * 4), o_it = seg(Bm * 4), o_pv = seg(Bm * 4);
  if (!h->d_stage) {
    CUDA_TRY(h, cudaSetDevice(h->device));
    CUDA_TRY(h, cudaMalloc(&h->d_stage, off));
    h->stage_bytes = off;
  }
  if (rows > N + 1) return fail(h, DMPC_ERR_ARG, "Last_xPredicted has more than N+1 rows%s");
  char* d = h->d_stage;
  if (!h->host_stream) CUDA_TRY(h, cudaStreamCreateWithFlags(&h->host_stream, cudaStreamNonBlocking));
  cudaStream_t s = h->host_stream;  // the call is synchronous and touches host buffers only: a private stream
#define H2D(dst_off, src, bytes) do { if (src) CUDA_TRY(h, cudaMemcpyAsync(d + (dst_off), (src), (bytes), cudaMemcpyHostToDevice, s)); } while (0)
#define D2H(dst, src_off, bytes) do { if (dst) CUDA_TRY(h, cudaMemcpyAsync((dst), d + (src_off), (bytes), cudaMemcpyDeviceToHost, s)); } while (0)
  H2D(o_x0, io->x0, (size_t)B * 9 * 8); H2D(o_xp, io->xprev, (size_t)B * rows * 9 * 8); H2D(o_up, io->uprev, (size_t)B * N * 2 * 8);
  H2D(o_uo, io->uold, (size_t)B * 2 * 8); H2D(o_nb, io->nbr_xy, (size_t)B * (N + 1) * nn * 2 * 8); H2D(o_po, io->pose, (size_t)B * (N + 1) * 2 * 8);
  H2D(o_nc, io->nbr_count, (size_t)B * 4);
  dmpc_batch_io dio = *io;
  dio.x0 = (const double*)(d + o_x0); dio.xprev = (const double*)(d + o_xp); dio.uprev = (const double*)(d + o_up);
  dio.uold = (const double*)(d + o_uo); dio.nbr_xy = (const double*)(d + o_nb); dio.pose = (const double*)(d + o_po);
  dio.nbr_count = io->nbr_count ? (const int32_t*)(d + o_nc) : nullptr;
  dio.z = io->z ? (double*)(d + o_z) : nullptr; dio.planes = io->planes ? (double*)(d + o_pl) : nullptr;
  dio.y_ineq = io->y_ineq ? (double*)(d + o_yi) : nullptr; dio.y_eq = io->y_eq ? (double*)(d + o_ye) : nullptr;
  dio.obj = io->obj ? (double*)(d + o_ob) : nullptr; dio.status = io->status ? (int32_t*)(d + o_st) : nullptr;
  dio.iters = io->iters ? (int32_t*)(d + o_it) : nullptr; dio.pivots = io->pivots ? (int32_t*)(d + o_pv) : nullptr;
  const int rc = dmpc_solve_batch(h, &dio, (void*)s);
  if (rc) return rc;
  D2H(io->z, o_z, (size_t)B * n * 8); D2H(io->planes, o_pl, (size_t)B * N * 3 * nn * 8); D2H(io->y_ineq, o_yi, (size_t)B * mi * 8);
  D2H(io->y_eq, o_ye, (size_t)B * me * 8); D2H(io->obj, o_ob, (size_t)B * 8); D2H(io->status, o_st, (size_t)B * 4);
  D2H(io->iters, o_it, (size_t)B * 4); D2H(io->pivots, o_pv, (size_t)B * 4);
#undef H2D
#undef D2H
  CUDA_TRY(h, cudaStreamSynchronize(s));
  return DMPC_OK;
}

// ---------------------------------------------------------------------------------------------------
// DMPC bookkeeping kernels (planner/scripts/LPV_HP_N_main.py:106,113-117; LPV_Planner.py:179-180).
// status != nullptr: the reference's stop rule (LPV_HP_N_main.py:102-104) - a QP whose solve was not feasible
// (LPV_Planner.py:246-248) keeps all its buffers (last valid plan) and its hold counter goes up; feasible resets it.
__global__ void shift_pack_kernel(int B, int N, const double* __restrict__ z, const int* __restrict__ status,
                                  double* __restrict__ x0, double* __restrict__ xprev, double* __restrict__ uprev,
                                  double* __restrict__ uold, double* __restrict__ xy, int* __restrict__ hold) {
  const int n = 16 * N + 12;
  const long long total = (long long)B * n;
  for (long long w = blockIdx.x * (long long)blockDim.x + threadIdx.x; w < total; w += (long long)gridDim.x * blockDim.x) {
    const int b = (int)(w / n), j = (int)(w % n);
    if (status) {
      const int st = status[b];
      const bool feas = st == DMPC_SOLVED || st == DMPC_SOLVED_INACCURATE || st == DMPC_MAX_ITER_REACHED;
      if (j == 0 && hold) hold[b] = feas ? 0 : hold[b] + 1;
      if (!feas) continue;
    }
    const double v = z[w];
    if (j < 12 * (N + 1)) {
      const int t = j / 12, r = j % 12;
      if (r >= 9) continue;
      if (t == 1 && x0) x0[(size_t)b * 9 + r] = v;                                  // r.x0 = x_pred[i][1,:]
      if (t >= 1 && xprev) xprev[((size_t)b * N + (t - 1)) * 9 + r] = v;            // x_old = x_pred[1:,:]
      if (r >= 7 && xy) xy[((size_t)b * (N + 1) + t) * 2 + (r - 7)] = v;            // agents[t,i,:] = xPred[t,7:9]
    } else if (j < 12 * (N + 1) + 2 * N) {
      const int q = j - 12 * (N + 1);
      if (uprev) uprev[(size_t)b * 2 * N + q] = v;                                  // u_old = u_pred (not shifted)
      if (q < 2 && uold) uold[(size_t)b * 2 + q] = v;                               // OldSteering / OldAccelera
    }
  }
}

// ---------------------------------------------------------------------------------------------------
// OCD dual update: planner/scripts/NL_EU_N_main.py:122-138 with eval_constraintEU
// (plan_lib/config/NL/config.py:19-23):  lambdas[s,i,j,k-1] += alpha * (dth - ||p_i,k - p_j,k||), i < j, k = 1..N.
// One thread per (s, i, j, k) of the upper triangle; consecutive threads walk k, so the lambda row (N doubles) and the
// two position rows are read and written as contiguous segments.  Roundings as numpy's: no fma contraction.
__global__ void ocd_lambda_kernel(int S, int A, int N, double alpha, double dth, const double* __restrict__ xy,
                                  double* __restrict__ lambdas) {
  const long long npair = (long long)A * (A - 1) / 2;
  const long long total = (long long)S * npair * N;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(e % N);
    long long r = e / N;
    const int pr = (int)(r % npair);
    const int s = (int)(r / npair);
    int i = 0, rem = pr;  // pair index -> (i, j), i < j, row-major over the upper triangle
    while (rem >= A - 1 - i) { rem -= A - 1 - i; ++i; }
    const int j = i + 1 + rem;
    const double* pi = xy + (((size_t)s * A + i) * (N + 1) + (k + 1)) * 2;
    const double* pj = xy + (((size_t)s * A + j) * (N + 1) + (k + 1)) * 2;
    const double dx = pi[0] - pj[0], dy = pi[1] - pj[1];
    const double d = sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
    double* l = lambdas + (((size_t)s * A + i) * A + j) * N + k;
    *l = __dadd_rn(*l, __dmul_rn(alpha, dth - d));
  }
}

extern "C" int dmpc_ocd_lambda_update(dmpc_handle* h, int32_t S, int32_t A, double alpha, double dth, const double* d_xy,
                                      double* d_lambdas, void* stream) {
  if (!h || S < 0 || A < 1 || (S > 0 && (!d_xy || !d_lambdas))) return DMPC_ERR_ARG;
  const long long total = (long long)S * A * (A - 1) / 2 * h->P.N;
  if (total == 0) return DMPC_OK;
  const long long want = (total + 255) / 256;
  const int blocks = (int)(want < 148LL * 8 ? want : 148LL * 8);
  ocd_lambda_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(S, A, h->P.N, alpha, dth, d_xy, d_lambdas);
  h->launches++;
  CUDA_TRY(h, cudaGetLastError());
  return DMPC_OK;
}

// Hyperplane form of the same update: planner/scripts/NL_HP_N_main.py:128-137 with eval_constraintHp
// (plan_lib/config/NL/config.py:25-29):  cost = -a_x x_j,k - a_y y_j,k - b + dth/2 with (a_x, a_y, b) =
// planes_raw[i][k-1, :, j], i < j, k = 1..N.  planes [S, A(i), N, 3, A(j)] is the stack of the planes every agent's
// solve returned.  Same thread mapping as above; the plane triple is a strided read (stride A doubles), the position
// and lambda rows are contiguous.  Operation order and roundings as numpy evaluates the expression, left to right.
__global__ void ocd_lambda_hp_kernel(int S, int A, int N, double alpha, double dth, const double* __restrict__ xy,
                                     const double* __restrict__ planes, double* __restrict__ lambdas) {
  const long long npair = (long long)A * (A - 1) / 2;
  const long long total = (long long)S * npair * N;
  const double half = dth / 2;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(e % N);
    long long r = e / N;
    const int pr = (int)(r % npair);
    const int s = (int)(r / npair);
    int i = 0, rem = pr;
    while (rem >= A - 1 - i) { rem -= A - 1 - i; ++i; }
    const int j = i + 1 + rem;
    const double* pj = xy + (((size_t)s * A + j) * (N + 1) + (k + 1)) * 2;
    const double* pl = planes + ((((size_t)s * A + i) * N + k) * 3) * A + j;
    double c = __dmul_rn(-pl[0], pj[0]);
    c = __dsub_rn(c, __dmul_rn(pl[A], pj[1]));
    c = __dsub_rn(c, pl[2 * (size_t)A]);
    c = __dadd_rn(c, half);
    double* l = lambdas + (((size_t)s * A + i) * A + j) * N + k;
    *l = __dadd_rn(*l, __dmul_rn(alpha, c));
  }
}

extern "C" int dmpc_ocd_lambda_update_hp(dmpc_handle* h, int32_t S, int32_t A, double alpha, double dth, const double* d_xy,
                                         const double* d_planes, double* d_lambdas, void* stream) {
  if (!h || S < 0 || A < 1 || (S > 0 && (!d_xy || !d_planes || !d_lambdas))) return DMPC_ERR_ARG;
  const long long total = (long long)S * A * (A - 1) / 2 * h->P.N;
  if (total == 0) return DMPC_OK;
  const long long want = (total + 255) / 256;
  const int blocks = (int)(want < 148LL * 8 ? want : 148LL * 8);
  ocd_lambda_hp_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(S, A, h->P.N, alpha, dth, d_xy, d_planes, d_lambdas);
  h->launches++;
  CUDA_TRY(h, cudaGetLastError());
  return DMPC_OK;
}

extern "C" int dmpc_sim_batch(dmpc_handle* h, int32_t B, const double* d_xpred, const double* d_u, double* d_out,
                              int32_t* d_status, void* stream) {
  if (!h || B < 0 || (B > 0 && (!d_xpred || !d_u || !d_out))) return DMPC_ERR_ARG;
  if (B == 0) return DMPC_OK;
  const size_t smem = sim_smem_bytes(h->P.N);
  if (smem > 48 * 1024 && !h->sim_attr) {
    CUDA_TRY(h, cudaFuncSetAttribute(lpv_sim_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    h->sim_attr = 1;
  }
  lpv_sim_kernel<<<(B + SIM_TILE - 1) / SIM_TILE, SIM_THREADS, smem, (cudaStream_t)stream>>>(h->P, B, d_xpred, d_u, d_out, d_status);
  h->launches++;
  CUDA_TRY(h, cudaGetLastError());
  return DMPC_OK;
}

static int launch_shift_pack(dmpc_handle* h, int32_t B, const double* d_z, const int32_t* d_status, double* d_x0,
                             double* d_xprev, double* d_uprev, double* d_uold, double* d_xy_send, int32_t* d_hold, void* stream) {
  if (!h || !d_z || B < 0) return DMPC_ERR_ARG;
  if (B == 0) return DMPC_OK;
  const long long total = (long long)B * (16 * h->P.N + 12);
  const int blocks = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
  shift_pack_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(B, h->P.N, d_z, d_status, d_x0, d_xprev, d_uprev, d_uold, d_xy_send, d_hold);
  h->launches++;
  CUDA_TRY(h, cudaGetLastError());
  return DMPC_OK;
}

extern "C" int dmpc_shift_pack(dmpc_handle* h, int32_t B, const double* d_z, double* d_x0, double* d_xprev, double* d_uprev,
                               double* d_uold, double* d_xy_send, void* stream) {
  return launch_shift_pack(h, B, d_z, nullptr, d_x0, d_xprev, d_uprev, d_uold, d_xy_send, nullptr, stream);
}

extern "C" int dmpc_shift_pack_hold(dmpc_handle* h, int32_t B, const double* d_z, const int32_t* d_status, double* d_x0,
                                    double* d_xprev, double* d_uprev, double* d_uold, double* d_xy_send, int32_t* d_hold,
                                    void* stream) {
  if (!d_status || !d_hold) return h ? fail(h, DMPC_ERR_ARG, "dmpc_shift_pack_hold needs status and hold%s") : DMPC_ERR_ARG;
  return launch_shift_pack(h, B, d_z, d_status, d_x0, d_xprev, d_uprev, d_uold, d_xy_send, d_hold, stream);
}

// xy_all is laid out as ncclAllGather leaves it: [G][S][A_local][T][2] with G = A_total / A_local ranks, i.e. global
// agent k of scenario s lives at (((k / A_local) * S + s) * A_local + k % A_local).  G = 1 is plain [S][A][T][2].
// rank_stride: doubles per rank in xy_all (S * A_local * T * 2, plus the flag tail of dmpc_gather_neighbours_stop).
// stop != nullptr: the tails [S * A_local] of all ranks are reduced into stop[s] (sticky OR over every agent of scenario s).
__global__ void gather_neighbours_kernel(int S, int A_total, int A_local, int agent_offset, int T, int nn,
                                         const double* __restrict__ xy_all, long long rank_stride, const int* __restrict__ nbr_idx,
                                         double* __restrict__ nbr_xy, double* __restrict__ pose, int* __restrict__ stop) {
  const long long per = (long long)T * (nn + 1);
  const long long total = (long long)S * A_local * per;
  if (stop) {
    const int G = A_total / A_local;
    const long long tail = (long long)S * A_local * T * 2;
    for (int s = blockIdx.x * blockDim.x + threadIdx.x; s < S; s += gridDim.x * blockDim.x) {
      int any = 0;
      for (int g = 0; g < G; ++g)
        for (int a = 0; a < A_local; ++a) any |= xy_all[(size_t)g * rank_stride + tail + (size_t)s * A_local + a] != 0.0;
      if (any) stop[s] = 1;
    }
  }
  for (long long w = blockIdx.x * (long long)blockDim.x + threadIdx.x; w < total; w += (long long)gridDim.x * blockDim.x) {
    const int b = (int)(w / per);
    const int rem = (int)(w % per), t = rem / (nn + 1), j = rem % (nn + 1);
    const int s = b / A_local, al = b % A_local, ag = agent_offset + al;
    const double* self = xy_all + (size_t)(ag / A_local) * rank_stride + ((((size_t)s) * A_local + ag % A_local) * T + t) * 2;
    if (j == nn) {
      pose[((size_t)b * T + t) * 2] = self[0]; pose[((size_t)b * T + t) * 2 + 1] = self[1];
    } else {
      const int k = nbr_idx[(size_t)b * nn + j];
      double vx, vy;
      if (k >= 0 && k < A_total) {
        const double* src = xy_all + (size_t)(k / A_local) * rank_stride + ((((size_t)s) * A_local + k % A_local) * T + t) * 2;
        vx = src[0]; vy = src[1];
      } else { vx = self[0] + 1e3; vy = self[1] + 1e3; }  // empty slot: geometrically inert, never enters the QP
      double* dst = nbr_xy + (((size_t)b * T + t) * nn + j) * 2;
      dst[0] = vx; dst[1] = vy;
    }
  }
}

static int launch_gather(dmpc_handle* h, int32_t S, int32_t A_total, int32_t A_local, int32_t agent_offset,
                         const double* d_xy_all, const int32_t* d_nbr_idx, double* d_nbr_xy, double* d_pose, int32_t* d_stop,
                         void* stream) {
  if (!h || !d_xy_all || !d_pose || S < 0 || A_local < 0 || A_total < A_local || agent_offset < 0 ||
      agent_offset + A_local > A_total || (A_local > 0 && A_total % A_local != 0))
    return DMPC_ERR_ARG;
  if (S == 0 || A_local == 0) return DMPC_OK;
  const int T = h->P.N + 1, nn = h->P.nn;
  if (nn > 0 && (!d_nbr_idx || !d_nbr_xy)) return DMPC_ERR_ARG;
  const long long total = (long long)S * A_local * T * (nn + 1);
  const long long rank_stride = (long long)S * A_local * (T * 2 + (d_stop ? 1 : 0));
  const int blocks = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
  gather_neighbours_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(S, A_total, A_local, agent_offset, T, nn, d_xy_all,
                                                                     rank_stride, d_nbr_idx, d_nbr_xy, d_pose, d_stop);
  h->launches++;
  CUDA_TRY(h, cudaGetLastError());
  return DMPC_OK;
}

extern "C" int dmpc_gather_neighbours(dmpc_handle* h, int32_t S, int32_t A_total, int32_t A_local, int32_t agent_offset,
                                      const double* d_xy_all, const int32_t* d_nbr_idx, double* d_nbr_xy, double* d_pose,
                                      void* stream) {
  return launch_gather(h, S, A_total, A_local, agent_offset, d_xy_all, d_nbr_idx, d_nbr_xy, d_pose, nullptr, stream);
}

extern "C" int dmpc_gather_neighbours_stop(dmpc_handle* h, int32_t S, int32_t A_total, int32_t A_local, int32_t agent_offset,
                                           const double* d_all, const int32_t* d_nbr_idx, double* d_nbr_xy, double* d_pose,
                                           int32_t* d_stop, void* stream) {
  if (!d_stop) return h ? fail(h, DMPC_ERR_ARG, "dmpc_gather_neighbours_stop needs d_stop%s") : DMPC_ERR_ARG;
  return launch_gather(h, S, A_total, A_local, agent_offset, d_all, d_nbr_idx, d_nbr_xy, d_pose, d_stop, stream);
}

// ---------------------------------------------------------------------------------------------------
// Allocation-free entry point (no counterpart in the reference).
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it):
// the stage-wise solve kernel with its global scratch - the slice of each QP where the interior-point iterate is parked
// between the phases of the mu ladder - in a file-scope device array instead of a handle-owned cudaMalloc block.
// The array is held by every GPU that loads this library until the process exits (24 MiB).  All calls share it, and it
// keeps the last call's contents: the kernel writes a QP's slice before it reads it (ipm_park3 saves before any warm
// phase loads), so nothing has to start at zero; calls that overlap on the GPU (two streams, two host threads) race on it.
#define DMPC_STATIC_MAX_DEVICES 64
static __device__ __align__(256) double dmpc_static_scratch[DMPC_STATIC_SCRATCH_DOUBLES];
// `dmpc_static_scratch` in HOST code is a stand-in for the symbol, not a device address: the kernel gets the address
// cudaGetSymbolAddress returns, looked up once per device with that device current.
static double* g_static_ptr[DMPC_STATIC_MAX_DEVICES];
static int g_static_optin[DMPC_STATIC_MAX_DEVICES];

static int static_fail(char* err, int32_t err_len, int code, const char* msg, const char* detail = "") {
  if (err && err_len > 0) snprintf(err, (size_t)err_len, "%s%s", msg, detail);
  return code;
}

extern "C" int dmpc_solve_batch_static(const dmpc_params* p, const dmpc_options* opt, const dmpc_batch_io* io, int32_t device,
                                       void* stream, char* err, int32_t err_len) {
  if (err && err_len > 0) err[0] = 0;
  if (!p || !io) return static_fail(err, err_len, DMPC_ERR_ARG, "null params / io");
  if (io->B == 0) return DMPC_OK;
  if (p->N < 1 || p->N > 511 || p->nn < 0 || p->nn > 123 || p->nseg < 1 || p->nseg > DMPC_MAX_SEG || !(p->dt > 0) || io->B < 0)
    return static_fail(err, err_len, DMPC_ERR_ARG, "bad N/nn/nseg/dt/B");
  if (!psd(p->Q, 9, false) || !psd(p->R, 2, false) || !psd(p->dR, 2, false))
    return static_fail(err, err_len, DMPC_ERR_ARG, "Q, R and dR must be positive semidefinite");
  if (!psd(p->Qs, 3, true)) return static_fail(err, err_len, DMPC_ERR_ARG, "Qs must be positive definite");
  {
    double rs[4];
    for (int i = 0; i < 4; ++i) rs[i] = p->R[i] + p->dR[i];
    if (!psd(rs, 2, true)) return static_fail(err, err_len, DMPC_ERR_ARG, "R + dR must be positive definite (strictly convex inputs)");
  }
  dmpc_options o;
  if (opt) o = *opt; else dmpc_default_options(&o);
  if (!options_ok(&o)) return static_fail(err, err_len, DMPC_ERR_ARG, "bad options");
  if (o.engine == DMPC_ENGINE_FAST) return static_fail(err, err_len, DMPC_ERR_UNSUPPORTED, "dmpc_solve_batch_static runs the stage-wise engine only");
  if (io->xprev_rows < p->N) return static_fail(err, err_len, DMPC_ERR_ARG, "Last_xPredicted needs at least N rows");
  if (!io->x0 || !io->xprev || !io->uprev || !io->uold || !io->pose || (p->nn > 0 && !io->nbr_xy))
    return static_fail(err, err_len, DMPC_ERR_ARG, "null input pointer");
  // the size limit: B QPs x the parked iterate of one QP must fit in the static array
  const size_t stride = (stage_park_doubles(p->N, p->nn) + 1) & ~(size_t)1;
  if ((size_t)io->B * stride > (size_t)DMPC_STATIC_SCRATCH_DOUBLES) {
    char msg[160];
    snprintf(msg, sizeof(msg), "batch of %d QPs x %zu scratch doubles per QP (N=%d, nn=%d) exceeds the static scratch of %d doubles: at most %zu QPs per call",
             io->B, stride, p->N, p->nn, (int)DMPC_STATIC_SCRATCH_DOUBLES, (size_t)DMPC_STATIC_SCRATCH_DOUBLES / stride);
    return static_fail(err, err_len, DMPC_ERR_NOMEM, msg);
  }
  cudaError_t e;
  if (device < 0) { if ((e = cudaGetDevice(&device)) != cudaSuccess) return static_fail(err, err_len, DMPC_ERR_CUDA, "CUDA: ", cudaGetErrorString(e)); }
  if (device >= DMPC_STATIC_MAX_DEVICES) return static_fail(err, err_len, DMPC_ERR_ARG, "device index beyond DMPC_STATIC_MAX_DEVICES");
  if ((e = cudaSetDevice(device)) != cudaSuccess) return static_fail(err, err_len, DMPC_ERR_CUDA, "CUDA: ", cudaGetErrorString(e));
  if (!g_static_ptr[device]) {  // set-up, once per device, that device current (not thread-safe, like the handles)
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return static_fail(err, err_len, DMPC_ERR_CUDA, "CUDA: ", cudaGetErrorString(e));
    if (prop.major < 10) return static_fail(err, err_len, DMPC_ERR_UNSUPPORTED, "this library is built for sm_100a (B200) only");
    int optin = 0;
    if ((e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device)) != cudaSuccess)
      return static_fail(err, err_len, DMPC_ERR_CUDA, "CUDA: ", cudaGetErrorString(e));
    double* ptr = nullptr;
    if ((e = cudaGetSymbolAddress((void**)&ptr, dmpc_static_scratch)) != cudaSuccess || !ptr)
      return static_fail(err, err_len, DMPC_ERR_CUDA, "cudaGetSymbolAddress(dmpc_static_scratch): ", cudaGetErrorString(e));
    g_static_optin[device] = optin;
    g_static_ptr[device] = ptr;
  }
  Ctx3 c3;
  const size_t smem = carve_workspace3(c3, nullptr, p->N, p->nn);
  if (smem + 1024 > (size_t)g_static_optin[device])
    return static_fail(err, err_len, DMPC_ERR_UNSUPPORTED, "horizon/neighbour count too large for the on-chip workspace");
  // the opt-in limit is a per-function, per-device setting shared with the handles: always raise it to the device maximum
  // (never below what a live handle asked for)
  if ((e = cudaFuncSetAttribute(lpv_qp_stage_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, g_static_optin[device])) != cudaSuccess)
    return static_fail(err, err_len, DMPC_ERR_CUDA, "CUDA: ", cudaGetErrorString(e));
  DevParams P;
  memset(&P, 0, sizeof(P));
  fill_dev_params(p, &o, P);
  BatchPtrs bp;
  memset(&bp, 0, sizeof(bp));
  bp.B = io->B; bp.xprev_rows = io->xprev_rows;
  bp.x0 = io->x0; bp.xprev = io->xprev; bp.uprev = io->uprev; bp.uold = io->uold; bp.nbr_xy = io->nbr_xy; bp.pose = io->pose;
  bp.nbr_count = io->nbr_count;
  bp.z = io->z; bp.planes = io->planes; bp.y_ineq = io->y_ineq; bp.y_eq = io->y_eq; bp.obj = io->obj;
  bp.status = io->status; bp.iters = io->iters; bp.pivots = io->pivots;
  bp.gws = g_static_ptr[device]; bp.gws_stride = (long long)stride;
  // nothing is carried between calls here: no launch order, no warm start (those live in handle-owned buffers)
  bp.bulk_in = bulk_in_enabled();
  lpv_qp_stage_kernel<<<io->B, QP3_THREADS, smem, (cudaStream_t)stream>>>(P, bp);
  if ((e = cudaGetLastError()) != cudaSuccess) return static_fail(err, err_len, DMPC_ERR_CUDA, "CUDA: ", cudaGetErrorString(e));
  return DMPC_OK;
}

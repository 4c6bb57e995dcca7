// TEST INFRASTRUCTURE ONLY (tests/hostsim): the stage-wise engine's DEVICE source (colaborativempc_b200/csrc/qp3_*.cuh)
// compiled for the host with one "lane" and no barriers (QP3_HOST), so that the algorithm - Riccati sweeps, active-set
// rounds, interior point, epilogue - is checked against the oracle in the CPU-only test tier, where no GPU exists.
// It is built by tests/test_stage_hostsim.py into tests/hostsim/_build/ and loaded by that test alone; the product
// (colaborativempc_b200/) never references it: the product path runs lpv_qp_stage_kernel on sm_100a or fails.
#define QP3_HOST 1
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../../colaborativempc_b200/csrc/qp3_kernel.cuh"
#include "../../colaborativempc_b200/csrc/qp_params.h"

extern "C" int qp3_host_workspace_bytes(int N, int nn) {
  Ctx3 c;
  return (int)carve_workspace3(c, nullptr, N, nn);
}

// counters [B,4]: factorisations, solves, active-set rounds, interior-point iterations
extern "C" int qp3_host_solve_batch(const dmpc_params* p, const dmpc_options* opt, const dmpc_batch_io* io,
                                    const dmpc_shift_io* sh, int* counters, unsigned char* warm, int warm_shift, int* cost) {
  DevParams P;
  memset(&P, 0, sizeof(P));
  fill_dev_params(p, opt, P);
  BatchPtrs bp;
  memset(&bp, 0, sizeof(bp));
  bp.B = io->B; bp.xprev_rows = io->xprev_rows;
  bp.x0 = io->x0; bp.xprev = io->xprev; bp.uprev = io->uprev; bp.uold = io->uold; bp.nbr_xy = io->nbr_xy; bp.pose = io->pose;
  bp.nbr_count = io->nbr_count;
  bp.z = io->z; bp.planes = io->planes; bp.y_ineq = io->y_ineq; bp.y_eq = io->y_eq; bp.obj = io->obj;
  bp.status = io->status; bp.iters = io->iters; bp.pivots = io->pivots;
  if (sh) { bp.nx_x0 = sh->x0; bp.nx_xprev = sh->xprev; bp.nx_uprev = sh->uprev; bp.nx_uold = sh->uold; bp.nx_xy = sh->xy_send; bp.hold = sh->hold;
            bp.flags = sh->flags; bp.stop = sh->stop; bp.stop_div = sh->stop_div > 0 ? sh->stop_div : 1; bp.laps = sh->laps; bp.end_s = sh->end_s; }
  bp.warm = warm; bp.warm_shift = warm_shift;
  std::vector<int> ident(io->B);
  for (int b = 0; b < io->B; ++b) ident[b] = b;
  if (cost) { bp.cost = cost; bp.order = ident.data(); }  // previous call's costs: the hard-QP hint is live
  const size_t park = stage_park_doubles(P.N, P.nn);
  std::vector<double> gws((size_t)io->B * park);
  bp.gws = gws.data(); bp.gws_stride = (long long)park;
  Ctx3 c0;
  const size_t bytes = carve_workspace3(c0, nullptr, P.N, P.nn);
#pragma omp parallel
  {
    char* ws = (char*)aligned_alloc(64, (bytes + 63) & ~(size_t)63);
    q3_host_smem = ws;
#pragma omp for schedule(dynamic, 1)
    for (int b = 0; b < io->B; ++b) {
      memset(ws, 0, bytes);
      if (bp.stop && bp.stop[b / bp.stop_div]) continue;
      solve_one_qp3(P, bp, b);
      if (counters) {
        const Hdr3* hd = (const Hdr3*)ws;
        counters[4 * b] = hd->n_fact; counters[4 * b + 1] = hd->n_solve; counters[4 * b + 2] = hd->n_rounds; counters[4 * b + 3] = hd->n_ipm;
      }
    }
    free(ws);
  }
  return 0;
}

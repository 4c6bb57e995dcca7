"""Per-iteration cost of the closed-loop DMPC step on one GPU (256 x 16-agent chain platoons, N=20): ms per iteration,
interior-point iterations and Schur rounds per QP, unsolved count - with the L2 flushed before every iteration and
without.  Shows how the cold-start transient differs from the settled loop.  Measurement aid."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from colaborativempc_b200 import scenarios as sc  # noqa: E402
from colaborativempc_b200.dmpc import ShardedDMPC  # noqa: E402
from colaborativempc_b200.planner import BatchedPlannerLPV  # noqa: E402
from colaborativempc_b200.track import TrackMap  # noqa: E402


def main(iters=80):
    cfg, tr = sc.shipped_config(), TrackMap("Highway")
    N, A, S = 20, 16, 256
    args = (cfg["Q"], cfg["Qs"], cfg["R"], cfg["dR"], N, cfg["dt"], tr, cfg["wq"], cfg["model_param"], cfg["sys_lim"])
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for do_flush in (1, 0):
        eng = BatchedPlannerLPV(*args, nn=2, max_batch=S * A)
        x0, xprev, uprev = sc.platoon_scenarios(S, A, N, cfg["dt"], tr, seed0=0)
        idx, cnt = sc.topology(A, "chain", nn=2)
        dm = ShardedDMPC(eng, S, A, idx, cnt, x0, xprev, uprev)
        for it in range(iters):
            if do_flush:
                flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = dm.iterate()
            e1.record()
            torch.cuda.synchronize()
            if it < 8 or it % 8 == 0:
                st = out["status"].cpu().numpy()
                print("flush %d it %3d  %.3f ms  ipm/QP %.3f  rounds/QP %.2f  unsolved %d" % (
                    do_flush, it, e0.elapsed_time(e1), out["iters"].double().mean().item(),
                    out["pivots"].double().mean().item(), int((st != 1).sum())), flush=True)
        st = out["status"].cpu().numpy()
        bad = np.flatnonzero(st != 1)
        z = out["z"].cpu().numpy()
        print("final status histogram", dict(zip(*np.unique(st, return_counts=True))))
        print("unsolved: agent index in platoon", np.bincount(bad % A, minlength=A).tolist(), " scenarios", len(np.unique(bad // A)),
              " NaN in z", int(np.isnan(z[bad]).any(1).sum()))
        x0 = dm.x0.cpu().numpy()
        print("unsolved x0 (vx, ey, epsi, s) sample", np.round(x0[bad[:6]][:, [0, 3, 4, 6]], 3).tolist())
        print("solved   x0 (vx, ey, epsi, s) sample", np.round(x0[:3][:, [0, 3, 4, 6]], 3).tolist())


if __name__ == "__main__":
    main()

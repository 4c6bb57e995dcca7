"""Closed loop of the DMPC bench workload through the host entry point; at the first iterations where a QP does not
come back with status 1, its inputs and outputs are saved (gpurun_out/bad_qp.npz) for analysis against the oracle."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from colaborativempc_b200 import scenarios as sc  # noqa: E402
from colaborativempc_b200.planner import BatchedPlannerLPV  # noqa: E402
from colaborativempc_b200.track import TrackMap  # noqa: E402

cfg, tr = sc.shipped_config(), TrackMap("Highway")
N, A, S = 20, 16, 256
args = (cfg["Q"], cfg["Qs"], cfg["R"], cfg["dR"], N, cfg["dt"], tr, cfg["wq"], cfg["model_param"], cfg["sys_lim"])
eng = BatchedPlannerLPV(*args, nn=2, max_batch=S * A)
x0, xprev, uprev = sc.platoon_scenarios(S, A, N, cfg["dt"], tr, seed0=0)
idx, cnt = sc.topology(A, "chain", nn=2)
rec = []  # run_dmpc stops at the first status outside {1, 2, -2}
sc.run_dmpc(lambda **kw: eng.solve_batch_host(**kw), x0, xprev, uprev, idx, cnt, 40, N, record=rec)
save = {}
for it, (ins, out) in enumerate(rec):
    bad = np.flatnonzero(out["status"] != 1)
    if len(bad) or it % 8 == 0: print("it", it, dict(zip(*np.unique(out["status"], return_counts=True))), bad[:12].tolist(), flush=True)
    for b in bad[:4]:
        if not np.isnan(ins["x0"][b]).any():
            for k, v in ins.items():
                save["it%d_b%d_in_%s" % (it, b, k)] = v[b]
            for k in ("z", "status", "iters", "pivots", "obj", "y_ineq"):
                save["it%d_b%d_out_%s" % (it, b, k)] = out[k][b]
            print("   b", b, "status", out["status"][b], "ipm iters", out["iters"][b], "rounds", out["pivots"][b],
                  "NaN in z", bool(np.isnan(out["z"][b]).any()), "obj", out["obj"][b])
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "bad_qp.npz"), **save)

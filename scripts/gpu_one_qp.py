"""The regression QP of tests/golden/closed_loop_ipm_start_qp.npz through both engines and a few option sets."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from colaborativempc_b200 import scenarios as sc  # noqa: E402
from colaborativempc_b200.planner import BatchedPlannerLPV  # noqa: E402
from colaborativempc_b200.track import TrackMap  # noqa: E402

KEYS = ("x0", "xprev", "uprev", "uold", "nbr_xy", "pose", "nbr_count")
g = np.load(os.path.join(ROOT, "tests", "golden", "closed_loop_ipm_start_qp.npz"))
cfg, tr = sc.shipped_config(), TrackMap("Highway")
args = (cfg["Q"], cfg["Qs"], cfg["R"], cfg["dR"], 20, cfg["dt"], tr, cfg["wq"], cfg["model_param"], cfg["sys_lim"])
ins = {k: g["in_" + k] for k in KEYS}
for opt in ({"engine": 1}, {"engine": 2}, {"engine": 2, "pdas_first": 8}, {"engine": 2, "pdas_first": 1},
            {"engine": 2, "ipm_mu0": 1e-3}, {"engine": 2, "ipm_mu0": 1e-9}, {"engine": 2, "order_hint": 0}):
    eng = BatchedPlannerLPV(*args, nn=2, max_batch=4, options=opt)
    out = eng.solve_batch_host(**ins)
    err = np.abs(out["z"] - g["out_z"]).max()
    print(opt, "status", out["status"], "ipm", out["iters"], "rounds", out["pivots"], "err", err, flush=True)
# perturbations of the inputs: how isolated is it?
eng = BatchedPlannerLPV(*args, nn=2, max_batch=64, options={"engine": 2})
rng = np.random.default_rng(0)
for eps in (1e-12, 1e-9, 1e-6, 1e-3):
    rep = {k: np.ascontiguousarray(np.repeat(v, 64, axis=0)) for k, v in ins.items()}
    rep["x0"] = rep["x0"] * (1 + eps * rng.normal(size=rep["x0"].shape))
    out = eng.solve_batch_host(**rep)
    print("x0 perturbed by", eps, dict(zip(*np.unique(out["status"], return_counts=True))), flush=True)

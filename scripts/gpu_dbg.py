"""Development aid: statuses of the warm-up DMPC iterations of the bench batch."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from colaborativempc_b200 import scenarios as sc
from colaborativempc_b200.planner import BatchedPlannerLPV
from colaborativempc_b200.track import TrackMap

cfg, track = sc.shipped_config(), TrackMap("Highway")
B = 4096
eng = BatchedPlannerLPV(*bench.planner_args(cfg, track, 20), nn=2, max_batch=B + 8)
S = (B + 2) // 3
x0, xprev, uprev = sc.platoon_scenarios(S, 3, 20, cfg["dt"], track, seed0=0)
idx, cnt = sc.topology(3, "full", nn=2)
rec = []
sc.run_dmpc(lambda **kw: eng.solve_batch_host(**kw), x0, xprev, uprev, idx, cnt, 4, 20, record=rec)
for k, (ins, out) in enumerate(rec):
    st = out["status"]
    print("iteration", k, "status counts", dict(zip(*np.unique(st, return_counts=True))), "nan z rows", int(np.isnan(out["z"]).any(1).sum()),
          "nan inputs", int(sum(np.isnan(v).sum() for v in ins.values() if v is not None and v.dtype.kind == "f")))
    bad = np.flatnonzero(st != 1)
    print("   not solved:", [(int(b), int(st[b]), int(out["iters"][b]), int(out["pivots"][b])) for b in bad[:12]])
    nz = np.flatnonzero(np.isnan(out["z"]).any(1))
    print("   nan z:", [(int(b), int(st[b]), int(out["iters"][b]), int(out["pivots"][b]), int(np.isnan(out["z"][b]).sum()), int(np.isnan(out["y_ineq"][b]).sum()), float(out["obj"][b])) for b in nz[:12]])
    if nz.size and st[nz[0]] == 1:
        b = nz[0]
        print("   z nan positions", np.flatnonzero(np.isnan(out["z"][b]))[:40])
        np.savez("gpurun_out/nan_qp.npz", **{kk: v[b:b + 1] for kk, v in ins.items() if v is not None})
    if k == 0 and bad.size:
        np.savez("gpurun_out/bad_qps.npz", idx=bad, **{kk: v[bad] for kk, v in ins.items() if v is not None})

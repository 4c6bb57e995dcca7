"""Development aid: cost split of the bench batch by solver path (Schur rounds only / penalty rounds / interior point)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from colaborativempc_b200 import scenarios as sc
from colaborativempc_b200.planner import BatchedPlannerLPV
from colaborativempc_b200.track import TrackMap

cfg, track = sc.shipped_config(), TrackMap("Highway")
B = 4096
eng = BatchedPlannerLPV(*bench.planner_args(cfg, track, 20), nn=2, max_batch=B + 8)
ins = bench.steady_state_inputs(lambda **kw: eng.solve_batch_host(duals=False, **kw), B, cfg, track, A=3)
out = eng.solve_batch_host(duals=False, **ins)
it, pv = out["iters"], out["pivots"]
print("ipm QPs %d (%.1f%%), mean its among them %.1f; rounds: mean %.2f p50 %d p90 %d p99 %d max %d" % (
    (it > 0).sum(), 100 * (it > 0).mean(), it[it > 0].mean() if (it > 0).any() else 0, pv.mean(), np.percentile(pv, 50),
    np.percentile(pv, 90), np.percentile(pv, 99), pv.max()))
print("rounds histogram", np.bincount(pv)[:20])
print("ipm QPs (index, ipm its, rounds, status):", [(int(i), int(it[i]), int(pv[i]), int(out["status"][i])) for i in np.flatnonzero(it > 0)])
np.savez("gpurun_out/hard_qps.npz", idx=np.flatnonzero(it > 0), **{k: v[it > 0] for k, v in ins.items()})

def timeit(sel, label):
    idx = np.flatnonzero(sel)
    if idx.size == 0:
        print(label, "empty"); return
    rep = np.resize(idx, B)
    d = {k: torch.from_numpy(np.ascontiguousarray(v[rep])).cuda() for k, v in ins.items()}
    o = eng.alloc_outputs(B, duals=False)
    for _ in range(3): eng.solve_batch(out=o, duals=False, **d)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): eng.solve_batch(out=o, duals=False, **d)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print("%-34s n=%5d  %.3f ms per 4096 -> %.0f QP/s" % (label, idx.size, ms, B / ms * 1e3))

timeit(np.ones(B, bool), "all")
timeit((it == 0) & (pv <= 1), "rounds<=1, no ipm")
timeit((it == 0) & (pv == 2), "rounds==2, no ipm")
timeit((it == 0) & (pv == 3), "rounds==3, no ipm")
timeit((it == 0) & (pv >= 4), "rounds>=4, no ipm")
timeit(it > 0, "ipm")

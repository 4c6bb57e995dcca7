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
opts = {k: (float(v) if "." in v or "e" in v else int(v)) for k, v in (a.split("=") for a in sys.argv[1:])}  # e.g. ipm_mu0=1e-8
print("options", opts)
eng = BatchedPlannerLPV(*bench.planner_args(cfg, track, 20), nn=2, max_batch=B + 8, options=opts)
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


def latency(sel, label):
    """Time of a batch holding exactly the selected QPs once (<= one CTA per SM: the latency of the slowest of them)."""
    idx = np.flatnonzero(sel)
    if idx.size == 0:
        return
    d = {k: torch.from_numpy(np.ascontiguousarray(v[idx])).cuda() for k, v in ins.items()}
    o = eng.alloc_outputs(idx.size, duals=False)
    for _ in range(3): eng.solve_batch(out=o, duals=False, **d)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): eng.solve_batch(out=o, duals=False, **d)
    e1.record(); torch.cuda.synchronize()
    print("%-34s n=%5d  latency %.3f ms" % (label, idx.size, e0.elapsed_time(e1) / 5))


latency(it > 0, "ipm QPs alone")
latency((it == 0) & (pv >= 12), "rounds>=12 alone")
worst = np.zeros(B, bool); worst[np.argmax(it * 24 + pv)] = True
latency(worst, "worst QP alone")

# ---- correctness of the interior-point QPs against the exact oracle, and a fingerprint of the batch
print("nan counts", {k: int(np.isnan(v).sum()) for k, v in ins.items() if v.dtype.kind == "f"}); print("fingerprint x0 %.12e xprev %.12e nbr %.12e" % (np.nansum(np.abs(ins["x0"])), np.nansum(np.abs(ins["xprev"])), np.nansum(np.abs(ins["nbr_xy"]))))
from colaborativempc_b200._abi import make_params
from oracle import lpv_oracle as lo
P = make_params(*bench.planner_args(cfg, track, 20), nn=2)
sel = np.flatnonzero(it > 0)
sub = {k: v[sel] for k, v in ins.items()}
o = lo.solve_batch(P, st=lo.settings(exact=True), nthreads=16, **sub)
g = eng.solve_batch_host(duals=False, **sub)
ok = o["status"] == 1
err = np.abs(g["z"] - o["z"]).max(1) / (1 + np.abs(o["z"]).max(1))
print("oracle certified %d of %d; rel err z vs oracle: max %.2e (over certified), obj rel %.2e" % (
    ok.sum(), sel.size, err[ok].max() if ok.any() else -1, (np.abs(g["obj"] - o["obj"]) / (1 + np.abs(o["obj"])))[ok].max() if ok.any() else -1))

# ---- what bounds the batch: the same batch without its interior-point QPs / without the long-round QPs
timeit(it == 0, "batch minus ipm QPs (replicated to 4096)")
timeit((it == 0) & (pv < 12), "batch minus ipm and rounds>=12")
timeit((it == 0) & (pv < 6), "batch minus ipm and rounds>=6")

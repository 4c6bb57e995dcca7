"""BASELINE.json configs 3 and 4 on one GPU: horizon sweep N in {10, 20, 50, 100} at nn=2 (B single-agent QPs out of
3-agent platoons) and the 64-vehicle platoon (chain nn<=2, fully connected nn=63).  For each case: 3 DMPC iterations
to steady state through the CUDA path, then the frozen batch is timed device-resident (CUDA events, L2 flushed), and a
sample of the batch is re-solved by the oracle's exact mode (parity: z / obj relative error, active set).
Measurement aid; prints one JSON line per case.  usage: gpu_sweep.py N nn topo A S [sample]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from colaborativempc_b200 import scenarios as sc  # noqa: E402
from colaborativempc_b200._abi import make_params  # noqa: E402
from colaborativempc_b200.planner import BatchedPlannerLPV  # noqa: E402
from colaborativempc_b200.track import TrackMap  # noqa: E402

KEYS = ("x0", "xprev", "uprev", "uold", "nbr_xy", "pose", "nbr_count")


def main(N, nn, topo, A, S, sample=24, steps=5):
    cfg, tr = sc.shipped_config(), TrackMap("Highway")
    args = (cfg["Q"], cfg["Qs"], cfg["R"], cfg["dR"], N, cfg["dt"], tr, cfg["wq"], cfg["model_param"], cfg["sys_lim"])
    B = S * A
    eng = BatchedPlannerLPV(*args, nn=nn, max_batch=B)
    x0, xprev, uprev = sc.platoon_scenarios(S, A, N, cfg["dt"], tr, seed0=0)
    idx, cnt = sc.topology(A, topo, nn=nn)
    st = sc.run_dmpc(lambda **kw: eng.solve_batch_host(duals=False, **kw), x0, xprev, uprev, idx, cnt, 3, N)
    ins = {k: np.ascontiguousarray(st[k]) for k in KEYS}
    dev = torch.device("cuda", 0)
    d_in = {k: torch.from_numpy(v).to(dev) for k, v in ins.items()}
    d_out = eng.alloc_outputs(B, duals=False)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for _ in range(3):
        eng.solve_batch(out=d_out, duals=False, **d_in)
    torch.cuda.synchronize()
    ms = []
    for _ in range(steps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        eng.solve_batch(out=d_out, duals=False, **d_in)
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    status = d_out["status"].cpu().numpy()
    line = {"N": N, "nn": nn, "topology": topo, "agents": A, "scenarios": S, "qp": B, "engine": "fast" if eng.fast else "generic",
            "ms_per_batch": float(np.mean(ms)), "qp_per_s": B / (np.mean(ms) * 1e-3),
            "status": {str(k): int(v) for k, v in zip(*np.unique(status, return_counts=True))},
            "ipm_iters_mean": float(d_out["iters"].double().mean().item())}
    if sample > 0:
        from oracle import lpv_oracle as lo  # checker only
        sel = np.linspace(0, B - 1, min(sample, B)).astype(int)
        out = eng.solve_batch_host(**{k: np.ascontiguousarray(v[sel]) for k, v in ins.items()})
        o = lo.solve_batch(make_params(*args, nn=nn), st=lo.settings(exact=True), nthreads=16,
                           **{k: np.ascontiguousarray(v[sel]) for k, v in ins.items()})
        both = (o["status"] == 1) & (out["status"] == 1)
        rel = lambda a, b: (np.abs(a - b).max(1) / (1 + np.abs(b).max(1)))
        m_in = eng.m_in
        sc_y = np.abs(o["y"]).max(1)
        act = lambda y: y > 1e-8 * (1 + sc_y[:, None])
        line["parity"] = {"sample": int(len(sel)), "both_solved": int(both.sum()),
                          "oracle_solved": int((o["status"] == 1).sum()), "cuda_solved": int((out["status"] == 1).sum()),
                          "z_rel_max": float(rel(out["z"], o["z"])[both].max()) if both.any() else None,
                          "obj_rel_max": float((np.abs(out["obj"] - o["obj"]) / (1 + np.abs(o["obj"])))[both].max()) if both.any() else None,
                          "y_rel_max": float((np.abs(out["y_ineq"] - o["y"][:, :m_in]).max(1) / (1 + sc_y))[both].max()) if both.any() else None,
                          "active_set_equal": bool(np.array_equal(act(out["y_ineq"])[both], act(o["y"][:, :m_in])[both]))}
    print(json.dumps(line))


if __name__ == "__main__":
    a = sys.argv[1:]
    main(int(a[0]), int(a[1]), a[2], int(a[3]), int(a[4]), int(a[5]) if len(a) > 5 else 24)

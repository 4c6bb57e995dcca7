"""Quick GPU sanity run: golden platoon fixture through the CUDA path, diffs vs the committed oracle output,
plus a rough device-side timing at B=4096.  Development aid; the real checks are tests/ -m gpu and bench.py."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from colaborativempc_b200 import scenarios as sc  # noqa: E402
from colaborativempc_b200.planner import BatchedPlannerLPV  # noqa: E402
from colaborativempc_b200.track import TrackMap  # noqa: E402


def main(engine=0):
    N = 20
    print("==== engine", engine)
    cfg = sc.shipped_config()
    tr = TrackMap("Highway")
    fx = np.load(os.path.join(ROOT, "tests", "golden", "platoon_N20.npz"))
    eng = BatchedPlannerLPV(cfg["Q"], cfg["Qs"], cfg["R"], cfg["dR"], N, cfg["dt"], tr, cfg["wq"], cfg["model_param"],
                            cfg["sys_lim"], nn=2, max_batch=4096, options={"engine": engine})
    for it in (0, 3):
        ins = {k: fx["it%d_%s" % (it, k)] for k in ("x0", "xprev", "uprev", "uold", "nbr_xy", "pose", "nbr_count")}
        out = eng.solve_batch_host(**ins)
        zg, yg, og = fx["it%d_out_z" % it], fx["it%d_out_y" % it], fx["it%d_out_obj" % it]
        m_in = eng.m_in
        dz = np.abs(out["z"] - zg).max(1) / (1 + np.abs(zg).max(1))
        dyi = np.abs(out["y_ineq"] - yg[:, :m_in]).max(1) / (1 + np.abs(yg).max(1))
        dye = np.abs(out["y_eq"] - yg[:, m_in:]).max(1) / (1 + np.abs(yg).max(1))
        dob = np.abs(out["obj"] - og) / (1 + np.abs(og))
        dpl = np.abs(out["planes"] - fx["it%d_out_planes" % it]).max()
        act_g = np.abs(yg[:, :m_in]) > 1e-9 * (1 + np.abs(yg).max(1, keepdims=True))
        act_o = np.abs(out["y_ineq"]) > 1e-9 * (1 + np.abs(yg).max(1, keepdims=True))
        print("it%d status %s ipm-its mean %.1f max %d rounds mean %.2f max %d" % (
            it, dict(zip(*np.unique(out["status"], return_counts=True))), out["iters"].mean(), out["iters"].max(),
            out["pivots"].mean(), out["pivots"].max()))
        print("   rel err z %.2e  y_ineq %.2e  y_eq %.2e  obj %.2e  planes %.1e  active-set mismatches %d" % (
            dz.max(), dyi.max(), dye.max(), dob.max(), dpl, int((act_g != act_o).sum())))
        worst = int(np.argmax(dz))
        print("   worst z row %d: dz %.2e status %d" % (worst, dz[worst], out["status"][worst]))
    # rough timing, device resident
    ins = {k: fx["it3_%s" % k] for k in ("x0", "xprev", "uprev", "uold", "nbr_xy", "pose", "nbr_count")}
    B = 4096
    rep = (B + ins["x0"].shape[0] - 1) // ins["x0"].shape[0]
    dev = {}
    for k, v in ins.items():
        t = np.concatenate([v] * rep, 0)[:B]
        dev[k] = torch.from_numpy(np.ascontiguousarray(t)).cuda()
    out = eng.alloc_outputs(B)
    for _ in range(2):
        eng.solve_batch(out=out, **dev)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    reps = 5
    for _ in range(reps):
        eng.solve_batch(out=out, **dev)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print("B=%d: %.3f ms per batch -> %.0f QP/s (device resident); status ok %d" % (
        B, ms, B / ms * 1e3, int((out["status"] == 1).sum().item())))
    t0 = time.time()
    host_out = eng.solve_batch_host(**{k: v.cpu().numpy() for k, v in dev.items()})
    t1 = time.time() - t0
    print("host path B=%d: %.3f ms -> %.0f QP/s" % (B, t1 * 1e3, B / t1))


if __name__ == "__main__":
    for e in ([int(a) for a in sys.argv[1:]] or [2, 1]):
        main(e)

"""Generates the committed golden fixtures.  Run in the BUILD container only (needs /root/reference):

    python tests/golden/make_golden.py

1. ref_startup_N20.npz   - the REFERENCE's own functions (imported unmodified from /root/reference, osqp stubbed)
                           run on the shipped 3-agent start-up case at N=20: planes, weights, A_k, B_k, ey, and the
                           OSQP-form QP (P diagonal, q, A, l, u) exactly as LPV_Planner.py:138-157,222-228 build them.
                           Pins oracle/lpv_oracle.c's assembly (bit-exact) and, through it, the CUDA path.
2. ref_map_tables.npz    - Map(...).PointAndTangent / halfWidth / TrackLength of the reference tracks.
3. ref_startup_pred.npz  - initialise_agents() of the reference for x0_database[0:3] at N=20 and N=125.
4. platoon_N20.npz       - 16 synthetic 3-agent platoons after 3 DMPC iterations driven by the oracle in
                           exact mode: inputs of iteration 4 + the oracle's certified solution (z, y, obj, status).
                           (Oracle output, not reference output: OSQP itself cannot run here.)
5. recorded_N125_it0.npz - first-iteration uPred of the recorded experiment LPV3r_agent_laptop (pck/*/u.pkl), the only
                           solver-level data the reference ships; loose (OSQP eps=1e-3) - used as a smoke bound only.
"""
import os
import pickle
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
HERE = os.path.dirname(os.path.abspath(__file__))

from oracle import ref_assembly as ra  # noqa: E402
from oracle import lpv_oracle as lo  # noqa: E402
from colaborativempc_b200 import scenarios as sc  # noqa: E402
from colaborativempc_b200._abi import make_params  # noqa: E402
from colaborativempc_b200.track import TrackMap  # noqa: E402


def main():
    ref = ra.load()
    S = ra.shipped_settings(np)
    # ---- 2. map tables
    maps = {}
    for name in ("Highway", "oval", "oval_mt", "TestOpenMap"):
        m = ref.Map(name)
        maps[name + "_pt"] = m.PointAndTangent
        maps[name + "_hw"] = np.asarray(m.halfWidth, dtype=np.float64)
        maps[name + "_len"] = np.asarray(m.TrackLength, dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, "ref_map_tables.npz"), **maps)
    # ---- 3. start-up predictions
    pred = {}
    mp = ref.Map("Highway")
    for N in (20, 125):
        agents, xs, us = ref.initialise_agents(ref.x0_database[0:3], N, S["dt"], [mp] * 3)
        pred["agents_N%d" % N] = agents
        pred["x_N%d" % N] = np.asarray(xs)
    np.savez_compressed(os.path.join(HERE, "ref_startup_pred.npz"), **pred)
    # ---- 1. reference assembly, N=20 start-up, plus a second DMPC-style call with shifted (N-row) history
    N = 20
    agents, x_old, u_old = ref.initialise_agents(ref.x0_database[0:3], N, S["dt"], [mp] * 3)
    ns = [[1, 2], [0, 2], [0, 1]]
    out = dict(agents=agents, x_old=np.asarray(x_old), u_old=np.asarray(u_old), ns=np.asarray(ns))
    rng = np.random.default_rng(7)
    for i in range(3):
        c = ref.PlannerLPV(S["Q"], S["Qs"], S["R"], S["dR"], N, S["dt"], mp, i, S["wq"], S["model_param"], S["sys_lim"])
        qp = ra.assemble(ref, c, x_old[i][0, :], x_old[i], u_old[i], agents[:, ns[i], :], ns[i], agents[:, i, :], [0.0, 0.0])
        # a second call on the same controller with N-row history, non-zero inputs and uOld (steady-state shape)
        xs2 = x_old[i][1:, :].copy()
        xs2[:, 0] += rng.uniform(-0.3, 0.3, N); xs2[:, 1] += rng.uniform(-0.05, 0.05, N); xs2[:, 4] += rng.uniform(-0.05, 0.05, N)
        xs2[3, 0] = 0.1  # exercises the low-speed branch (LPV_Planner.py:505)
        us2 = np.stack([rng.uniform(-0.2, 0.2, N), rng.uniform(-2, 2, N)], 1)
        uo2 = [0.05, -0.7]
        qp2 = ra.assemble(ref, c, xs2[0], xs2, us2, agents[:, ns[i], :], ns[i], agents[:, i, :], uo2)
        for tag, q_, extra in (("a%d" % i, qp, {}), ("b%d" % i, qp2, dict(xs=xs2, us=us2, uold=np.asarray(uo2)))):
            assert np.count_nonzero(q_["P"] - np.diag(np.diag(q_["P"]))) == 0
            out.update({tag + "_Pdiag": np.diag(q_["P"]).copy(), tag + "_q": q_["q"], tag + "_A": q_["A"], tag + "_l": q_["l"],
                        tag + "_u": q_["u"], tag + "_planes": q_["planes"], tag + "_weights": q_["weights"], tag + "_Ak": q_["Ak"],
                        tag + "_Bk": q_["Bk"], tag + "_ey": q_["ey"]})
            out.update({tag + "_" + k: v for k, v in extra.items()})
    np.savez_compressed(os.path.join(HERE, "ref_startup_N20.npz"), **out)
    # ---- 4. platoon fixture from the oracle (exact mode)
    cfg = sc.shipped_config()
    tr = TrackMap("Highway")
    P = make_params(cfg["Q"], cfg["Qs"], cfg["R"], cfg["dR"], N, cfg["dt"], tr, cfg["wq"], cfg["model_param"], cfg["sys_lim"], nn=2)
    Sn, A = 16, 3
    x0, xprev, uprev = sc.platoon_scenarios(Sn, A, N, cfg["dt"], tr, seed0=0)
    idx, cnt = sc.topology(A, "full")
    st = lo.settings(exact=True)

    def solve(**kw):
        return lo.solve_batch(P, kw["x0"], kw["xprev"], kw["uprev"], kw["uold"], kw["nbr_xy"], kw["pose"], kw["nbr_count"], st=st, nthreads=8)

    rec = []
    sc.run_dmpc(solve, x0, xprev, uprev, idx, cnt, 4, N, record=rec)
    fx = {}
    for it in (0, 3):
        ins, o = rec[it]
        assert np.all(o["status"] == 1) and o["cert"].max() < 1e-8, (o["status"], o["cert"].max(0))
        fx.update({"it%d_%s" % (it, k): v for k, v in ins.items()})
        fx.update({"it%d_out_%s" % (it, k): o[k] for k in ("z", "y", "obj", "status", "planes")})
    np.savez_compressed(os.path.join(HERE, "platoon_N20.npz"), **fx)
    # ---- 5. recorded first iteration
    base = "/root/reference/planner/scripts/data/experiments_paper/LPV3r_agent_laptop/pck"
    recd = {}
    for i in range(3):
        with open(os.path.join(base, str(i), "u.pkl"), "rb") as f:
            u = pickle.load(f)
        recd["u_it0_agent%d" % i] = np.asarray(u[0])
        recd["u_it1_agent%d" % i] = np.asarray(u[1])
    np.savez_compressed(os.path.join(HERE, "recorded_N125_it0.npz"), **recd)
    for f in sorted(os.listdir(HERE)):
        print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()

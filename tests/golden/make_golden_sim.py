"""Golden fixture for SURVEY 8(f2): the REFERENCE's LPV_Model.sim (LPV_val.py:77-94), imported unmodified from
/root/reference, run on the shipped start-up predictions.  Run in the BUILD container only:

    python tests/golden/make_golden_sim.py   ->  tests/golden/ref_lpv_model_sim_N20.npz

Three agents (x0_database[0:3], Highway map, shipped vehicle parameters), 6 consecutive sim() calls each with seeded
random inputs; every call's xPred before, u, and returned xPred are stored (18 cases)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
HERE = os.path.dirname(os.path.abspath(__file__))

from oracle import ref_assembly as ra  # noqa: E402


def main():
    ref = ra.load()
    S = ra.shipped_settings(np)
    N = 20
    mp = ref.Map("Highway")
    agents, x_old, u_old = ref.initialise_agents(ref.x0_database[0:3], N, S["dt"], [mp] * 3)
    rng = np.random.default_rng(11)
    xin, uin, xout = [], [], []
    for i in range(3):
        model = ref.LPV_Model(N, S["dt"], mp, model_param=S["model_param"], sys_lim=S["sys_lim"])
        model.xPred = np.array(x_old[i], dtype=np.float64)
        for call in range(6):
            u = np.stack([rng.uniform(-0.25, 0.25, N), rng.uniform(-2.0, 2.0, N)], 1)
            xin.append(np.array(model.xPred, dtype=np.float64))
            uin.append(u)
            xout.append(np.array(model.sim(u), dtype=np.float64))
    np.savez_compressed(os.path.join(HERE, "ref_lpv_model_sim_N20.npz"), xpred=np.stack(xin), u=np.stack(uin), out=np.stack(xout))
    print("cases", len(xin), "max |x|", np.abs(np.stack(xout)).max())


if __name__ == "__main__":
    main()

"""Golden fixture for SURVEY 8(f4): the dual update of the reference's OCD loop (NL_EU_N_main.py:122-138) executed with
the REFERENCE's own eval_constraintEU / eval_constraintHp (NL_HP_N_main.py:128-137) / get_alpha (plan_lib/config/NL/config.py:5-28, imported unmodified).
Run in the BUILD container only:

    python tests/golden/make_golden_ocd.py   ->  tests/golden/ref_ocd_lambda.npz
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference/planner/lib")
from plan_lib.config.NL import config as nl  # noqa: E402


def reference_update(lambdas, agents, dth):
    """The loop of NL_EU_N_main.py:125-136, verbatim in structure: agents (N+1, n_agents, 2), lambdas (n, n, N)."""
    n_agents, N = lambdas.shape[0], lambdas.shape[2]
    cost = np.zeros((n_agents, n_agents, N))
    for k in range(1, N + 1):
        for i in range(0, n_agents):
            for j in range(0, n_agents):
                if (i != j) and i < j:
                    cost[i, j, k - 1] = nl.eval_constraintEU(agents[k, i, :], agents[k, j, :], dth)
    return lambdas + nl.get_alpha() * cost


def reference_update_hp(lambdas, agents, planes_raw, dth):
    """The loop of NL_HP_N_main.py:128-137: planes_raw[i] (N, 3, n_agents) is what agent i's solve returned."""
    n_agents, N = lambdas.shape[0], lambdas.shape[2]
    cost = np.zeros((n_agents, n_agents, N))
    planes = np.zeros((N, n_agents, n_agents, 3))
    for k in range(1, N + 1):
        for i in range(0, n_agents):
            for j in range(0, n_agents):
                if (i != j) and i < j:
                    planes[k - 1, i, j, :] = planes_raw[i][k - 1, :, j]
                    cost[i, j, k - 1] = nl.eval_constraintHp(agents[k, j, :], planes[k - 1, i, j, :], dth)
    return lambdas + nl.get_alpha() * cost


def main():
    rng = np.random.default_rng(21)
    out = {"alpha": np.array(nl.get_alpha())}
    for tag, (S, A, N, dth) in dict(a=(6, 3, 20, 0.25), b=(3, 5, 10, 0.4)).items():
        xy = rng.normal(size=(S, A, N + 1, 2)) * 0.5
        lam0 = rng.uniform(0, 1, size=(S, A, A, N))
        lam1 = np.stack([reference_update(lam0[s], np.swapaxes(xy[s], 0, 1), dth) for s in range(S)])
        lam2 = np.stack([reference_update(lam1[s], np.swapaxes(xy[s], 0, 1), dth) for s in range(S)])
        out.update({tag + "_xy": xy, tag + "_lam0": lam0, tag + "_lam1": lam1, tag + "_lam2": lam2, tag + "_dth": np.array(dth)})
    for tag, (S, A, N, dth) in dict(hp_a=(5, 3, 20, 0.25), hp_b=(2, 6, 10, 0.4)).items():
        xy = rng.normal(size=(S, A, N + 1, 2)) * 0.5
        pl = rng.normal(size=(S, A, N, 3, A))
        pl[:, :, :, :2] /= np.sqrt((pl[:, :, :, :2] ** 2).sum(3, keepdims=True))  # unit normals, as compute_hyperplane
        lam0 = rng.uniform(0, 1, size=(S, A, A, N))
        lam1 = np.stack([reference_update_hp(lam0[s], np.swapaxes(xy[s], 0, 1), pl[s], dth) for s in range(S)])
        lam2 = np.stack([reference_update_hp(lam1[s], np.swapaxes(xy[s], 0, 1), pl[s], dth) for s in range(S)])
        out.update({tag + "_xy": xy, tag + "_planes": pl, tag + "_lam0": lam0, tag + "_lam1": lam1, tag + "_lam2": lam2,
                    tag + "_dth": np.array(dth)})
    np.savez_compressed(os.path.join(HERE, "ref_ocd_lambda.npz"), **out)
    print("ok", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()

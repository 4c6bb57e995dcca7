"""Closed-loop run bookkeeping of the reference drivers, host side of `ShardedDMPC`:

* `check_end`      - the loop condition of LPV_HP_N_main.py:99 (utilities/misc.py:28-48) for S scenarios at once;
* `slack_violated` - the "slack violated" warning of LPV_HP_N_main.py:46-49;
* `RunLog`         - what `experiment_utilities.save` keeps per agent and `save_to_csv` writes (config/base_class.py:47-90):
                     `<path_csv>csv/<id>/{states,u,plan_dist,time}.dat`, '%.5e', space separated, one row per iteration,
                     so the reference's MATLAB / plotting scripts read the files of a batched run unchanged;
* `run_closed_loop` - LPV_HP_N_main.py:96-120 around `ShardedDMPC.iterate` (the loop body runs on the device).
"""
import os
import time
import warnings

import numpy as np


def check_end(s_now, track_length, laps=1):
    """misc.py:28-48 for every scenario: s_now [S,A] = x_pred[i][0,-3] of each agent (row 0 of the current prediction).
    A scenario has ended when ANY of its agents is within 0.15 m of the end of the track (np.isclose, atol=0.15) or
    beyond it, on lap number `laps` (track_initialization.py:319-323: floor(s / TrackLength)).  Returns bool [S]."""
    s = np.atleast_2d(np.asarray(s_now, dtype=np.float64))
    L = float(np.asarray(track_length, dtype=np.float64).reshape(-1)[0])
    near = np.isclose(s, L, atol=0.15) | (s > L)
    return np.any(near & (np.floor(s / L) == laps), axis=1)


def slack_violated(s_pred, thr=0.1):
    """LPV_HP_N_main.py:46: (Controller.sPred[:,1] >= 0.1).any() per QP.  s_pred [..., N, 3] -> bool [...]."""
    return np.any(np.asarray(s_pred)[..., 1] >= thr, axis=-1)


class RunLog(object):
    """Per-agent history of one scenario, as `experiment_utilities` keeps it (base_class.py:47-62)."""

    def __init__(self, agent_id, path_csv=""):
        self.id, self.path_csv = agent_id, path_csv
        self.states, self.u, self.look_ahead, self.planes, self.status, self.time_op = [], [], [], [], [], []

    def save(self, xPred, uPred=None, feas=None, planes=None):
        self.states.append(np.array(xPred[0, :]))
        self.look_ahead.append(xPred[-1, 6] - xPred[0, 6])
        if uPred is not None:
            self.u.append(np.array(uPred[0, :]))
        if planes is not None:
            self.planes.append(np.array(planes[0, :]))
        if feas is not None:
            self.status.append(feas)

    def save_to_csv(self):
        path = os.path.join(self.path_csv + "csv/", str(self.id))
        os.makedirs(path, exist_ok=True)
        kw = dict(fmt="%.5e", delimiter=" ")
        np.savetxt(path + "/states.dat", self.states, **kw)
        np.savetxt(path + "/u.dat", self.u, **kw)
        np.savetxt(path + "/plan_dist.dat", self.look_ahead, **kw)
        np.savetxt(path + "/time.dat", self.time_op, **kw)
        return path


def run_closed_loop(dm, max_it, logs=None, log_scenario=0, warn_slack=True):
    """LPV_HP_N_main.py:96-120 around a `ShardedDMPC`: iterate until `max_it`, until ANY agent of the logged scenario - on
    any rank - has a QP that is not feasible (the reference breaks on the first `not feas`, before it records the
    iteration), or until any of them is at the end of the track (`checkEnd` with the track length `dm.end_s`, evaluated
    like the reference on row 0 of the last prediction, before the next iteration).  Both conditions are read from the
    flags that travelled with the exchange, so every rank takes the same decision on the same iteration.
    logs: list of `RunLog`, one per local agent of scenario `log_scenario` (rows [log_scenario*A_local, +A_local) of the
    batch); per iteration only that scenario's solution (A_local x (16N+12) doubles) crosses PCIe.
    Returns (iterations done, ended, error)."""
    N, al = dm.N, dm.a_local
    rows = slice(log_scenario * al, (log_scenario + 1) * al)
    it, error, ended = 0, False, False
    while it < max_it and not ended:
        tic = time.time()
        out = dm.iterate()
        z = out["z"][rows].cpu().numpy()  # synchronises the stream
        st = out["status"][rows].cpu().numpy()
        pl = out["planes"][rows].cpu().numpy()
        dt_op = (time.time() - tic) / max(dm.B, 1)
        ended_now, failed = dm.scenario_flags(log_scenario)   # of this iteration, over all ranks
        if failed:
            error = True
            break
        feas = np.isin(st, (1, 2, -2))  # LPV_Planner.py:246-248 (all true here: a local failure raises the flag)
        xs = z[:, :12 * (N + 1)].reshape(al, N + 1, 12)
        x_pred, s_pred = xs[:, :, :9], xs[:, 1:, 9:]
        u_pred = z[:, 12 * (N + 1):12 * (N + 1) + 2 * N].reshape(al, N, 2)
        if warn_slack and slack_violated(s_pred).any():
            warnings.warn("WARNING slack violated !")
        if logs is not None:
            for a, lg in enumerate(logs):
                lg.time_op.append(dt_op)
                lg.save(x_pred[a], u_pred[a], int(feas[a]), pl[a])
        it += 1
        ended = ended_now
    return it, ended, error

"""Golden fixture for SURVEY 8(f1): the files the REFERENCE's experiment_utilities.save / save_to_csv write
(plan_lib/config/base_class.py:47-90, imported unmodified) for a seeded history, and the truth table of the reference's
checkEnd (plan_lib/utilities/misc.py:28-48) on the reference's own Map("Highway").  Run in the BUILD container only:

    python tests/golden/make_golden_runlog.py   ->  tests/golden/ref_runlog.npz
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference/planner/lib")
from plan_lib.config.base_class import experiment_utilities  # noqa: E402
from plan_lib.mapManager import Map  # noqa: E402
from plan_lib.utilities.misc import checkEnd  # noqa: E402


class _Agent(experiment_utilities):
    """The bookkeeping half of `agent` (LPV_HP_N_main.py:20-36): the object is its own `data`."""

    def __init__(self, settings, id):
        experiment_utilities.__init__(self, self, settings)
        self.id, self.states, self.u, self.planes, self.status, self.time_op = id, [], [], [], [], []


def main():
    rng = np.random.default_rng(5)
    N, nn, T = 20, 2, 7
    xs = rng.normal(size=(T, N + 1, 9)) * np.array([3, .1, .1, .2, .1, 1, 20, 10, 10])
    us = rng.normal(size=(T, N, 2)) * 0.3
    pl = rng.normal(size=(T, N, 3, nn))
    tm = rng.uniform(1e-4, 5e-2, size=T)
    out = dict(xs=xs, us=us, pl=pl, tm=tm)
    with tempfile.TemporaryDirectory() as d:
        eu = data = _Agent({"path_csv": d + "/", "path_pck": d + "/", "vx_ref": 3.0}, 4)
        for t in range(T):
            eu.save(xs[t], us[t], 1, pl[t])
            data.time_op.append(tm[t])
        eu.save_to_csv()
        for name in ("states", "u", "plan_dist", "time"):
            with open(os.path.join(d, "csv", "4", name + ".dat"), "rb") as f:
                out["file_" + name] = np.frombuffer(f.read(), dtype=np.uint8)
    m = Map("Highway")
    L = float(np.asarray(m.TrackLength).reshape(-1)[0])
    s = np.concatenate([np.linspace(-1, 2.2 * L, 97), L + np.array([-0.16, -0.15, -0.1499, -1e-9, 0, 1e-9, 0.15, 0.2]),
                        2 * L + np.array([-0.15, -0.01, 0, 0.01])])
    s3 = rng.choice(s, size=(40, 3))
    def ref(svals, laps):
        xl = []
        for v in svals:
            x = np.zeros((N + 1, 9)); x[0, 6] = v
            xl.append(x)
        return bool(checkEnd(xl, [m] * len(xl), laps))
    out.update(track_length=np.array(L), s1=s, end1_l1=np.array([ref([v], 1) for v in s]),
               end1_l2=np.array([ref([v], 2) for v in s]), s3=s3, end3_l1=np.array([ref(r, 1) for r in s3]))
    np.savez_compressed(os.path.join(HERE, "ref_runlog.npz"), **out)
    print("ok", L, int(out["end1_l1"].sum()), int(out["end1_l2"].sum()), int(out["end3_l1"].sum()))


if __name__ == "__main__":
    main()

"""SURVEY 8(f1) host side: the run log / .dat writer and the end-of-track check against files and answers produced by
the REFERENCE's own code (tests/golden/make_golden_runlog.py)."""
import os

import numpy as np

from colaborativempc_b200.runlog import RunLog, check_end, slack_violated

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_runlog.npz")


def test_dat_files_byte_identical_with_reference(tmp_path):
    g = np.load(GOLD)
    lg = RunLog(4, str(tmp_path) + "/")
    for t in range(g["xs"].shape[0]):
        lg.save(g["xs"][t], g["us"][t], 1, g["pl"][t])
        lg.time_op.append(g["tm"][t])
    path = lg.save_to_csv()
    assert path == os.path.join(str(tmp_path), "csv", "4")
    for name in ("states", "u", "plan_dist", "time"):
        with open(os.path.join(path, name + ".dat"), "rb") as f:
            assert f.read() == g["file_" + name].tobytes(), name
    assert np.array_equal(np.asarray(lg.planes), g["pl"][:, 0]) and lg.status == [1] * g["xs"].shape[0]


def test_check_end_matches_reference_truth_table():
    g = np.load(GOLD)
    L = g["track_length"]
    assert np.array_equal(check_end(g["s1"][:, None], L, 1), g["end1_l1"])
    assert np.array_equal(check_end(g["s1"][:, None], L, 2), g["end1_l2"])
    assert np.array_equal(check_end(g["s3"], [float(L)], 1), g["end3_l1"])
    assert g["end1_l1"].any() and not g["end1_l1"].all() and g["end1_l2"].any()


def test_slack_warning_rule():
    s = np.zeros((3, 20, 3))
    s[1, 7, 1] = 0.1
    s[2, 7, 0] = 5.0  # only the second slack (lane) is looked at, LPV_HP_N_main.py:46
    assert slack_violated(s).tolist() == [False, True, False]

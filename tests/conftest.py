"""pytest configuration: `gpu` marker, repo root on sys.path, shared fixtures."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def built():
    import __graft_entry__ as g
    g.build()
    return True


@pytest.fixture(scope="session")
def cfg():
    from colaborativempc_b200 import scenarios as sc
    return sc.shipped_config()


@pytest.fixture(scope="session")
def track():
    from colaborativempc_b200.track import TrackMap
    return TrackMap("Highway")


def planner_args(cfg, track, N):
    return (cfg["Q"], cfg["Qs"], cfg["R"], cfg["dR"], N, cfg["dt"], track, cfg["wq"], cfg["model_param"], cfg["sys_lim"])


def golden(name):
    return np.load(os.path.join(GOLDEN, name))


INPUT_KEYS = ("x0", "xprev", "uprev", "uold", "nbr_xy", "pose", "nbr_count")


def rel_err(a, b):
    """max-abs error per row relative to 1 + max|b| of that row."""
    a = np.asarray(a).reshape(len(b), -1)
    b = np.asarray(b).reshape(len(b), -1)
    return np.abs(a - b).max(1) / (1.0 + np.abs(b).max(1))

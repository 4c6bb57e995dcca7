"""CPU tier: the stage-wise engine's device source (csrc/qp3_*.cuh) compiled for the host (tests/hostsim, one lane, no
barriers) against the oracle and the committed golden vectors.  This checks the ALGORITHM the sm_100a kernel runs -
Riccati sweeps, active-set rounds, interior point, epilogue, fused bookkeeping - where no GPU exists; the kernel itself
is checked on the B200 by tests/test_gpu_parity.py.  The host build is test infrastructure: the product never loads it.

Parity bar (BASELINE.json north_star): same active set; primal / dual solution and objective within 1e-6 relative."""
import numpy as np
import pytest

from conftest import INPUT_KEYS, golden, planner_args, rel_err

TOL = 1e-6


@pytest.fixture(scope="module")
def hs(built):
    from hostsim import hostsim
    hostsim.build()
    return hostsim


def _params(cfg, track, N, nn):
    from colaborativempc_b200._abi import make_params
    return make_params(*planner_args(cfg, track, N), nn=nn)


def _oracle(cfg, track, ins, N=20, nn=2, nthreads=8):
    from oracle import lpv_oracle as lo
    return lo.solve_batch(_params(cfg, track, N, nn), st=lo.settings(exact=True), nthreads=nthreads, **ins)


def _active(y_in, scale):
    return y_in > 1e-8 * (1.0 + scale[:, None])


def _compare(out, z, y, obj, planes, m_in, tol=TOL):
    assert np.all(out["status"] == 1), np.unique(out["status"], return_counts=True)
    scale = np.abs(y).max(1)
    assert rel_err(out["z"], z).max() < tol
    assert rel_err(out["obj"][:, None], obj[:, None]).max() < tol
    assert (np.abs(out["y_ineq"] - y[:, :m_in]).max(1) / (1 + scale)).max() < tol
    assert (np.abs(out["y_eq"] - y[:, m_in:]).max(1) / (1 + scale)).max() < tol
    assert np.array_equal(_active(out["y_ineq"], scale), _active(y[:, :m_in], scale))
    if planes is not None and planes.size:
        assert np.abs(out["planes"] - planes).max() < 1e-12


@pytest.mark.parametrize("it", [0, 3])
def test_golden_platoon(hs, cfg, track, it):
    fx = golden("platoon_N20.npz")
    ins = {k: fx["it%d_%s" % (it, k)] for k in INPUT_KEYS}
    out = hs.solve_batch(_params(cfg, track, 20, 2), **ins)
    _compare(out, fx["it%d_out_z" % it], fx["it%d_out_y" % it], fx["it%d_out_obj" % it], fx["it%d_out_planes" % it], 200, tol=1e-7)


def test_low_speed_qp(hs, cfg, track):
    """The QP of the 16-agent closed loop whose condensed Hessian cannot be factorised (vx 0.57 -> 0.22, lateral dynamics
    unstable, prediction matrices ~1e11): the stage-wise solve must certify it and match the oracle's optimum."""
    g = golden("closed_loop_ipm_start_qp.npz")
    out = hs.solve_batch(_params(cfg, track, 20, 2), **{k: g["in_" + k] for k in INPUT_KEYS})
    _compare(out, g["out_z"], g["out_y"], g["out_obj"], g["out_planes"], 200, tol=1e-8)


def test_reference_startup_case(hs, cfg, track):
    """KKT certificate on the matrices the REFERENCE assembled (golden, generated from the unmodified reference)."""
    g = golden("ref_startup_N20.npz")
    ins = dict(x0=g["x_old"][:, 0], xprev=g["x_old"], uprev=g["u_old"], uold=np.zeros((3, 2)),
               nbr_xy=np.stack([g["agents"][:, g["ns"][i], :] for i in range(3)]),
               pose=np.stack([g["agents"][:, i, :] for i in range(3)]), nbr_count=None)
    out = hs.solve_batch(_params(cfg, track, 20, 2), **ins)
    assert np.all(out["status"] == 1)
    for i in range(3):
        assert np.array_equal(out["planes"][i], g["a%d_planes" % i])
        z = out["z"][i]
        y = np.concatenate([out["y_ineq"][i], out["y_eq"][i]])
        Pd, q, A, l, u = (g["a%d_%s" % (i, k)] for k in ("Pdiag", "q", "A", "l", "u"))
        Az = A @ z
        assert max((Az - u).max(), (l - Az).max()) < 1e-9
        assert np.abs(Pd * z + q + A.T @ y).max() < 1e-6 * (1 + np.abs(q).max())
        assert out["y_ineq"][i].min() >= 0
        assert abs(0.5 * z @ (Pd * z) + q @ z - out["obj"][i]) < 1e-9 * abs(out["obj"][i])
    want_obj = (-1396.897816823358, -1490.2691025771865, -1402.3910432237897)
    assert np.abs(out["obj"] - want_obj).max() < 1e-6 * 1400


@pytest.mark.parametrize("Nh,nn,topo,S,A,iters", [(10, 2, "full", 8, 3, 4), (20, 2, "chain", 6, 6, 4), (20, 5, "full", 4, 6, 4),
                                                  (50, 2, "full", 4, 3, 3), (20, 0, "full", 8, 1, 4), (100, 2, "full", 2, 3, 2),
                                                  (20, 15, "full", 1, 16, 2)])
def test_closed_loop_vs_oracle(hs, cfg, track, Nh, nn, topo, S, A, iters):
    """DMPC iterations driven by the stage-wise engine; every iteration's batch is re-solved by the oracle's exact mode
    (an independent uncondensed sparse interior point + active-set polish + KKT certificate)."""
    from colaborativempc_b200 import scenarios as sc
    P = _params(cfg, track, Nh, nn)
    x0, xprev, uprev = sc.platoon_scenarios(S, A, Nh, cfg["dt"], track, seed0=100)
    idx, cnt = sc.topology(A, topo, nn=nn)
    rec = []
    sc.run_dmpc(lambda **kw: hs.solve_batch(P, **kw), x0, xprev, uprev, idx, cnt, iters, Nh, record=rec)
    assert len(rec) == iters
    n_cmp = 0
    for ins, out in rec:
        assert np.all(out["status"] == 1), np.unique(out["status"], return_counts=True)
        o = _oracle(cfg, track, ins, N=Nh, nn=nn)
        ok = o["status"] == 1
        assert ok.mean() >= 0.6
        _compare({k: v[ok] for k, v in out.items() if k not in ("next",)}, o["z"][ok], o["y"][ok], o["obj"][ok], o["planes"][ok],
                 (8 + nn) * Nh)
        n_cmp += int(ok.sum())
    assert n_cmp > 0


def test_recorded_reference_run_horizon_125(hs, cfg, track):
    """The reference's shipped horizon (N=125, config_files/config_LPV.py:24) from its own cold start: certified, equal to
    the oracle's optimum, and within 0.1 of the recorded un-polished OSQP iterate (see test_oracle_golden)."""
    from colaborativempc_b200 import scenarios as sc
    rec = golden("recorded_N125_it0.npz")
    Nh = 125
    P = _params(cfg, track, Nh, 2)
    agents, xs, us = sc.initialise_agents(sc.X0_DATABASE[0:3], Nh, cfg["dt"], track)
    idx, cnt = sc.topology(3, "full", nn=2)
    log = []
    sc.run_dmpc(lambda **kw: hs.solve_batch(P, **kw), np.stack([x[0] for x in xs])[None], np.stack(xs)[None],
                np.stack(us)[None], idx, cnt, 1, Nh, record=log)
    ins, out = log[0]
    assert np.all(out["status"] == 1)
    o = _oracle(cfg, track, ins, N=Nh, nn=2, nthreads=3)
    assert np.all(o["status"] == 1)
    _compare(out, o["z"], o["y"], o["obj"], o["planes"], 10 * Nh)
    u0 = 12 * (Nh + 1)
    for i in range(3):
        u = out["z"][i, u0:u0 + 2 * Nh].reshape(Nh, 2)
        assert np.abs(u - rec["u_it0_agent%d" % i]).max() < 0.1


def test_edge_cases_and_fused_bookkeeping(hs, cfg, track):
    """Ragged neighbour counts; failed map lookup (the reference raises) and coincident agents; and the epilogue's
    bookkeeping (LPV_HP_N_main.py:102-117): feasible QPs are shifted exactly as the host loop shifts them, failed QPs keep
    their buffers and count in `hold`."""
    from colaborativempc_b200 import scenarios as sc
    N = 20
    fx = golden("platoon_N20.npz")
    P = _params(cfg, track, N, 2)
    ins = {k: fx["it3_%s" % k][:6].copy() for k in INPUT_KEYS}
    L = float(np.ravel(track.TrackLength)[0])
    ins["xprev"][1, 5, 6] = L                      # curvature() finds no segment (misc.py:97)
    ins["nbr_xy"][2, 4, 0] = ins["pose"][2, 4]     # compute_plane.py:52 divides by zero
    ins["nbr_count"] = np.array([2, 2, 2, 1, 0, 2], np.int32)
    out = hs.solve_batch(P, shift=True, **ins)
    assert out["status"][1] == -20 and np.all(np.isnan(out["z"][1]))
    assert out["status"][2] == -21
    keep = np.array([0, 3, 4, 5])
    o = _oracle(cfg, track, {k: v[keep] for k, v in ins.items()})
    sub = {k: v[keep] for k, v in out.items() if k != "next"}
    _compare(sub, o["z"], o["y"], o["obj"], None, 200)
    yi = out["y_ineq"][4][:(4 + 2) * N].reshape(N, 6)
    assert np.all(yi[:, 4:] == 0)                  # rows of unused neighbour slots carry zero multipliers
    nx = out["next"]
    x0n, xpn, upn, uon, xyn = sc.dmpc_shift(out["z"][keep], N)
    assert np.array_equal(nx["x0"][keep], x0n) and np.array_equal(nx["xprev"][keep], xpn)
    assert np.array_equal(nx["uprev"][keep], upn) and np.array_equal(nx["uold"][keep], uon)
    assert np.array_equal(nx["xy_send"][keep], xyn)
    assert np.array_equal(nx["hold"], [0, 1, 1, 0, 0, 0])
    for b in (1, 2):                               # failed: last valid plan kept, nothing non-finite leaves the QP
        assert np.array_equal(nx["x0"][b], ins["x0"][b]) and np.array_equal(nx["uprev"][b], ins["uprev"][b])
        assert np.array_equal(nx["xprev"][b], ins["xprev"][b]) and np.all(nx["xy_send"][b] == 0)


def test_host_build_is_not_on_the_product_path():
    """No Python module of the product imports or loads the host build, and no product source names its library."""
    import os
    root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "colaborativempc_b200")
    for dp, _, fs in os.walk(root):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "libqp3_host" not in txt and "QP3_HOST 1" not in txt, f
                if f.endswith(".py"):
                    assert "hostsim" not in txt, f

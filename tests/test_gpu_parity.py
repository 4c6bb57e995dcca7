"""GPU: the CUDA path (through the C ABI) against the oracle and the committed golden vectors.

Parity bar (BASELINE.json north_star): same active set; primal / dual solution and objective within 1e-6 relative
at FP64.  rel_err = max|a-b| / (1 + max|b|) per QP."""
import numpy as np
import pytest

from conftest import INPUT_KEYS, golden, planner_args, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-6
N = 20


def _engine(cfg, track, N=20, nn=2, max_batch=64, **kw):
    from colaborativempc_b200.planner import BatchedPlannerLPV
    return BatchedPlannerLPV(*planner_args(cfg, track, N), nn=nn, max_batch=max_batch, **kw)


def _oracle(cfg, track, ins, N=20, nn=2, nthreads=8):
    from colaborativempc_b200._abi import make_params
    from oracle import lpv_oracle as lo
    P = make_params(*planner_args(cfg, track, N), nn=nn)
    o = lo.solve_batch(P, st=lo.settings(exact=True), nthreads=nthreads, **ins)
    return o


def _active(y_in, scale):
    return y_in > 1e-8 * (1.0 + scale[:, None])


def _compare(out, z, y, obj, planes, m_in):
    assert np.all(out["status"] == 1), np.unique(out["status"], return_counts=True)
    scale = np.abs(y).max(1)
    assert rel_err(out["z"], z).max() < TOL
    assert rel_err(out["obj"][:, None], obj[:, None]).max() < TOL
    assert (np.abs(out["y_ineq"] - y[:, :m_in]).max(1) / (1 + scale)).max() < TOL
    assert (np.abs(out["y_eq"] - y[:, m_in:]).max(1) / (1 + scale)).max() < TOL
    assert np.array_equal(_active(out["y_ineq"], scale), _active(y[:, :m_in], scale))
    if planes is not None and planes.size:
        assert np.abs(out["planes"] - planes).max() < 1e-12


def _kkt_certificate(cfg, track, ins, out, b, N, nn, tol=TOL):
    """Independent optimality check of QP b of a batch against the matrices of the reference's own assembly
    (oracle.assemble_dense is pinned bit-exact to the reference's _buildMat* by tests/test_oracle_golden.py):
    primal feasibility, dual sign, stationarity and complementarity.  The QP is strictly convex in the inputs, so a
    point passing this IS the reference QP's optimum, whatever route a solver took to it."""
    from colaborativempc_b200._abi import make_params
    from oracle import lpv_oracle as lo
    P = make_params(*planner_args(cfg, track, N), nn=nn)
    d = lo.assemble_dense(P, ins["x0"][b], ins["xprev"][b], ins["uprev"][b], ins["uold"][b], ins["nbr_xy"][b], ins["pose"][b])
    A, l, u, Pm, q = d["A"], d["l"], d["u"], d["P"], d["q"]
    z = out["z"][b]
    y = np.concatenate([out["y_ineq"][b], out["y_eq"][b]])
    Az = A @ z
    zsc = 1 + np.abs(z).max()  # the parity bar is 1e-6 RELATIVE: states reach |s| ~ 47 m, and a 50-step roll-out of the
    assert max((Az - u).max(), (l - Az).max()) < tol * zsc  # LPV dynamics amplifies rounding to ~1e-5 absolute
    m_in = out["y_ineq"].shape[1]
    ysc = 1 + np.abs(y).max()
    assert out["y_ineq"][b].min() >= -1e-9 * ysc
    assert np.abs(Pm @ z + q + A.T @ y).max() < tol * (ysc + np.abs(q).max())
    assert np.abs(y[:m_in] * (Az[:m_in] - u[:m_in])).max() < tol * ysc * zsc
    assert abs(0.5 * z @ Pm @ z + q @ z - out["obj"][b]) < 1e-8 * (1 + abs(out["obj"][b]))


@pytest.mark.parametrize("engine", [3, 2])
@pytest.mark.parametrize("it", [0, 3])
def test_golden_platoon(built, cfg, track, it, engine):
    """Committed oracle solutions of 16 three-agent platoons (first call: N+1 history rows; later call: N rows),
    on the stage-wise (Riccati, engine 3) and on the condensed (engine 2) engine."""
    fx = golden("platoon_N20.npz")
    eng = _engine(cfg, track, options={"engine": engine})
    ins = {k: fx["it%d_%s" % (it, k)] for k in INPUT_KEYS}
    out = eng.solve_batch_host(**ins)
    _compare(out, fx["it%d_out_z" % it], fx["it%d_out_y" % it], fx["it%d_out_obj" % it], fx["it%d_out_planes" % it], eng.m_in)


def test_reference_startup_case(built, cfg, track):
    """The reference's shipped 3-agent start-up at N=20 (inputs from the reference's initialise_agents, golden)."""
    g = golden("ref_startup_N20.npz")
    eng = _engine(cfg, track)
    ins = dict(x0=g["x_old"][:, 0], xprev=g["x_old"], uprev=g["u_old"], uold=np.zeros((3, 2)),
               nbr_xy=np.stack([g["agents"][:, g["ns"][i], :] for i in range(3)]),
               pose=np.stack([g["agents"][:, i, :] for i in range(3)]), nbr_count=None)
    out = eng.solve_batch_host(**ins)
    for i in range(3):
        assert np.array_equal(out["planes"][i], g["a%d_planes" % i])
    # KKT certificate against the matrices the reference assembled
    for i in range(3):
        z = out["z"][i]
        y = np.concatenate([out["y_ineq"][i], out["y_eq"][i]])
        Pd, q, A, l, u = (g["a%d_%s" % (i, k)] for k in ("Pdiag", "q", "A", "l", "u"))
        Az = A @ z
        assert max((Az - u).max(), (l - Az).max()) < 1e-9
        assert np.abs(Pd * z + q + A.T @ y).max() < 1e-6 * (1 + np.abs(q).max())
        assert out["y_ineq"][i].min() >= 0
        assert abs(0.5 * z @ (Pd * z) + q @ z - out["obj"][i]) < 1e-9 * abs(out["obj"][i])
    o = _oracle(cfg, track, ins)
    _compare(out, o["z"], o["y"], o["obj"], o["planes"], eng.m_in)


@pytest.mark.parametrize("engine", [3, 2])
def test_low_speed_qp_is_solved(built, cfg, track, engine):
    """A QP of the 16-agent closed loop (scenario 231, agent 3, second DMPC iteration; inputs captured on the GPU in round
    1, solution = the oracle's certified optimum).  The vehicle crawls (vx 0.57 -> 0.22 over the horizon) where the
    forward-Euler lateral dynamics of _EstimateABC are unstable (|1 - dt (Cf+Cr)/(m vx)| up to 7): the condensed
    prediction matrices grow to ~1e11 and the condensed Hessian cannot be factorised in FP64.  The reference hands the
    UNCONDENSED QP to OSQP (LPV_Planner.py:222-239) and solves it; so must this path: the stage-wise engine directly, the
    condensed engine by re-solving stage-wise what it reports as DMPC_UNSOLVED, inside the same call."""
    g = golden("closed_loop_ipm_start_qp.npz")
    eng = _engine(cfg, track, options={"engine": engine})
    out = eng.solve_batch_host(**{k: g["in_" + k] for k in INPUT_KEYS})
    _compare(out, g["out_z"], g["out_y"], g["out_obj"], g["out_planes"], eng.m_in)


def test_recorded_reference_run_horizon_125(built, cfg, track):
    """The reference's recorded experiment (LPV3r_agent_laptop, N=125): its first DMPC iteration from the reference's
    cold start through the CUDA path (stage-wise engine, 125 stages in shared memory).  Parity with the oracle's certified
    optimum at 1e-6; against the recorded uPred - an un-polished OSQP iterate at eps 1e-3, see
    test_restated_osqp_reproduces_the_reference_recorded_run - the optimum sits within 0.1 (9.1e-2 / 5.9e-2 / 4.2e-2,
    SURVEY 8c), and the recorded point's objective is not better than ours once its constraint violation is paid for."""
    from colaborativempc_b200 import scenarios as sc
    rec = golden("recorded_N125_it0.npz")
    Nh = 125
    eng = _engine(cfg, track, N=Nh, nn=2, max_batch=3)
    agents, xs, us = sc.initialise_agents(sc.X0_DATABASE[0:3], Nh, cfg["dt"], track)
    idx, cnt = sc.topology(3, "full", nn=2)
    log = []
    sc.run_dmpc(lambda **kw: eng.solve_batch_host(**kw), np.stack([x[0] for x in xs])[None], np.stack(xs)[None],
                np.stack(us)[None], idx, cnt, 1, Nh, record=log)
    ins, out = log[0]
    o = _oracle(cfg, track, ins, N=Nh, nn=2, nthreads=3)
    assert np.all(o["status"] == 1)
    _compare(out, o["z"], o["y"], o["obj"], o["planes"], eng.m_in)
    u0 = 12 * (Nh + 1)
    for i in range(3):
        u = out["z"][i, u0:u0 + 2 * Nh].reshape(Nh, 2)
        assert np.abs(u - rec["u_it0_agent%d" % i]).max() < 0.1
        _kkt_certificate(cfg, track, ins, out, i, Nh, 2)


@pytest.mark.parametrize("Nh,nn,topo,S,A,iters", [(10, 2, "full", 8, 3, 4), (20, 2, "chain", 4, 6, 4), (20, 5, "full", 4, 6, 4),
                                                  (50, 2, "full", 4, 3, 4), (20, 0, "full", 8, 1, 4), (100, 2, "full", 2, 3, 3),
                                                  (20, 15, "full", 2, 16, 3)])
def test_closed_loop_vs_oracle(built, cfg, track, Nh, nn, topo, S, A, iters):
    """DMPC iterations driven by the CUDA path; every iteration's batch is re-solved by the oracle.  Every QP the oracle's
    exact mode certifies must come back certified (status 1) and is compared at 1e-6 with the same active set; a QP the
    oracle's exact mode does not certify (its sparse interior point gives up on some long-horizon QPs with near-dependent
    active rows) may come back 1 or 2 and is held to the independent KKT certificate on the reference-assembled matrices."""
    from colaborativempc_b200 import scenarios as sc
    eng = _engine(cfg, track, N=Nh, nn=nn, max_batch=S * A)
    x0, xprev, uprev = sc.platoon_scenarios(S, A, Nh, cfg["dt"], track, seed0=100)
    idx, cnt = sc.topology(A, topo, nn=nn)
    rec = []
    sc.run_dmpc(lambda **kw: eng.solve_batch_host(**kw), x0, xprev, uprev, idx, cnt, iters, Nh, record=rec)
    assert len(rec) == iters
    n_cmp = 0
    for ins, out in rec:
        o = _oracle(cfg, track, ins, N=Nh, nn=nn)
        ok = o["status"] == 1
        assert ok.mean() >= 0.6
        # every QP the oracle's exact mode certifies must come back certified; status 2 ("solved inaccurate": the
        # interior-point iterate, no working set certified - near-dependent active rows on long horizons) is accepted
        # only where the oracle's exact mode does not certify either
        assert np.all(out["status"][ok] == 1), np.unique(out["status"][ok], return_counts=True)
        assert np.all(np.isin(out["status"][~ok], (1, 2)))
        _compare({k: v[ok] for k, v in out.items()}, o["z"][ok], o["y"][ok], o["obj"][ok], o["planes"][ok], eng.m_in)
        n_cmp += int(ok.sum())
        if nn > 0 and np.all(ins["nbr_count"] == nn):  # independent KKT check on the reference-assembled matrices
            # status 2 = "solved inaccurate", the interior-point iterate of a QP on which the oracle's exact mode gives up as
            # well (multipliers ~1e8: slacks of metres, a platoon the plant model has already lost): KKT residuals within
            # 2e-2 of the multiplier scale, flagged as inaccurate to the caller exactly like OSQP's status 2
            for b in np.flatnonzero(~ok):
                _kkt_certificate(cfg, track, ins, out, b, Nh, nn, tol=TOL if out["status"][b] == 1 else 2e-2)
    assert n_cmp > 0


def _compare_certified(cfg, track, ins, out, Nh, nn, rows=None, min_ok=0.9):
    """Oracle (exact mode) on `rows` of a batch: every QP it certifies must be certified here and equal at 1e-6 with the same
    active set; returns the number compared."""
    rows = np.arange(ins["x0"].shape[0]) if rows is None else np.asarray(rows)
    sub = {k: (None if v is None else np.ascontiguousarray(v[rows])) for k, v in ins.items()}
    o = _oracle(cfg, track, sub, N=Nh, nn=nn)
    ok = o["status"] == 1
    assert ok.mean() >= min_ok, ok.mean()
    got = {k: v[rows][ok] for k, v in out.items()}
    _compare(got, o["z"][ok], o["y"][ok], o["obj"][ok], o["planes"][ok], (8 + nn) * Nh)
    return int(ok.sum())


def test_bench_workload_sample_vs_oracle(built, cfg, track):
    """BASELINE config[1] exactly as bench.py builds it (same generator, same seeds): 256 evenly spaced QPs of the first
    TIMED batch (closed-loop iteration 6 of the 1 366 platoons) against the oracle; the whole batch certified."""
    import bench
    eng = _engine(cfg, track, max_batch=bench.stream_capacity(bench.BATCH))
    stream = bench.closed_loop_stream(lambda **kw: eng.solve_batch_host(duals=False, **kw), bench.BATCH, 4, cfg, track)
    ins = stream[3]
    assert ins["x0"].shape[0] == 4096
    out = eng.solve_batch_host(**ins)
    assert np.all(out["status"] == 1), np.unique(out["status"], return_counts=True)
    assert _compare_certified(cfg, track, ins, out, N, 2, rows=np.arange(0, 4096, 16)) >= 250


@pytest.mark.parametrize("A,topo,nn,S,its", [(16, "chain", 2, 8, (0, 3, 40)), (64, "chain", 2, 2, (0, 3, 40)),
                                             (64, "full", 63, 1, (0, 3)), (16, "full", 15, 2, (0, 3, 12))])
def test_platoon_configs_vs_oracle(built, cfg, track, A, topo, nn, S, its):
    """BASELINE configs 3 and 5: the 16-agent chain the multi-GPU bench times, the 64-vehicle platoon as a chain and fully
    connected (nn = 63), 16 agents fully connected - closed loop on the device (`ShardedDMPC`, warm start and launch
    order live), the batches of iterations `its` re-solved by the oracle."""
    import torch
    from colaborativempc_b200 import scenarios as sc
    from colaborativempc_b200.dmpc import ShardedDMPC
    x0, xprev, uprev = sc.platoon_scenarios(S, A, N, cfg["dt"], track, seed0=8)
    idx, cnt = sc.topology(A, topo, nn=nn)
    eng = _engine(cfg, track, nn=nn, max_batch=S * A)
    dm = ShardedDMPC(eng, S, A, idx, cnt, x0, xprev, uprev, end_s=1e9)
    dm.out = eng.alloc_outputs(dm.B, duals=True)
    n_cmp = 0
    for it in range(max(its) + 1):
        keep = it in its
        if keep:
            ins = dict(x0=dm.x0.cpu().numpy(), xprev=(dm.xprev_first if dm.first else dm.xprev).cpu().numpy(),
                       uprev=dm.uprev.cpu().numpy(), uold=dm.uold.cpu().numpy(), nbr_count=dm.nbr_cnt.cpu().numpy())
        e = dm.eng   # the body of ShardedDMPC.iterate with the duals switched on
        e.gather_neighbours(dm.all, dm.nbr_idx, dm.nbr_xy, dm.pose, dm.S, dm.A, dm.a_local, dm.off, stop=dm.stop)
        e.solve_batch(dm.x0, dm.xprev_first if dm.first else dm.xprev, dm.uprev, dm.uold, dm.nbr_xy, dm.pose, dm.nbr_cnt,
                      out=dm.out, duals=True, shift=dm._shift())
        dm.first = False
        dm.exchange()
        torch.cuda.synchronize()
        st = dm.out["status"].cpu().numpy()
        run = ~np.repeat(dm.stop.cpu().numpy().astype(bool), A)
        assert not np.isin(st[run], (-10, -20, -21)).any()
        if keep:
            ins.update(nbr_xy=dm.nbr_xy.cpu().numpy(), pose=dm.pose.cpu().numpy())
            out = {k: v.cpu().numpy() for k, v in dm.out.items()}
            rows = np.flatnonzero(run & (st == 1))
            assert rows.size >= 0.95 * run.sum()
            n_cmp += _compare_certified(cfg, track, ins, out, N, nn, rows=rows, min_ok=0.85)
    assert n_cmp >= 0.8 * len(its) * S * A


def test_config1_closed_loop_to_check_end(built, cfg, track):
    """BASELINE config 1: the reference's own experiment (planner/scripts/LPV_HP_N_main.py:57-122) at N = 20 - the three
    vehicles of x0_database[0:3], fully connected, from the reference's cold start, driven through the drop-in path until
    checkEnd (utilities/misc.py:28-48) fires - with EVERY iteration's three QPs re-solved by the oracle."""
    from colaborativempc_b200 import scenarios as sc
    from colaborativempc_b200.runlog import check_end
    eng = _engine(cfg, track, max_batch=3, options={"warm_start": 2})
    agents, xs, us = sc.initialise_agents(sc.X0_DATABASE[0:3], N, cfg["dt"], track)
    idx, cnt = sc.topology(3, "full", nn=2)
    L = float(np.ravel(track.TrackLength)[0])
    x0 = np.stack([x[0] for x in xs])
    xprev, uprev, uold = np.stack(xs), np.stack(us), np.zeros((3, 2))
    xy = xprev[None, :, :, 7:9].copy()
    rec, s_now, it = [], x0[None, :, 6], 0
    while it < 2000 and not check_end(s_now, L)[0]:
        nbr, pose = sc.gather_neighbours(xy, idx)
        ins = dict(x0=x0, xprev=xprev, uprev=uprev, uold=uold, nbr_xy=nbr.reshape(3, N + 1, 2, 2), pose=pose.reshape(3, N + 1, 2),
                   nbr_count=np.asarray(cnt, np.int32))
        out = eng.solve_batch_host(**ins)
        assert np.all(np.isin(out["status"], (1, 2))), (it, out["status"])     # feasible throughout, as in the recorded runs
        rec.append((ins, out))
        x0, xprev, uprev, uold, xyb = sc.dmpc_shift(out["z"], N)
        xy = xyb[None]
        s_now = out["z"][:, :12 * (N + 1)].reshape(3, N + 1, 12)[None, :, 0, 6]
        it += 1
    assert 200 < it < 2000 and check_end(s_now, L)[0]          # the run ENDED at the end of the track (47 m at ~3 m/s)
    n = _compare_certified(cfg, track, rec[0][0], rec[0][1], N, 2, min_ok=1.0)   # first call: N + 1 history rows
    big_in = {k: np.concatenate([r[0][k] for r in rec[1:]], 0) for k in INPUT_KEYS}
    big_out = {k: np.concatenate([r[1][k] for r in rec[1:]], 0) for k in rec[0][1]}
    assert (big_out["status"] == 1).mean() > 0.99
    n += _compare_certified(cfg, track, big_in, big_out, N, 2, rows=np.flatnonzero(big_out["status"] == 1), min_ok=0.97)
    assert n > 0.97 * 3 * it


def test_device_and_host_entry_points_agree(built, cfg, track):
    import torch
    fx = golden("platoon_N20.npz")
    eng = _engine(cfg, track, options={"warm_start": 0, "order_hint": 0})  # bit-for-bit: every call starts from the empty guess
    ins = {k: fx["it3_%s" % k] for k in INPUT_KEYS}
    h = eng.solve_batch_host(**ins)
    dev = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in ins.items()}
    d = eng.solve_batch(**dev)
    torch.cuda.synchronize()
    for k in ("z", "y_ineq", "y_eq", "obj", "status", "planes"):
        assert np.array_equal(d[k].cpu().numpy(), h[k]), k
    assert eng.launches == 2  # stage-wise engine: one launch per call, the whole of PlannerLPV.solve


def test_page_locked_host_buffers_are_used_in_place(built, cfg, track):
    """dmpc_options.host_direct: with page-locked buffers the kernel reads/writes host memory in place over PCIe (no
    staging copies); the results are bit-identical with the staged path, also when only some buffers are page-locked
    (the whole call is then staged) and for a view at an offset inside a page-locked allocation."""
    import torch
    fx = golden("platoon_N20.npz")
    ins = {k: fx["it3_%s" % k] for k in INPUT_KEYS}
    staged = _engine(cfg, track, options={"host_direct": 0, "warm_start": 0, "order_hint": 0}).solve_batch_host(**ins)
    eng = _engine(cfg, track, options={"warm_start": 0, "order_hint": 0})  # bit-for-bit: nothing carried between calls
    B = ins["x0"].shape[0]

    def pinned(a, pad=0):  # a copy of `a` living `pad` elements into a page-locked allocation
        t = torch.empty((a.size + pad,), dtype=torch.from_numpy(a).dtype).pin_memory()
        v = t.numpy()[pad:].reshape(a.shape)
        v[...] = a
        return v
    for pad, partial in ((0, False), (3, False), (0, True)):
        p_in = {k: (None if v is None else (np.ascontiguousarray(v) if partial and k == "xprev" else pinned(np.ascontiguousarray(v), pad)))
                for k, v in ins.items()}
        p_out = {k: pinned(np.full_like(v, -7), pad) for k, v in staged.items()}
        got = eng.solve_batch_host(out=p_out, **p_in)
        for k in staged:
            assert np.array_equal(got[k], staged[k]), (k, pad, partial)


def test_cost_ordered_launch_changes_nothing(built, cfg, track):
    """Condensed engine (engine 2).  order_hint only permutes which CTA solves which QP (most expensive of the previous call first): outputs are
    bit-identical with the hint off, on the first call (identity order) and on later calls (sorted order), also when
    the batch size changes between calls."""
    from colaborativempc_b200 import scenarios as sc
    S, A = 64, 6
    x0, xprev, uprev = sc.platoon_scenarios(S, A, N, cfg["dt"], track, seed0=7)
    idx, cnt = sc.topology(A, "chain", nn=2)
    rec = []
    ref = _engine(cfg, track, max_batch=S * A, options={"order_hint": 0, "engine": 2})
    sc.run_dmpc(lambda **kw: ref.solve_batch_host(**kw), x0, xprev, uprev, idx, cnt, 3, N, record=rec)
    eng = _engine(cfg, track, max_batch=S * A, options={"engine": 2})
    keys = ("z", "y_ineq", "y_eq", "obj", "status", "planes", "iters", "pivots")
    for rep in range(2):
        for ins, out in rec:
            got = eng.solve_batch_host(**ins)
            for k in keys:
                assert np.array_equal(got[k], out[k], equal_nan=k not in ("status", "iters", "pivots")), (rep, k)
    half = {k: (v[: S * A // 2] if v is not None else None) for k, v in rec[1][0].items()}
    got = eng.solve_batch_host(**half)
    for k in keys:
        assert np.array_equal(got[k], rec[1][1][k][: S * A // 2], equal_nan=k not in ("status", "iters", "pivots")), k
    # condensed engine: solve (split in two from the second call of a batch size on) + launch-order kernel + the
    # stage-wise pass over what it could not factorise
    assert ref.launches == 6 and eng.launches == 26


def test_drop_in_planner_matches_reference_call_surface(built, cfg, track):
    """PlannerLPV(...).solve(...) as planner/scripts/LPV_HP_N_main.py:41 calls it; attributes of LPV_Planner.py:173-180."""
    from colaborativempc_b200.planner import PlannerLPV
    from colaborativempc_b200 import scenarios as sc
    g = golden("ref_startup_N20.npz")
    ctrls = [PlannerLPV(cfg["Q"], cfg["Qs"], cfg["R"], cfg["dR"], N, cfg["dt"], track, i, cfg["wq"], cfg["model_param"],
                        cfg["sys_lim"]) for i in range(3)]
    agents, x_old, u_old = sc.initialise_agents(sc.X0_DATABASE[0:3], N, cfg["dt"], track)
    ns = [[1, 2], [0, 2], [0, 1]]
    want_obj = (-1396.897816823358, -1490.2691025771865, -1402.3910432237897)
    for it in range(2):
        x_pred, u_pred = [], []
        for i, c in enumerate(ctrls):
            feas, sol, planes = c.solve(x_old[i][0, :], x_old[i], u_old[i], agents[:, ns[i], :], ns[i], agents[:, i, :])
            assert feas == 1 and sol.shape == (16 * N + 12,) and planes.shape == (N, 3, 2)
            assert c.xPred.shape == (N + 1, 9) and c.uPred.shape == (N, 2) and c.sPred.shape == (N, 3)
            assert c.duPred.shape == (N, 2) and c.raw_States.shape == (N + 1, 12)
            if it == 0:
                assert abs(c.obj_val - want_obj[i]) < 1e-6 * abs(want_obj[i])
                assert np.array_equal(planes, g["a%d_planes" % i])
            assert c.OldSteering[0] == c.uPred[0, 0] and c.OldAccelera[0] == c.uPred[0, 1]
            x_pred.append(c.xPred)
            u_pred.append(c.uPred)
        u_old = u_pred
        x_old = [x[1:, :] for x in x_pred]
        agents = np.swapaxes(np.asarray(x_pred)[:, :, -2:], 0, 1)


def test_edge_cases(built, cfg, track):
    """Empty batch; ragged neighbour counts; failed map lookup (the reference raises) and coincident agents."""
    fx = golden("platoon_N20.npz")
    eng = _engine(cfg, track)
    ins = {k: fx["it3_%s" % k][:6].copy() for k in INPUT_KEYS}
    empty = eng.solve_batch_host(**{k: v[:0] for k, v in ins.items()})
    assert empty["z"].shape == (0, 16 * N + 12)
    L = float(np.ravel(track.TrackLength)[0])
    ins["xprev"][1, 5, 6] = L                      # curvature() finds no segment (misc.py:97)
    ins["nbr_xy"][2, 4, 0] = ins["pose"][2, 4]     # compute_plane.py:52 divides by zero
    ins["nbr_count"] = np.array([2, 2, 2, 1, 0, 2], np.int32)
    out = eng.solve_batch_host(**ins)
    assert out["status"][1] == -20 and np.all(np.isnan(out["z"][1]))
    assert out["status"][2] == -21
    keep = np.array([0, 3, 4, 5])
    o = _oracle(cfg, track, {k: v[keep] for k, v in ins.items()})
    sub = {k: v[keep] for k, v in out.items()}
    _compare(sub, o["z"], o["y"], o["obj"], None, eng.m_in)
    # rows of unused neighbour slots carry zero multipliers
    yi = out["y_ineq"][4][:(4 + 2) * N].reshape(N, 6)
    assert np.all(yi[:, 4:] == 0)


def test_input_staging_is_alignment_independent(built, cfg, track):
    """The stage-wise kernel stages the per-QP inputs in shared memory with bulk async copies (TMA) of the 16-byte aligned
    interior of each array and plain loads of an odd head / tail double.  Every alignment class must give the same bits:
    caller pointers that are only 8-byte aligned (views starting one double into a buffer), Last_xPredicted with N and
    with N+1 rows (72-byte rows: with 21 of them odd QPs start 8 bytes off), a single QP."""
    import torch
    fx = golden("platoon_N20.npz")
    eng = _engine(cfg, track, options={"warm_start": 0, "order_hint": 0})  # bit-for-bit: nothing carried between calls
    ins = {k: np.ascontiguousarray(fx["it3_%s" % k]) for k in INPUT_KEYS}
    assert ins["xprev"].shape[1] == 20

    def dev_off(a, off):
        flat = torch.empty(a.size + off, dtype=torch.float64 if a.dtype == np.float64 else torch.int32, device="cuda")
        v = flat[off:].view(*a.shape)
        v.copy_(torch.from_numpy(a))
        return v

    def run(d, off):
        return eng.solve_batch(**{k: dev_off(v, off if v.dtype == np.float64 else 0) for k, v in d.items()})

    ref = {k: v.cpu().numpy() for k, v in run(ins, 0).items()}
    assert np.all(ref["status"] == 1)
    keys = ("z", "y_ineq", "y_eq", "obj", "planes", "status")
    for off in (1, 3):
        d = {k: dev_off(v, off if v.dtype == np.float64 else 0) for k, v in ins.items()}
        assert d["xprev"].data_ptr() % 16 == 8 and d["pose"].data_ptr() % 16 == 8
        out = eng.solve_batch(**d)
        for k in keys:
            assert np.array_equal(out[k].cpu().numpy(), ref[k]), (off, k)
    # N + 1 rows (the extra row only adds its track lookup): 1512-byte slices, every odd QP starts 8 bytes off
    ins21 = dict(ins, xprev=np.ascontiguousarray(np.concatenate([ins["xprev"], ins["xprev"][:, -1:]], axis=1)))
    for off in (0, 1):
        out = run(ins21, off)
        for k in keys:
            assert np.array_equal(out[k].cpu().numpy(), ref[k]), ("21 rows", off, k)
    # one QP alone, from an odd position
    one = {k: np.ascontiguousarray(v[5:6]) for k, v in ins21.items()}
    o = run(one, 1)
    assert np.array_equal(o["z"].cpu().numpy()[0], ref["z"][5]) and np.array_equal(o["y_eq"].cpu().numpy()[0], ref["y_eq"][5])


def test_full_size_batch_properties(built, cfg, track):
    """BASELINE config 2 at full size (4096 scenario QPs): all certified, batch-position independent (bitwise),
    replicas identical, objective = 1/2 z'Pz + q'z recomputed on the host."""
    import torch
    from colaborativempc_b200 import scenarios as sc
    S, A = 64, 3
    eng = _engine(cfg, track, max_batch=4096, options={"warm_start": 0, "order_hint": 0})  # bitwise properties: nothing carried
    x0, xprev, uprev = sc.platoon_scenarios(S, A, N, cfg["dt"], track, seed0=1000)
    idx, cnt = sc.topology(A, "full")
    st = sc.run_dmpc(lambda **kw: eng.solve_batch_host(**kw), x0, xprev, uprev, idx, cnt, 3, N)
    ins = {k: st[k] for k in INPUT_KEYS}
    B0 = S * A
    rep = (4096 + B0 - 1) // B0
    big = {k: np.concatenate([v] * rep, 0)[:4096] for k, v in ins.items()}
    perm = np.random.default_rng(0).permutation(4096)
    out = eng.solve_batch_host(**big)
    outp = eng.solve_batch_host(**{k: v[perm] for k, v in big.items()})
    assert np.all(out["status"] == 1)
    assert np.array_equal(out["z"][perm], outp["z"])
    assert np.array_equal(out["z"][:B0], out["z"][B0:2 * B0])
    o = _oracle(cfg, track, {k: v[:B0] for k, v in big.items()})
    _compare({k: v[:B0] for k, v in out.items()}, o["z"], o["y"], o["obj"], o["planes"], eng.m_in)
    u0 = 12 * (N + 1)
    u = out["z"][:, u0:u0 + 2 * N].reshape(-1, N, 2)
    assert u[:, :, 0].max() <= 0.3 + 1e-9 and u[:, :, 0].min() >= -0.3 - 1e-9
    assert u[:, :, 1].max() <= 5.0 + 1e-9 and u[:, :, 1].min() >= -10.0 - 1e-9


def test_warm_start_and_launch_order_change_the_route_not_the_optimum(built, cfg, track):
    """dmpc_options.warm_start / order_hint: the previous call's certified working sets are the first guesses and its costs
    the launch order.  Replaying a batch then certifies every QP in ONE round without the interior point; a permuted and a
    perturbed batch (stale guesses) still certify; all agree with the cold solve to the certificate tolerance."""
    from colaborativempc_b200 import scenarios as sc
    S, A = 128, 3
    cold = _engine(cfg, track, max_batch=S * A, options={"warm_start": 0, "order_hint": 0})
    x0, xprev, uprev = sc.platoon_scenarios(S, A, N, cfg["dt"], track, seed0=2000)
    idx, cnt = sc.topology(A, "full")
    st = sc.run_dmpc(lambda **kw: cold.solve_batch_host(**kw), x0, xprev, uprev, idx, cnt, 3, N)
    ins = {k: st[k] for k in INPUT_KEYS}
    ref = cold.solve_batch_host(**ins)
    assert np.all(ref["status"] == 1)
    eng = _engine(cfg, track, max_batch=S * A)
    first = eng.solve_batch_host(**ins)
    for k in ("z", "y_ineq", "y_eq", "obj", "status"):
        assert np.array_equal(first[k], ref[k]), k      # first call of a batch size: empty guesses, identity order
    again = eng.solve_batch_host(**ins)
    assert np.all(again["status"] == 1) and np.all(again["pivots"] == 1) and np.all(again["iters"] == 0)
    assert ref["pivots"].mean() > 2.0
    perm = np.random.default_rng(1).permutation(S * A)
    stale = eng.solve_batch_host(**{k: v[perm] for k, v in ins.items()})   # every QP gets another QP's working set
    moved = dict(ins)
    moved["x0"] = ins["x0"] + np.random.default_rng(2).normal(scale=1e-3, size=ins["x0"].shape) * np.array([1, 1, 1, 1, 1, 0, 0, 0, 0.0])
    ref_moved = cold.solve_batch_host(**moved)
    got_moved = eng.solve_batch_host(**moved)
    for got, want in ((again, ref), ({k: v[np.argsort(perm)] for k, v in stale.items()}, ref), (got_moved, ref_moved)):
        assert np.all(got["status"] == 1) and np.all(want["status"] == 1)
        scale = np.abs(np.concatenate([want["y_ineq"], want["y_eq"]], 1)).max(1)
        assert rel_err(got["z"], want["z"]).max() < 1e-8
        assert rel_err(got["obj"][:, None], want["obj"][:, None]).max() < 1e-8
        assert (np.abs(got["y_ineq"] - want["y_ineq"]).max(1) / (1 + scale)).max() < 1e-6
        assert np.array_equal(_active(got["y_ineq"], scale), _active(want["y_ineq"], scale))


def test_device_loop_and_run_log(built, cfg, track, tmp_path):
    """`ShardedDMPC.iterate` (gather -> solve -> shift/pack -> exchange, all on the device) reproduces the
    host-driven loop of LPV_HP_N_main.py:96-117 bit for bit, and `run_closed_loop` logs what the reference's agents log
    (base_class.py:47-62) and stops on the reference's end-of-track rule (misc.py:28-48)."""
    import torch
    from colaborativempc_b200 import scenarios as sc
    from colaborativempc_b200.dmpc import ShardedDMPC
    from colaborativempc_b200.runlog import RunLog, run_closed_loop
    S, A, iters = 4, 3, 5
    x0, xprev, uprev = sc.platoon_scenarios(S, A, N, cfg["dt"], track, seed0=300)
    idx, cnt = sc.topology(A, "full")
    # the host loop on a handle with the same carried state as ShardedDMPC's (warm start moved one stage per step)
    rec, host_eng = [], _engine(cfg, track, max_batch=S * A, options={"warm_start": 2})
    sc.run_dmpc(lambda **kw: host_eng.solve_batch_host(**kw), x0, xprev, uprev, idx, cnt, iters + 3, N, record=rec)
    dm = ShardedDMPC(_engine(cfg, track, max_batch=S * A), S, A, idx, cnt, x0, xprev, uprev)
    for it in range(iters):
        out = dm.iterate()
        torch.cuda.synchronize()
        assert np.array_equal(out["z"].cpu().numpy(), rec[it][1]["z"]), it
        assert np.array_equal(out["status"].cpu().numpy(), rec[it][1]["status"])
    assert not dm.stopped().any() and int(dm.hold.sum()) == 0
    # the same iteration as ONE CUDA graph (gather -> solve + bookkeeping -> launch order -> exchange): same bits
    dm.capture()   # runs one more iteration eagerly (on the capture stream), then records the next one
    torch.cuda.synchronize()
    assert np.array_equal(dm.out["z"].cpu().numpy(), rec[iters][1]["z"])
    for it in range(iters + 1, iters + 3):
        out = dm.iterate()
        torch.cuda.synchronize()
        assert np.array_equal(out["z"].cpu().numpy(), rec[it][1]["z"]), it
    # the logged run: scenario 1, a track "ending" 0.6 m ahead of the leading agent so that the end rule fires
    L = float(x0[1, :, 6].max()) + 0.6
    dm = ShardedDMPC(_engine(cfg, track, max_batch=S * A), S, A, idx, cnt, x0, xprev, uprev, end_s=L)
    logs = [RunLog(a, str(tmp_path) + "/") for a in range(A)]
    n_it, ended, error = run_closed_loop(dm, 200, logs=logs, log_scenario=1)
    assert ended and not error and 0 < n_it < 200
    zs = [r[1]["z"].reshape(S, A, -1)[1] for r in rec]
    for a, lg in enumerate(logs):
        assert len(lg.states) == len(lg.u) == len(lg.look_ahead) == len(lg.time_op) == n_it
        for t in range(min(n_it, iters)):
            xs = zs[t][a][:12 * (N + 1)].reshape(N + 1, 12)
            assert np.array_equal(lg.states[t], xs[0, :9]) and lg.look_ahead[t] == xs[-1, 6] - xs[0, 6]
            assert np.array_equal(lg.u[t], zs[t][a][12 * (N + 1):12 * (N + 1) + 2])
            assert np.array_equal(lg.planes[t], rec[t][1]["planes"].reshape(S, A, N, 3, -1)[1, a, 0])
        path = lg.save_to_csv()
        assert np.loadtxt(path + "/states.dat").reshape(n_it, 9).shape == (n_it, 9)
    s_last = max(lg.states[-1][6] for lg in logs)
    assert s_last < L + 0.5  # stopped at the first iteration whose row 0 passed L - 0.15


def test_failed_solve_holds_the_plan_and_stops_the_scenario(built, cfg, track):
    """LPV_HP_N_main.py:102-104,110-111: the reference stops an experiment at the first QP that is not feasible.  Here a
    failed QP (its map lookup is made to fail, where the reference raises) keeps its last valid plan, nothing non-finite
    enters the exchange, its scenario is flagged and frozen from the next iteration on - on every rank, because the flag
    travels with the all-gather - and the other scenarios run on."""
    import torch
    from colaborativempc_b200 import scenarios as sc
    from colaborativempc_b200.dmpc import ShardedDMPC
    S, A = 6, 3
    x0, xprev, uprev = sc.platoon_scenarios(S, A, N, cfg["dt"], track, seed0=400)
    idx, cnt = sc.topology(A, "full")
    dm = ShardedDMPC(_engine(cfg, track, max_batch=S * A), S, A, idx, cnt, x0, xprev, uprev)
    for _ in range(3):
        dm.iterate()
    assert not dm.stopped().any()
    bad = 2 * A + 1                                   # scenario 2, agent 1
    L = float(np.ravel(track.TrackLength)[0])
    keep = {k: getattr(dm, k).clone() for k in ("x0", "xprev", "uprev", "uold", "xy_send")}
    dm.xprev[bad, 5, 6] = L                           # curvature() finds no segment for s == TrackLength (misc.py:97)
    out = dm.iterate()
    torch.cuda.synchronize()
    st = out["status"].cpu().numpy()
    assert st[bad] == -20 and np.all(np.delete(st, bad) == 1)
    assert int(dm.hold[bad]) == 1 and int(dm.hold.sum()) == 1
    for k in ("x0", "uprev", "uold", "xy_send"):      # the failed agent keeps its last valid plan ...
        assert torch.equal(getattr(dm, k)[bad], keep[k][bad]), k
    assert torch.isfinite(dm.send).all() and torch.isfinite(dm.all).all()   # ... and nothing non-finite is exchanged
    assert dm.scenario_flags(2) == (False, True) and dm.scenario_flags(1) == (False, False)
    stop = dm.stopped()
    assert stop.tolist() == [False, False, True, False, False, False]
    frozen = {k: getattr(dm, k).clone() for k in ("x0", "xprev", "uprev", "uold", "xy_send")}
    for _ in range(3):
        out = dm.iterate()
    torch.cuda.synchronize()
    rows = slice(2 * A, 3 * A)
    for k, v in frozen.items():                       # the stopped scenario does not move any more, the others do
        assert torch.equal(getattr(dm, k)[rows], v[rows]), k
        assert not torch.equal(getattr(dm, k)[:A], v[:A]), k
    st = out["status"].cpu().numpy()
    assert np.all(np.delete(st, np.arange(2 * A, 3 * A)) == 1)
    assert dm.stopped().tolist() == [False, False, True, False, False, False]


def test_closed_loop_infeasible_qps_are_infeasible(built, cfg, track):
    """The 256 x 16 closed loop the bench times does produce QPs that are not feasible (crawling vehicles whose linearised
    lateral model has spun out: hard rows vx >= min_vel against the dynamics).  Every such verdict (-3) is checked
    against the oracle's exact mode, which must find the same QP infeasible; no QP may end in a numerical breakdown
    (-10), and every other QP of the sampled iterations is certified."""
    import torch
    from colaborativempc_b200 import scenarios as sc
    from colaborativempc_b200.dmpc import ShardedDMPC
    S, A = 256, 16
    x0, xprev, uprev = sc.platoon_scenarios(S, A, N, cfg["dt"], track, seed0=0)
    idx, cnt = sc.topology(A, "chain", nn=2)
    eng = _engine(cfg, track, max_batch=S * A)
    dm = ShardedDMPC(eng, S, A, idx, cnt, x0, xprev, uprev, end_s=1e9)
    seen = 0
    for it in range(45):
        ins = dict(x0=dm.x0.clone(), xprev=(dm.xprev_first if dm.first else dm.xprev).clone(), uprev=dm.uprev.clone(), uold=dm.uold.clone())
        out = dm.iterate()
        torch.cuda.synchronize()
        st = out["status"].cpu().numpy()
        run = ~np.repeat(dm.stop.cpu().numpy().astype(bool), A)
        assert not np.isin(st[run], (-10, -20, -21)).any(), np.unique(st[run], return_counts=True)
        assert np.all(np.isin(st[run], (1, 2, -3)))
        for b in np.flatnonzero(run & (st == -3)):
            one = {k: v[b:b + 1].cpu().numpy() for k, v in ins.items()}
            one.update(nbr_xy=dm.nbr_xy[b:b + 1].cpu().numpy(), pose=dm.pose[b:b + 1].cpu().numpy(), nbr_count=dm.nbr_cnt[b:b + 1].cpu().numpy())
            o = _oracle(cfg, track, one, nthreads=1)
            assert o["status"][0] == -3, (it, b, o["status"])
            seen += 1
    assert seen >= 1 and int(dm.stopped().sum()) == seen   # each infeasible QP stopped exactly its own scenario


def test_model_rollout_vs_reference_and_oracle(built, cfg, track):
    """SURVEY 8(f2), LPV_Model.sim (LPV_val.py:77-94): the CUDA roll-out through the C ABI against (1) the reference's own
    outputs (golden, 18 cases), (2) the oracle on 1024 synthetic platoon predictions, (3) the drop-in class chained like
    scripts/eval_LPV_model.py:50-56, (4) a failed map lookup.  FP64; sincos and FMA contraction differ from numpy's
    libm / matmul order: tolerance 1e-12 relative (achieved ~1e-15)."""
    import torch
    from colaborativempc_b200 import scenarios as sc
    from colaborativempc_b200._abi import make_params
    from colaborativempc_b200.planner import LPV_Model
    from oracle import lpv_oracle as lo
    g = golden("ref_lpv_model_sim_N20.npz")
    eng = _engine(cfg, track, nn=0, max_batch=1)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    out, st = eng.sim_batch(dev(g["xpred"]), dev(g["u"]))
    assert np.all(st.cpu().numpy() == 0)
    assert np.abs(out.cpu().numpy() - g["out"]).max() <= 1e-12 * (1 + np.abs(g["out"]).max())
    # (2) a batch: states of synthetic platoons, inputs perturbed
    S, A = 256, 4
    x0, xprev, uprev = sc.platoon_scenarios(S, A, N, cfg["dt"], track, seed0=5)
    xp = np.ascontiguousarray(xprev.reshape(S * A, N + 1, 9))
    rng = np.random.default_rng(3)
    u = np.ascontiguousarray(uprev.reshape(S * A, N, 2) + rng.uniform(-0.05, 0.05, (S * A, N, 2)))
    P = make_params(*planner_args(cfg, track, N), nn=0)
    ref, rst = lo.sim_batch(P, xp, u)
    out, st = eng.sim_batch(dev(xp), dev(u))
    assert np.array_equal(st.cpu().numpy(), rst) and np.all(rst == 0)
    assert np.abs(out.cpu().numpy() - ref).max() <= 1e-12 * (1 + np.abs(ref).max())
    # (2b) the tile paths of the kernel: a ragged last tile whose byte range is not a multiple of 16 (plain loads / stores
    # instead of the bulk copies) and caller pointers that are only 8-byte aligned (a view starting at model 1)
    full = out.cpu().numpy()
    xd, ud = dev(xp), dev(u)
    out, st = eng.sim_batch(xd[:1021], ud[:1021])
    assert np.all(st.cpu().numpy() == 0) and np.array_equal(out.cpu().numpy(), full[:1021])
    out, st = eng.sim_batch(xd[1:1000], ud[1:1000])
    assert xd[1:].data_ptr() % 16 == 8
    assert np.all(st.cpu().numpy() == 0) and np.array_equal(out.cpu().numpy(), full[1:1000])
    # (3) drop-in class, chained calls
    m = LPV_Model(N, cfg["dt"], track, model_param=cfg["model_param"], sys_lim=cfg["sys_lim"])
    m.xPred = g["xpred"][0].copy()
    for call in range(6):
        ret = m.sim(g["u"][call])
        assert np.abs(ret - g["out"][call]).max() <= 1e-12 * (1 + np.abs(g["out"][call]).max())
    # (4) no segment for a state's s: the reference raises inside curvature()
    bad = g["xpred"][:2].copy()
    bad[1, 7, 6] = np.nan
    out, st = eng.sim_batch(dev(bad), dev(g["u"][:2]))
    st = st.cpu().numpy()
    assert st[0] == 0 and st[1] == -20 and np.all(np.isnan(out[1].cpu().numpy())) and not np.isnan(out[0].cpu().numpy()).any()


def test_ocd_dual_update_bit_exact(built, cfg, track):
    """the CUDA dual update through the C ABI, bit-exact against the reference-generated golden (N=20, the
    handle's horizon), against the oracle on a larger batch, and a single-agent batch (no pair: nothing changes)."""
    import torch
    from oracle import lpv_oracle as lo
    g = golden("ref_ocd_lambda.npz")
    eng = _engine(cfg, track, nn=0, max_batch=1)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    lam = dev(g["a_lam0"])
    xy = dev(g["a_xy"])
    eng.ocd_lambda_update(lam, xy, float(g["a_dth"]))
    assert np.array_equal(lam.cpu().numpy(), g["a_lam1"])
    eng.ocd_lambda_update(lam, xy, float(g["a_dth"]), alpha=float(g["alpha"]))
    assert np.array_equal(lam.cpu().numpy(), g["a_lam2"])
    rng = np.random.default_rng(9)
    S, A = 300, 7
    xy = rng.normal(size=(S, A, N + 1, 2))
    lam0 = rng.uniform(-1, 1, size=(S, A, A, N))
    ref = lo.ocd_lambda_update(lam0, xy, 0.25, 0.3)
    got = eng.ocd_lambda_update(dev(lam0), dev(xy), 0.3).cpu().numpy()
    assert np.array_equal(got, ref)
    one = dev(rng.uniform(size=(4, 1, 1, N)))
    keep = one.clone()
    eng.ocd_lambda_update(one, dev(rng.normal(size=(4, 1, N + 1, 2))), 0.3)
    assert torch.equal(one, keep)
    # hyperplane form (NL_HP_N_main.py:128-137): reference-generated golden, then a larger batch against the oracle
    lam = dev(g["hp_a_lam0"])
    xy, pl = dev(g["hp_a_xy"]), dev(g["hp_a_planes"])
    eng.ocd_lambda_update_hp(lam, xy, pl, float(g["hp_a_dth"]))
    assert np.array_equal(lam.cpu().numpy(), g["hp_a_lam1"])
    eng.ocd_lambda_update_hp(lam, xy, pl, float(g["hp_a_dth"]), alpha=float(g["alpha"]))
    assert np.array_equal(lam.cpu().numpy(), g["hp_a_lam2"])
    xy = rng.normal(size=(S, A, N + 1, 2))
    pl = rng.normal(size=(S, A, N, 3, A))
    ref = lo.ocd_lambda_update_hp(lam0, xy, pl, 0.25, 0.3)
    got = eng.ocd_lambda_update_hp(dev(lam0), dev(xy), dev(pl), 0.3).cpu().numpy()
    assert np.array_equal(got, ref)

"""CPU: the oracle (oracle/lpv_oracle.c) against the golden vectors generated from the REFERENCE's own functions
(tests/golden/make_golden.py imports /root/reference unmodified) and against its own committed exact solutions."""
import numpy as np
import pytest

from conftest import INPUT_KEYS, golden, planner_args, rel_err
from colaborativempc_b200._abi import make_params, map_table
from colaborativempc_b200 import scenarios as sc
from colaborativempc_b200.track import TrackMap
from oracle import lpv_oracle as lo

N = 20


@pytest.fixture(scope="module")
def P(cfg, track):
    return make_params(*planner_args(cfg, track, N), nn=2)


def test_track_table_matches_reference_map():
    """TrackMap reproduces Map(...).PointAndTangent / halfWidth (track_initialization.py:98-300)."""
    g = golden("ref_map_tables.npz")
    for name in ("Highway", "oval", "oval_mt", "TestOpenMap"):
        tr = TrackMap(name)
        pt = g[name + "_pt"]
        mine = np.asarray(tr.PointAndTangent)
        assert mine.shape[:2] == pt.shape[:2]
        np.testing.assert_allclose(mine[:, :, 0], pt[:, :, 0], rtol=0, atol=1e-12)
        np.testing.assert_allclose(np.ravel(tr.TrackLength)[0], np.ravel(g[name + "_len"])[0], atol=1e-12)
        hw = g[name + "_hw"]
        np.testing.assert_allclose(np.asarray(tr.halfWidth)[:len(hw)], hw, atol=0)


def test_startup_prediction_matches_reference(cfg, track):
    """initialise_agents / predicted_vectors_generation (misc.py:155-210)."""
    g = golden("ref_startup_pred.npz")
    for n in (20, 125):
        agents, xs, us = sc.initialise_agents(sc.X0_DATABASE[0:3], n, cfg["dt"], track)
        np.testing.assert_allclose(agents, g["agents_N%d" % n], rtol=0, atol=1e-12)
        np.testing.assert_allclose(np.asarray(xs), g["x_N%d" % n], rtol=0, atol=1e-12)
        assert all(np.all(u == 0) for u in us)


@pytest.mark.parametrize("tag", ["a0", "a1", "a2", "b0", "b1", "b2"])
def test_assembly_bit_exact_vs_reference(P, tag):
    """planes, weights, A_k, B_k, ey, P, q, A, l, u exactly as the reference builds them
    (LPV_Planner.py:138-157, 222-228).  'a*' = first call (N+1 history rows, uOld=0), 'b*' = later call
    (N rows, non-zero inputs, low-speed branch hit)."""
    g = golden("ref_startup_N20.npz")
    i = int(tag[1])
    ns = g["ns"][i]
    nbr, pose = g["agents"][:, ns, :], g["agents"][:, i, :]
    if tag[0] == "a":
        x0, xs, us, uold = g["x_old"][i][0], g["x_old"][i], g["u_old"][i], np.zeros(2)
    else:
        xs, us, uold = g[tag + "_xs"], g[tag + "_us"], g[tag + "_uold"]
        x0 = xs[0]
    o = lo.assemble_dense(P, x0, xs, us, uold, nbr, pose)
    assert np.array_equal(o["planes"], g[tag + "_planes"])
    assert np.array_equal(o["weights"], g[tag + "_weights"])
    assert np.array_equal(o["Ak"], g[tag + "_Ak"])
    assert np.array_equal(o["Bk"], g[tag + "_Bk"])
    assert np.array_equal(o["ey"], g[tag + "_ey"][:N])
    assert np.array_equal(np.diag(o["P"]), g[tag + "_Pdiag"])
    assert np.count_nonzero(o["P"] - np.diag(np.diag(o["P"]))) == 0
    assert np.array_equal(o["q"], g[tag + "_q"])
    assert np.array_equal(o["A"], g[tag + "_A"])
    assert np.array_equal(o["u"], g[tag + "_u"])
    lg = g[tag + "_l"]
    fin = np.isfinite(lg)
    assert np.array_equal(o["l"][fin], lg[fin])
    assert np.all(o["l"][~fin] <= -1e20)


def test_qp_sizes(P):
    g = golden("ref_startup_N20.npz")
    assert g["a0_A"].shape == ((8 + 2) * N + 14 * N + 12, 16 * N + 12)  # m = m_in + m_eq, n = 16N+12


def test_exact_oracle_reproduces_committed_solutions(P):
    """Determinism of the exact mode + KKT certificate of every committed solution."""
    fx = golden("platoon_N20.npz")
    for it in (0, 3):
        ins = {k: fx["it%d_%s" % (it, k)] for k in INPUT_KEYS}
        o = lo.solve_batch(P, st=lo.settings(exact=True), nthreads=4, **ins)
        assert np.all(o["status"] == 1)
        assert o["cert"].max() < 1e-8
        assert rel_err(o["z"], fx["it%d_out_z" % it]).max() < 1e-10
        assert rel_err(o["y"], fx["it%d_out_y" % it]).max() < 1e-8
        assert rel_err(o["obj"][:, None], fx["it%d_out_obj" % it][:, None]).max() < 1e-12


def test_startup_known_answers(P):
    """SURVEY 8(c) probe facts: objective and active-set size of the shipped 3-agent start-up case at N=20."""
    g = golden("ref_startup_N20.npz")
    want_obj = (-1396.878194, -1490.269101, -1402.391041)
    want_act = (17, 1, 0)
    for i in range(3):
        ns = g["ns"][i]
        o = lo.solve_batch(P, g["x_old"][i][0][None], g["x_old"][i][None], g["u_old"][i][None], np.zeros((1, 2)),
                           g["agents"][:, ns, :][None], g["agents"][:, i, :][None], st=lo.settings(exact=True))
        assert o["status"][0] == 1 and o["cert"].max() < 1e-8
        # the survey probe's own active-set solve was good to ~1e-5 relative on agent 0
        assert abs(o["obj"][0] - want_obj[i]) < 2e-5 * abs(want_obj[i])
        y_in = o["y"][0, :o["m_in"]]
        assert int((y_in > 1e-9 * (1 + np.abs(o["y"]).max())).sum()) == want_act[i]
        # independent KKT certificate on the matrices the REFERENCE assembled (golden), in numpy
        z, y = o["z"][0], o["y"][0]
        Pd, q, A, l, u = (g["a%d_%s" % (i, k)] for k in ("Pdiag", "q", "A", "l", "u"))
        Az = A @ z
        assert max((Az - u).max(), (l - Az).max()) < 1e-10
        assert np.abs(Pd * z + q + A.T @ y).max() < 1e-8 * (1 + np.abs(q).max())
        assert y_in.min() >= 0.0 and np.abs(y_in * (Az - u)[:o["m_in"]]).max() < 1e-8
        assert abs(0.5 * z @ (Pd * z) + q @ z - o["obj"][0]) < 1e-9 * abs(o["obj"][0])


def test_osqp_restatement_agrees_with_exact_when_polished(P):
    """Reference settings (eps 1e-3, polish on, LPV_Planner.py:233): a successful polish lands on the exact optimum;
    an unsuccessful one stays within OSQP's loose tolerance (this is what the recorded runs show)."""
    fx = golden("platoon_N20.npz")
    ins = {k: fx["it3_%s" % k][:12] for k in INPUT_KEYS}
    o = lo.solve_batch(P, st=lo.settings(), nthreads=4, **ins)
    zg = fx["it3_out_z"][:12]
    assert np.all(np.isin(o["status"], (1, 2)))
    e = rel_err(o["z"], zg)
    ok = o["polish"] == 1
    assert ok.any()
    # OSQP's "polish successful" only means the polished residuals improved: the guessed active set is usually,
    # not always, the optimal one
    assert np.median(e[ok]) < 1e-9 and e[ok].max() < 1e-3
    u0 = 12 * (N + 1)
    eu = np.abs(o["z"][:, u0:u0 + 2 * N] - zg[:, u0:u0 + 2 * N]).max(1)
    assert eu.max() < 2.0  # loose iterate: bounded by the input box (|delta|<=0.3, a in [-10,5]) scale


def test_c_and_numpy_osqp_agree(P):
    """Two independent restatements of the OSQP algorithm (C sparse LDL vs numpy dense) on the same QP."""
    from oracle import osqp_np
    g = golden("ref_startup_N20.npz")
    Pm, q, A, l, u = np.diag(g["a1_Pdiag"]), g["a1_q"], g["a1_A"], g["a1_l"], g["a1_u"]
    l = np.where(np.isfinite(l), l, -1e30)
    a = lo.osqp_dense(Pm, q, A, l, u, st=lo.settings(exact=True))
    assert a["status"] == 1 and max(a["cert"]) < 1e-8
    b = osqp_np.solve_exact(Pm, q, A, l, u)
    assert max(b.cert) < 1e-6
    assert np.abs(b.x - a["x"]).max() / (1 + np.abs(a["x"]).max()) < 1e-7
    assert abs(b.obj_val - a["obj"]) < 1e-8 * abs(a["obj"])
    assert np.array_equal(osqp_np.active_set(A, l, u, b.x, b.y), osqp_np.active_set(A, l, u, a["x"], a["y"]))


def test_map_lookup_edge_cases(P, cfg, track):
    """curvature()/get_ey() semantics (misc.py:78-126): s == TrackLength finds no segment -> status -20;
    s < 0 clamps to 0; s > L wraps."""
    fx = golden("platoon_N20.npz")
    ins = {k: fx["it3_%s" % k][:1].copy() for k in INPUT_KEYS}
    L = float(np.ravel(track.TrackLength)[0])
    bad = {k: v.copy() for k, v in ins.items()}
    bad["xprev"][0, 5, 6] = L
    o = lo.solve_batch(P, st=lo.settings(exact=True), **bad)
    assert o["status"][0] == -20 and np.all(np.isnan(o["z"]))
    neg = {k: v.copy() for k, v in ins.items()}
    neg["xprev"][0, 2, 6] = -3.0
    clamp = {k: v.copy() for k, v in ins.items()}
    clamp["xprev"][0, 2, 6] = 0.0
    o1 = lo.solve_batch(P, st=lo.settings(exact=True), **neg)
    o2 = lo.solve_batch(P, st=lo.settings(exact=True), **clamp)
    assert o1["status"][0] == 1 and np.array_equal(o1["z"], o2["z"])
    s0, ln, cv, hw = map_table(track)
    assert len(s0) == 10 and ln[0] == 0.0 and ln[-1] == 0.0  # Highway has two zero-length segments


def test_model_rollout_matches_reference_lpv_model(P):
    """oracle restatement of LPV_Model.sim against the REFERENCE's own LPV_Model.sim (fixture made by
    tests/golden/make_golden_sim.py from the unmodified reference): 3 agents x 6 chained calls.  Only the summation
    order of the 9-term row products differs (numpy matmul vs a plain loop): 1e-13 relative."""
    g = golden("ref_lpv_model_sim_N20.npz")
    out, st = lo.sim_batch(P, g["xpred"], g["u"])
    assert np.all(st == 0)
    assert np.abs(out - g["out"]).max() <= 1e-13 * (1 + np.abs(g["out"]).max())
    assert np.all(out[:, -1, :] == 1.0)                       # np.vstack(..., np.ones((1, 9)))  (LPV_val.py:92)
    assert np.array_equal(g["xpred"][1], g["out"][0])         # the fixture chains: next call's xPred = previous return
    bad = g["xpred"][:1].copy()
    bad[0, 5, 6] = np.nan                                     # curvature() finds no segment: the reference raises
    out, st = lo.sim_batch(P, bad, g["u"][:1])
    assert st[0] == -20 and np.all(np.isnan(out))


def test_ocd_dual_update_bit_exact_vs_reference():
    """oracle restatement of the dual update of NL_EU_N_main.py:122-138 against the loop executed with the
    REFERENCE's eval_constraintEU / get_alpha (tests/golden/make_golden_ocd.py), two chained updates, bit-exact."""
    g = golden("ref_ocd_lambda.npz")
    alpha = float(g["alpha"])
    assert alpha == 0.25
    for t in "ab":
        l1 = lo.ocd_lambda_update(g[t + "_lam0"], g[t + "_xy"], alpha, float(g[t + "_dth"]))
        l2 = lo.ocd_lambda_update(l1, g[t + "_xy"], alpha, float(g[t + "_dth"]))
        assert np.array_equal(l1, g[t + "_lam1"]) and np.array_equal(l2, g[t + "_lam2"])
        A = l1.shape[1]
        low = np.tril_indices(A)                                   # i >= j: untouched (the reference adds zero)
        assert np.array_equal(l2[:, low[0], low[1], :], g[t + "_lam0"][:, low[0], low[1], :])


def test_ocd_dual_update_hyperplane_form_bit_exact_vs_reference():
    """Hyperplane form (NL_HP_N_main.py:128-137, eval_constraintHp): same fixture script, same bar."""
    g = golden("ref_ocd_lambda.npz")
    alpha = float(g["alpha"])
    for t in ("hp_a", "hp_b"):
        l1 = lo.ocd_lambda_update_hp(g[t + "_lam0"], g[t + "_xy"], g[t + "_planes"], alpha, float(g[t + "_dth"]))
        l2 = lo.ocd_lambda_update_hp(l1, g[t + "_xy"], g[t + "_planes"], alpha, float(g[t + "_dth"]))
        assert np.array_equal(l1, g[t + "_lam1"]) and np.array_equal(l2, g[t + "_lam2"])
        low = np.tril_indices(l1.shape[1])
        assert np.array_equal(l2[:, low[0], low[1], :], g[t + "_lam0"][:, low[0], low[1], :])


def test_restated_osqp_reproduces_the_reference_recorded_run(cfg, track):
    """The one solver-level output the reference ships: uPred of the recorded 3-agent run LPV3r_agent_laptop (N=125,
    pck/<agent>/u.pkl, written from PlannerLPV.solve -> real OSQP; fixture recorded_N125_it0.npz holds iterations 0, 1).
    The restated ADMM (Ruiz scaling, per-row rho, adaptive rho, scaled termination at eps 1e-3) started from the
    reference's cold start lands on the recorded iterate: <= 1.3e-7 at iteration 0 for all three agents, and after the
    caller's shift (LPV_HP_N_main.py:106-117) <= 2.1e-4 at iteration 1 (39 375 / 150 / 55 150 ADMM iterations).  The
    recorded points are un-polished iterates (6e-3 constraint violation, 5e-2 away from the QP's optimum), so the
    comparison runs with polish off; with polish on, agents 0 and 1 at iteration 0 are rejected polishes and identical."""
    from colaborativempc_b200 import scenarios as sc
    from colaborativempc_b200._abi import make_params
    rec = golden("recorded_N125_it0.npz")
    Nh = 125
    P = make_params(cfg["Q"], cfg["Qs"], cfg["R"], cfg["dR"], Nh, cfg["dt"], track, cfg["wq"], cfg["model_param"],
                    cfg["sys_lim"], nn=2)
    agents, xs, us = sc.initialise_agents(sc.X0_DATABASE[0:3], Nh, cfg["dt"], track)
    idx, cnt = sc.topology(3, "full", nn=2)
    st = lo.settings(polish=0)
    log = []
    sc.run_dmpc(lambda **k: lo.solve_batch(P, st=st, nthreads=3, **k), np.stack([x[0] for x in xs])[None],
                np.stack(xs)[None], np.stack(us)[None], idx, cnt, 2, Nh, record=log)
    u0 = 12 * (Nh + 1)
    for it, tol in ((0, 1e-6), (1, 1e-3)):
        out = log[it][1]
        assert np.all(out["status"] == 1)
        for i in range(3):
            u = out["z"][i, u0:u0 + 2 * Nh].reshape(Nh, 2)
            assert np.abs(u - rec["u_it%d_agent%d" % (it, i)]).max() < tol, (it, i)
    ins = log[0][0]
    o = lo.solve_batch(P, st=lo.settings(), nthreads=3, **ins)
    for i in (0, 1):
        assert o["polish"][i] == -1
        assert np.abs(o["z"][i, u0:u0 + 2 * Nh].reshape(Nh, 2) - rec["u_it0_agent%d" % i]).max() < 1e-6

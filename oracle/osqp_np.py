"""TEST INFRASTRUCTURE ONLY (oracle/): dense numpy restatement of the OSQP algorithm.

The reference's QP arithmetic lives in the third-party package ``osqp>=0.6.2.post5``
(/root/reference/requirements.txt:5; call site LPV_Planner.py:222-249), which is neither
vendored nor installable here.  This file restates the *published* algorithm
(Stellato et al., "OSQP: an operator splitting solver for quadratic programs", 2020, and the
0.6.x documentation defaults) with dense linear algebra so it can be used for small cases:

  Ruiz equilibration (10 passes, cost scaling) -> ADMM with sigma=1e-6, alpha=1.6, rho=0.1,
  rho*1e3 on equality rows, adaptive rho (residual-ratio rule, 5x tolerance),
  termination on unscaled residuals every 25 iterations, polish (delta=1e-6, 3 refinements).

OSQP's time-based adaptive-rho interval is replaced by a fixed interval (default 25 = one
termination-check period, what the time rule rounds to on small problems): **parity is
unpinned at the iterate level**; only converged / polished solutions are compared, and every
solution is independently certified by `kkt_certificate` below.

It also provides `solve_exact`: ADMM to tight tolerance + active-set KKT solve + certificate,
the parity anchor for the CUDA path.  Slow (dense) - the C restatement in lpv_oracle.c is the
timed CPU baseline; this file cross-checks it.
"""
import numpy as np
import scipy.linalg as sla

OSQP_INFTY = 1e30
MIN_SCALING, MAX_SCALING = 1e-4, 1e4
RHO_MIN, RHO_MAX, RHO_TOL, RHO_EQ_FACTOR = 1e-6, 1e6, 1e-4, 1e3

SOLVED, SOLVED_INACCURATE, MAX_ITER, PRIMAL_INF, DUAL_INF = 1, 2, -2, -3, -4


def _limit(v):
    v = np.where(v < MIN_SCALING, 1.0, v)
    return np.minimum(v, MAX_SCALING)


def ruiz(P, q, A, l, u, passes=10):
    n, m = P.shape[0], A.shape[0]
    P, q, A = P.copy(), q.copy(), A.copy()
    D, E, c = np.ones(n), np.ones(m), 1.0
    for _ in range(passes):
        dn = np.maximum(np.abs(P).max(axis=0), np.abs(A).max(axis=0) if m else 0.0)
        en = np.abs(A).max(axis=1) if m else np.zeros(0)
        dt = 1.0 / np.sqrt(_limit(dn))
        et = 1.0 / np.sqrt(_limit(en))
        P = dt[:, None] * P * dt[None, :]
        A = et[:, None] * A * dt[None, :]
        q = dt * q
        D *= dt
        E *= et
        cn = float(_limit(np.array([np.abs(P).max(axis=0).mean()]))[0])
        qn = float(_limit(np.array([np.abs(q).max()]))[0])
        ct = 1.0 / max(cn, qn)
        P *= ct
        q *= ct
        c *= ct
    return P, q, A, l * E, u * E, D, E, c


class Result(object):
    pass


def solve(P, q, A, l, u, eps_abs=1e-3, eps_rel=1e-3, max_iter=500000, polish=True, rho0=0.1,
          sigma=1e-6, alpha=1.6, check=25, adapt_interval=25, scaling=10, delta=1e-6, refine=3):
    n, m = P.shape[0], A.shape[0]
    l = np.maximum(l, -OSQP_INFTY)
    u = np.minimum(u, OSQP_INFTY)
    Ps, qs, As, ls, us, D, E, c = ruiz(P, q, A, l, u, scaling)
    eq = (us - ls) < RHO_TOL
    loose = (ls <= -OSQP_INFTY * MIN_SCALING) & (us >= OSQP_INFTY * MIN_SCALING)

    def rho_vec(rho):
        r = np.full(m, rho)
        r[eq] = RHO_EQ_FACTOR * rho
        r[loose] = RHO_MIN
        return r

    def factor(rv):
        K = np.block([[Ps + sigma * np.eye(n), As.T], [As, -np.diag(1.0 / rv)]])
        return sla.lu_factor(K)

    rho = rho0
    rv = rho_vec(rho)
    lu = factor(rv)
    x, z, y = np.zeros(n), np.zeros(m), np.zeros(m)
    res = Result()
    res.rho_updates = 0
    status = MAX_ITER
    it = 0
    for it in range(1, max_iter + 1):
        xp, zp = x, z
        sol = sla.lu_solve(lu, np.concatenate([sigma * xp - qs, zp - y / rv]))
        xt = sol[:n]
        zt = zp + (sol[n:] - y) / rv
        x = alpha * xt + (1 - alpha) * xp
        zr = alpha * zt + (1 - alpha) * zp
        z = np.clip(zr + y / rv, ls, us)
        y = y + rv * (zr - z)
        do_check = (it % check == 0)
        do_adapt = adapt_interval and (it % adapt_interval == 0)
        if not (do_check or do_adapt):
            continue
        Ax, Px, Aty = As @ x, Ps @ x, As.T @ y
        if do_check:
            pri = np.abs((Ax - z) / E).max() if m else 0.0
            dua = np.abs((Px + qs + Aty) / D).max() / c
            ep = eps_abs + eps_rel * max(np.abs(Ax / E).max(), np.abs(z / E).max()) if m else eps_abs
            ed = eps_abs + eps_rel * max(np.abs(Px / D).max(), np.abs(Aty / D).max(), np.abs(qs / D).max()) / c
            if pri <= ep and dua <= ed:
                status = SOLVED
                break
            # infeasibility certificates (OSQP eps_prim_inf = eps_dual_inf = 1e-4)
            dy = y - getattr(res, "_yprev", np.zeros(m))
            dx = x - xp
            # (kept minimal: the LPV QPs are always feasible thanks to the slacks)
        if do_adapt:
            pr = np.abs(z - Ax).max() / (max(np.abs(z).max(), np.abs(Ax).max()) + 1e-10)
            dr = np.abs(Px + qs + Aty).max() / (max(np.abs(qs).max(), np.abs(Aty).max(), np.abs(Px).max()) + 1e-10)
            rho_new = min(max(rho * np.sqrt(pr / (dr + 1e-10)), RHO_MIN), RHO_MAX)
            if rho_new > 5 * rho or rho_new < rho / 5:
                rho = rho_new
                rv = rho_vec(rho)
                lu = factor(rv)
                res.rho_updates += 1
    res.iter = it
    res.status_val = status
    Ax, Px, Aty = As @ x, Ps @ x, As.T @ y
    pri = np.abs((Ax - z) / E).max() if m else 0.0
    dua = np.abs((Px + qs + Aty) / D).max() / c
    res.status_polish = 0
    if polish and status == SOLVED:
        low = (z - ls) < -y
        upp = (~low) & ((us - z) < y)
        act = low | upp
        Ar = As[act]
        na = Ar.shape[0]
        Kt = np.block([[Ps, Ar.T], [Ar, np.zeros((na, na))]])
        Kr = Kt + np.diag(np.concatenate([delta * np.ones(n), -delta * np.ones(na)]))
        rhs = np.concatenate([-qs, np.where(low, ls, us)[act]])
        f = sla.lu_factor(Kr)
        s = sla.lu_solve(f, rhs)
        for _ in range(refine):
            s = s + sla.lu_solve(f, rhs - Kt @ s)
        xpol = s[:n]
        ypol = np.zeros(m)
        ypol[act] = s[n:]
        zpol = np.clip(As @ xpol, ls, us)
        ppri = np.abs((As @ xpol - zpol) / E).max() if m else 0.0
        pdua = np.abs((Ps @ xpol + qs + As.T @ ypol) / D).max() / c
        ok = (ppri < pri and pdua < dua) or (ppri < pri and dua < 1e-10) or (pdua < dua and pri < 1e-10)
        if ok:
            x, y, z, pri, dua = xpol, ypol, zpol, ppri, pdua
            res.status_polish = 1
        else:
            res.status_polish = -1
    res.x = D * x
    res.y = E * y / c
    res.z = z / E
    res.pri_res, res.dua_res = pri, dua
    res.obj_val = 0.5 * res.x @ (P @ res.x) + q @ res.x
    return res


def kkt_certificate(P, q, A, l, u, x, y):
    """Independent optimality certificate for a convex QP: returns (primal infeasibility,
    stationarity residual, complementarity/sign violation), all absolute inf-norms."""
    Ax = A @ x
    pinf = max(0.0, float(np.max(np.maximum(l - Ax, Ax - u)))) if len(l) else 0.0
    stat = float(np.abs(P @ x + q + A.T @ y).max())
    tol = 1e-7 * (1.0 + np.abs(Ax))
    at_l = np.abs(Ax - l) <= tol
    at_u = np.abs(Ax - u) <= tol
    bad = np.where(y > 0, np.where(at_u, 0.0, y), np.where(at_l, 0.0, -y))
    return pinf, stat, float(bad.max()) if len(bad) else 0.0


def active_set(A, l, u, x, y, ytol=1e-9):
    """+1 upper-active, -1 lower-active, 0 inactive, for INEQUALITY rows (l<u); equality rows 2."""
    s = np.zeros(len(l), dtype=np.int8)
    s[y > ytol] = 1
    s[y < -ytol] = -1
    s[(u - l) < RHO_TOL] = 2
    return s


def solve_exact(P, q, A, l, u, max_rounds=8):
    """Exact optimum: ADMM at tight eps -> active-set guess -> KKT solve with refinement,
    repeated with primal-dual active-set corrections until the KKT certificate holds."""
    n, m = P.shape[0], A.shape[0]
    r = solve(P, q, A, l, u, eps_abs=1e-7, eps_rel=1e-7, polish=True, max_iter=200000)
    x, y = r.x, r.y
    lo = np.maximum(l, -OSQP_INFTY)
    up = np.minimum(u, OSQP_INFTY)
    eq = (up - lo) < RHO_TOL
    for _ in range(max_rounds):
        Ax = A @ x
        upp = (~eq) & ((y > 1e-10) | (Ax - up > 1e-9))
        low = (~eq) & ((y < -1e-10) | (lo - Ax > 1e-9)) & ~upp
        act = eq | upp | low
        Ar = A[act]
        na = Ar.shape[0]
        Kt = np.block([[P, Ar.T], [Ar, np.zeros((na, na))]])
        Kr = Kt + np.diag(np.concatenate([1e-9 * np.ones(n), -1e-9 * np.ones(na)]))
        rhs = np.concatenate([-q, np.where(low, lo, up)[act]])
        f = sla.lu_factor(Kr)
        s = sla.lu_solve(f, rhs)
        for _ in range(6):
            s = s + sla.lu_solve(f, rhs - Kt @ s)
        x = s[:n]
        y = np.zeros(m)
        y[act] = s[n:]
        # drop wrong-signed multipliers on inequality rows
        y[upp & (y < 0)] = 0.0
        y[low & (y > 0)] = 0.0
        pinf, stat, comp = kkt_certificate(P, q, A, lo, up, x, y)
        if pinf < 1e-8 and stat < 1e-6 and comp < 1e-8:
            break
    res = Result()
    res.x, res.y = x, y
    res.obj_val = 0.5 * x @ (P @ x) + q @ x
    res.cert = kkt_certificate(P, q, A, lo, up, x, y)
    res.admm = r
    return res

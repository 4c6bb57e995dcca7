"""TEST INFRASTRUCTURE ONLY (oracle/): ctypes wrapper of oracle/_ref/liblpv_oracle.so.

The library is the CPU restatement of PlannerLPV.solve (see the header of lpv_oracle.c for the
reference file:line of every function).  Importers allowed: tests/, __graft_entry__.smoke(),
bench.py's cpu_baseline / --impl reference legs.  The product never imports this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from colaborativempc_b200._abi import DmpcParams, make_params  # struct layout shared with include/dmpc_b200.h

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_ref", "liblpv_oracle.so")


class OracleSettings(C.Structure):
    _fields_ = [("rho", C.c_double), ("sigma", C.c_double), ("alpha", C.c_double), ("eps_abs", C.c_double),
                ("eps_rel", C.c_double), ("max_iter", C.c_int32), ("check_every", C.c_int32),
                ("adapt_interval", C.c_int32), ("scaling", C.c_int32), ("polish", C.c_int32),
                ("polish_refine", C.c_int32), ("delta", C.c_double), ("exact", C.c_int32), ("reserved", C.c_int32)]


class OracleIO(C.Structure):
    _fields_ = [("B", C.c_int32), ("xprev_rows", C.c_int32)] + \
               [(k, C.c_void_p) for k in ("x0", "xprev", "uprev", "uold", "nbr_xy", "pose", "nbr_count",
                                          "z", "planes", "y", "obj", "status", "iters", "polish",
                                          "rho_updates", "cert")]


def build(force=False):
    if force or not os.path.exists(LIB) or any(
            os.path.getmtime(os.path.join(HERE, f)) > os.path.getmtime(LIB)
            for f in ("lpv_oracle.c", "osqp_restated.inc", "exact_ipm.inc", "api.inc")):
        subprocess.check_call(["make", "-C", HERE, "-B", "-s"])
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(LIB)
        _lib.oracle_default_settings.argtypes = [C.POINTER(OracleSettings)]
        _lib.oracle_solve_batch.argtypes = [C.POINTER(DmpcParams), C.POINTER(OracleSettings), C.POINTER(OracleIO), C.c_int]
        _lib.oracle_solve_batch.restype = C.c_int
        _lib.oracle_assemble_dense.restype = C.c_int
        _lib.oracle_osqp_dense.restype = C.c_int
        assert _lib.oracle_sizeof_params() == C.sizeof(DmpcParams)
    return _lib


def settings(exact=False, **kw):
    """OSQP defaults as the reference passes them (LPV_Planner.py:233); exact=True adds the
    active-set corrections + KKT certificate (the parity anchor)."""
    s = OracleSettings()
    lib().oracle_default_settings(C.byref(s))
    if exact:
        s.exact = 1
        s.eps_abs = s.eps_rel = 1e-5
        s.max_iter = 4000
    for k, v in kw.items():
        setattr(s, k, v)
    return s


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f64(a, shape):
    a = np.ascontiguousarray(np.asarray(a, dtype=np.float64))
    assert a.shape == tuple(shape), (a.shape, shape)
    return a


def solve_batch(params, x0, xprev, uprev, uold, nbr_xy, pose, nbr_count=None, st=None, nthreads=1):
    """Returns dict(z, planes, y, obj, status, iters, polish, rho_updates, cert)."""
    N, nn = params.N, params.nn
    x0 = np.ascontiguousarray(np.asarray(x0, dtype=np.float64))
    B = x0.shape[0]
    xprev = np.ascontiguousarray(np.asarray(xprev, dtype=np.float64))
    rows = xprev.shape[1]
    uprev = _f64(uprev, (B, N, 2)); uold = _f64(uold, (B, 2))
    nbr_xy = _f64(nbr_xy, (B, N + 1, nn, 2)); pose = _f64(pose, (B, N + 1, 2))
    nc = None if nbr_count is None else np.ascontiguousarray(np.asarray(nbr_count, dtype=np.int32))
    n, m = 16 * N + 12, (8 + nn) * N + 14 * N + 12
    out = dict(z=np.zeros((B, n)), planes=np.zeros((B, N, 3, nn)), y=np.zeros((B, m)), obj=np.zeros(B),
               status=np.zeros(B, np.int32), iters=np.zeros(B, np.int32), polish=np.zeros(B, np.int32),
               rho_updates=np.zeros(B, np.int32), cert=np.zeros((B, 3)))
    io = OracleIO()
    io.B, io.xprev_rows = B, rows
    for k, a in (("x0", x0), ("xprev", xprev), ("uprev", uprev), ("uold", uold), ("nbr_xy", nbr_xy),
                 ("pose", pose), ("nbr_count", nc)):
        setattr(io, k, _ptr(a))
    for k in out:
        setattr(io, k, _ptr(out[k]))
    st = st or settings()
    rc = lib().oracle_solve_batch(C.byref(params), C.byref(st), C.byref(io), int(nthreads))
    if rc:
        raise RuntimeError("oracle_solve_batch failed: %d" % rc)
    out["m_in"] = (8 + nn) * N
    return out


def assemble_dense(params, x0, xprev, uprev, uold, nbr_xy, pose):
    """Dense (P, q, A, l, u) + planes/weights/Ak/Bk/ey for ONE agent, as the reference builds them."""
    N, nn = params.N, params.nn
    n, m = 16 * N + 12, (8 + nn) * N + 14 * N + 12
    xprev = np.ascontiguousarray(np.asarray(xprev, dtype=np.float64))
    o = dict(planes=np.zeros((N, 3, nn)), weights=np.zeros((N, nn)), Ak=np.zeros((N, 9, 9)), Bk=np.zeros((N, 9, 2)),
             ey=np.zeros(N), P=np.zeros((n, n)), q=np.zeros(n), A=np.zeros((m, n)), l=np.zeros(m), u=np.zeros(m))
    args = [_f64(x0, (9,)), xprev, _f64(uprev, (N, 2)), _f64(uold, (2,)), _f64(nbr_xy, (N + 1, nn, 2)), _f64(pose, (N + 1, 2))]
    rc = lib().oracle_assemble_dense(C.byref(params), C.c_int(xprev.shape[0]), *[_ptr(a) for a in args],
                                     *[_ptr(o[k]) for k in ("planes", "weights", "Ak", "Bk", "ey", "P", "q", "A", "l", "u")])
    if rc:
        raise RuntimeError("oracle_assemble_dense failed: %d" % rc)
    return o


def osqp_dense(P, q, A, l, u, st=None):
    n, m = P.shape[0], A.shape[0]
    x, y, info = np.zeros(n), np.zeros(m), np.zeros(8)
    st = st or settings()
    arrs = [np.ascontiguousarray(a, dtype=np.float64) for a in (P, q, A, l, u)]
    rc = lib().oracle_osqp_dense(C.c_int(n), C.c_int(m), *[_ptr(a) for a in arrs], C.byref(st), _ptr(x), _ptr(y), _ptr(info))
    if rc:
        raise RuntimeError("oracle_osqp_dense failed")
    return dict(x=x, y=y, status=int(info[0]), iters=int(info[1]), polish=int(info[2]), obj=info[3], cert=info[4:7],
                rho_updates=int(info[7]))


def sim_batch(params, xpred, u):
    """LPV_val.py:77-94 (LPV_Model.sim) for a batch: xpred (B, N+1, 9), u (B, N, 2) -> (rolled-out and shifted xPred, status)."""
    N = params.N
    xpred = np.ascontiguousarray(np.asarray(xpred, dtype=np.float64))
    B = xpred.shape[0]
    u = _f64(u, (B, N, 2))
    out = np.zeros((B, N + 1, 9))
    status = np.zeros(B, dtype=np.int32)
    f = lib().oracle_sim_batch
    f.restype = C.c_int
    rc = f(C.byref(params), C.c_int(B), _ptr(xpred), _ptr(u), _ptr(out), status.ctypes.data_as(C.c_void_p))
    if rc:
        raise RuntimeError("oracle_sim_batch failed: %d" % rc)
    return out, status


def ocd_lambda_update(lambdas, xy, alpha, dth):
    """NL_EU_N_main.py:122-138 (Euclidean form): xy (S, A, N+1, 2), lambdas (S, A, A, N) -> updated copy."""
    xy = np.ascontiguousarray(np.asarray(xy, dtype=np.float64))
    S, A, N1, _ = xy.shape
    lam = np.array(lambdas, dtype=np.float64, order="C", copy=True)
    assert lam.shape == (S, A, A, N1 - 1)
    f = lib().oracle_ocd_lambda_update
    f.restype = C.c_int
    f.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_void_p, C.c_void_p]
    rc = f(S, A, N1 - 1, float(alpha), float(dth), _ptr(xy), _ptr(lam))
    if rc:
        raise RuntimeError("oracle_ocd_lambda_update failed")
    return lam


def ocd_lambda_update_hp(lambdas, xy, planes, alpha, dth):
    """NL_HP_N_main.py:128-137 (hyperplane form): xy (S, A, N+1, 2), planes (S, A, N, 3, A), lambdas (S, A, A, N)."""
    xy = np.ascontiguousarray(np.asarray(xy, dtype=np.float64))
    planes = np.ascontiguousarray(np.asarray(planes, dtype=np.float64))
    S, A, N1, _ = xy.shape
    lam = np.array(lambdas, dtype=np.float64, order="C", copy=True)
    assert lam.shape == (S, A, A, N1 - 1) and planes.shape == (S, A, N1 - 1, 3, A)
    f = lib().oracle_ocd_lambda_update_hp
    f.restype = C.c_int
    f.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p]
    rc = f(S, A, N1 - 1, float(alpha), float(dth), _ptr(xy), _ptr(planes), _ptr(lam))
    if rc:
        raise RuntimeError("oracle_ocd_lambda_update_hp failed")
    return lam


def unpack(z, N):
    """LPV_Planner.py:162-180, including the duPred indexing quirk (:175)."""
    z = np.asarray(z)
    raw = z[: 12 * (N + 1)].reshape(N + 1, 12)
    xPred = raw[:, :9].copy()
    sPred = raw[1:, 9:12].copy()
    base = 12 * (N + 1)
    uPred = z[base + np.arange(2 * N)].reshape(N, 2)
    duPred = z[base + 2 * np.arange(2 * N)].reshape(N, 2)
    return xPred, uPred, sPred, duPred, raw.copy()

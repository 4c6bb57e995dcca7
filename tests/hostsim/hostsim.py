"""TEST INFRASTRUCTURE ONLY: builds and loads tests/hostsim/_build/libqp3_host.so - the stage-wise engine's device
source compiled for the host (see qp3_host.cpp).  Imported by tests only; the product package never loads it."""
import ctypes as C
import os
import subprocess

import numpy as np

from colaborativempc_b200._abi import DmpcBatchIO, DmpcOptions, DmpcParams

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "..", "..", "colaborativempc_b200", "csrc")
LIB = os.path.join(HERE, "_build", "libqp3_host.so")
SRCS = [os.path.join(HERE, "qp3_host.cpp")] + [os.path.join(CSRC, f) for f in
                                               ("qp3_engine.cuh", "qp3_solve.cuh", "qp3_kernel.cuh", "qp_types.h", "qp_params.h")]


class ShiftIO(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ("x0", "xprev", "uprev", "uold", "xy_send", "hold", "flags", "stop")] + \
               [("end_s", C.c_double), ("laps", C.c_int32), ("stop_div", C.c_int32)]


def build(force=False):
    if force or not os.path.exists(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in SRCS):
        os.makedirs(os.path.dirname(LIB), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-fopenmp", "-std=c++17", "-shared", "-fPIC", "-D_GNU_SOURCE",
                               "-Wno-unknown-pragmas", "-o", LIB, SRCS[0]])
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.qp3_host_solve_batch.restype = C.c_int
        _lib.qp3_host_workspace_bytes.restype = C.c_int
    return _lib


def default_options():
    o = DmpcOptions()
    o.polish_delta, o.cert_tol, o.ipm_mu0 = 1e-9, 1e-9, 1e-6
    o.polish_refine, o.pdas_first, o.pdas_after, o.ipm_max_iter = 4, 4, 2, 60
    o.engine, o.order_hint, o.host_direct, o.warm_start = 3, 0, 0, 0
    return o


def solve_batch(params, x0, xprev, uprev, uold, nbr_xy, pose, nbr_count=None, options=None, shift=False, warm=None,
                warm_shift=0, end_s=0.0, stop=None, stop_div=1, cost=None):
    """warm: uint8 [B, (8+nn)N] working sets, read as first guesses and overwritten with the certified ones."""
    N, nn = params.N, params.nn
    f = lambda a: np.ascontiguousarray(a, dtype=np.float64)
    x0, xprev, uprev, uold, nbr_xy, pose = f(x0), f(xprev), f(uprev), f(uold), f(nbr_xy), f(pose)
    B, rows = x0.shape[0], xprev.shape[1]
    nc = None if nbr_count is None else np.ascontiguousarray(nbr_count, dtype=np.int32)
    n, mi, me = 16 * N + 12, (8 + nn) * N, 14 * N + 12
    out = dict(z=np.zeros((B, n)), planes=np.zeros((B, N, 3, nn)), y_ineq=np.zeros((B, mi)), y_eq=np.zeros((B, me)),
               obj=np.zeros(B), status=np.zeros(B, np.int32), iters=np.zeros(B, np.int32), pivots=np.zeros(B, np.int32))
    io = DmpcBatchIO()
    io.B, io.xprev_rows = B, rows
    p = lambda a: None if a is None else a.ctypes.data
    for k, a in (("x0", x0), ("xprev", xprev), ("uprev", uprev), ("uold", uold), ("nbr_xy", nbr_xy), ("pose", pose), ("nbr_count", nc)):
        setattr(io, k, p(a))
    for k in ("z", "planes", "y_ineq", "y_eq", "obj", "status", "iters", "pivots"):
        setattr(io, k, p(out[k]))
    counters = np.zeros((B, 4), np.int32)
    sh = None
    if shift:
        nx = dict(x0=x0.copy(), xprev=np.zeros((B, N, 9)), uprev=uprev.copy(), uold=uold.copy(),
                  xy_send=np.zeros((B, N + 1, 2)), hold=np.zeros(B, np.int32))
        if rows == N:
            nx["xprev"][:] = xprev
        sh = ShiftIO()
        if end_s > 0:
            nx["flags"] = np.full(B, -1.0)
        if stop is not None:
            nx["stop"] = np.ascontiguousarray(stop, dtype=np.int32)
        for k, a in nx.items():
            setattr(sh, k, p(a))
        sh.end_s, sh.laps, sh.stop_div = float(end_s), 1, int(stop_div)
        out["next"] = nx
    o = options or default_options()
    if warm is not None:
        assert warm.dtype == np.uint8 and warm.flags.c_contiguous and warm.shape == (B, mi)
    rc = lib().qp3_host_solve_batch(C.byref(params), C.byref(o), C.byref(io), C.byref(sh) if sh is not None else None,
                                    counters.ctypes.data_as(C.c_void_p), None if warm is None else warm.ctypes.data_as(C.c_void_p),
                                    C.c_int(int(warm_shift)), None if cost is None else cost.ctypes.data_as(C.c_void_p))
    assert rc == 0
    out["counters"] = counters
    return out

"""ctypes mirror of include/dmpc_b200.h (structs + library loader).

No numerics live here; this is the thin C-ABI binding layer named by BASELINE.json's north_star
("Python, calling CUDA through a thin C-ABI/ctypes layer").
"""
import ctypes as C
import os

import numpy as np

DMPC_MAX_SEG = 32
NS, NSLACK, NEXP, NU = 9, 3, 12, 2

# per-QP status values (OSQP status_val compatible; LPV_Planner.py:246-248)
SOLVED, SOLVED_INACCURATE, MAX_ITER_REACHED = 1, 2, -2
PRIMAL_INFEASIBLE, DUAL_INFEASIBLE, UNSOLVED = -3, -4, -10
MAP_LOOKUP_FAILED, DEGENERATE_PLANE = -20, -21

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)


class DmpcParams(C.Structure):
    _fields_ = [
        ("N", C.c_int32), ("nn", C.c_int32), ("nseg", C.c_int32), ("reserved", C.c_int32),
        ("dt", C.c_double), ("wq", C.c_double),
        ("Q", C.c_double * 81), ("Qs", C.c_double * 9), ("R", C.c_double * 4), ("dR", C.c_double * 4),
        ("lf", C.c_double), ("lr", C.c_double), ("m", C.c_double), ("I", C.c_double),
        ("Cf", C.c_double), ("Cr", C.c_double), ("mu", C.c_double),
        ("vx_ref", C.c_double), ("min_dist", C.c_double), ("max_vel", C.c_double), ("min_vel", C.c_double),
        ("max_rs", C.c_double), ("max_ls", C.c_double), ("max_ac", C.c_double), ("max_dc", C.c_double),
        ("seg_s0", C.c_double * DMPC_MAX_SEG), ("seg_len", C.c_double * DMPC_MAX_SEG),
        ("seg_curv", C.c_double * DMPC_MAX_SEG), ("seg_hw", C.c_double * DMPC_MAX_SEG),
    ]


class DmpcOptions(C.Structure):
    _fields_ = [
        ("polish_delta", C.c_double), ("cert_tol", C.c_double), ("ipm_mu0", C.c_double),
        ("polish_refine", C.c_int32), ("pdas_first", C.c_int32), ("pdas_after", C.c_int32),
        ("ipm_max_iter", C.c_int32), ("engine", C.c_int32), ("order_hint", C.c_int32), ("host_direct", C.c_int32),
        ("warm_start", C.c_int32),
    ]


class DmpcBatchIO(C.Structure):
    _fields_ = [
        ("B", C.c_int32), ("xprev_rows", C.c_int32),
        ("x0", C.c_void_p), ("xprev", C.c_void_p), ("uprev", C.c_void_p), ("uold", C.c_void_p),
        ("nbr_xy", C.c_void_p), ("pose", C.c_void_p), ("nbr_count", C.c_void_p),
        ("z", C.c_void_p), ("planes", C.c_void_p), ("y_ineq", C.c_void_p), ("y_eq", C.c_void_p),
        ("obj", C.c_void_p), ("status", C.c_void_p), ("iters", C.c_void_p), ("pivots", C.c_void_p),
    ]


class DmpcShiftIO(C.Structure):
    """dmpc_shift_io: the next DMPC iteration's buffers, written by the solve's own epilogue."""
    _fields_ = [(k, C.c_void_p) for k in ("x0", "xprev", "uprev", "uold", "xy_send", "hold", "flags", "stop")] + \
               [("end_s", C.c_double), ("laps", C.c_int32), ("stop_div", C.c_int32)]


ENGINE_AUTO, ENGINE_FAST, ENGINE_STAGE = 0, 2, 3

# Defaults of PlannerLPV.__init__ when model_param / sys_lim are None (LPV_Planner.py:33-57)
DEFAULT_MODEL = dict(lf=0.12, lr=0.14, m=2.250, I=0.06, Cf=60.0, Cr=60.0, mu=0.05)
DEFAULT_LIMITS = dict(vx_ref=6.0, min_dist=0.45, max_vel=5.5, min_vel=0.2, max_rs=0.45, max_ls=0.45,
                      max_ac=4.0, max_dc=3.0, sm=0.9)


def map_table(map_obj):
    """(s0, len, curv, halfwidth) rows of the reference-style map object, exactly what
    curvature()/get_ey() read (utilities/misc.py:84,109,125): PointAndTangent[:, :, lane], halfWidth."""
    pt = np.asarray(map_obj.PointAndTangent, dtype=np.float64)
    if pt.ndim == 3:
        pt = pt[:, :, int(getattr(map_obj, "lane", 0))]
    hw = np.asarray(map_obj.halfWidth, dtype=np.float64).ravel()
    nseg = pt.shape[0]
    if nseg > DMPC_MAX_SEG:
        raise ValueError("map has %d segments, DMPC_MAX_SEG is %d" % (nseg, DMPC_MAX_SEG))
    if hw.shape[0] < nseg:  # get_ey indexes halfWidth[i] with the PointAndTangent row index
        hw = np.concatenate([hw, np.full(nseg - hw.shape[0], np.nan)])
    return pt[:, 3].copy(), pt[:, 4].copy(), pt[:, 5].copy(), hw[:nseg].copy()


def make_params(Q, Qs, R, dR, N, dt, map_obj, wq=0.0, model_param=None, sys_lim=None, nn=0):
    mp = dict(DEFAULT_MODEL if model_param is None else model_param)
    sl = dict(DEFAULT_LIMITS if sys_lim is None else sys_lim)
    p = DmpcParams()
    p.N, p.nn, p.reserved = int(N), int(nn), 0
    p.dt, p.wq = float(dt), float(wq)
    for name, arr, k in (("Q", Q, 81), ("Qs", Qs, 9), ("R", R, 4), ("dR", dR, 4)):
        flat = np.ascontiguousarray(np.asarray(arr, dtype=np.float64)).ravel()
        if flat.shape[0] != k:
            raise ValueError("%s must have %d entries" % (name, k))
        getattr(p, name)[:] = flat.tolist()
    for k in ("lf", "lr", "m", "I", "Cf", "Cr", "mu"):
        setattr(p, k, float(mp[k]))
    for k in ("vx_ref", "min_dist", "max_vel", "min_vel", "max_rs", "max_ls", "max_ac", "max_dc"):
        setattr(p, k, float(sl[k]))
    s0, ln, cv, hw = map_table(map_obj)
    p.nseg = len(s0)
    for name, arr in (("seg_s0", s0), ("seg_len", ln), ("seg_curv", cv), ("seg_hw", hw)):
        a = getattr(p, name)
        for i, v in enumerate(arr):
            a[i] = float(v)
    return p


def default_lib_path():
    if os.environ.get("DMPC_B200_LIB"):  # development: A/B of two builds of the same C ABI in one GPU session
        return os.environ["DMPC_B200_LIB"]
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "libdmpc_b200.so")


EXPORTS = ["dmpc_default_options", "dmpc_version", "dmpc_create", "dmpc_destroy", "dmpc_last_error",
           "dmpc_set_options", "dmpc_solve_batch", "dmpc_solve_batch_static", "dmpc_solve_shift_batch", "dmpc_solve_batch_host", "dmpc_shift_pack",
           "dmpc_shift_pack_hold",
           "dmpc_gather_neighbours", "dmpc_gather_neighbours_stop", "dmpc_launch_count", "dmpc_smem_bytes", "dmpc_engine", "dmpc_sim_batch",
           "dmpc_ocd_lambda_update", "dmpc_ocd_lambda_update_hp"]

_lib = None


def load_library(path=None):
    """Loads the CUDA library.  There is NO fallback: a missing library is an ImportError."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = path or default_lib_path()
    if not os.path.exists(path):
        raise ImportError("CUDA library %s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback)" % path)
    lib = C.CDLL(path)
    lib.dmpc_default_options.argtypes = [C.POINTER(DmpcOptions)]
    lib.dmpc_default_options.restype = None
    lib.dmpc_version.restype = C.c_int
    lib.dmpc_create.argtypes = [C.POINTER(DmpcParams), C.POINTER(DmpcOptions), C.c_int32, C.c_int32,
                                C.POINTER(C.c_void_p)]
    lib.dmpc_create.restype = C.c_int
    lib.dmpc_destroy.argtypes = [C.c_void_p]
    lib.dmpc_destroy.restype = None
    lib.dmpc_last_error.argtypes = [C.c_void_p]
    lib.dmpc_last_error.restype = C.c_char_p
    lib.dmpc_set_options.argtypes = [C.c_void_p, C.POINTER(DmpcOptions)]
    lib.dmpc_set_options.restype = C.c_int
    lib.dmpc_solve_batch.argtypes = [C.c_void_p, C.POINTER(DmpcBatchIO), C.c_void_p]
    lib.dmpc_solve_batch.restype = C.c_int
    lib.dmpc_solve_shift_batch.argtypes = [C.c_void_p, C.POINTER(DmpcBatchIO), C.POINTER(DmpcShiftIO), C.c_void_p]
    lib.dmpc_solve_shift_batch.restype = C.c_int
    lib.dmpc_shift_pack_hold.argtypes = [C.c_void_p, C.c_int32] + [C.c_void_p] * 10
    lib.dmpc_shift_pack_hold.restype = C.c_int
    lib.dmpc_solve_batch_static.argtypes = [C.POINTER(DmpcParams), C.POINTER(DmpcOptions), C.POINTER(DmpcBatchIO), C.c_int32,
                                            C.c_void_p, C.c_char_p, C.c_int32]
    lib.dmpc_solve_batch_static.restype = C.c_int
    lib.dmpc_solve_batch_host.argtypes = [C.c_void_p, C.POINTER(DmpcBatchIO)]
    lib.dmpc_solve_batch_host.restype = C.c_int
    lib.dmpc_shift_pack.argtypes = [C.c_void_p, C.c_int32] + [C.c_void_p] * 7
    lib.dmpc_shift_pack.restype = C.c_int
    lib.dmpc_gather_neighbours.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32] + [C.c_void_p] * 5
    lib.dmpc_gather_neighbours.restype = C.c_int
    lib.dmpc_gather_neighbours_stop.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32] + [C.c_void_p] * 6
    lib.dmpc_gather_neighbours_stop.restype = C.c_int
    lib.dmpc_engine.argtypes = [C.c_void_p]
    lib.dmpc_engine.restype = C.c_int
    lib.dmpc_sim_batch.argtypes = [C.c_void_p, C.c_int32] + [C.c_void_p] * 5
    lib.dmpc_sim_batch.restype = C.c_int
    lib.dmpc_ocd_lambda_update.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_double, C.c_double, C.c_void_p, C.c_void_p,
                                           C.c_void_p]
    lib.dmpc_ocd_lambda_update.restype = C.c_int
    lib.dmpc_ocd_lambda_update_hp.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_double, C.c_double, C.c_void_p, C.c_void_p,
                                              C.c_void_p, C.c_void_p]
    lib.dmpc_ocd_lambda_update_hp.restype = C.c_int
    lib.dmpc_smem_bytes.argtypes = [C.c_void_p]
    lib.dmpc_smem_bytes.restype = C.c_int64
    lib.dmpc_launch_count.argtypes = [C.c_void_p]
    lib.dmpc_launch_count.restype = C.c_int64
    if path == default_lib_path():
        _lib = lib
    return lib

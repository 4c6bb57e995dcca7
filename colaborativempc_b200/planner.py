"""Host-side mirror of the reference planner interface over the CUDA C ABI.

* ``PlannerLPV``      drop-in for ``plan_lib.distributedPlanner.PlannerLPV`` (LPV_Planner.py:15-182): same
                      constructor, same ``solve(x0, Last_xPredicted, uPred, x_agents, agents_id, pose)``, same result
                      attributes (xPred, uPred, duPred, sPred, raw_States, planes), same ``feasible`` rule (:246-248).
* ``BatchedPlannerLPV`` the batched entry point: B = scenarios x agents QPs per call on device tensors.

PyTorch is used for device memory and streams only.  All numerics run in csrc/ (hand-written sm_100a kernels)
behind include/dmpc_b200.h; there is no CPU fallback - importing without the built library raises ImportError.
"""
import ctypes as C
import warnings

import numpy as np

from . import _abi
from ._abi import DmpcBatchIO, DmpcOptions, DmpcShiftIO, make_params

FEASIBLE_STATUS = (_abi.SOLVED, _abi.SOLVED_INACCURATE, _abi.MAX_ITER_REACHED)  # LPV_Planner.py:246-248


class DmpcError(RuntimeError):
    pass


def _torch():
    import torch
    return torch


class BatchedPlannerLPV(object):
    """B independent agent QPs per call.  Tensors are CUDA float64 / int32, C-contiguous, caller-owned."""

    def __init__(self, Q, Qs, R, dR, N, dt, map, wq=0.0, model_param=None, sys_lim=None, nn=0, max_batch=1,
                 device=None, options=None):
        torch = _torch()
        self.lib = _abi.load_library()
        if not torch.cuda.is_available():
            raise DmpcError("no CUDA device: the LPV-MPC path has no CPU fallback")
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None else device)
        self.N, self.nn, self.max_batch = int(N), int(nn), int(max_batch)
        self.params = make_params(Q, Qs, R, dR, N, dt, map, wq, model_param, sys_lim, nn=nn)
        self.options = DmpcOptions()
        self.lib.dmpc_default_options(C.byref(self.options))
        for k, v in (options or {}).items():
            setattr(self.options, k, v)
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            rc = self.lib.dmpc_create(C.byref(self.params), C.byref(self.options), self.max_batch, self.device.index,
                                      C.byref(self._h))
        if rc != 0:
            msg = self.lib.dmpc_last_error(self._h).decode() if self._h else "allocation failed"
            self.lib.dmpc_destroy(self._h)
            self._h = None
            raise DmpcError("dmpc_create failed (%d): %s" % (rc, msg))
        self.engine = self.lib.dmpc_engine(self._h)      # _abi.ENGINE_STAGE (qp3_*.cuh) or _abi.ENGINE_FAST (qp2_*.cuh)
        self.fast = self.engine == _abi.ENGINE_FAST
        self.n = 16 * self.N + 12
        ns = int(self.params.nseg)
        self.track_length = float(self.params.seg_s0[ns - 1] + self.params.seg_len[ns - 1])  # Map.TrackLength
        self.m_in = (8 + self.nn) * self.N
        self.m_eq = 14 * self.N + 12

    def close(self):
        if getattr(self, "_h", None):
            self.lib.dmpc_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise DmpcError("dmpc call failed (%d): %s" % (rc, self.lib.dmpc_last_error(self._h).decode()))

    @property
    def launches(self):
        return int(self.lib.dmpc_launch_count(self._h))

    def stage_smem_bytes(self):
        """Dynamic shared memory per QP (one CTA) of the handle's engine."""
        return int(self.lib.dmpc_smem_bytes(self._h))

    def set_options(self, **kw):
        for k, v in kw.items():
            setattr(self.options, k, v)
        self._check(self.lib.dmpc_set_options(self._h, C.byref(self.options)))

    # ---- device path -------------------------------------------------------------------------------------
    def alloc_outputs(self, B, duals=True):
        torch = _torch()
        kw = dict(device=self.device)
        out = dict(z=torch.empty((B, self.n), dtype=torch.float64, **kw),
                   planes=torch.empty((B, self.N, 3, self.nn), dtype=torch.float64, **kw),
                   obj=torch.empty((B,), dtype=torch.float64, **kw),
                   status=torch.empty((B,), dtype=torch.int32, **kw),
                   iters=torch.empty((B,), dtype=torch.int32, **kw),
                   pivots=torch.empty((B,), dtype=torch.int32, **kw))
        if duals:
            out["y_ineq"] = torch.empty((B, self.m_in), dtype=torch.float64, **kw)
            out["y_eq"] = torch.empty((B, self.m_eq), dtype=torch.float64, **kw)
        return out

    def _io(self, B, rows, ins, outs, ptr):
        io = DmpcBatchIO()
        io.B, io.xprev_rows = B, rows
        for k in ("x0", "xprev", "uprev", "uold", "nbr_xy", "pose", "nbr_count"):
            setattr(io, k, ptr(ins.get(k)))
        for k in ("z", "planes", "y_ineq", "y_eq", "obj", "status", "iters", "pivots"):
            setattr(io, k, ptr(outs.get(k)))
        return io

    def solve_batch(self, x0, xprev, uprev, uold, nbr_xy, pose, nbr_count=None, out=None, stream=None, duals=True,
                    shift=None):
        """Asynchronous on `stream` (default: torch's current stream).  Returns the dict of output tensors.
        shift: dict(x0, xprev [B,N,9], uprev, uold, xy_send [B,N+1,2], hold int32 [B]) of CUDA tensors - the next DMPC
        iteration's buffers, written by the solve's own epilogue (dmpc_solve_shift_batch; LPV_HP_N_main.py:102-117);
        optional flags [B] float64 (with end_s, laps) and stop int32 [B / stop_div] (with stop_div): dmpc_shift_io."""
        torch = _torch()
        B, rows = x0.shape[0], xprev.shape[1]
        ins = dict(x0=x0, xprev=xprev, uprev=uprev, uold=uold, nbr_xy=nbr_xy, pose=pose, nbr_count=nbr_count)
        shapes = dict(x0=(B, 9), xprev=(B, rows, 9), uprev=(B, self.N, 2), uold=(B, 2),
                      nbr_xy=(B, self.N + 1, self.nn, 2), pose=(B, self.N + 1, 2), nbr_count=(B,))
        for k, t in ins.items():
            if t is None:
                continue
            want = torch.int32 if k == "nbr_count" else torch.float64
            if not (t.is_cuda and t.dtype == want and t.is_contiguous() and tuple(t.shape) == shapes[k]):
                raise DmpcError("%s must be a contiguous CUDA %s tensor of shape %s" % (k, want, shapes[k]))
        out = out if out is not None else self.alloc_outputs(B, duals)
        s = torch.cuda.current_stream(self.device) if stream is None else stream
        io = self._io(B, rows, ins, out, lambda t: None if t is None else t.data_ptr())
        if shift is None:
            self._check(self.lib.dmpc_solve_batch(self._h, C.byref(io), C.c_void_p(s.cuda_stream)))
            return out
        div = int(shift.get("stop_div", 1))
        shapes = dict(x0=(B, 9), xprev=(B, self.N, 9), uprev=(B, self.N, 2), uold=(B, 2), xy_send=(B, self.N + 1, 2), hold=(B,),
                      flags=(B,), stop=((B + div - 1) // div,))
        sh = DmpcShiftIO()
        for k, shp in shapes.items():
            t = shift.get(k)
            if t is None and k in ("flags", "stop"):
                continue
            want = torch.int32 if k in ("hold", "stop") else torch.float64
            if not (t.is_cuda and t.dtype == want and t.is_contiguous() and tuple(t.shape) == shp):
                raise DmpcError("shift[%s] must be a contiguous CUDA %s tensor of shape %s" % (k, want, shp))
            setattr(sh, k, t.data_ptr())
        sh.end_s, sh.laps, sh.stop_div = float(shift.get("end_s", 0.0)), int(shift.get("laps", 1)), div
        self._check(self.lib.dmpc_solve_shift_batch(self._h, C.byref(io), C.byref(sh), C.c_void_p(s.cuda_stream)))
        return out

    # ---- host path (numpy in / numpy out; H2D + solve + D2H inside the C call) -----------------------------
    def solve_batch_host(self, x0, xprev, uprev, uold, nbr_xy, pose, nbr_count=None, out=None, duals=True):
        f = lambda a: np.ascontiguousarray(a, dtype=np.float64)
        x0, xprev, uprev, uold, nbr_xy, pose = f(x0), f(xprev), f(uprev), f(uold), f(nbr_xy), f(pose)
        B, rows = x0.shape[0], xprev.shape[1]
        nc = None if nbr_count is None else np.ascontiguousarray(nbr_count, dtype=np.int32)
        if out is None:
            out = dict(z=np.empty((B, self.n)), planes=np.empty((B, self.N, 3, self.nn)), obj=np.empty(B),
                       status=np.empty(B, np.int32), iters=np.empty(B, np.int32), pivots=np.empty(B, np.int32))
            if duals:
                out["y_ineq"] = np.empty((B, self.m_in))
                out["y_eq"] = np.empty((B, self.m_eq))
        ins = dict(x0=x0, xprev=xprev, uprev=uprev, uold=uold, nbr_xy=nbr_xy, pose=pose, nbr_count=nc)
        io = self._io(B, rows, ins, out, lambda a: None if a is None else a.ctypes.data)
        self._check(self.lib.dmpc_solve_batch_host(self._h, C.byref(io)))
        return out

    # ---- DMPC bookkeeping on device (LPV_HP_N_main.py:106,113-117) -----------------------------------------
    def shift_pack(self, z, x0, xprev, uprev, uold, xy_send, stream=None):
        torch = _torch()
        s = torch.cuda.current_stream(self.device) if stream is None else stream
        p = lambda t: None if t is None else t.data_ptr()
        self._check(self.lib.dmpc_shift_pack(self._h, z.shape[0], p(z), p(x0), p(xprev), p(uprev), p(uold), p(xy_send),
                                             C.c_void_p(s.cuda_stream)))

    def shift_pack_hold(self, z, status, x0, xprev, uprev, uold, xy_send, hold, stream=None):
        """shift_pack with the reference's stop rule (LPV_HP_N_main.py:102-104): QPs whose status is not feasible keep their
        buffers (last valid plan) and count in `hold`."""
        torch = _torch()
        s = torch.cuda.current_stream(self.device) if stream is None else stream
        p = lambda t: None if t is None else t.data_ptr()
        self._check(self.lib.dmpc_shift_pack_hold(self._h, z.shape[0], p(z), p(status), p(x0), p(xprev), p(uprev), p(uold),
                                                  p(xy_send), p(hold), C.c_void_p(s.cuda_stream)))

    def gather_neighbours(self, xy_all, nbr_idx, nbr_xy, pose, S, A_total, A_local, agent_offset, stream=None, stop=None):
        """stop (int32 [S]): xy_all is the all-gathered payload WITH the flag tails (dmpc_gather_neighbours_stop)."""
        torch = _torch()
        s = torch.cuda.current_stream(self.device) if stream is None else stream
        p = lambda t: None if t is None else t.data_ptr()
        if stop is None:
            self._check(self.lib.dmpc_gather_neighbours(self._h, S, A_total, A_local, agent_offset, p(xy_all), p(nbr_idx),
                                                        p(nbr_xy), p(pose), C.c_void_p(s.cuda_stream)))
        else:
            self._check(self.lib.dmpc_gather_neighbours_stop(self._h, S, A_total, A_local, agent_offset, p(xy_all), p(nbr_idx),
                                                             p(nbr_xy), p(pose), p(stop), C.c_void_p(s.cuda_stream)))


    def sim_batch(self, xpred, u, out=None, status=None, stream=None):
        """Batched LPV_Model.sim (LPV_val.py:77-94) on device tensors: xpred [B,N+1,9], u [B,N,2] -> [B,N+1,9], status [B]."""
        torch = _torch()
        B = xpred.shape[0]
        assert xpred.is_cuda and xpred.dtype == torch.float64 and xpred.is_contiguous() and tuple(xpred.shape) == (B, self.N + 1, 9)
        assert u.is_cuda and u.dtype == torch.float64 and u.is_contiguous() and tuple(u.shape) == (B, self.N, 2)
        if out is None:
            out = torch.empty_like(xpred)
        if status is None:
            status = torch.empty((B,), dtype=torch.int32, device=xpred.device)
        s = torch.cuda.current_stream(self.device) if stream is None else stream
        self._check(self.lib.dmpc_sim_batch(self._h, B, xpred.data_ptr(), u.data_ptr(), out.data_ptr(), status.data_ptr(),
                                            C.c_void_p(s.cuda_stream)))
        return out, status


    def ocd_lambda_update(self, lambdas, xy, dth, alpha=0.25, stream=None):
        """In-place dual update of the OCD loop (NL_EU_N_main.py:122-138, Euclidean form) on device tensors:
        xy [S,A,N+1,2], lambdas [S,A,A,N]; alpha defaults to the reference's get_alpha()."""
        torch = _torch()
        S, A = xy.shape[0], xy.shape[1]
        assert xy.is_cuda and xy.dtype == torch.float64 and xy.is_contiguous() and tuple(xy.shape) == (S, A, self.N + 1, 2)
        assert lambdas.is_cuda and lambdas.dtype == torch.float64 and lambdas.is_contiguous()
        assert tuple(lambdas.shape) == (S, A, A, self.N)
        s = torch.cuda.current_stream(self.device) if stream is None else stream
        self._check(self.lib.dmpc_ocd_lambda_update(self._h, S, A, float(alpha), float(dth), xy.data_ptr(), lambdas.data_ptr(),
                                                    C.c_void_p(s.cuda_stream)))
        return lambdas

    def ocd_lambda_update_hp(self, lambdas, xy, planes, dth, alpha=0.25, stream=None):
        """In-place dual update of the OCD loop, hyperplane form (NL_HP_N_main.py:128-137 with eval_constraintHp) on
        device tensors: xy [S,A,N+1,2], planes [S,A,N,3,A] (agent i's planes_raw stacked over i), lambdas [S,A,A,N]."""
        torch = _torch()
        S, A = xy.shape[0], xy.shape[1]
        for t, shp in ((xy, (S, A, self.N + 1, 2)), (planes, (S, A, self.N, 3, A)), (lambdas, (S, A, A, self.N))):
            assert t.is_cuda and t.dtype == torch.float64 and t.is_contiguous() and tuple(t.shape) == shp
        s = torch.cuda.current_stream(self.device) if stream is None else stream
        self._check(self.lib.dmpc_ocd_lambda_update_hp(self._h, S, A, float(alpha), float(dth), xy.data_ptr(),
                                                       planes.data_ptr(), lambdas.data_ptr(), C.c_void_p(s.cuda_stream)))
        return lambdas


def unpack_solution(z, N):
    """Views of res.x as PlannerLPV.solve unpacks it (LPV_Planner.py:162-180), incl. the duPred indexing of :175."""
    z = np.asarray(z)
    raw = z[:12 * (N + 1)].reshape(N + 1, 12)
    base = 12 * (N + 1)
    xPred = raw[:, :9].copy()
    sPred = raw[1:, 9:12].copy()
    uPred = z[base:base + 2 * N].reshape(N, 2).copy()
    duPred = z[base + 2 * np.arange(2 * N)].reshape(N, 2).copy()  # every second entry over [u | du] - reference quirk
    return xPred, uPred, sPred, duPred, raw.copy()


class PlannerLPV(object):
    """Single-agent controller with the reference's call surface (internally a batch of one)."""

    def __init__(self, Q, Qs, R, dR, N, dt, map, id, wq=0, model_param=None, sys_lim=None):
        self.n_s, self.slack, self.n_u = 9, 3, 2
        self.n_exp = self.n_s + self.slack
        self.id, self.dt, self.map, self.N = id, dt, map, N
        self.first_it = True
        self.OldSteering = [0.0]
        self.OldAccelera = [0.0]
        mp = dict(_abi.DEFAULT_MODEL if model_param is None else model_param)
        sl = dict(_abi.DEFAULT_LIMITS if sys_lim is None else sys_lim)
        for k in ("lf", "lr", "m", "I", "Cf", "Cr", "mu"):
            setattr(self, k, mp[k])
        for k in ("vx_ref", "min_dist", "max_vel", "min_vel", "max_rs", "max_ls", "max_ac", "max_dc", "sm"):
            setattr(self, k, sl[k])
        self._mp, self._sl = mp, sl
        Q, Qs, R, dR = (np.asarray(a, dtype=np.float64) for a in (Q, Qs, R, dR))
        # wrong-shape gains fall back to identity with a warning (LPV_Planner.py:84-105)
        if Q.shape[0] != self.n_s:
            warnings.warn("Q has not the correct shape!, defaulting to identity of " + str(self.n_s))
            Q = np.eye(self.n_s)
        if Qs.shape[0] != self.slack:
            warnings.warn("Qs has not the correct shape!, defaulting to identity of " + str(self.slack))
            Qs = np.eye(self.slack)
        if R.shape[0] != self.n_u:
            warnings.warn("R has not the correct shape!, defaulting to identity of " + str(self.n_u))
            R = np.eye(self.n_u)
        self.Q, self.Qs, self.R, self.dR, self.wq = Q, Qs, R, dR, wq
        self._engine = None

    def solve(self, x0, Last_xPredicted, uPred, x_agents, agents_id, pose):
        self.agent_list = agents_id
        if self.first_it:
            self.first_it = False
            self.n_agents = len(agents_id)
            self._engine = BatchedPlannerLPV(self.Q, self.Qs, self.R, self.dR, self.N, self.dt, self.map, self.wq,
                                             self._mp, self._sl, nn=self.n_agents, max_batch=1)
        N, nn = self.N, self.n_agents
        xprev = np.asarray(Last_xPredicted, dtype=np.float64)
        if x_agents is None:
            if nn != 0:
                raise DmpcError("x_agents=None with %d neighbours: the reference's placeholder planes (N,nn,3) are "
                                "inconsistent with its own (N,3,nn) indexing (LPV_Planner.py:132-135)" % nn)
            x_agents = np.zeros((N + 1, 0, 2))
        uold = np.array([[self.OldSteering[0], self.OldAccelera[0]]], dtype=np.float64)
        out = self._engine.solve_batch_host(np.asarray(x0, dtype=np.float64)[None], xprev[None],
                                            np.asarray(uPred, dtype=np.float64)[None], uold,
                                            np.asarray(x_agents, dtype=np.float64)[None],
                                            np.asarray(pose, dtype=np.float64)[None])
        status = int(out["status"][0])
        if status in (_abi.MAP_LOOKUP_FAILED, _abi.DEGENERATE_PLANE):
            # the reference raises inside curvature()/get_ey() (misc.py:97,123) resp. divides by zero (compute_plane.py:52)
            raise DmpcError("LPV solve failed before the QP: status %d" % status)
        feasible = 1 if status in FEASIBLE_STATUS else 0
        if status != _abi.SOLVED:
            print("OSQP exited with status '%s'" % status)
        if feasible == 0:
            print('QUIT...')
        Solution = out["z"][0]
        self.status_val, self.obj_val = status, float(out["obj"][0])
        self.y_ineq, self.y_eq = out["y_ineq"][0], out["y_eq"][0]
        self.planes = out["planes"][0]
        self.xPred, self.uPred, self.sPred, self.duPred, self.raw_States = unpack_solution(Solution, N)
        self.OldSteering = [self.uPred[0, 0]]
        self.OldAccelera = [self.uPred[0, 1]]
        return feasible, Solution, self.planes


class LPV_Model(object):
    """Drop-in for plan_lib.distributedPlanner.LPV_Model (LPV_val.py:7-94): same constructor, `xPred` attribute and
    `sim(u)`; the roll-out runs in `lpv_sim_kernel` (a batch of one; `BatchedPlannerLPV.sim_batch` is the batched form)."""

    # LPV_val.py:27-35,45-53: this class has its own defaults, different from PlannerLPV's
    _MODEL = dict(lf=0.125, lr=0.125, m=1.98, I=0.06, Cf=60.0, Cr=60.0, mu=0.05)
    _LIMITS = dict(vx_ref=4.5, max_vel=5, min_vel=0.2, max_rs=0.45, max_ls=0.45, max_ac=4.0, max_dc=3.0, sm=0.9)

    def __init__(self, N, dt, map, model_param=None, sys_lim=None):
        self.n_s, self.slack, self.n_u = 9, 3, 2
        self.xPred = np.ones((N + 1, self.n_s))
        self.dt, self.map, self.N = dt, map, N
        self.first_it = True
        self.OldSteering = [0.0]
        self.OldAccelera = [0.0]
        mp = dict(self._MODEL if model_param is None else model_param)
        sl = dict(_abi.DEFAULT_LIMITS)
        sl.update(self._LIMITS if sys_lim is None else sys_lim)
        for k in ("lf", "lr", "m", "I", "Cf", "Cr", "mu"):
            setattr(self, k, mp[k])
        for k in ("vx_ref", "max_vel", "min_vel", "max_rs", "max_ls", "max_ac", "max_dc", "sm"):
            setattr(self, k, sl[k])
        self._engine = BatchedPlannerLPV(np.eye(9), np.eye(3), np.eye(2), np.eye(2), N, dt, map, 0, mp, sl, nn=0, max_batch=1)

    def sim(self, u):
        torch = _torch()
        dev = self._engine.device
        xp = torch.from_numpy(np.ascontiguousarray(self.xPred, dtype=np.float64)[None]).to(dev)
        uu = torch.from_numpy(np.ascontiguousarray(u, dtype=np.float64)[None]).to(dev)
        out, st = self._engine.sim_batch(xp, uu)
        if int(st.cpu()[0]) != 0:
            raise DmpcError("curvature()/get_ey() find no segment (misc.py:97,123): status %d" % int(st.cpu()[0]))
        self.xPred = out[0].cpu().numpy()
        return self.xPred

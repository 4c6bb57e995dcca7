"""TEST INFRASTRUCTURE ONLY (oracle/): import the reference's own QP-assembly functions.

This module imports the *unmodified* reference package from ``/root/reference/planner/lib``
(read-only mount, present only in the build container) so that golden fixtures can be
generated from the reference's own code.  ``osqp`` (requirements.txt:5) is not installed
and cannot be installed here, so a stub module is registered under ``sys.modules['osqp']``
before the import; every reference function except ``osqp_solve_qp``
(LPV_Planner.py:192-249) then runs as shipped.

Nothing under ``colaborativempc_b200`` imports this file.  It is used by
``tests/golden/make_golden.py`` (fixture generator) and by CPU tests that are skipped
when ``/root/reference`` is absent (e.g. on the GPU box).
"""
import os
import sys
import types
import warnings

REF_LIB = "/root/reference/planner/lib"


def available():
    return os.path.isdir(os.path.join(REF_LIB, "plan_lib"))


def load():
    """Returns a namespace with the reference's hot-path callables."""
    if not available():
        raise RuntimeError("reference tree not present at %s" % REF_LIB)
    if "osqp" not in sys.modules:
        try:
            import osqp  # noqa: F401
        except Exception:
            stub = types.ModuleType("osqp")

            class OSQP(object):  # placeholder: the solve itself is restated in oracle/
                def __init__(self, *a, **k):
                    raise RuntimeError("osqp is not installed; use oracle/ restatement")

            stub.OSQP = OSQP
            sys.modules["osqp"] = stub
    if REF_LIB not in sys.path:
        sys.path.insert(0, REF_LIB)
    warnings.filterwarnings("ignore", category=DeprecationWarning)
    import numpy as np
    fmt = np.get_printoptions()
    from plan_lib.distributedPlanner import LPV_Planner as P
    from plan_lib.distributedPlanner import LPV_val as V
    from plan_lib.planes import hyperplane_separator
    from plan_lib.utilities import misc
    from plan_lib.mapManager.track_initialization import Map
    from plan_lib.config import x0_database
    np.set_printoptions(**{k: fmt[k] for k in ("formatter",)})  # reference clobbers it on import
    ns = types.SimpleNamespace(
        PlannerLPV=P.PlannerLPV, EstimateABC=P._EstimateABC, buildMatIneqConst=P._buildMatIneqConst,
        buildMatEqConst=P._buildMatEqConst, buildMatCost=P._buildMatCost,
        hyperplane_separator=hyperplane_separator, compute_weights=misc.compute_weights,
        curvature=misc.curvature, get_ey=misc.get_ey, initialise_agents=misc.initialise_agents,
        predicted_vectors_generation=misc.predicted_vectors_generation, checkEnd=misc.checkEnd,
        Map=Map, x0_database=x0_database, LPV_Model=V.LPV_Model)
    return ns


# Values the shipped experiments use (scripts/config_files/config_LPV.py:6-11,
# plan_lib/config/base_class.py:19-42).  Restated as data, not copied code.
def shipped_settings(np):
    Q = np.diag([10.0, 0.0, 0.0, 25.0, 10.0, 0.0, 0.0, 0.0, 0.0])
    return dict(
        Q=Q, Qs=1e7 * np.eye(3), R=np.zeros((2, 2)), dR=50.0 * np.eye(2), wq=5.0, dt=0.025,
        model_param=dict(lf=0.125, lr=0.125, m=1.98, I=0.09, Cf=70.0, Cr=70.0, mu=0.05),
        sys_lim=dict(vx_ref=3.0, min_dist=0.25, max_vel=5.5, min_vel=0.0, max_rs=0.3, max_ls=0.3,
                     max_ac=5.0, max_dc=10.0, sm=0.9, LPV=True))


def assemble(ref, ctrl, x0, Last_xPredicted, uPred, x_agents, agents_id, pose, uOld):
    """Runs steps 1-7 of PlannerLPV.solve (LPV_Planner.py:124-153) on `ctrl` and returns the
    OSQP-form QP (P, q, A, l, u) exactly as osqp_solve_qp (:222-228) would hand it to OSQP."""
    import numpy as np
    ctrl.agent_list = agents_id
    if ctrl.first_it:
        ctrl.first_it = False
        ctrl.n_agents = len(agents_id)
        ctrl.plane_comp = ref.hyperplane_separator(ctrl.n_agents, ctrl.N)
    ctrl.planes = ctrl.plane_comp.compute_hyperplane(x_agents, pose, ctrl.id, agents_id, keep_sign=True)
    ctrl.weights, ctrl.dist = ref.compute_weights(pose, x_agents, ctrl.min_dist)
    ctrl.A, ctrl.B, ctrl.C, ey = ref.EstimateABC(ctrl, Last_xPredicted, uPred)
    F, b = ref.buildMatIneqConst(ctrl, ey)
    G, E, L, Eu, Eoa = ref.buildMatEqConst(ctrl)
    M, q = ref.buildMatCost(ctrl)
    beq = np.add(np.dot(E, x0), np.dot(Eu, uOld))
    A = np.vstack([F, G])
    l = np.hstack([-np.inf * np.ones(len(b)), beq])
    u = np.hstack([b, beq])
    return dict(P=M, q=q, A=A, l=l, u=u, F=F, b=b, G=G, beq=beq, planes=ctrl.planes.copy(),
                weights=ctrl.weights.copy(), Ak=np.array(ctrl.A), Bk=np.array(ctrl.B), ey=np.array(ey))

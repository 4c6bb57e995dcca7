"""Start-up predictions, synthetic platoon scenarios and the host-side DMPC bookkeeping.

Host (numpy) code only: everything here runs once per experiment or is pure index bookkeeping; the
per-iteration arithmetic is in csrc/.  Reference semantics restated:
  startup_prediction   plan_lib/utilities/misc.py:168-210 (predicted_vectors_generation)
  initialise_agents    plan_lib/utilities/misc.py:155-165
  shipped_config       planner/scripts/config_files/config_LPV.py:6-32, plan_lib/config/base_class.py:19-42,
                       plan_lib/config/__init__.py:3-7 (x0_database)
  dmpc_shift/exchange  planner/scripts/LPV_HP_N_main.py:106,113-117
"""
import numpy as np

# x0_database (plan_lib/config/__init__.py:3-7), [vx vy psidot e_y e_psi theta s x y]
X0_DATABASE = [
    [1.3, -0.16, 0.00, 0.0, 0, 0.0, 0, 0.0, 1.0],
    [1.3, -0.16, 0.00, -0.25, 0, 0.0, 0, 0.0, 1.0],
    [1.3, -0.16, 0.00, 0.45, 0, 0.0, 0, 0.0, 1.45],
    [1.3, -0.16, 0.00, 0.25, 0, 0.0, 0.25, 0.0, 1.5],
]


def shipped_config():
    """Gains, vehicle and limits of the shipped LPV experiments."""
    return dict(
        Q=np.diag([10.0, 0.0, 0.0, 25.0, 10.0, 0.0, 0.0, 0.0, 0.0]), Qs=1e7 * np.eye(3), R=np.zeros((2, 2)),
        dR=50.0 * np.eye(2), wq=5.0, dt=0.025, N=125, map_type="Highway",
        model_param=dict(lf=0.125, lr=0.125, m=1.98, I=0.09, Cf=70.0, Cr=70.0, mu=0.05),
        sys_lim=dict(vx_ref=3.0, min_dist=0.25, max_vel=5.5, min_vel=0.0, max_rs=0.3, max_ls=0.3, max_ac=5.0,
                     max_dc=10.0, sm=0.9, LPV=True))


def startup_prediction(N, x0, dt, track, accel_rate=0.0, s0=0.0):
    """(N+1,9) state guess and (N,2) zero inputs.  With s0=0 this is the reference's start-up:
    s starts at 0 whatever x0[6] says, vy/psidot/e_y/e_psi are held, vx ramps with 1 m/s^2, and the
    Cartesian pose of row i+1 is evaluated at (s_i, e_y_i) (one row late, misc.py:205)."""
    x0 = np.asarray(x0, dtype=np.float64)
    xx = np.zeros((N + 1, 9))
    xx[:, 1], xx[:, 2], xx[:, 3], xx[:, 4] = x0[1], x0[2], x0[3], x0[4]
    acc = 1.0 + accel_rate * np.arange(N)
    xx[0, 0], xx[0, 6] = x0[0], s0
    for i in range(N):
        xx[i + 1, 0] = xx[i, 0] + acc[i] * dt
        xx[i + 1, 6] = xx[i, 6] + xx[i, 0] * dt
    px, py, th = track.getGlobalPosition(xx[0, 6], xx[0, 3])
    xx[0, 7], xx[0, 8], xx[0, 5] = px, py, th
    for i in range(N):
        px, py, th = track.getGlobalPosition(xx[i, 6], xx[i, 3])
        xx[i + 1, 7], xx[i + 1, 8], xx[i + 1, 5] = px, py, th
    return xx, np.zeros((N, 2))


def initialise_agents(x0_list, N, dt, tracks, s0_list=None):
    """agents (N+1, n, 2), x_pred list of (N+1,9), u_pred list of (N,2)."""
    n = len(x0_list)
    agents = np.zeros((N + 1, n, 2))
    xs, us = [], []
    for i, x0 in enumerate(x0_list):
        tr = tracks[i] if isinstance(tracks, (list, tuple)) else tracks
        xx, uu = startup_prediction(N, x0, dt, tr, s0=0.0 if s0_list is None else s0_list[i])
        agents[:, i, :] = xx[:, 7:9]
        xs.append(xx)
        us.append(uu)
    return agents, xs, us


def platoon_scenarios(S, A, N, dt, track, seed0=0):
    """S synthetic platoons of A agents: seed = seed0 + scenario index.
    Returns x0 [S,A,9], xprev [S,A,N+1,9], uprev [S,A,N,2] (cold start)."""
    x0 = np.zeros((S, A, 9))
    xprev = np.zeros((S, A, N + 1, 9))
    uprev = np.zeros((S, A, N, 2))
    for s in range(S):
        rng = np.random.default_rng(seed0 + s)
        L = float(track.TrackLength[track.lane]) - 3.0  # keep the whole horizon on the (open) track
        gap_hi = 1.0 if A <= 1 else min(1.0, max(0.5, (L - 1.0) / (A - 1)))
        tail = rng.uniform(0.0, max(0.0, min(30.0, L - (A - 1) * gap_hi)))
        gaps = rng.uniform(0.45, gap_hi, size=A)
        gaps[0] = 0.0
        spos = tail + gaps.sum() - np.cumsum(gaps)  # agent 0 leads, followers behind, tail at `tail`
        ey = rng.uniform(-0.45, 0.45, size=A)
        vx = rng.uniform(0.5, 3.5, size=A)
        vy = rng.uniform(-0.2, 0.2, size=A)
        wz = rng.uniform(-0.3, 0.3, size=A)
        ep = rng.uniform(-0.1, 0.1, size=A)
        for a in range(A):
            st = [vx[a], vy[a], wz[a], ey[a], ep[a], 0.0, 0.0, 0.0, 0.0]
            xx, uu = startup_prediction(N, st, dt, track, s0=float(spos[a]))
            xprev[s, a], uprev[s, a] = xx, uu
            x0[s, a] = xx[0]
    return x0, xprev, uprev


def topology(A, kind="full", nn=None):
    """Neighbour id lists per agent (agents_id of PlannerLPV.solve).  'full' is the reference convention
    (LPV_HP_N_main.py:82-85); 'chain' keeps the predecessor/successor only.  Returns (idx [A,nn] int32 with
    -1 padding, count [A])."""
    lists = []
    for a in range(A):
        if kind == "full":
            lists.append([j for j in range(A) if j != a])
        elif kind == "chain":
            lists.append([j for j in (a - 1, a + 1) if 0 <= j < A])
        else:
            raise ValueError(kind)
    width = max(len(l) for l in lists) if nn is None else nn
    idx = -np.ones((A, max(width, 1)), dtype=np.int32)
    cnt = np.zeros(A, dtype=np.int32)
    for a, l in enumerate(lists):
        idx[a, :len(l)] = l
        cnt[a] = len(l)
    return idx[:, :max(width, 0)] if width > 0 else idx[:, :0], cnt


def gather_neighbours(xy_all, nbr_idx):
    """Host restatement of the exchange indexing (LPV_HP_N_main.py:100,117):
    xy_all [S,A,N+1,2] -> nbr_xy [S,A,N+1,nn,2], pose [S,A,N+1,2].  Empty slots copy the ego pose shifted by
    +1e3 m so that they are geometrically inert if ever read."""
    S, A, T, _ = xy_all.shape
    nn = nbr_idx.shape[1]
    nbr = np.zeros((S, A, T, nn, 2))
    for a in range(A):
        for j in range(nn):
            k = int(nbr_idx[a, j])
            nbr[:, a, :, j, :] = xy_all[:, k] if k >= 0 else xy_all[:, a] + 1e3
    return nbr, xy_all.copy()


def dmpc_shift(z, N):
    """Caller bookkeeping after one solve for a batch z [B,16N+12] (LPV_HP_N_main.py:106,113-115 and
    LPV_Planner.py:179-180): returns x0 [B,9], xprev [B,N,9] (shifted), uprev [B,N,2] (NOT shifted),
    uold [B,2], xy [B,N+1,2]."""
    z = np.asarray(z)
    B = z.shape[0]
    raw = z[:, :12 * (N + 1)].reshape(B, N + 1, 12)
    xpred = raw[:, :, :9]
    upred = z[:, 12 * (N + 1):12 * (N + 1) + 2 * N].reshape(B, N, 2)
    return (xpred[:, 1].copy(), xpred[:, 1:].copy(), upred.copy(), upred[:, 0].copy(), xpred[:, :, 7:9].copy())


def run_dmpc(solve_fn, x0, xprev, uprev, nbr_idx, nbr_cnt, iters, N, uold=None, record=None):
    """Host-driven Jacobi DMPC loop over S scenarios x A agents (LPV_HP_N_main.py:96-117).
    solve_fn(x0 [B,9], xprev [B,rows,9], uprev [B,N,2], uold [B,2], nbr_xy [B,N+1,nn,2], pose [B,N+1,2],
             nbr_count [B]) -> dict with 'z' [B,16N+12] and 'status' [B];  B = S*A, scenario-major.
    Returns the state dict to feed the next iteration (and is what fixtures store)."""
    S, A = x0.shape[:2]
    B = S * A
    x0 = x0.reshape(B, 9).copy()
    xprev = xprev.reshape(B, -1, 9).copy()
    uprev = uprev.reshape(B, N, 2).copy()
    uold = np.zeros((B, 2)) if uold is None else uold.reshape(B, 2).copy()
    xy = xprev[:, :, 7:9].reshape(S, A, -1, 2)[:, :, :N + 1].copy()
    cnt = np.tile(np.asarray(nbr_cnt, dtype=np.int32), S)
    out = None
    for it in range(iters):
        nbr, pose = gather_neighbours(xy, nbr_idx)
        nbr = nbr.reshape(B, N + 1, -1, 2)
        pose = pose.reshape(B, N + 1, 2)
        inputs = dict(x0=x0, xprev=xprev, uprev=uprev, uold=uold, nbr_xy=nbr, pose=pose, nbr_count=cnt)
        out = solve_fn(**inputs)
        if record is not None:
            record.append((inputs, out))
        if not np.all(np.isin(out["status"], (1, 2, -2))):
            break
        x0, xprev, uprev, uold, xyb = dmpc_shift(out["z"], N)
        xy = xyb.reshape(S, A, N + 1, 2)
    nbr, pose = gather_neighbours(xy, nbr_idx)
    return dict(x0=x0, xprev=xprev, uprev=uprev, uold=uold, nbr_xy=nbr.reshape(B, N + 1, -1, 2),
                pose=pose.reshape(B, N + 1, 2), nbr_count=cnt, last=out)

"""Sharded DMPC iteration: the loop body of planner/scripts/LPV_HP_N_main.py:96-117 for S scenarios x A agents,
agents partitioned across the ranks of a torch.distributed group (one process per GPU).

One iteration =  gather_neighbours (CUDA; also reduces the stop flags every rank received)
                 ->  solve (CUDA, one warp per agent QP; its epilogue shifts the horizon, packs the (x,y) predictions
                     for the exchange, holds the last valid plan of a failed QP and raises the stop flags)
                 ->  all_gather of [packed predictions | stop flags] (NCCL over NVLink; the one exchange step).

The exchange replaces `agents = swapaxes(asarray(x_pred)[:,:,-2:],0,1)` (LPV_HP_N_main.py:117) and the ROS
publish/callback pair (LPV_ROS_main.py:71-77).  The reference's two ways out of its loop - any agent of the experiment
at the end of the track (`checkEnd`, utilities/misc.py:28-48, LPV_HP_N_main.py:96) and any agent's QP not feasible
(:102-104,110-111) - travel as one flag per agent in the tail of the same all-gather, so every rank derives the same
per-scenario `stop` from the same bytes and no rank can leave the loop alone.  Layout conventions are shared with the
host mirrors below so that the world_size-2 gloo tests exercise the same index arithmetic.
"""
import numpy as np


def shard_layout(A, world, rank):
    """Agents are split in contiguous blocks: rank r owns agents [r*A_local, (r+1)*A_local)."""
    if A % world != 0:
        raise ValueError("agents per scenario (%d) must be divisible by the number of ranks (%d)" % (A, world))
    a_local = A // world
    return rank * a_local, a_local


def allgather_index(k, s, S, a_local):
    """Offset (in trajectories) of global agent k, scenario s inside the all-gathered buffer [G,S,A_local,...]."""
    return ((k // a_local) * S + s) * a_local + (k % a_local)


def payload_doubles(S, a_local, T):
    """Doubles one rank sends per iteration: packed (x,y) predictions [S,A_local,T,2] + one stop flag per agent."""
    return S * a_local * (2 * T + 1)


def host_exchange_layout(xy_all_ranks, nbr_idx_local, agent_offset):
    """numpy mirror of dmpc_gather_neighbours on the all-gather layout.
    xy_all_ranks [G,S,A_local,T,2]; nbr_idx_local [A_local,nn] global ids (-1 empty).
    Returns nbr_xy [S*A_local,T,nn,2], pose [S*A_local,T,2]."""
    G, S, a_local, T, _ = xy_all_ranks.shape
    nn = nbr_idx_local.shape[1]
    flat = xy_all_ranks.reshape(G * S * a_local, T, 2)
    nbr = np.zeros((S, a_local, T, nn, 2))
    pose = np.zeros((S, a_local, T, 2))
    for s in range(S):
        for al in range(a_local):
            me = flat[allgather_index(agent_offset + al, s, S, a_local)]
            pose[s, al] = me
            for j in range(nn):
                k = int(nbr_idx_local[al, j])
                nbr[s, al, :, j, :] = flat[allgather_index(k, s, S, a_local)] if k >= 0 else me + 1e3
    return nbr.reshape(S * a_local, T, nn, 2), pose.reshape(S * a_local, T, 2)


def host_split_payload(all_ranks, S, a_local, T):
    """numpy mirror of the payload layout of dmpc_gather_neighbours_stop: all_ranks [G, S*a_local*(2T+1)] ->
    (xy [G,S,A_local,T,2], flags [G,S,A_local])."""
    G = all_ranks.shape[0]
    n_xy = S * a_local * T * 2
    return (all_ranks[:, :n_xy].reshape(G, S, a_local, T, 2), all_ranks[:, n_xy:].reshape(G, S, a_local))


def host_stop_reduce(stop, flags_all_ranks):
    """numpy mirror of the stop reduction: stop [S] (sticky) |= any flag of any agent of the scenario on any rank."""
    return np.asarray(stop, dtype=bool) | np.any(np.asarray(flags_all_ranks) != 0.0, axis=(0, 2))


def end_flag(s, track_length, laps=1):
    """utilities/misc.py:28-48 for one agent: s (row 0 of its prediction) within 0.15 m of the end (np.isclose) or beyond
    it, on lap `laps` (track_initialization.py:319-323)."""
    L = float(track_length)
    return bool((np.isclose(s, L, atol=0.15) or s > L) and np.floor(s / L) == laps)


class ShardedDMPC(object):
    """Device-resident DMPC state of this rank: S scenarios x A_local agents (B = S*A_local QPs per iteration)."""

    def __init__(self, engine, S, A, nbr_idx, nbr_cnt, x0, xprev, uprev, group=None, rank=0, world=1, end_s=None, laps=1):
        """engine: BatchedPlannerLPV with max_batch >= S*A_local.  nbr_idx [A,nn] / nbr_cnt [A]: topology (global ids).
        x0 [S,A,9], xprev [S,A,N+1,9], uprev [S,A,N,2]: cold start of ALL agents (numpy); this rank keeps its slice.
        end_s: track length of the end rule (default: the engine's map)."""
        import torch
        self.torch = torch
        self.eng, self.S, self.A, self.rank, self.world, self.group = engine, int(S), int(A), rank, world, group
        self.N, self.nn = engine.N, engine.nn
        self.off, self.a_local = shard_layout(A, world, rank)
        self.B = self.S * self.a_local
        dev = engine.device
        sl = slice(self.off, self.off + self.a_local)
        f = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        N, T = self.N, self.N + 1
        self.fused = not engine.fast  # the stage-wise engine's epilogue does the bookkeeping; the condensed one needs a kernel
        if self.fused:
            engine.set_options(warm_start=2)  # consecutive closed-loop steps: previous working set, one stage on
        self.x0 = f(x0[:, sl].reshape(self.B, 9))
        self.xprev_first = f(xprev[:, sl].reshape(self.B, T, 9))      # N+1 rows on the first call only
        self.xprev = torch.zeros((self.B, N, 9), dtype=torch.float64, device=dev)
        self.uprev = f(uprev[:, sl].reshape(self.B, N, 2))
        self.uold = torch.zeros((self.B, 2), dtype=torch.float64, device=dev)
        self.nbr_idx = f(np.tile(np.asarray(nbr_idx, np.int32)[sl], (self.S, 1)).reshape(self.B, -1))
        self.nbr_cnt = f(np.tile(np.asarray(nbr_cnt, np.int32)[sl], self.S))
        # one contiguous payload per rank: [xy_send | flags]; the views below are what the kernels write
        n_xy = self.B * T * 2
        self.send = torch.zeros((payload_doubles(self.S, self.a_local, T),), dtype=torch.float64, device=dev)
        self.xy_send = self.send[:n_xy].view(self.B, T, 2)
        self.flags = self.send[n_xy:]
        self.xy_send.copy_(f(xprev[:, sl, :, 7:9].reshape(self.B, T, 2)))
        self.all = torch.zeros((world, self.send.numel()), dtype=torch.float64, device=dev)
        self.hold = torch.zeros((self.B,), dtype=torch.int32, device=dev)
        self.stop = torch.zeros((self.S,), dtype=torch.int32, device=dev)
        self.nbr_xy = torch.zeros((self.B, T, max(self.nn, 0), 2), dtype=torch.float64, device=dev)
        self.pose = torch.zeros((self.B, T, 2), dtype=torch.float64, device=dev)
        self.out = engine.alloc_outputs(self.B, duals=False)
        self.end_s = float(end_s) if end_s is not None else float(engine.track_length)
        self.laps = int(laps)
        self.first = True
        self.graph = None
        self.launches_per_iteration = 3 if self.fused else 6
        self.exchange()

    def _shift(self):
        d = dict(x0=self.x0, xprev=self.xprev, uprev=self.uprev, uold=self.uold, xy_send=self.xy_send, hold=self.hold)
        if self.fused:
            d.update(flags=self.flags, stop=self.stop, end_s=self.end_s, laps=self.laps, stop_div=self.a_local)
        return d

    def exchange(self):
        """The one collective of a DMPC iteration: all ranks' [packed (x,y) predictions | stop flags] -> self.all [G, .]."""
        if self.world == 1:
            self.all.view(-1).copy_(self.send)
        else:
            self.torch.distributed.all_gather_into_tensor(self.all.view(-1), self.send, group=self.group)

    def _body(self, xp):
        e = self.eng
        e.gather_neighbours(self.all, self.nbr_idx, self.nbr_xy, self.pose, self.S, self.A, self.a_local, self.off, stop=self.stop)
        if self.fused:
            e.solve_batch(self.x0, xp, self.uprev, self.uold, self.nbr_xy, self.pose, self.nbr_cnt, out=self.out, duals=False,
                          shift=self._shift())
        else:
            e.solve_batch(self.x0, xp, self.uprev, self.uold, self.nbr_xy, self.pose, self.nbr_cnt, out=self.out, duals=False)
            e.shift_pack_hold(self.out["z"], self.out["status"], self.x0, self.xprev, self.uprev, self.uold, self.xy_send, self.hold)
        self.exchange()

    def capture(self):
        """Captures one iteration (gather -> solve+bookkeeping -> launch-order kernel -> all-gather) into a CUDA graph;
        `iterate` replays it from then on.  Call after at least two eager iterations (the first one reads the N+1-row
        history, and the handle settles its launch-order / warm-start state on the first call of a batch size).  The call
        itself advances the loop by ONE iteration (the warm-up run CUDA graphs need), whose outputs are in `self.out`."""
        torch = self.torch
        if self.first:
            raise RuntimeError("capture() after the first iteration")
        s = torch.cuda.Stream(device=self.eng.device)
        s.wait_stream(torch.cuda.current_stream(self.eng.device))
        n0 = self.eng.launches
        with torch.cuda.stream(s):
            self._body(self.xprev)
        self.launches_per_iteration = self.eng.launches - n0   # kernels of this library in one iteration (= in one replay)
        torch.cuda.current_stream(self.eng.device).wait_stream(s)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._body(self.xprev)
        self.graph = g
        return g

    def release(self):
        """Drops the captured graph (it holds NCCL work: release it BEFORE torch.distributed.destroy_process_group, which
        otherwise waits forever on the communicator the graph still references).  Synchronises."""
        self.graph = None
        self.torch.cuda.synchronize()

    def iterate(self):
        """One DMPC iteration; asynchronous on the current stream.  Returns the output dict of the solve."""
        if self.graph is not None:
            self.graph.replay()
            return self.out
        self._body(self.xprev_first if self.first else self.xprev)
        self.first = False
        return self.out

    def iterate_phases(self):
        """One iteration launched eagerly with CUDA events between its phases; returns the device times in ms of
        (gather + stop reduce, solve + bookkeeping + launch order, exchange).  Synchronises.  For measurements."""
        torch, e = self.torch, self.eng
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        xp = self.xprev_first if self.first else self.xprev
        ev[0].record()
        e.gather_neighbours(self.all, self.nbr_idx, self.nbr_xy, self.pose, self.S, self.A, self.a_local, self.off, stop=self.stop)
        ev[1].record()
        if self.fused:
            e.solve_batch(self.x0, xp, self.uprev, self.uold, self.nbr_xy, self.pose, self.nbr_cnt, out=self.out, duals=False,
                          shift=self._shift())
        else:
            e.solve_batch(self.x0, xp, self.uprev, self.uold, self.nbr_xy, self.pose, self.nbr_cnt, out=self.out, duals=False)
            e.shift_pack_hold(self.out["z"], self.out["status"], self.x0, self.xprev, self.uprev, self.uold, self.xy_send, self.hold)
        ev[2].record()
        self.exchange()
        ev[3].record()
        self.first = False
        torch.cuda.synchronize()
        return [ev[i].elapsed_time(ev[i + 1]) for i in range(3)]

    def scenario_flags(self, s):
        """(ended, failed) of scenario s as of the last exchange, over ALL its agents on all ranks: the reference's loop
        condition (checkEnd, LPV_HP_N_main.py:96) and its break (:102-104).  Host copy of G*A_local doubles; synchronises."""
        n_xy = self.B * (self.N + 1) * 2
        fl = self.all[:, n_xy + s * self.a_local:n_xy + (s + 1) * self.a_local].cpu().numpy()
        return bool(np.any(np.mod(fl, 2.0) >= 1.0)), bool(np.any(fl >= 2.0))

    def stopped(self):
        """Per-scenario stop flags as every rank sees them after the last exchange (host copy, synchronises)."""
        e = self.eng
        # the flags of the LAST solve are reduced by the next gather; reduce them now for the caller's loop condition
        e.gather_neighbours(self.all, self.nbr_idx, self.nbr_xy, self.pose, self.S, self.A, self.a_local, self.off, stop=self.stop)
        return self.stop.cpu().numpy().astype(bool)

"""Sharded DMPC iteration: the loop body of planner/scripts/LPV_HP_N_main.py:96-117 for S scenarios x A agents,
agents partitioned across the ranks of a torch.distributed group (one process per GPU).

One iteration =  gather_neighbours (CUDA)  ->  solve_batch (CUDA, one CTA per agent QP)  ->  shift_pack (CUDA)
                 ->  all_gather of the packed (x,y) predictions (NCCL over NVLink; the one exchange step).

The exchange replaces `agents = swapaxes(asarray(x_pred)[:,:,-2:],0,1)` (LPV_HP_N_main.py:117) and the ROS
publish/callback pair (LPV_ROS_main.py:71-77).  Layout conventions are shared with the host mirror in
`host_exchange_layout` so the world_size-2 gloo tests exercise the same index arithmetic.
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


class ShardedDMPC(object):
    """Device-resident DMPC state of this rank: S scenarios x A_local agents (B = S*A_local QPs per iteration)."""

    def __init__(self, engine, S, A, nbr_idx, nbr_cnt, x0, xprev, uprev, group=None, rank=0, world=1):
        """engine: BatchedPlannerLPV with max_batch >= S*A_local.  nbr_idx [A,nn] / nbr_cnt [A]: topology (global ids).
        x0 [S,A,9], xprev [S,A,N+1,9], uprev [S,A,N,2]: cold start of ALL agents (numpy); this rank keeps its slice."""
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
        self.x0 = f(x0[:, sl].reshape(self.B, 9))
        self.xprev_first = f(xprev[:, sl].reshape(self.B, T, 9))      # N+1 rows on the first call only
        self.xprev = torch.zeros((self.B, N, 9), dtype=torch.float64, device=dev)
        self.uprev = f(uprev[:, sl].reshape(self.B, N, 2))
        self.uold = torch.zeros((self.B, 2), dtype=torch.float64, device=dev)
        self.nbr_idx = f(np.tile(np.asarray(nbr_idx, np.int32)[sl], (self.S, 1)).reshape(self.B, -1))
        self.nbr_cnt = f(np.tile(np.asarray(nbr_cnt, np.int32)[sl], self.S))
        self.xy_send = f(xprev[:, sl, :, 7:9].reshape(self.B, T, 2))
        self.xy_all = torch.zeros((world, self.S, self.a_local, T, 2), dtype=torch.float64, device=dev)
        self.nbr_xy = torch.zeros((self.B, T, max(self.nn, 0), 2), dtype=torch.float64, device=dev)
        self.pose = torch.zeros((self.B, T, 2), dtype=torch.float64, device=dev)
        self.out = engine.alloc_outputs(self.B, duals=False)
        self.first = True
        self.exchange()

    def exchange(self):
        """The one collective of a DMPC iteration: all ranks' packed (x,y) predictions -> xy_all [G,S,A_local,T,2]."""
        if self.world == 1:
            self.xy_all.view(-1).copy_(self.xy_send.view(-1))
        else:
            self.torch.distributed.all_gather_into_tensor(self.xy_all.view(-1), self.xy_send.view(-1), group=self.group)

    def iterate(self):
        """One DMPC iteration; asynchronous on the current stream.  Returns the output dict of the solve."""
        e = self.eng
        e.gather_neighbours(self.xy_all, self.nbr_idx, self.nbr_xy, self.pose, self.S, self.A, self.a_local, self.off)
        xp = self.xprev_first if self.first else self.xprev
        e.solve_batch(self.x0, xp, self.uprev, self.uold, self.nbr_xy, self.pose, self.nbr_cnt, out=self.out,
                      duals=False)
        e.shift_pack(self.out["z"], self.x0, self.xprev, self.uprev, self.uold, self.xy_send)
        self.first = False
        self.exchange()
        return self.out

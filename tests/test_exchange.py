"""CPU, world_size 2 over gloo: the one exchange step of a DMPC iteration (LPV_HP_N_main.py:117) - every
rank all-gathers its agents' packed (x,y) predictions, then indexes its neighbours out of the gathered buffer.  The
sharded result must equal the single-process exchange (`scenarios.gather_neighbours`) for every topology."""
import os
import socket

import numpy as np
import pytest

from colaborativempc_b200 import scenarios as sc
from colaborativempc_b200.dmpc import allgather_index, host_exchange_layout, shard_layout


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, S, A, T, nn, kind, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        xy = np.random.default_rng(3).normal(size=(S, A, T, 2))  # the same on every rank; a rank SENDS only its slice
        idx, cnt = sc.topology(A, kind, nn=nn)
        off, al = shard_layout(A, world, rank)
        send = torch.from_numpy(np.ascontiguousarray(xy[:, off:off + al]))
        allb = torch.zeros((world, S, al, T, 2), dtype=torch.float64)
        dist.all_gather_into_tensor(allb.view(-1), send.view(-1))
        nbr, pose = host_exchange_layout(allb.numpy(), np.asarray(idx)[off:off + al], off)
        q.put((rank, nbr, pose))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("kind,A,nn", [("chain", 6, 2), ("full", 4, 3)])
def test_sharded_exchange_equals_single_process(kind, A, nn):
    import torch.multiprocessing as mp
    S, T, world = 3, 11, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, world, port, S, A, T, nn, kind, q)) for r in range(world)]
    for p in ps:
        p.start()
    got = dict()
    for _ in range(world):
        r, nbr, pose = q.get(timeout=120)
        got[r] = (nbr, pose)
    for p in ps:
        p.join(60)
        assert p.exitcode == 0
    xy = np.random.default_rng(3).normal(size=(S, A, T, 2))
    idx, cnt = sc.topology(A, kind, nn=nn)
    ref_nbr, ref_pose = sc.gather_neighbours(xy, np.asarray(idx))
    for r in range(world):
        off, al = shard_layout(A, world, r)
        nbr, pose = got[r]
        assert np.array_equal(pose.reshape(S, al, T, 2), ref_pose[:, off:off + al])
        want = ref_nbr[:, off:off + al]
        valid = (np.asarray(idx)[off:off + al] >= 0)[None, :, None, :, None]
        assert np.array_equal(np.where(valid, nbr.reshape(S, al, T, nn, 2), 0), np.where(valid, want, 0))
        assert np.array_equal(nbr.reshape(S, al, T, nn, 2), want)  # empty slots too: ego pose + 1e3


def _stop_worker(rank, world, port, S, A, T, q):
    """Termination over the exchange: the agent that reaches the end of the track (scenario 1) and the agent whose QP fails
    (scenario 3) both live on rank 1; rank 0 sees neither locally.  Both ranks must stop the same scenarios after the
    same all-gather (LPV_HP_N_main.py:96,102-104; the flags ride in the payload tail)."""
    import torch
    import torch.distributed as dist
    from colaborativempc_b200.dmpc import end_flag, host_split_payload, host_stop_reduce, payload_doubles
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        off, al = shard_layout(A, world, rank)
        L = 47.0
        s_now = np.full((S, A), 10.0)
        s_now[1, A - 1] = L + 0.05           # past the end, lap 1: checkEnd fires (agent on the last rank); L - 0.1 is
        s_now[2, 0] = L - 0.1                # 'close' but still lap 0, and the reference's rule wants lap == laps: no stop
        failed = np.zeros((S, A), bool)
        failed[3, A - 2] = True              # infeasible QP (also on the last rank)
        xy = np.random.default_rng(5).normal(size=(S, A, T, 2))
        stop = np.zeros(S, bool)
        hist = []
        for it in range(3):
            flags = np.array([[1.0 * end_flag(s_now[s, a], L) + 2.0 * failed[s, a] for a in range(off, off + al)] for s in range(S)])
            if it == 0:
                flags[:] = 0.0               # nothing is flagged before the first solve
            send = torch.from_numpy(np.concatenate([xy[:, off:off + al].ravel(), flags.ravel()]))
            assert send.numel() == payload_doubles(S, al, T)
            allb = torch.zeros((world, send.numel()), dtype=torch.float64)
            dist.all_gather_into_tensor(allb.view(-1), send)
            xy_all, fl_all = host_split_payload(allb.numpy(), S, al, T)
            stop = host_stop_reduce(stop, fl_all)
            hist.append(stop.copy())
            assert np.array_equal(xy_all[rank], xy[:, off:off + al])
        q.put((rank, np.array(hist)))
    finally:
        dist.destroy_process_group()


def test_termination_flags_reach_every_rank():
    import torch.multiprocessing as mp
    S, A, T, world = 5, 4, 6, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_stop_worker, args=(r, world, port, S, A, T, q)) for r in range(world)]
    for p in ps:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(world))
    for p in ps:
        p.join(60)
        assert p.exitcode == 0
    assert np.array_equal(got[0], got[1])                          # identical decisions on both ranks, every iteration
    assert not got[0][0].any()
    assert got[0][1].tolist() == [False, True, False, True, False]  # ended (scenario 1) and failed (scenario 3), sticky
    assert got[0][2].tolist() == [False, True, False, True, False]


def test_layout_arithmetic():
    assert shard_layout(16, 4, 2) == (8, 4)
    with pytest.raises(ValueError):
        shard_layout(6, 4, 0)
    S, al = 5, 3
    seen = {allgather_index(k, s, S, al) for k in range(12) for s in range(S)}
    assert seen == set(range(12 * S))

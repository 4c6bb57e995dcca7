"""GPU: the allocation-free entry point dmpc_solve_batch_static (static device scratch, no handle, no cudaMalloc) against
the existing entry point dmpc_solve_batch, through the C ABI, on two calls in a row.  Collected last on purpose (file
name).  No counterpart in the reference."""
import ctypes as C

import numpy as np
import pytest

from conftest import INPUT_KEYS, golden, planner_args, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-6   # the parity tolerance of tests/test_gpu_parity.py


def _io(B, rows, ins, outs):
    from colaborativempc_b200._abi import DmpcBatchIO
    io = DmpcBatchIO()
    io.B, io.xprev_rows = B, rows
    for k in INPUT_KEYS:
        setattr(io, k, None if ins.get(k) is None else ins[k].data_ptr())
    for k in ("z", "planes", "y_ineq", "y_eq", "obj", "status", "iters", "pivots"):
        setattr(io, k, outs[k].data_ptr())
    return io


def test_static_scratch_entry_point_matches_the_handle_path(built, cfg, track):
    import torch
    from colaborativempc_b200 import _abi
    from colaborativempc_b200.planner import BatchedPlannerLPV
    N, nn = 20, 2
    lib = _abi.load_library()
    # the existing entry point, nothing carried between calls (what the static variant computes)
    eng = BatchedPlannerLPV(*planner_args(cfg, track, N), nn=nn, max_batch=64, options={"warm_start": 0, "order_hint": 0})
    params = _abi.make_params(*planner_args(cfg, track, N), nn=nn)
    opt = _abi.DmpcOptions()
    lib.dmpc_default_options(C.byref(opt))
    fx = golden("platoon_N20.npz")
    low = golden("closed_loop_ipm_start_qp.npz")
    dev = torch.device("cuda", torch.cuda.current_device())
    err = C.create_string_buffer(256)
    stream = torch.cuda.current_stream(dev)
    # two calls in a row: a later DMPC iteration (N history rows, interior-point QPs that park their iterate in the scratch),
    # then the first iteration (N + 1 rows) with the crawling-vehicle QP appended - the second call finds the first one's
    # contents in the shared array
    batches = []
    for it in (3, 0):
        b = {k: np.ascontiguousarray(fx["it%d_%s" % (it, k)]) for k in INPUT_KEYS}
        if it == 3:
            b = {k: np.concatenate([v, np.ascontiguousarray(low["in_" + k])], 0) for k, v in b.items()}
        batches.append(b)
    for call, b in enumerate(batches):
        ins = {k: torch.from_numpy(v).to(dev) for k, v in b.items()}
        B, rows = ins["x0"].shape[0], ins["xprev"].shape[1]
        want = eng.solve_batch(**ins)
        got = eng.alloc_outputs(B, duals=True)
        for t in got.values():
            t.fill_(-7)
        torch.cuda.synchronize()
        rc = lib.dmpc_solve_batch_static(C.byref(params), C.byref(opt), C.byref(_io(B, rows, ins, got)), dev.index,
                                         C.c_void_p(stream.cuda_stream), err, len(err))
        assert rc == 0, (rc, err.value)
        torch.cuda.synchronize()
        w = {k: v.cpu().numpy() for k, v in want.items()}
        g = {k: v.cpu().numpy() for k, v in got.items()}
        assert np.all(w["status"] == 1) and np.array_equal(g["status"], w["status"]), call
        scale = np.abs(np.concatenate([w["y_ineq"], w["y_eq"]], 1)).max(1)
        assert rel_err(g["z"], w["z"]).max() < TOL
        assert rel_err(g["obj"][:, None], w["obj"][:, None]).max() < TOL
        assert (np.abs(g["y_ineq"] - w["y_ineq"]).max(1) / (1 + scale)).max() < TOL
        assert (np.abs(g["y_eq"] - w["y_eq"]).max(1) / (1 + scale)).max() < TOL
        assert np.abs(g["planes"] - w["planes"]).max() < 1e-12
        assert np.array_equal(g["iters"], w["iters"]) and np.array_equal(g["pivots"], w["pivots"])
        if call == 0:
            assert g["iters"].max() > 0   # the scratch was really used: some QP went through the interior point
    # beyond the static array: a clear error, nothing launched, outputs untouched
    ins = {k: torch.from_numpy(v).to(dev) for k, v in batches[0].items()}
    got = eng.alloc_outputs(ins["x0"].shape[0], duals=True)
    got["status"].fill_(-7)
    io = _io(ins["x0"].shape[0], ins["xprev"].shape[1], ins, got)
    io.B = 5000                                          # 5000 x 680 doubles > 3 Mi doubles (checked before any access)
    rc = lib.dmpc_solve_batch_static(C.byref(params), C.byref(opt), C.byref(io), dev.index, C.c_void_p(stream.cuda_stream), err, len(err))
    torch.cuda.synchronize()
    assert rc == -4 and b"exceeds the static scratch" in err.value and b"4626" in err.value, (rc, err.value)
    assert int(got["status"][0]) == -7

"""bench.py - batched DMPC QP solves/sec at N=20 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

N=1 workload  = BASELINE config[1]: 4096 independent single-agent LPV-MPC QPs (horizon 20, 2 neighbour slots), synthetic
                platoon scenarios (seed = scenario index).  The timed steps solve a STREAM of consecutive
                closed-loop batches (step i = the 4096 QPs of DMPC iteration 3+i of the same 1366 platoons), so whatever a
                step carries over (launch order, first working-set guess) comes from the PREVIOUS, different batch, as in
                a live MPC loop.  `replay` (the same frozen batch again) and `cold` (nothing carried, first sight of a
                batch) are reported beside it.  A step = one pass of the hot path (planes -> linearisation -> QP solve ->
                unpack, one kernel launch) over the 4096 QPs.
N>1 workload  = the DMPC iteration with its exchange: 256*N scenarios x 16-agent chain platoons, agents sharded over
                the N ranks (4096 QPs per GPU per iteration, weak scaling); a step = gather -> solve (+ shift/pack in its
                epilogue) -> NCCL all-gather of [(x,y) predictions | stop flags].  `weak_baseline` = the same workload
                (256 x 16) on rank 0 alone, timed in the same run.
--impl reference = the reference's CPU path for the same QPs (oracle restatement of PlannerLPV.solve with OSQP at the
                reference's settings) on all host threads, bounded sample per step.
The run FAILS (exit code 3) on any numerical breakdown (status -10/-20/-21) and if more than 1 % of the timed QPs are not
feasible; a genuinely infeasible QP stops its scenario (as the reference stops the experiment) and is reported.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

HORIZON, NN, BATCH = 20, 2, 4096
INPUT_KEYS = ("x0", "xprev", "uprev", "uold", "nbr_xy", "pose", "nbr_count")
METRIC = "batched DMPC QP solves/sec at N=20"


def planner_args(cfg, track, N):
    return (cfg["Q"], cfg["Qs"], cfg["R"], cfg["dR"], N, cfg["dt"], track, cfg["wq"], cfg["model_param"], cfg["sys_lim"])


def algorithmic_bytes(N, nn, duals):
    """inputs x0, Last_xPredicted (N rows), uPred, uOld, x_agents, pose (+count); outputs Solution, planes,
    obj/status/iters (+ full dual y when asked)."""
    b_in = 8 * (9 + 9 * N + 2 * N + 2 + 2 * (N + 1) * (nn + 1)) + 4
    b_out = 8 * ((16 * N + 12) + 3 * N * nn + 1) + 12
    if duals:
        b_out += 8 * ((8 + nn) * N + 14 * N + 12)
    return b_in, b_out


def load_json(*rel):
    try:
        return json.load(open(os.path.join(ROOT, *rel)))
    except Exception:
        return {}


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU while the timed region runs (B200_PROFILING.md clocks line)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                 "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                 "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                 "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap,
                 "hw_power_brake": nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.02)

    def result(self):
        self.stop_flag = True
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def stream_capacity(n_qp, A=3):
    """QPs per iteration the workload generator simulates for n_qp kept QPs (3 % spare platoons)."""
    S_need = (n_qp + A - 1) // A
    return (S_need + max(4, S_need // 32)) * A


def closed_loop_stream(solve_fn, n_qp, n_batches, cfg, track, N=HORIZON, nn=NN, A=3, topo="full", skip=3, seed0=0):
    """`n_batches` consecutive DMPC iterations of synthetic platoons driven by `solve_fn`, the first `skip` (the cold
    start) dropped; returns the list of input dicts (n_qp QPs per iteration).  A platoon whose experiment the reference
    would have stopped inside the window (a QP that is not feasible: crawling vehicles whose linearised model has spun out
    do produce genuinely infeasible QPs, the oracle agrees) is left out: 3 % spare platoons are simulated, a stopped one
    is frozen (LPV_HP_N_main.py:102-104) and the first ceil(n_qp/A) platoons that run through the window are kept."""
    from colaborativempc_b200 import scenarios as sc
    S_need = (n_qp + A - 1) // A
    B = stream_capacity(n_qp, A)
    S = B // A
    x0, xprev, uprev = sc.platoon_scenarios(S, A, N, cfg["dt"], track, seed0=seed0)
    idx, cnt = sc.topology(A, topo, nn=nn)
    x0 = x0.reshape(B, 9).copy()
    xprev = xprev.reshape(B, -1, 9).copy()
    uprev = uprev.reshape(B, N, 2).copy()
    uold = np.zeros((B, 2))
    xy = xprev[:, :, 7:9].reshape(S, A, -1, 2)[:, :, :N + 1].copy()
    cntb = np.tile(np.asarray(cnt, dtype=np.int32), S)
    stop = np.zeros(S, dtype=bool)
    rec = []
    for it in range(skip + n_batches):
        nbr, pose = sc.gather_neighbours(xy, idx)
        ins = dict(x0=x0.copy(), xprev=xprev.copy(), uprev=uprev.copy(), uold=uold.copy(),
                   nbr_xy=nbr.reshape(B, N + 1, -1, 2).copy(), pose=pose.reshape(B, N + 1, 2).copy(), nbr_count=cntb)
        out = solve_fn(**ins)
        rec.append(ins)
        feas = np.isin(out["status"], (1, 2, -2))
        stop |= (~feas).reshape(S, A).any(1)
        ok = feas & ~np.repeat(stop, A)
        nx0, nxp, nup, nuo, nxy = sc.dmpc_shift(out["z"], N)
        if xprev.shape[1] != N:
            xprev = np.ascontiguousarray(xprev[:, 1:])
        x0[ok], xprev[ok], uprev[ok], uold[ok] = nx0[ok], nxp[ok], nup[ok], nuo[ok]
        xy.reshape(B, N + 1, 2)[ok] = nxy[ok]
    alive = np.flatnonzero(~stop)
    if alive.size < S_need:
        raise SystemExit("workload generation: only %d of %d platoons ran through the window" % (alive.size, S))
    rows = (alive[:S_need, None] * A + np.arange(A)[None, :]).ravel()[:n_qp]
    return [{k: np.ascontiguousarray(ins[k][rows]) for k in INPUT_KEYS} for ins in rec[skip:]]


def steady_state_inputs(solve_fn, n_qp, cfg, track, A=3, topo="full", warm_iters=3, seed0=0, N=HORIZON, nn=NN):
    """Config-2 inputs: ceil(n_qp/A) synthetic platoons, `warm_iters` DMPC iterations with `solve_fn`, then frozen."""
    return closed_loop_stream(solve_fn, n_qp, 1, cfg, track, N=N, nn=nn, A=A, topo=topo, skip=warm_iters, seed0=seed0)[0]


# ------------------------------------------------------------------------------------------------------------
# CPU legs (the only places bench.py executes oracle/): cpu_baseline of the GPU arm and the --impl reference arm.
def oracle_solver(cfg, track, nthreads, exact, N=HORIZON, nn=NN):
    from colaborativempc_b200._abi import make_params
    from oracle import lpv_oracle as lo
    P = make_params(*planner_args(cfg, track, N), nn=nn)
    st = lo.settings(exact=exact)  # exact=False: OSQP at the reference's settings (LPV_Planner.py:233)

    def solve(**kw):
        return lo.solve_batch(P, st=st, nthreads=nthreads, **kw)
    return solve


def cpu_reference_rate(ins, cfg, track, nthreads, sample, N=HORIZON, nn=NN):
    """QP/s of the reference CPU path (restated PlannerLPV.solve + OSQP, reference settings) on `sample` QPs."""
    solve = oracle_solver(cfg, track, nthreads, exact=False, N=N, nn=nn)
    sub = {k: v[:sample] for k, v in ins.items()}
    t0 = time.perf_counter()
    o = solve(**sub)
    dt = time.perf_counter() - t0
    return sample / dt, dt, o


CPU_NOTE = ("OSQP restatement at the reference's settings (eps 1e-3, polish, max_iter 500000) with C assembly; the port "
            "caches the KKT elimination ordering across calls (the reference redoes OSQP's whole setup per call and "
            "spends ~15 ms per QP in Python assembly on top), so this baseline is stronger than the reference itself")


def run_reference(args):
    """Reference arm: rank 0 only, all host threads, bounded sample of the config-2 workload per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from colaborativempc_b200 import scenarios as sc
    from colaborativempc_b200.track import TrackMap
    cfg, track = sc.shipped_config(), TrackMap("Highway")
    cores = len(os.sched_getaffinity(0))
    sample = int(os.environ.get("DMPC_REF_SAMPLE", str(max(48, min(BATCH, 256 * cores)))))
    # same seeds as the GPU arm: the first `sample` QPs of its first timed batch, warmed up by the oracle's certified solves
    ins = steady_state_inputs(oracle_solver(cfg, track, cores, exact=True), sample, cfg, track)
    solve = oracle_solver(cfg, track, cores, exact=False)
    for _ in range(max(args.warmup, 1)):
        solve(**{k: v[:min(sample, 2 * cores)] for k, v in ins.items()})
    times = []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        o = solve(**ins)
        times.append(time.perf_counter() - t0)
    total = float(np.sum(times))
    val = sample * args.steps / total
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "QP/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "config[1]: single-agent LPV-MPC QP, N=20, nn=2, synthetic platoon scenarios, "
                                   "steady state after 3 DMPC iterations", "batch": BATCH, "sample_per_step": sample},
            "cpu_baseline": {"value": val, "unit": "QP/s", "cores": cores, "kind": "port",
                             "sample": "%d QPs per step x %d steps; %s" % (sample, args.steps, CPU_NOTE)},
            "e2e": {"value": val, "unit": "QP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "osqp_status": {str(k): int(v) for k, v in zip(*np.unique(o["status"], return_counts=True))},
            "osqp_iters_mean": float(o["iters"].mean())}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------------------
def gpu_env(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d: launch with torch.distributed.run --nproc-per-node %d" % (args.gpus, world, args.gpus))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    return torch, dist, world, rank, local, dev


def roofline_objects(qp_per_s, kern_ms, kern_qps, eng, workload_key):
    """`roofline`: the roof that binds - on-chip FP64 - from the ncu-measured FP64 instruction counts of one launch of this
    workload (profiles/r02_roofline_counts.json, scripts/gpu_roofline.sh) and the FP64-FMA peak measured on this pool's
    B200 (profiles/r02_peaks.json, scripts/ubench/peaks.cu).  `roofline_hbm`: the HBM line SURVEY 8(d) also asks for."""
    b_in, b_out = algorithmic_bytes(HORIZON, NN, duals=False)
    peaks = load_json("MEASURED_PEAKS.json")
    ub = load_json("profiles", "r02_peaks.json")
    cnt = load_json("profiles", "r02_roofline_counts.json").get(workload_key, {})
    kernel = "lpv_qp_fast_kernel" if eng.fast else "lpv_qp_stage_kernel"
    hbm_peak = float(peaks.get("hbm_gbs", 6550.0))
    hbm = (b_in + b_out) * kern_qps / (kern_ms * 1e-3) / 1e9
    roof_hbm = {"bound": "hbm", "achieved": hbm, "peak": hbm_peak, "unit": "GB/s", "frac": hbm / hbm_peak,
                "traffic": cnt.get("dram_bytes_per_launch"), "algorithmic_bytes_per_qp": b_in + b_out,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (burst)" if "hbm_gbs" in peaks else "B200_PROFILING.md fallback",
                "note": "compulsory HBM traffic is 6.5 KB/QP: three orders below the roof, it does not bind"}
    fpq = cnt.get("fp64_flops_per_qp")
    fp_peak = ub.get("fp64_fma_tflops")
    if fpq and fp_peak:
        ach = fpq * kern_qps / (kern_ms * 1e-3) / 1e12
        roof = {"bound": "fp64", "achieved": ach, "peak": float(fp_peak), "unit": "TFLOP/s", "frac": ach / float(fp_peak),
                "traffic": cnt.get("dram_bytes_per_launch"), "kernel": kernel, "kernel_ms": kern_ms,
                "fp64_flops_per_qp": fpq, "flops_source": "ncu smsp__sass_thread_inst_executed_op_{dfma x2,dadd,dmul}_pred_on of one "
                "launch of this workload (%s) / QPs in the launch" % cnt.get("source", "profiles/"),
                "peak_source": "profiles/r02_peaks.json fp64_fma_tflops (scripts/ubench/peaks.cu on this pool's B200)",
                "note": "latency-bound, not throughput-bound: one warp per QP walks N dependent Riccati stages (DESIGN.md 7)"}
    else:
        roof = dict(roof_hbm, kernel=kernel, kernel_ms=kern_ms)
    return roof, roof_hbm


def _time_calls(torch, fn, n, flush=None):
    """Mean device time (ms) of n calls of fn, CUDA events on the current stream, L2 flushed before each."""
    evs = []
    for _ in range(n):
        if flush is not None:
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        evs.append((e0, e1))
    torch.cuda.synchronize()
    return [a.elapsed_time(b) for a, b in evs]


def horizon_sweep(torch, dev, cfg, track, flush, cores, quick):
    """BASELINE config 4: N = 10 / 20 / 50 / 100, nn = 2.  Per horizon: GPU QP/s on a never-seen batch (`cold`) and on the
    same batch again (`replay`, previous working sets as first guesses), the CPU port on a bounded sample, and the parity of
    a sample against the oracle's certified optimum."""
    from colaborativempc_b200.planner import BatchedPlannerLPV
    out = []
    for N, B in ((10, 4096), (20, 4096), (50, 4096), (100, 1024)):
        if quick:
            B = min(B, 512)
        eng = BatchedPlannerLPV(*planner_args(cfg, track, N), nn=NN, max_batch=stream_capacity(B))
        ins = steady_state_inputs(lambda **kw: eng.solve_batch_host(duals=False, **kw), B, cfg, track, N=N)
        d_in = {k: torch.from_numpy(v).to(dev) for k, v in ins.items()}
        d_out = eng.alloc_outputs(B, duals=False)
        half = {k: v[:B // 2].contiguous() for k, v in d_in.items()}
        eng.solve_batch(duals=False, **half)       # another batch size: the next full-size call starts from nothing
        torch.cuda.synchronize()
        cold = _time_calls(torch, lambda: eng.solve_batch(out=d_out, duals=False, **d_in), 1, flush)[0]
        st_cold = d_out["status"].cpu().numpy()
        rep = float(np.mean(_time_calls(torch, lambda: eng.solve_batch(out=d_out, duals=False, **d_in), 5, flush)))
        st = d_out["status"].cpu().numpy()
        sample = min(B, 4 * cores if N >= 50 else 16 * cores)
        rate, secs, _ = cpu_reference_rate(ins, cfg, track, cores, sample, N=N)
        npar = 8
        o = oracle_solver(cfg, track, min(cores, npar), exact=True, N=N)(**{k: v[:npar] for k, v in ins.items()})
        z = d_out["z"][:npar].cpu().numpy()
        ok = o["status"] == 1
        err = float((np.abs(z - o["z"]).max(1) / (1 + np.abs(o["z"]).max(1)))[ok].max()) if ok.any() else None
        out.append({"N": N, "nn": NN, "qp": B, "smem_bytes_per_qp": eng.stage_smem_bytes(),
                    "gpu_cold_qp_per_s": B / (cold * 1e-3), "gpu_cold_ms": cold, "gpu_replay_qp_per_s": B / (rep * 1e-3),
                    "gpu_replay_ms": rep, "status_not_1": int((st != 1).sum() + (st_cold != 1).sum()),
                    "cpu_port_qp_per_s": rate, "cpu_cores": cores, "cpu_sample": sample, "cpu_seconds": secs,
                    "gpu_cold_over_cpu": B / (cold * 1e-3) / rate,
                    "parity_sample": {"qps": int(ok.sum()), "oracle_certified": int(ok.sum()), "max_rel_err_z": err}})
        eng.close()
    return out


def run_single(args):
    """N = 1: BASELINE config[1]."""
    torch, dist, world, rank, local, dev = gpu_env(args)
    from colaborativempc_b200 import scenarios as sc
    from colaborativempc_b200.planner import BatchedPlannerLPV
    from colaborativempc_b200.track import TrackMap
    cfg, track = sc.shipped_config(), TrackMap("Highway")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    W = max(args.warmup, 3)
    nb = W + args.steps
    opts = {"engine": args.engine, "warm_start": 2 if args.warm else 0}
    mk = lambda o: BatchedPlannerLPV(*planner_args(cfg, track, HORIZON), nn=NN, max_batch=stream_capacity(BATCH), options=o)
    gen = mk({"engine": args.engine})
    stream = closed_loop_stream(lambda **kw: gen.solve_batch_host(duals=False, **kw), BATCH, nb, cfg, track)
    gen.close()
    eng, eng_h = mk(opts), mk(opts)       # device-resident arm / host-buffer arm: each sees the same stream of batches
    d_in = [{k: torch.from_numpy(v).to(dev) for k, v in b.items()} for b in stream]
    d_out = eng.alloc_outputs(BATCH, duals=False)
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
    h_in = [{k: pin(v) for k, v in b.items()} for b in stream]
    h_out = dict(z=pin(np.empty((BATCH, eng.n))), planes=pin(np.empty((BATCH, HORIZON, 3, NN))), obj=pin(np.empty(BATCH)),
                 status=pin(np.empty(BATCH, np.int32)), iters=pin(np.empty(BATCH, np.int32)), pivots=pin(np.empty(BATCH, np.int32)))
    h2d = int(sum(v.nbytes for v in h_in[0].values()))
    d2h = int(sum(v.nbytes for v in h_out.values()))
    n_bad = torch.zeros((), dtype=torch.int64, device=dev)
    n_one = torch.zeros((), dtype=torch.int64, device=dev)
    it_sum = torch.zeros((), dtype=torch.float64, device=dev)
    rd_sum = torch.zeros((), dtype=torch.float64, device=dev)
    # hard-QP hint (the one-bit "ended in the interior point last step" carried with the launch order): how well does the
    # previous, different batch predict this one?  Counted on the device, read after the timed loop.
    prev_hard = torch.zeros((BATCH,), dtype=torch.bool, device=dev)
    hint_cnt = torch.zeros((4,), dtype=torch.int64, device=dev)  # hard now, hard now & hinted, hinted, hinted & not hard now
    sampler = ClockSampler(local)
    pairs, e2e_s, bad_h = [], [], 0
    for i in range(nb):
        timed = i >= W
        if i == W:
            torch.cuda.synchronize()
            launches0 = eng.launches
            sampler.start()
            t_wall0 = time.perf_counter()
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        eng.solve_batch(out=d_out, duals=False, **d_in[i])
        e1.record()
        if timed:
            pairs.append((e0, e1))
            n_bad += (~((d_out["status"] == 1) | (d_out["status"] == 2))).sum()
            n_one += (d_out["status"] == 1).sum()
            it_sum += d_out["iters"].double().sum()
            rd_sum += d_out["pivots"].double().sum()
            hard = d_out["iters"] > 0
            hint_cnt += torch.stack([hard.sum(), (hard & prev_hard).sum(), prev_hard.sum(), (prev_hard & ~hard).sum()])
        prev_hard = d_out["iters"] > 0
        # the same batch end to end through the host-buffer C-ABI call, on the same iteration
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        eng_h.solve_batch_host(out=h_out, duals=False, **h_in[i])
        dt = time.perf_counter() - t0
        if timed:
            e2e_s.append(dt)
            bad_h += int((~np.isin(h_out["status"], (1, 2))).sum())
    torch.cuda.synchronize()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.result()
    launches = eng.launches - launches0
    per_step = np.array([a.elapsed_time(b) for a, b in pairs])
    total_ms = float(per_step.sum())
    value = BATCH * args.steps / (total_ms * 1e-3)
    qp_timed = BATCH * args.steps
    solved, infeasible = int(n_one.item()), int(n_bad.item()) + bad_h
    e2e = {"value": BATCH * args.steps / float(np.sum(e2e_s)), "unit": "QP/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
           "ms_per_step": 1e3 * float(np.mean(e2e_s)),
           "path": "dmpc_solve_batch_host on page-locked buffers, read/written in place by the kernel over PCIe (host_direct=1), "
                   "same stream of batches, measured on the same iterations as `value` (interleaved, own handle)"}
    # ---- the same frozen batch again (previous working sets are exact) and a never-seen batch (nothing carried)
    last = d_in[-1]
    rep_eng = mk({"engine": args.engine, "warm_start": 1})   # same batch again: the working sets as certified, not shifted
    rep_eng.solve_batch(out=d_out, duals=False, **last)
    replay = float(np.mean(_time_calls(torch, lambda: rep_eng.solve_batch(out=d_out, duals=False, **last), 6, flush)[1:]))
    cold_eng = mk({"engine": args.engine, "warm_start": 0, "order_hint": 0})
    cold_eng.solve_batch(out=d_out, duals=False, **last)
    cold = float(np.mean(_time_calls(torch, lambda: cold_eng.solve_batch(out=d_out, duals=False, **last), 5, flush)))
    eng_h.set_options(host_direct=0)
    t0 = time.perf_counter()
    for _ in range(3):
        eng_h.solve_batch_host(out=h_out, duals=False, **h_in[-1])
    e2e["staged_replay_value"] = BATCH * 3 / (time.perf_counter() - t0)
    roofline, roof_hbm = roofline_objects(value, total_ms / args.steps, BATCH, eng, "stream")
    # ---- CPU baseline: the reference path restated, on the box's host cores, bounded samples (all cores and one core)
    cpu, sweep = None, None
    if not args.no_cpu_baseline:
        cores = len(os.sched_getaffinity(0))
        sample = max(48, min(BATCH, 128 * cores))
        rate, secs, o = cpu_reference_rate(stream[W], cfg, track, cores, sample)
        s1 = max(16, min(sample, 96))
        rate1, secs1, _ = cpu_reference_rate(stream[W], cfg, track, 1, s1)
        cpu = {"value": rate, "unit": "QP/s", "cores": cores, "kind": "port",
               "sample": "first %d QPs of the first timed batch, %.1f s wall; %s" % (sample, secs, CPU_NOTE),
               "osqp_iters_mean": float(o["iters"].mean()),
               "one_core": {"value": rate1, "unit": "QP/s", "cores": 1, "sample": "first %d QPs, %.1f s wall" % (s1, secs1)}}
        if not args.no_sweep:
            sweep = horizon_sweep(torch, dev, cfg, track, flush, cores, args.quick)
    line = {"metric": METRIC, "value": value, "unit": "QP/s", "n_gpus": 1, "steps": args.steps, "warmup": W,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": "config[1]: 4096 independent single-agent LPV-MPC QPs, N=20, nn=2, synthetic platoon scenarios (seed = "
                                   "scenario index); step i solves the QPs of closed-loop DMPC iteration 3+i (a stream of consecutive, "
                                   "different batches; inputs resident in HBM)",
                       "horizon": HORIZON, "nn": NN, "qp_per_step": BATCH, "parallelism": "1 GPU, one warp (one-warp CTA) per QP",
                       "l2": "flushed between timed steps (256 MiB memset)", "engine": "stage-wise" if not eng.fast else "condensed",
                       "outputs": "Solution z, planes, obj, status (duals off; parity tests run with duals on)",
                       "carried_between_steps": "launch order and first working-set guess from the previous (different) batch "
                                                "(dmpc_options.order_hint, warm_start=2); `cold` = nothing carried"},
            "replay": {"ms_per_step": replay, "value": BATCH / (replay * 1e-3), "unit": "QP/s", "what": "the last batch again: previous working sets are exact"},
            "cold": {"ms_per_step": cold, "value": BATCH / (cold * 1e-3), "unit": "QP/s", "what": "warm_start=0, order_hint=0: every QP from the empty working set, QP index = CTA index"},
            "roofline": roofline, "roofline_hbm": roof_hbm, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
            "p50_step_ms": float(np.median(per_step)), "wall_ms_per_step": 1e3 * t_wall / args.steps,
            "solved": solved, "qp_timed": qp_timed, "solved_inaccurate": qp_timed - solved - int(n_bad.item()), "infeasible": infeasible,
            "ipm_iters_mean": float(it_sum.item()) / qp_timed, "polish_rounds_mean": float(rd_sum.item()) / qp_timed, "sweep": sweep}
    hc = [int(v) for v in hint_cnt.tolist()]
    line["hint"] = {"interior_point_qps": hc[0], "of_them_hinted": hc[1], "hit_rate": hc[1] / max(hc[0], 1), "hinted_qps": hc[2],
                    "hinted_but_certified_by_rounds": hc[3], "share_of_batch": hc[0] / qp_timed,
                    "what": "live stream: a QP is hinted when the QP of the same index in the PREVIOUS (different) batch ended in the "
                            "interior point; hit_rate = interior-point QPs of a step that were hinted (started first, one round only)"}
    print(json.dumps(line))
    if infeasible:
        sys.stderr.write("FAIL: %d QPs of the timed steps did not come back feasible\n" % infeasible)
        sys.exit(3)


FAIL_STATUS = (-10, -20, -21)   # numerical breakdown / failed map lookup / degenerate plane: never acceptable in a bench run


def _dmpc_loop(torch, dm, steps, flush, e2e_io=True):
    """`steps` closed-loop iterations; per step the device time of the iteration (CUDA events around gather -> solve ->
    exchange) and the wall time including the host traffic of a closed loop (measured state up, first move + status down).
    Only the QPs of RUNNING scenarios count: a scenario whose experiment has stopped (a QP that was not feasible) is
    skipped by the kernel and by the throughput.  Returns (device ms per step, wall s per step, counts) with counts =
    [QPs solved in running scenarios, status 1, status 2, not feasible (each failure once), of which breakdowns,
     interior-point iterations, active-set rounds]."""
    dev = dm.eng.device
    h_x0 = torch.empty((dm.B, 9), dtype=torch.float64).pin_memory()
    h_x0.copy_(dm.x0)
    h_u0 = torch.empty((dm.B, 2), dtype=torch.float64).pin_memory()
    h_st = torch.empty((dm.B,), dtype=torch.int32).pin_memory()
    acc = torch.zeros((7,), dtype=torch.float64, device=dev)
    pairs, walls = [], []
    for _ in range(steps):
        if flush is not None:
            flush.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        if e2e_io:
            dm.x0.copy_(h_x0, non_blocking=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        dm.iterate()
        e1.record()
        if e2e_io:
            h_u0.copy_(dm.uold, non_blocking=True)
            h_st.copy_(dm.out["status"], non_blocking=True)
        torch.cuda.synchronize()
        walls.append(time.perf_counter() - t0)
        pairs.append((e0, e1))
        st = dm.out["status"]
        run = (dm.stop == 0).repeat_interleave(dm.a_local)   # as the solve of this iteration saw it (set by its gather)
        feas = (st == 1) | (st == 2) | (st == -2)
        brk = torch.zeros_like(run)
        for v in FAIL_STATUS:
            brk |= st == v
        zero = torch.zeros((), dtype=torch.float64, device=dev)
        acc += torch.stack([run.sum().double(), (run & (st == 1)).sum().double(), (run & (st == 2)).sum().double(),
                            (run & ~feas).sum().double(), (run & brk).sum().double(),
                            torch.where(run, dm.out["iters"].double(), zero).sum(), torch.where(run, dm.out["pivots"].double(), zero).sum()])
        if e2e_io:
            h_x0.copy_(dm.x0)  # the plant is the planner's own prediction (LPV_HP_N_main.py:106)
    return np.array([p.elapsed_time(q) for p, q in pairs]), np.array(walls), acc.cpu().numpy()


def run_dmpc(args):
    """N >= 1: the sharded closed loop (config 5 shape: 256 scenarios per GPU x 16-agent chain platoons)."""
    torch, dist, world, rank, local, dev = gpu_env(args)
    from colaborativempc_b200 import scenarios as sc
    from colaborativempc_b200.dmpc import ShardedDMPC
    from colaborativempc_b200.planner import BatchedPlannerLPV
    from colaborativempc_b200.track import TrackMap
    cfg, track = sc.shipped_config(), TrackMap("Highway")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    W = max(args.warmup, 3)
    opts = {"engine": args.engine}

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    made = []   # every ShardedDMPC of this run: their graphs are released before the process group is torn down

    def make(S, A, w, r, topo="chain", nn=NN, seed0=0, group=None):
        eng = BatchedPlannerLPV(*planner_args(cfg, track, HORIZON), nn=nn, max_batch=S * A // w, options=opts)
        x0, xprev, uprev = sc.platoon_scenarios(S, A, HORIZON, cfg["dt"], track, seed0=seed0)
        idx, cnt = sc.topology(A, topo, nn=nn)
        dm = ShardedDMPC(eng, S, A, idx, cnt, x0, xprev, uprev, rank=r, world=w, end_s=1e9, group=group)  # synthetic run: no end of track
        if not args.warm:
            eng.set_options(warm_start=0)
        made.append(dm)
        return eng, dm

    def settle(dm, n):
        for _ in range(n):
            dm.iterate()
        if not args.no_graph:
            dm.capture()

    def pipelined(S_total, w, r, group_b, sync):
        """The same scenarios as TWO independent groups, each with its own state, handle, CUDA graph, stream and (w > 1)
        communicator, interleaved: while one group waits for its all-gather - i.e. for the slowest rank's hardest QP of
        that group - the other group's solve has the GPU, and the tail of one solve kernel is filled by the bulk of the
        other.  Every scenario still does one solve and one exchange per iteration; the iterations of a group stay in
        order.  Times `steps` iterations of ALL scenarios between two synchronisations (no L2 flush in between)."""
        half = S_total // 2
        sa, sb = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
        grp = []
        for st_, seed, g in ((sa, 0, None), (sb, half, group_b)):
            with torch.cuda.stream(st_):
                e_, d_ = make(half, A, w, r, seed0=seed, group=g)
                settle(d_, args.settle)
            grp.append((st_, e_, d_, torch.zeros((5,), dtype=torch.float64, device=dev)))
        torch.cuda.synchronize()

        def run(n, count):
            main = torch.cuda.current_stream(dev)
            sync()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(main)
            for st_, _, _, _ in grp:
                st_.wait_stream(main)
            for _ in range(n):
                for st_, _, d_, acc_ in grp:
                    with torch.cuda.stream(st_):
                        d_.iterate()
                        if count:
                            stt = d_.out["status"]
                            runm = (d_.stop == 0).repeat_interleave(d_.a_local)
                            feas = (stt == 1) | (stt == 2) | (stt == -2)
                            brk = torch.zeros_like(runm)
                            for v in FAIL_STATUS:
                                brk |= stt == v
                            acc_ += torch.stack([runm.sum().double(), (runm & (stt == 1)).sum().double(), (runm & (stt == 2)).sum().double(),
                                                 (runm & ~feas).sum().double(), (runm & brk).sum().double()])
            for st_, _, _, _ in grp:
                main.wait_stream(st_)
            e1.record(main)
            torch.cuda.synchronize()
            sync()
            return e0.elapsed_time(e1)

        run(W, False)
        ms_tot = run(args.steps, True)
        cnt = sum(g[3] for g in grp).cpu().numpy()
        for _, e_, d_, _ in grp:
            d_.release()
        return ms_tot, cnt

    A = 16
    # ---- the same workload on ONE GPU (rank 0 alone, no collective), in the same run: the weak-scaling baseline
    weak = None
    if world > 1:
        if rank == 0:
            e1, d1 = make(256, A, 1, 0)
            settle(d1, args.settle)
            _dmpc_loop(torch, d1, W, flush)
            ms1, w1, c1 = _dmpc_loop(torch, d1, args.steps, flush)
            weak = {"value": c1[0] / (ms1.sum() * 1e-3), "unit": "QP/s", "ms_per_step": float(ms1.mean()), "n_gpus": 1,
                    "qp_timed": int(c1[0]), "not_feasible": int(c1[3]), "breakdowns": int(c1[4]), "e2e_value": c1[0] / float(w1.sum()),
                    "what": "256 scenarios x 16-agent chain platoons on rank 0 alone, same settle / warm-up / steps"}
            e1.close()
            if args.pipelined:
                msp1, cp1 = pipelined(256, 1, 0, None, torch.cuda.synchronize)
                weak["pipelined_value"] = float(cp1[0]) / (msp1 * 1e-3)
        barrier()
    S = 256 * world
    eng, dm = make(S, A, world, rank)
    settle(dm, args.settle)
    _dmpc_loop(torch, dm, W, flush)
    barrier()
    launches0 = eng.launches
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    t_wall0 = time.perf_counter()
    ms, walls, cn = _dmpc_loop(torch, dm, args.steps, flush)
    barrier()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.result()
    launches = (eng.launches - launches0) if args.no_graph else args.steps * dm.launches_per_iteration
    stopped = int(dm.stopped().sum())
    # ---- where a step goes: 8 more closed-loop iterations launched eagerly with CUDA events between the three phases
    # (gather + stop reduce | solve + bookkeeping + launch order | exchange), and the collective alone on the real buffer
    ph = np.array([dm.iterate_phases() for _ in range(8)])[2:].mean(0)
    k_ms = float(ph[1])
    coll = None
    if world > 1:
        coll = float(np.mean(_time_calls(torch, dm.exchange, 50)))
    pipe = None
    if args.pipelined:
        gb = dist.new_group(ranks=list(range(world))) if world > 1 else None
        msp, cp = pipelined(S, world, rank, gb, barrier)
        tp = torch.tensor([msp], dtype=torch.float64, device=dev)
        cpt = torch.tensor(cp, dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tp, op=dist.ReduceOp.MAX)
            dist.all_reduce(cpt)
        cpn = cpt.cpu().numpy()
        pipe = {"value": float(cpn[0]) / (float(tp.item()) * 1e-3), "unit": "QP/s", "ms_per_iteration": float(tp.item()) / args.steps,
                "qp_timed": int(cpn[0]), "solved": int(cpn[1]), "solved_inaccurate": int(cpn[2]), "not_feasible": int(cpn[3]),
                "breakdowns": int(cpn[4]),
                "what": "the same scenarios as two independent groups (own state, handle, CUDA graph, stream, communicator) "
                        "interleaved on two streams; %d iterations of every scenario between two synchronisations, max over "
                        "ranks, no L2 flush in between" % args.steps}
    # ---- reduce over ranks: a step is as slow as its slowest rank
    t = torch.tensor(np.concatenate([ms, walls * 1e3]), dtype=torch.float64, device=dev)
    cnts = torch.tensor(list(cn) + [ph[0], ph[1], ph[2]], dtype=torch.float64, device=dev)
    if world > 1:
        allt = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allt, t)
        allc = [torch.zeros_like(cnts) for _ in range(world)]
        dist.all_gather(allc, cnts)
        allt, allc = torch.stack(allt).cpu().numpy(), torch.stack(allc).cpu().numpy()
    else:
        allt, allc = t.cpu().numpy()[None], cnts.cpu().numpy()[None]
    dev_ms, wall_ms = allt[:, :args.steps], allt[:, args.steps:]
    step_ms = dev_ms.max(0)
    qp_timed = float(allc[:, 0].sum())
    value = qp_timed / (step_ms.sum() * 1e-3)
    e2e_val = qp_timed / (wall_ms.max(0).sum() * 1e-3)
    not_feasible, breakdowns = int(allc[:, 3].sum()), int(allc[:, 4].sum())
    config3 = None
    if not args.no_config3 and 64 % world == 0:
        config3 = {}
        for name, S3, topo, nn3 in (("chain_1_platoon", 1, "chain", 2), ("chain_64_platoons", 64, "chain", 2), ("full_nn63_8_platoons", 8, "full", 63)):
            for seed in range(7, 19):   # a single platoon can be one the reference would have stopped: take the first that
                e3, d3 = make(S3, 64, world, rank, topo=topo, nn=nn3, seed0=seed)   # runs through the timed window
                settle(d3, 10)
                _dmpc_loop(torch, d3, 3, None, e2e_io=False)
                barrier()
                m3, _, c3 = _dmpc_loop(torch, d3, 20, None, e2e_io=False)
                if S3 > 1 or not d3.stopped().any():
                    break
                e3.close()
            tt = torch.tensor(m3, dtype=torch.float64, device=dev)
            bb = torch.tensor(c3, dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                dist.all_reduce(bb)
            m3, c3 = tt.cpu().numpy(), bb.cpu().numpy()
            config3[name] = {"scenarios": S3, "agents": 64, "agents_per_gpu": 64 // world, "nn": nn3, "qp_per_iter": S3 * 64,
                             "ms_per_iter_p50": float(np.median(m3)), "qp_per_s": float(c3[0]) / (m3.sum() * 1e-3),
                             "qp_timed": int(c3[0]), "solved": int(c3[1]), "solved_inaccurate": int(c3[2]),
                             "not_feasible": int(c3[3]), "breakdowns": int(c3[4]), "stopped_scenarios": int(d3.stopped().sum())}
            breakdowns += int(c3[4])
            e3.close()
    if rank == 0:
        roofline, roof_hbm = roofline_objects(value, k_ms, dm.B, eng, "dmpc")
        line = {"metric": METRIC, "value": value, "unit": "QP/s", "n_gpus": world, "steps": args.steps, "warmup": W,
                "ms_per_step": float(step_ms.mean()), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": "DMPC iteration: %d scenarios x 16-agent chain platoon (nn<=2), N=20, agents sharded %d per GPU (4096 QPs "
                                       "per GPU), one all-gather of [(x,y) predictions | stop flags] per iteration, closed loop, timed after %d "
                                       "settling iterations + warm-up; only the QPs of running scenarios count (a scenario whose experiment "
                                       "the reference would have stopped is frozen)" % (S, A // world, args.settle),
                           "horizon": HORIZON, "nn": NN, "qp_per_step": dm.B * world,
                           "parallelism": "agents sharded over %d GPUs, one process per GPU, NCCL all-gather per iteration" % world,
                           "l2": "flushed between timed steps (256 MiB memset)", "cuda_graph": not args.no_graph,
                           "iteration": "gather+stop-reduce kernel -> solve kernel (epilogue: shift, pack, hold, flags) -> launch-order kernel -> all-gather"},
                "weak_baseline": weak,
                "efficiency_same_workload": (value / (world * weak["value"])) if weak else None,
                "step_breakdown_ms": {"step_max_over_ranks": float(step_ms.mean()), "rank_step_min": float(dev_ms.mean(1).min()),
                                      "rank_step_median": float(np.median(dev_ms.mean(1))), "rank_step_max": float(dev_ms.mean(1).max()),
                                      "eager_phase_gather_per_rank": [float(v) for v in allc[:, 7]],
                                      "eager_phase_solve_per_rank": [float(v) for v in allc[:, 8]],
                                      "eager_phase_exchange_per_rank": [float(v) for v in allc[:, 9]], "all_gather_alone": coll,
                                      "all_gather_bytes_per_rank": int(dm.send.numel() * 8),
                                      "wall_minus_device": float(wall_ms.max(0).mean() - step_ms.mean())},
                "roofline": roofline, "roofline_hbm": roof_hbm, "cpu_baseline": None,
                "e2e": {"value": e2e_val, "unit": "QP/s", "h2d_bytes_per_step": int(dm.B * 72 * world), "d2h_bytes_per_step": int(dm.B * 20 * world),
                        "ms_per_step": float(wall_ms.max(0).mean()),
                        "path": "closed loop on the SAME iterations as `value`: measured state x0 up from pinned memory, iteration, first "
                                "move + status down, synchronise; wall clock, max over ranks"},
                "gpu_launches": int(launches), "clocks": clocks, "p50_step_ms": float(np.median(step_ms)),
                "wall_ms_per_step": 1e3 * t_wall / args.steps,
                "qp_timed": int(qp_timed), "solved": int(allc[:, 1].sum()), "solved_inaccurate": int(allc[:, 2].sum()),
                "not_feasible": not_feasible, "breakdowns": breakdowns, "stopped_scenarios_rank0": stopped,
                "solved_per_rank": [int(v) for v in allc[:, 1]], "qp_timed_per_rank": [int(v) for v in allc[:, 0]],
                "ipm_iters_mean": float(allc[:, 5].sum()) / max(qp_timed, 1), "polish_rounds_mean": float(allc[:, 6].sum()) / max(qp_timed, 1),
                "config3": config3}
        if pipe is not None:
            if weak and weak.get("pipelined_value"):
                pipe["weak_baseline_pipelined"] = weak["pipelined_value"]
                pipe["efficiency_same_workload"] = pipe["value"] / (world * weak["pipelined_value"])
            line["pipelined"] = pipe
        print(json.dumps(line))
    rc = 3 if (breakdowns or not_feasible > 0.01 * qp_timed) else 0
    if rc:
        sys.stderr.write("FAIL: %d breakdowns, %d of %d timed QPs not feasible\n" % (breakdowns, not_feasible, int(qp_timed)))
    if world > 1:
        # captured graphs hold NCCL work: drop them first, then tear the group down with a hard limit (the result line is
        # out; a teardown that does not come back must not keep the job alive)
        for m in made:
            m.release()
        import gc
        gc.collect()
        barrier()
        sys.stdout.flush()
        sys.stderr.flush()
        t = threading.Thread(target=dist.destroy_process_group, daemon=True)
        t.start()
        t.join(20.0)
        if t.is_alive():
            os._exit(rc)
    sys.exit(rc)   # a genuinely infeasible QP stops its scenario (reported above); a numerical breakdown, or more than
                   # 1 % of the timed QPs not feasible, fails the run


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sweep", action="store_true", help="N=1: skip the horizon sweep (config 4)")
    ap.add_argument("--no-config3", action="store_true", help="DMPC workload: skip the 64-vehicle platoon lines (config 3)")
    ap.add_argument("--pipelined", action="store_true", help="DMPC workload: also time the scenarios as TWO interleaved groups on two streams (extra key `pipelined`)")
    ap.add_argument("--no-graph", action="store_true", help="DMPC workload: launch the iteration eagerly instead of as a CUDA graph")
    ap.add_argument("--quick", action="store_true", help="smaller sweep batches (development)")
    ap.add_argument("--warm", type=int, default=1, help="0: dmpc_options.warm_start off")
    ap.add_argument("--engine", type=int, default=0, help="dmpc_options.engine: 0 auto (stage-wise), 2 condensed, 3 stage-wise")
    ap.add_argument("--settle", type=int, default=30, help="DMPC workload: closed-loop iterations before warm-up")
    ap.add_argument("--workload", default="batch", choices=["batch", "dmpc"],
                    help="N=1 only: 'dmpc' runs the multi-GPU step (16-agent platoons, closed loop) on one GPU")
    args = ap.parse_args()
    if os.environ.get("DMPC_BENCH_WATCHDOG"):   # development: dump every thread's stack if the run is still going after N s
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ["DMPC_BENCH_WATCHDOG"]), exit=True)
    if args.impl == "reference":
        run_reference(args)
    elif int(os.environ.get("WORLD_SIZE", "1")) == 1 and args.workload == "batch":
        run_single(args)
    else:
        run_dmpc(args)


if __name__ == "__main__":
    main()

"""bench.py - batched DMPC QP solves/sec at N=20 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

N=1 workload  = BASELINE config[1]: 4096 independent single-agent LPV-MPC QPs (horizon 20, 2 neighbour slots), synthetic
                platoon scenarios (seed = scenario index), brought to steady state by 3 DMPC iterations, then
                frozen.  A step = one pass of the hot path (planes -> linearisation -> condensation -> QP solve -> unpack,
                one kernel launch) over the 4096 QPs.
N>1 workload  = the DMPC iteration with its exchange: 256*N scenarios x 16-agent chain platoons, agents sharded over
                the N ranks (4096 QPs per GPU per iteration, weak scaling); a step = gather -> solve -> shift/pack ->
                NCCL all-gather of the packed (x,y) predictions.
--impl reference = the reference's CPU path for the same QPs (oracle restatement of PlannerLPV.solve with OSQP at the
                reference's settings) on all host threads, bounded sample per step.
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


def steady_state_inputs(solve_fn, n_qp, cfg, track, A=3, topo="full", warm_iters=3, seed0=0):
    """Config-2 inputs: ceil(n_qp/A) synthetic platoons, `warm_iters` DMPC iterations with `solve_fn`, then frozen."""
    from colaborativempc_b200 import scenarios as sc
    S = (n_qp + A - 1) // A
    x0, xprev, uprev = sc.platoon_scenarios(S, A, HORIZON, cfg["dt"], track, seed0=seed0)
    idx, cnt = sc.topology(A, topo, nn=NN)
    st = sc.run_dmpc(solve_fn, x0, xprev, uprev, idx, cnt, warm_iters, HORIZON)
    return {k: np.ascontiguousarray(st[k][:n_qp]) for k in INPUT_KEYS}


# ------------------------------------------------------------------------------------------------------------
# CPU legs (the only places bench.py executes oracle/): cpu_baseline of the GPU arm and the --impl reference arm.
def oracle_solver(cfg, track, nthreads, exact):
    from colaborativempc_b200._abi import make_params
    from oracle import lpv_oracle as lo
    P = make_params(*planner_args(cfg, track, HORIZON), nn=NN)
    st = lo.settings(exact=exact)  # exact=False: OSQP at the reference's settings (LPV_Planner.py:233)

    def solve(**kw):
        return lo.solve_batch(P, st=st, nthreads=nthreads, **kw)
    return solve


def cpu_reference_rate(ins, cfg, track, nthreads, sample, repeats=1):
    """QP/s of the reference CPU path (restated PlannerLPV.solve + OSQP, reference settings) on `sample` QPs."""
    solve = oracle_solver(cfg, track, nthreads, exact=False)
    sub = {k: v[:sample] for k, v in ins.items()}
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        o = solve(**sub)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return sample / best, best, o


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
    # same seeds as the GPU arm: the first `sample` QPs of its batch, warmed up by the oracle's certified solves
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
                             "sample": "%d QPs per step x %d steps, OSQP restatement at reference settings "
                                       "(eps 1e-3, polish, fresh setup per call), C assembly" % (sample, args.steps)},
            "e2e": {"value": val, "unit": "QP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "osqp_status": {str(k): int(v) for k, v in zip(*np.unique(o["status"], return_counts=True))},
            "osqp_iters_mean": float(o["iters"].mean())}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------------------
def _flush_l2(torch, buf):
    buf.zero_()  # 256 MiB > 126 MB L2: evicts the previous step's inputs/outputs


def run_gpu(args):
    import torch
    import torch.distributed as dist
    from colaborativempc_b200 import scenarios as sc
    from colaborativempc_b200.dmpc import ShardedDMPC
    from colaborativempc_b200.planner import BatchedPlannerLPV
    from colaborativempc_b200.track import TrackMap

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d: launch with torch.distributed.run --nproc-per-node %d" % (args.gpus, world, args.gpus))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cfg, track = sc.shipped_config(), TrackMap("Highway")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ev = lambda: torch.cuda.Event(enable_timing=True)
    sampler = ClockSampler(local)
    extra = {}

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    single = world == 1 and args.workload == "batch"  # --workload dmpc: the N>1 step (sharded closed loop) on one GPU
    if single:
        A = 3
        S = (BATCH + A - 1) // A
        eng = BatchedPlannerLPV(*planner_args(cfg, track, HORIZON), nn=NN, max_batch=S * A, options={"engine": args.engine, "warm_start": args.warm})
        ins = steady_state_inputs(lambda **kw: eng.solve_batch_host(duals=False, **kw), BATCH, cfg, track, A=A)
        d_in = {k: torch.from_numpy(v).to(dev) for k, v in ins.items()}
        d_out = eng.alloc_outputs(BATCH, duals=False)
        step = lambda: eng.solve_batch(out=d_out, duals=False, **d_in)
        qp_per_step = BATCH
        workload = ("config[1]: 4096 independent single-agent LPV-MPC QPs, N=20, nn=2, synthetic platoon scenarios "
                    "(seed = scenario index), steady state after 3 DMPC iterations, frozen")
        par = "1 GPU, one CTA per QP"
    else:
        A, S = 16, 256 * world
        eng = BatchedPlannerLPV(*planner_args(cfg, track, HORIZON), nn=NN, max_batch=S * A // world, options={"engine": args.engine, "warm_start": args.warm})
        x0, xprev, uprev = sc.platoon_scenarios(S, A, HORIZON, cfg["dt"], track, seed0=0)
        idx, cnt = sc.topology(A, "chain", nn=NN)
        dm = ShardedDMPC(eng, S, A, idx, cnt, x0, xprev, uprev, rank=rank, world=world)
        step = dm.iterate
        # cold-start platoons first close their gaps (more saturated QPs per iteration than in cruise): let the closed loop
        # settle before the warm-up, as config[1]'s batch is taken "at steady state"; --settle 0 times the transient
        for _ in range(args.settle):
            step()
        qp_per_step = dm.B * world
        d_out = dm.out
        workload = ("DMPC iteration: %d scenarios x 16-agent chain platoon (nn<=2), N=20, agents sharded %d per GPU "
                    "(4096 QPs per GPU), one NCCL all-gather of (x,y) predictions per iteration, closed loop, timed after %d settling "
                    "iterations + warm-up" % (S, A // world, args.settle))
        par = "agents sharded over %d GPUs, all-gather per iteration" % world

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    launches0 = eng.launches
    sampler.start()
    pairs = []
    barrier()
    t_wall0 = time.perf_counter()
    for _ in range(args.steps):
        _flush_l2(torch, flush)
        e0, e1 = ev(), ev()
        e0.record()
        step()
        e1.record()
        pairs.append((e0, e1))
    barrier()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.result()
    launches = eng.launches - launches0
    per_step = torch.tensor([a.elapsed_time(b) for a, b in pairs], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(per_step, op=dist.ReduceOp.MAX)
    per_step = per_step.cpu().numpy()
    total_ms = float(per_step.sum())
    value = qp_per_step * args.steps / (total_ms * 1e-3)
    status = d_out["status"].cpu().numpy()
    n_ok = int((status == 1).sum())
    iters_mean = float(d_out["iters"].double().mean().item())
    rounds_mean = float(d_out["pivots"].double().mean().item())

    # ---- the same steps with the cost-ordered launch off (blockIdx = QP index): what a cold, never-seen batch costs
    unordered = None
    if single and eng.fast:
        eng.set_options(order_hint=0)
        step()
        torch.cuda.synchronize()
        up = []
        for _ in range(args.steps):
            _flush_l2(torch, flush)
            e0, e1 = ev(), ev()
            e0.record()
            step()
            e1.record()
            up.append((e0, e1))
        torch.cuda.synchronize()
        ums = float(np.mean([a.elapsed_time(b) for a, b in up]))
        unordered = {"ms_per_step": ums, "value": BATCH / (ums * 1e-3), "unit": "QP/s"}
        eng.set_options(order_hint=1)
        step()
        torch.cuda.synchronize()

    # ---- dominant kernel (lpv_qp_solve_kernel) timed alone on its launch stream, for the roofline object
    if single:
        kern_ms = total_ms / args.steps
        kern_qps = BATCH
    else:
        ks = []
        xp = dm.xprev
        # inputs of the NEXT iteration, consistent with each other: neighbour rows and own pose gathered from the last exchange
        eng.gather_neighbours(dm.xy_all, dm.nbr_idx, dm.nbr_xy, dm.pose, dm.S, dm.A, dm.a_local, dm.off)
        for _ in range(2):  # the launch order follows the previous call's cost: let it settle on this frozen batch
            eng.solve_batch(dm.x0, xp, dm.uprev, dm.uold, dm.nbr_xy, dm.pose, dm.nbr_cnt, out=dm.out, duals=False)
        for _ in range(5):
            _flush_l2(torch, flush)
            e0, e1 = ev(), ev()
            e0.record()
            eng.solve_batch(dm.x0, xp, dm.uprev, dm.uold, dm.nbr_xy, dm.pose, dm.nbr_cnt, out=dm.out, duals=False)
            e1.record()
            ks.append((e0, e1))
        torch.cuda.synchronize()
        kern_ms = float(np.mean([a.elapsed_time(b) for a, b in ks]))
        kern_qps = dm.B
    b_in, b_out = algorithmic_bytes(HORIZON, NN, duals=False)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6550.0))  # fallback: B200_PROFILING.md measured copy bandwidth
    achieved = (b_in + b_out) * kern_qps / (kern_ms * 1e-3) / 1e9
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json"))).get("dram_bytes_per_launch")
    except Exception:
        pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "kernel": "lpv_qp_fast_kernel" if eng.fast else "lpv_qp_stage_kernel", "kernel_ms": kern_ms,
                "algorithmic_bytes_per_qp": b_in + b_out,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (burst)" if "hbm_gbs" in peaks else "B200_PROFILING.md fallback",
                "note": "compulsory HBM traffic is 6.6 KB/QP; the kernel is bound by on-chip FP64 work (DESIGN.md), "
                        "so the HBM fraction is small by construction"}

    # ---- end to end through the host-buffer C-ABI call: pinned host inputs -> H2D -> solve -> D2H, every step
    if single:
        pin = lambda a: torch.from_numpy(a).pin_memory().numpy()
        h_in = {k: pin(v) for k, v in ins.items()}
        h_out = dict(z=pin(np.empty((BATCH, eng.n))), planes=pin(np.empty((BATCH, HORIZON, 3, NN))), obj=pin(np.empty(BATCH)),
                     status=pin(np.empty(BATCH, np.int32)), iters=pin(np.empty(BATCH, np.int32)), pivots=pin(np.empty(BATCH, np.int32)))
        h2d = int(sum(v.nbytes for v in h_in.values()))
        d2h = int(sum(v.nbytes for v in h_out.values()))
        e2e_step = lambda: eng.solve_batch_host(out=h_out, duals=False, **h_in)
    else:
        # closed loop: the measured states go up, the first control move and the status come back, every iteration
        h_x0 = torch.empty((dm.B, 9), dtype=torch.float64).pin_memory()
        h_x0.copy_(dm.x0)
        h_u0 = torch.empty((dm.B, 2), dtype=torch.float64).pin_memory()
        h_st = torch.empty((dm.B,), dtype=torch.int32).pin_memory()
        h2d, d2h = h_x0.numel() * 8 * world, (h_u0.numel() * 8 + h_st.numel() * 4) * world

        def e2e_step():
            dm.x0.copy_(h_x0, non_blocking=True)
            dm.iterate()
            h_u0.copy_(dm.uold, non_blocking=True)
            h_st.copy_(dm.out["status"], non_blocking=True)
            torch.cuda.synchronize()
            h_x0.copy_(dm.x0)  # the plant is the planner's own prediction (LPV_HP_N_main.py:106)
    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    t_e2e = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
    e2e = {"value": qp_per_step * args.steps / float(t_e2e.item()), "unit": "QP/s", "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": d2h, "ms_per_step": 1e3 * float(t_e2e.item()) / args.steps}
    if single:
        # the same call with the page-locked buffers copied through device staging instead of used in place
        eng.set_options(host_direct=0)
        e2e_step()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_step()
        e2e["staged_value"] = qp_per_step * args.steps / (time.perf_counter() - t0)
        eng.set_options(host_direct=1)
        e2e["path"] = ("dmpc_solve_batch_host on page-locked buffers, read/written in place by the kernel over PCIe "
                       "(host_direct=1); staged_value: same call with H2D/D2H staging copies")

    # ---- CPU baseline (rank 0, N=1 only): the reference path restated, on the box's host cores, bounded sample
    cpu = None
    if single and not args.no_cpu_baseline:
        cores = len(os.sched_getaffinity(0))
        sample = max(48, min(BATCH, 256 * cores))
        rate, secs, o = cpu_reference_rate(ins, cfg, track, cores, sample)
        cpu = {"value": rate, "unit": "QP/s", "cores": cores, "kind": "port",
               "sample": "first %d of the 4096 QPs, %.1f s wall, OSQP restatement at the reference's settings (eps 1e-3, "
                         "polish, fresh setup per call) with C assembly" % (sample, secs),
               "osqp_iters_mean": float(o["iters"].mean())}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "QP/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": workload, "horizon": HORIZON, "nn": NN, "qp_per_step": qp_per_step,
                           "parallelism": par, "l2": "flushed between timed steps (256 MiB memset)",
                           "outputs": "Solution z, planes, obj, status (duals off; parity tests run with duals on)",
                           "launch_order": "QPs start in decreasing order of the previous step's solve cost (dmpc_options."
                                           "order_hint, scheduling only: every QP is solved from scratch every step); "
                                           "`unordered` = same steps with the hint off"},
                "unordered": unordered,
                "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
                "p50_step_ms": float(np.median(per_step)), "wall_ms_per_step": 1e3 * t_wall / args.steps,
                "solved": n_ok, "qp_local": int(status.size), "ipm_iters_mean": iters_mean, "polish_rounds_mean": rounds_mean}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--warm", type=int, default=1, help="dmpc_options.warm_start")
    ap.add_argument("--engine", type=int, default=0, help="dmpc_options.engine: 0 auto (stage-wise), 2 condensed, 3 stage-wise")
    ap.add_argument("--settle", type=int, default=30, help="DMPC workload: closed-loop iterations before warm-up")
    ap.add_argument("--workload", default="batch", choices=["batch", "dmpc"],
                    help="N=1 only: 'dmpc' runs the multi-GPU step (16-agent platoons, closed loop) on one GPU")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()

#!/bin/bash
# Round-2 final 1-GPU evidence: GPU tests, full N=1 line, reference arm, DMPC workload on one GPU, sim bench, roofline
# counts, launch list, ncu --set full of the stage kernel.  $1 = tag
T=${1:-r02f}
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log)
tail -n 5 gpurun_out/${T}_pytest.log | cut -c1-250
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/${T}_bench_n1.json 2> gpurun_out/${T}_bench_n1.err; echo "bench n1 rc=$?"; tail -3 gpurun_out/${T}_bench_n1.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${T}_bench_ref_n1.json 2> gpurun_out/${T}_bench_ref_n1.err; echo "bench ref rc=$?"
timeout 600 python bench.py --workload dmpc --steps 20 --warmup 3 > gpurun_out/${T}_bench_dmpc1.json 2> gpurun_out/${T}_bench_dmpc1.err; echo "bench dmpc rc=$?"; tail -3 gpurun_out/${T}_bench_dmpc1.err
timeout 300 python scripts/gpu_sim_bench.py > gpurun_out/${T}_sim.json 2> gpurun_out/${T}_sim.err; echo "sim rc=$?"; cat gpurun_out/${T}_sim.json
bash scripts/gpu_roofline.sh
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-sweep > gpurun_out/${T}_launches.log 2>&1; echo "launch list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lpv_qp_stage_kernel --launch-skip 6 -c 1 -o gpurun_out/${T}_stage python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-sweep > gpurun_out/${T}_ncu.log 2>&1; echo "ncu rc=$?"
python - "$T" <<'PY'
import json,glob,sys
for f in sorted(glob.glob("gpurun_out/%s_bench_*.json" % sys.argv[1])):
    try: d=json.loads(open(f).read().strip().split("\n")[-1])
    except Exception as e: print(f, "unparsable", e); continue
    print(f, {k: d.get(k) for k in ("value","ms_per_step","solved","qp_timed","infeasible","ipm_iters_mean","polish_rounds_mean","gpu_launches")})
    for k in ("replay","cold","e2e","roofline","cpu_baseline"):
        if d.get(k): print("   ", k, json.dumps(d[k])[:500])
    if d.get("sweep"):
        for s in d["sweep"]: print("    sweep", json.dumps(s)[:500])
PY

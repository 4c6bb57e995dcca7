#!/bin/bash
# One GPU-box session (round 2): GPU tests, bench N=1 (full line), bench DMPC on one GPU, roofline counts.
T=${1:-r02a}
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log)
tail -n 25 gpurun_out/${T}_pytest.log | cut -c1-250
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/${T}_bench_n1.json 2> gpurun_out/${T}_bench_n1.err; echo "bench n1 rc=$?"; tail -3 gpurun_out/${T}_bench_n1.err
timeout 600 python bench.py --workload dmpc --steps 20 --warmup 3 > gpurun_out/${T}_bench_dmpc1.json 2> gpurun_out/${T}_bench_dmpc1.err; echo "bench dmpc rc=$?"; tail -3 gpurun_out/${T}_bench_dmpc1.err
bash scripts/gpu_roofline.sh
python - "$T" <<'PY'
import json,glob,sys
for f in sorted(glob.glob("gpurun_out/%s_bench_*.json" % sys.argv[1] if len(sys.argv)>1 else "gpurun_out/*_bench_*.json")):
    try: d=json.loads(open(f).read().strip().split("\n")[-1])
    except Exception as e: print(f, "unparsable", e); continue
    print(f, {k: d.get(k) for k in ("value","ms_per_step","solved","qp_timed","infeasible","ipm_iters_mean","polish_rounds_mean","gpu_launches")})
    for k in ("replay","cold","e2e","weak_baseline","step_breakdown_ms","config3","cpu_baseline"): 
        if d.get(k): print("   ", k, json.dumps(d[k])[:700])
    if d.get("sweep"):
        for s in d["sweep"]: print("    sweep", json.dumps(s)[:600])
PY

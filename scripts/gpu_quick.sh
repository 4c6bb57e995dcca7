#!/bin/bash
# GPU tests + the N=1 bench line (no CPU baseline, no sweep unless $2 = full).  $1 = tag
T=${1:-q}
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log)
tail -n 6 gpurun_out/${T}_pytest.log | cut -c1-250
if [ "$2" = "full" ]; then EXTRA=""; else EXTRA="--no-cpu-baseline --no-sweep"; fi
timeout 900 python bench.py --steps 20 --warmup 3 $EXTRA > gpurun_out/${T}_bench_n1.json 2> gpurun_out/${T}_bench_n1.err; echo "bench n1 rc=$?"; tail -3 gpurun_out/${T}_bench_n1.err
python - "$T" <<'PY'
import json,sys
d=json.loads(open("gpurun_out/%s_bench_n1.json" % sys.argv[1]).read().strip().split("\n")[-1])
print({k: d.get(k) for k in ("value","ms_per_step","solved","qp_timed","infeasible","ipm_iters_mean","polish_rounds_mean","p50_step_ms")})
for k in ("replay","cold","e2e"):
    if d.get(k): print("   ", k, json.dumps(d[k])[:300])
for s in d.get("sweep") or []: print("    sweep", json.dumps(s)[:400])
PY

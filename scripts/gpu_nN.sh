#!/bin/bash
# N-GPU bench with the driver's own command line, under a watchdog: $1 = N, $2 = tag
N=${1:-2}; T=${2:-r02}
mkdir -p gpurun_out
t0=$(date +%s)
DMPC_BENCH_WATCHDOG=280 timeout 320 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/${T}_bench_n$N.json 2> gpurun_out/${T}_bench_n$N.err
echo "n=$N rc=$? wall=$(( $(date +%s) - t0 ))s"; grep -i "FAIL\|Error\|Timeout" gpurun_out/${T}_bench_n$N.err | head -5
python - "$N" "$T" <<'PY'
import json, sys
d = json.loads(open("gpurun_out/%s_bench_n%s.json" % (sys.argv[2], sys.argv[1])).read().strip().split("\n")[-1])
for k in ("value", "ms_per_step", "efficiency_same_workload", "weak_baseline", "step_breakdown_ms", "e2e", "qp_timed", "solved", "solved_inaccurate", "not_feasible", "breakdowns", "gpu_launches", "roofline", "config3"):
    print(k, json.dumps(d.get(k))[:1100])
PY

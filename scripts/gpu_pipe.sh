#!/bin/bash
# N-GPU DMPC bench with the two-group pipelined measurement: $1 = N, $2 = tag
N=${1:-2}; T=${2:-r02k}
mkdir -p gpurun_out
DMPC_BENCH_WATCHDOG=280 timeout 320 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 --pipelined --no-config3 > gpurun_out/${T}_pipe_n$N.json 2> gpurun_out/${T}_pipe_n$N.err
echo "n=$N rc=$?"; grep -i "FAIL\|Error\|Timeout" gpurun_out/${T}_pipe_n$N.err | head -5
python - "$N" "$T" <<'PY'
import json, sys
d = json.loads(open("gpurun_out/%s_pipe_n%s.json" % (sys.argv[2], sys.argv[1])).read().strip().split("\n")[-1])
print("value", d["value"], "ms", d["ms_per_step"], "eff", d["efficiency_same_workload"], "weak", d["weak_baseline"]["value"], d["weak_baseline"].get("pipelined_value"))
print("pipelined", d.get("pipelined"))
PY

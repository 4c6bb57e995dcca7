#!/bin/bash
# One GPU-box session (round 2): GPU tests, config[1] bench on both engines / warm start on-off.  Logs under gpurun_out/.
T=${1:-r02a}
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log)
for cfgs in "3 1" "3 0" "2 1"; do set -- $cfgs
timeout 300 python bench.py --steps 20 --warmup 3 --engine $1 --warm $2 --no-cpu-baseline > gpurun_out/${T}_bench_e$1_w$2.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_bench_e$1_w$2.log
done
tail -n 30 gpurun_out/${T}_pytest.log | cut -c1-250
for f in gpurun_out/${T}_bench_e*.log; do echo $f; tail -n 2 $f | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: print(l[:300]); continue
    print({k: d[k] for k in ('value','ms_per_step','solved','qp_local','ipm_iters_mean','polish_rounds_mean','gpu_launches')}, 'e2e', d['e2e']['value'], 'unordered', d.get('unordered'))
"; done

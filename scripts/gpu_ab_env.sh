#!/bin/bash
# A/B of one environment switch of the product build in one GPU session: GPU tests, then the quick N=1 bench with the
# switch on and off.  $1 = tag, $2 = variable name
T=${1:-abe}; V=${2:-DMPC_B200_BULK_IN}
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log)
tail -n 15 gpurun_out/${T}_pytest.log | cut -c1-300
for v in 1 0 1 0; do
  export $V=$v
  timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-sweep > gpurun_out/${T}_bench_$v.json 2> gpurun_out/${T}_bench_$v.err; echo "bench $V=$v rc=$?"; tail -3 gpurun_out/${T}_bench_$v.err
  python - "$T" $v <<'PY'
import json,sys
d=json.loads(open("gpurun_out/%s_bench_%s.json" % (sys.argv[1], sys.argv[2])).read().strip().split("\n")[-1])
print(sys.argv[2], {k: d.get(k) for k in ("value","ms_per_step","solved","qp_timed","p50_step_ms")}, {k: (d[k].get("value"), d[k].get("ms_per_step")) for k in ("replay","cold","e2e") if d.get(k)})
PY
done

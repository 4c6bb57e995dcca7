#!/bin/bash
# A/B of two builds of the C ABI in one GPU session: GPU tests on the product build, then the quick N=1 bench on both.
# $1 = tag, $2 = path of the other build
T=${1:-ab}
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log)
tail -n 15 gpurun_out/${T}_pytest.log | cut -c1-300
for v in new old; do
  if [ $v = old ]; then export DMPC_B200_LIB=$2; fi
  timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-sweep > gpurun_out/${T}_bench_$v.json 2> gpurun_out/${T}_bench_$v.err; echo "bench $v rc=$?"; tail -3 gpurun_out/${T}_bench_$v.err
  python - "$T" $v <<'PY'
import json,sys
d=json.loads(open("gpurun_out/%s_bench_%s.json" % (sys.argv[1], sys.argv[2])).read().strip().split("\n")[-1])
print(sys.argv[2], {k: d.get(k) for k in ("value","ms_per_step","solved","qp_timed","ipm_iters_mean","polish_rounds_mean","p50_step_ms")})
for k in ("replay","cold","e2e"):
    if d.get(k): print("   ", k, {kk: d[k].get(kk) for kk in ("value", "ms_per_step")})
PY
done

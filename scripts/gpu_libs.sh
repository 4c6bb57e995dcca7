#!/bin/bash
# quick N=1 bench of several builds of the C ABI in one GPU session.  $1 = tag, rest = library paths ("-" = product build)
T=$1; shift
mkdir -p gpurun_out
i=0
for L in "$@"; do
  i=$((i+1))
  if [ "$L" = "-" ]; then unset DMPC_B200_LIB; else export DMPC_B200_LIB=$L; fi
  timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-sweep > gpurun_out/${T}_bench_$i.json 2> gpurun_out/${T}_bench_$i.err; echo "bench $L rc=$?"; tail -3 gpurun_out/${T}_bench_$i.err
  python - "$T" $i "$L" <<'PY'
import json,sys
d=json.loads(open("gpurun_out/%s_bench_%s.json" % (sys.argv[1], sys.argv[2])).read().strip().split("\n")[-1])
print(sys.argv[3], {k: d.get(k) for k in ("value","ms_per_step","solved","p50_step_ms")}, {k: (round(d[k].get("value")), round(d[k].get("ms_per_step"),3)) for k in ("replay","cold","e2e") if d.get(k)})
PY
done

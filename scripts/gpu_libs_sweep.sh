#!/bin/bash
# GPU tests on the product build, then N=1 bench WITH the horizon sweep for several builds.  $1 = tag, rest = libs ("-" = product)
T=$1; shift
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log)
tail -n 4 gpurun_out/${T}_pytest.log | cut -c1-300
i=0
for L in "$@"; do
  i=$((i+1))
  if [ "$L" = "-" ]; then unset DMPC_B200_LIB; else export DMPC_B200_LIB=$L; fi
  timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_$i.json 2> gpurun_out/${T}_bench_$i.err; echo "bench $L rc=$?"; tail -3 gpurun_out/${T}_bench_$i.err
  python - "$T" $i "$L" <<'PY'
import json,sys
d=json.loads(open("gpurun_out/%s_bench_%s.json" % (sys.argv[1], sys.argv[2])).read().strip().split("\n")[-1])
print(sys.argv[3], {k: d.get(k) for k in ("value","ms_per_step","solved","p50_step_ms")}, {k: (round(d[k].get("value")), round(d[k].get("ms_per_step"),3)) for k in ("replay","cold","e2e") if d.get(k)})
print("   sweep cold/replay:", [(s["N"], round(s["gpu_cold_qp_per_s"]), round(s["gpu_replay_qp_per_s"]), s["status_not_1"]) for s in d.get("sweep") or []])
PY
done

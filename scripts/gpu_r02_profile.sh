#!/bin/bash
# Round-2 evidence session: GPU tests, the N=1 bench line, the ncu launch list of the same command, and one
# `ncu --set full` capture of the stage kernel on a timed step.  $1 = tag
T=${1:-r02p}
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log)
tail -n 6 gpurun_out/${T}_pytest.log | cut -c1-250
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/${T}_bench_n1.json 2> gpurun_out/${T}_bench_n1.err; echo "bench n1 rc=$?"; tail -3 gpurun_out/${T}_bench_n1.err
cut -c1-600 gpurun_out/${T}_bench_n1.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-sweep > gpurun_out/${T}_launches.log 2>&1; echo "launch list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lpv_qp_stage_kernel --launch-skip 6 -c 1 -o gpurun_out/${T}_stage python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-sweep > gpurun_out/${T}_ncu.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/

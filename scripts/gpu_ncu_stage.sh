#!/bin/bash
# one `ncu --set full` capture of the stage kernel on a timed stream step.  $1 = tag
T=${1:-ncu}
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lpv_qp_stage_kernel --launch-skip 6 -c 1 -o gpurun_out/${T}_stage python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-sweep > gpurun_out/${T}_ncu.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/${T}_stage.ncu-rep

#!/bin/bash
# bench (stage engine) then one ncu --set full capture of the stage kernel.  $1 = tag
T=${1:-r02b}
mkdir -p gpurun_out
timeout 300 python bench.py --steps 20 --warmup 3 --engine 3 --no-cpu-baseline > gpurun_out/${T}_bench_stage.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_bench_stage.log
tail -n 2 gpurun_out/${T}_bench_stage.log | cut -c1-400
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lpv_qp_stage_kernel --launch-skip 6 -c 1 -o gpurun_out/${T}_stage python bench.py --steps 2 --warmup 3 --engine 3 --no-cpu-baseline > gpurun_out/${T}_ncu.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/${T}_stage.ncu-rep

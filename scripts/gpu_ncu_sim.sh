#!/bin/bash
# one `ncu --set full` capture of the roll-out kernel.  $1 = tag
T=${1:-sim}
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lpv_sim_kernel --launch-skip 5 -c 1 -o gpurun_out/${T}_sim python scripts/gpu_sim_bench.py > gpurun_out/${T}_ncu_sim.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/${T}_sim.ncu-rep

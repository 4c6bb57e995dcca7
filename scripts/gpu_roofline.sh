#!/bin/bash
# ncu counts of one timed launch of the solve kernel for the two bench workloads -> gpurun_out/r02_counts_{stream,dmpc}.csv
M=smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,smsp__inst_executed.sum
mkdir -p gpurun_out
timeout 600 ncu --metrics $M --clock-control none -k regex:lpv_qp_stage_kernel --launch-skip 30 -c 1 --csv --log-file gpurun_out/r02_counts_stream.csv python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/r02_counts_stream.log 2>&1; echo "stream rc=$?"
timeout 600 ncu --metrics $M --clock-control none -k regex:lpv_qp_stage_kernel --launch-skip 45 -c 1 --csv --log-file gpurun_out/r02_counts_dmpc.csv python bench.py --workload dmpc --no-graph --no-config3 --steps 8 --warmup 3 > gpurun_out/r02_counts_dmpc.log 2>&1; echo "dmpc rc=$?"

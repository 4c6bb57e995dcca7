#!/bin/bash
# One GPU-box session: GPU tests, the DMPC step on one GPU, then the BASELINE config 3/4 sweep.  Logs under gpurun_out/.
mkdir -p gpurun_out
(timeout 600 python -m pytest tests -m gpu -q > gpurun_out/s28_pytest2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s28_pytest2.log)
timeout 300 python bench.py --workload dmpc --steps 20 --warmup 3 > gpurun_out/s28_bench_dmpc1.log 2>&1
: > gpurun_out/s28_sweep.jsonl
for c in "10 2 full 3 1366" "20 2 full 3 1366" "50 2 full 3 1366" "100 2 full 3 342 12" "20 2 chain 64 64" "20 63 full 64 8 12" "20 15 full 16 64 12"; do
  timeout 280 python scripts/gpu_sweep.py $c >> gpurun_out/s28_sweep.jsonl 2>> gpurun_out/s28_sweep.err
  echo "case $c rc=$?" >> gpurun_out/s28_sweep.err
done
tail -n 40 gpurun_out/s28_pytest2.log | cut -c1-300
tail -n 1 gpurun_out/s28_bench_dmpc1.log | cut -c1-400
cut -c1-600 gpurun_out/s28_sweep.jsonl
tail -n 12 gpurun_out/s28_sweep.err

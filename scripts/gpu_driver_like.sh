#!/bin/bash
# what the driver runs at round end, in its order: GPU tests, smoke(), reference arm, own arm (defaults)
mkdir -p gpurun_out
python -m pytest tests/ -x -q -m gpu 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py --impl reference > gpurun_out/drv_ref.json 2> gpurun_out/drv_ref.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/drv_ref.json
t0=$(date +%s); timeout 900 python bench.py > gpurun_out/drv_own.json 2> gpurun_out/drv_own.err; echo "own rc=$? wall=$(( $(date +%s) - t0 ))s"; cut -c1-200 gpurun_out/drv_own.json; tail -2 gpurun_out/drv_own.err

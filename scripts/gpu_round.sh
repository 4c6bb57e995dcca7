#!/bin/bash
# GPU tests on the product build, sim-kernel bench, then the quick bench of several builds.  $1 = tag, rest = library paths
T=$1
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log)
tail -n 8 gpurun_out/${T}_pytest.log | cut -c1-300
timeout 300 python scripts/gpu_sim_bench.py > gpurun_out/${T}_sim.json 2> gpurun_out/${T}_sim.err; echo "sim rc=$?"; cat gpurun_out/${T}_sim.json; tail -3 gpurun_out/${T}_sim.err
bash scripts/gpu_libs.sh "$@"

#!/bin/bash
# sim-kernel parity tests on the product build, then the roll-out bench of several builds.  $1 = tag, rest = libs ("-" = product)
T=$1; shift
mkdir -p gpurun_out
(timeout 600 python -m pytest tests -m gpu -q -x -k "sim or Model" > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log)
tail -n 4 gpurun_out/${T}_pytest.log | cut -c1-300
for L in "$@"; do
  if [ "$L" = "-" ]; then unset DMPC_B200_LIB; else export DMPC_B200_LIB=$L; fi
  echo "== $L"; timeout 300 python scripts/gpu_sim_bench.py 2>&1 | tail -1 | cut -c1-400
done

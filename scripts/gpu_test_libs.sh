#!/bin/bash
# GPU tests on the product build, then the quick bench of several builds.  $1 = tag, rest = library paths ("-" = product)
T=$1
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log)
tail -n 8 gpurun_out/${T}_pytest.log | cut -c1-300
bash scripts/gpu_libs.sh "$@"

#!/bin/bash
# GPU parity tests AND quick bench of a variant build of the C ABI next to the product build.  $1 = tag, $2 = variant library
T=$1; L=$2
mkdir -p gpurun_out
(DMPC_B200_LIB=$L timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest_variant.log 2>&1; echo "pytest(variant) rc=$?" >> gpurun_out/${T}_pytest_variant.log)
tail -n 6 gpurun_out/${T}_pytest_variant.log | cut -c1-300
bash scripts/gpu_libs.sh $T - $L - $L

#!/bin/bash
# 2-GPU bench variants under short limits (development): $1 = extra flags tag
mkdir -p gpurun_out
run() { tag=$1; shift
  DMPC_BENCH_WATCHDOG=150 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 "$@" > gpurun_out/r02_n2_$tag.json 2> gpurun_out/r02_n2_$tag.err
  echo "$tag rc=$?"; grep -v "^\s*$" gpurun_out/r02_n2_$tag.err | grep -i "File \"/root/repo\|Error\|error\|Thread\|FAIL" | head -30
  tail -c 600 gpurun_out/r02_n2_$tag.json; echo; }
run eager --no-graph --no-config3
run graph --no-config3

#!/bin/bash
# tcgen05 acting kernel on the GPU box: its tests (bounded: a barrier mistake hangs the kernel), then both layouts timed
TAG=${1:-umma}
mkdir -p gpurun_out
timeout 180 python -m pytest tests/test_policy_gpu.py -m gpu -x -q -k "sw128" > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -15 gpurun_out/${TAG}_tests.log
nvidia-smi --query-gpu=name,memory.used --format=csv,noheader
for L in rowpad sw128; do
  timeout 240 python tools/policy_rollout.py --worlds 65536 --steps 20 --fused $L > gpurun_out/${TAG}_roll_$L.json 2> gpurun_out/${TAG}_roll_$L.err; echo "roll $L rc=$?"; tail -2 gpurun_out/${TAG}_roll_$L.json
done

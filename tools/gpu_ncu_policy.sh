#!/bin/bash
# ncu --set full of the fused policy kernel (one launch)
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:policy_fused_kernel -s 8 -c 1 -o gpurun_out/${1:-p}_policy -f python tools/policy_rollout.py --worlds 65536 --steps 3 --warmup 3 --fused > gpurun_out/${1:-p}_ncu.log 2>&1
tail -2 gpurun_out/${1:-p}_ncu.log | cut -c1-200

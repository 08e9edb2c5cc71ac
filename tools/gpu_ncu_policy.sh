#!/bin/bash
# ncu --set full of the fused policy kernel (one launch):  gpu_ncu_policy.sh TAG [sw128|rowpad]
L=${2:-sw128}; K=policy_umma_kernel; [ "$L" = rowpad ] && K=policy_fused_kernel
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$K -s 8 -c 1 -o gpurun_out/${1:-p}_policy -f python tools/policy_rollout.py --worlds 65536 --steps 3 --warmup 3 --fused $L > gpurun_out/${1:-p}_ncu.log 2>&1
tail -2 gpurun_out/${1:-p}_ncu.log | cut -c1-200

#!/bin/bash
# ncu --set full of the FAST export kernel (one launch); TAG names the report
TAG=${1:-x}
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"export_kernel" -s 8 -c 2 -o gpurun_out/${TAG}_export -f python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu.log 2>&1
tail -2 gpurun_out/${TAG}_ncu.log | cut -c1-200

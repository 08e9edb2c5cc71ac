#!/bin/bash
# ncu --set full + source of one launch of the export kernel variant chosen by MBOTS_EXPORT_WPW; TAG names the report
TAG=${1:-x}; V=${2:-1}; KRE=${3:-export_lanes_kernel}
mkdir -p gpurun_out
MBOTS_EXPORT_WPW=$V timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$KRE" -s 8 -c 1 -o gpurun_out/${TAG}_lanes -f python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu.log 2>&1
tail -2 gpurun_out/${TAG}_ncu.log | cut -c1-200

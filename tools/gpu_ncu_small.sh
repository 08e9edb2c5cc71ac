#!/bin/bash
# why is the export kernel at 8,192 worlds 1.5x its share of the 65,536-world time? ncu --set full at 8,192 / 1,024 worlds
TAG=${1:-r2f}
mkdir -p gpurun_out
for W in 8192 1024; do
 for S in 0 1; do
  MBOTS_GRAPHS=0 MBOTS_EXPORT_SPLIT_LOG2=$S timeout 300 ncu --set full --clock-control none -k export_kernel -s 12 -c 2 -o gpurun_out/${TAG}_export_${W}w_s$S -f python bench.py --worlds $W --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_${W}_$S.log 2>&1
  tail -1 gpurun_out/${TAG}_ncu_${W}_$S.log | cut -c1-160
 done
done

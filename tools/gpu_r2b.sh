#!/bin/bash
# export_lanes_kernel: parity under every variant, then A/B timing against the round-1 FAST kernel
TAG=${1:-r2b}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; tail -5 gpurun_out/${TAG}_tests.log
for V in 1 2 4; do
  MBOTS_EXPORT_WPW=$V timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "small or ragged or limit or golden or 2500 or capacity or large" > gpurun_out/${TAG}_tests_wpw$V.log 2>&1; tail -2 gpurun_out/${TAG}_tests_wpw$V.log
done
for W in 65536 8192 1024; do
  for V in 0 1 2 4; do
    MBOTS_EXPORT_WPW=$V timeout 300 python bench.py --worlds $W --no-cpu-baseline --steps 30 > gpurun_out/${TAG}_bench_${W}w_v$V.json 2> gpurun_out/${TAG}_bench_${W}w_v$V.err
    python -c "
import json;d=json.load(open('gpurun_out/${TAG}_bench_${W}w_v$V.json'));print($W, 'variant', $V, round(d['value']/1e9,3),round(d['ms_per_step'],4),{k:round(v['ms'],4) for k,v in d['kernels'].items()})"
  done
done

#!/bin/bash
TAG=${1:-r2i}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; tail -3 gpurun_out/${TAG}_tests.log
for W in 65536 16384 8192 1024; do
  timeout 300 python bench.py --worlds $W --no-cpu-baseline > gpurun_out/${TAG}_bench_${W}w.json 2> gpurun_out/${TAG}_bench_${W}w.err || tail -5 gpurun_out/${TAG}_bench_${W}w.err
  python -c "
import json;d=json.load(open('gpurun_out/${TAG}_bench_${W}w.json'));print($W, round(d['value']/1e9,3),'G', round(d['ms_per_step'],4),'ms | sync/step', round(d['value_sync_per_step']['ms_per_step'],4), '| e2e', round(d['e2e']['value']/1e6,1),'M | launches/step', d['launches_per_step'], d['timing']['repeats']);print({k:round(v['ms'],4) for k,v in d['kernels'].items()})"
done

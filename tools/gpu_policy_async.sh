#!/bin/bash
# acting path with device-resolved row ranges: tests, then the policy leg of the bench at two sizes
TAG=${1:-pa}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_policy_gpu.py -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; tail -4 gpurun_out/${TAG}_tests.log
for W in 65536 8192; do
  timeout 400 python bench.py --worlds $W --policy --no-cpu-baseline --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_policy_${W}w.json 2> gpurun_out/${TAG}_bench_policy_${W}w.err || tail -5 gpurun_out/${TAG}_bench_policy_${W}w.err
  python -c "
import json;d=json.load(open('gpurun_out/${TAG}_bench_policy_${W}w.json'));p=d['policy_rollout']
print($W, 'sim', round(d['ms_per_step'],4), '| blocking', round(p['value']/1e9,3),'G', round(p['ms_per_step'],4),'ms (policy', round(p['policy_ms_per_step'],4), ') | async', round(p['async']['value']/1e9,3),'G', round(p['async']['ms_per_step'],4),'ms | tensor frac', round(p['roofline']['frac'],3))"
done

#!/bin/bash
# final pass of a round: all GPU tests, the checked build, bench lines at the four sizes, ncu evidence of the same code
TAG=${1:-final}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; tail -3 gpurun_out/${TAG}_tests.log
bash tools/gpu_checked.sh ${TAG}
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; tail -1 gpurun_out/${TAG}_smoke.log
timeout 400 python bench.py > gpurun_out/${TAG}_bench_65536w.json 2> gpurun_out/${TAG}_bench_65536w.err || tail -5 gpurun_out/${TAG}_bench_65536w.err
for W in 16384 8192 1024; do
  timeout 300 python bench.py --worlds $W --no-cpu-baseline > gpurun_out/${TAG}_bench_${W}w.json 2> gpurun_out/${TAG}_bench_${W}w.err || tail -5 gpurun_out/${TAG}_bench_${W}w.err
done
for W in 65536 16384 8192 1024; do
  python -c "
import json;d=json.load(open('gpurun_out/${TAG}_bench_${W}w.json'));print($W, round(d['value']/1e9,3),'G', round(d['ms_per_step'],4),'ms | sync/step', round(d['value_sync_per_step']['ms_per_step'],4), '| e2e', round(d['e2e']['value']/1e6,1),'M', round(d['e2e']['link_frac'],3), '| roofline', round(d['roofline']['frac'],3), d['roofline']['issue_frac'], d['roofline']['traffic_source'][:30]);print({k:round(v['ms'],4) for k,v in d['kernels'].items()}, d.get('cpu_baseline',{}).get('value'))"
done
timeout 300 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/${TAG}_reference_arm.json 2>/dev/null; cut -c1-300 gpurun_out/${TAG}_reference_arm.json
bash tools/gpu_ncu_capture.sh ${TAG} 65536

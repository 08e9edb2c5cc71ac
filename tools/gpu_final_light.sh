#!/bin/bash
# final pass without the ncu capture (kernel sources unchanged since the committed capture): all GPU tests, the checked
# build, smoke, the default bench line + the reference arm, and the policy leg
TAG=${1:-final}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; tail -3 gpurun_out/${TAG}_tests.log
MBOTS_LIB=$PWD/madrona-bots_b200/lib/variants/libmbots_checked.so timeout 1200 python -m pytest tests/test_parity_gpu.py tests/test_checkpoint_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/${TAG}_checked_tests.log 2>&1; tail -2 gpurun_out/${TAG}_checked_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; tail -1 gpurun_out/${TAG}_smoke.log
timeout 300 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/${TAG}_reference_arm.json 2>/dev/null; cut -c1-200 gpurun_out/${TAG}_reference_arm.json
timeout 500 python bench.py --policy > gpurun_out/${TAG}_bench_65536w.json 2> gpurun_out/${TAG}_bench_65536w.err || tail -5 gpurun_out/${TAG}_bench_65536w.err
python -c "
import json;d=json.load(open('gpurun_out/${TAG}_bench_65536w.json'));p=d.get('policy_rollout') or {}
print(round(d['value']/1e9,3),'G', round(d['ms_per_step'],4),'ms | sync/step', round(d['value_sync_per_step']['ms_per_step'],4), '| e2e', round(d['e2e']['value']/1e6,1),'M', round(d['e2e']['link_frac'],3), '| roofline', round(d['roofline']['frac'],3), d['roofline']['issue_frac'], d['roofline']['traffic_source'][:40], '| cpu', d.get('cpu_baseline',{}).get('value'), '| policy async', round((p.get('async') or {}).get('value',0)/1e9,3))
print({k:round(v['ms'],4) for k,v in d['kernels'].items()}, d['clocks'])"

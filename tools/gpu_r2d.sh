#!/bin/bash
# graphs + splits: parity (default + forced split levels + graphs off), then the bench at the strong-scaling sizes
TAG=${1:-r2d}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; tail -5 gpurun_out/${TAG}_tests.log
for V in 0 1 2; do
  MBOTS_EXPORT_SPLIT_LOG2=$V timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "small or ragged or limit or golden or 2500 or capacity or large or graph or reference_text" > gpurun_out/${TAG}_tests_split$V.log 2>&1; tail -2 gpurun_out/${TAG}_tests_split$V.log
done
MBOTS_GRAPHS=0 timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "small or graph or async" > gpurun_out/${TAG}_tests_nograph.log 2>&1; tail -2 gpurun_out/${TAG}_tests_nograph.log
for W in 65536 16384 8192 1024; do
  timeout 300 python bench.py --worlds $W --no-cpu-baseline > gpurun_out/${TAG}_bench_${W}w.json 2> gpurun_out/${TAG}_bench_${W}w.err || tail -5 gpurun_out/${TAG}_bench_${W}w.err
  python -c "
import json;d=json.load(open('gpurun_out/${TAG}_bench_${W}w.json'));print($W, round(d['value']/1e9,3),'G', round(d['ms_per_step'],4),'ms | sync/step', round(d['value_sync_per_step']['ms_per_step'],4), '| e2e', round(d['e2e']['value']/1e6,1),'M link', round(d['e2e']['link_gbs_measured'],1), '| launches/step', d['launches_per_step'], d['timing']['repeats']);print({k:round(v['ms'],4) for k,v in d['kernels'].items()})"
done
for S in 0 1; do
  for W in 8192; do
    MBOTS_EXPORT_SPLIT_LOG2=$S timeout 300 python bench.py --worlds $W --no-cpu-baseline > gpurun_out/${TAG}_bench_${W}w_s$S.json 2> /dev/null
    python -c "
import json;d=json.load(open('gpurun_out/${TAG}_bench_${W}w_s$S.json'));print($W,'split_log2',$S, round(d['ms_per_step'],4),'ms', {k:round(v['ms'],4) for k,v in d['kernels'].items()})"
  done
done

#!/bin/bash
# quick A/B pass: parity subset, bench at 65,536 and 8,192 worlds, instruction counts of one export / step launch
TAG=${1:-ab}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "small or ragged or limit or golden or 2500 or capacity or large or graph or reference_text or 16384" > gpurun_out/${TAG}_tests.log 2>&1; tail -2 gpurun_out/${TAG}_tests.log
for W in 65536 8192; do
  timeout 300 python bench.py --worlds $W --no-cpu-baseline > gpurun_out/${TAG}_bench_${W}w.json 2> gpurun_out/${TAG}_bench_${W}w.err || tail -5 gpurun_out/${TAG}_bench_${W}w.err
  python -c "
import json;d=json.load(open('gpurun_out/${TAG}_bench_${W}w.json'));print($W, round(d['value']/1e9,3),'G', round(d['ms_per_step'],4),'ms', {k:round(v['ms'],4) for k,v in d['kernels'].items()})"
done
MBOTS_GRAPHS=0 timeout 300 ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio --clock-control none -k regex:"export_kernel|step_kernel" -s 40 -c 3 --csv --log-file gpurun_out/${TAG}_inst.csv python bench.py --worlds 65536 --steps 5 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open('gpurun_out/${TAG}_inst.csv')) if len(r)>10]
h=rows[0]
for r in rows[1:]:
    d=dict(zip(h,r)); print(d['Kernel Name'][:60], d['Metric Name'], d['Metric Value'])
PY

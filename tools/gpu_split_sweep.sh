#!/bin/bash
# sweep of the export kernel's split settings over launch sizes (tools/split_sweep.py), parity subset first
TAG=${1:-sweep}
mkdir -p gpurun_out
OUT=gpurun_out/${TAG}_sweep.jsonl
: > $OUT
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "small or ragged or limit or golden or 2500 or capacity or large or graph or reference_text" > gpurun_out/${TAG}_tests.log 2>&1; tail -2 gpurun_out/${TAG}_tests.log
for W in 8192 16384 32768 65536; do
  for CFG in "0 0" "1 0" "2 0"; do   # (parts log2, heavy64 -- the latter only in the r3c experiment)
    set -- $CFG
    MBOTS_EXPORT_SPLIT_LOG2=$1 MBOTS_EXPORT_SPLIT_HEAVY64=$2 timeout 200 python tools/split_sweep.py $W 50 >> $OUT 2>> gpurun_out/${TAG}_sweep.err
  done
done
python - <<PY
import json
for l in open('$OUT'):
    d=json.loads(l); print(d['worlds'], d['split_log2'], d['heavy64'], d['ms_per_step'], d['ms_min'], d['kernels_ms']['export_sense'], d['kernels_ms']['step'], d['n_mean'], d['n_std'], d['n_max'])
PY

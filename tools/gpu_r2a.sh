#!/bin/bash
# round-2 baseline pass: parity tests (incl. the new sizes), bench at the strong-scaling per-GPU sizes, sanitizer
TAG=${1:-r2a}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/${TAG}_tests.log 2>&1; tail -15 gpurun_out/${TAG}_tests.log
for W in 65536 16384 8192 1024; do
  timeout 300 python bench.py --worlds $W --no-cpu-baseline > gpurun_out/${TAG}_bench_${W}w.json 2> gpurun_out/${TAG}_bench_${W}w.err
  python -c "
import json;d=json.load(open('gpurun_out/${TAG}_bench_${W}w.json'));print($W, d['value'],d['ms_per_step'],d['e2e']['value'],d['gpu_launches']);print({k:round(v['ms'],4) for k,v in d['kernels'].items()})"
done
sed -i 's/timeout 900 \$SAN/timeout 300 $SAN/' tools/gpu_sanitize.sh
bash tools/gpu_sanitize.sh ${TAG}

#!/bin/bash
# quick GPU-box pass: parity tests + bench (no CPU baseline); TAG names the outputs
TAG=${1:-q}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; tail -3 gpurun_out/${TAG}_tests.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python -c "
import json;d=json.load(open('gpurun_out/${TAG}_bench.json'));print(d['value'],d['ms_per_step'],d['e2e']['value'],d['gpu_launches']);print({k:round(v['ms'],4) for k,v in d['kernels'].items()})"; tail -3 gpurun_out/${TAG}_bench.err

#!/bin/bash
# bench.py per-kernel times for every library build under lib/variants (A/B timing of kernel variants)
mkdir -p gpurun_out
for lib in default madrona-bots_b200/lib/variants/*.so; do
  if [ "$lib" = default ]; then unset MBOTS_LIB; else export MBOTS_LIB=$PWD/$lib; fi
  timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);k=d['kernels']
print('$lib', 'step_ms', round(d['ms_per_step'],4), {n:round(v['ms'],4) for n,v in k.items()}, 'e2e', round(d['e2e']['value']/1e6,1))"
done

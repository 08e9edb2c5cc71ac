#!/bin/bash
# A/B: bench every library build under madrona-bots_b200/lib/variants (plus the default one)
mkdir -p gpurun_out
for lib in default madrona-bots_b200/lib/variants/*.so; do
  if [ "$lib" = default ]; then unset MBOTS_LIB; else export MBOTS_LIB=$PWD/$lib; fi
  python bench.py --no-cpu-baseline --steps 30 > gpurun_out/var.json 2>/dev/null
  python -c "
import json;d=json.load(open('gpurun_out/var.json'));print('$lib', round(d['ms_per_step'],4),{k:round(v['ms'],4) for k,v in d['kernels'].items()})"
done

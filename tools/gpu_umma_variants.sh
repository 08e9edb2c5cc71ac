#!/bin/bash
# time the acting kernel of every library build under lib/variants (timing experiments; results may be wrong by design)
mkdir -p gpurun_out
for lib in default madrona-bots_b200/lib/variants/*.so; do
  if [ "$lib" = default ]; then unset MBOTS_LIB; else export MBOTS_LIB=$PWD/$lib; fi
  timeout 200 python tools/policy_rollout.py --worlds 65536 --steps 20 --fused sw128 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('$lib', round(d['policy_ms_per_step'],4))"
done

#!/bin/bash
# scaling pass on an N-GPU box, launched the way the driver launches bench.py: strong scaling (65,536 worlds in total),
# the weak companion leg and the policy leg ride along in the same JSON line. TAG names the outputs.
TAG=${1:-scale}; STEPS=${2:-20}; WARM=${3:-5}
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
for N in ${NS:-1 2 4 8}; do        # NS="8" runs one point only
  [ $N -le $NG ] || continue
  if [ $N -eq 1 ]; then
    timeout 500 python bench.py --gpus 1 --steps $STEPS --warmup $WARM --no-cpu-baseline --policy > gpurun_out/${TAG}_n$N.json 2> gpurun_out/${TAG}_n$N.err
  else
    timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500+N)) bench.py --gpus $N --steps $STEPS --warmup $WARM --policy > gpurun_out/${TAG}_n$N.json 2> gpurun_out/${TAG}_n$N.err
  fi
  python -c "
import json;d=json.load(open('gpurun_out/${TAG}_n$N.json'));w=d.get('weak_65536_per_gpu') or {};p=d.get('policy_rollout') or {}
print($N, 'strong', round(d['value']/1e9,3),'G', round(d['ms_per_step'],4),'ms | sync/step', round(d['value_sync_per_step']['value']/1e9,3), '| e2e', round(d['e2e']['value']/1e6,1),'M link', round(d['e2e']['link_gbs_all_ranks'],1), '| weak', round(w.get('value',0)/1e9,3), '| policy', round(p.get('value',0)/1e9,3),'G', round(p.get('policy_ms_per_step',0),3),'ms', round((p.get('roofline') or {}).get('frac',0),3), '| policy async', round((p.get('async') or {}).get('value',0)/1e9,3),'G', round((p.get('async') or {}).get('ms_per_step',0),4),'ms')" || tail -5 gpurun_out/${TAG}_n$N.err
done

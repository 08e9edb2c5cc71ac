#!/bin/bash
# scaling pass on an N-GPU box: ours + reference arm at every N <= available
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
for N in 1 2 4 8; do
  [ $N -le $NG ] || continue
  if [ $N -eq 1 ]; then
    timeout 400 python bench.py --gpus 1 --steps 50 --warmup 10 --no-cpu-baseline > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
  else
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500+N)) bench.py --gpus $N --steps 50 --warmup 10 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
  fi
  python -c "
import json;d=json.load(open('gpurun_out/scale_n$N.json'));print($N, d['value'], d['ms_per_step'], d['e2e']['value'])" || tail -3 gpurun_out/scale_n$N.err
done

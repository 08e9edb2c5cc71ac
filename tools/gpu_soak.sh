#!/bin/bash
# parity soak of the final kernels: blocking path and fused graph path, long rollouts against the oracle
TAG=${1:-r2}
mkdir -p gpurun_out
: > gpurun_out/${TAG}_soak.jsonl
timeout 1500 python tools/soak.py --worlds 256 --steps 3000 --mode bits --check-every 250 >> gpurun_out/${TAG}_soak.jsonl 2>> gpurun_out/${TAG}_soak.err
timeout 1500 python tools/soak.py --worlds 256 --steps 3000 --mode bits --check-every 250 --fused >> gpurun_out/${TAG}_soak.jsonl 2>> gpurun_out/${TAG}_soak.err
timeout 1500 python tools/soak.py --worlds 2048 --steps 1000 --mode onehot --check-every 100 --fused >> gpurun_out/${TAG}_soak.jsonl 2>> gpurun_out/${TAG}_soak.err
timeout 1500 python tools/soak.py --worlds 64 --steps 1500 --mode bits --check-every 100 --cap 64 --fused >> gpurun_out/${TAG}_soak.jsonl 2>> gpurun_out/${TAG}_soak.err
timeout 1500 python tools/soak.py --worlds 16384 --steps 300 --mode onehot --check-every 100 --fused >> gpurun_out/${TAG}_soak.jsonl 2>> gpurun_out/${TAG}_soak.err
cat gpurun_out/${TAG}_soak.jsonl; tail -3 gpurun_out/${TAG}_soak.err

#!/bin/bash
TAG=${1:-r2g}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; tail -3 gpurun_out/${TAG}_tests.log
bash tools/gpu_checked.sh ${TAG}
bash tools/gpu_scale.sh ${TAG}_scale 20 5

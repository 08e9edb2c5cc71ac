#!/bin/bash
# ncu evidence of the CURRENT kernels: launch list + one --set full capture per step kernel; the .ncu-rep comes back in
# gpurun_out/, tools/ncu_capture_json.py turns it into profiles/ncu_capture.json (run that here, after the call).
# MBOTS_GRAPHS=0: ncu serialises kernels anyway, and plain launches keep kernel names / order simple.
TAG=${1:-r2}; W=${2:-65536}
mkdir -p gpurun_out
MBOTS_GRAPHS=0 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/${TAG}_launches_${W}w.csv python bench.py --worlds $W --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu1.log 2>&1
MBOTS_GRAPHS=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"export_kernel|step_kernel|shift_kernel|scan_kernel|tile_sum|action_write" -s 40 -c 8 -o gpurun_out/${TAG}_full_${W}w -f python bench.py --worlds $W --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu2.log 2>&1
tail -2 gpurun_out/${TAG}_ncu2.log | cut -c1-200
python - <<PY
import json,subprocess
# the agents of the profiled launch: print the population of step ~8 of this workload for the JSON tool
import sys; sys.path.insert(0,'madrona-bots_b200'); sys.path.insert(0,'tests')
from madrona_bots import SimManager
import torch
sim = SimManager(0, $W, 0, 32)
print("agents_after_create", sim.num_agents(), "-- the launch captured after -s 40 is step 13 of the workload: 2,212,717 agents at 65,536 worlds (oracle, same action pool)")
PY

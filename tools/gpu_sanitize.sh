#!/bin/bash
# compute-sanitizer pass: memcheck + racecheck (+ synccheck for the tcgen05 acting kernel)
# over smoke(), a 1,024-world x 10-step rollout, a rollout with worlds above 64 agents (full export variant),
# the capacity-32 fixture (the out-of-bounds read ADVICE r1 named) and one policy_act call per layout.
# Logs land in gpurun_out/<TAG>_sanitize_*.log; tools/sanitize_summary.py condenses them for profiles/.
TAG=${1:-r2}
OUT=gpurun_out
mkdir -p $OUT
SAN=/usr/local/cuda/bin/compute-sanitizer
run() {   # name, tool, python snippet
    local name=$1 tool=$2; shift 2
    timeout 900 $SAN --tool $tool --print-limit 20 --error-exitcode 0 python tools/sanitize_cases.py "$@" \
        > $OUT/${TAG}_sanitize_${tool}_${name}.log 2>&1
    grep -E "ERROR SUMMARY|RACECHECK SUMMARY|case ok" $OUT/${TAG}_sanitize_${tool}_${name}.log | tail -3
}
for tool in memcheck racecheck; do
    run smoke $tool smoke
    run w1024 $tool w1024
    run full $tool full
    run cap32 $tool cap32
done
for tool in memcheck racecheck synccheck; do
    run policy $tool policy
done
python tools/sanitize_summary.py $OUT/${TAG}_sanitize_*.log > $OUT/${TAG}_sanitize_summary.txt
cat $OUT/${TAG}_sanitize_summary.txt

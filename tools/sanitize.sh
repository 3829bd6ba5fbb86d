#!/bin/bash
# compute-sanitizer over every kernel at small sizes (tools/sanitize_cases.py); summaries -> gpurun_out/<tag>_sanitizer_*.txt
# usage (under gpurun, from the repo root): bash tools/sanitize.sh <tag>
TAG=${1:-r2}
OUT=gpurun_out
mkdir -p $OUT
for tool in memcheck racecheck synccheck; do
  for group in ransac pipeline ukf assoc ols aug; do
    log=$OUT/${TAG}_sanitizer_${tool}_${group}.txt
    timeout 900 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_cases.py $group > $log 2>&1
    echo "$tool $group rc=$? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $log | tail -1) $(grep -c ' ok' $log) ok-lines"
  done
done

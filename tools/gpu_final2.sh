#!/bin/bash
# Closing launch list + full capture of the headline kernel at HEAD (the ncu passes of tools/gpu_round.sh only).
# usage (from the repo root, under gpurun): bash tools/gpu_final2.sh <tag>
set -u
TAG=${1:-r2ad}
OUT=gpurun_out
mkdir -p $OUT
echo "== ncu launches"
CMD="python bench.py --steps 2 --warmup 3 --scans 262144 --no-cpu --no-e2e --no-aug --no-legs --no-points"
timeout 200 $CMD > $OUT/${TAG}_plain.log 2>&1 && \
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu_launches.log 2>&1
echo "launches rc=$?"
echo "== ncu full"
CMD2="python tools/ransac_one.py c2 262144"
timeout 200 $CMD2 > $OUT/${TAG}_plain2.log 2>&1 && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:ransac_ -s 2 -c 1 -f -o $OUT/${TAG}_tile_c2 $CMD2 > $OUT/${TAG}_ncu_full.log 2>&1
echo "full rc=$?" ; cat $OUT/${TAG}_plain2.log

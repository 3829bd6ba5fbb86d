#!/bin/bash
# RANSAC kernel iteration on the GPU: parity tests of both kernels, launch-shape sweep, one ncu capture of the
# warp-specialised kernel.  usage: bash tools/gpu_ws.sh <tag> [extra ransac_one options, e.g. scorers=12 posts=8]
set -u
TAG=${1:-ws}; shift
OUT=gpurun_out
mkdir -p $OUT
echo "== pytest ransac" ; timeout 900 python -m pytest tests/test_gpu_ransac.py -x -q > $OUT/${TAG}_pytest.log 2>&1 ; echo "pytest rc=$?" ; tail -3 $OUT/${TAG}_pytest.log
echo "== sweep" ; timeout 1200 python tools/ransac_sweep.py > $OUT/${TAG}_sweep.log 2>&1 ; echo "sweep rc=$?" ; tail -3 $OUT/${TAG}_sweep.log
cp $OUT/ransac_sweep.json $OUT/${TAG}_sweep.json
CMD2="python tools/ransac_one.py c2 262144 kernel=2 $*"
timeout 300 $CMD2 > $OUT/${TAG}_plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ransac_ -s 2 -c 1 -f -o $OUT/${TAG}_ransac_c2 $CMD2 > $OUT/${TAG}_ncu_full.log 2>&1
echo "full rc=$?" ; cat $OUT/${TAG}_plain2.log

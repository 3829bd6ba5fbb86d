#!/bin/bash
# Short closing GPU session (bounded to a few minutes): parity tests, the no-allocation test alone, smoke, bench (both arms),
# then ncu captures of the two kernels changed last (OLS-flavour RANSAC, association).
# usage (from the repo root, under gpurun): bash tools/gpu_final.sh <tag>
set -u
TAG=${1:-r2ad}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv > $OUT/${TAG}_smi.txt 2>&1
echo "== pytest" ; timeout 240 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1 ; echo "pytest rc=$?" ; tail -4 $OUT/${TAG}_pytest.log
echo "== no-alloc test alone" ; timeout 120 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $OUT/${TAG}_noalloc.log 2>&1 ; echo "noalloc rc=$?" ; tail -3 $OUT/${TAG}_noalloc.log
echo "== smoke" ; timeout 120 python __graft_entry__.py --smoke > $OUT/${TAG}_smoke.log 2>&1 ; echo "smoke rc=$?" ; tail -3 $OUT/${TAG}_smoke.log
echo "== bench" ; timeout 300 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err ; echo "bench rc=$?" ; cat $OUT/${TAG}_bench.json ; tail -3 $OUT/${TAG}_bench.err
echo "== ncu ols"
CMD="python tools/ols_one.py 65536"
timeout 120 $CMD > $OUT/${TAG}_ols_plain.log 2>&1 && \
timeout 240 ncu --set full --clock-control none --import-source on -k regex:ols -s 2 -c 1 -f -o $OUT/${TAG}_ols $CMD > $OUT/${TAG}_ncu_ols.log 2>&1
echo "ols rc=$?" ; cat $OUT/${TAG}_ols_plain.log
echo "== ncu assoc"
CMD="python tools/assoc_one.py 1048576"
timeout 120 $CMD > $OUT/${TAG}_assoc_plain.log 2>&1 && \
timeout 240 ncu --set full --clock-control none --import-source on -k regex:assoc -s 2 -c 1 -f -o $OUT/${TAG}_assoc $CMD > $OUT/${TAG}_ncu_assoc.log 2>&1
echo "assoc rc=$?" ; cat $OUT/${TAG}_assoc_plain.log
echo "== bench reference" ; timeout 200 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_ref.json 2>> $OUT/${TAG}_bench.err ; cat $OUT/${TAG}_bench_ref.json

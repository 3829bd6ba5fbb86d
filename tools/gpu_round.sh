#!/bin/bash
# One GPU session: parity tests, smoke, bench, sweep, ncu launch list + full capture of the RANSAC kernel.
# usage (from the repo root, under gpurun): bash tools/gpu_round.sh <tag> [skip-sweep]
set -u
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv > $OUT/${TAG}_smi.txt 2>&1
echo "== pytest" ; timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1 ; echo "pytest rc=$?" ; tail -5 $OUT/${TAG}_pytest.log
echo "== smoke" ; timeout 300 python __graft_entry__.py --smoke > $OUT/${TAG}_smoke.log 2>&1 ; echo "smoke rc=$?" ; tail -3 $OUT/${TAG}_smoke.log
echo "== bench" ; timeout 600 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err ; echo "bench rc=$?" ; cat $OUT/${TAG}_bench.json ; tail -3 $OUT/${TAG}_bench.err
echo "== bench reference" ; timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_ref.json 2>> $OUT/${TAG}_bench.err ; cat $OUT/${TAG}_bench_ref.json
if [ "${2:-}" != "skip-sweep" ]; then
  echo "== sweep" ; timeout 900 python tools/ransac_sweep.py > $OUT/${TAG}_sweep.log 2>&1 ; echo "sweep rc=$?" ; tail -5 $OUT/${TAG}_sweep.log
fi
echo "== ncu launches"
CMD="python bench.py --steps 2 --warmup 3 --scans 262144 --no-cpu --no-e2e"
timeout 300 $CMD > $OUT/${TAG}_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu_launches.log 2>&1
echo "launches rc=$?"
echo "== ncu full"
CMD2="python tools/ransac_one.py c2 262144"
timeout 300 $CMD2 > $OUT/${TAG}_plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ransac_ -s 2 -c 1 -f -o $OUT/${TAG}_ransac_c2 $CMD2 > $OUT/${TAG}_ncu_full.log 2>&1
echo "full rc=$?" ; cat $OUT/${TAG}_plain2.log
CMD3="python tools/ukf_one.py 1048576 4"
timeout 300 $CMD3 > $OUT/${TAG}_plain3.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ukf_kernel -s 2 -c 1 -f -o $OUT/${TAG}_ukf $CMD3 > $OUT/${TAG}_ncu_full_ukf.log 2>&1
echo "full ukf rc=$?" ; cat $OUT/${TAG}_plain3.log
CMD4="python tools/ukf_one.py 1048576 8"
timeout 300 $CMD4 > $OUT/${TAG}_plain4.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ukf_kernel -s 2 -c 1 -f -o $OUT/${TAG}_ukf_L8 $CMD4 > $OUT/${TAG}_ncu_full_ukf8.log 2>&1
echo "full ukf L8 rc=$?" ; cat $OUT/${TAG}_plain4.log
timeout 300 python tools/ukf_one.py 131072 32
echo "== no-alloc test alone" ; timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -3

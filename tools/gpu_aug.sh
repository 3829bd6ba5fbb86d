#!/bin/bash
# Augmented-UKF GPU session: timing at the C4 shape, ncu launch list, one full capture of the Pbar GEMM.
# usage (under gpurun): bash tools/gpu_aug.sh <tag>
set -u
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
echo "== aug timing"
timeout 600 python tools/ukf_aug_one.py 200 1024 3 > $OUT/${TAG}_aug_time.log 2>&1 ; cat $OUT/${TAG}_aug_time.log
timeout 300 python tools/ukf_aug_one.py 200 256 3 >> $OUT/${TAG}_aug_time.log 2>&1 ; tail -1 $OUT/${TAG}_aug_time.log
timeout 300 python tools/ukf_aug_one.py 8 16384 3 >> $OUT/${TAG}_aug_time.log 2>&1 ; tail -1 $OUT/${TAG}_aug_time.log
echo "== aug launch list"
CMD="python tools/ukf_aug_one.py 200 256 1"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/${TAG}_aug_launches.csv $CMD > $OUT/${TAG}_aug_ncu_launches.log 2>&1
echo "launches rc=$?" ; python tools/aug_launch_summary.py $OUT/${TAG}_aug_launches.csv | tee $OUT/${TAG}_aug_launch_summary.txt
echo "== aug ncu full (Pzx GEMM = 27th GEMM launch)"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:aug_gemm_tma -s 26 -c 1 -f -o $OUT/${TAG}_aug_gemm $CMD > $OUT/${TAG}_aug_ncu_full.log 2>&1
echo "full rc=$?"
echo "== aug ncu full (a SOLVE GEMM late in a factorisation, the diagonal-block kernel)"
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:aug_gemm_tma_kernel<32, 64, 1, 8, 3, 1" -s 30 -c 1 -f -o $OUT/${TAG}_aug_solve $CMD > $OUT/${TAG}_aug_ncu_solve.log 2>&1
echo "solve rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:aug_chol_diag -s 20 -c 1 -f -o $OUT/${TAG}_aug_diag $CMD > $OUT/${TAG}_aug_ncu_diag.log 2>&1
echo "diag rc=$?"

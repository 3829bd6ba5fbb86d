#!/bin/bash
# One multi-GPU session (gpurun --gpus 8): peer-gather parity under torchrun, the host-copy ceiling at 1/2/4/8 ranks,
# bench.py at N = 2 and 8 (weak headline + strong object + gather check).  usage: bash tools/gpu_multi.sh <tag>
TAG=${1:-r2}
OUT=gpurun_out
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
echo "== dist tests"; timeout 600 python -m pytest tests/test_gpu_dist.py -m gpu -x -q -s > $OUT/${TAG}_dist_tests.log 2>&1; echo "rc=$?"; tail -4 $OUT/${TAG}_dist_tests.log
echo "== host copy probe"
timeout 200 python tools/probe/host_copy_probe.py > $OUT/${TAG}_host_copy_probe.jsonl 2>> $OUT/${TAG}_multi.err
for n in 2 4 8; do timeout 300 $TR --nproc-per-node $n --master-port $((29500 + n)) tools/probe/host_copy_probe.py >> $OUT/${TAG}_host_copy_probe.jsonl 2>> $OUT/${TAG}_multi.err; done
cat $OUT/${TAG}_host_copy_probe.jsonl
for n in 2 8; do
  echo "== bench N=$n"
  timeout 900 $TR --nproc-per-node $n --master-port $((29600 + n)) bench.py --gpus $n --steps 10 --warmup 3 > $OUT/${TAG}_bench_n$n.json 2>> $OUT/${TAG}_multi.err
  echo "rc=$?"; cat $OUT/${TAG}_bench_n$n.json | head -c 3000; echo
done
tail -5 $OUT/${TAG}_multi.err

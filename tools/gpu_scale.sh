#!/bin/bash
# Multi-GPU bench line: bash tools/gpu_scale.sh <N> <tag>   (run under gpurun --gpus N)
N=$1; TAG=${2:-r01i}; OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --no-cpu-baseline > $OUT/bench_${TAG}_n$N.json 2> $OUT/bench_${TAG}_n$N.err
echo "rc=$?"; cut -c1-330 $OUT/bench_${TAG}_n$N.json; tail -3 $OUT/bench_${TAG}_n$N.err

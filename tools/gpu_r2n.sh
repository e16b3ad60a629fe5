#!/bin/bash
# Round-2 multi-GPU visit: bash tools/gpu_r2n.sh <tag> <N>   (gpurun --gpus N)
# parity tests (incl. the 2-process exchange test), smoke, N=1 bench with the driver's flags, then the N-rank bench.
TAG=${1:-r02a}; N=${2:-2}
OUT=gpurun_out; mkdir -p $OUT
nvidia-smi -L | head -8
timeout 1200 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_fullsize.py > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -25 $OUT/pytest_$TAG.log
timeout 200 python __graft_entry__.py --smoke 2>&1 | tail -3
timeout 600 python bench.py --steps 20 --warmup 5 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"; cut -c1-1200 $OUT/bench_$TAG.json; tail -5 $OUT/bench_$TAG.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps 20 --warmup 5 > $OUT/bench_${TAG}_n$N.json 2> $OUT/bench_${TAG}_n$N.err
echo "benchN rc=$?"; cut -c1-1200 $OUT/bench_${TAG}_n$N.json; tail -8 $OUT/bench_${TAG}_n$N.err

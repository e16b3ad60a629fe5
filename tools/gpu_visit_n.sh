#!/bin/bash
# Round-2 multi-GPU visit: bash tools/gpu_visit_n.sh <tag> <N>   (gpurun --gpus N)
# the 2-process exchange test, then the N-rank bench with the driver's flags (and N=1 beside it for the efficiency).
TAG=${1:-r02a}; N=${2:-2}
OUT=gpurun_out; mkdir -p $OUT
nvidia-smi -L | head -8
timeout 600 python -m pytest tests/test_gpu_exchange.py tests/test_gpu_stft_ex.py -m gpu -q -x > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -12 $OUT/pytest_$TAG.log
timeout 400 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extra > $OUT/bench_${TAG}_n1.json 2> $OUT/bench_${TAG}_n1.err; echo "bench1 rc=$?"; cut -c1-400 $OUT/bench_${TAG}_n1.json
for n in 2 4 8; do
  if [ $n -le $N ]; then
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n \
      bench.py --gpus $n --steps 20 --warmup 5 --no-cpu-baseline > $OUT/bench_${TAG}_n$n.json 2> $OUT/bench_${TAG}_n$n.err
    echo "bench n=$n rc=$?"; cut -c1-400 $OUT/bench_${TAG}_n$n.json; tail -4 $OUT/bench_${TAG}_n$n.err
    python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_${TAG}_n$n.json")); print("n=$n ms", d["ms_per_step"], "p50", d.get("ms_per_step_p50"), "p95", d.get("ms_per_step_p95"), "parity_n", d.get("parity_n"), "e2e", d["e2e"]["value"])
except Exception as e: print("no json", e)
PY
  fi
done

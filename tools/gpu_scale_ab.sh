#!/bin/bash
# A/B of the NCCL stream priority at N GPUs: bash tools/gpu_scale_ab.sh <N>
N=$1; OUT=gpurun_out; mkdir -p $OUT
for hp in 0 1; do
  TAC_NCCL_HIPRI=$hp timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$hp \
    bench.py --gpus $N --no-cpu-baseline > $OUT/bench_hp${hp}_n$N.json 2> $OUT/bench_hp${hp}_n$N.err
  echo "hipri=$hp rc=$?"; python -c "
import json; d=json.load(open('$OUT/bench_hp${hp}_n$N.json')); print(d['n_gpus'], d['value'], d['ms_per_step'])"
  tail -2 $OUT/bench_hp${hp}_n$N.err
done

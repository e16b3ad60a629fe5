#!/bin/bash
# 1 -> N scaling lines with the driver's flags: bash tools/gpu_scale.sh <tag> <N> [extra bench flags]   (gpurun --gpus N)
TAG=${1:-r02e}; N=${2:-8}; shift 2
OUT=gpurun_out; mkdir -p $OUT
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extra "$@" 2> $OUT/bench_${TAG}_n1.err | grep "^{" > $OUT/bench_${TAG}_n1.json; echo "n=1 rc=$?"
for n in 2 4 8; do
  if [ $n -le $N ]; then
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n \
      bench.py --gpus $n --steps 20 --warmup 5 --no-cpu-baseline --no-extra "$@" 2> $OUT/bench_${TAG}_n$n.err | grep "^{" > $OUT/bench_${TAG}_n$n.json
    echo "n=$n rc=$?"; grep -v "^\*\*\*\|OMP_NUM\|^$" $OUT/bench_${TAG}_n$n.err | tail -3
  fi
done
python - <<PY
import json
base=None
for n in (1,2,4,8):
    try:
        d=json.loads(open("$OUT/bench_${TAG}_n%d.json"%n).read())
    except Exception as e:
        print(n, "no json", e); continue
    if n==1: base=d["ms_per_step"]
    print("n=%d ms %.4f p50 %.4f p95 %.4f max %.4f eff %.3f parity_n %s e2e_ms %.3f"%(n, d["ms_per_step"], d["ms_per_step_p50"], d["ms_per_step_p95"], d["ms_per_step_max"], base/d["ms_per_step"] if base else 0, (d.get("parity_n") or {}).get("verdict"), d["e2e"]["ms_per_step"]))
PY

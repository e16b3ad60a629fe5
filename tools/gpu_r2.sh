#!/bin/bash
# Round-2 GPU visit (1 GPU): parity tests, smoke, bench line (driver's flags), optional ncu captures.
#   gpurun --timeout 1500 -- 'bash tools/gpu_r2.sh r02a [ncu]'
TAG=${1:-r02a}; NCU=${2:-}
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_fullsize.py > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -15 $OUT/pytest_$TAG.log
timeout 200 python __graft_entry__.py --smoke 2>&1 | tail -2
timeout 600 python bench.py --steps 20 --warmup 5 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"; cut -c1-600 $OUT/bench_$TAG.json; tail -5 $OUT/bench_$TAG.err
if [ -n "$NCU" ]; then
  BCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra"
  for k in dpca_gram_tc_kernel; do
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -f -o $OUT/prof_${k}_$TAG $BCMD > $OUT/ncu_full_$k.log 2>&1
    echo "ncu full $k rc=$?"
  done
fi
ls -la $OUT | tail -8

#!/bin/bash
# Round-2 1-GPU visit: full parity suite, smoke, bench line, launch list, ncu --set full of named kernels.
#   gpurun --timeout 1500 -- 'bash tools/gpu_r2f.sh <tag> "<kernel regexes, space separated>"'
TAG=${1:-r02f}; KERNELS=${2:-dpca_gram_tc_kernel}
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_fullsize.py > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -6 $OUT/pytest_$TAG.log
timeout 200 python __graft_entry__.py --smoke 2>&1 | tail -1
timeout 600 python bench.py --steps 20 --warmup 5 2> $OUT/bench_$TAG.err | grep "^{" > $OUT/bench_$TAG.json; echo "bench rc=$?"; cut -c1-300 $OUT/bench_$TAG.json; tail -3 $OUT/bench_$TAG.err
BCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/ncu_launches_$TAG.csv $BCMD > $OUT/ncu_launches_$TAG.log 2>&1; echo "launch list rc=$?"
for k in $KERNELS; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -f -o $OUT/prof_${k}_$TAG $BCMD > $OUT/ncu_full_${k}_$TAG.log 2>&1
  echo "ncu full $k rc=$?"
done
ls -la $OUT | tail -6

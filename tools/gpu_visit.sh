#!/bin/bash
# Round-2 1-GPU visit (re-entry): full parity suite, A/B of variants/lib_*.so against the in-tree library, bench line,
# ncu --set full (+ source) of named kernels under the bench command.
#   gpurun --timeout 1500 -- 'bash tools/gpu_visit.sh <tag> "<kernel regexes>"'
TAG=${1:-r02g}; KERNELS=${2:-mrstft_stream_kernel}
OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
timeout 900 python -m pytest tests -m gpu -q -x > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/pytest_$TAG.log
( echo "=== in-tree"; timeout 300 python tools/time_legs.py 2>&1 | grep -E "fused|fwd|frontend|sums"; timeout 300 python tools/time_scales.py 2>&1 | grep -E "n_fft"
  for so in variants/lib_*.so; do
    [ -f "$so" ] || continue
    echo "=== $so"
    TAC_SPECTRAL_LIB=$PWD/$so timeout 300 python tools/time_legs.py 2>&1 | grep -E "fused|fwd|frontend|sums"
    TAC_SPECTRAL_LIB=$PWD/$so timeout 300 python tools/time_scales.py 2>&1 | grep -E "n_fft"
  done ) 2>&1 | tee $OUT/ab_$TAG.log
timeout 600 python bench.py --steps 20 --warmup 5 2> $OUT/bench_$TAG.err | grep "^{" > $OUT/bench_$TAG.json; echo "bench rc=$?"; cut -c1-300 $OUT/bench_$TAG.json; tail -3 $OUT/bench_$TAG.err
BCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra"
for k in $KERNELS; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -f -o $OUT/prof_${k}_$TAG $BCMD > $OUT/ncu_full_${k}_$TAG.log 2>&1
  echo "ncu full $k rc=$?"
done
ls -la $OUT | tail -8

#!/bin/bash
# One GPU-box visit: parity tests, a bench line, the ncu launch list of the same bench command and one
# `--set full` capture of each hot kernel.  Everything lands in gpurun_out/ (copied into profiles/ by hand).
#   gpurun --timeout 1500 -- 'bash tools/gpu_round.sh [tag]'
TAG=${1:-r01}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi_$TAG.txt 2>&1
timeout 600 python -m pytest tests -m gpu -x -q > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_$TAG.log
tail -3 $OUT/pytest_$TAG.log
timeout 600 python bench.py > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"
cat $OUT/bench_$TAG.json
BCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 $BCMD > $OUT/bench_short_$TAG.json 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches_$TAG.csv $BCMD > $OUT/ncu_launch_$TAG.log 2>&1
echo "launch list rc=$?"
for leg in fused frontend; do
  timeout 200 python tools/prof_fused.py $leg > $OUT/prof_plain_$leg.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'mrstft|stft_tiled|stft_planes' -s 1 -c 2 -f -o $OUT/prof_${leg}_$TAG python tools/prof_fused.py $leg > $OUT/ncu_full_$leg.log 2>&1
  echo "ncu full $leg rc=$?"
done
ls -la $OUT

#!/bin/bash
# Round-end GPU visit: parity tests, smoke, the bench line, the ncu launch list of the same bench command and `--set full`
# captures of the hot kernels taken UNDER the bench command (inputs larger than L2, graph replay).
#   gpurun --timeout 1800 -- 'bash tools/gpu_round2.sh r01e'
TAG=${1:-r01e}
OUT=gpurun_out
mkdir -p $OUT
timeout 700 python -m pytest tests -m gpu -x -q > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/pytest_$TAG.log
timeout 200 python __graft_entry__.py --smoke 2>&1 | tail -2
timeout 600 python bench.py > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"; cut -c1-400 $OUT/bench_$TAG.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_ref_$TAG.json 2>> $OUT/bench_$TAG.err; echo "ref rc=$?"; cut -c1-300 $OUT/bench_ref_$TAG.json
BCMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 $BCMD > $OUT/bench_short_$TAG.json 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches_$TAG.csv $BCMD > $OUT/ncu_launch_$TAG.log 2>&1
echo "launch list rc=$?"
for k in mrstft_stream_kernel stft_frontend_kernel dpca_eig_tri_kernel; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$k -s 4 -c 1 -f -o $OUT/prof_${k}_$TAG $BCMD > $OUT/ncu_full_$k.log 2>&1
  echo "ncu full $k rc=$?"
done
timeout 200 python tools/prof_istft.py > /dev/null 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:istft_kernel -s 1 -c 1 -f -o $OUT/prof_istft_$TAG python tools/prof_istft.py > $OUT/ncu_full_istft.log 2>&1
echo "ncu full istft rc=$?"
timeout 300 python tools/time_istft.py > $OUT/istft_$TAG.jsonl 2>&1; tail -3 $OUT/istft_$TAG.jsonl | cut -c1-200
timeout 300 python tools/time_planes.py > $OUT/planes_$TAG.jsonl 2>&1; head -3 $OUT/planes_$TAG.jsonl | cut -c1-200
ls -la $OUT

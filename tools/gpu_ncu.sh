#!/bin/bash
# One --set full capture of a leg: bash tools/gpu_ncu.sh <leg> <tag> [kernel-regex]
LEG=${1:-fused}; TAG=${2:-x}; RX=${3:-mrstft|stft_tiled|stft_planes}
OUT=gpurun_out; mkdir -p $OUT
timeout 200 python tools/prof_fused.py $LEG > $OUT/prof_plain_$LEG.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$RX" -s 1 -c 1 -f -o $OUT/prof_${LEG}_$TAG python tools/prof_fused.py $LEG > $OUT/ncu_full_$LEG.log 2>&1
echo "ncu full $LEG rc=$?"

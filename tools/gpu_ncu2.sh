#!/bin/bash
# bash tools/gpu_ncu2.sh <script.py> <tag> <kernel-regex>
S=$1; TAG=$2; RX=$3; OUT=gpurun_out; mkdir -p $OUT
timeout 200 python $S > $OUT/prof_plain_$TAG.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$RX" -s 1 -c 1 -f -o $OUT/prof_$TAG python $S > $OUT/ncu_full_$TAG.log 2>&1
echo "ncu rc=$?"

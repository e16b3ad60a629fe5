#!/bin/bash
# A/B of variants/lib_*.so: fused-loss leg + per-scale times
OUT=gpurun_out; mkdir -p $OUT
for so in variants/lib_*.so; do
  echo "=== $so"
  TAC_SPECTRAL_LIB=$PWD/$so timeout 300 python tools/time_legs.py 2>&1 | grep -E "fused|sums"
  TAC_SPECTRAL_LIB=$PWD/$so timeout 300 python tools/time_scales.py 2>&1 | grep -E "2048|\(1024,\)|\(512,\)"
done 2>&1 | tee $OUT/ab_time.log

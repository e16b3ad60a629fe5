#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for so in variants/lib_*.so; do
  echo "=== $so"
  TAC_SPECTRAL_LIB=$PWD/$so timeout 300 python tools/time_legs.py 2>&1 | grep -E "fused|fwd|frontend"
  TAC_SPECTRAL_LIB=$PWD/$so timeout 300 python tools/time_planes.py 2>&1 | grep -E '"cfg4"' | cut -c1-200
done 2>&1 | tee $OUT/ab_time.log

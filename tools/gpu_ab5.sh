#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_loss.py tests/test_gpu_fullsize.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python tools/check_eig.py 2>&1 | tail -12 | tee $OUT/ab_eig.log
for so in variants/lib_*.so; do
  echo "=== $so"
  TAC_SPECTRAL_LIB=$PWD/$so timeout 300 python tools/time_legs.py 2>&1 | grep -E "fused"
  TAC_SPECTRAL_LIB=$PWD/$so timeout 300 python tools/time_scales.py 2>&1 | grep -E "2048|\(1024,\)|\(512,\)"
done 2>&1 | tee $OUT/ab_time.log

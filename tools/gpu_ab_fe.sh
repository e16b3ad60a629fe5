#!/bin/bash
# A/B of front-end variants (variants/lib_fe_*.so against the in-tree library): front-end leg time on cfg 2
OUT=gpurun_out; mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_stft.py -m gpu -q -x 2>&1 | tail -3
( echo "=== default"; timeout 300 python tools/time_legs.py 2>&1 | grep -E "frontend|fused"
for so in variants/lib_*.so; do
  echo "=== $so"
  TAC_SPECTRAL_LIB=$PWD/$so timeout 300 python tools/time_legs.py 2>&1 | grep -E "frontend|fused"
done ) 2>&1 | tee $OUT/ab_fe.log

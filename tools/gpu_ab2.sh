#!/bin/bash
# Parity of the loss tests + A/B timing of variants/lib_*.so in one GPU visit.
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_loss.py -m gpu -x -q > $OUT/ab_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $OUT/ab_pytest.log
for so in variants/lib_*.so; do
  echo "=== $so"
  TAC_SPECTRAL_LIB=$PWD/$so timeout 300 python tools/time_legs.py 2>&1 | tail -7
  TAC_SPECTRAL_LIB=$PWD/$so timeout 300 python tools/time_scales.py 2>&1 | tail -7
done 2>&1 | tee $OUT/ab_time.log

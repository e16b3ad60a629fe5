#!/bin/bash
# A/B timing of kernel variants built with build.py --out=variants/lib_<name>.so (one GPU visit for all of them).
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/ab_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $OUT/ab_pytest.log
for so in variants/lib_*.so; do
  echo "=== $so"
  TAC_SPECTRAL_LIB=$PWD/$so timeout 300 python tools/time_legs.py 2>&1 | tail -7
done 2>&1 | tee $OUT/ab_time.log
timeout 300 python tools/check_eig.py 2>&1 | tee $OUT/ab_eig.log | tail -12
for so in variants/lib_*.so; do TAC_SPECTRAL_LIB=$PWD/$so timeout 300 python tools/time_scales.py 2>&1 | tail -8; done | tee $OUT/ab_scales.log

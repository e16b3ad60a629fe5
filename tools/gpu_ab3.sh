#!/bin/bash
# STFT/iSTFT/fullsize parity + A/B of K1 kernels across variants/lib_*.so
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_stft.py tests/test_gpu_istft.py tests/test_gpu_loss.py -m gpu -x -q > $OUT/ab_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $OUT/ab_pytest.log
for so in variants/lib_*.so; do
  echo "=== $so"
  TAC_SPECTRAL_LIB=$PWD/$so timeout 300 python tools/time_legs.py 2>&1 | grep -E "fused|frontend"
  TAC_SPECTRAL_LIB=$PWD/$so timeout 300 python tools/time_planes.py 2>&1 | grep -E "cfg4|cfg2" | cut -c1-260
done 2>&1 | tee $OUT/ab_time.log

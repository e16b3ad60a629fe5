#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_dpca.py tests/test_gpu_fullsize.py tests/test_gpu_latent.py -m gpu -x -q 2>&1 | tail -5
timeout 300 python tools/check_eig.py 2>&1 | tail -30 | tee $OUT/ab_eig.log
timeout 300 python tools/time_legs.py 2>&1 | tail -6

#!/bin/bash
# Full GPU test suite + leg timings (one GPU visit)
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-r01g}
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/pytest_$TAG.log
timeout 300 python tools/time_legs.py 2>&1 | tail -6 | tee $OUT/legs_$TAG.log
timeout 300 python tools/time_planes.py 2>&1 | grep -E '"cfg4"|cfg2-disc' | cut -c1-250 | tee $OUT/planes_$TAG.jsonl
timeout 300 python tools/time_istft.py 2>&1 | cut -c1-250 | tee $OUT/istft_$TAG.jsonl

#!/bin/bash
# Round-2 1-GPU visit: parity suite, A/B of variants, front-end / chain timings, DRAM bytes of the chain under ncu.
TAG=${1:-r02h}
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q -x > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/pytest_$TAG.log
( echo "=== in-tree"; timeout 300 python tools/time_legs.py 2>&1 | grep -E "fused|fwd|frontend"; timeout 300 python tools/time_scales.py 2>&1 | grep -E "n_fft"
  for so in variants/lib_*.so; do
    [ -f "$so" ] || continue
    echo "=== $so"
    TAC_SPECTRAL_LIB=$PWD/$so timeout 300 python tools/time_legs.py 2>&1 | grep -E "fused|fwd|frontend"
    TAC_SPECTRAL_LIB=$PWD/$so timeout 300 python tools/time_scales.py 2>&1 | grep -E "n_fft"
  done ) 2>&1 | tee $OUT/ab_$TAG.log
timeout 300 python tools/time_planes.py 2>&1 | tee $OUT/planes_$TAG.jsonl | cut -c1-300
timeout 300 python tools/time_chain.py 128 2>&1 | tee $OUT/chain_$TAG.jsonl
timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:"stft_planes|istft" --csv --log-file $OUT/chain_dram_$TAG.csv python tools/time_chain.py 128 once > $OUT/chain_ncu_$TAG.log 2>&1; echo "chain ncu rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2> $OUT/bench_$TAG.err | grep "^{" > $OUT/bench_$TAG.json; echo "bench rc=$?"; cut -c1-250 $OUT/bench_$TAG.json

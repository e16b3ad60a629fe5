#!/bin/bash
# compute-sanitizer over the product's kernels (SURVEY.md section 5 "race detection / sanitizers"), through the GPU tests.
#   gpurun --timeout 3000 -- 'bash tools/sanitize.sh'      -> gpurun_out/sanitizer/*.log + summary.txt
# memcheck / initcheck: every test file except the full-size ones (minutes under instrumentation) and the multi-process
# ones; torch's caching allocator is switched off so that every tensor is its own cudaMalloc (out-of-bounds accesses inside
# a cached block would otherwise go unseen).  racecheck / synccheck: the shared-memory kernels (planes, persistent planes,
# front end, loss stream / small / chunked, iSTFT, stft_ex, DPCA Gram / eig / project) at their smallest test shapes.
set -u
OUT=gpurun_out/sanitizer
mkdir -p "$OUT"
CS=/usr/local/cuda/bin/compute-sanitizer
export PYTORCH_NO_CUDA_MEMORY_CACHING=1
ALL="tests/test_gpu_stft.py tests/test_gpu_stft_ex.py tests/test_gpu_loss.py tests/test_gpu_dpca.py tests/test_gpu_istft.py tests/test_gpu_latent.py tests/test_gpu_tail.py tests/test_gpu_exchange.py tests/test_gpu_step.py tests/test_zz_no_alloc.py"
SKIP="not two_gpu and not two_devices and not long_clip and not timeout"
RACE_K="five_scales_one_launch or frontend_mag_phase or persist or forward_default_scales or (no_kinks and 0.5) or spec_losses_golden or golden_disc or sgmse_front_end or db_front_end or data_pca_golden or dpca_block_golden or istft_golden or batch_reconstruct or push_pop or fused_stats or static_scratch"
run() {  # tool, log name, pytest -k expression, files...
    local tool=$1 name=$2 kexpr=$3; shift 3
    echo "== $tool ($name)" | tee -a "$OUT/summary.txt"
    timeout 900 $CS --tool "$tool" --error-exitcode 66 --print-limit 30 \
        python -m pytest "$@" -q -m gpu -p no:cacheprovider -k "$kexpr" > "$OUT/$name.log" 2>&1
    echo "exit code $?" | tee -a "$OUT/summary.txt"
    grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|error" "$OUT/$name.log" | tail -n 6 | tee -a "$OUT/summary.txt"
}
: > "$OUT/summary.txt"
run memcheck memcheck "$SKIP" $ALL
run initcheck initcheck "$SKIP" $ALL
run synccheck synccheck "$RACE_K" $ALL
run racecheck racecheck "$RACE_K" $ALL
cat "$OUT/summary.txt"

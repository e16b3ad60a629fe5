// Chain  decoder spectrogram -> waveform -> discriminator STFT inputs  without re-reading the waveform from HBM:
// train_pkl_file_task_aware.py:290 (batch_reconstruct_waveform of the clean magnitude / phase), :333 (the waveform itself
// feeds the MSE, so it IS written), :346-353 (unsqueeze(1), disc(...)) -> msstftd.py:188-191 (loop over the scale
// discriminators) -> :151-153 (spec_transform, cat(real, imag), 'b c w t -> b c t w').
//
// The two kernels exist (istft.cu: K4, stft.cu: K1 planes); this entry point enqueues them back to back on the caller's stream
// for row groups of the batch, writes every output once and never synchronises with the host (graph-capturable).  Results are
// bit-identical to tac_istft followed by tac_stft_multi on the whole batch (same kernels, same per-row work).
//
// Group size, measured on B200 (tools/time_chain.py, 128 x [2, 1025, 600] decoder outputs -> 306 688-sample waveforms ->
// 3 scales; profiles/chain_r02i.jsonl): the waveform is 157 MB of the chain's 2.67 GB of traffic, so keeping it in the 126 MB
// L2 between the kernels can save 6 % at best, while every extra group adds two partial last waves (K4 runs one CTA per SM,
// K1 two): whole batch 1.85 ms, groups of 64 / 32 / 16 / 8 rows 2.10 / 2.29 / 2.55 / 3.23 ms.  rows_per_group = 0 therefore
// means ONE group; tac_chain_rows_per_group gives the L2-resident group size for callers that want the waveform hot for
// further consumers (the MSE of :333, a second discriminator pass) and accept the tails.
#include "tac_common.cuh"

using namespace tac;

extern "C" int64_t tac_chain_rows_per_group(int64_t rows, int64_t out_len, int64_t l2_budget_bytes) {
    if (rows < 1 || out_len < 1) return 0;
    const int64_t budget = l2_budget_bytes > 0 ? l2_budget_bytes : (40ll << 20);
    int64_t g = budget / (out_len * (int64_t)sizeof(float));
    if (g < 1) g = 1;
    if (g > rows) g = rows;
    // even groups: ceil(rows / ngroups) rows each, so that the last group is not a sliver
    const int64_t ng = (rows + g - 1) / g;
    return (rows + ng - 1) / ng;
}

extern "C" int tac_istft_disc_inputs(const float* a, const float* b, int in_mode, int64_t rows, int64_t row_stride, int64_t ld,
                                     int64_t F, int n_fft, int hop, int win, int center, int normalized, float* wave,
                                     int64_t out_len, int n_scales, const int* d_n_fft, const int* d_hop, const int* d_win,
                                     int d_normalized, void* const* planes, int64_t rows_per_group, tac_stream_t stream) {
    TAC_REQUIRE(a && wave && d_n_fft && d_hop && planes, "null argument");
    TAC_REQUIRE(rows >= 1 && out_len >= 1, "rows and out_len must be positive");
    TAC_REQUIRE(n_scales >= 1 && n_scales <= TAC_MAX_SCALES, "n_scales must be in [1, %d]", TAC_MAX_SCALES);
    TAC_REQUIRE(in_mode == TAC_IN_POLAR || in_mode == TAC_IN_PLANES || in_mode == TAC_IN_COMPLEX, "unknown in_mode %d", in_mode);
    TAC_REQUIRE(in_mode == TAC_IN_COMPLEX || b, "b is null");
    int64_t g = rows_per_group > 0 ? rows_per_group : rows;
    if (g > rows) g = rows;
    int64_t Fs[TAC_MAX_SCALES];
    for (int s = 0; s < n_scales; ++s) {
        TAC_REQUIRE(planes[s] != nullptr, "planes[%d] is null", s);
        Fs[s] = num_frames(out_len, d_n_fft[s], d_hop[s], 0);
        TAC_REQUIRE(Fs[s] >= 1, "waveform of %lld samples is shorter than n_fft=%d", (long long)out_len, d_n_fft[s]);
    }
    const int64_t in_unit = in_mode == TAC_IN_COMPLEX ? 2 : 1;          // floats per element of `a`
    for (int64_t r0 = 0; r0 < rows; r0 += g) {
        const int64_t nr = rows - r0 < g ? rows - r0 : g;
        float* w = wave + r0 * out_len;
        int rc = tac_istft(a + r0 * row_stride * in_unit, b ? b + r0 * row_stride : nullptr, in_mode, nr, row_stride, ld, F, n_fft,
                           hop, win, center, normalized, w, out_len, stream);
        if (rc) return rc;
        void* outs[TAC_MAX_SCALES];
        for (int s = 0; s < n_scales; ++s)                             // [rows, 2, F_s, K_s], one channel
            outs[s] = reinterpret_cast<float*>(planes[s]) + r0 * 2 * Fs[s] * (int64_t)(d_n_fft[s] / 2 + 1);
        rc = tac_stft_multi(w, nr, 1, out_len, n_scales, d_n_fft, d_hop, d_win, 0, TAC_PAD_ZERO, d_normalized, TAC_OUT_PLANES, 0,
                            outs, stream);
        if (rc) return rc;
    }
    return TAC_OK;
}

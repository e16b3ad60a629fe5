// Device helpers shared by the loss kernels (loss.cu: chunk/round kernels; loss_stream.cu: streaming kernels).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace tac {

// Single-MUFU approximations (<= 2 ulp; the loss tolerance is rtol 1e-4): no Newton fix-up, no slow-path branch.
__device__ __forceinline__ float fast_sqrt(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float fast_rsqrt(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float fast_lg2(float x) { float r; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
// clamp(v * 2^100, -c, c) == c * sign(v) for every v a float subtraction can produce (0 stays 0): three instructions,
// no predicates.
__device__ __forceinline__ float signed_coef(float v, float c) { return fminf(fmaxf(v * 0x1p100f, -c), c); }
constexpr float kTinyMag2 = 1e-30f;     // |Y|^2 below this is treated as |Y| == 0 (zero gradient, like torch.abs at 0)
// |Z| from |Z|^2 -- the SAME formula for X and Y, so that identical inputs give a loss of exactly 0
__device__ __forceinline__ float fast_mag(float m2) { return m2 >= kTinyMag2 ? m2 * fast_rsqrt(m2) : 0.0f; }

// ---- per-bin loss terms from the SQUARED magnitudes (round 2: 24 instructions per bin against 33) ------------------
// With a = |X|^2, b = |Y|^2 (same arithmetic for both, so X == Y gives exactly 0 everywhere below):
//   |X| = sqrt.approx(a), |Y| = sqrt.approx(b)                              one MUFU each, no guard (sqrt(0) = 0)
//   |ln max(|Y|, eps) - ln max(|X|, eps)| = (ln 2 / 2) |lg2 max(b, eps^2) - lg2 max(a, eps^2)|   (kLogHalf folds the factor)
//   sign(|Y| - |X|) = sign(b - a), and wherever the log term has a gradient (|Y| >= eps) its sign is sign(b - a) too
//   (sqrt and log o max are monotone), so ONE sign carries both terms:
//   dL/dY[k] = Y[k] * sign(b - a) * (c_l1 + c_log [|Y| >= eps] / |Y|) / |Y|
// min.xorsign.abs(u, g) = sign(u) sign(g) min(|u|, |g|): with u = (b - a) 2^100 (inf is fine) and g >= 0 that is
// sign(b - a) g, and exactly 0 when a == b.  1 / |Y| = rsqrt(max(b, tiny)): for b == 0 the factor Y[k] is 0 itself.
#ifndef TAC_EPI_MUFU
#define TAC_EPI_MUFU 5
#endif
constexpr float kLogHalf = 0.34657359027997265471f;     // ln 2 / 2: scale of the accumulated |lg2 b~ - lg2 a~|
__device__ __forceinline__ float min_xorsign_abs(float a, float b) { float r; asm("min.xorsign.abs.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float eps_squared(float eps) { return fmaxf(eps * eps, 1.17549435e-38f); }
template <bool WITH_SUMS>
__device__ __forceinline__ void loss_bin_value(float m2x, float m2y, float eps2, float& l1, float& ll) {
    if (WITH_SUMS) {
#if TAC_EPI_MUFU == 4       // |Z| = |Z|^2 rsqrt(max(|Z|^2, tiny)): shares the reciprocal root with the gradient (4 MUFU per bin, 3 more ALU)
        l1 += fabsf(m2y * fast_rsqrt(fmaxf(m2y, kTinyMag2)) - m2x * fast_rsqrt(fmaxf(m2x, kTinyMag2)));
#else
        l1 += fabsf(fast_sqrt(m2y) - fast_sqrt(m2x));
#endif
        ll += fabsf(fast_lg2(fmaxf(m2y, eps2)) - fast_lg2(fmaxf(m2x, eps2)));
    }
}
// returns g with dL/dY[k] = g Y[k]
template <bool WITH_SUMS>
__device__ __forceinline__ float loss_bin_grad(float m2x, float m2y, float eps2, float c_l1, float c_log, float& l1, float& ll) {
    loss_bin_value<WITH_SUMS>(m2x, m2y, eps2, l1, ll);
    const float inv = fast_rsqrt(fmaxf(m2y, kTinyMag2));
    const float invm = m2y >= eps2 ? inv : 0.0f;
    const float g = fmaf(c_log, invm, c_l1) * inv;
    return min_xorsign_abs((m2y - m2x) * 0x1p100f, g);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}


// Streaming fused loss (loss_stream.cu): applicable when every scale has hop == n_fft / 4 and T is even.
bool stream_applicable(int64_t T, int n_scales, const int* n_fft, const int* hop, const int* win);
size_t stream_workspace_bytes(int64_t rows, int64_t T, int n_scales, const int* n_fft, const int* hop);
int launch_stream(const float* x, const float* y, int64_t rows, int64_t T, int n_scales, const int* n_fft, const int* hop,
                  const int* win, float eps, const float* coef, double* sums, float* grad_y, void* workspace,
                  size_t workspace_bytes, cudaStream_t st);

}  // namespace tac

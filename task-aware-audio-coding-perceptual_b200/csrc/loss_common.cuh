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

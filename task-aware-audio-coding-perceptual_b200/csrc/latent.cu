// Latent regulariser next to the DPCA block (second half): the pairwise cosine loss
//   cos_loss = mean_b sum_{i<j} cos(z_i[b], z_j[b])
// of pkl_file_models.py:390-391 (two views), :761-763 and pkl_file_model_updated.py:429-430, 513-514 (four views:
// six torch.nn.CosineSimilarity() calls, a sum and a mean -- ~20 small kernels that re-read every view three times).
// One warp owns a batch row: it reads the row of every view once, keeps the nv norms and nv(nv-1)/2 dot products in
// registers and folds the pair sum; CTA partials are float64 and are added in a fixed order (bit-stable).  The backward
// kernel recomputes the same scalars and writes d/dz for every view in the same pass.
// Algorithmic bytes: forward 4 * nv * B * d; backward 8 * nv * B * d.
#include "tac_common.cuh"

namespace tac {

constexpr int kCosMaxViews = 8;
constexpr int kCosWarps = 8;

struct CosPtrs {
    const float* z[kCosMaxViews];
    float* g[kCosMaxViews];
};

__device__ __forceinline__ float cos_warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// nrm[i] = |z_i|^2, dot[p] = <z_i, z_j> for the pairs in (0,1), (0,2), ..., (nv-2, nv-1) order -- all lanes hold the totals
template <int NV>
__device__ __forceinline__ void cos_row_scalars(const CosPtrs& p, long long row, int d, int lane, float (&nrm)[NV],
                                                float (&dot)[NV * (NV - 1) / 2 + 1]) {
#pragma unroll
    for (int i = 0; i < NV; ++i) nrm[i] = 0.0f;
#pragma unroll
    for (int q = 0; q < NV * (NV - 1) / 2 + 1; ++q) dot[q] = 0.0f;
    for (int c = lane; c < d; c += 32) {
        float v[NV];
#pragma unroll
        for (int i = 0; i < NV; ++i) v[i] = __ldg(p.z[i] + row * d + c);
        int q = 0;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            nrm[i] = fmaf(v[i], v[i], nrm[i]);
#pragma unroll
            for (int j = i + 1; j < NV; ++j) { dot[q] = fmaf(v[i], v[j], dot[q]); ++q; }
        }
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) nrm[i] = cos_warp_sum(nrm[i]);
#pragma unroll
    for (int q = 0; q < NV * (NV - 1) / 2; ++q) dot[q] = cos_warp_sum(dot[q]);
}

// torch.nn.CosineSimilarity(dim=1, eps): x.y / (max(|x|, eps) * max(|y|, eps))
template <int NV>
__global__ void __launch_bounds__(kCosWarps * 32) cosine_fwd_kernel(const CosPtrs p, long long B, int d, float eps,
                                                                    double* __restrict__ partial) {
    __shared__ double s_part[kCosWarps];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double acc = 0.0;
    for (long long row = (long long)blockIdx.x * kCosWarps + warp; row < B; row += (long long)gridDim.x * kCosWarps) {
        float nrm[NV], dot[NV * (NV - 1) / 2 + 1];
        cos_row_scalars<NV>(p, row, d, lane, nrm, dot);
        float inv[NV];
#pragma unroll
        for (int i = 0; i < NV; ++i) inv[i] = 1.0f / fmaxf(sqrtf(nrm[i]), eps);
        float s = 0.0f;
        int q = 0;
#pragma unroll
        for (int i = 0; i < NV; ++i)
#pragma unroll
            for (int j = i + 1; j < NV; ++j) { s += dot[q] * inv[i] * inv[j]; ++q; }
        acc += (double)s;
    }
    if (lane == 0) s_part[warp] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double v = 0.0;
        for (int w = 0; w < kCosWarps; ++w) v += s_part[w];
        partial[blockIdx.x] = v;
    }
}

__global__ void __launch_bounds__(256) cosine_final_kernel(const double* __restrict__ partial, int n, double* __restrict__ out) {
    __shared__ double sh[256];
    double v = 0.0;
    for (int i = threadIdx.x; i < n; i += 256) v += partial[i];
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[0] = sh[0];
}

// d/dz_i[b] of scale * sum_{i<j} cos_ij:  scale * sum_{j != i} ( z_j / (n_i n_j) - cos_ij z_i / n_i^2 ), n = max(|z|, eps)
// (a clamped norm is a constant, so its second term vanishes -- what autograd gives for the clamp)
template <int NV>
__global__ void __launch_bounds__(kCosWarps * 32) cosine_bwd_kernel(const CosPtrs p, long long B, int d, float eps,
                                                                    const float* __restrict__ gscale, float scale) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float sc = scale * (gscale ? __ldg(gscale) : 1.0f);
    for (long long row = (long long)blockIdx.x * kCosWarps + warp; row < B; row += (long long)gridDim.x * kCosWarps) {
        float nrm[NV], dot[NV * (NV - 1) / 2 + 1];
        cos_row_scalars<NV>(p, row, d, lane, nrm, dot);
        float inv[NV], self[NV];       // self[i] = - sum_j cos_ij / n_i^2 (0 where the norm is clamped)
        float cross[NV][NV];           // cross[i][j] = 1 / (n_i n_j)
#pragma unroll
        for (int i = 0; i < NV; ++i) { inv[i] = 1.0f / fmaxf(sqrtf(nrm[i]), eps); self[i] = 0.0f; }
        int q = 0;
#pragma unroll
        for (int i = 0; i < NV; ++i)
#pragma unroll
            for (int j = i + 1; j < NV; ++j) {
                const float w = inv[i] * inv[j];
                cross[i][j] = cross[j][i] = w;
                const float c = dot[q] * w;
                self[i] -= c; self[j] -= c;
                ++q;
            }
#pragma unroll
        for (int i = 0; i < NV; ++i) self[i] = sqrtf(nrm[i]) >= eps ? self[i] * inv[i] * inv[i] : 0.0f;
        for (int c = lane; c < d; c += 32) {
            float v[NV];
#pragma unroll
            for (int i = 0; i < NV; ++i) v[i] = __ldg(p.z[i] + row * d + c);
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                float g = self[i] * v[i];
#pragma unroll
                for (int j = 0; j < NV; ++j)
                    if (j != i) g = fmaf(cross[i][j], v[j], g);
                p.g[i][row * d + c] = sc * g;
            }
        }
    }
}

static int cos_grid(long long B) {
    long long g = (B + kCosWarps - 1) / kCosWarps;
    if (g > 4 * kNumSMs) g = 4 * kNumSMs;
    return (int)(g < 1 ? 1 : g);
}

}  // namespace tac

using namespace tac;

extern "C" size_t tac_cosine_pairs_workspace_bytes(int64_t B) { return (size_t)cos_grid(B) * sizeof(double); }

extern "C" int tac_cosine_pairs_fwd(const float* const* z, int nv, int64_t B, int d, float eps, double* sum, void* workspace,
                                    size_t workspace_bytes, tac_stream_t stream) {
    TAC_REQUIRE(z && sum, "null argument");
    TAC_REQUIRE(nv >= 2 && nv <= kCosMaxViews && d >= 1 && B >= 1, "bad nv / d / B");
    TAC_REQUIRE(nv == 2 || nv == 3 || nv == 4, "pairwise cosine loss is built for 2, 3 or 4 views (got %d)", nv);
    const size_t need = tac_cosine_pairs_workspace_bytes(B);
    if (!workspace || workspace_bytes < need) return set_error(TAC_EWORKSPACE, "workspace needs %zu bytes, got %zu", need, workspace_bytes);
    CosPtrs p{};
    for (int i = 0; i < nv; ++i) { TAC_REQUIRE(z[i], "z[%d] is null", i); p.z[i] = z[i]; }
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    double* partial = reinterpret_cast<double*>(workspace);
    const int grid = cos_grid(B);
    switch (nv) {
        case 2: cosine_fwd_kernel<2><<<grid, kCosWarps * 32, 0, st>>>(p, B, d, eps, partial); break;
        case 3: cosine_fwd_kernel<3><<<grid, kCosWarps * 32, 0, st>>>(p, B, d, eps, partial); break;
        default: cosine_fwd_kernel<4><<<grid, kCosWarps * 32, 0, st>>>(p, B, d, eps, partial); break;
    }
    TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    cosine_final_kernel<<<1, 256, 0, st>>>(partial, grid, sum);
    TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

extern "C" int tac_cosine_pairs_bwd(const float* const* z, int nv, int64_t B, int d, float eps, const float* gscale, float scale,
                                    float* const* grad_z, tac_stream_t stream) {
    TAC_REQUIRE(z && grad_z, "null argument");
    TAC_REQUIRE(nv == 2 || nv == 3 || nv == 4, "pairwise cosine loss is built for 2, 3 or 4 views (got %d)", nv);
    TAC_REQUIRE(d >= 1 && B >= 1, "bad d / B");
    CosPtrs p{};
    for (int i = 0; i < nv; ++i) { TAC_REQUIRE(z[i] && grad_z[i], "z[%d] / grad_z[%d] is null", i, i); p.z[i] = z[i]; p.g[i] = grad_z[i]; }
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const int grid = cos_grid(B);
    switch (nv) {
        case 2: cosine_bwd_kernel<2><<<grid, kCosWarps * 32, 0, st>>>(p, B, d, eps, gscale, scale); break;
        case 3: cosine_bwd_kernel<3><<<grid, kCosWarps * 32, 0, st>>>(p, B, d, eps, gscale, scale); break;
        default: cosine_bwd_kernel<4><<<grid, kCosWarps * 32, 0, st>>>(p, B, d, eps, gscale, scale); break;
    }
    TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

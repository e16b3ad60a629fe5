// Two small streaming kernels either side of the spectral path.
//
// (1) The bottleneck that is live in the SpectralRes models -- pkl_file_models.py:1108-1127 (E2D2),
//     :692-718 (E4D1), :335-354 (E2D1): every view's latent z_i [B, d] is split at column `split` into a private and a
//     shared half, the shared halves of a view PAIR (1,2), (3,4) are mixed with a per-row alpha ~ U(0,1)
//     (random_bottle_neck), both halves are truncated to `trunc` columns and concatenated.  The reference spends ~20
//     slice / mul / add / cat kernels on it; here one thread writes one output element of
//         out [B, nv * (tp + ts)],   tp = min(trunc, min(split, d)),   ts = min(trunc, max(d - split, 0)),
//     view i occupying columns [i (tp + ts), (i + 1)(tp + ts)) -- z_sample of :348 / :711, or z1_sample | z2_sample of
//     :1122-1123 as two column ranges of the same buffer.  The backward kernel is its adjoint.
//
// (2) The loss tail of task_aware() (row a5-i) -- train_pkl_file_task_aware.py:307-364:
//       * F.interpolate(shorter.unsqueeze(1), size=longer, mode='linear', align_corners=False) followed by
//         F.mse_loss (:311-333): interp_mse_kernel never materialises the resampled waveform;
//       * sum over 15 feature-map pairs of F.l1_loss, plus F.mse_loss(y_enh, 1) + F.mse_loss(y_gt, 0) over the 3 logit
//         pairs (:352-361): ONE multi-tensor reduction over a segment table instead of ~60 launches.
//     Partials are float64 per CTA, added in a fixed order by a single-CTA kernel (bit-stable).
#include "tac_common.cuh"

namespace tac {

constexpr int kMixMaxViews = 8;

struct MixArgs {
    const float* z[kMixMaxViews];
    const float* alpha[kMixMaxViews / 2];      // per pair, [B] (null: no mixing for that pair)
    float* g[kMixMaxViews];                    // backward: gradient w.r.t. z_i [B, d] (fully written, zeros outside the kept columns)
    long long B;
    int nv, d, split, tp, ts;
};

__global__ void __launch_bounds__(256) bottleneck_mix_fwd_kernel(const MixArgs a, float* __restrict__ out) {
    const int w = a.tp + a.ts;
    const long long total = a.B * a.nv * w;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
        const long long b = i / ((long long)a.nv * w);
        const int r = (int)(i - b * a.nv * w);
        const int v = r / w, c = r - v * w;
        float val;
        if (c < a.tp) {
            val = __ldg(a.z[v] + b * a.d + c);
        } else {
            const int cs = a.split + (c - a.tp);
            const float* al = a.alpha[v >> 1];
            const int v0 = v & ~1, v1 = v0 + 1;
            if (al != nullptr && v1 < a.nv) {
                const float t = __ldg(al + b);
                // :346 / :700-701 / :1117 -- two rounded products and a rounded sum like the reference's three torch kernels
                // (no FMA contraction): bit-equal results
                val = __fadd_rn(__fmul_rn(t, __ldg(a.z[v0] + b * a.d + cs)), __fmul_rn(__fsub_rn(1.0f, t), __ldg(a.z[v1] + b * a.d + cs)));
            } else {
                val = __ldg(a.z[v] + b * a.d + cs);
            }
        }
        out[i] = val;
    }
}

// grad z_i [B, d] from grad out [B, nv * w]
__global__ void __launch_bounds__(256) bottleneck_mix_bwd_kernel(const MixArgs a, const float* __restrict__ gout) {
    const int w = a.tp + a.ts;
    const long long total = a.B * a.nv * a.d;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
        const long long b = i / ((long long)a.nv * a.d);
        const int r = (int)(i - b * a.nv * a.d);
        const int v = r / a.d, c = r - v * a.d;
        const float* grow = gout + b * a.nv * w;
        float val = 0.0f;
        if (c < a.tp) val = grow[v * w + c];
        const int cs = c - a.split;
        if (cs >= 0 && cs < a.ts) {
            const float* al = a.alpha[v >> 1];
            const int v0 = v & ~1, v1 = v0 + 1;
            if (al != nullptr && v1 < a.nv) {
                const float t = __ldg(al + b);
                const float gs = grow[v0 * w + a.tp + cs] + grow[v1 * w + a.tp + cs];      // both views of the pair carry the mix
                val += (v == v0 ? t : 1.0f - t) * gs;
            } else {
                val += grow[v * w + a.tp + cs];
            }
        }
        a.g[v][b * a.d + c] = val;
    }
}

// ------------------------------------------------------------------------------------------ loss tail
constexpr int kRedChunk = 8192;          // elements per CTA
constexpr int kRedMaxSeg = 48;

struct RedSeg { const float* a; const float* b; long long n; float c; int mode; };   // mode 0: |a-b|, 1: (a-b)^2, 2: (a-c)^2
struct RedArgs {
    RedSeg seg[kRedMaxSeg];
    int first_chunk[kRedMaxSeg + 1];
    int nseg;
};

__device__ __forceinline__ double block_sum_to_thread0(double v, double* sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
    if (threadIdx.x == 0)
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += sh[w];
    return t;
}

__global__ void __launch_bounds__(256) multi_reduce_kernel(const RedArgs a, double* __restrict__ partial) {
    __shared__ double sh[8];
    int s = 0;
    while ((int)blockIdx.x >= a.first_chunk[s + 1]) ++s;
    const RedSeg& g = a.seg[s];
    const long long lo = (long long)(blockIdx.x - a.first_chunk[s]) * kRedChunk;
    const long long hi = lo + kRedChunk < g.n ? lo + kRedChunk : g.n;
    float acc = 0.0f;
    for (long long i = lo + threadIdx.x; i < hi; i += 256) {
        const float d = __ldg(g.a + i) - (g.mode == 2 ? g.c : __ldg(g.b + i));
        acc += g.mode == 0 ? fabsf(d) : d * d;
    }
    const double t = block_sum_to_thread0((double)acc, sh);
    if (threadIdx.x == 0) partial[blockIdx.x] = t;
}

// out[0] = sum_s mean_s, out[1 + s] = mean_s   (fixed order)
__global__ void __launch_bounds__(256) multi_reduce_final_kernel(const RedArgs a, const double* __restrict__ partial, double* __restrict__ out) {
    __shared__ double sh[8];
    __shared__ double s_total;
    if (threadIdx.x == 0) s_total = 0.0;
    __syncthreads();
    for (int s = 0; s < a.nseg; ++s) {
        double v = 0.0;
        for (int c = a.first_chunk[s] + threadIdx.x; c < a.first_chunk[s + 1]; c += 256) v += partial[c];
        const double t = block_sum_to_thread0(v, sh);
        if (threadIdx.x == 0) {
            const double m = t / (double)a.seg[s].n;
            out[1 + s] = m;
            s_total += m;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) out[0] = s_total;
}

// sum over rows and i < Ll of (long[b][i] - lerp(short[b], i))^2 with torch's align_corners=False source index
//   src = max((i + 0.5) * Ls / Ll - 0.5, 0), evaluated exactly in integers: src = max((2 i + 1) Ls - Ll, 0) / (2 Ll).
__global__ void __launch_bounds__(256) interp_mse_kernel(const float* __restrict__ lng, long long Ll, const float* __restrict__ sht,
                                                        long long Ls, long long B, double* __restrict__ partial) {
    __shared__ double sh[8];
    const long long total = B * Ll;
    float acc = 0.0f;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
        const long long b = i / Ll, j = i - b * Ll;
        long long num = (2 * j + 1) * Ls - Ll;
        num = num < 0 ? 0 : num;
        const long long den = 2 * Ll;
        const long long i0 = num / den;
        const long long i1 = i0 + 1 < Ls ? i0 + 1 : Ls - 1;
        const float lam = (float)((double)(num - i0 * den) / (double)den);
        const float* sr = sht + b * Ls;
        const float v = (1.0f - lam) * __ldg(sr + i0) + lam * __ldg(sr + i1);
        const float d = __ldg(lng + i) - v;
        acc = fmaf(d, d, acc);
    }
    const double t = block_sum_to_thread0((double)acc, sh);
    if (threadIdx.x == 0) partial[blockIdx.x] = t;
}

__global__ void __launch_bounds__(256) sum_partials_kernel(const double* __restrict__ partial, int n, double scale, double* __restrict__ out) {
    __shared__ double sh[8];
    double v = 0.0;
    for (int c = threadIdx.x; c < n; c += 256) v += partial[c];
    const double t = block_sum_to_thread0(v, sh);
    if (threadIdx.x == 0) out[0] = t * scale;
}

constexpr int kInterpGrid = 592;          // 4 CTAs per SM

}  // namespace tac

using namespace tac;

static int mix_args(MixArgs* a, const float* const* z, int nv, int64_t B, int d, int split, int trunc, const float* const* alpha) {
    TAC_REQUIRE(z != nullptr, "null argument");
    TAC_REQUIRE(nv >= 1 && nv <= kMixMaxViews, "nv must be in [1, %d]", kMixMaxViews);
    TAC_REQUIRE(B >= 1 && d >= 1 && split >= 0 && trunc >= 0, "bad shape B=%lld d=%d split=%d trunc=%d", (long long)B, d, split, trunc);
    memset(a, 0, sizeof(*a));
    for (int i = 0; i < nv; ++i) {
        TAC_REQUIRE(z[i] != nullptr, "z[%d] is null", i);
        a->z[i] = z[i];
    }
    for (int p = 0; p < (nv + 1) / 2; ++p) a->alpha[p] = alpha ? alpha[p] : nullptr;
    const int sp = split < d ? split : d;
    a->B = B; a->nv = nv; a->d = d; a->split = split;
    a->tp = trunc < sp ? trunc : sp;
    a->ts = trunc < d - sp ? trunc : d - sp;
    return TAC_OK;
}

extern "C" int tac_bottleneck_width(int d, int split, int trunc) {
    const int sp = split < d ? split : d;
    return (trunc < sp ? trunc : sp) + (trunc < d - sp ? trunc : d - sp);
}

extern "C" int tac_bottleneck_mix(const float* const* z, int nv, int64_t B, int d, int split, int trunc, const float* const* alpha,
                                  float* out, tac_stream_t stream) {
    MixArgs a;
    int rc = mix_args(&a, z, nv, B, d, split, trunc, alpha);
    if (rc) return rc;
    TAC_REQUIRE(out != nullptr, "null argument");
    const long long total = (long long)B * nv * (a.tp + a.ts);
    if (total == 0) return TAC_OK;
    long long grid = (total + 255) / 256;
    grid = grid > 148 * 8 ? 148 * 8 : grid;
    bottleneck_mix_fwd_kernel<<<(unsigned)grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a, out); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

extern "C" int tac_bottleneck_mix_bwd(const float* grad_out, int nv, int64_t B, int d, int split, int trunc, const float* const* alpha,
                                      float* const* grad_z, tac_stream_t stream) {
    MixArgs a;
    TAC_REQUIRE(grad_out && grad_z, "null argument");
    int rc = mix_args(&a, reinterpret_cast<const float* const*>(grad_z), nv, B, d, split, trunc, alpha);
    if (rc) return rc;
    for (int i = 0; i < nv; ++i) { a.g[i] = grad_z[i]; a.z[i] = nullptr; }
    const long long total = (long long)B * nv * d;
    long long grid = (total + 255) / 256;
    grid = grid > 148 * 8 ? 148 * 8 : grid;
    bottleneck_mix_bwd_kernel<<<(unsigned)grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a, grad_out); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

static int red_args(RedArgs* r, int nseg, const float* const* a, const float* const* b, const int64_t* n, const int* mode, const float* c) {
    TAC_REQUIRE(a && n && mode, "null argument");
    TAC_REQUIRE(nseg >= 1 && nseg <= kRedMaxSeg, "nseg must be in [1, %d]", kRedMaxSeg);
    memset(r, 0, sizeof(*r));
    long long chunks = 0;
    for (int s = 0; s < nseg; ++s) {
        TAC_REQUIRE(a[s] != nullptr && n[s] >= 1, "segment %d: null pointer or empty", s);
        TAC_REQUIRE(mode[s] >= 0 && mode[s] <= 2, "segment %d: unknown mode %d", s, mode[s]);
        TAC_REQUIRE(mode[s] == 2 || (b && b[s] != nullptr), "segment %d: second operand is null", s);
        r->seg[s].a = a[s]; r->seg[s].b = (b && mode[s] != 2) ? b[s] : nullptr; r->seg[s].n = n[s];
        r->seg[s].c = c ? c[s] : 0.0f; r->seg[s].mode = mode[s];
        r->first_chunk[s] = (int)chunks;
        chunks += (n[s] + kRedChunk - 1) / kRedChunk;
        TAC_REQUIRE(chunks < (1ll << 30), "too many elements");
    }
    r->first_chunk[nseg] = (int)chunks;
    r->nseg = nseg;
    return TAC_OK;
}

extern "C" size_t tac_multi_reduce_workspace_bytes(int nseg, const int64_t* n) {
    size_t chunks = 0;
    for (int s = 0; s < nseg; ++s) chunks += (size_t)((n[s] + kRedChunk - 1) / kRedChunk);
    return chunks * sizeof(double) + 16;
}

extern "C" int tac_multi_reduce(int nseg, const float* const* a, const float* const* b, const int64_t* n, const int* mode, const float* c,
                                double* out, void* workspace, size_t workspace_bytes, tac_stream_t stream) {
    RedArgs r;
    int rc = red_args(&r, nseg, a, b, n, mode, c);
    if (rc) return rc;
    TAC_REQUIRE(out && workspace, "null argument");
    const int chunks = r.first_chunk[nseg];
    if (workspace_bytes < (size_t)chunks * sizeof(double))
        return set_error(TAC_EWORKSPACE, "workspace needs %zu bytes", (size_t)chunks * sizeof(double));
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    double* partial = reinterpret_cast<double*>(workspace);
    multi_reduce_kernel<<<(unsigned)chunks, 256, 0, st>>>(r, partial); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    multi_reduce_final_kernel<<<1, 256, 0, st>>>(r, partial, out); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

extern "C" size_t tac_interp_mse_workspace_bytes(void) { return (size_t)kInterpGrid * sizeof(double); }

extern "C" int tac_interp_mse(const float* a, int64_t La, const float* b, int64_t Lb, int64_t B, double* out, void* workspace,
                              size_t workspace_bytes, tac_stream_t stream) {
    TAC_REQUIRE(a && b && out && workspace, "null argument");
    TAC_REQUIRE(B >= 1 && La >= 1 && Lb >= 1 && La < (1ll << 30) && Lb < (1ll << 30), "bad shape B=%lld La=%lld Lb=%lld", (long long)B, (long long)La, (long long)Lb);
    if (workspace_bytes < tac_interp_mse_workspace_bytes())
        return set_error(TAC_EWORKSPACE, "workspace needs %zu bytes", tac_interp_mse_workspace_bytes());
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    double* partial = reinterpret_cast<double*>(workspace);
    const float* lng = La >= Lb ? a : b;
    const float* sht = La >= Lb ? b : a;
    const long long Ll = La >= Lb ? La : Lb, Ls = La >= Lb ? Lb : La;
    long long grid = (B * Ll + 255) / 256;
    grid = grid > kInterpGrid ? kInterpGrid : grid;
    interp_mse_kernel<<<(unsigned)grid, 256, 0, st>>>(lng, Ll, sht, Ls, B, partial); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    sum_partials_kernel<<<1, 256, 0, st>>>(partial, (int)grid, 1.0 / ((double)B * (double)Ll), out); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

// K2 -- fused multi-resolution STFT loss (forward and backward) and K2' -- the reference's
// spectral MSE / SNR reductions.  No spectrogram ever reaches HBM: every warp transforms a frame
// of x and the same frame of y in registers, compares magnitudes and folds the two per-scale sums
// with warp shuffles -> shared memory -> one float64 partial per CTA -> a fixed-order final
// reduction (bit-stable run to run).  The backward recomputes both transforms, forms dL/dY in the
// frequency domain, runs the adjoint transform (fft_core.cuh: warp_fft_adj) and overlap-adds the
// windowed frames into a shared-memory accumulator in conflict-free "rounds" (frames f, f + R,
// f + 2R ... with R = ceil(n/hop) never overlap), so the result is deterministic too.
#include "tac_common.cuh"
#include "loss_common.cuh"
#include "fft_core_pk.cuh"

namespace tac {

struct LossScale {
    int n, hop, F, K, R;     // R = ceil(n / hop): frames f and f + R do not overlap
    FftTables t;
};

struct LossArgs {
    const float* x;
    const float* y;
    long long T;
    int nscales, chunk, nchunks, tile_len, vec_ok;
    float eps;
    double* partial;          // fwd: [grid][2*nscales]
    const float* coef;        // bwd: [2*nscales] device
    float* grad;              // bwd: [rows][T]
    LossScale sc[TAC_MAX_SCALES];
};

__device__ __forceinline__ int nfr_of(int n) { return n >= 2048 ? 1 : (n >= 512 ? 2 : (n >= 128 ? 4 : 8)); }

// ------------------------------------------------------------------------------------------ forward
constexpr int kFwdWarps = 6;
constexpr int kFwdThreads = kFwdWarps * 32;

// tile[j] = x[g0 + j] (zero behind the row end), g0 >= 0: float4 copies for the aligned in-range part even when the chunk
// hangs over the row end (the last chunk of every row), scalar for the <= 3 samples behind it.
__device__ __forceinline__ void stage_plain(float* tile, const float* __restrict__ xr, long long g0, int len,
                                            long long T, bool vec_ok) {
    const long long hi_ll = T - g0;
    const int hi = (int)(hi_ll < 0 ? 0 : (hi_ll < len ? hi_ll : len));      // in-range samples: j in [0, hi)
    const int jb = (vec_ok && (g0 & 3) == 0) ? (hi & ~3) : 0;
    if (jb > 0) {
        const float4* src = reinterpret_cast<const float4*>(xr + g0);
        float4* dst = reinterpret_cast<float4*>(tile);
        for (int i = threadIdx.x; i < (jb >> 2); i += blockDim.x) dst[i] = __ldg(src + i);
    }
    for (int j = jb + threadIdx.x; j < len; j += blockDim.x) tile[j] = (j < hi) ? __ldg(xr + g0 + j) : 0.0f;
}

template <class P>
__device__ __forceinline__ void loss_fwd_task(const LossScale& sc, const float* tx, const float* ty, float2* xbuf,
                                              int lane, int f_first, int f_end, int origin, float eps,
                                              float& acc_l1, float& acc_log) {
    cf z[P::BQ];
    float mx[P::BQ + 1];
    const float2* __restrict__ win2 = sc.t.win2;
    const int hop = sc.hop;
    const float* cur = tx;
    auto load = [&](auto, int fr, int m, auto) -> cf {
        int f = f_first + fr;
        f = f < f_end ? f : f_first;
        return mul2(*reinterpret_cast<const cf*>(cur + (f * hop - origin) + 2 * m), ldg_cf(win2 + m));   // plain product
    };
    warp_fft_fwd_pk<P>(z, xbuf, sc.t.tw, lane, load);
    warp_fft_post_pk<P>(z, sc.t.pt, lane, [&](int, int, cf v, int slot) {
        const float vr = lo(v), vi = hi(v);
        mx[slot] = fmaf(vr, vr, vi * vi);                               // |X|^2
    });
    cur = ty;
    warp_fft_fwd_pk<P>(z, xbuf, sc.t.tw, lane, load);
    const bool valid = (f_first + lane / P::A) < f_end;
    const float eps2 = eps_squared(eps);
    float l1 = 0.0f, ll = 0.0f;
    warp_fft_post_pk<P>(z, sc.t.pt, lane, [&](int, int, cf v, int slot) {
        const float vr = lo(v), vi = hi(v);
        loss_bin_value<true>(mx[slot], fmaf(vr, vr, vi * vi), eps2, l1, ll);   // same arithmetic as the fused kernels; exact 0 when X == Y
    });
    if (valid) { acc_l1 += l1; acc_log += ll * kLogHalf; }
}

__global__ void __launch_bounds__(kFwdThreads, 2) mrstft_fwd_kernel(const LossArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* tx = reinterpret_cast<float*>(smem_raw);
    float* ty = tx + a.tile_len;
    float2* xbuf_all = reinterpret_cast<float2*>(ty + a.tile_len);
    __shared__ int s_task_begin[TAC_MAX_SCALES + 1];
    __shared__ int s_f_lo[TAC_MAX_SCALES], s_f_hi[TAC_MAX_SCALES];
    __shared__ double s_acc[kFwdWarps][2 * TAC_MAX_SCALES];

    const int c = blockIdx.x % a.nchunks;
    const long long row = blockIdx.x / a.nchunks;
    const long long c0 = (long long)c * a.chunk;
    stage_plain(tx, a.x + row * a.T, c0, a.tile_len, a.T, a.vec_ok != 0);
    stage_plain(ty, a.y + row * a.T, c0, a.tile_len, a.T, a.vec_ok != 0);
    if (threadIdx.x == 0) {
        int acc = 0;
        for (int s = 0; s < a.nscales; ++s) {
            const int h = a.sc[s].hop;
            long long lo = (c0 + h - 1) / h, hi = (c0 + a.chunk + h - 1) / h;
            if (hi > a.sc[s].F) hi = a.sc[s].F;
            if (lo > hi) lo = hi;
            const int nfr = nfr_of(a.sc[s].n);
            s_f_lo[s] = (int)lo; s_f_hi[s] = (int)hi;
            s_task_begin[s] = acc;
            acc += (int)((hi - lo + nfr - 1) / nfr);
        }
        s_task_begin[a.nscales] = acc;
    }
    for (int i = threadIdx.x; i < kFwdWarps * 2 * TAC_MAX_SCALES; i += blockDim.x) (&s_acc[0][0])[i] = 0.0;
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float2* xbuf = xbuf_all + warp * kXbufFloat2;
    const int ntasks = s_task_begin[a.nscales];
    for (int t = warp; t < ntasks; t += kFwdWarps) {          // static assignment -> deterministic sums
        int s = 0;
        while (t >= s_task_begin[s + 1]) ++s;
        const LossScale& sc = a.sc[s];
        const int f_first = s_f_lo[s] + (t - s_task_begin[s]) * nfr_of(sc.n);
        const int f_end = s_f_hi[s];
        float l1 = 0.0f, ll = 0.0f;
        switch (sc.n) {
            case 2048: loss_fwd_task<Plan2048>(sc, tx, ty, xbuf, lane, f_first, f_end, (int)c0, a.eps, l1, ll); break;
            case 1024: loss_fwd_task<Plan1024>(sc, tx, ty, xbuf, lane, f_first, f_end, (int)c0, a.eps, l1, ll); break;
            case 512: loss_fwd_task<Plan512>(sc, tx, ty, xbuf, lane, f_first, f_end, (int)c0, a.eps, l1, ll); break;
            case 256: loss_fwd_task<Plan256>(sc, tx, ty, xbuf, lane, f_first, f_end, (int)c0, a.eps, l1, ll); break;
            case 128: loss_fwd_task<Plan128>(sc, tx, ty, xbuf, lane, f_first, f_end, (int)c0, a.eps, l1, ll); break;
            default: loss_fwd_task<Plan64>(sc, tx, ty, xbuf, lane, f_first, f_end, (int)c0, a.eps, l1, ll); break;
        }
        l1 = warp_sum(l1);
        ll = warp_sum(ll);
        if (lane == 0) { s_acc[warp][2 * s] += (double)l1; s_acc[warp][2 * s + 1] += (double)ll; }
    }
    __syncthreads();
    if (threadIdx.x < 2 * a.nscales) {
        double v = 0.0;
        for (int w = 0; w < kFwdWarps; ++w) v += s_acc[w][threadIdx.x];
        a.partial[(long long)blockIdx.x * (2 * a.nscales) + threadIdx.x] = v;
    }
}

// Fixed-order reduction of the per-CTA partials: out[j] = sum_i partial[i][j].
__global__ void __launch_bounds__(256) partial_reduce_kernel(const double* __restrict__ partial, long long n, int width,
                                                            double* __restrict__ out) {
    __shared__ double sh[256];
    const int j = blockIdx.x;
    double v = 0.0;
    for (long long i = threadIdx.x; i < n; i += 256) v += partial[i * width + j];
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[j] = sh[0];
}

// ------------------------------------------------------------------------------------------ backward
#ifndef TAC_BWD_OCC
#define TAC_BWD_OCC 3
#endif
#ifndef TAC_BWD_WARPS
#define TAC_BWD_WARPS 4
#endif
constexpr int kBwdWarps = TAC_BWD_WARPS;
constexpr int kBwdThreads = kBwdWarps * 32;

template <class P, bool WITH_SUMS>
__device__ __forceinline__ void loss_bwd_task(const LossScale& sc, const float* __restrict__ xr, const float* __restrict__ yr,
                                              float* acc, float2* xbuf, int lane, int f_base, int f_stride, int f_end,
                                              long long c0, float eps, float c_l1, float c_log, float& acc_l1, float& acc_log) {
    cf z[P::BQ];
    float mx[P::BQ + 1];
    const float2* __restrict__ win2 = sc.t.win2;
    const int hop = sc.hop;
    const float* cur = xr;
    // frames of this task: f_base + i * f_stride, i in [0, NFR)
    auto load = [&](auto, int fr, int m, auto) -> cf {
        int f = f_base + fr * f_stride;
        f = f < f_end ? f : f_base;
        return mul2(__ldg(reinterpret_cast<const cf*>(cur + (long long)f * hop) + m), ldg_cf(win2 + m));   // plain product
    };
    warp_fft_fwd_pk<P>(z, xbuf, sc.t.tw, lane, load);
    warp_fft_post_pk<P>(z, sc.t.pt, lane, [&](int, int, cf v, int slot) {
        const float vr = lo(v), vi = hi(v);
        mx[slot] = fmaf(vr, vr, vi * vi);                            // |X|^2
    });
    cur = yr;
    warp_fft_fwd_pk<P>(z, xbuf, sc.t.tw, lane, load);
    const float eps2 = eps_squared(eps);
    float l1 = 0.0f, ll = 0.0f;
    // post-processing, loss terms and the adjoint's first stage in one loop over the bin pairs (fft_core_pk.cuh)
    cf zo[P::BQ], uo[P::BQ / 2], gh;
    warp_fft_post_adj_head_pk<P>(z, sc.t.pt, lane, [&](int, int, cf v, int slot) -> cf {
        const float vr = lo(v), vi = hi(v);
        const float g = loss_bin_grad<WITH_SUMS>(mx[slot], fmaf(vr, vr, vi * vi), eps2, c_l1, c_log, l1, ll);
        return mul2(v, bc(g));                                        // odd slots: conj(dL/dY[M-k]), as the adjoint expects
    }, zo, uo, gh);
    if (WITH_SUMS && (f_base + (lane / P::A) * f_stride) < f_end) { acc_l1 += l1; acc_log += ll * kLogHalf; }
    warp_fft_adj_tail_pk<P>(zo, uo, gh, xbuf, sc.t.tw, lane,
        [&](auto, int fr, auto ac, int b, cf g) {
            const int m = P::BQ * decltype(ac)::value + b;
            const int f = f_base + fr * f_stride;
            if (f < f_end) {
                cf* p = reinterpret_cast<cf*>(acc + ((long long)f * hop - c0)) + m;
                *p = fma2(g, ldg_cf(win2 + m), *p);
            }
        });
}

template <bool WITH_SUMS>
__global__ void __launch_bounds__(kBwdThreads, TAC_BWD_OCC) mrstft_bwd_kernel(const LossArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* acc = reinterpret_cast<float*>(smem_raw);
    float2* xbuf_all = reinterpret_cast<float2*>(acc + a.tile_len);
    __shared__ double s_acc[kBwdWarps][2 * TAC_MAX_SCALES];
    if (WITH_SUMS)
        for (int i = threadIdx.x; i < kBwdWarps * 2 * TAC_MAX_SCALES; i += blockDim.x) (&s_acc[0][0])[i] = 0.0;

    const int c = blockIdx.x % a.nchunks;
    const long long row = blockIdx.x / a.nchunks;
    const long long c0 = (long long)c * a.chunk;
    const float* xr = a.x + row * a.T;
    const float* yr = a.y + row * a.T;
    for (int j = threadIdx.x; j < a.tile_len; j += blockDim.x) acc[j] = 0.0f;
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float2* xbuf = xbuf_all + warp * kXbufFloat2;
    for (int s = 0; s < a.nscales; ++s) {
        const LossScale& sc = a.sc[s];
        const int h = sc.hop, R = sc.R, nfr = nfr_of(sc.n);
        long long lo = (c0 + h - 1) / h, hi = (c0 + a.chunk + h - 1) / h;
        if (hi > sc.F) hi = sc.F;
        const float c_l1 = __ldg(a.coef + 2 * s), c_log = __ldg(a.coef + 2 * s + 1);
        float l1 = 0.0f, ll = 0.0f;
        for (int r = 0; r < R; ++r) {
            // frames lo <= f < hi with f % R == r, grouped nfr at a time
            const int first = (int)lo + (int)(((r - lo % R) % R + R) % R);
            const int cnt = first < hi ? (int)((hi - first + R - 1) / R) : 0;
            const int ntasks = (cnt + nfr - 1) / nfr;
            for (int t = warp; t < ntasks; t += kBwdWarps) {
                const int f_base = first + t * nfr * R;
                switch (sc.n) {
                    case 2048: loss_bwd_task<Plan2048, WITH_SUMS>(sc, xr, yr, acc, xbuf, lane, f_base, R, (int)hi, c0, a.eps, c_l1, c_log, l1, ll); break;
                    case 1024: loss_bwd_task<Plan1024, WITH_SUMS>(sc, xr, yr, acc, xbuf, lane, f_base, R, (int)hi, c0, a.eps, c_l1, c_log, l1, ll); break;
                    case 512: loss_bwd_task<Plan512, WITH_SUMS>(sc, xr, yr, acc, xbuf, lane, f_base, R, (int)hi, c0, a.eps, c_l1, c_log, l1, ll); break;
                    case 256: loss_bwd_task<Plan256, WITH_SUMS>(sc, xr, yr, acc, xbuf, lane, f_base, R, (int)hi, c0, a.eps, c_l1, c_log, l1, ll); break;
                    case 128: loss_bwd_task<Plan128, WITH_SUMS>(sc, xr, yr, acc, xbuf, lane, f_base, R, (int)hi, c0, a.eps, c_l1, c_log, l1, ll); break;
                    default: loss_bwd_task<Plan64, WITH_SUMS>(sc, xr, yr, acc, xbuf, lane, f_base, R, (int)hi, c0, a.eps, c_l1, c_log, l1, ll); break;
                }
            }
            __syncthreads();
        }
        if (WITH_SUMS) {                                             // fixed task -> warp assignment: deterministic sums
            l1 = warp_sum(l1); ll = warp_sum(ll);
            if (lane == 0) { s_acc[warp][2 * s] = (double)l1; s_acc[warp][2 * s + 1] = (double)ll; }
        }
    }
    if (WITH_SUMS) {
        __syncthreads();
        if (threadIdx.x < 2 * a.nscales) {
            double v = 0.0;
            for (int w = 0; w < kBwdWarps; ++w) v += s_acc[w][threadIdx.x];
            a.partial[(long long)blockIdx.x * (2 * a.nscales) + threadIdx.x] = v;
        }
    }
    // write-out: [c0, c0 + chunk) is this CTA's, but its first (tile_len - chunk) samples also receive
    // the halo of chunk c-1 and its halo lands in chunk c+1 -> those two regions use atomics (at most
    // two contributions per sample, so the sum is order independent); grad was zeroed by the caller.
    float* gout = a.grad + row * a.T + c0;
    const int halo = a.tile_len - a.chunk;
    if (a.nchunks == 1) {
        for (int j = threadIdx.x; j < a.tile_len; j += blockDim.x)
            if (c0 + j < a.T) gout[j] = acc[j];
    } else {
        for (int j = threadIdx.x; j < a.tile_len; j += blockDim.x) {
            if (c0 + j >= a.T) break;
            if (j < halo || j >= a.chunk) atomicAdd(gout + j, acc[j]);
            else gout[j] = acc[j];
        }
    }
}

// ------------------------------------------------------------------------------------------ K2'
// One pass over (obs, dec) [B, 2, K, F]: per (item, channel) sum of squared error and of obs^2.
constexpr int kSpecThreads = 256;
constexpr int kSpecSplit = 8;          // CTAs per (item, channel) plane

__global__ void __launch_bounds__(kSpecThreads) spec_partial_kernel(const float* __restrict__ obs, const float* __restrict__ dec,
                                                                   long long plane, double* __restrict__ partial) {
    __shared__ double sh[2][kSpecThreads / 32];
    const long long pl = blockIdx.x / kSpecSplit;
    const int part = blockIdx.x % kSpecSplit;
    const float* o = obs + pl * plane;
    const float* d = dec + pl * plane;
    const long long per = (((plane + kSpecSplit - 1) / kSpecSplit) + 3) & ~3ll;   // keep 16-byte alignment
    const long long beg = (part * per < plane) ? part * per : plane, end = (beg + per < plane) ? beg + per : plane;
    float se = 0.0f, so = 0.0f;
    double dse = 0.0, dso = 0.0;
    const bool vec = (((uintptr_t)(o + beg) | (uintptr_t)(d + beg)) & 15u) == 0;
    long long i = beg + (long long)threadIdx.x * 4;
    int it = 0;
    if (vec) {
        for (; i + 3 < end; i += (long long)kSpecThreads * 4) {
            const float4 a4 = __ldcs(reinterpret_cast<const float4*>(o + i));
            const float4 b4 = __ldcs(reinterpret_cast<const float4*>(d + i));
            float e;
            e = a4.x - b4.x; se = fmaf(e, e, se); so = fmaf(a4.x, a4.x, so);
            e = a4.y - b4.y; se = fmaf(e, e, se); so = fmaf(a4.y, a4.y, so);
            e = a4.z - b4.z; se = fmaf(e, e, se); so = fmaf(a4.z, a4.z, so);
            e = a4.w - b4.w; se = fmaf(e, e, se); so = fmaf(a4.w, a4.w, so);
            if (++it == 16) { dse += se; dso += so; se = so = 0.0f; it = 0; }   // bound fp32 run length
        }
    }
    // tail (and the unaligned case): scalar
    long long j0 = vec ? (beg + ((end - beg) / 4) * 4) : beg;
    for (long long j = j0 + threadIdx.x; j < end; j += kSpecThreads) {
        const float av = __ldg(o + j), bv = __ldg(d + j);
        const float e = av - bv;
        se = fmaf(e, e, se); so = fmaf(av, av, so);
    }
    dse += se; dso += so;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        dse += __shfl_xor_sync(0xffffffffu, dse, off);
        dso += __shfl_xor_sync(0xffffffffu, dso, off);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { sh[0][warp] = dse; sh[1][warp] = dso; }
    __syncthreads();
    if (threadIdx.x < 2) {
        double v = 0.0;
        for (int w = 0; w < kSpecThreads / 32; ++w) v += sh[threadIdx.x][w];
        partial[(long long)blockIdx.x * 2 + threadIdx.x] = v;
    }
}

// partial: [B][2 channels][kSpecSplit][2]  ->  out5 + per_sample
__global__ void __launch_bounds__(256) spec_final_kernel(const double* __restrict__ partial, long long B, long long plane,
                                                        float* __restrict__ out5, float* __restrict__ per_sample) {
    __shared__ double sh[3][256];
    double e0 = 0.0, e1 = 0.0, ob = 0.0;
    for (long long b = threadIdx.x; b < B; b += 256) {
        double se[2] = {0.0, 0.0}, so = 0.0;
        for (int ch = 0; ch < 2; ++ch)
            for (int p = 0; p < kSpecSplit; ++p) {
                const double* q = partial + (((b * 2 + ch) * kSpecSplit) + p) * 2;
                se[ch] += q[0]; so += q[1];
            }
        if (per_sample) per_sample[b] = (float)(0.5 * (se[0] + se[1]) / (2.0 * (double)plane));
        e0 += se[0]; e1 += se[1]; ob += so;
    }
    sh[0][threadIdx.x] = e0; sh[1][threadIdx.x] = e1; sh[2][threadIdx.x] = ob;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) {
            sh[0][threadIdx.x] += sh[0][threadIdx.x + o];
            sh[1][threadIdx.x] += sh[1][threadIdx.x + o];
            sh[2][threadIdx.x] += sh[2][threadIdx.x + o];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double n1 = (double)B * (double)plane;
        const double tot = (sh[0][0] + sh[1][0]) / (2.0 * n1);
        out5[0] = (float)(sh[0][0] / n1);
        out5[1] = (float)(sh[1][0] / n1);
        out5[2] = (float)tot;
        out5[3] = (float)(-10.0 * log10(tot / (sh[2][0] / (2.0 * n1))));
        out5[4] = (float)(0.5 * tot);
    }
}


// Backward of K2': grad_dec[b, c, :, :] = w[b, c] * (dec - obs); the caller folds every output's
// upstream gradient and normaliser into the [B, 2] weight table w (device memory, no sync).
__global__ void __launch_bounds__(256) spec_bwd_kernel(const float* __restrict__ obs, const float* __restrict__ dec,
                                                      long long plane, const float* __restrict__ w, float* __restrict__ grad) {
    const long long pl = blockIdx.y;
    const float wv = __ldg(w + pl);
    const float* o = obs + pl * plane;
    const float* d = dec + pl * plane;
    float* g = grad + pl * plane;
    const bool vec = ((((uintptr_t)o | (uintptr_t)d | (uintptr_t)g) & 15u) == 0);
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (vec) {
        const long long n4 = plane >> 2;
        for (long long i = tid; i < n4; i += stride) {
            const float4 a = __ldcs(reinterpret_cast<const float4*>(o) + i);
            const float4 b = __ldcs(reinterpret_cast<const float4*>(d) + i);
            __stcs(reinterpret_cast<float4*>(g) + i, make_float4(wv * (b.x - a.x), wv * (b.y - a.y), wv * (b.z - a.z), wv * (b.w - a.w)));
        }
        for (long long i = (n4 << 2) + tid; i < plane; i += stride) g[i] = wv * (__ldg(d + i) - __ldg(o + i));
    } else {
        for (long long i = tid; i < plane; i += stride) g[i] = wv * (__ldg(d + i) - __ldg(o + i));
    }
}

static int fill_loss_args(LossArgs& a, const float* x, const float* y, int64_t B, int64_t T, int n_scales,
                          const int* n_fft, const int* hop, const int* win, float eps, int chunk) {
    TAC_REQUIRE(x && y && n_fft && hop, "null argument");
    TAC_REQUIRE(B > 0 && T > 0, "B and T must be positive");
    TAC_REQUIRE(n_scales >= 1 && n_scales <= TAC_MAX_SCALES, "n_scales must be in [1, %d]", TAC_MAX_SCALES);
    TAC_REQUIRE(eps > 0.0f, "eps must be positive");
    TAC_REQUIRE(T < (1ll << 30), "T too large");
    a.x = x; a.y = y; a.T = T; a.nscales = n_scales; a.eps = eps; a.chunk = chunk;
    a.vec_ok = ((((uintptr_t)x | (uintptr_t)y) & 15u) == 0 && (T & 3) == 0) ? 1 : 0;
    int nmax = 0;
    long long last_start = 0;
    for (int s = 0; s < n_scales; ++s) {
        LossScale& sc = a.sc[s];
        TAC_REQUIRE(supported_nfft(n_fft[s]), "unsupported n_fft %d", n_fft[s]);
        TAC_REQUIRE(hop[s] >= 2 && (hop[s] & 1) == 0 && hop[s] <= n_fft[s], "hop %d must be even and in [2, n_fft]", hop[s]);
        sc.n = n_fft[s]; sc.hop = hop[s];
        sc.F = (int)num_frames(T, sc.n, sc.hop, 0);
        TAC_REQUIRE(sc.F >= 1, "signal of %lld samples is shorter than n_fft=%d", (long long)T, sc.n);
        sc.K = sc.n / 2 + 1;
        sc.R = (sc.n + sc.hop - 1) / sc.hop;
        int rc = get_tables(sc.n, win ? win[s] : sc.n, 1, &sc.t);
        if (rc) return rc;
        nmax = sc.n > nmax ? sc.n : nmax;
        const long long ls = (long long)(sc.F - 1) * sc.hop;
        last_start = ls > last_start ? ls : last_start;
    }
    a.nchunks = (int)(last_start / chunk) + 1;
    a.tile_len = chunk + nmax;
    return TAC_OK;
}

constexpr int kFwdChunk = 2048;
#ifndef TAC_BWD_CHUNK
#define TAC_BWD_CHUNK 8192
#endif
constexpr int kBwdChunk = TAC_BWD_CHUNK;

}  // namespace tac

using namespace tac;

extern "C" size_t tac_mrstft_workspace_bytes(int64_t B, int64_t T, int n_scales, const int* n_fft, const int* hop) {
    if (B <= 0 || T <= 0 || n_scales <= 0 || !n_fft || !hop) return 0;
    long long last_start = 0;
    for (int s = 0; s < n_scales; ++s) {
        const long long F = num_frames(T, n_fft[s], hop[s], 0);
        const long long ls = F > 0 ? (F - 1) * hop[s] : 0;
        last_start = ls > last_start ? ls : last_start;
    }
    const long long nchunks = last_start / kFwdChunk + 1;
    size_t need = (size_t)(B * nchunks * 2 * n_scales) * sizeof(double);
#ifndef TAC_NO_STREAM
    if (n_scales <= TAC_MAX_SCALES && stream_applicable(T, n_scales, n_fft, hop, nullptr)) {
        const size_t sn = stream_workspace_bytes(B, T, n_scales, n_fft, hop);
        need = sn > need ? sn : need;
    }
#endif
    return need;
}

extern "C" int tac_mrstft_loss_fwd(const float* x, const float* y, int64_t B, int64_t T, int n_scales, const int* n_fft,
                                   const int* hop, const int* win, float eps, double* sums, void* workspace,
                                   size_t workspace_bytes, tac_stream_t stream) {
    LossArgs a;
    int rc = fill_loss_args(a, x, y, B, T, n_scales, n_fft, hop, win, eps, kFwdChunk);
    if (rc) return rc;
    TAC_REQUIRE(sums && workspace, "null output / workspace");
    const long long grid = B * a.nchunks;
    TAC_REQUIRE(grid < (1ll << 31), "grid too large");
    const size_t need = (size_t)grid * 2 * n_scales * sizeof(double);
    if (workspace_bytes < need) return set_error(TAC_EWORKSPACE, "workspace needs %zu bytes, got %zu", need, workspace_bytes);
    TAC_ALIGNED16(workspace, "workspace");
    a.partial = reinterpret_cast<double*>(workspace);
    a.coef = nullptr; a.grad = nullptr;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const size_t smem = 2 * (size_t)a.tile_len * sizeof(float) + (size_t)kFwdWarps * kXbufFloat2 * sizeof(float2);
    TAC_CHECK_CUDA(cudaFuncSetAttribute(mrstft_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
    mrstft_fwd_kernel<<<(unsigned)grid, kFwdThreads, smem, st>>>(a); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    partial_reduce_kernel<<<2 * n_scales, 256, 0, st>>>(a.partial, grid, 2 * n_scales, sums); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

static int launch_bwd(const float* x, const float* y, int64_t B, int64_t T, int n_scales, const int* n_fft, const int* hop,
                      const int* win, float eps, const float* coef, float* grad_y, double* sums, void* workspace,
                      size_t workspace_bytes, tac_stream_t stream) {
    LossArgs a;
    int rc = fill_loss_args(a, x, y, B, T, n_scales, n_fft, hop, win, eps, kBwdChunk);
    if (rc) return rc;
    TAC_REQUIRE(coef && grad_y, "null coef / grad_y");
#ifndef TAC_NO_STREAM
    // hop == n_fft / 4 everywhere (the msstftd.py:174-175 framing): barrier-free streaming kernel (loss_stream.cu)
    if (stream_applicable(T, n_scales, n_fft, hop, win) && workspace &&
        workspace_bytes >= stream_workspace_bytes(B, T, n_scales, n_fft, hop))
        return launch_stream(x, y, B, T, n_scales, n_fft, hop, win, eps, coef, sums, grad_y, workspace, workspace_bytes,
                             reinterpret_cast<cudaStream_t>(stream));
#endif
    TAC_REQUIRE((((uintptr_t)x | (uintptr_t)y) & 7u) == 0 && (T & 1) == 0, "backward needs 8-byte aligned x, y and even T");
    a.partial = nullptr; a.coef = coef; a.grad = grad_y;
    const long long grid = B * a.nchunks;
    TAC_REQUIRE(grid < (1ll << 31), "grid too large");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (sums) {
        const size_t need = (size_t)grid * 2 * n_scales * sizeof(double);
        if (!workspace || workspace_bytes < need)
            return set_error(TAC_EWORKSPACE, "workspace needs %zu bytes, got %zu", need, workspace_bytes);
        TAC_ALIGNED16(workspace, "workspace");
        a.partial = reinterpret_cast<double*>(workspace);
    }
    if (a.nchunks > 1) TAC_CHECK_CUDA(cudaMemsetAsync(grad_y, 0, (size_t)B * T * sizeof(float), st));
    const size_t smem = (size_t)a.tile_len * sizeof(float) + (size_t)kBwdWarps * kXbufFloat2 * sizeof(float2);
    TAC_SMEM_OPT_IN(mrstft_bwd_kernel<false>, 75 * 1024);
    TAC_SMEM_OPT_IN(mrstft_bwd_kernel<true>, 75 * 1024);
    if (sums) {
        mrstft_bwd_kernel<true><<<(unsigned)grid, kBwdThreads, smem, st>>>(a); TAC_LAUNCHED(1);
        TAC_CHECK_CUDA(cudaGetLastError());
        partial_reduce_kernel<<<2 * n_scales, 256, 0, st>>>(a.partial, grid, 2 * n_scales, sums); TAC_LAUNCHED(1);
    } else {
        mrstft_bwd_kernel<false><<<(unsigned)grid, kBwdThreads, smem, st>>>(a); TAC_LAUNCHED(1);
    }
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

extern "C" int tac_mrstft_loss_bwd(const float* x, const float* y, int64_t B, int64_t T, int n_scales, const int* n_fft,
                                   const int* hop, const int* win, float eps, const float* coef, float* grad_y,
                                   void* workspace, size_t workspace_bytes, tac_stream_t stream) {
    return launch_bwd(x, y, B, T, n_scales, n_fft, hop, win, eps, coef, grad_y, nullptr, workspace, workspace_bytes, stream);
}

extern "C" int tac_mrstft_loss_fwd_bwd(const float* x, const float* y, int64_t B, int64_t T, int n_scales, const int* n_fft,
                                       const int* hop, const int* win, float eps, const float* coef, double* sums,
                                       float* grad_y, void* workspace, size_t workspace_bytes, tac_stream_t stream) {
    TAC_REQUIRE(sums != nullptr, "null sums");
    return launch_bwd(x, y, B, T, n_scales, n_fft, hop, win, eps, coef, grad_y, sums, workspace, workspace_bytes, stream);
}

extern "C" size_t tac_spec_losses_workspace_bytes(int64_t B) {
    return B > 0 ? (size_t)B * 2 * kSpecSplit * 2 * sizeof(double) : 0;
}

extern "C" int tac_spec_losses(const float* obs, const float* dec, int64_t B, int64_t K, int64_t F, float* out5,
                               float* per_sample, void* workspace, size_t workspace_bytes, tac_stream_t stream) {
    TAC_REQUIRE(obs && dec && out5 && workspace, "null argument");
    TAC_REQUIRE(B > 0 && K > 0 && F > 0, "B, K, F must be positive");
    const size_t need = tac_spec_losses_workspace_bytes(B);
    if (workspace_bytes < need) return set_error(TAC_EWORKSPACE, "workspace needs %zu bytes, got %zu", need, workspace_bytes);
    TAC_REQUIRE(B * 2 * kSpecSplit < (1ll << 31), "grid too large");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    double* partial = reinterpret_cast<double*>(workspace);
    spec_partial_kernel<<<(unsigned)(B * 2 * kSpecSplit), kSpecThreads, 0, st>>>(obs, dec, K * F, partial); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    spec_final_kernel<<<1, 256, 0, st>>>(partial, B, K * F, out5, per_sample); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

extern "C" int tac_spec_losses_bwd(const float* obs, const float* dec, int64_t B, int64_t K, int64_t F, const float* w,
                                   float* grad_dec, tac_stream_t stream) {
    TAC_REQUIRE(obs && dec && w && grad_dec, "null argument");
    TAC_REQUIRE(B > 0 && K > 0 && F > 0, "B, K, F must be positive");
    TAC_REQUIRE(B * 2 < 65536, "B too large for one launch (max 32767)");
    const long long plane = K * F;
    long long gx = (plane / 4 + 255) / 256;
    gx = gx < 1 ? 1 : (gx > 64 ? 64 : gx);
    spec_bwd_kernel<<<dim3((unsigned)gx, (unsigned)(B * 2)), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(obs, dec, plane, w, grad_dec); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

// ------------------------------------------------------------------------------------------ K2' without a workspace
// tac_spec_losses_static: the same two kernels as tac_spec_losses, with the per-CTA partial sums parked in memory that
// belongs to the library image instead of a buffer the caller allocates.  It exists for hosts that hold device pointers
// to their data but have no way to obtain more device memory at the call site: a plugin running inside another
// framework's memory pool whose allocator it must not bypass, a real-time audio thread or a signal/callback context in
// which cudaMalloc's implicit device synchronisation and driver locks are off limits, or a process whose memory budget
// was fixed up front.  The price is a cap on B and one array per GPU for the lifetime of the process.
//
// kSpecStaticMaxB items => kSpecStaticMaxB * 2 channels * kSpecSplit CTAs * 2 sums * 8 bytes = 1 MiB per device.
constexpr long long kSpecStaticMaxB = TAC_SPEC_LOSSES_STATIC_MAX_B;
static __device__ __align__(256) double g_spec_static_partial[kSpecStaticMaxB * 2 * kSpecSplit * 2];

// In host code `g_spec_static_partial` is only a handle for the runtime's symbol table, never a device address: the
// address differs per device and has to be asked for with that device current.  Looked up once per device and kept.
static int spec_static_scratch(double** out) {
    constexpr int kMaxDev = 64;
    static std::atomic<double*> addr[kMaxDev];
    int dev = 0;
    TAC_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDev) return set_error(TAC_EINVAL, "device ordinal %d outside 0..%d", dev, kMaxDev - 1);
    double* p = addr[dev].load(std::memory_order_acquire);
    if (!p) {
        void* q = nullptr;
        TAC_CHECK_CUDA(cudaGetSymbolAddress(&q, g_spec_static_partial));
        if (!q) return set_error(TAC_ECUDA, "cudaGetSymbolAddress returned a null address for the static scratch");
        p = static_cast<double*>(q);
        addr[dev].store(p, std::memory_order_release);   // two racing first calls store the same value
    }
    *out = p;
    return TAC_OK;
}

extern "C" int tac_spec_losses_static(const float* obs, const float* dec, int64_t B, int64_t K, int64_t F, float* out5,
                                      float* per_sample, tac_stream_t stream) {
    TAC_REQUIRE(obs && dec && out5, "null argument");
    TAC_REQUIRE(B > 0 && K > 0 && F > 0, "B, K, F must be positive");
    if (B > kSpecStaticMaxB)
        return set_error(TAC_EINVAL, "tac_spec_losses_static holds scratch for at most %lld items, got B = %lld: split "
                         "the batch or use tac_spec_losses with a caller-allocated workspace",
                         kSpecStaticMaxB, (long long)B);
    double* partial = nullptr;
    const int rc = spec_static_scratch(&partial);
    if (rc) return rc;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    // Nothing to clear: spec_partial_kernel's CTA i overwrites slots 2i, 2i+1 and spec_final_kernel reads only the
    // B * 2 * kSpecSplit * 2 slots this very launch pair wrote, so what an earlier call left behind is never read.
    spec_partial_kernel<<<(unsigned)(B * 2 * kSpecSplit), kSpecThreads, 0, st>>>(obs, dec, K * F, partial); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    spec_final_kernel<<<1, 256, 0, st>>>(partial, B, K * F, out5, per_sample); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

// K1x -- single-scale STFT for ANY n_fft with an output epilogue fused into the write-out (tac_stft_ex).
//
// The reference's remaining front ends do not use power-of-two frames:
//   * spectrogram.ipynb cell 9 (raw :2894-2913): librosa.stft(n_fft = 2046, hop 512, win 2046) -> abs, angle, amplitude_to_db
//   * sgmse/sgmse/data_module.py:109-116,162-169,212-214: torch.stft(n_fft = 510, hop 128, hann, center=True) followed by
//     abs(z)^0.5 * exp(1j angle(z)) * 0.15, frames zero-padded to a multiple of 64 (sgmse/sgmse/util/other.py:76-90)
//   * testing_scripts.ipynb:246-249: torchaudio.transforms.Spectrogram() (n_fft = 400, hop 200, power 2, reflect) and
//     10 * log10(S + 1e-10)
// A CTA owns TF consecutive frames of one row.  The samples are staged once; frames whose length the warp core
// (fft_core_pk.cuh) has a plan for are transformed by warps, any other length by Bluestein's chirp-z identity
//     X[k] = conj(c[k]) * sum_j (x[j] w[j] conj(c[j])) c[k - j],        c[m] = exp(+j pi m^2 / n),
// i.e. a circular convolution of size M = max(1024, 2^ceil(log2(n + K - 1))) done on the packed warp core: M = 1024 by ONE warp
// per frame (forward 1024-point complex transform = the inner transform of warp_fft_fwd_pk<Plan2048>, product with the
// pre-transformed chirp, warp_cfft_inv_pk); M = 2048 / 4096 by R = 2 / 4 warps per frame, one 1024-point sub-transform of
// the decimated sequence y[R m + r] each, with everything between the forward and the inverse sub-transforms pointwise in
// k' < 1024 (twiddle, radix-R, filter, inverse radix-R, twiddle) in a shared [R][1024] buffer between two named barriers.
// (Round 2: this replaced a CTA-wide radix-2 convolution -- 2 log2 M block barriers and ~330 KB of shared-memory traffic per
// frame: SGMSE front end 1.94 -> 0.34 ms, n = 400 1.27 -> 0.21 ms, n = 2046 3.33 -> 0.89 ms on 512 x 1 s.)
// Every table (window * conj chirp, FFT of the chirp filter, twiddles) is generated on the host in
// float64, with m^2 reduced mod 2n in integers.  Both transforms leave (re, im) in a [K][TF + 1] shared-memory tile and
// share ONE write-out that applies the epilogue on the way to HBM: complex, power-law compression, |z| and angle,
// |z|^p, p1 * log10(|z|^p + eps), dB against a fixed reference, a mel filter bank -- each (k) row leaves as a
// contiguous run of TF frames, the frames beyond F are zero-filled by the CTA of the last tile (pad_tensor / pad_spec).
#include <math.h>
#include <complex>
#include <map>
#include <mutex>
#include <tuple>
#include <vector>

#include "tac_common.cuh"
#include "fft_core_pk.cuh"

namespace tac {

constexpr int kExThreads = 256;
constexpr int kExWarps = kExThreads / 32;
constexpr int kBluMaxM = 4096;

struct BluTables {
    const float2* a;      // [n]    window[j] * scale * conj(c[j])
    const float2* bf;     // [M]    FFT_M(chirp filter) / M, natural order
    const float2* post;   // [K]    conj(c[k])
    const float2* tw;     // [M/2]  exp(-2 pi j i / M)
    int M, logM;
};

namespace {
struct BluEntry { void* base; BluTables t; };
std::mutex g_blu_mu;
std::map<std::tuple<int, int, int, int, int>, BluEntry> g_blu_cache;   // (device, n, win, normalized, window kind)

void host_fft(std::vector<std::complex<double>>& v) {                  // in-place forward radix-2, natural order
    const size_t n = v.size();
    for (size_t i = 1, j = 0; i < n; ++i) {
        size_t bit = n >> 1;
        for (; j & bit; bit >>= 1) j ^= bit;
        j ^= bit;
        if (i < j) std::swap(v[i], v[j]);
    }
    for (size_t len = 2; len <= n; len <<= 1) {
        for (size_t i = 0; i < n; i += len)
            for (size_t k = 0; k < len / 2; ++k) {
                const double ang = -2.0 * M_PI * (double)k / (double)len;
                const std::complex<double> w(cos(ang), sin(ang));
                const std::complex<double> u = v[i + k], t = v[i + k + len / 2] * w;
                v[i + k] = u + t;
                v[i + k + len / 2] = u - t;
            }
    }
}

inline std::complex<double> chirp(long long m, int n) {                // exp(+j pi m^2 / n), m^2 reduced mod 2n exactly
    const long long e = (m * m) % (2ll * n);
    const double ang = M_PI * (double)e / (double)n;
    return std::complex<double>(cos(ang), sin(ang));
}
}  // namespace

// window kinds: 0 periodic Hann, 1 sqrt of the periodic Hann (sgmse/sgmse/data_module.py:13-17)
// may_build false (the caller's stream is being captured): a missing table is an error instead of a cudaMalloc + synchronous
// upload in the middle of the capture -- tac_prepare_stft_ex builds it ahead of time.
static int get_blu_tables(int n, int win, int normalized, int window, BluTables* out, bool may_build = true) {
    int dev = 0;
    TAC_CHECK_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(g_blu_mu);
    auto key = std::make_tuple(dev, n, win, normalized ? 1 : 0, window);
    auto it = g_blu_cache.find(key);
    if (it != g_blu_cache.end()) { *out = it->second.t; return TAC_OK; }
    if (!may_build)
        return set_error(TAC_EINVAL, "chirp-z tables for n_fft=%d win=%d normalized=%d window=%d do not exist yet and the stream is "
                         "being captured: call tac_prepare_stft_ex(n_fft, win, normalized, window) before the capture",
                         n, win, normalized ? 1 : 0, window);
    const int K = n / 2 + 1;
    int M = 1, logM = 0;
    while (M < n + K - 1 || M < 1024) { M <<= 1; ++logM; }     // the warp transforms are 1024 points wide
    if (M > kBluMaxM) return set_error(TAC_EINVAL, "n_fft %d too large for the chirp-z path (n_fft + n_fft/2 <= %d)", n, kBluMaxM);
    std::vector<double> w(n, 0.0);
    const int left = (n - win) / 2;
    double ss = 0.0;
    for (int i = 0; i < win; ++i) {
        double v = 0.5 - 0.5 * cos(2.0 * M_PI * (double)i / (double)win);
        if (window == 1) v = sqrt(v);
        w[left + i] = v;
        ss += v * v;
    }
    const double scale = normalized ? 1.0 / sqrt(ss) : 1.0;
    std::vector<float2> host((size_t)n + M + K + M / 2);
    float2* ha = host.data();
    float2* hbf = ha + n;
    float2* hpost = hbf + M;
    float2* htw = hpost + K;
    for (int j = 0; j < n; ++j) {
        const std::complex<double> v = std::conj(chirp(j, n)) * (w[j] * scale);
        ha[j] = make_float2((float)v.real(), (float)v.imag());
    }
    std::vector<std::complex<double>> b((size_t)M, std::complex<double>(0.0, 0.0));
    for (int m = 0; m < K; ++m) b[m] = chirp(m, n);
    for (int m = 1; m < n; ++m) b[M - m] = chirp(m, n);                 // c is even: c[-m] = c[m]
    host_fft(b);
    // FFT_M(chirp filter) / M in natural order, rows of 32 consecutive bins per sub-transform: [q][k2][k1] <- Bf[k1 + 32 k2 + 1024 q]
    for (int i = 0; i < M; ++i) {
        const std::complex<double> v = b[i] / (double)M;
        hbf[i] = make_float2((float)v.real(), (float)v.imag());
    }
    for (int k = 0; k < K; ++k) {
        const std::complex<double> v = std::conj(chirp(k, n));
        hpost[k] = make_float2((float)v.real(), (float)v.imag());
    }
    for (int i = 0; i < M / 2; ++i) {
        const double ang = -2.0 * M_PI * (double)i / (double)M;
        htw[i] = make_float2((float)cos(ang), (float)sin(ang));
    }
    void* base = nullptr;
    TAC_CHECK_CUDA(cudaMalloc(&base, host.size() * sizeof(float2)));
    cudaError_t e = cudaMemcpy(base, host.data(), host.size() * sizeof(float2), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        cudaFree(base);
        return set_error(TAC_ECUDA, "table upload failed: %s", cudaGetErrorString(e));
    }
    BluEntry en;
    en.base = base;
    en.t.a = reinterpret_cast<const float2*>(base);
    en.t.bf = en.t.a + n;
    en.t.post = en.t.bf + M;
    en.t.tw = en.t.post + K;
    en.t.M = M; en.t.logM = logM;
    g_blu_cache[key] = en;
    *out = en.t;
    return TAC_OK;
}

void free_blu_tables() {                                                // tac_shutdown()
    std::lock_guard<std::mutex> lk(g_blu_mu);
    int cur = 0;
    cudaGetDevice(&cur);
    for (auto& kv : g_blu_cache) {
        cudaSetDevice(std::get<0>(kv.first));
        cudaFree(kv.second.base);
    }
    g_blu_cache.clear();
    cudaSetDevice(cur);
}

struct ExArgs {
    const float* x;
    long long T;
    int n, hop, pad, F, K, TF, ntiles, tile_len, pad_mode, vec_ok;
    long long f_pad;                 // time extent of the output rows (>= F)
    FftTables t;                     // warp-core path
    BluTables bt;                    // chirp-z path
    tac_epilogue epi;
    void* out;
};

__device__ __forceinline__ float ex_padded_sample(const float* __restrict__ xr, long long g, long long T, int pad_mode) {
    if (g >= 0 && g < T) return __ldg(xr + g);
    if (pad_mode == TAC_PAD_REFLECT) {
        long long r = g < 0 ? -g : 2 * (T - 1) - g;
        if (r >= 0 && r < T) return __ldg(xr + r);
    }
    return 0.0f;
}

__device__ __forceinline__ float2 cmulf(float2 a, float2 b) {
    return make_float2(fmaf(a.x, b.x, -a.y * b.y), fmaf(a.x, b.y, a.y * b.x));
}

// |z|^p for the exponents the reference uses without a powf (1, 2, 0.5); anything else through powf.
__device__ __forceinline__ float ex_pow_mag(float re, float im, float p) {
    const float m2 = fmaf(re, re, im * im);
    if (p == 2.0f) return m2;
    const float m = sqrtf(m2);
    if (p == 1.0f) return m;
    if (p == 0.5f) return sqrtf(m);
    return m > 0.0f ? powf(m, p) : 0.0f;
}

// The shared write-out: tile (re, im) [K][TS] -> HBM through the epilogue.  Called by all threads after a __syncthreads.
__device__ void ex_write_out(const ExArgs& a, float* o0, float* o1, int TS, long long row, int f0, int nf, bool last_tile) {
    const tac_epilogue& e = a.epi;
    const int K = a.K, kind = e.kind;
    const long long fp = a.f_pad;
    const int total = K * nf;
    float lmax = 0.0f;
    if (kind == TAC_EPI_MEL) {
        for (int i = threadIdx.x; i < total; i += blockDim.x) {
            const int k = i / nf, tt = i - k * nf;
            o0[k * TS + tt] = ex_pow_mag(o0[k * TS + tt], o1[k * TS + tt], e.p0);
        }
        __syncthreads();
        float* out = reinterpret_cast<float*>(a.out) + row * (long long)e.n_mels * fp + f0;
        for (int i = threadIdx.x; i < e.n_mels * nf; i += blockDim.x) {
            const int m = i / nf, tt = i - m * nf;
            const int klo = __ldg(e.mel_range + 2 * m), khi = __ldg(e.mel_range + 2 * m + 1);
            float acc = 0.0f;
            for (int k = klo; k < khi; ++k) acc = fmaf(__ldg(e.mel_fb + (long long)k * e.n_mels + m), o0[k * TS + tt], acc);
            if (e.p1 != 0.0f) acc = e.p1 * log10f(acc + e.p2);
            __stcs(out + (long long)m * fp + tt, acc);
        }
    } else if (kind == TAC_EPI_COMPLEX || kind == TAC_EPI_COMPRESS) {
        float2* out = reinterpret_cast<float2*>(a.out) + row * (long long)K * fp + f0;
        for (int i = threadIdx.x; i < total; i += blockDim.x) {
            const int k = i / nf, tt = i - k * nf;
            float re = o0[k * TS + tt], im = o1[k * TS + tt];
            if (kind == TAC_EPI_COMPRESS) {                      // |z|^p0 e^{j angle z} p1 = z |z|^(p0 - 1) p1
                const float m2 = fmaf(re, re, im * im);
                float s = 0.0f;
                if (m2 > 0.0f) {
                    const float m = sqrtf(m2);
                    s = (e.p0 == 0.5f ? rsqrtf(m) : (e.p0 == 1.0f ? 1.0f : powf(m, e.p0 - 1.0f))) * e.p1;
                }
                re *= s; im *= s;
            }
            __stcs(out + (long long)k * fp + tt, make_float2(re, im));
        }
    } else if (kind == TAC_EPI_MAGPHASE || kind == TAC_EPI_MAGDB) {
        float* out0 = reinterpret_cast<float*>(a.out) + (row * 2) * (long long)K * fp + f0;
        float* out1 = out0 + (long long)K * fp;
        for (int i = threadIdx.x; i < total; i += blockDim.x) {
            const int k = i / nf, tt = i - k * nf;
            const float re = o0[k * TS + tt], im = o1[k * TS + tt];
            const float mag = sqrtf(fmaf(re, re, im * im));
            lmax = fmaxf(lmax, mag);
            __stcs(out0 + (long long)k * fp + tt, mag);
            if (kind == TAC_EPI_MAGPHASE) __stcs(out1 + (long long)k * fp + tt, atan2f(im, re));
            else __stcs(out1 + (long long)k * fp + tt, fmaxf(20.0f * log10f(fmaxf(mag, e.p0)) - e.p1, e.p2));
        }
    } else {                                                     // one real plane
        float* out = reinterpret_cast<float*>(a.out) + row * (long long)K * fp + f0;
        for (int i = threadIdx.x; i < total; i += blockDim.x) {
            const int k = i / nf, tt = i - k * nf;
            const float re = o0[k * TS + tt], im = o1[k * TS + tt];
            float v;
            if (kind == TAC_EPI_DB) {                            // amplitude_to_db against a fixed reference
                const float mag = sqrtf(fmaf(re, re, im * im));
                lmax = fmaxf(lmax, mag);
                v = fmaxf(20.0f * log10f(fmaxf(mag, e.p0)) - e.p1, e.p2);
            } else {
                v = ex_pow_mag(re, im, e.p0);
                if (e.p0 == 1.0f) lmax = fmaxf(lmax, v);
                if (kind == TAC_EPI_LOGPOWER) v = e.p1 * log10f(v + e.p2);
            }
            __stcs(out + (long long)k * fp + tt, v);
        }
    }
    if (e.rowmax != nullptr) {                                   // running max |z| of the row (float bits order like unsigned)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) lmax = fmaxf(lmax, __shfl_xor_sync(0xffffffffu, lmax, o));
        if ((threadIdx.x & 31) == 0 && lmax > 0.0f) atomicMax(e.rowmax + row, __float_as_uint(lmax));
    }
    if (last_tile && fp > a.F) {                                 // zero-fill frames [F, f_pad) of every output row
        int nrows, esz;
        switch (kind) {
            case TAC_EPI_MEL: nrows = e.n_mels; esz = 1; break;
            case TAC_EPI_COMPLEX: case TAC_EPI_COMPRESS: nrows = K; esz = 2; break;
            case TAC_EPI_MAGPHASE: case TAC_EPI_MAGDB: nrows = 2 * K; esz = 1; break;
            default: nrows = K; esz = 1; break;
        }
        const int npad = (int)(fp - a.F) * esz;
        float* base = reinterpret_cast<float*>(a.out) + (row * nrows * fp + a.F) * esz;
        const long long tot = (long long)nrows * npad;
        for (long long i = threadIdx.x; i < tot; i += blockDim.x) {
            const long long r = i / npad, c = i - r * npad;
            base[r * fp * esz + c] = 0.0f;
        }
    }
}

struct WBluPlan {};      // chirp-z with M == 1024: the convolution by ONE WARP per frame on the packed warp core (n_fft <= 682)
template <int R> struct WBluSplit {};   // chirp-z with M == 1024 R (R = 2, 4): R warps per frame, one 1024-point sub-transform each
template <class P> struct is_split { static constexpr int R = 0; };
template <int R_> struct is_split<WBluSplit<R_>> { static constexpr int R = R_; };

template <class P>
__global__ void __launch_bounds__(kExThreads, std::is_same<P, WBluPlan>::value ? 2 : 1) stft_ex_kernel(const ExArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* tile = reinterpret_cast<float*>(smem_raw);
    const int TS = a.TF + 1;
    float* o0 = tile + a.tile_len;
    float* o1 = o0 + (size_t)a.K * TS;
    unsigned char* work = reinterpret_cast<unsigned char*>(tile + a.tile_len + ((2 * (size_t)a.K * TS + 3) & ~(size_t)3));
    __shared__ int s_next;

    const int ft = blockIdx.x % a.ntiles;
    const long long row = blockIdx.x / a.ntiles;
    const float* xr = a.x + row * a.T;
    const int f0 = ft * a.TF;
    const int f_end = min(a.F, f0 + a.TF);
    const int nf = f_end - f0;
    const long long g0 = (long long)f0 * a.hop - a.pad;
    {   // in-range, 16-byte aligned middle of the tile with float4 loads even when the tile hangs over a row end (centre padding in
        // front, the last tile behind); the padded samples and the <= 3 + 3 around the middle take the scalar path (as stft.cu)
        const int len = a.tile_len;
        const long long lo_ll = g0 < 0 ? -g0 : 0, hi_ll = a.T - g0;
        const int lo = (int)(lo_ll < len ? lo_ll : len);
        const int hi = (int)(hi_ll < lo ? lo : (hi_ll < len ? hi_ll : len));
        int ja = lo, jb = lo;
        if (a.vec_ok && (g0 & 3) == 0) {
            ja = (lo + 3) & ~3; jb = hi & ~3;
            if (jb < ja) ja = jb = lo;
        }
        if (jb > ja) {
            const float4* src = reinterpret_cast<const float4*>(xr) + (g0 >> 2);
            float4* dst = reinterpret_cast<float4*>(tile);
            for (int i = (ja >> 2) + threadIdx.x; i < (jb >> 2); i += blockDim.x) dst[i] = __ldg(src + i);
        }
        for (int j = threadIdx.x; j < ja; j += blockDim.x) tile[j] = ex_padded_sample(xr, g0 + j, a.T, a.pad_mode);
        for (int j = jb + threadIdx.x; j < len; j += blockDim.x) tile[j] = ex_padded_sample(xr, g0 + j, a.T, a.pad_mode);
    }
    if (threadIdx.x == 0) s_next = 0;
    __syncthreads();

    if constexpr (std::is_same<P, WBluPlan>::value) {
        // X[k] = conj(c_k) sum_j (x_j w_j conj(c_j)) c_{k-j}: a circular convolution of 1024 points, done per frame by one warp as
        // forward complex transform (the inner transform of warp_fft_fwd_pk<Plan2048>: 1024 complex points, lane k1 / slot k2 <->
        // bin k1 + 32 k2) -> product with the pre-transformed chirp (table in that layout) -> inverse transform
        // (warp_cfft_inv_pk); only the outputs m < K are kept.  Against the CTA-wide radix-2 form below (20 block barriers and
        // ~330 KB of shared-memory traffic per frame): one exchange round trip per transform, no block barrier.
        using PW = Plan2048;
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        float2* xbuf = reinterpret_cast<float2*>(work) + warp * kXbufFloat2;
        const float2* __restrict__ ta = a.bt.a;
        const float2* __restrict__ tbw = a.bt.bf;                 // M == 1024: Bf[k1 + 32 k2] at [k2][k1]
        const float2* __restrict__ tp = a.bt.post;
        const int n = a.n, K = a.K;
        for (int fr = warp; fr < nf; fr += kExWarps) {
            const float* src = tile + fr * a.hop;
            cf z[PW::BQ];
            auto load = [&](auto, int, int m, auto) -> cf {
                if (m >= n) return 0ull;
                const float sv = src[m];
                return mul2(ldg_cf(ta + m), bc(sv));
            };
            warp_fft_fwd_pk<PW>(z, xbuf, a.t.tw, lane, load);
            static_for<0, PW::BQ>([&](auto kc) {
                constexpr int k2 = decltype(kc)::value;
                const cf w = ldg_cf(tbw + k2 * 32 + lane);
                z[k2] = fma2(rotp(z[k2]), bc(hi(w)), mul2(z[k2], bc(lo(w))));           // z * w
            });
            warp_cfft_inv_pk<PW>(z, xbuf, a.t.tw, lane, [&](auto, int, auto ac, int b, cf v) {
                const int m = PW::BQ * decltype(ac)::value + b;
                if (m < K) {
                    const cf w = ldg_cf(tp + m);
                    const cf r = fma2(rotp(v), bc(hi(w)), mul2(v, bc(lo(w))));
                    o0[m * TS + fr] = lo(r);
                    o1[m * TS + fr] = hi(r);
                }
            });
        }
        __syncthreads();
    } else if constexpr (is_split<P>::R > 0) {
        // M = 1024 R points: decimation in time into R sub-sequences y[R m + r], one 1024-point warp transform each (warp r of a
        // group of R); everything between the forward and the inverse sub-transforms is POINTWISE in k' < 1024 --
        //   e_r = W_M^{r k'} E_r[k'],  Z_q = sum_r W_R^{r q} e_r  (bin k' + 1024 q),  V_q = Z_q Bf[k' + 1024 q],
        //   U_r'[k'] = conj(W_M^{r' k'}) sum_q W_R^{-r' q} V_q,   v[R m' + r'] = IDFT_1024(U_r')[m']
        // -- and runs in place in the group's [R][1024] buffer between two named barriers.  Only outputs m < K are kept.
        constexpr int R = is_split<P>::R;
        using PW = Plan2048;
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        constexpr int NG = kExWarps / R;
        const int group = warp / R, r = warp % R;
        float2* xbuf = reinterpret_cast<float2*>(work) + warp * kXbufFloat2;
        float2* ebuf = reinterpret_cast<float2*>(work) + kExWarps * kXbufFloat2 + group * (R * 1024);
        cf* ebc = reinterpret_cast<cf*>(ebuf);
        const float2* __restrict__ ta = a.bt.a;
        const float2* __restrict__ tbf = a.bt.bf;
        const float2* __restrict__ twm = a.bt.tw;                // exp(-2 pi j i / M), i < M / 2
        const float2* __restrict__ tp = a.bt.post;
        const int n = a.n, K = a.K;
        constexpr int M = 1024 * R;
        auto wm = [&](int idx) -> float2 {                       // exp(-2 pi j idx / M), idx < M
            if (idx < M / 2) return __ldg(twm + idx);
            const float2 w = __ldg(twm + idx - M / 2);
            return make_float2(-w.x, -w.y);
        };
        for (int fr0 = 0; fr0 < nf; fr0 += NG) {
            const int fr = fr0 + group;
            if (fr >= nf) continue;                              // uniform over the group: its barriers are skipped together
            const float* src = tile + fr * a.hop;
            {
                cf z[PW::BQ];
                auto load = [&](auto, int, int m, auto) -> cf {
                    const int j = R * m + r;
                    if (j >= n) return 0ull;
                    return mul2(ldg_cf(ta + j), bc(src[j]));
                };
                warp_fft_fwd_pk<PW>(z, xbuf, a.t.tw, lane, load);
                static_for<0, PW::BQ>([&](auto kc) { constexpr int k2 = decltype(kc)::value; ebc[r * 1024 + k2 * 32 + lane] = z[k2]; });
            }
            asm volatile("bar.sync %0, %1;" :: "r"(1 + group), "r"(R * 32) : "memory");
            for (int kp = r * 32 + lane; kp < 1024; kp += R * 32) {
                float2 e[R], V[R];
#pragma unroll
                for (int i = 0; i < R; ++i) { e[i] = ebuf[i * 1024 + kp]; if (i > 0) e[i] = cmulf(e[i], wm(i * kp)); }
                if constexpr (R == 2) {
                    V[0] = make_float2(e[0].x + e[1].x, e[0].y + e[1].y);
                    V[1] = make_float2(e[0].x - e[1].x, e[0].y - e[1].y);
                } else {                                          // W_4 = -j: forward radix-4
                    const float2 s02 = make_float2(e[0].x + e[2].x, e[0].y + e[2].y), d02 = make_float2(e[0].x - e[2].x, e[0].y - e[2].y);
                    const float2 s13 = make_float2(e[1].x + e[3].x, e[1].y + e[3].y), d13 = make_float2(e[1].x - e[3].x, e[1].y - e[3].y);
                    V[0] = make_float2(s02.x + s13.x, s02.y + s13.y);
                    V[2] = make_float2(s02.x - s13.x, s02.y - s13.y);
                    V[1] = make_float2(d02.x + d13.y, d02.y - d13.x);       // d02 - j d13
                    V[3] = make_float2(d02.x - d13.y, d02.y + d13.x);       // d02 + j d13
                }
#pragma unroll
                for (int q = 0; q < R; ++q) V[q] = cmulf(V[q], __ldg(tbf + kp + 1024 * q));
                float2 u[R];
                if constexpr (R == 2) {
                    u[0] = make_float2(V[0].x + V[1].x, V[0].y + V[1].y);
                    u[1] = make_float2(V[0].x - V[1].x, V[0].y - V[1].y);
                } else {                                          // inverse radix-4 (W_4^{-1} = +j)
                    const float2 s02 = make_float2(V[0].x + V[2].x, V[0].y + V[2].y), d02 = make_float2(V[0].x - V[2].x, V[0].y - V[2].y);
                    const float2 s13 = make_float2(V[1].x + V[3].x, V[1].y + V[3].y), d13 = make_float2(V[1].x - V[3].x, V[1].y - V[3].y);
                    u[0] = make_float2(s02.x + s13.x, s02.y + s13.y);
                    u[2] = make_float2(s02.x - s13.x, s02.y - s13.y);
                    u[1] = make_float2(d02.x - d13.y, d02.y + d13.x);       // d02 + j d13
                    u[3] = make_float2(d02.x + d13.y, d02.y - d13.x);       // d02 - j d13
                }
#pragma unroll
                for (int i = 0; i < R; ++i) {
                    if (i > 0) { const float2 w = wm(i * kp); u[i] = cmulf(u[i], make_float2(w.x, -w.y)); }
                    ebuf[i * 1024 + kp] = u[i];
                }
            }
            asm volatile("bar.sync %0, %1;" :: "r"(1 + group), "r"(R * 32) : "memory");
            {
                cf z[PW::BQ];
                static_for<0, PW::BQ>([&](auto kc) { constexpr int k2 = decltype(kc)::value; z[k2] = ebc[r * 1024 + k2 * 32 + lane]; });
                warp_cfft_inv_pk<PW>(z, xbuf, a.t.tw, lane, [&](auto, int, auto ac, int b, cf v) {
                    const int m = R * (PW::BQ * decltype(ac)::value + b) + r;
                    if (m < K) {
                        const cf w = ldg_cf(tp + m);
                        const cf o = fma2(rotp(v), bc(hi(w)), mul2(v, bc(lo(w))));
                        o0[m * TS + fr] = lo(o);
                        o1[m * TS + fr] = hi(o);
                    }
                });
            }
        }
        __syncthreads();
    } else {
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        float2* xbuf = reinterpret_cast<float2*>(work) + warp * kXbufFloat2;
        const int ntasks = (nf + P::NFR - 1) / P::NFR;
        for (;;) {
            int t = 0;
            if (lane == 0) t = atomicAdd(&s_next, 1);
            t = __shfl_sync(0xffffffffu, t, 0);
            if (t >= ntasks) break;
            const int f_first = f0 + t * P::NFR;
            cf z[P::BQ];
            const float2* __restrict__ win2 = a.t.win2;
            auto load = [&](auto, int fr, int m, auto) -> cf {
                int f = f_first + fr;
                f = f < f_end ? f : f_first;
                return mul2(*reinterpret_cast<const cf*>(tile + (f - f0) * a.hop + 2 * m), ldg_cf(win2 + m));
            };
            warp_fft_fwd_pk<P>(z, xbuf, a.t.tw, lane, load);
            const int myf = f_first + lane / P::A;
            const bool valid = myf < f_end;
            const int col = myf - f0;
            warp_fft_post_pk<P, true>(z, a.t.pt, lane, [&](int, int k, cf v, int) {
                if (valid) { o0[k * TS + col] = lo(v); o1[k * TS + col] = hi(v); }
            });
        }
        __syncthreads();
    }
    ex_write_out(a, o0, o1, TS, row, f0, nf, ft == a.ntiles - 1);
}

template <class P>
static int launch_ex(const ExArgs& a, long long rows, size_t smem, cudaStream_t st) {
    auto kern = stft_ex_kernel<P>;
    TAC_SMEM_OPT_IN(kern, 200 * 1024);
    kern<<<(unsigned)(rows * a.ntiles), kExThreads, smem, st>>>(a); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

// Second phase of librosa.amplitude_to_db(S, ref=np.max) (spectrogram.ipynb cell 9 L31-63): the row maxima come out of
// the STFT epilogue (tac_epilogue.rowmax), so the magnitude plane is read once.
__global__ void __launch_bounds__(256) db_from_rowmax_kernel(const float* __restrict__ mag, long long row_stride, int K, int F, long long ld,
                                                            const unsigned* __restrict__ rowmax, float amin, float top_db,
                                                            float* __restrict__ out) {
    const long long row = blockIdx.y;
    const float* m = mag + row * row_stride;
    float* o = out + row * (long long)K * F;
    const float ref = 20.0f * log10f(fmaxf(__uint_as_float(rowmax[row]), amin));
    const long long total = (long long)K * F;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
        const long long k = i / F, t = i - k * F;
        o[i] = fmaxf(20.0f * log10f(fmaxf(__ldg(m + k * ld + t), amin)) - ref, -top_db);
    }
}

}  // namespace tac

using namespace tac;

extern "C" int tac_stft_ex_supported(int n_fft) {
    return n_fft >= 8 && n_fft + n_fft / 2 <= kBluMaxM ? 1 : 0;
}

extern "C" int tac_db_from_rowmax(const float* mag, int64_t rows, int64_t row_stride, int K, int F, int64_t ld, const unsigned* rowmax,
                                  float amin, float top_db, float* out, tac_stream_t stream) {
    TAC_REQUIRE(mag && out && rowmax, "null argument");
    TAC_REQUIRE(rows >= 1 && rows < 65536 && K >= 1 && F >= 1 && ld >= F, "bad shape (rows < 65536, ld >= F)");
    TAC_REQUIRE(amin > 0.0f && top_db >= 0.0f, "amin must be positive, top_db non-negative");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    long long gx = ((long long)K * F + 255) / 256;
    gx = gx < 1 ? 1 : (gx > 64 ? 64 : gx);
    db_from_rowmax_kernel<<<dim3((unsigned)gx, (unsigned)rows), 256, 0, st>>>(mag, row_stride, K, F, ld, rowmax, amin, top_db, out); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

extern "C" int tac_prepare_stft_ex(int n_fft, int win, int normalized, int window) {
    TAC_REQUIRE(tac_stft_ex_supported(n_fft), "n_fft=%d unsupported by the general STFT (8 <= n_fft, n_fft + n_fft/2 <= %d)", n_fft, kBluMaxM);
    TAC_REQUIRE(win >= 1 && win <= n_fft, "win_length %d must be in [1, n_fft]", win);
    TAC_REQUIRE(window == 0 || window == 1, "unknown window kind %d (0 hann, 1 sqrt-hann)", window);
    FftTables t;
    int rc = TAC_OK;
    if (window == 0 && supported_nfft(n_fft)) {                      // even hops run on the plain core, odd ones on chirp-z: both
        rc = get_tables(n_fft, win, normalized, &t);
        if (rc) return rc;
    }
    BluTables bt;
    rc = get_blu_tables(n_fft, win, normalized, window, &bt);
    if (rc) return rc;
    return get_tables(2048, 2048, 0, &t);
}

extern "C" int tac_stft_ex(const float* x, int64_t rows, int64_t T, int n_fft, int hop, int win, int window, int center,
                           int pad_mode, int normalized, const tac_epilogue* epi, int64_t f_pad, void* out, tac_stream_t stream) {
    TAC_REQUIRE(x && epi && out, "null argument");
    TAC_REQUIRE(rows > 0 && T > 0 && rows * T < (1ll << 40), "bad shape rows=%lld T=%lld", (long long)rows, (long long)T);
    TAC_REQUIRE(tac_stft_ex_supported(n_fft), "unsupported n_fft %d (8 <= n_fft, n_fft + n_fft/2 <= %d)", n_fft, kBluMaxM);
    TAC_REQUIRE(hop >= 1 && hop <= n_fft, "hop %d must be in [1, n_fft]", hop);
    if (win <= 0) win = n_fft;
    TAC_REQUIRE(win <= n_fft, "win_length %d must be in [1, n_fft]", win);
    TAC_REQUIRE(window == 0 || window == 1, "unknown window kind %d (0 hann, 1 sqrt-hann)", window);
    TAC_REQUIRE(pad_mode == TAC_PAD_ZERO || pad_mode == TAC_PAD_REFLECT, "unknown pad_mode %d", pad_mode);
    TAC_REQUIRE(epi->kind >= TAC_EPI_COMPLEX && epi->kind <= TAC_EPI_MEL, "unknown epilogue kind %d", epi->kind);
    TAC_REQUIRE((reinterpret_cast<uintptr_t>(x) & 3u) == 0 && (reinterpret_cast<uintptr_t>(out) & 7u) == 0, "x must be 4-byte, out 8-byte aligned");
    if (center && pad_mode == TAC_PAD_REFLECT) TAC_REQUIRE(T > n_fft / 2, "reflect padding needs T > n_fft/2");
    if (epi->kind == TAC_EPI_MEL)
        TAC_REQUIRE(epi->n_mels >= 1 && epi->mel_fb && epi->mel_range, "mel epilogue needs n_mels, mel_fb [K][n_mels] and mel_range [n_mels][2]");
    if (epi->kind == TAC_EPI_DB || epi->kind == TAC_EPI_MAGDB) TAC_REQUIRE(epi->p0 > 0.0f, "dB epilogue: p0 (amin) must be positive");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);

    ExArgs a;
    memset(&a, 0, sizeof(a));
    a.x = x; a.T = T; a.n = n_fft; a.hop = hop; a.pad = center ? n_fft / 2 : 0;
    const int64_t F = num_frames(T, n_fft, hop, center);
    TAC_REQUIRE(F >= 1, "signal of %lld samples is shorter than n_fft=%d", (long long)T, n_fft);
    TAC_REQUIRE(F < (1ll << 30), "too many frames");
    a.F = (int)F;
    a.K = n_fft / 2 + 1;
    a.pad_mode = pad_mode;
    a.vec_ok = ((reinterpret_cast<uintptr_t>(x) & 15u) == 0 && (T & 3) == 0) ? 1 : 0;
    a.f_pad = f_pad > 0 ? f_pad : F;
    if (a.f_pad < a.F) a.F = (int)a.f_pad;                                    // crop (pad_tensor with negative padding)
    a.epi = *epi;
    a.out = out;
    const bool core = window == 0 && supported_nfft(n_fft) && (hop & 1) == 0;
    size_t work = 0;
    if (core) {
        int rc = get_tables(n_fft, win, normalized, &a.t);
        if (rc) return rc;
        work = (size_t)kExWarps * kXbufFloat2 * sizeof(float2);
    } else {
        cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
        TAC_CHECK_CUDA(cudaStreamIsCapturing(st, &cap));
        int rc = get_blu_tables(n_fft, win, normalized, window, &a.bt, cap == cudaStreamCaptureStatusNone);
        if (rc) return rc;
        rc = get_tables(2048, 2048, 0, &a.t);                                 // inter-pass twiddles of the 1024-point complex transform
        if (rc) return rc;
        work = (size_t)kExWarps * kXbufFloat2 * sizeof(float2);               // one warp per 1024-point (sub-)transform
        if (a.bt.M > 1024) work += (size_t)kExWarps * 1024 * sizeof(float2);  // + the groups' [R][1024] buffers
    }
    size_t smem = 0;
    for (a.TF = a.K > 512 ? 8 : (a.K > 128 ? 16 : 32); a.TF >= 1; a.TF >>= 1) {
        a.tile_len = ((a.TF - 1) * hop + n_fft + 3) & ~3;
        smem = (size_t)a.tile_len * 4 + ((2 * (size_t)a.K * (a.TF + 1) + 3) & ~(size_t)3) * 4 + work;
        if (smem <= 200 * 1024) break;
    }
    TAC_REQUIRE(a.TF >= 1, "configuration needs %zu bytes of shared memory", smem);
    a.ntiles = (a.F + a.TF - 1) / a.TF;
    TAC_REQUIRE(rows * a.ntiles < (1ll << 31), "grid too large");
    if (!core) return a.bt.M == 1024 ? launch_ex<WBluPlan>(a, rows, smem, st)
                    : (a.bt.M == 2048 ? launch_ex<WBluSplit<2>>(a, rows, smem, st) : launch_ex<WBluSplit<4>>(a, rows, smem, st));
    switch (n_fft) {
        case 2048: return launch_ex<Plan2048>(a, rows, smem, st);
        case 1024: return launch_ex<Plan1024>(a, rows, smem, st);
        case 512: return launch_ex<Plan512>(a, rows, smem, st);
        case 256: return launch_ex<Plan256>(a, rows, smem, st);
        case 128: return launch_ex<Plan128>(a, rows, smem, st);
        default: return launch_ex<Plan64>(a, rows, smem, st);
    }
}

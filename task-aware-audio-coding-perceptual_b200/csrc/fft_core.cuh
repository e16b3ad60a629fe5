// Warp-level real-input FFT building blocks for the STFT kernels (sm_100a).
//
// One warp transforms a "task" of NFR real frames of length n = 2*M at once.  A frame is packed
// as M complex points z[m] = x[2m] + i*x[2m+1]; the M-point complex DFT is split M = A * BQ:
//
//   pass 1  lane <-> (frame, b):  A-point DFT over a of z[BQ*a + b]          (registers)
//   twiddle W_M^(b*k1), exchange through a padded per-warp shared-memory tile (one round trip)
//   pass 2  lane <-> (frame, k1): BQ-point DFT over b  -> Z[k1 + A*k2]       (registers)
//   post    pairs (k, M-k) via one warp shuffle -> X[k], k = 0..M (one-sided real spectrum)
//
// A*NFR == 32 so that every lane owns exactly one pass-2 sub-transform; BQ % A == 0 so that
// every lane owns BQ/A pass-1 sub-transforms.  The butterflies are fully unrolled radix-2 DIF
// networks with compile-time twiddles (outputs land bit-reversed in the register array, which is
// a free renaming).  The adjoint (for the loss backward) runs the same network backwards.
//
// All tables (window*scale, pass twiddles, post twiddles) are generated on the host in float64.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <type_traits>

namespace tac {

// ------------------------------------------------------------------ compile-time trigonometry
constexpr double kPi = 3.14159265358979323846264338327950288;

__host__ __device__ constexpr double cx_sin_small(double x) {   // |x| <= pi/4
    double x2 = x * x, term = x, sum = x;
    for (int i = 1; i < 12; ++i) { term *= -x2 / ((2 * i) * (2 * i + 1)); sum += term; }
    return sum;
}
__host__ __device__ constexpr double cx_cos_small(double x) {
    double x2 = x * x, term = 1.0, sum = 1.0;
    for (int i = 1; i < 12; ++i) { term *= -x2 / ((2 * i - 1) * (2 * i)); sum += term; }
    return sum;
}
// cos / sin of 2*pi*i/n for 0 <= i <= n/2, exact symmetries first
__host__ __device__ constexpr double cx_cos2pi(int i, int n) {
    if (8 * i <= n) return cx_cos_small(2.0 * kPi * i / n);
    if (4 * i <= n) return cx_sin_small(2.0 * kPi * (n - 4 * i) / (4.0 * n));
    if (8 * i <= 3 * n) return -cx_sin_small(2.0 * kPi * (4 * i - n) / (4.0 * n));
    return -cx_cos_small(2.0 * kPi * (n - 2 * i) / (2.0 * n));
}
__host__ __device__ constexpr double cx_sin2pi(int i, int n) {
    if (8 * i <= n) return cx_sin_small(2.0 * kPi * i / n);
    if (4 * i <= n) return cx_cos_small(2.0 * kPi * (n - 4 * i) / (4.0 * n));
    if (8 * i <= 3 * n) return cx_cos_small(2.0 * kPi * (4 * i - n) / (4.0 * n));
    return cx_sin_small(2.0 * kPi * (n - 2 * i) / (2.0 * n));
}

__host__ __device__ constexpr int cx_log2(int n) { int l = 0; while ((1 << l) < n) ++l; return l; }
__host__ __device__ constexpr int brev(int k, int n) {
    int r = 0;
    for (int b = cx_log2(n); b > 0; --b) { r = (r << 1) | (k & 1); k >>= 1; }
    return r;
}

// Compile-time loop: f(std::integral_constant<int, I>) for I in [0, N).
template <int I, int N, class F>
__device__ __forceinline__ void static_for(F&& f) {
    if constexpr (I < N) {
        f(std::integral_constant<int, I>{});
        static_for<I + 1, N>(f);
    }
}

// ------------------------------------------------------------------ radix-2 DIF network
// (r, i) *= W_N^I = exp(-2*pi*j*I/N), 0 <= I < N/2, with the trivial cases folded away.
template <int N, int I>
__device__ __forceinline__ void twmul(float& r, float& i) {
    if constexpr (I == 0) {
    } else if constexpr (4 * I == N) {            // * (-j)
        float t = r; r = i; i = -t;
    } else if constexpr (8 * I == N) {            // * (1 - j)/sqrt2
        constexpr float h = 0.70710678118654752440f;
        float t = (r + i) * h; i = (i - r) * h; r = t;
    } else if constexpr (8 * I == 3 * N) {        // * (-1 - j)/sqrt2
        constexpr float h = 0.70710678118654752440f;
        float t = (i - r) * h; i = -(r + i) * h; r = t;
    } else {
        constexpr float c = (float)cx_cos2pi(I, N);
        constexpr float s = (float)cx_sin2pi(I, N);
        float t = fmaf(i, s, r * c);               // (r + j i)(c - j s)
        i = fmaf(-r, s, i * c);
        r = t;
    }
}

template <int N, int OFF, int I>
struct DifRow {
    template <class Arr>
    static __device__ __forceinline__ void run(Arr& re, Arr& im) {
        constexpr int H = N / 2;
        float ar = re[OFF + I], ai = im[OFF + I], br = re[OFF + I + H], bi = im[OFF + I + H];
        re[OFF + I] = ar + br; im[OFF + I] = ai + bi;
        float dr = ar - br, di = ai - bi;
        twmul<N, I>(dr, di);
        re[OFF + I + H] = dr; im[OFF + I + H] = di;
        if constexpr (I + 1 < H) DifRow<N, OFF, I + 1>::run(re, im);
    }
};

// In-place forward DFT of re/im[OFF .. OFF+N): natural-order input, bit-reversed output
// (X[k] ends in slot OFF + brev(k, N)).  Passing (im, re) instead of (re, im) gives the
// unnormalised inverse DFT.
template <int N, int OFF = 0>
struct Dif {
    template <class Arr>
    static __device__ __forceinline__ void run(Arr& re, Arr& im) {
        if constexpr (N >= 2) {
            DifRow<N, OFF, 0>::run(re, im);
            Dif<N / 2, OFF>::run(re, im);
            Dif<N / 2, OFF + N / 2>::run(re, im);
        }
    }
};

// ------------------------------------------------------------------ plans
template <int A_, int BQ_>
struct Plan {
    static constexpr int A = A_;            // pass-1 radix
    static constexpr int BQ = BQ_;          // pass-2 radix = complex points per lane
    static constexpr int M = A_ * BQ_;      // complex points per frame
    static constexpr int N = 2 * M;         // real frame length (n_fft)
    static constexpr int NFR = 32 / A_;     // frames per warp task
    static constexpr int NS1 = BQ_ / A_;    // pass-1 sub-transforms per lane
#ifdef TAC_XBUF_ODD
    static constexpr int RS = BQ_ + 1;      // exchange row stride (float2), odd -> conflict free
#else
    // exchange row stride (float2): BQ + 2 keeps every row 16-byte aligned so that a lane moves its row with 128-bit
    // accesses (quarter-warp bank pattern 4*lane mod 32: conflict free); the small plans keep the odd stride, which is
    // what makes their 64-bit column accesses conflict free when a half-warp spans two frames
    static constexpr int RS = BQ_ >= 16 ? BQ_ + 2 : BQ_ + 1;
#endif
    static constexpr bool kVec4 = (RS % 2) == 0;
    static_assert(32 % A_ == 0 && BQ_ % A_ == 0 && BQ_ >= A_, "unsupported split");
};
using Plan2048 = Plan<32, 32>;
using Plan1024 = Plan<16, 32>;
using Plan512 = Plan<16, 16>;
using Plan256 = Plan<8, 16>;
using Plan128 = Plan<8, 8>;
using Plan64 = Plan<4, 8>;

constexpr int kXbufFloat2 = 32 * 34;        // per-warp exchange tile (largest plan), 8704 B

struct FftTables {                          // device pointers, one set per (n_fft, win, normalized)
    const float2* win2;                     // [M]   (w[2m], w[2m+1]) * 0.5 * norm
    const float2* tw;                       // [A][BQ] (cos, sin)(2*pi*b*k1/M) at k1*BQ + b
    const float2* pt;                       // [M]   (cos, sin)(2*pi*k/n)
};

__device__ __forceinline__ float2 ldg2(const float2* p) { return __ldg(p); }

// ------------------------------------------------------------------ forward transform
// load(integral_constant s, fr, m) returns the windowed packed sample z_fr[m] (float2); s is the lane's pass-1
// sub-transform index (compile-time), fr = (lane + 32 s) / BQ the frame of the task it belongs to.  On return lane
// (fr = lane / A, k1 = lane % A) holds Z_fr[k1 + A*k2] in slot brev(k2, BQ) of (zr, zi).
template <class P, class Load>
__device__ __forceinline__ void warp_fft_fwd(float (&zr)[P::BQ], float (&zi)[P::BQ], float2* xbuf,
                                             const float2* __restrict__ tw, int lane, Load load) {
    constexpr int A = P::A, BQ = P::BQ, RS = P::RS;
    static_for<0, P::NS1>([&](auto sc) {
        constexpr int s = decltype(sc)::value;
        const int q = lane + 32 * s;
        const int fr = q / BQ, b = q % BQ;
        float ar[A], ai[A];
#pragma unroll
        for (int a = 0; a < A; ++a) {
            float2 v = load(sc, fr, BQ * a + b);
            ar[a] = v.x; ai[a] = v.y;
        }
        Dif<A>::run(ar, ai);
        float2* row = xbuf + (fr * A) * RS + b;
#pragma unroll
        for (int k1 = 0; k1 < A; ++k1) {
            float r = ar[brev(k1, A)], i = ai[brev(k1, A)];
            if (k1 > 0) {
                float2 w = ldg2(tw + k1 * BQ + b);          // W_M^(b*k1) = w.x - j w.y
                float t = fmaf(i, w.y, r * w.x);
                i = fmaf(-r, w.y, i * w.x);
                r = t;
            }
            row[k1 * RS] = make_float2(r, i);
        }
    });
    __syncwarp();
    const float2* rd = xbuf + lane * RS;
    if constexpr (P::kVec4) {
        const float4* rd4 = reinterpret_cast<const float4*>(rd);
#pragma unroll
        for (int b = 0; b < BQ / 2; ++b) {
            const float4 v = rd4[b];
            zr[2 * b] = v.x; zi[2 * b] = v.y; zr[2 * b + 1] = v.z; zi[2 * b + 1] = v.w;
        }
    } else {
#pragma unroll
        for (int b = 0; b < BQ; ++b) {
            float2 v = rd[b];
            zr[b] = v.x; zi[b] = v.y;
        }
    }
    __syncwarp();
    Dif<BQ>::run(zr, zi);
}

// (cos, sin)(2 pi (k1 + A j) / n) from the per-lane base (cos, sin)(2 pi k1 / n) and the compile-time rotation
// by 2 pi j / (2 BQ)  (n = 2 A BQ): four FP instructions instead of a dependent table load per bin pair.
template <int BQ, int J>
__device__ __forceinline__ float2 post_twiddle(const float2 base) {
    if constexpr (J == 0) {
        return base;
    } else {
        constexpr float cj = (float)cx_cos2pi(J, 2 * BQ), sj = (float)cx_sin2pi(J, 2 * BQ);
        return make_float2(fmaf(base.x, cj, -base.y * sj), fmaf(base.y, cj, base.x * sj));
    }
}

// Post-process pairs (k, M-k) of the packed transform into the one-sided real spectrum.
// emit(fr, k, re, im, slot) is called exactly once for every k in [0, M] of every frame of the task;
// slot is a compile-time-resolvable index in [0, BQ] (2j, 2j+1 for pair j; BQ for the k = M/2 bin of
// the k1 == 0 lanes) that callers use to keep per-bin state in registers.
template <class P, class Emit>
__device__ __forceinline__ void warp_fft_post(const float (&zr)[P::BQ], const float (&zi)[P::BQ],
                                              const float2* __restrict__ pt, int lane, Emit emit) {
    constexpr int A = P::A, BQ = P::BQ, M = P::M;
    const int k1 = lane % A, fr = lane / A;
    const int plane = fr * A + ((A - k1) % A);
    const float2 base = ldg2(pt + k1);                      // W_n^k1; W_n^(k1 + A j) = base * W_(2 BQ)^j (compile-time)
    static_for<0, BQ / 2>([&](auto jc) {
        constexpr int j = decltype(jc)::value;
        const float mr = zr[brev(j, BQ)], mi = zi[brev(j, BQ)];
        float pr = __shfl_sync(0xffffffffu, zr[brev(BQ - 1 - j, BQ)], plane);
        float pi = __shfl_sync(0xffffffffu, zi[brev(BQ - 1 - j, BQ)], plane);
        if (k1 == 0) { pr = zr[brev((BQ - j) % BQ, BQ)]; pi = zi[brev((BQ - j) % BQ, BQ)]; }
        const float sr = mr + pr, si = mi - pi;             // Z[k] + conj(Z[M-k])
        const float dr = mr - pr, di = mi + pi;             // Z[k] - conj(Z[M-k])
        const int k = k1 + A * j;
        const float2 cs = post_twiddle<BQ, j>(base);
        const float tr = fmaf(cs.x, di, -cs.y * dr);        // T = -j W_n^k (Z[k] - conj(Z[M-k]))
        const float ti = -fmaf(cs.x, dr, cs.y * di);
        emit(fr, k, sr + tr, si + ti, 2 * j);
        emit(fr, M - k, sr - tr, ti - si, 2 * j + 1);
    });
    if (k1 == 0) emit(fr, M / 2, 2.0f * zr[brev(BQ / 2, BQ)], -2.0f * zi[brev(BQ / 2, BQ)], BQ);
}

// ------------------------------------------------------------------ adjoint transform
// grad(fr, k, slot) returns dL/dX_fr[k] (float2, re/im parts) for k in [0, M]; it is called exactly
// once per (frame, k) by the lane that emitted X_fr[k] in warp_fft_post, with the same slot.  store(fr, m, g)
// receives dL/dz_fr[m], m = BQ * a + b, *before* the window (caller multiplies by win2[m]); it is called as
// store(integral_constant s, f2, integral_constant a, b, value) with s the pass-1 sub-transform index of the lane.
template <class P, class Grad, class Store>
__device__ __forceinline__ void warp_fft_adj(float2* xbuf, const float2* __restrict__ tw,
                                             const float2* __restrict__ pt, int lane, Grad grad, Store store) {
    constexpr int A = P::A, BQ = P::BQ, M = P::M, RS = P::RS;
    const int k1 = lane % A, fr = lane / A;
    const int plane = fr * A + ((A - k1) % A);
    float zr[BQ], zi[BQ];
    float qr[BQ / 2], qi[BQ / 2];                           // gradients owed to the partner lane
    const float2 base = ldg2(pt + k1);
    static_for<0, BQ / 2>([&](auto jc) {
        constexpr int j = decltype(jc)::value;
        const int k = k1 + A * j;
        const float2 g = grad(fr, k, 2 * j);
        const float2 gp = grad(fr, M - k, 2 * j + 1);
        const float sr = g.x + gp.x, si = g.y - gp.y;       // G[k] + conj(G[M-k])
        const float dr = g.x - gp.x, di = g.y + gp.y;       // G[k] - conj(G[M-k])
        const float2 cs = post_twiddle<BQ, j>(base);
        // T' = conj(u) * D with u = -j W_n^k = (-s, -c)  ->  conj(u) = (-s, +c)
        const float tr = -fmaf(cs.x, di, cs.y * dr);
        const float ti = fmaf(cs.x, dr, -cs.y * di);
        zr[j] = sr + tr; zi[j] = si + ti;                   // dL/dZ[k]
        qr[j] = sr - tr; qi[j] = ti - si;                   // dL/dZ[M-k] = conj(S - T')
    });
#pragma unroll
    for (int j = 0; j < BQ / 2; ++j) {
        // my slot BQ-1-j receives what the partner lane computed for its pair j
        float rr = __shfl_sync(0xffffffffu, qr[j], plane);
        float ri = __shfl_sync(0xffffffffu, qi[j], plane);
        zr[BQ - 1 - j] = rr; zi[BQ - 1 - j] = ri;
    }
    if (k1 == 0) {
        // lanes with k1 == 0 pair inside themselves: slot j <-> slot (BQ - j) % BQ
        const float2 gh = grad(fr, M / 2, BQ);
        float t0r = zr[0] + qr[0], t0i = zi[0] + qi[0];     // Z[0] feeds both X[0] and X[M]
#pragma unroll
        for (int j = 1; j < BQ / 2; ++j) { zr[BQ - j] = qr[j]; zi[BQ - j] = qi[j]; }
        zr[0] = t0r; zi[0] = t0i;
        zr[BQ / 2] = 2.0f * gh.x; zi[BQ / 2] = -2.0f * gh.y;
    }
    // inverse BQ-point DFT over k2 (natural in, bit-reversed out): swap re/im roles
    Dif<BQ>::run(zi, zr);
    float2* wr = xbuf + lane * RS;
    float pr_ = 0.0f, pi_ = 0.0f;
#pragma unroll
    for (int b = 0; b < BQ; ++b) {
        float r = zr[brev(b, BQ)], i = zi[brev(b, BQ)];
        const float2 w = ldg2(tw + k1 * BQ + b);            // conj(W_M^(b*k1)) = w.x + j w.y
        if (b > 0) {
            float t = fmaf(-i, w.y, r * w.x);
            i = fmaf(r, w.y, i * w.x);
            r = t;
        }
        if constexpr (P::kVec4) {
            if (b & 1) reinterpret_cast<float4*>(wr)[b >> 1] = make_float4(pr_, pi_, r, i);
            else { pr_ = r; pi_ = i; }
        } else {
            wr[b] = make_float2(r, i);
        }
    }
    __syncwarp();
    static_for<0, P::NS1>([&](auto sc) {
        constexpr int s = decltype(sc)::value;
        const int q = lane + 32 * s;
        const int f2 = q / BQ, b = q % BQ;
        float ar[A], ai[A];
        const float2* col = xbuf + (f2 * A) * RS + b;
#pragma unroll
        for (int kk = 0; kk < A; ++kk) {
            float2 v = col[kk * RS];
            ar[kk] = v.x; ai[kk] = v.y;
        }
        Dif<A>::run(ai, ar);
        // sample pair m = BQ * a + b of frame f2; s and a are compile-time (integral constants) for the callee
        static_for<0, A>([&](auto ac) {
            constexpr int a = decltype(ac)::value;
            store(sc, f2, ac, b, make_float2(ar[brev(a, A)], ai[brev(a, A)]));
        });
    });
    __syncwarp();
}

}  // namespace tac

// Packed-complex form of the warp-level real-input FFT (fft_core.cuh) for sm_100a.
//
// Blackwell has two-wide FP32 instructions (FADD2 / FMUL2 / FFMA2; PTX add/mul/fma.rn.f32x2) that operate on an
// aligned register pair, take ONE issue slot, have the latency of the scalar instruction (4 cycles, measured:
// tools/ubench/f32x2_bench.cu) and accept per-operand modifiers -- swap of the halves (.LO_HI), negation of one half
// (.NP / .PN), broadcast of a 32-bit register or immediate to both halves.  A complex number held as (re, im) in such
// a pair therefore costs one instruction per complex add, and  a +- W b  (W a compile-time twiddle) costs THREE:
//
//     W b = c (b + t rot(b)),  rot(b) = -j b = (b.im, -b.re),  t = s / c        (or the s-factored form when |s| > |c|)
//     u = fma2(rot(b), t, b);   a' = fma2(u, +c, a);   b' = fma2(u, -c, a)
//
// against eight scalar instructions for the radix-2 butterfly of fft_core.cuh.  The STFT loss kernel is bound by
// instruction issue (ncu: profiles/ncu_full_mrstft_stream_kernel_r01f.txt), not by the FP32 pipe, so halving the
// instruction stream is what shortens it.
//
// Same decomposition, lane mappings, tables and exchange tile as fft_core.cuh; the differences are
//   * radix-2 decimation in TIME (input slot brev(i) holds x[i], output slot k holds X[k]): the twiddle sits on the
//     butterfly input, which is what allows the three-instruction form;
//   * post-processing hands the caller conj(X[M-k]) in the odd slots (the sign flip is free on the consumer side and
//     the adjoint wants exactly conj(G[M-k])).
#pragma once
#include "fft_core.cuh"

namespace tac {

typedef unsigned long long cf;          // packed complex: re in the low half, im in the high half of a register pair

__device__ __forceinline__ cf pk(float a, float b) { cf r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float lo(cf v) { float a, b; asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); (void)b; return a; }
__device__ __forceinline__ float hi(cf v) { float a, b; asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); (void)a; return b; }
__device__ __forceinline__ cf fma2(cf a, cf b, cf c) { cf r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ cf mul2(cf a, cf b) { cf r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ cf add2(cf a, cf b) { cf r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ cf sub2(cf a, cf b) { cf r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
// ptxas folds these three into operand modifiers of the consuming packed instruction (checked in the SASS)
__device__ __forceinline__ cf bc(float s) { return pk(s, s); }                      // broadcast
__device__ __forceinline__ cf rotm(cf v) { return pk(hi(v), -lo(v)); }             // -j v
__device__ __forceinline__ cf rotp(cf v) { return pk(-hi(v), lo(v)); }             // +j v
__device__ __forceinline__ cf cnj(cf v) { return pk(lo(v), -hi(v)); }              // conj(v)

__device__ __forceinline__ cf ldg_cf(const float2* p) { return __ldg(reinterpret_cast<const cf*>(p)); }

// ------------------------------------------------------------------ radix-2 DIT network
// a' = a + W b, b' = a - W b,  W = exp(-+ 2 pi j K / N) (upper sign: forward), 0 <= K < N/2.
template <int N, int K, bool INV>
__device__ __forceinline__ void dit_bfly(cf& a, cf& b) {
    if constexpr (K == 0) {
        const cf t = b;
        b = sub2(a, t); a = add2(a, t);
    } else if constexpr (4 * K == N) {
        const cf r = INV ? rotp(b) : rotm(b);
        b = sub2(a, r); a = add2(a, r);
    } else {
        constexpr double c = cx_cos2pi(K, N), s = cx_sin2pi(K, N);     // s > 0
        constexpr double sg = INV ? -1.0 : 1.0;                        // W b = c b + sg s rot(b)
        if constexpr ((c >= 0 ? c : -c) >= s) {
            constexpr float t = (float)(sg * s / c), cc = (float)c;
            const cf u = fma2(rotm(b), bc(t), b);
            b = fma2(u, bc(-cc), a); a = fma2(u, bc(cc), a);
        } else {                                                       // W b = sg s rot(u), u = b - sg (c / s) rot(b)
            constexpr float r = (float)(-sg * c / s), ss = (float)(sg * s);
            const cf u = fma2(rotm(b), bc(r), b);
            const cf ru = rotm(u);
            b = fma2(ru, bc(-ss), a); a = fma2(ru, bc(ss), a);
        }
    }
}

// In-place DFT of z[OFF .. OFF+N): slot OFF + brev(i, N) holds x[i] on entry, slot OFF + k holds X[k] on return
// (INV: the unnormalised inverse).
template <int N, bool INV, int OFF = 0>
struct Dit {
    template <class Arr>
    static __device__ __forceinline__ void run(Arr& z) {
        if constexpr (N >= 2) {
            Dit<N / 2, INV, OFF>::run(z);
            Dit<N / 2, INV, OFF + N / 2>::run(z);
            static_for<0, N / 2>([&](auto kc) {
                constexpr int k = decltype(kc)::value;
                dit_bfly<N, k, INV>(z[OFF + k], z[OFF + k + N / 2]);
            });
        }
    }
};

// ------------------------------------------------------------------ forward transform
// load(integral_constant s, fr, m, integral_constant a) returns the windowed packed sample pair z_fr[m] = (x[2m], x[2m+1]),
// m = BQ a + b; the compile-time a tells a hop = n/4 caller which quarter of the frame it reads (cache policy).
// On return lane
// (fr = lane / A, k1 = lane % A) holds Z_fr[k1 + A*k2] in slot k2 of z.
template <class P, class Load>
__device__ __forceinline__ void warp_fft_fwd_pk(cf (&z)[P::BQ], float2* xbuf, const float2* __restrict__ tw, int lane,
                                                Load load) {
    constexpr int A = P::A, BQ = P::BQ, RS = P::RS;
    static_for<0, P::NS1>([&](auto sc) {
        constexpr int s = decltype(sc)::value;
        const int q = lane + 32 * s;
        const int fr = q / BQ, b = q % BQ;
        cf a[A];
        static_for<0, A>([&](auto ac) {
            constexpr int av = decltype(ac)::value;
            a[brev(av, A)] = load(sc, fr, BQ * av + b, ac);
        });
        // Inter-pass twiddles W_M^(b k1), k1 = 8 q + r, as W^(b r) W^(8 b q): 7 + A/8 - 1 table reads per sub-transform instead
        // of A - 1 (the loss kernel is bound by the L1TEX data pipe, and the tables were 38 % of its bytes), at two more
        // packed instructions for the elements that need both factors.
        float2 wr_[8], wq_[A / 8 > 0 ? A / 8 : 1];
        static_for<1, (A < 8 ? A : 8)>([&](auto rc) { constexpr int r = decltype(rc)::value; wr_[r] = ldg2(tw + r * BQ + b); });
        static_for<1, A / 8>([&](auto qc) { constexpr int q = decltype(qc)::value; wq_[q] = ldg2(tw + 8 * q * BQ + b); });
        Dit<A, false>::run(a);
        cf* row = reinterpret_cast<cf*>(xbuf) + (fr * A) * RS + b;
        static_for<0, A>([&](auto kc) {
            constexpr int k1 = decltype(kc)::value;
            constexpr int r = k1 % 8, q = k1 / 8;
            cf v = a[k1];
            if constexpr (r > 0) v = fma2(rotm(v), bc(wr_[r].y), mul2(v, bc(wr_[r].x)));     // W = w.x - j w.y
            if constexpr (q > 0) v = fma2(rotm(v), bc(wq_[q].y), mul2(v, bc(wq_[q].x)));
            row[k1 * RS] = v;
        });
    });
    __syncwarp();
    const cf* rd = reinterpret_cast<const cf*>(xbuf) + lane * RS;
    if constexpr (P::kVec4) {
        const ulonglong2* rd2 = reinterpret_cast<const ulonglong2*>(rd);
        static_for<0, BQ / 2>([&](auto bc_) {
            constexpr int b = decltype(bc_)::value;
            const ulonglong2 v = rd2[b];
            z[brev(2 * b, BQ)] = v.x; z[brev(2 * b + 1, BQ)] = v.y;
        });
    } else {
        static_for<0, BQ>([&](auto bc_) {
            constexpr int b = decltype(bc_)::value;
            z[brev(b, BQ)] = rd[b];
        });
    }
    __syncwarp();
    Dit<BQ, false>::run(z);
}

// (cos, sin)(2 pi (k1 + A j) / n) from the per-lane base (cos, sin)(2 pi k1 / n): base * exp(+j 2 pi J / (2 BQ)).
template <int BQ, int J>
__device__ __forceinline__ cf post_twiddle_pk(const cf base) {
    if constexpr (J == 0) {
        return base;
    } else {
        constexpr float cj = (float)cx_cos2pi(J, 2 * BQ), sj = (float)cx_sin2pi(J, 2 * BQ);
        return fma2(rotp(base), bc(sj), mul2(base, bc(cj)));
    }
}

// One-sided real spectrum from the packed transform.  emit(fr, k, v, slot) is called once for every k in [0, M]:
// EVEN slots 2j carry X[k], k = k1 + A j; ODD slots 2j + 1 carry conj(X[M - k]) (TRUE_ODD: X[M - k] itself, for callers
// that write the spectrum out); slot BQ (k1 == 0 lanes) carries X[M/2].
template <class P, bool TRUE_ODD = false, class Emit>
__device__ __forceinline__ void warp_fft_post_pk(const cf (&z)[P::BQ], const float2* __restrict__ pt, int lane, Emit emit) {
    constexpr int A = P::A, BQ = P::BQ, M = P::M;
    const int k1 = lane % A, fr = lane / A;
    const int plane = fr * A + ((A - k1) % A);
    const cf base = ldg_cf(pt + k1);
    static_for<0, BQ / 2>([&](auto jc) {
        constexpr int j = decltype(jc)::value;
        const cf m = z[j];
        cf p = __shfl_sync(0xffffffffu, z[BQ - 1 - j], plane);
        if (k1 == 0) p = z[(BQ - j) % BQ];
        const cf pc = cnj(p);
        const cf S = add2(m, pc), D = sub2(m, pc);          // Z[k] +- conj(Z[M-k])
        const cf cs = post_twiddle_pk<BQ, j>(base);
        // T = -j W_n^k D = cs.x rot(D) - cs.y D
        const cf T = fma2(rotm(D), bc(lo(cs)), mul2(D, bc(-hi(cs))));
        const int k = k1 + A * j;
        emit(fr, k, add2(S, T), 2 * j);
        if constexpr (TRUE_ODD) emit(fr, M - k, sub2(cnj(S), cnj(T)), 2 * j + 1);
        else emit(fr, M - k, sub2(S, T), 2 * j + 1);
    });
    if (k1 == 0) emit(fr, M / 2, mul2(cnj(z[BQ / 2]), bc(2.0f)), BQ);
}

// ------------------------------------------------------------------ two forward transforms at once
// The loss kernels transform the SAME frame positions of two signals (target x, decoded y).  Doing both in one call makes
// the butterflies of the two transforms independent instruction streams inside the same basic blocks (ILP 2 for a kernel
// that runs two warps per scheduler and waits on fixed-latency dependencies), and the window, the inter-pass twiddles,
// the post-processing twiddles and the exchange tile are loaded / computed ONCE for both: y's twiddled pass-1 outputs
// wait in registers while x goes through the tile.
// load(integral_constant s, fr, m, integral_constant a, vx, vy) fills both windowed packed sample pairs.
template <class P, class Load>
__device__ __forceinline__ void warp_fft_fwd2_pk(cf (&zx)[P::BQ], cf (&zy)[P::BQ], float2* xbuf, const float2* __restrict__ tw,
                                                 int lane, Load load) {
    constexpr int A = P::A, BQ = P::BQ, RS = P::RS, NS1 = P::NS1;
    cf ay[NS1][A];
    static_for<0, NS1>([&](auto sc) {
        constexpr int s = decltype(sc)::value;
        const int q = lane + 32 * s;
        const int fr = q / BQ, b = q % BQ;
        cf ax[A];
        static_for<0, A>([&](auto ac) {
            constexpr int av = decltype(ac)::value;
            load(sc, fr, BQ * av + b, ac, ax[brev(av, A)], ay[s][brev(av, A)]);
        });
        float2 wr_[8], wq_[A / 8 > 0 ? A / 8 : 1];
        static_for<1, (A < 8 ? A : 8)>([&](auto rc) { constexpr int r = decltype(rc)::value; wr_[r] = ldg2(tw + r * BQ + b); });
        static_for<1, A / 8>([&](auto qc) { constexpr int qq = decltype(qc)::value; wq_[qq] = ldg2(tw + 8 * qq * BQ + b); });
        Dit<A, false>::run(ax);
        Dit<A, false>::run(ay[s]);
        cf* row = reinterpret_cast<cf*>(xbuf) + (fr * A) * RS + b;
        static_for<0, A>([&](auto kc) {
            constexpr int k1 = decltype(kc)::value;
            constexpr int r = k1 % 8, qq = k1 / 8;
            cf vx = ax[k1], vy = ay[s][k1];
            if constexpr (r > 0) {
                vx = fma2(rotm(vx), bc(wr_[r].y), mul2(vx, bc(wr_[r].x)));     // W = w.x - j w.y
                vy = fma2(rotm(vy), bc(wr_[r].y), mul2(vy, bc(wr_[r].x)));
            }
            if constexpr (qq > 0) {
                vx = fma2(rotm(vx), bc(wq_[qq].y), mul2(vx, bc(wq_[qq].x)));
                vy = fma2(rotm(vy), bc(wq_[qq].y), mul2(vy, bc(wq_[qq].x)));
            }
            row[k1 * RS] = vx;
            ay[s][k1] = vy;
        });
    });
    __syncwarp();
    const cf* rd = reinterpret_cast<const cf*>(xbuf) + lane * RS;
    auto read_tile = [&](cf (&z)[BQ]) {
        if constexpr (P::kVec4) {
            const ulonglong2* rd2 = reinterpret_cast<const ulonglong2*>(rd);
            static_for<0, BQ / 2>([&](auto bc_) {
                constexpr int b = decltype(bc_)::value;
                const ulonglong2 v = rd2[b];
                z[brev(2 * b, BQ)] = v.x; z[brev(2 * b + 1, BQ)] = v.y;
            });
        } else {
            static_for<0, BQ>([&](auto bc_) {
                constexpr int b = decltype(bc_)::value;
                z[brev(b, BQ)] = rd[b];
            });
        }
    };
    read_tile(zx);
    __syncwarp();
    static_for<0, NS1>([&](auto sc) {
        constexpr int s = decltype(sc)::value;
        const int q = lane + 32 * s;
        const int fr = q / BQ, b = q % BQ;
        cf* row = reinterpret_cast<cf*>(xbuf) + (fr * A) * RS + b;
        static_for<0, A>([&](auto kc) { constexpr int k1 = decltype(kc)::value; row[k1 * RS] = ay[s][k1]; });
    });
    __syncwarp();
    read_tile(zy);
    __syncwarp();
    Dit<BQ, false>::run(zx);
    Dit<BQ, false>::run(zy);
}

// One-sided spectra of both transforms; emit(fr, k, X, Y, slot) with the slot convention of warp_fft_post_pk (odd slots carry
// the conjugates of X[M - k], Y[M - k]).  The post-processing twiddle of a slot is formed once for both.
template <class P, class Emit>
__device__ __forceinline__ void warp_fft_post2_pk(const cf (&zx)[P::BQ], const cf (&zy)[P::BQ], const float2* __restrict__ pt, int lane,
                                                  Emit emit) {
    constexpr int A = P::A, BQ = P::BQ, M = P::M;
    const int k1 = lane % A, fr = lane / A;
    const int plane = fr * A + ((A - k1) % A);
    const cf base = ldg_cf(pt + k1);
    static_for<0, BQ / 2>([&](auto jc) {
        constexpr int j = decltype(jc)::value;
        cf px = __shfl_sync(0xffffffffu, zx[BQ - 1 - j], plane);
        cf py = __shfl_sync(0xffffffffu, zy[BQ - 1 - j], plane);
        if (k1 == 0) { px = zx[(BQ - j) % BQ]; py = zy[(BQ - j) % BQ]; }
        const cf pxc = cnj(px), pyc = cnj(py);
        const cf Sx = add2(zx[j], pxc), Dx = sub2(zx[j], pxc);
        const cf Sy = add2(zy[j], pyc), Dy = sub2(zy[j], pyc);
        const cf cs = post_twiddle_pk<BQ, j>(base);
        const cf Tx = fma2(rotm(Dx), bc(lo(cs)), mul2(Dx, bc(-hi(cs))));
        const cf Ty = fma2(rotm(Dy), bc(lo(cs)), mul2(Dy, bc(-hi(cs))));
        const int k = k1 + A * j;
        emit(fr, k, add2(Sx, Tx), add2(Sy, Ty), 2 * j);
        emit(fr, M - k, sub2(Sx, Tx), sub2(Sy, Ty), 2 * j + 1);
    });
    if (k1 == 0) emit(fr, M / 2, mul2(cnj(zx[BQ / 2]), bc(2.0f)), mul2(cnj(zy[BQ / 2]), bc(2.0f)), BQ);
}

// ------------------------------------------------------------------ post-processing fused with the adjoint's first stage
// The loss kernels turn every spectrum bin into a gradient bin and feed it straight back into the adjoint, whose first stage
// pairs dL/dX[k] with conj(dL/dX[M-k]) exactly as the post-processing paired Z[k] with conj(Z[M-k]).  Doing both in ONE loop
// over the pairs means the gradient never exists as an array: the forward registers z[j], z[BQ-1-j] die as the adjoint
// registers zo[j], u[j] are born (a separate post + adjoint keeps 2 BQ + 2 more packed values alive between the two loops, which
// at 255 registers cost the streaming kernel its scheduling freedom and ~0.3 M local-memory instructions per launch), and the
// post-processing twiddle of a pair is formed once for both directions.
//   bin(fr, k, v, slot) -> gradient of that slot (even slots / slot BQ: dL/dX[k] given v = X[k]; odd slots: conj(dL/dX[M-k])
//   given v = conj(X[M-k])) -- for the magnitude losses that is simply v * g(|v|^2).
// On return zo[j] = dL/dZ[k1 + A j] and u[j] = conj(dL/dZ[M - k]) for j < BQ / 2, gh = gradient of bin M / 2 (k1 == 0 lanes):
// the inputs of warp_fft_adj_tail_pk.
template <class P, class Bin>
__device__ __forceinline__ void warp_fft_post_adj_head_pk(const cf (&z)[P::BQ], const float2* __restrict__ pt, int lane, Bin bin,
                                                          cf (&zo)[P::BQ], cf (&u)[P::BQ / 2], cf& gh) {
    constexpr int A = P::A, BQ = P::BQ, M = P::M;
    const int k1 = lane % A, fr = lane / A;
    const int plane = fr * A + ((A - k1) % A);
    const cf base = ldg_cf(pt + k1);
    static_for<0, BQ / 2>([&](auto jc) {
        constexpr int j = decltype(jc)::value;
        cf p = __shfl_sync(0xffffffffu, z[BQ - 1 - j], plane);
        if (k1 == 0) p = z[(BQ - j) % BQ];
        const cf pc = cnj(p);
        const cf S = add2(z[j], pc), D = sub2(z[j], pc);    // Z[k] +- conj(Z[M-k])
        const cf cs = post_twiddle_pk<BQ, j>(base);
        const cf T = fma2(rotm(D), bc(lo(cs)), mul2(D, bc(-hi(cs))));     // -j W_n^k D
        const int k = k1 + A * j;
        const cf g = bin(fr, k, add2(S, T), 2 * j);
        const cf gpc = bin(fr, M - k, sub2(S, T), 2 * j + 1);
        const cf Sg = add2(g, gpc), Dg = sub2(g, gpc);      // G[k] +- conj(G[M-k])
        const cf Tn = fma2(rotm(Dg), bc(lo(cs)), mul2(Dg, bc(hi(cs))));   // -conj(-j W_n^k) Dg
        zo[j] = sub2(Sg, Tn);                                // dL/dZ[k]
        u[j] = add2(Sg, Tn);                                 // conj(dL/dZ[M-k])
    });
    gh = 0ull;
    if (k1 == 0) gh = bin(fr, M / 2, mul2(cnj(z[BQ / 2]), bc(2.0f)), BQ);
}

// Two forward transforms (target x, decoded y) into ONE gradient: bin(fr, k, X, Y, slot) -> gradient slot of y.
template <class P, class Bin>
__device__ __forceinline__ void warp_fft_post2_adj_head_pk(const cf (&zx)[P::BQ], const cf (&zy)[P::BQ], const float2* __restrict__ pt,
                                                           int lane, Bin bin, cf (&zo)[P::BQ], cf (&u)[P::BQ / 2], cf& gh) {
    constexpr int A = P::A, BQ = P::BQ, M = P::M;
    const int k1 = lane % A, fr = lane / A;
    const int plane = fr * A + ((A - k1) % A);
    const cf base = ldg_cf(pt + k1);
    static_for<0, BQ / 2>([&](auto jc) {
        constexpr int j = decltype(jc)::value;
        cf px = __shfl_sync(0xffffffffu, zx[BQ - 1 - j], plane);
        cf py = __shfl_sync(0xffffffffu, zy[BQ - 1 - j], plane);
        if (k1 == 0) { px = zx[(BQ - j) % BQ]; py = zy[(BQ - j) % BQ]; }
        const cf pxc = cnj(px), pyc = cnj(py);
        const cf Sx = add2(zx[j], pxc), Dx = sub2(zx[j], pxc);
        const cf Sy = add2(zy[j], pyc), Dy = sub2(zy[j], pyc);
        const cf cs = post_twiddle_pk<BQ, j>(base);
        const cf Tx = fma2(rotm(Dx), bc(lo(cs)), mul2(Dx, bc(-hi(cs))));
        const cf Ty = fma2(rotm(Dy), bc(lo(cs)), mul2(Dy, bc(-hi(cs))));
        const int k = k1 + A * j;
        const cf g = bin(fr, k, add2(Sx, Tx), add2(Sy, Ty), 2 * j);
        const cf gpc = bin(fr, M - k, sub2(Sx, Tx), sub2(Sy, Ty), 2 * j + 1);
        const cf Sg = add2(g, gpc), Dg = sub2(g, gpc);
        const cf Tn = fma2(rotm(Dg), bc(lo(cs)), mul2(Dg, bc(hi(cs))));
        zo[j] = sub2(Sg, Tn);
        u[j] = add2(Sg, Tn);
    });
    gh = 0ull;
    if (k1 == 0) gh = bin(fr, M / 2, mul2(cnj(zx[BQ / 2]), bc(2.0f)), mul2(cnj(zy[BQ / 2]), bc(2.0f)), BQ);
}

// Unnormalised inverse complex DFT of M = A * BQ points per frame -- the adjoint of the complex transform inside
// warp_fft_fwd_pk: lane (fr, k1) passes V_fr[k1 + A k2] in slot k2 of z; store(integral_constant s, f2, integral_constant a, b, v)
// receives v = sum_k V_f2[k] exp(+2 pi j m k / M) for m = BQ a + b.  (Second half of every adjoint; with warp_fft_fwd_pk and a
// pointwise product it is also the circular convolution of the chirp-z transform in stft_ex.cu.)
template <class P, class Store>
__device__ __forceinline__ void warp_cfft_inv_pk(const cf (&z)[P::BQ], float2* xbuf, const float2* __restrict__ tw, int lane, Store store) {
    constexpr int A = P::A, BQ = P::BQ, RS = P::RS;
    const int k1 = lane % A;
    cf y[BQ];                                                // inverse BQ-point DFT over k2
    static_for<0, BQ>([&](auto kc) {
        constexpr int k2 = decltype(kc)::value;
        y[brev(k2, BQ)] = z[k2];
    });
    float2 wr_[8], wq_[BQ / 8];                             // conj(W_M^(b k1)) = conj(W^(r k1)) conj(W^(8 q k1)), b = 8 q + r
    static_for<1, 8>([&](auto rc) { constexpr int r = decltype(rc)::value; wr_[r] = ldg2(tw + k1 * BQ + r); });
    static_for<1, BQ / 8>([&](auto qc) { constexpr int q = decltype(qc)::value; wq_[q] = ldg2(tw + k1 * BQ + 8 * q); });
    Dit<BQ, true>::run(y);
    cf* wr = reinterpret_cast<cf*>(xbuf) + lane * RS;
    static_for<0, BQ>([&](auto bc_) {
        constexpr int b = decltype(bc_)::value;
        constexpr int r = b % 8, q = b / 8;
        if constexpr (r > 0) y[b] = fma2(rotp(y[b]), bc(wr_[r].y), mul2(y[b], bc(wr_[r].x)));   // conj(W) = w.x + j w.y
        if constexpr (q > 0) y[b] = fma2(rotp(y[b]), bc(wq_[q].y), mul2(y[b], bc(wq_[q].x)));
        if constexpr (P::kVec4) {
            if constexpr (b & 1) reinterpret_cast<ulonglong2*>(wr)[b >> 1] = make_ulonglong2(y[b - 1], y[b]);
        } else {
            wr[b] = y[b];
        }
    });
    __syncwarp();
    static_for<0, P::NS1>([&](auto sc) {
        constexpr int s = decltype(sc)::value;
        const int q = lane + 32 * s;
        const int f2 = q / BQ, b = q % BQ;
        cf a[A];
        const cf* col = reinterpret_cast<const cf*>(xbuf) + (f2 * A) * RS + b;
        static_for<0, A>([&](auto kc) {
            constexpr int kk = decltype(kc)::value;
            a[brev(kk, A)] = col[kk * RS];
        });
        Dit<A, true>::run(a);
        static_for<0, A>([&](auto ac) {
            constexpr int av = decltype(ac)::value;
            store(sc, f2, ac, b, a[av]);
        });
    });
    __syncwarp();
}

// The adjoint from its second stage on: zo / u / gh as left by the two functions above (zo[BQ/2 ..) is filled here).
template <class P, class Store>
__device__ __forceinline__ void warp_fft_adj_tail_pk(cf (&z)[P::BQ], const cf (&u)[P::BQ / 2], const cf gh, float2* xbuf,
                                                     const float2* __restrict__ tw, int lane, Store store) {
    constexpr int A = P::A, BQ = P::BQ, RS = P::RS;
    const int k1 = lane % A, fr = lane / A;
    const int plane = fr * A + ((A - k1) % A);
    static_for<0, BQ / 2>([&](auto jc) {
        constexpr int j = decltype(jc)::value;
        z[BQ - 1 - j] = cnj(__shfl_sync(0xffffffffu, u[j], plane));
    });
    if (k1 == 0) {                                           // these lanes pair inside themselves: slot j <-> (BQ - j) % BQ
        const cf t0 = add2(z[0], cnj(u[0]));                 // Z[0] feeds both X[0] and X[M]
        static_for<1, BQ / 2>([&](auto jc) {
            constexpr int j = decltype(jc)::value;
            z[BQ - j] = cnj(u[j]);
        });
        z[0] = t0;
        z[BQ / 2] = mul2(cnj(gh), bc(2.0f));
    }
    warp_cfft_inv_pk<P>(z, xbuf, tw, lane, store);
}

// ------------------------------------------------------------------ adjoint transform
// grad(fr, k, slot) returns, for the slot convention of warp_fft_post_pk, dL/dX[k] (even slots, slot BQ) or
// conj(dL/dX[M-k]) (odd slots).  store(integral_constant s, f2, integral_constant a, b, v) receives the gradient of the
// packed sample pair m = BQ * a + b of frame f2 *before* the window.
template <class P, class Grad, class Store>
__device__ __forceinline__ void warp_fft_adj_pk(float2* xbuf, const float2* __restrict__ tw, const float2* __restrict__ pt,
                                                int lane, Grad grad, Store store) {
    constexpr int A = P::A, BQ = P::BQ, M = P::M, RS = P::RS;
    const int k1 = lane % A, fr = lane / A;
    const int plane = fr * A + ((A - k1) % A);
    cf z[BQ];                                                // natural order: slot k2 holds dL/dZ[k1 + A k2]
    cf u[BQ / 2];                                            // S - T' of pair j; its conjugate is owed to the partner lane
    const cf base = ldg_cf(pt + k1);
    static_for<0, BQ / 2>([&](auto jc) {
        constexpr int j = decltype(jc)::value;
        const int k = k1 + A * j;
        const cf g = grad(fr, k, 2 * j);
        const cf gpc = grad(fr, M - k, 2 * j + 1);
        const cf S = add2(g, gpc), D = sub2(g, gpc);        // G[k] +- conj(G[M-k])
        const cf cs = post_twiddle_pk<BQ, j>(base);
        // T' = conj(-j W_n^k) D = -(cs.x rot(D) + cs.y D)
        const cf Tn = fma2(rotm(D), bc(lo(cs)), mul2(D, bc(hi(cs))));
        z[j] = sub2(S, Tn);                                  // dL/dZ[k]
        u[j] = add2(S, Tn);                                  // conj(dL/dZ[M-k])
    });
    static_for<0, BQ / 2>([&](auto jc) {
        constexpr int j = decltype(jc)::value;
        z[BQ - 1 - j] = cnj(__shfl_sync(0xffffffffu, u[j], plane));
    });
    if (k1 == 0) {                                           // these lanes pair inside themselves: slot j <-> (BQ - j) % BQ
        const cf gh = grad(fr, M / 2, BQ);
        const cf t0 = add2(z[0], cnj(u[0]));                 // Z[0] feeds both X[0] and X[M]
        static_for<1, BQ / 2>([&](auto jc) {
            constexpr int j = decltype(jc)::value;
            z[BQ - j] = cnj(u[j]);
        });
        z[0] = t0;
        z[BQ / 2] = mul2(cnj(gh), bc(2.0f));
    }
    cf y[BQ];                                                // inverse BQ-point DFT over k2
    static_for<0, BQ>([&](auto kc) {
        constexpr int k2 = decltype(kc)::value;
        y[brev(k2, BQ)] = z[k2];
    });
    float2 wr_[8], wq_[BQ / 8];                             // conj(W_M^(b k1)) = conj(W^(r k1)) conj(W^(8 q k1)), b = 8 q + r
    static_for<1, 8>([&](auto rc) { constexpr int r = decltype(rc)::value; wr_[r] = ldg2(tw + k1 * BQ + r); });
    static_for<1, BQ / 8>([&](auto qc) { constexpr int q = decltype(qc)::value; wq_[q] = ldg2(tw + k1 * BQ + 8 * q); });
    Dit<BQ, true>::run(y);
    cf* wr = reinterpret_cast<cf*>(xbuf) + lane * RS;
    static_for<0, BQ>([&](auto bc_) {
        constexpr int b = decltype(bc_)::value;
        constexpr int r = b % 8, q = b / 8;
        if constexpr (r > 0) y[b] = fma2(rotp(y[b]), bc(wr_[r].y), mul2(y[b], bc(wr_[r].x)));   // conj(W) = w.x + j w.y
        if constexpr (q > 0) y[b] = fma2(rotp(y[b]), bc(wq_[q].y), mul2(y[b], bc(wq_[q].x)));
        if constexpr (P::kVec4) {
            if constexpr (b & 1) reinterpret_cast<ulonglong2*>(wr)[b >> 1] = make_ulonglong2(y[b - 1], y[b]);
        } else {
            wr[b] = y[b];
        }
    });
    __syncwarp();
    static_for<0, P::NS1>([&](auto sc) {
        constexpr int s = decltype(sc)::value;
        const int q = lane + 32 * s;
        const int f2 = q / BQ, b = q % BQ;
        cf a[A];
        const cf* col = reinterpret_cast<const cf*>(xbuf) + (f2 * A) * RS + b;
        static_for<0, A>([&](auto kc) {
            constexpr int kk = decltype(kc)::value;
            a[brev(kk, A)] = col[kk * RS];
        });
        Dit<A, true>::run(a);
        static_for<0, A>([&](auto ac) {
            constexpr int av = decltype(ac)::value;
            store(sc, f2, ac, b, a[av]);
        });
    });
    __syncwarp();
}

}  // namespace tac

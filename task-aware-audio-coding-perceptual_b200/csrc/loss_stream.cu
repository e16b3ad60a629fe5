// K2, streaming form -- fused multi-resolution STFT loss value + gradient for the hop = n_fft / 4 framing of
// msstftd.py:174-175, with NO block barrier, NO shared accumulator tile and NO atomics.
//
// A warp walks a run of CONSECUTIVE frames of one row and one scale.  Per frame it transforms x and y in registers
// (fft_core.cuh), folds the two loss sums, forms dL/dY in the frequency domain and runs the adjoint transform.  The
// adjoint leaves sample pair m = BQ * a + b of the frame in lane b; a hop (n / 4 samples) is a shift of A / 4 in `a`
// with the SAME b, so the overlap-add of the four frames that cover a hop block is LANE-PRIVATE: every lane keeps its
// slice of the four pending hop blocks in a per-warp shared-memory ring that only it touches (no conflict, no
// synchronisation), and after each frame the oldest block is complete and leaves with one coalesced store.
//
// Rows are cut into runs of ~5 frames so that there are enough independent warp-runs to fill the GPU.  A run's
// first three hop blocks also need the last three frames of the previous run: the previous run flushes its pending
// partial sums for them into g, the run itself writes its own partial sums into h, and the combine kernel -- which
// exists anyway to add the scales in a fixed order -- adds h on exactly those blocks.  Every output sample is
// therefore written once per scale with plain stores, in a fixed order: the gradient is bit-stable run to run.
#include "tac_common.cuh"
#include "loss_common.cuh"
#include "fft_core_pk.cuh"

namespace tac {

#ifndef TAC_STREAM_WARPS
#define TAC_STREAM_WARPS 4
#endif
constexpr int kStreamWarps = TAC_STREAM_WARPS;
constexpr int kStreamThreads = kStreamWarps * 32;
#ifndef TAC_STREAM_PK
#define TAC_STREAM_PK 1       // packed-complex core (FFMA2); 0 = the scalar core of fft_core.cuh
#endif
#ifndef TAC_STREAM_JOINT
#define TAC_STREAM_JOINT 1    // x and y frames through the forward transform together (fft_core_pk.cuh: warp_fft_fwd2_pk): 1 = plans below n = 2048, 2 = all, 0 = none
#endif
#ifndef TAC_STREAM_PF
#define TAC_STREAM_PF 1      // d > 0: prefetch.global.L1 of the new quarter of the frame d steps ahead (A/B on B200, cfg 2: off 0.386, 1: 0.381 ms)
#endif
#ifndef TAC_RUN_TARGET
#define TAC_RUN_TARGET 5       // A/B on B200, cfg 2 (fused loss, ms): 3: 0.409, 4: 0.391, 5: 0.390, 6: 0.392, 7: 0.396, 8: 0.411, 10: 0.402, 14: 0.408
#endif
constexpr int kRunTarget = TAC_RUN_TARGET;          // frames per run (work unit of one lane group)
// per-warp overlap-add ring.  Scalar kernel: four hop blocks, NFR * M float2 <= 1024 (8 KiB).  Packed kernel: THREE --
// the block that completes in a frame (q == 0) and the one that opens in it (q == 3) belong to the same lane and share a
// slot -- so a CTA needs 58 KiB instead of 66, two CTAs fit the 132 KiB shared-memory configuration and the L1 that
// serves the 3/4-overlapping frame reads grows from 64 to 96 KiB.
constexpr int kOlaFloat2 = TAC_STREAM_PK ? 768 : 1024;
constexpr int kStreamWarpFloat2 = kXbufFloat2 + kOlaFloat2;

struct StreamScale {
    int n, hop, F, NR;           // NR runs per row
    int cta_begin, wpr;          // blockIdx range of this scale starts at cta_begin; wpr = warp-runs per row
    long long nwr;               // warp-runs of this scale = rows * wpr (a CTA takes kStreamWarps consecutive ones)
    int orig;                    // index of the scale in the caller's lists (sums / coef slots)
    int l2hop, small;            // hop == 1 << l2hop; small: F * NR < 2^31, the run arithmetic fits 32 bits
    FftTables t;
    float* g;                    // [rows][T] complete blocks + flushed partial sums
    float* h;                    // [rows][T] partial sums of the first three blocks of runs r >= 1
};

struct StreamArgs {
    const float* x;
    const float* y;
    long long T, rows;
    int nscales, total_ctas, cta_base;   // cta_base: global index of this launch's CTA 0 (the n = 512 scales run as a second launch)
    float eps;
    const float* coef;           // [2 * nscales] device
    double* partial;             // [total_ctas][kStreamWarps][2] or null
    float* grad;                 // [rows][T] (combine kernel)
    StreamScale sc[TAC_MAX_SCALES];
};

__host__ __device__ __forceinline__ int run_begin(int r, int F, int NR) { return (int)(((long long)r * F) / NR); }

// Frame reads: allocating read-only loads.  Consecutive frames of a run share 3/4 of their samples, and at 2 CTAs per SM
// the L1 keeps ~90 KiB, enough for the frames of the resident warps (measured: 0.68 ms against 0.72 ms with
// L1::no_allocate on cfg 2).
__device__ __forceinline__ float2 ld_stream2(const float* p) {
#ifdef TAC_STREAM_NOALLOC
    float2 v;
    asm("ld.global.nc.L1::no_allocate.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(p));
    return v;
#else
    return __ldg(reinterpret_cast<const float2*>(p));
#endif
}

// One CTA = kStreamWarps warps, each walking NFR consecutive runs of one row (one run per lane group).
template <class P, bool WITH_SUMS>
__device__ __forceinline__ void stream_cta(const StreamArgs& a, const StreamScale& sc, int cta, float2* smem) {
    constexpr int A = P::A, BQ = P::BQ, M = P::M, NFR = P::NFR, NS1 = P::NS1, QA = A / 4;
    static_assert(A % 4 == 0 && (TAC_STREAM_PK || NFR * M <= kOlaFloat2), "plan does not fit the streaming overlap-add ring");
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float2* xbuf = smem + warp * kStreamWarpFloat2;
    float2* ola = xbuf + kXbufFloat2;                      // [NFR][4 quarters][QA][BQ]
    const long long wr = (long long)cta * kStreamWarps + warp;    // warp-run index inside the scale
    if (wr >= sc.nwr) {                                            // spare warp of the scale's last CTA
        if (WITH_SUMS && lane < 2) a.partial[((long long)(blockIdx.x + a.cta_base) * kStreamWarps + warp) * 2 + lane] = 0.0;
        return;
    }
    const long long row = wr / sc.wpr;
    const int run0 = (int)(wr % sc.wpr) * NFR;
    const int F = sc.F, NR = sc.NR, hop = sc.hop;
    const float* xr = a.x + row * a.T;
    const float* yr = a.y + row * a.T;
    float* grow = sc.g + row * a.T;
    float* hrow = sc.h + row * a.T;
    const float2* __restrict__ win2 = sc.t.win2;

    auto run_lo = [&](int g) { const int r = run0 + g; return r < NR ? run_begin(r, F, NR) : 0; };
    auto run_len = [&](int g) { const int r = run0 + g; return r < NR ? run_begin(r + 1, F, NR) - run_begin(r, F, NR) : 0; };
    int maxlen = 0;
#pragma unroll
    for (int g = 0; g < NFR; ++g) maxlen = max(maxlen, run_len(g));
    // lane-group parameters in the two lane <-> frame mappings of fft_core.cuh
    const int len_post = run_len(lane / A);                // post-processing: frame group lane / A
    int lo_s[NS1], len_s[NS1], f2_s[NS1];                  // pass 1 / adjoint output: group (lane + 32 s) / BQ
#pragma unroll
    for (int s = 0; s < NS1; ++s) {
        f2_s[s] = (lane + 32 * s) / BQ;
        lo_s[s] = run_lo(f2_s[s]);
        len_s[s] = run_len(f2_s[s]);
    }
    const int b_lane = lane % BQ;

    {   // empty ring
        float4* o4 = reinterpret_cast<float4*>(ola);
        for (int i = lane; i < NFR * M / 2; i += 32) o4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __syncwarp();

    const float c_l1 = __ldg(a.coef + 2 * sc.orig), c_log = __ldg(a.coef + 2 * sc.orig + 1);
    const float eps2 = eps_squared(a.eps);
    float l1 = 0.0f, ll = 0.0f;
    for (int t = 0; t < maxlen; ++t) {
        float zr[BQ], zi[BQ];
        float gr[BQ + 1], gi[BQ + 1];
        const float* fx[NS1];
        const float* fy[NS1];
#pragma unroll
        for (int s = 0; s < NS1; ++s) {                    // finished (or empty) groups recompute a valid frame, masked below
            const int f = lo_s[s] + min(t, max(len_s[s] - 1, 0));
            fx[s] = xr + (long long)f * hop;
            fy[s] = yr + (long long)f * hop;
        }
        auto load_x = [&](auto scn, int, int m) -> float2 {
            const float2 v = ld_stream2(fx[decltype(scn)::value] + 2 * m);
            const float2 w = ldg2(win2 + m);
            return make_float2(__fmul_rn(v.x, w.x), __fmul_rn(v.y, w.y));
        };
        auto load_y = [&](auto scn, int, int m) -> float2 {
            const float2 v = ld_stream2(fy[decltype(scn)::value] + 2 * m);
            const float2 w = ldg2(win2 + m);
            return make_float2(__fmul_rn(v.x, w.x), __fmul_rn(v.y, w.y));
        };
        warp_fft_fwd<P>(zr, zi, xbuf, sc.t.tw, lane, load_x);
        warp_fft_post<P>(zr, zi, sc.t.pt, lane, [&](int, int, float vr, float vi, int slot) {
            gr[slot] = fmaf(vr, vr, vi * vi);                             // |X|^2
        });
        warp_fft_fwd<P>(zr, zi, xbuf, sc.t.tw, lane, load_y);
        float tl1 = 0.0f, tll = 0.0f;
        warp_fft_post<P>(zr, zi, sc.t.pt, lane, [&](int, int, float vr, float vi, int slot) {
            const float g = loss_bin_grad<WITH_SUMS>(gr[slot], fmaf(vr, vr, vi * vi), eps2, c_l1, c_log, tl1, tll);
            gr[slot] = g * vr; gi[slot] = g * vi;
        });
        if (WITH_SUMS && t < len_post) { l1 += tl1; ll += tll; }
        warp_fft_adj<P>(xbuf, sc.t.tw, sc.t.pt, lane,
            [&](int, int, int slot) -> float2 { return make_float2(gr[slot], gi[slot]); },
            [&](auto scn, int f2, auto ac, int b, float2 gv) {
                constexpr int s = decltype(scn)::value, av = decltype(ac)::value;
                constexpr int q = av / QA, ap = av % QA;                  // hop block of the frame, pair index inside it
                const bool active = t < len_s[s];
                const int f = lo_s[s] + t;
                const float2 w = ldg2(win2 + BQ * av + b);
                const int p = (q + f) & 3;                                // ring quarter that holds hop block f + q
                float2* slot = ola + ((f2 * 4 + p) * QA + ap) * BQ + b;
                if (q == 3) {                                             // newest block: nothing pending yet
                    if (active) *slot = make_float2(gv.x * w.x, gv.y * w.y);
                } else if (q == 0) {                                      // oldest block: complete after this frame
                    const float2 o = *slot;
                    float* dst = (t < 3 && run0 + f2 > 0) ? hrow : grow;  // first three blocks of a run r >= 1 are partial
                    if (active)
                        *reinterpret_cast<float2*>(dst + (long long)f * hop + 2 * (BQ * ap + b)) =
                            make_float2(fmaf(gv.x, w.x, o.x), fmaf(gv.y, w.y, o.y));
                } else {
                    float2 o = *slot;
                    o.x = fmaf(gv.x, w.x, o.x); o.y = fmaf(gv.y, w.y, o.y);
                    if (active) *slot = o;
                }
            });
    }
    // flush the three pending hop blocks behind the last frame of every run (partial sums unless the row ends here)
#pragma unroll
    for (int s = 0; s < NS1; ++s) {
        if (len_s[s] > 0) {
            const int f = lo_s[s] + len_s[s] - 1;
#pragma unroll
            for (int q = 1; q < 4; ++q) {
                const int p = (q + f) & 3;
#pragma unroll
                for (int ap = 0; ap < QA; ++ap) {
                    const float2 o = ola[((f2_s[s] * 4 + p) * QA + ap) * BQ + b_lane];
                    *reinterpret_cast<float2*>(grow + (long long)(f + q) * hop + 2 * (BQ * ap + b_lane)) = o;
                }
            }
        }
    }
    if (WITH_SUMS) {                                       // fixed lane -> warp order, one partial per warp: deterministic sums
        l1 = warp_sum(l1); ll = warp_sum(ll);
        if (lane == 0) {
            double* pp = a.partial + ((long long)(blockIdx.x + a.cta_base) * kStreamWarps + warp) * 2;
            pp[0] = (double)l1; pp[1] = (double)ll * (double)kLogHalf;
        }
    }
}


// Cache policy of the packed kernel.  A sample is read by the four consecutive frames that cover it: the first quarter
// of a frame is the LAST use of its data (evict first), everything else is wanted again one step later.  The gradient
// blocks are written once and read by the combine kernel only: streaming stores keep them out of the way of the frames.
#ifndef TAC_LD_HINT
#define TAC_LD_HINT 2         // 0: plain read-only loads; 1: frame hints; 2: + tables evict-last (A/B on B200: 0.570 -> 0.559 ms with 2 + streaming stores)
#endif
#ifndef TAC_ST_CS
#define TAC_ST_CS 1
#endif
template <int QUARTER>
__device__ __forceinline__ cf ld_frame(const float* p) {
#if TAC_LD_HINT >= 1
    cf v;
    if constexpr (QUARTER == 0) asm("ld.global.nc.L1::evict_first.b64 %0, [%1];" : "=l"(v) : "l"(p));
    else asm("ld.global.nc.L1::evict_last.b64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
#else
    return __ldg(reinterpret_cast<const cf*>(p));
#endif
}
__device__ __forceinline__ cf ld_table(const float2* p) {
#if TAC_LD_HINT >= 2
    cf v;
    asm("ld.global.nc.L1::evict_last.b64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
#else
    return ldg_cf(p);
#endif
}
__device__ __forceinline__ void st_grad(float* p, cf v) {
#if TAC_ST_CS
    asm volatile("st.global.cs.b64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
#else
    *reinterpret_cast<cf*>(p) = v;
#endif
}

// The same walk on the packed-complex core (fft_core_pk.cuh): every complex value is one aligned register pair and the
// butterflies, twiddles, window products and overlap-adds are two-wide FP32 instructions -- about half the instruction
// stream of stream_cta for the same arithmetic.  Only the per-bin loss terms (|.|, log, sign) stay scalar.
// COMPACT: the per-warp shared memory is sized for THIS plan (exchange tile 32 x RS, ring NFR * M * 3 / 4) instead of the largest
// one -- the n = 512 launch needs 30 KB per CTA instead of 58, so that three CTAs per SM leave the L1 156 KB instead of ~60 for
// the 3/4-overlapping frame reads (its hit rate was 39 % and the frame-load phase held 23 % of that kernel's stall samples).
template <class P> constexpr int kCompactXbuf = 32 * P::RS;
template <class P> constexpr int kCompactWarpFloat2 = kCompactXbuf<P> + P::NFR * P::M * 3 / 4;
template <class P, bool WITH_SUMS, bool COMPACT = false>
__device__ __forceinline__ void stream_cta_pk(const StreamArgs& a, const StreamScale& sc, int cta, float2* smem) {
    constexpr int A = P::A, BQ = P::BQ, M = P::M, NFR = P::NFR, NS1 = P::NS1, QA = A / 4;
    constexpr int kWarpF2 = COMPACT ? kCompactWarpFloat2<P> : kStreamWarpFloat2, kXb = COMPACT ? kCompactXbuf<P> : kXbufFloat2;
    static_assert((kWarpF2 % 2) == 0 && (kXb % 2) == 0, "16-byte alignment of the per-warp tiles");
    constexpr bool kJoint = TAC_STREAM_JOINT != 0 && (TAC_STREAM_JOINT == 2 || M < 1024);
    static_assert(A % 4 == 0 && NFR * M * 3 / 4 <= kOlaFloat2, "plan does not fit the streaming overlap-add ring");
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float2* xbuf = smem + warp * kWarpF2;
    cf* ola = reinterpret_cast<cf*>(xbuf + kXb);          // [NFR][3 slots][QA][BQ]: hop block j of a frame group lives in slot j % 3
    const long long wr = (long long)cta * kStreamWarps + warp;
    if (wr >= sc.nwr) {
        if (WITH_SUMS && lane < 2) a.partial[((long long)(blockIdx.x + a.cta_base) * kStreamWarps + warp) * 2 + lane] = 0.0;
        return;
    }
    const long long row = wr / sc.wpr;
    const int run0 = (int)(wr % sc.wpr) * NFR;
    const int F = sc.F, NR = sc.NR, hop = sc.hop;
    const float* xr = a.x + row * a.T;
    const float* yr = a.y + row * a.T;
    float* grow = sc.g + row * a.T;
    float* hrow = sc.h + row * a.T;
    const float2* __restrict__ win2 = sc.t.win2;

    auto run_lo = [&](int g) { const int r = run0 + g; return r < NR ? run_begin(r, F, NR) : 0; };
    auto run_len = [&](int g) { const int r = run0 + g; return r < NR ? run_begin(r + 1, F, NR) - run_begin(r, F, NR) : 0; };
    int maxlen = 0;
#pragma unroll
    for (int g = 0; g < NFR; ++g) maxlen = max(maxlen, run_len(g));
    const int len_post = run_len(lane / A);
    int lo_s[NS1], len_s[NS1], f2_s[NS1], fm_s[NS1];       // fm_s: (current frame) % 3
#pragma unroll
    for (int s = 0; s < NS1; ++s) {
        f2_s[s] = (lane + 32 * s) / BQ;
        lo_s[s] = run_lo(f2_s[s]);
        len_s[s] = run_len(f2_s[s]);
        fm_s[s] = lo_s[s] % 3;
    }
    const int b_lane = lane % BQ;
    {
        float4* o4 = reinterpret_cast<float4*>(ola);
        for (int i = lane; i < NFR * M * 3 / 8; i += 32) o4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __syncwarp();

    auto window = [&](auto, auto, int m) -> cf { return ld_table(win2 + m); };
    const float c_l1 = __ldg(a.coef + 2 * sc.orig), c_log = __ldg(a.coef + 2 * sc.orig + 1);
    const float eps2 = eps_squared(a.eps);
    float l1 = 0.0f, ll = 0.0f;
    for (int t = 0; t < maxlen; ++t) {
        cf zo[BQ], uo[BQ / 2], gh;                           // adjoint inputs (fft_core_pk.cuh: warp_fft_post_adj_head_pk)
        const float* fx[NS1];
        const float* fy[NS1];
#pragma unroll
        for (int s = 0; s < NS1; ++s) {
            const int f = lo_s[s] + min(t, max(len_s[s] - 1, 0));
            fx[s] = xr + (long long)f * hop;
            fy[s] = yr + (long long)f * hop;
#if TAC_STREAM_PF
            // the one quarter of the NEXT frame that is not in the L1 yet (hop samples behind this frame's end): lane b fetches
            // line b of it now, a whole frame step ahead of its use
            const int bq = (lane + 32 * s) % BQ;
            if (t + TAC_STREAM_PF < len_s[s] && bq * 32 < hop) {
                asm volatile("prefetch.global.L1 [%0];" :: "l"(fx[s] + (3 + TAC_STREAM_PF) * hop + 32 * bq));
                asm volatile("prefetch.global.L1 [%0];" :: "l"(fy[s] + (3 + TAC_STREAM_PF) * hop + 32 * bq));
            }
#endif
        }
        float tl1 = 0.0f, tll = 0.0f;
        if constexpr (kJoint) {
        // x and y frames through the transform TOGETHER: one window / twiddle load for both, two independent butterfly streams
        // (A/B on B200, cfg 2: n = 512 scale 0.172 -> 0.157 ms, n = 1024 0.182 -> 0.180 ms; n = 2048 0.159 -> 0.160 ms -- its 64
        // complex values per lane leave no room for a second transform -- so the largest plan keeps the sequential form)
        cf zx[BQ], zy[BQ];
        auto load_xy = [&](auto scn, int, int m, auto ac, cf& vx, cf& vy) {
            const cf w = window(scn, ac, m);
            vx = mul2(ld_frame<(4 * decltype(ac)::value) / A>(fx[decltype(scn)::value] + 2 * m), w);
            vy = mul2(ld_frame<(4 * decltype(ac)::value) / A>(fy[decltype(scn)::value] + 2 * m), w);
        };
        warp_fft_fwd2_pk<P>(zx, zy, xbuf, sc.t.tw, lane, load_xy);
        warp_fft_post2_adj_head_pk<P>(zx, zy, sc.t.pt, lane, [&](int, int, cf vx, cf v, int) -> cf {
            const float xr_ = lo(vx), xi_ = hi(vx), vr = lo(v), vi = hi(v);
            const float g = loss_bin_grad<WITH_SUMS>(fmaf(xr_, xr_, xi_ * xi_), fmaf(vr, vr, vi * vi), eps2, c_l1, c_log, tl1, tll);
            return mul2(v, bc(g));                                        // dL/dY[k]  (odd slots: its conjugate)
        }, zo, uo, gh);
        } else {
        cf z[BQ];
        float mx[BQ + 1];
        auto load_x = [&](auto scn, int, int m, auto ac) -> cf {
            return mul2(ld_frame<(4 * decltype(ac)::value) / A>(fx[decltype(scn)::value] + 2 * m), window(scn, ac, m));
        };
        auto load_y = [&](auto scn, int, int m, auto ac) -> cf {
            return mul2(ld_frame<(4 * decltype(ac)::value) / A>(fy[decltype(scn)::value] + 2 * m), window(scn, ac, m));
        };
        warp_fft_fwd_pk<P>(z, xbuf, sc.t.tw, lane, load_x);
        warp_fft_post_pk<P>(z, sc.t.pt, lane, [&](int, int, cf v, int slot) {
            const float vr = lo(v), vi = hi(v);
            mx[slot] = fmaf(vr, vr, vi * vi);                             // |X|^2
        });
        warp_fft_fwd_pk<P>(z, xbuf, sc.t.tw, lane, load_y);
        warp_fft_post_adj_head_pk<P>(z, sc.t.pt, lane, [&](int, int, cf v, int slot) -> cf {
            const float vr = lo(v), vi = hi(v);
            const float g = loss_bin_grad<WITH_SUMS>(mx[slot], fmaf(vr, vr, vi * vi), eps2, c_l1, c_log, tl1, tll);
            return mul2(v, bc(g));                                        // dL/dY[k]  (odd slots: its conjugate)
        }, zo, uo, gh);
        }
        if (WITH_SUMS && t < len_post) { l1 += tl1; ll += tll; }
        // post-processing, per-bin loss terms and the adjoint's first stage ran as one loop above (the gradient never exists as an
        // array); the adjoint from its second stage on:
        warp_fft_adj_tail_pk<P>(zo, uo, gh, xbuf, sc.t.tw, lane,
            [&](auto scn, int f2, auto ac, int b, cf gv) {
                constexpr int s = decltype(scn)::value, av = decltype(ac)::value;
                constexpr int q = av / QA, ap = av % QA;
                const bool active = t < len_s[s];
                const int f = lo_s[s] + t;
                const cf w = window(scn, ac, BQ * av + b);
                int p = fm_s[s] + (q == 3 ? 0 : q);                  // slot of hop block f + q; q == 3 reuses the slot q == 0 just left
                p = p >= 3 ? p - 3 : p;
                cf* slot = ola + ((f2 * 3 + p) * QA + ap) * BQ + b;
                if (q == 3) {
                    if (active) *slot = mul2(gv, w);
                } else if (q == 0) {
                    const cf o = *slot;
                    float* dst = (t < 3 && run0 + f2 > 0) ? hrow : grow;
                    if (active) st_grad(dst + (long long)f * hop + 2 * (BQ * ap + b), fma2(gv, w, o));
                } else {
                    const cf o = fma2(gv, w, *slot);
                    if (active) *slot = o;
                }
            });
#pragma unroll
        for (int s = 0; s < NS1; ++s) fm_s[s] = fm_s[s] == 2 ? 0 : fm_s[s] + 1;
    }
#pragma unroll
    for (int s = 0; s < NS1; ++s) {
        if (len_s[s] > 0) {
            const int f = lo_s[s] + len_s[s] - 1;
#pragma unroll
            for (int q = 1; q < 4; ++q) {
                const int p = (q + f) % 3;
#pragma unroll
                for (int ap = 0; ap < QA; ++ap) {
                    const cf o = ola[((f2_s[s] * 3 + p) * QA + ap) * BQ + b_lane];
                    st_grad(grow + (long long)(f + q) * hop + 2 * (BQ * ap + b_lane), o);
                }
            }
        }
    }
    if (WITH_SUMS) {
        l1 = warp_sum(l1); ll = warp_sum(ll);
        if (lane == 0) {
            double* pp = a.partial + ((long long)(blockIdx.x + a.cta_base) * kStreamWarps + warp) * 2;
            pp[0] = (double)l1; pp[1] = (double)ll * (double)kLogHalf;
        }
    }
}

#ifndef TAC_STREAM_OCC
#define TAC_STREAM_OCC 2      // 255 registers; 3 (168 registers) spills and measured no faster
#endif

#ifndef TAC_STREAM_SPLIT
#define TAC_STREAM_SPLIT 1    // the n = 512 scales as a second, programmatically dependent launch at 3 CTAs per SM (168 registers, no spills)
#endif

template <bool WITH_SUMS>
__global__ void __launch_bounds__(kStreamThreads, TAC_STREAM_OCC) mrstft_stream_kernel(const StreamArgs a) {
    extern __shared__ __align__(16) unsigned char stream_smem_raw[];
    float2* smem = reinterpret_cast<float2*>(stream_smem_raw);
#if TAC_STREAM_SPLIT
    // the second launch (n = 512 scales) may start as soon as every CTA of this grid has been scheduled: its CTAs fill the
    // SMs that this grid's last, partial wave leaves idle
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#endif
    const int gcta = (int)blockIdx.x + a.cta_base;
    int s = 0;
    while (s + 1 < a.nscales && gcta >= a.sc[s + 1].cta_begin) ++s;
    const StreamScale& sc = a.sc[s];
    const int cta = gcta - sc.cta_begin;
#if TAC_STREAM_PK
    switch (sc.n) {                                                                // stream_applicable: n >= 512
        case 2048: stream_cta_pk<Plan2048, WITH_SUMS>(a, sc, cta, smem); break;
        case 1024: stream_cta_pk<Plan1024, WITH_SUMS>(a, sc, cta, smem); break;
        default: stream_cta_pk<Plan512, WITH_SUMS>(a, sc, cta, smem); break;
    }
#else
    switch (sc.n) {
        case 2048: stream_cta<Plan2048, WITH_SUMS>(a, sc, cta, smem); break;
        case 1024: stream_cta<Plan1024, WITH_SUMS>(a, sc, cta, smem); break;
        default: stream_cta<Plan512, WITH_SUMS>(a, sc, cta, smem); break;
    }
#endif
}

// The n = 512 plan keeps 16 complex values per lane: it compiles to 168 registers without spills, so three CTAs fit an SM
// (12 warps instead of 8 for a kernel that is bound by dependent-instruction latency).  Launched behind the kernel above
// with programmatic stream serialisation; before a CTA exits it waits for the primary grid (griddepcontrol.wait), so that
// the completion of THIS grid implies the completion of both -- the combine kernel behind it reads the planes of all scales.
template <bool WITH_SUMS>
__global__ void __launch_bounds__(kStreamThreads, 3) mrstft_stream_small_kernel(const StreamArgs a) {
    extern __shared__ __align__(16) unsigned char stream_smem_raw[];
    float2* smem = reinterpret_cast<float2*>(stream_smem_raw);
    const int gcta = (int)blockIdx.x + a.cta_base;
    int s = 0;
    while (s + 1 < a.nscales && gcta >= a.sc[s + 1].cta_begin) ++s;
    const StreamScale& sc = a.sc[s];
    stream_cta_pk<Plan512, WITH_SUMS, true>(a, sc, gcta - sc.cta_begin, smem);
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

// grad[row][t] = sum over scales (fixed order) of g_s[t] (+ h_s[t] on the first three hop blocks of runs r >= 1);
// samples behind the last frame of a scale get nothing from it.
// A warp step covers 32 * VEC consecutive samples of one row, starting on a multiple of 32 * VEC <= 128 <= every hop, so the
// hop block -- and with it "is this block a partial one" -- is the same for the whole warp: lane s works it out for scale s
// (one 32-bit division) and two ballots hand the answer to everybody.  (Round 2: the per-thread version -- a 64-bit row
// division and a run search per scale and thread -- issued 380 instructions per float4 and ran at 67 % issue utilisation,
// 37 us for 131 MB; this one is bound by its loads.)
constexpr int kCombineIter = 4;                 // warp steps of a CTA along its row
template <int VEC>
__global__ void __launch_bounds__(256) mrstft_combine_kernel(const StreamArgs a, const unsigned ctas_per_row) {
    // the sums kernel behind this one does not read what this one writes: it may start at once (it waits for this grid before it exits)
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned row = blockIdx.x / ctas_per_row;
    const unsigned span0 = (blockIdx.x - row * ctas_per_row) * (256u * VEC * kCombineIter);
    const long long rbase = (long long)row * a.T;
    const unsigned T = (unsigned)a.T;           // stream_applicable: T < 2^30
#pragma unroll
    for (int it = 0; it < kCombineIter; ++it) {
        const unsigned t0 = span0 + (unsigned)it * (256u * VEC) + warp * (32u * VEC);
        if (t0 >= T) break;                                               // warp-uniform
        bool valid = false, partial = false;
        if ((int)lane < a.nscales) {
            const StreamScale& sc = a.sc[lane];
            valid = t0 < (unsigned)(sc.F - 1) * (unsigned)sc.hop + (unsigned)sc.n;     // else behind the last frame
            const unsigned j = t0 >> sc.l2hop;
            if (valid && j < (unsigned)sc.F) {
                // run that owns hop block j: r = max { r : floor(r F / NR) <= j }
                if (sc.small) {
                    const unsigned F = (unsigned)sc.F, NR = (unsigned)sc.NR;
                    unsigned r = (j * NR) / F;
                    while (r + 1 < NR && ((r + 1) * F) / NR <= j) ++r;
                    while (r > 0 && (r * F) / NR > j) --r;
                    partial = r > 0 && j - (r * F) / NR < 3u;
                } else {
                    int r = (int)(((long long)j * sc.NR) / sc.F);
                    while (r + 1 < sc.NR && run_begin(r + 1, sc.F, sc.NR) <= (int)j) ++r;
                    while (r > 0 && run_begin(r, sc.F, sc.NR) > (int)j) --r;
                    partial = r > 0 && (int)j - run_begin(r, sc.F, sc.NR) < 3;
                }
            }
        }
        const unsigned mvalid = __ballot_sync(0xffffffffu, valid), mpart = __ballot_sync(0xffffffffu, partial);
        const unsigned t = t0 + lane * VEC;
        if (t >= T) continue;                                             // T is a multiple of VEC
        const long long off = rbase + t;
        float acc[VEC];
#pragma unroll
        for (int v = 0; v < VEC; ++v) acc[v] = 0.0f;
        for (int s = 0; s < a.nscales; ++s) {
            if (!((mvalid >> s) & 1u)) continue;
            const StreamScale& sc = a.sc[s];
            float v[VEC];
            if constexpr (VEC == 4) { const float4 q = __ldcs(reinterpret_cast<const float4*>(sc.g + off)); v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w; }
            else { const float2 q = __ldcs(reinterpret_cast<const float2*>(sc.g + off)); v[0] = q.x; v[1] = q.y; }
            if ((mpart >> s) & 1u) {
                if constexpr (VEC == 4) { const float4 q = __ldcs(reinterpret_cast<const float4*>(sc.h + off)); v[0] += q.x; v[1] += q.y; v[2] += q.z; v[3] += q.w; }
                else { const float2 q = __ldcs(reinterpret_cast<const float2*>(sc.h + off)); v[0] += q.x; v[1] += q.y; }
            }
#pragma unroll
            for (int e = 0; e < VEC; ++e) acc[e] += v[e];
        }
        float* dst = a.grad + off;
        if constexpr (VEC == 4) *reinterpret_cast<float4*>(dst) = make_float4(acc[0], acc[1], acc[2], acc[3]);
        else *reinterpret_cast<float2*>(dst) = make_float2(acc[0], acc[1]);
    }
}

// sums[2 * orig + i] = sum over the CTAs of that scale (fixed order) of partial[cta][i].  grid = 2 * nscales.
__global__ void __launch_bounds__(256) mrstft_stream_sums_kernel(const StreamArgs a, double* __restrict__ sums) {
    __shared__ double sh[256];
    const int s = blockIdx.x >> 1, i = blockIdx.x & 1;
    const StreamScale& sc = a.sc[s];
    const long long c0 = (long long)sc.cta_begin * kStreamWarps;
    const long long n = ((sc.nwr + kStreamWarps - 1) / kStreamWarps) * kStreamWarps;   // warp slots of this scale (spares hold 0)
    double v = 0.0;
    for (long long c = threadIdx.x; c < n; c += 256) v += a.partial[(c0 + c) * 2 + i];
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) sums[2 * sc.orig + i] = sh[0];
    // launched with programmatic stream serialisation behind the combine kernel (it only needs the stream kernels' partials, which
    // are complete before the combine kernel starts): completion of THIS grid must still imply completion of the combine kernel
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

static inline size_t align16(size_t v) { return (v + 15) & ~(size_t)15; }

bool stream_applicable(int64_t T, int n_scales, const int* n_fft, const int* hop, const int* win) {
    if ((T & 1) != 0 || T >= (1ll << 30)) return false;
    for (int s = 0; s < n_scales; ++s) {
        if (!supported_nfft(n_fft[s]) || hop[s] * 4 != n_fft[s]) return false;
        if (n_fft[s] < 512) return false;      // small plans (4-8 frames per warp task): the chunk/round kernel is faster
        if (win && win[s] != n_fft[s] && win[s] <= 0) return false;
        if (num_frames(T, n_fft[s], hop[s], 0) < 1) return false;
    }
    return true;
}

// Workspace: per scale g and h ([rows][T] float each), then [total_ctas][2] float64 partials.
static void stream_layout(int64_t rows, int64_t T, int n_scales, const int* n_fft, const int* hop, int* order, int* NR, int* cpr,
                          long long* total_ctas) {
    // heavy scales first: their CTAs are the longest, the small ones fill the tail
    for (int s = 0; s < n_scales; ++s) order[s] = s;
    for (int i = 0; i < n_scales; ++i)
        for (int j = i + 1; j < n_scales; ++j)
            if (n_fft[order[j]] > n_fft[order[i]]) { int t = order[i]; order[i] = order[j]; order[j] = t; }
    long long tot = 0;
    for (int s = 0; s < n_scales; ++s) {
        const int F = (int)num_frames(T, n_fft[s], hop[s], 0);
        int nr = (F + kRunTarget / 2) / kRunTarget;
        if (nr < 1) nr = 1;
        const int nfr = n_fft[s] >= 2048 ? 1 : (n_fft[s] >= 512 ? 2 : (n_fft[s] >= 128 ? 4 : 8));
        NR[s] = nr;
        cpr[s] = (nr + nfr - 1) / nfr;                                  // warp-runs per row
        tot += (rows * cpr[s] + kStreamWarps - 1) / kStreamWarps;
    }
    *total_ctas = tot;
}

size_t stream_workspace_bytes(int64_t rows, int64_t T, int n_scales, const int* n_fft, const int* hop) {
    int order[TAC_MAX_SCALES], NR[TAC_MAX_SCALES], cpr[TAC_MAX_SCALES];
    long long tot;
    stream_layout(rows, T, n_scales, n_fft, hop, order, NR, cpr, &tot);
    return (size_t)n_scales * 2 * align16((size_t)rows * T * sizeof(float)) + (size_t)tot * kStreamWarps * 2 * sizeof(double);
}

int launch_stream(const float* x, const float* y, int64_t rows, int64_t T, int n_scales, const int* n_fft, const int* hop,
                  const int* win, float eps, const float* coef, double* sums, float* grad_y, void* workspace,
                  size_t workspace_bytes, cudaStream_t st) {
    const size_t need = stream_workspace_bytes(rows, T, n_scales, n_fft, hop);
    if (!workspace || workspace_bytes < need)
        return set_error(TAC_EWORKSPACE, "workspace needs %zu bytes, got %zu", need, workspace_bytes);
    TAC_ALIGNED16(workspace, "workspace");
    TAC_REQUIRE((((uintptr_t)x | (uintptr_t)y | (uintptr_t)grad_y) & 7u) == 0, "x, y, grad_y must be 8-byte aligned");
    int order[TAC_MAX_SCALES], NR[TAC_MAX_SCALES], cpr[TAC_MAX_SCALES];
    long long tot;
    stream_layout(rows, T, n_scales, n_fft, hop, order, NR, cpr, &tot);
    TAC_REQUIRE(tot < (1ll << 31), "grid too large");
    StreamArgs a;
    a.x = x; a.y = y; a.T = T; a.rows = rows; a.nscales = n_scales; a.total_ctas = (int)tot; a.eps = eps; a.coef = coef;
    a.grad = grad_y;
    unsigned char* ws = reinterpret_cast<unsigned char*>(workspace);
    const size_t buf = align16((size_t)rows * T * sizeof(float));
    a.partial = sums ? reinterpret_cast<double*>(ws + (size_t)n_scales * 2 * buf) : nullptr;
    long long begin = 0;
    for (int i = 0; i < n_scales; ++i) {
        const int s = order[i];
        StreamScale& sc = a.sc[i];
        sc.n = n_fft[s]; sc.hop = hop[s]; sc.F = (int)num_frames(T, sc.n, sc.hop, 0); sc.NR = NR[s];
        sc.cta_begin = (int)begin; sc.wpr = cpr[s]; sc.nwr = rows * cpr[s]; sc.orig = s;
        sc.l2hop = 0;
        while ((1 << sc.l2hop) < sc.hop) ++sc.l2hop;                      // hop = n_fft / 4, a power of two (stream_applicable)
        sc.small = ((long long)sc.F * sc.NR < (1ll << 31)) ? 1 : 0;
        sc.g = reinterpret_cast<float*>(ws + (size_t)(2 * i) * buf);
        sc.h = reinterpret_cast<float*>(ws + (size_t)(2 * i + 1) * buf);
        int rc = get_tables(sc.n, win ? win[s] : sc.n, 1, &sc.t);
        if (rc) return rc;
        begin += (rows * cpr[s] + kStreamWarps - 1) / kStreamWarps;
    }
    const size_t smem = (size_t)kStreamWarps * kStreamWarpFloat2 * sizeof(float2);
    void (*kern)(const StreamArgs) = sums ? mrstft_stream_kernel<true> : mrstft_stream_kernel<false>;
    if (sums) TAC_SMEM_OPT_IN(mrstft_stream_kernel<true>, smem);
    else TAC_SMEM_OPT_IN(mrstft_stream_kernel<false>, smem);
    // scales are ordered heavy first: [0, n_big) CTAs belong to n >= 1024, the rest to n = 512
    long long n_big = tot;
    for (int i = 0; i < n_scales; ++i)
        if (a.sc[i].n == 512) { n_big = a.sc[i].cta_begin; break; }
    a.cta_base = 0;
    if (TAC_STREAM_SPLIT && TAC_STREAM_PK && n_big > 0 && n_big < tot) {
        void (*small)(const StreamArgs) = sums ? mrstft_stream_small_kernel<true> : mrstft_stream_small_kernel<false>;
        const size_t smem_small = (size_t)kStreamWarps * kCompactWarpFloat2<Plan512> * sizeof(float2);
        if (sums) TAC_SMEM_OPT_IN(mrstft_stream_small_kernel<true>, smem_small);
        else TAC_SMEM_OPT_IN(mrstft_stream_small_kernel<false>, smem_small);
        // three CTAs need 3 x (30 + 1) KB: ask for the 100 KB shared-memory configuration, the rest of the 256 KB is L1
        if (sums) TAC_CARVEOUT_ONCE(mrstft_stream_small_kernel<true>, 44);
        else TAC_CARVEOUT_ONCE(mrstft_stream_small_kernel<false>, 44);
        kern<<<(unsigned)n_big, kStreamThreads, smem, st>>>(a);
        TAC_LAUNCHED(1);
        TAC_CHECK_CUDA(cudaGetLastError());
        StreamArgs b = a;
        b.cta_base = (int)n_big;
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.gridDim = dim3((unsigned)(tot - n_big)); cfg.blockDim = dim3(kStreamThreads); cfg.dynamicSmemBytes = smem_small; cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        TAC_CHECK_CUDA(cudaLaunchKernelEx(&cfg, small, b));
        TAC_LAUNCHED(1);
    } else {
        kern<<<(unsigned)tot, kStreamThreads, smem, st>>>(a);
        TAC_LAUNCHED(1);
        TAC_CHECK_CUDA(cudaGetLastError());
    }
    const bool v4 = (T % 4 == 0) && ((reinterpret_cast<uintptr_t>(grad_y) & 15u) == 0);
    const long long span = 256ll * (v4 ? 4 : 2) * kCombineIter;                 // samples of a row per CTA
    const long long ctas_row = (T + span - 1) / span;
    TAC_REQUIRE(rows * ctas_row < (1ll << 31), "grid too large");
    if (v4) mrstft_combine_kernel<4><<<(unsigned)(rows * ctas_row), 256, 0, st>>>(a, (unsigned)ctas_row);
    else mrstft_combine_kernel<2><<<(unsigned)(rows * ctas_row), 256, 0, st>>>(a, (unsigned)ctas_row);
    TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    if (sums) {
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.gridDim = dim3((unsigned)(2 * n_scales)); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = 0; cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        TAC_CHECK_CUDA(cudaLaunchKernelEx(&cfg, mrstft_stream_sums_kernel, a, sums));
        TAC_LAUNCHED(1);
    }
    return TAC_OK;
}

}  // namespace tac

// K1 -- batched multi-scale STFT (framing + Hann window + real FFT + output epilogue, fused).
//
//  * stft_planes_kernel: ALL scales of a signal chunk in one CTA.  The chunk (+ halo) is staged in
//    shared memory once with 16-byte loads; every warp pulls "tasks" (NFR frames of one scale) from a
//    shared counter, transforms them in registers (fft_core.cuh) and writes the real / imaginary
//    planes in the frequency-innermost layout [B, 2C, F, K] of msstftd.py:152-153 -- one fully
//    coalesced 128-byte store per warp instruction.  HBM traffic = read x once + write the planes.
//  * stft_tiled_kernel: one scale, a tile of TF consecutive frames per CTA, for the time-innermost
//    layouts (complex [B,C,K,F] of torchaudio; magnitude/phase [B,2,K,f_pad] of the notebook front
//    end).  Results are transposed through a padded shared-memory tile so that each (k) row is
//    written as a contiguous run of TF frames.
#include "tac_common.cuh"
#include "fft_core_pk.cuh"

namespace tac {

constexpr int kWarps = 8;
constexpr int kThreads = kWarps * 32;

struct ScaleArg {
    int n, hop, pad, F, K;
    FftTables t;
    float* out;
};

struct PlanesArgs {
    const float* x;
    long long T;
    int C, nscales, chunk, nchunks, pmax, tile_len, pad_mode, vec_ok;
    ScaleArg sc[TAC_MAX_SCALES];
};

// Original-signal value at (possibly out-of-range) index g, honouring the centre padding mode.
__device__ __forceinline__ float padded_sample(const float* __restrict__ xr, long long g, long long T, int pad_mode) {
    if (g >= 0 && g < T) return __ldg(xr + g);
    if (pad_mode == TAC_PAD_REFLECT) {
        long long r = g < 0 ? -g : 2 * (T - 1) - g;
        if (r >= 0 && r < T) return __ldg(xr + r);
    }
    return 0.0f;
}

// tile[j] = x_padded[g0 + j], j in [0, len).  The in-range, 16-byte aligned middle of the tile is copied with float4 loads even
// when the tile hangs over either end of the row (centre padding in front, the last tile / chunk behind): only the <= 3 + 3
// samples around it and the padded samples take the scalar path.  (Round 2: tiles that touched a row end used to be staged
// sample by sample -- half of the front end's CTAs on 1 s clips -- and that loop held 43 % of the kernel's stall samples.)
__device__ __forceinline__ void stage_tile(float* tile, const float* __restrict__ xr, long long g0, int len,
                                           long long T, int pad_mode, bool vec_ok) {
    const long long lo_ll = g0 < 0 ? -g0 : 0, hi_ll = T - g0;
    const int lo = (int)(lo_ll < len ? lo_ll : len);                        // in-range samples: j in [lo, hi)
    const int hi = (int)(hi_ll < lo ? lo : (hi_ll < len ? hi_ll : len));
    int ja = lo, jb = lo;                                                   // [ja, jb): 16-byte copies
    if (vec_ok && (g0 & 3) == 0) {
        ja = (lo + 3) & ~3; jb = hi & ~3;
        if (jb < ja) ja = jb = lo;
    }
    if (jb > ja) {
        const float4* src = reinterpret_cast<const float4*>(xr) + (g0 >> 2);   // g0 is a multiple of 4 (possibly negative)
        float4* dst = reinterpret_cast<float4*>(tile);
        for (int i = (ja >> 2) + threadIdx.x; i < (jb >> 2); i += blockDim.x) dst[i] = __ldg(src + i);
    }
    for (int j = threadIdx.x; j < ja; j += blockDim.x) tile[j] = padded_sample(xr, g0 + j, T, pad_mode);
    for (int j = jb + threadIdx.x; j < len; j += blockDim.x) tile[j] = padded_sample(xr, g0 + j, T, pad_mode);
}

#ifndef TAC_PLANES_PERSIST
#define TAC_PLANES_PERSIST 1  // calls with up to this many scales run one persistent 16-warp CTA per SM with a double-buffered, bulk-copied sample tile (0: never)
#endif
#ifndef TAC_PLANES_PF
#define TAC_PLANES_PF 2       // one-shot kernel: prefetch the tile of unit blockIdx.x + PF * SMs into the L2 (0: off)
#endif
#ifndef TAC_PLANES_BULK
#define TAC_PLANES_BULK 1     // 1: a task's two plane spans leave through shared memory as bulk async copies; 0: 4-byte streaming stores
#endif

__device__ __forceinline__ void bulk_wait_read() {            // this thread's bulk copies have finished READING shared memory
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

// One contiguous span of `len` floats, staged at stg[mis .. mis + len) with mis = (global float address) & 3 so that the
// 16-byte aligned middle of the span is 16-byte aligned in shared memory too: lane 0 hands the middle to the copy engine
// (cp.async.bulk.global.shared::cta, issued by lane `issuer`), lanes 8.. store the <= 3 + 3 head / tail floats.  Called by the
// whole warp after a fence.proxy.async + __syncwarp.
__device__ __forceinline__ void bulk_store_span(float* gdst, const float* stg, int mis, int len, int lane, int issuer) {
    const int head = min((4 - mis) & 3, len);
    const int mid = (len - head) & ~3;
    const int tail = len - head - mid;
    if (lane == issuer && mid > 0) {
        const unsigned src = (unsigned)__cvta_generic_to_shared(stg + mis + head);
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                     :: "l"(gdst + head), "r"(src), "r"(mid * 4) : "memory");
    }
    const int e = lane - 8;
    if (e >= 0 && e < head) __stcs(gdst + e, stg[mis + e]);
    else if (e >= head && e < head + tail) __stcs(gdst + mid + e, stg[mis + mid + e]);
}

// One-plane epilogues of the multi-scale kernel (tac_stft_multi_epi): |z|^p0, or p1 * log10(|z|^p0 + p2).
struct MagEpi {
    int kind;          // TAC_EPI_POWER | TAC_EPI_LOGPOWER
    float p0, p1, p2;
};
// EP 1: |z| -- EP 2: |z|^2 -- EP 3: anything else (general exponent, log form); compile-time so that the two common forms carry
// neither the powf / log10f bodies nor a branch per bin.  `e` lives in registers (a by-value copy made once per thread: a
// pointer to the kernel parameter would turn every field read into a local-memory load on the L1TEX pipe that bounds K1).
template <int EP>
__device__ __forceinline__ float mag_epilogue(const MagEpi& e, float re, float im) {
    const float m2 = fmaf(re, re, im * im);                // the arithmetic of stft_ex.cu: ex_pow_mag
    if constexpr (EP == 2) return m2;
    else if constexpr (EP == 1) return sqrtf(m2);
    else {
        float v;
        if (e.p0 == 2.0f) v = m2;
        else {
            const float m = sqrtf(m2);
            v = e.p0 == 1.0f ? m : (e.p0 == 0.5f ? sqrtf(m) : (m > 0.0f ? powf(m, e.p0) : 0.0f));
        }
        return e.kind == TAC_EPI_LOGPOWER ? e.p1 * log10f(v + e.p2) : v;
    }
}

// One warp task of the planes kernel: frames [f_first, f_first + NFR) of one scale.  MAG: ONE plane per (item, channel)
// holding mag_epilogue(z) instead of the re / im pair (half the bytes written; the `im` staging and spans are not touched).
template <class P, int EP = 0>
__device__ __forceinline__ void planes_task(const ScaleArg& sc, const float* tile, float2* xbuf, int lane,
                                            int f_first, int f_end, int tile_origin /* f*hop coordinate of tile[0] */,
                                            float* out_re_row, long long im_off, const MagEpi me = MagEpi()) {
    constexpr bool MAG = EP != 0;
    cf z[P::BQ];
    const float2* __restrict__ win2 = sc.t.win2;
    const int hop = sc.hop;
    auto load = [&](auto, int fr, int m, auto) -> cf {
        int f = f_first + fr;
        f = f < f_end ? f : f_first;                       // invalid frames recompute the first one
        // a plain product (mul.rn, never contracted): X(x) is bit-reproducible
        return mul2(*reinterpret_cast<const cf*>(tile + (f * hop - tile_origin) + 2 * m), ldg_cf(win2 + m));
    };
    // How a task's output leaves (A/B per plan on B200, cfg 4 single-scale, ms: direct 4-byte stores / one span per task /
    // one row per frame): n = 2048: 0.628 / 0.472 / =; 1024: 0.613 / 0.546 / 0.491; 512: 0.597 / 0.563 / 0.563;
    // 256: 0.616 / 0.569 / 0.665; 128: 0.669 / 0.704 / 1.000 -- the copy engine wins once a row is ~2 KB.
    constexpr int kMode = !TAC_PLANES_BULK ? 0 : (P::M >= 512 ? 2 : (P::M >= 128 ? 1 : 0));
#if TAC_PLANES_BULK
    if (lane < 8) bulk_wait_read();                        // the previous task's rows (any scale: up to 8 issuing lanes) have left
    __syncwarp();                                          // the exchange tile
#endif
    warp_fft_fwd_pk<P>(z, xbuf, sc.t.tw, lane, load);
    const int myf = f_first + lane / P::A;
    const bool valid = myf < f_end;
    const int k1 = lane % P::A;
    if constexpr (kMode != 0) {
    // The exchange tile is free after the forward transform: it stages the task's output rows (one frame of one plane = K
    // contiguous floats in HBM, because K is the innermost dimension; the frames of a task follow each other).  A row (mode 2)
    // or the whole span of the task (mode 1) is staged at its global misalignment (address & 3 floats) so that its 16-byte
    // aligned middle is aligned in shared memory too and can be handed to the copy engine.  Mode 2: rows of different frames
    // sit SF floats apart with SF = A (mod 32), so the A lanes of every frame hit their own bank range (K = 1 (mod 32) itself
    // puts frame g one bank off frame g - 1: 2- to 4-way conflicts, 45 M extra wavefronts on cfg 4 -- accepted in mode 1,
    // where per-row copies of 0.25 - 0.5 KB cost more than the conflicts).
    constexpr int SF = kMode == 1 ? P::M + 1 : (P::NFR == 1 ? ((P::M + 1 + 3 + 3) & ~3) : ((P::M + 1 + 3 - P::A + 31) / 32 * 32 + P::A));
    constexpr int R = kMode == 1 ? ((P::NFR * (P::M + 1) + 3 + 3) & ~3) : P::NFR * SF;   // floats per staged plane
    static_assert(2 * R * sizeof(float) <= kXbufFloat2 * sizeof(float2), "staged rows do not fit the exchange tile");
    float* stg_re = reinterpret_cast<float*>(xbuf);
    float* stg_im = stg_re + R;
    float* g_re0 = out_re_row + (long long)f_first * sc.K;
    const int fr = lane / P::A;
    const long long goff = kMode == 1 ? 0 : (long long)fr * sc.K;          // mode 1: one shift for the task's span
    const int mis_re = (int)((reinterpret_cast<uintptr_t>(g_re0 + goff) >> 2) & 3u);
    const int mis_im = (int)((reinterpret_cast<uintptr_t>(g_re0 + im_off + goff) >> 2) & 3u);
    float* s_re = stg_re + fr * SF + mis_re + k1;                          // bins k1 + A j
    float* s_im = stg_im + fr * SF + mis_im + k1;
    float* h_re = stg_re + fr * SF + mis_re + (P::M - k1);                 // bins M - k1 - A j (consecutive lanes: consecutive banks)
    float* h_im = stg_im + fr * SF + mis_im + (P::M - k1);
    warp_fft_post_pk<P, true>(z, sc.t.pt, lane, [&](int, int, cf v, int slot) {
        float vr = lo(v);
        const float vi = hi(v);
        if constexpr (MAG) vr = mag_epilogue<EP>(me, vr, vi);
        if (valid) {
            if (slot == P::BQ) {                       // k = M/2 (k1 == 0 lanes only)
                s_re[P::M / 2] = vr;
                if constexpr (!MAG) s_im[P::M / 2] = vi;
            } else if (slot & 1) {                     // k = M - k1 - A*j
                h_re[-P::A * (slot >> 1)] = vr;
                if constexpr (!MAG) h_im[-P::A * (slot >> 1)] = vi;
            } else {                                   // k = k1 + A*j
                s_re[P::A * (slot >> 1)] = vr;
                if constexpr (!MAG) s_im[P::A * (slot >> 1)] = vi;
            }
        }
    });
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy stores -> visible to the copy engine
    __syncwarp();
    const int nv = min(P::NFR, f_end - f_first);
    if constexpr (kMode == 1) {
        const int mr = (int)((reinterpret_cast<uintptr_t>(g_re0) >> 2) & 3u), mi = (int)((reinterpret_cast<uintptr_t>(g_re0 + im_off) >> 2) & 3u);
        bulk_store_span(g_re0, stg_re, mr, nv * sc.K, lane, 0);
        if constexpr (!MAG) bulk_store_span(g_re0 + im_off, stg_im, mi, nv * sc.K, lane, 0);
    } else {
#pragma unroll
        for (int g = 0; g < P::NFR; ++g) {
            if (g < nv) {
                float* gr = g_re0 + (long long)g * sc.K;
                float* gi = gr + im_off;
                bulk_store_span(gr, stg_re + g * SF, (int)((reinterpret_cast<uintptr_t>(gr) >> 2) & 3u), sc.K, lane, g);
                if constexpr (!MAG) bulk_store_span(gi, stg_im + g * SF, (int)((reinterpret_cast<uintptr_t>(gi) >> 2) & 3u), sc.K, lane, g);
            }
        }
    }
    if (lane < P::NFR) asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    } else {
    // per-lane base pointers so that every store address is base + compile-time offset
    float* re_lo = out_re_row + (long long)myf * sc.K + k1;
    float* re_hi = out_re_row + (long long)myf * sc.K + (P::M - k1);
    float* im_lo = re_lo + im_off;
    float* im_hi = re_hi + im_off;
    warp_fft_post_pk<P, true>(z, sc.t.pt, lane, [&](int, int, cf v, int slot) {
        float vr = lo(v);
        const float vi = hi(v);
        if constexpr (MAG) vr = mag_epilogue<EP>(me, vr, vi);
        if (valid) {
            if (slot == P::BQ) {                       // k = M/2 (k1 == 0 lanes only)
                __stcs(re_lo + P::M / 2, vr);
                if constexpr (!MAG) __stcs(im_lo + P::M / 2, vi);
            } else if (slot & 1) {                     // k = M - k1 - A*j
                __stcs(re_hi - P::A * (slot >> 1), vr);
                if constexpr (!MAG) __stcs(im_hi - P::A * (slot >> 1), vi);
            } else {                                   // k = k1 + A*j
                __stcs(re_lo + P::A * (slot >> 1), vr);
                if constexpr (!MAG) __stcs(im_lo + P::A * (slot >> 1), vi);
            }
        }
    });
    }
}

// All tasks of one (row, chunk) unit that belong to this warp.  Tasks (NFR frames of one scale) are numbered scale by scale and
// dealt round-robin: warp w of W takes the global task numbers t = w (mod W).  What depends on the scale only -- frame range,
// task count, plane pointers -- is computed ONCE per unit, lane s for scale s (32-bit arithmetic), and broadcast by shuffles;
// the plan switch sits outside the task loop.  (Round 2 history, cfg 4, 128 tasks per unit: a shared atomic counter + shuffle +
// scale search + 64-bit plane arithmetic PER TASK were 17 % of the kernel's instructions; hoisted per scale with 64-bit divisions
// they were still ~250 instructions per scale and warp.)
template <int W, int EP = 0>
__device__ __forceinline__ void planes_unit(const PlanesArgs& a, const float* tile, float2* xbuf, int lane, int warp,
                                            unsigned row, unsigned c0, const MagEpi me = MagEpi()) {
    constexpr bool MAG = EP != 0;
    static_assert((W & (W - 1)) == 0, "W must be a power of two");
    static_assert(Plan2048::NFR == 1 && Plan1024::NFR == 2 && Plan512::NFR == 2 && Plan256::NFR == 4 && Plan128::NFR == 4 && Plan64::NFR == 8, "nfr");
    int my_lo = 0, my_hi = 0, my_nt = 0;
    float* my_out = nullptr;
    long long my_im = 0;
    if (lane < a.nscales) {
        const ScaleArg& sc = a.sc[lane];
        const unsigned h = (unsigned)sc.hop;
        unsigned lo = (c0 + h - 1u) / h, hi = (c0 + (unsigned)a.chunk + h - 1u) / h;
        if (hi > (unsigned)sc.F) hi = (unsigned)sc.F;
        if (lo > hi) lo = hi;
        const int l2nfr = sc.n >= 2048 ? 0 : (sc.n >= 512 ? 1 : (sc.n >= 128 ? 2 : 3));
        my_lo = (int)lo; my_hi = (int)hi;
        my_nt = (int)((hi - lo + (1u << l2nfr) - 1u) >> l2nfr);
        const unsigned b = row / (unsigned)a.C, ch = row - b * (unsigned)a.C;
        const long long plane = (long long)sc.F * sc.K;
        my_out = sc.out + ((long long)(b * (MAG ? 1u : 2u)) * a.C + ch) * plane;
        my_im = (long long)a.C * plane;
    }
    int my_tb = my_nt;                                       // exclusive prefix sum over the (<= 8) scales
#pragma unroll
    for (int o = 1; o < TAC_MAX_SCALES; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, my_tb, o);
        if (lane >= o) my_tb += v;
    }
    my_tb -= my_nt;
    for (int s = 0; s < a.nscales; ++s) {
        const int nt = __shfl_sync(0xffffffffu, my_nt, s);
        const int first = (warp - __shfl_sync(0xffffffffu, my_tb, s)) & (W - 1);   // first local task with global number = warp (mod W)
        if (first >= nt) continue;                           // warp-uniform
        const int f_lo = __shfl_sync(0xffffffffu, my_lo, s), f_end = __shfl_sync(0xffffffffu, my_hi, s);
        float* out_re_row = reinterpret_cast<float*>(__shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(my_out), s));
        const long long im_off = __shfl_sync(0xffffffffu, my_im, s);
        const ScaleArg& sc = a.sc[s];
        // tile[0] is padded-coordinate (c0 - pmax) of the widest scale == (c0 - pmax + pad_s) of scale s
        const int origin = (int)c0 - a.pmax + sc.pad;
        auto my_tasks = [&](auto&& body) {
            for (int t = first; t < nt; t += W) body(t);
        };
        switch (sc.n) {
            case 2048: my_tasks([&](int t) { planes_task<Plan2048, EP>(sc, tile, xbuf, lane, f_lo + t * 1, f_end, origin, out_re_row, im_off, me); }); break;
            case 1024: my_tasks([&](int t) { planes_task<Plan1024, EP>(sc, tile, xbuf, lane, f_lo + t * 2, f_end, origin, out_re_row, im_off, me); }); break;
            case 512: my_tasks([&](int t) { planes_task<Plan512, EP>(sc, tile, xbuf, lane, f_lo + t * 2, f_end, origin, out_re_row, im_off, me); }); break;
            case 256: my_tasks([&](int t) { planes_task<Plan256, EP>(sc, tile, xbuf, lane, f_lo + t * 4, f_end, origin, out_re_row, im_off, me); }); break;
            case 128: my_tasks([&](int t) { planes_task<Plan128, EP>(sc, tile, xbuf, lane, f_lo + t * 4, f_end, origin, out_re_row, im_off, me); }); break;
            default: my_tasks([&](int t) { planes_task<Plan64, EP>(sc, tile, xbuf, lane, f_lo + t * 8, f_end, origin, out_re_row, im_off, me); }); break;
        }
    }
}

__global__ void __launch_bounds__(kThreads, 2) stft_planes_kernel(const PlanesArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* tile = reinterpret_cast<float*>(smem_raw);
    float2* xbuf_all = reinterpret_cast<float2*>(smem_raw + (size_t)a.tile_len * sizeof(float));

    const unsigned c = blockIdx.x % (unsigned)a.nchunks;
    const unsigned row = blockIdx.x / (unsigned)a.nchunks;
    const float* xr = a.x + (long long)row * a.T;
    const unsigned c0 = c * (unsigned)a.chunk;               // chunk origin in padded frame-start coordinates (T < 2^31 - 2^16)

#if TAC_PLANES_PF
    {   // L2 prefetch (evict-last) of the tile of the CTA that will take this one's place: CTAs are dispatched in blockIdx order,
        // 2 per SM, so unit blockIdx.x + 2 * SMs starts about when this one ends.  Its staging loads then hit the L2 instead of
        // queueing behind the plane writes in DRAM (the staging wait + barrier were 14 % of the warp-time samples on cfg 4).
        const unsigned u2 = blockIdx.x + TAC_PLANES_PF * kNumSMs;
        if (u2 < gridDim.x) {
            const unsigned row2 = u2 / (unsigned)a.nchunks;
            const long long g2 = (long long)(u2 - row2 * (unsigned)a.nchunks) * a.chunk - a.pmax;
            const long long lo = g2 < 0 ? 0 : g2;
            long long hi = g2 + a.tile_len; hi = hi > a.T ? a.T : hi;
            const char* p = reinterpret_cast<const char*>(a.x + (long long)row2 * a.T + lo);
            const long long bytes = (hi - lo) * 4;
            for (long long off = (long long)threadIdx.x * 128; off < bytes; off += (long long)kThreads * 128)
                asm volatile("prefetch.global.L2::evict_last [%0];" :: "l"(p + off));
        }
    }
#endif
    stage_tile(tile, xr, (long long)c0 - a.pmax, a.tile_len, a.T, a.pad_mode, a.vec_ok != 0);
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    planes_unit<kWarps>(a, tile, xbuf_all + warp * kXbufFloat2, lane, warp, row, c0);
#if TAC_PLANES_BULK
    if (lane < 8) bulk_wait_read();                        // shared memory must outlive the copy engine's reads
#endif
}

// The same unit with the one-plane epilogue (tac_stft_multi_epi): [B, C, F, K] of |z|^p or p1 log10(|z|^p + p2) per scale.
template <int EP>
__global__ void __launch_bounds__(kThreads, 2) stft_planes_mag_kernel(const PlanesArgs a, const MagEpi me_param) {
    const MagEpi me = me_param;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* tile = reinterpret_cast<float*>(smem_raw);
    float2* xbuf_all = reinterpret_cast<float2*>(smem_raw + (size_t)a.tile_len * sizeof(float));
    const unsigned c = blockIdx.x % (unsigned)a.nchunks;
    const unsigned row = blockIdx.x / (unsigned)a.nchunks;
    const unsigned c0 = c * (unsigned)a.chunk;
    stage_tile(tile, a.x + (long long)row * a.T, (long long)c0 - a.pmax, a.tile_len, a.T, a.pad_mode, a.vec_ok != 0);
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    planes_unit<kWarps, EP>(a, tile, xbuf_all + warp * kXbufFloat2, lane, warp, row, c0, me);
#if TAC_PLANES_BULK
    if (lane < 8) bulk_wait_read();                        // shared memory must outlive the copy engine's reads
#endif
}

// ------------------------------------------------------------------------------------------
// Persistent form of the planes kernel: ONE 16-warp CTA per SM walks the (row, chunk) units blockIdx.x, + gridDim.x, ...; the
// sample tile is double-buffered and filled by the copy engine (cp.async.bulk global -> shared, completion on an mbarrier), so
// the tile of unit i + 1 (and, once every warp has released unit i's buffer, of unit i + 2) lands while unit i is being
// transformed.  full[b]: 1 arrival (the producer's, carrying the byte count); empty[b]: one arrival per warp.  Warps drift:
// a warp that finishes its tasks of unit i starts unit i + 1 as soon as that tile has landed, no block barrier in the loop.
// Why: in the one-shot kernel every CTA began with a DRAM read that queues behind 3 TB/s of plane writes (~9 us of a 63 us
// CTA life; with the barrier behind it 14 % of the warp-time samples of cfg 4).
// Used when the tile can be bulk-copied: 16-byte aligned rows (vec_ok), zero (or no) padding.
constexpr int kPWarps = 16;
constexpr int kPThreads = kPWarps * 32;

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {     // bounded: traps instead of hanging the GPU; the
                                                                                 // hint lets a waiting warp sleep up to 2 us per probe
    const unsigned addr = smem_u32(bar);
    for (unsigned it = 0; it < (1u << 26); ++it) {
        unsigned done;
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, 2000;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                     : "=r"(done) : "r"(addr), "r"(parity) : "memory");
        if (done) return;
    }
    __trap();
}

__global__ void __launch_bounds__(kPThreads, 1) stft_planes_persist_kernel(const PlanesArgs a, const long long nunits) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* tiles = reinterpret_cast<float*>(smem_raw);                                        // [2][tile_len]
    float2* xbuf_all = reinterpret_cast<float2*>(smem_raw + 2 * (size_t)a.tile_len * sizeof(float));
    __shared__ __align__(8) uint64_t s_full[2], s_empty[2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float2* xbuf = xbuf_all + warp * kXbufFloat2;
    if (threadIdx.x == 0) {
        for (int b = 0; b < 2; ++b) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(&s_full[b])), "r"(1) : "memory");
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(&s_empty[b])), "r"(kPWarps) : "memory");
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int n_my = (int)((nunits - blockIdx.x + gridDim.x - 1) / gridDim.x);

    // warp 0: tile of this CTA's i-th unit -> buffer buf (zero outside the row, bulk copy of the in-range part)
    auto issue = [&](int i, int buf) {
        const long long u = blockIdx.x + (long long)i * gridDim.x;
        const long long row = u / a.nchunks;
        const long long g0 = (u % a.nchunks) * a.chunk - a.pmax;
        const int len = a.tile_len;
        const long long lo_ll = g0 < 0 ? -g0 : 0, hi_ll = a.T - g0;
        const int lo = (int)(lo_ll < len ? lo_ll : len);
        const int hi = (int)(hi_ll < lo ? lo : (hi_ll < len ? hi_ll : len));                  // g0, T, len are multiples of 4: so are lo, hi
        float4* t4 = reinterpret_cast<float4*>(tiles + (size_t)buf * len);
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int j = lane; j < (lo >> 2); j += 32) t4[j] = z;
        for (int j = (hi >> 2) + lane; j < (len >> 2); j += 32) t4[j] = z;
        __syncwarp();
        if (lane == 0) {
            const unsigned bar = smem_u32(&s_full[buf]);
            const unsigned bytes = (unsigned)(hi - lo) * 4u;
            if (bytes) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             :: "r"(smem_u32(tiles + (size_t)buf * len + lo)), "l"(a.x + row * a.T + g0 + lo), "r"(bytes), "r"(bar) : "memory");
            } else {
                asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(bar) : "memory");
            }
        }
    };
    if (warp == 0) {
        issue(0, 0);
        if (n_my > 1) issue(1, 1);
    }
    for (int i = 0; i < n_my; ++i) {
        const int buf = i & 1;
        const unsigned par = (unsigned)(i >> 1) & 1u;
        const unsigned u = blockIdx.x + (unsigned)i * gridDim.x;                              // nunits < 2^31
        const unsigned row = u / (unsigned)a.nchunks;
        const unsigned c0 = (u - row * (unsigned)a.nchunks) * (unsigned)a.chunk;
        mbar_wait(&s_full[buf], par);
        planes_unit<kPWarps>(a, tiles + (size_t)buf * a.tile_len, xbuf, lane, warp, row, c0);
        __syncwarp();
        if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(&s_empty[buf])) : "memory");
        if (warp == 0 && i + 2 < n_my) {
            mbar_wait(&s_empty[buf], par);                   // every warp has finished unit i: its buffer takes unit i + 2
            issue(i + 2, buf);
        }
    }
#if TAC_PLANES_BULK
    if (lane < 8) bulk_wait_read();                        // shared memory must outlive the copy engine's reads
#endif
}

// ------------------------------------------------------------------------------------------
struct TiledArgs {
    const float* x;
    long long T;
    int n, hop, pad, F, K, TF, ntiles, tile_len, pad_mode, vec_ok;
    long long f_pad;                 // time extent of the output rows (>= F)
    FftTables t;
    float* out;
};

template <class P, int MODE>
__global__ void __launch_bounds__(kThreads, 1) stft_tiled_kernel(const TiledArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* tile = reinterpret_cast<float*>(smem_raw);
    float* otile = tile + a.tile_len;                               // [2][K][TF+1]
    float2* xbuf_all = reinterpret_cast<float2*>(otile + ((2 * (size_t)a.K * (a.TF + 1) + 3) & ~(size_t)3));   // 16-byte aligned
    __shared__ int s_next;

    const int ft = blockIdx.x % a.ntiles;
    const long long row = blockIdx.x / a.ntiles;
    const float* xr = a.x + row * a.T;
    const int f0 = ft * a.TF;
    const int f_end = min(a.F, f0 + a.TF);
    const long long g0 = (long long)f0 * a.hop - a.pad;
    stage_tile(tile, xr, g0, a.tile_len, a.T, a.pad_mode, a.vec_ok != 0);
    if (threadIdx.x == 0) s_next = 0;
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float2* xbuf = xbuf_all + warp * kXbufFloat2;
    const int ntasks = (f_end - f0 + P::NFR - 1) / P::NFR;
    const int TS = a.TF + 1;
    float* o0 = otile;
    float* o1 = otile + (size_t)a.K * TS;
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
            return mul2(*reinterpret_cast<const cf*>(tile + (f - f0) * a.hop + 2 * m), ldg_cf(win2 + m));   // plain product
        };
        warp_fft_fwd_pk<P>(z, xbuf, a.t.tw, lane, load);
        const int myf = f_first + lane / P::A;
        const bool valid = myf < f_end;
        const int col = myf - f0;
        warp_fft_post_pk<P, true>(z, a.t.pt, lane, [&](int, int k, cf v, int) {
            const float vr = lo(v), vi = hi(v);
            if (valid) {
                if (MODE == TAC_OUT_MAGPHASE) {
                    o0[k * TS + col] = sqrtf(fmaf(vr, vr, vi * vi));
                    o1[k * TS + col] = atan2f(vi, vr);
                } else {
                    o0[k * TS + col] = vr;
                    o1[k * TS + col] = vi;
                }
            }
        });
    }
    __syncthreads();
    const int nf = f_end - f0;
    const int total = a.K * nf;
    if (MODE == TAC_OUT_MAGPHASE) {
        float* out0 = a.out + (row * 2) * (long long)a.K * a.f_pad + f0;
        float* out1 = out0 + (long long)a.K * a.f_pad;
        for (int i = threadIdx.x; i < total; i += blockDim.x) {
            const int k = i / nf, tt = i - k * nf;
            __stcs(out0 + (long long)k * a.f_pad + tt, o0[k * TS + tt]);
            __stcs(out1 + (long long)k * a.f_pad + tt, o1[k * TS + tt]);
        }
    } else {
        float2* outc = reinterpret_cast<float2*>(a.out) + row * (long long)a.K * a.F + f0;
        for (int i = threadIdx.x; i < total; i += blockDim.x) {
            const int k = i / nf, tt = i - k * nf;
            __stcs(outc + (long long)k * a.F + tt, make_float2(o0[k * TS + tt], o1[k * TS + tt]));
        }
    }
}

// ------------------------------------------------------------------------------------------
// Front-end kernel (TAC_OUT_MAGPHASE): kFeWarps warps per CTA, one warp task (NFR frames) per warp, so a CTA covers
// TF = kFeWarps * NFR consecutive frames of one row.  The per-warp exchange tiles are only needed during the transform and
// the two [K][TF + 1] transposition tiles (magnitude, phase) only after it, so they live in the SAME shared memory; both
// planes leave in one pass, every (k) row as a contiguous run of TF frames.  With 8 warps a CTA needs 96 KiB and two are
// resident per SM: one computes while the other stages its samples or writes out.  (Recorded negative results, A/B on B200:
// 128-bit write-out stores, 4 frames of one row per instruction, 0.109 -> 0.114 ms; 4-warp CTAs, four per SM, 0.118 ms --
// the kernel waits on its phase latencies, not on the store path.)  atan2 is a degree-15 odd minimax
// polynomial (1.2e-7 abs. error in fp32) instead of the library call.
#ifndef TAC_FE_WARPS
#define TAC_FE_WARPS 8          // 8-warp CTAs (96 KiB, two per SM): one CTA computes while the other stages / writes out -- 0.136 -> 0.109 ms on cfg 2 against 16-warp CTAs
#endif
constexpr int kFeWarps = TAC_FE_WARPS;
constexpr int kFeOcc = kFeWarps <= 4 ? 4 : (kFeWarps <= 8 ? 2 : 1);   // 8-warp CTAs: 96 KiB each, two per SM
constexpr int kFeThreads = kFeWarps * 32;

__device__ __forceinline__ float fe_rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float fe_rsqrt(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

__device__ __forceinline__ float fast_atan2(float y, float x) {
    const float ax = fabsf(x), ay = fabsf(y);
    const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
    const float a = mx > 0.0f ? mn * fe_rcp(mx) : 0.0f;           // in [0, 1]
    const float s = a * a;
    float p = -0.004054559860378504f;
    p = fmaf(p, s, 0.021862933412194252f);
    p = fmaf(p, s, -0.05591228976845741f);
    p = fmaf(p, s, 0.09642194956541061f);
    p = fmaf(p, s, -0.1390862911939621f);
    p = fmaf(p, s, 0.19946564733982086f);
    p = fmaf(p, s, -0.33329859375953674f);
    p = fmaf(p, s, 0.9999993443489075f);
    float r = p * a;                                              // atan(min / max)
    r = ay > ax ? 1.57079632679489661923f - r : r;
    r = x < 0.0f ? 3.14159265358979323846f - r : r;
    return copysignf(r, y);
}

template <class P>
__global__ void __launch_bounds__(kFeThreads, kFeOcc) stft_frontend_kernel(const TiledArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* tile = reinterpret_cast<float*>(smem_raw);
    unsigned char* shared2 = smem_raw + (size_t)a.tile_len * sizeof(float);      // exchange tiles, then the transposition tile
    float2* xbuf_all = reinterpret_cast<float2*>(shared2);
    float* otile = reinterpret_cast<float*>(shared2);                            // [K][TF + 1]

    constexpr int TF = kFeWarps * P::NFR, TS = TF + 1;
    const int ft = blockIdx.x % a.ntiles;
    const long long row = blockIdx.x / a.ntiles;
    const float* xr = a.x + row * a.T;
    const int f0 = ft * TF;
    const int f_end = min(a.F, f0 + TF);
    stage_tile(tile, xr, (long long)f0 * a.hop - a.pad, a.tile_len, a.T, a.pad_mode, a.vec_ok != 0);
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float2* xbuf = xbuf_all + warp * kXbufFloat2;
    const int f_first = f0 + warp * P::NFR;
    const bool task = f_first < f_end;                             // warp-uniform
    cf z[P::BQ];
    const float2* __restrict__ win2 = a.t.win2;
    if (task) {
        auto load = [&](auto, int fr, int m, auto) -> cf {
            int f = f_first + fr;
            f = f < f_end ? f : f_first;
            return mul2(*reinterpret_cast<const cf*>(tile + (f - f0) * a.hop + 2 * m), ldg_cf(win2 + m));   // plain product
        };
        warp_fft_fwd_pk<P>(z, xbuf, a.t.tw, lane, load);
    }
    __syncthreads();                                               // every warp is done with its exchange tile
    const int myf = f_first + lane / P::A;
    const bool valid = task && myf < f_end;
    const int col = myf - f0;
    const int nf = f_end - f0;
    float* out0 = a.out + (row * 2) * (long long)a.K * a.f_pad + f0;
    float* otile1 = otile + (size_t)a.K * TS;                      // phase plane next to the magnitude plane
    if (task) {
        warp_fft_post_pk<P, true>(z, a.t.pt, lane, [&](int, int k, cf c, int) {
            const float vr = lo(c), vi = hi(c);
            const float m2 = fmaf(vr, vr, vi * vi);
            const float mag = m2 >= 1e-30f ? m2 * fe_rsqrt(m2) : 0.0f;   // |X| (<= 2 ulp); |X| < 1e-15 reads as 0
            const float ph = fast_atan2(vi, vr);
            if (valid) { otile[k * TS + col] = mag; otile1[k * TS + col] = ph; }
        });
    }
    __syncthreads();
    // write-out: thread -> (frame tt, bins k0, k0 + RPI, ...): the frame index and both row pointers are loop invariants, the
    // shared-memory offsets compile-time constants (it was 23 % of the kernel's instructions as i / TF, i % TF, 64-bit address
    // arithmetic per element)
    constexpr int RPI = kFeThreads / TF;                           // bins per pass of the CTA
    static_assert(kFeThreads % TF == 0, "TF must divide the CTA size");
    const int tt = threadIdx.x % TF, k0 = threadIdx.x / TF;
    if (tt < nf) {
        const float* s0 = otile + k0 * TS + tt;
        const float* s1 = otile1 + k0 * TS + tt;
        float* g0p = out0 + (long long)k0 * a.f_pad + tt;
        float* g1p = g0p + (long long)a.K * a.f_pad;
        const long long gstep = (long long)RPI * a.f_pad;
        const int K = a.K;
#pragma unroll 4
        for (int k = k0; k < K; k += RPI) {
            __stcs(g0p, *s0); __stcs(g1p, *s1);
            s0 += RPI * TS; s1 += RPI * TS; g0p += gstep; g1p += gstep;
        }
    }
}

template <class P>
static int launch_frontend(TiledArgs a, long long rows, cudaStream_t st, bool* done) {
    const int TF = kFeWarps * P::NFR;
    a.TF = TF;
    a.ntiles = (a.F + TF - 1) / TF;
    a.tile_len = ((TF - 1) * a.hop + a.n + 3) & ~3;
    const size_t xb = (size_t)kFeWarps * kXbufFloat2 * sizeof(float2);
    const size_t ot = (size_t)a.K * (TF + 1) * sizeof(float);
    const size_t smem = (size_t)a.tile_len * 4 + (xb > 2 * ot ? xb : 2 * ot);     // both planes leave through one pass
    *done = false;
    if (smem > 200 * 1024 || rows * a.ntiles >= (1ll << 31)) return TAC_OK;      // caller falls back to the generic kernel
    auto kern = stft_frontend_kernel<P>;
    TAC_SMEM_OPT_IN(kern, 200 * 1024);
    kern<<<(unsigned)(rows * a.ntiles), kFeThreads, smem, st>>>(a); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    *done = true;
    return TAC_OK;
}

static int dispatch_frontend(const TiledArgs& a, long long rows, cudaStream_t st, bool* done) {
    switch (a.n) {
        case 2048: return launch_frontend<Plan2048>(a, rows, st, done);
        case 1024: return launch_frontend<Plan1024>(a, rows, st, done);
        case 512: return launch_frontend<Plan512>(a, rows, st, done);
        case 256: return launch_frontend<Plan256>(a, rows, st, done);
        case 128: return launch_frontend<Plan128>(a, rows, st, done);
        default: return launch_frontend<Plan64>(a, rows, st, done);
    }
}

template <class P, int MODE>
static int launch_tiled(const TiledArgs& a, long long rows, size_t smem, cudaStream_t st) {
    auto kern = stft_tiled_kernel<P, MODE>;
    TAC_SMEM_OPT_IN(kern, 200 * 1024);                 // per instantiation and device
    const long long grid = rows * a.ntiles;
    kern<<<(unsigned)grid, kThreads, smem, st>>>(a); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

template <int MODE>
static int dispatch_tiled(const TiledArgs& a, long long rows, size_t smem, cudaStream_t st) {
    switch (a.n) {
        case 2048: return launch_tiled<Plan2048, MODE>(a, rows, smem, st);
        case 1024: return launch_tiled<Plan1024, MODE>(a, rows, smem, st);
        case 512: return launch_tiled<Plan512, MODE>(a, rows, smem, st);
        case 256: return launch_tiled<Plan256, MODE>(a, rows, smem, st);
        case 128: return launch_tiled<Plan128, MODE>(a, rows, smem, st);
        default: return launch_tiled<Plan64, MODE>(a, rows, smem, st);
    }
}

// ------------------------------------------------------------------------------------------ dB channel
// librosa.amplitude_to_db(S, ref=np.max) per utterance (spectrogram.ipynb cell 9 L31-63, raw :2894-2913):
//   db = 20 log10(max(S, amin)) - 20 log10(max(amin, max S)),  floored at max(db) - top_db.
// Pass 1: per-row maximum of the magnitude plane (non-negative floats order like their bit patterns: atomicMax on
// unsigned).  Pass 2: element-wise.  max(db) = 20 log10(max(max S, amin)) - ref = 0 whenever max S >= amin.
__global__ void __launch_bounds__(256) db_rowmax_kernel(const float* __restrict__ mag, long long row_stride, int K, int F, long long ld,
                                                       unsigned* __restrict__ rowmax) {
    const long long row = blockIdx.y;
    const float* m = mag + row * row_stride;
    const long long total = (long long)K * F;
    float v = 0.0f;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
        const long long k = i / F, t = i - k * F;
        v = fmaxf(v, __ldg(m + k * ld + t));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    if ((threadIdx.x & 31) == 0) atomicMax(rowmax + row, __float_as_uint(v));
}

__global__ void __launch_bounds__(256) db_apply_kernel(const float* __restrict__ mag, long long row_stride, int K, int F, long long ld,
                                                      const unsigned* __restrict__ rowmax, float amin, float top_db,
                                                      float* __restrict__ out) {
    const long long row = blockIdx.y;
    const float* m = mag + row * row_stride;
    float* o = out + row * (long long)K * F;
    const float mx = fmaxf(__uint_as_float(rowmax[row]), amin);
    const float ref = 20.0f * log10f(mx);
    const float floor_db = -top_db;                                  // max(db) - top_db with max(db) = 0
    const long long total = (long long)K * F;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
        const long long k = i / F, t = i - k * F;
        const float db = 20.0f * log10f(fmaxf(__ldg(m + k * ld + t), amin)) - ref;
        o[i] = fmaxf(db, floor_db);
    }
}

}  // namespace tac

using namespace tac;

extern "C" int tac_amplitude_to_db(const float* mag, int64_t rows, int64_t row_stride, int K, int F, int64_t ld, float amin,
                                   float top_db, float* out, void* workspace, size_t workspace_bytes, tac_stream_t stream) {
    TAC_REQUIRE(mag && out && workspace, "null argument");
    TAC_REQUIRE(rows >= 1 && rows < 65536 && K >= 1 && F >= 1 && ld >= F, "bad shape (rows < 65536, ld >= F)");
    TAC_REQUIRE(amin > 0.0f && top_db >= 0.0f, "amin must be positive, top_db non-negative");
    if (workspace_bytes < (size_t)rows * sizeof(unsigned)) return set_error(TAC_EWORKSPACE, "workspace needs %zu bytes", (size_t)rows * sizeof(unsigned));
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    unsigned* rowmax = reinterpret_cast<unsigned*>(workspace);
    TAC_CHECK_CUDA(cudaMemsetAsync(rowmax, 0, (size_t)rows * sizeof(unsigned), st));
    long long gx = ((long long)K * F + 255) / 256;
    gx = gx < 1 ? 1 : (gx > 32 ? 32 : gx);
    db_rowmax_kernel<<<dim3((unsigned)gx, (unsigned)rows), 256, 0, st>>>(mag, row_stride, K, F, ld, rowmax); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    db_apply_kernel<<<dim3((unsigned)gx, (unsigned)rows), 256, 0, st>>>(mag, row_stride, K, F, ld, rowmax, amin, top_db, out); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}


static int stft_multi_impl(const float* x, int64_t B, int64_t C, int64_t T, int n_scales, const int* n_fft,
                           const int* hop, const int* win, int center, int pad_mode, int normalized,
                           int out_mode, int64_t f_pad, void* const* out, tac_stream_t stream, const MagEpi* me);

extern "C" int tac_stft_multi(const float* x, int64_t B, int64_t C, int64_t T, int n_scales, const int* n_fft,
                              const int* hop, const int* win, int center, int pad_mode, int normalized,
                              int out_mode, int64_t f_pad, void* const* out, tac_stream_t stream) {
    return stft_multi_impl(x, B, C, T, n_scales, n_fft, hop, win, center, pad_mode, normalized, out_mode, f_pad, out, stream, nullptr);
}

extern "C" int tac_stft_multi_epi(const float* x, int64_t B, int64_t C, int64_t T, int n_scales, const int* n_fft,
                                  const int* hop, const int* win, int center, int pad_mode, int normalized,
                                  const tac_epilogue* epi, void* const* out, tac_stream_t stream) {
    TAC_REQUIRE(epi != nullptr, "null epilogue");
    TAC_REQUIRE(epi->kind == TAC_EPI_POWER || epi->kind == TAC_EPI_LOGPOWER,
                "tac_stft_multi_epi forms TAC_EPI_POWER or TAC_EPI_LOGPOWER planes (kind %d: use tac_stft_ex per scale)", epi->kind);
    TAC_REQUIRE(epi->p0 > 0.0f, "exponent p0 must be positive");
    TAC_REQUIRE(epi->rowmax == nullptr, "rowmax is not produced by the multi-scale kernel");
    MagEpi me;
    me.kind = epi->kind; me.p0 = epi->p0; me.p1 = epi->p1; me.p2 = epi->p2;
    return stft_multi_impl(x, B, C, T, n_scales, n_fft, hop, win, center, pad_mode, normalized, TAC_OUT_PLANES, 0, out, stream, &me);
}

static int stft_multi_impl(const float* x, int64_t B, int64_t C, int64_t T, int n_scales, const int* n_fft,
                           const int* hop, const int* win, int center, int pad_mode, int normalized,
                           int out_mode, int64_t f_pad, void* const* out, tac_stream_t stream, const MagEpi* me) {
    TAC_REQUIRE(x && n_fft && hop && out, "null argument");
    TAC_REQUIRE(B > 0 && C > 0 && T > 0, "B, C, T must be positive (got %lld, %lld, %lld)", (long long)B, (long long)C, (long long)T);
    TAC_REQUIRE(n_scales >= 1 && n_scales <= TAC_MAX_SCALES, "n_scales must be in [1, %d]", TAC_MAX_SCALES);
    TAC_REQUIRE(out_mode >= 0 && out_mode <= 2, "unknown out_mode %d", out_mode);
    TAC_REQUIRE(pad_mode == TAC_PAD_ZERO || pad_mode == TAC_PAD_REFLECT, "unknown pad_mode %d", pad_mode);
    TAC_REQUIRE(B * C * T < (1ll << 40) && T < (1ll << 31) - 65536, "input too large");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const long long rows = B * C;
    const int vec_ok = ((reinterpret_cast<uintptr_t>(x) & 15u) == 0 && (T & 3) == 0) ? 1 : 0;
    TAC_REQUIRE((reinterpret_cast<uintptr_t>(x) & 3u) == 0, "x must be 4-byte aligned");

    int nmax = 0;
    for (int s = 0; s < n_scales; ++s) {
        const int w = win ? win[s] : n_fft[s];
        TAC_REQUIRE(supported_nfft(n_fft[s]), "unsupported n_fft %d", n_fft[s]);
        TAC_REQUIRE(hop[s] >= 2 && (hop[s] & 1) == 0 && hop[s] <= n_fft[s], "hop %d must be even and in [2, n_fft]", hop[s]);
        TAC_REQUIRE(w >= 1 && w <= n_fft[s], "win_length %d must be in [1, n_fft]", w);
        TAC_REQUIRE(out[s] != nullptr, "out[%d] is null", s);
        TAC_REQUIRE((reinterpret_cast<uintptr_t>(out[s]) & 7u) == 0, "out[%d] must be 8-byte aligned", s);
        if (center && pad_mode == TAC_PAD_REFLECT)
            TAC_REQUIRE(T > n_fft[s] / 2, "reflect padding needs T > n_fft/2");
        TAC_REQUIRE(num_frames(T, n_fft[s], hop[s], center) >= 1, "signal of %lld samples is shorter than n_fft=%d", (long long)T, n_fft[s]);
        nmax = n_fft[s] > nmax ? n_fft[s] : nmax;
    }

    if (out_mode == TAC_OUT_PLANES) {
        PlanesArgs a;
        a.x = x; a.T = T; a.C = (int)C; a.nscales = n_scales; a.pad_mode = pad_mode; a.vec_ok = vec_ok;
        // samples of a row per CTA: 8192 when that still leaves every SM a few CTAs per row batch (long clips: the tile load and
        // the launch are amortised over twice the tasks, cfg 4 2.20 -> 2.08 ms), else 4096 (1 s clips: 0.138 against 0.160 ms)
#ifdef TAC_PLANES_CHUNK
        a.chunk = TAC_PLANES_CHUNK;
#else
        a.chunk = (T >= 4 * 8192 && rows * ((T + 8191) / 8192) >= 8ll * 2 * kNumSMs) ? 8192 : 4096;
#endif
        a.pmax = center ? nmax / 2 : 0;
        long long last_start = 0;
        for (int s = 0; s < n_scales; ++s) {
            ScaleArg& sc = a.sc[s];
            sc.n = n_fft[s]; sc.hop = hop[s]; sc.pad = center ? n_fft[s] / 2 : 0;
            sc.F = (int)num_frames(T, sc.n, sc.hop, center);
            sc.K = sc.n / 2 + 1;
            sc.out = reinterpret_cast<float*>(out[s]);
            int rc = get_tables(sc.n, win ? win[s] : sc.n, normalized, &sc.t);
            if (rc) return rc;
            const long long ls = (long long)(sc.F - 1) * sc.hop;
            last_start = ls > last_start ? ls : last_start;
        }
        a.nchunks = (int)(last_start / a.chunk) + 1;
        a.tile_len = a.chunk + nmax;
        const size_t smem = (size_t)a.tile_len * sizeof(float) + (size_t)kWarps * kXbufFloat2 * sizeof(float2);
        const long long grid = rows * a.nchunks;
        TAC_REQUIRE(grid < (1ll << 31), "grid too large");
        if (me) {                                           // one-plane epilogue: the one-shot kernel, every scale count
            const int ep = me->kind == TAC_EPI_POWER ? (me->p0 == 1.0f ? 1 : (me->p0 == 2.0f ? 2 : 3)) : 3;
            if (ep == 1) {
                TAC_SMEM_OPT_IN(stft_planes_mag_kernel<1>, 111 * 1024);
                stft_planes_mag_kernel<1><<<(unsigned)grid, kThreads, smem, st>>>(a, *me);
            } else if (ep == 2) {
                TAC_SMEM_OPT_IN(stft_planes_mag_kernel<2>, 111 * 1024);
                stft_planes_mag_kernel<2><<<(unsigned)grid, kThreads, smem, st>>>(a, *me);
            } else {
                TAC_SMEM_OPT_IN(stft_planes_mag_kernel<3>, 111 * 1024);
                stft_planes_mag_kernel<3><<<(unsigned)grid, kThreads, smem, st>>>(a, *me);
            }
            TAC_LAUNCHED(1);
            TAC_CHECK_CUDA(cudaGetLastError());
            return TAC_OK;
        }
#if TAC_PLANES_PERSIST
        // persistent double-buffered form: the tile must be bulk-copyable (16-byte aligned rows, zero or no padding) and there
        // must be enough units to give every SM several
        // Measured on B200 (cfg 4 shapes, ms, one-shot -> persistent): single scale n = 2048 / 1024 / 512 / 256 / 128: 0.452 / 0.444 /
        // 0.450 / 0.471 / 0.537 -> 0.363 / 0.370 / 0.391 / 0.418 / 0.490 (68 ... 52 % of the HBM peak); all five scales in one
        // launch: 1.88 -> 2.04 (ncu: instruction-fetch stalls 1.09 per issue against < 0.2 -- sixteen warps that drift through
        // five unrolled plan bodies, 150 KB of code; in-phase starts, block-cyclic deals and rotated scale orders all measured
        // worse), so several scales keep the one-shot kernel (+ L2 prefetch of the successor's tile).
        if (n_scales <= TAC_PLANES_PERSIST && vec_ok && (a.pmax == 0 || pad_mode == TAC_PAD_ZERO) && (a.pmax & 3) == 0 && grid >= 4ll * kNumSMs) {
            const size_t psmem = 2 * (size_t)a.tile_len * sizeof(float) + (size_t)kPWarps * kXbufFloat2 * sizeof(float2);
            if (psmem <= 226 * 1024) {
                TAC_SMEM_OPT_IN(stft_planes_persist_kernel, 226 * 1024);
                stft_planes_persist_kernel<<<kNumSMs, kPThreads, psmem, st>>>(a, grid); TAC_LAUNCHED(1);
                TAC_CHECK_CUDA(cudaGetLastError());
                return TAC_OK;
            }
        }
#endif
        TAC_SMEM_OPT_IN(stft_planes_kernel, 111 * 1024);
        stft_planes_kernel<<<(unsigned)grid, kThreads, smem, st>>>(a); TAC_LAUNCHED(1);
        TAC_CHECK_CUDA(cudaGetLastError());
        return TAC_OK;
    }

    // time-innermost layouts: one launch per scale
    for (int s = 0; s < n_scales; ++s) {
        TiledArgs a;
        a.x = x; a.T = T; a.n = n_fft[s]; a.hop = hop[s]; a.pad = center ? a.n / 2 : 0;
        a.F = (int)num_frames(T, a.n, a.hop, center);
        a.K = a.n / 2 + 1;
        a.pad_mode = pad_mode; a.vec_ok = vec_ok;
        a.TF = a.n >= 2048 ? 8 : (a.n >= 1024 ? 16 : 32);
        a.ntiles = (a.F + a.TF - 1) / a.TF;
        a.tile_len = (a.TF - 1) * a.hop + a.n;
        a.tile_len = (a.tile_len + 3) & ~3;
        a.f_pad = a.F;
        a.out = reinterpret_cast<float*>(out[s]);
        int rc = get_tables(a.n, win ? win[s] : a.n, normalized, &a.t);
        if (rc) return rc;
        const size_t smem = (size_t)a.tile_len * 4 + ((2 * (size_t)a.K * (a.TF + 1) + 3) & ~(size_t)3) * 4 + (size_t)kWarps * kXbufFloat2 * 8;
        TAC_REQUIRE(rows * a.ntiles < (1ll << 31), "grid too large");
        if (out_mode == TAC_OUT_MAGPHASE) {
            if (f_pad > 0) {
                a.f_pad = f_pad;
                if (f_pad < a.F) {               // crop (pad_tensor with negative padding)
                    a.F = (int)f_pad;
                    a.ntiles = (a.F + a.TF - 1) / a.TF;
                } else if (f_pad > a.F) {
                    TAC_CHECK_CUDA(cudaMemsetAsync(a.out, 0, (size_t)rows * 2 * a.K * (size_t)f_pad * sizeof(float), st));
                }
            }
            bool done = false;
#ifndef TAC_NO_FRONTEND16
            rc = dispatch_frontend(a, rows, st, &done);
            if (rc) return rc;
#endif
            if (!done) rc = dispatch_tiled<TAC_OUT_MAGPHASE>(a, rows, smem, st);
        } else {
            rc = dispatch_tiled<TAC_OUT_COMPLEX>(a, rows, smem, st);
        }
        if (rc) return rc;
    }
    return TAC_OK;
}

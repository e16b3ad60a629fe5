// Tensor-core (tcgen05) Gram matrix of the centred latent batch, 3xTF32 split precision.
//
//   G = Zc^T Zc,  Zc = Z - mean  ([B, d] float32, d <= 128 features zero-padded to 128)
//
// Zc is split as hi + lo with hi = tf32(Zc), lo = tf32(Zc - hi); the kernel accumulates
// hi^T hi + hi^T lo + lo^T hi (the dropped lo^T lo term is ~2^-22 relative) with
// tcgen05.mma.kind::tf32, M = N = 128, K = 8 per instruction, fp32 accumulator in TMEM.
// Both operands are the SAME shared-memory tile (A = B = Zc^T, GEMM-K = batch rows), stored in the
// canonical no-swizzle K-major layout (cute/arch/mma_sm100_desc.hpp, "INTERLEAVE"):
//   element (feature f, row k) at  (k/8)*4096 + (f/8)*SBO + ((k%8)/4)*LBO + (f%8)*16 + (k%4)*4  bytes,
//   SBO = 256 (next group of 8 features), LBO = 128 (next 16-byte K core matrix),
// i.e. a core matrix is 8 features x 4 rows, so staging transposes Z on the fly: a thread gathers 4
// consecutive rows of one feature (4 coalesced scalar loads across the warp) and issues one 16-byte,
// conflict-free shared store.  (The MN-major no-swizzle form -- which would need no transpose --
// was measured on B200 to leave the accumulator untouched for kind::tf32; see DESIGN.md.)
// Centring with the local mean first keeps the fp32 accumulation well conditioned; the caller-
// visible statistic sum z z^T is re-assembled in float64 as G + n * mean mean^T.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "xchg.cuh"

namespace tac {

constexpr int kTcFeat = 128;                 // padded feature count = UMMA M = UMMA N
#ifndef TAC_TC_ROWS
#define TAC_TC_ROWS 64            // 4 CTAs per 256 rows: the statistics sit in front of the collective on the critical path (0.039 -> 0.023 ms at B = 256)
#endif
constexpr int kTcRowsPerCta = TAC_TC_ROWS;   // batch rows (GEMM K) per CTA (multiple of kTcChunk)
constexpr int kTcChunk = 64;                 // rows staged per pipeline stage
constexpr int kTcStages = 2;
constexpr int kTcThreads = 128;
constexpr uint32_t kTcSBO = 256;                                  // 8-feature group stride
constexpr uint32_t kTcLBO = 128;                                  // K core-matrix stride
constexpr uint32_t kTcKGroup = (kTcFeat / 8) * kTcSBO;            // 4096 bytes per 8 rows (one MMA, K = 8)
constexpr int kTcTileBytes = (kTcChunk / 8) * (int)kTcKGroup;     // 32 KiB per (hi | lo) tile
constexpr int kTcSmemBytes = kTcStages * 2 * kTcTileBytes + 1024; // + alignment slack

__device__ __forceinline__ uint32_t tc_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }


__device__ __forceinline__ uint64_t tc_make_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFFu);
    d |= (uint64_t)((kTcLBO >> 4) & 0x3FFFu) << 16;
    d |= (uint64_t)((kTcSBO >> 4) & 0x3FFFu) << 32;
    d |= (uint64_t)1 << 46;                                       // descriptor version 1 (sm_100)
    return d;                                                     // base offset 0, SWIZZLE_NONE
}

// kind::tf32 (a/b format 2), fp32 accumulate (c format 1), A and B both K-major, M = 128, N = 128
constexpr uint32_t kTcIdesc = (1u << 4) | (2u << 7) | (2u << 10) |
                              ((uint32_t)(kTcFeat >> 3) << 17) | ((uint32_t)(kTcFeat >> 4) << 24);

__device__ __forceinline__ void tc_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}\n"
        :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(kTcIdesc), "r"(accumulate) : "memory");
}

__device__ __forceinline__ void tc_mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(tc_smem_u32(bar)), "r"(count) : "memory");
}

// Bounded parity wait: traps instead of hanging the GPU if the barrier never completes.
__device__ __forceinline__ void tc_mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = tc_smem_u32(bar);
    for (uint32_t it = 0; it < (1u << 24); ++it) {
        uint32_t done;
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}\n"
            : "=r"(done) : "r"(addr), "r"(parity) : "memory");
        if (done) return;
    }
    __trap();
}

__device__ __forceinline__ float tc_to_tf32(float v) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
    return __uint_as_float(r);
}

// grid (nsplit, nv); block 128.  partial[v][split][128][128] fp32 = Gram of rows [split*256, +256).
__global__ void __launch_bounds__(kTcThreads, 1)
dpca_gram_tc_kernel(const float* z0, const float* z1, const float* z2,
                    const float* z3, const float* z4, const float* z5, const float* z6, const float* z7,
                    long long B, int d, float* __restrict__ partial, double* __restrict__ lsum, float* __restrict__ lmean) {
    extern __shared__ unsigned char tc_smem_raw[];
    __shared__ __align__(8) uint64_t s_bar[kTcStages];
    __shared__ uint32_t s_tmem;
    __shared__ float s_mean[kTcFeat];

    const int v = blockIdx.y, split = blockIdx.x, nsplit = gridDim.x;
    const float* zs[8] = {z0, z1, z2, z3, z4, z5, z6, z7};
    const float* z = zs[v];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long r_begin = (long long)split * kTcRowsPerCta;
    const long long r_end = (r_begin + kTcRowsPerCta < B) ? r_begin + kTcRowsPerCta : B;

    // 1024-byte aligned operand tiles: [stage][hi|lo][kTcTileBytes]
    const uint32_t raw = tc_smem_u32(tc_smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    unsigned char* tiles = tc_smem_raw + (base - raw);

    // Column sums of THIS CTA's rows (float64) and their mean as the centring shift -- any shift c gives the raw statistic
    // back exactly (sum z z^T = G_c + c s^T + s c^T - n c c^T, re-assembled in float64 by dpca_gram_finish_kernel), so the
    // separate column-sum launch in front of this kernel is gone (the statistics sit in front of the exchange at N > 1).
    if (threadIdx.x < kTcFeat) {
        const int f = threadIdx.x;
        double acc = 0.0;
        if (f < d)
            for (long long r = r_begin; r < r_end; ++r) acc += (double)__ldg(z + r * d + f);
        const float mean = r_end > r_begin ? (float)(acc / (double)(r_end - r_begin)) : 0.0f;
        s_mean[f] = mean;
        lsum[((size_t)v * nsplit + split) * kTcFeat + f] = acc;
        lmean[((size_t)v * nsplit + split) * kTcFeat + f] = mean;
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < kTcStages; ++s) tc_mbar_init(&s_bar[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     :: "r"(tc_smem_u32(&s_tmem)), "r"((uint32_t)kTcFeat) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = s_tmem;

    const int nchunks = (int)((r_end - r_begin + kTcChunk - 1) / kTcChunk);
    uint32_t phase[kTcStages] = {0, 0};
    for (int c = 0; c < nchunks; ++c) {
        const int stage = c % kTcStages;
        if (c >= kTcStages) {                           // the MMAs that read this stage must have retired
            tc_mbar_wait(&s_bar[stage], phase[stage]);
            phase[stage] ^= 1u;
        }
        unsigned char* hi = tiles + (size_t)stage * 2 * kTcTileBytes;
        unsigned char* lo = hi + kTcTileBytes;
        // item = (row quad q, feature f): 4 consecutive rows of one feature -> one 16-byte store per tile.
        // A warp covers 32 consecutive features of one quad: coalesced 128-byte loads, conflict-free stores.
        for (int it = threadIdx.x; it < (kTcChunk / 4) * kTcFeat; it += kTcThreads) {
            const int f = it % kTcFeat, q = it / kTcFeat;
            const long long row0 = r_begin + (long long)c * kTcChunk + 4 * q;
            float a[4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
                a[i] = (f < d && row0 + i < r_end) ? __ldg(z + (row0 + i) * d + f) - s_mean[f] : 0.0f;
            float4 h, l;
            h.x = tc_to_tf32(a[0]); h.y = tc_to_tf32(a[1]); h.z = tc_to_tf32(a[2]); h.w = tc_to_tf32(a[3]);
            l.x = tc_to_tf32(a[0] - h.x); l.y = tc_to_tf32(a[1] - h.y); l.z = tc_to_tf32(a[2] - h.z); l.w = tc_to_tf32(a[3] - h.w);
            const size_t off = (size_t)(q >> 1) * kTcKGroup + (size_t)(f >> 3) * kTcSBO + (size_t)(q & 1) * kTcLBO + (size_t)(f & 7) * 16;
            *reinterpret_cast<float4*>(hi + off) = h;
            *reinterpret_cast<float4*>(lo + off) = l;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy stores -> async proxy
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (threadIdx.x == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t hi_addr = tc_smem_u32(hi), lo_addr = tc_smem_u32(lo);
#pragma unroll
            for (int g = 0; g < kTcChunk / 8; ++g) {
                const uint64_t dh = tc_make_desc(hi_addr + g * kTcKGroup);
                const uint64_t dl = tc_make_desc(lo_addr + g * kTcKGroup);
                tc_mma(tmem, dh, dh, (c > 0 || g > 0) ? 1u : 0u);
                tc_mma(tmem, dh, dl, 1u);
                tc_mma(tmem, dl, dh, 1u);
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                         :: "r"(tc_smem_u32(&s_bar[stage])) : "memory");
        }
    }
    // drain: wait for the last commit of every stage that was used
    for (int s = 0; s < kTcStages; ++s) {
        const int uses = (nchunks - s + kTcStages - 1) / kTcStages;     // chunks that used stage s
        if (s < nchunks && uses > 0) tc_mbar_wait(&s_bar[s], phase[s]);
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    // epilogue: warp w owns TMEM lanes 32w .. 32w+31 (= Gram rows); thread <-> one row, 32 columns at a time
    const int row = 32 * warp + lane;
    float* prow = partial + (((size_t)v * nsplit + split) * kTcFeat + row) * kTcFeat;
#pragma unroll
    for (int c0 = 0; c0 < kTcFeat; c0 += 32) {
        uint32_t r[32];
        const uint32_t taddr = tmem + ((uint32_t)(32 * warp) << 16) + (uint32_t)c0;
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
            "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
            : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
              "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
              "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
              "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
            : "r"(taddr) : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
        for (int j = 0; j < 32; j += 4)
            *reinterpret_cast<float4*>(prow + c0 + j) =
                make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"((uint32_t)kTcFeat) : "memory");
}

// stats: n = B, sum z = sum_c s_c, sum z z^T = sum_c [G_c + m_c s_c^T + s_c m_c^T - n_c m_c m_c^T]   (float64; c = row split,
// m_c = the float32 shift CTA c centred with, s_c its float64 column sums).  grid (d, nv), block 128.
// With an exchange (xd.world > 1) the kernel that completes the statistics also SENDS them: every value is peer-stored
// into this rank's slot on all ranks as it is formed, and the last CTA raises the flags (xchg.cuh) -- compute and
// collective are one kernel, the transfer overlaps the re-assembly row by row.
__global__ void __launch_bounds__(kTcFeat)
dpca_gram_finish_kernel(const float* __restrict__ partial, const double* __restrict__ lsum, const float* __restrict__ lmean,
                        int nsplit, long long B, int d, double* __restrict__ stats, const XchgDev xd) {
    const int v = blockIdx.y, i = blockIdx.x, j = threadIdx.x;
    const size_t stride = 1 + (size_t)d + (size_t)d * d;
    double* st = stats + (size_t)v * stride;
    const bool push = xd.world > 1;
    const unsigned seq = push ? xchg_push_seq(xd) : 0u;
    if (j < d) {
        double acc = 0.0, sj = 0.0;
        for (int s = 0; s < nsplit; ++s) {
            // symmetrise: (i, j) and (j, i) were accumulated by different instruction orders
            const size_t c = (size_t)v * nsplit + s;
            const float* p = partial + c * kTcFeat * kTcFeat;
            const long long r0 = (long long)s * kTcRowsPerCta;
            const double nc = (double)((r0 + kTcRowsPerCta < B ? r0 + kTcRowsPerCta : B) - r0);
            const double mi = (double)lmean[c * kTcFeat + i], mj = (double)lmean[c * kTcFeat + j];
            const double si = lsum[c * kTcFeat + i];
            const double sjc = lsum[c * kTcFeat + j];
            acc += 0.5 * ((double)p[(size_t)i * kTcFeat + j] + (double)p[(size_t)j * kTcFeat + i]) + (mi * sjc + si * mj - nc * mi * mj);
            sj += sjc;
        }
        st[1 + d + (size_t)i * d + j] = acc;
        if (i == 0) {
            st[1 + j] = sj;
            if (j == 0) st[0] = (double)B;
        }
        if (push) {
            xchg_store(xd, seq, (size_t)v * stride + 1 + d + (size_t)i * d + j, acc);
            if (i == 0) {                                   // n and the column sums ride along
                xchg_store(xd, seq, (size_t)v * stride + 1 + j, sj);
                if (j == 0) xchg_store(xd, seq, (size_t)v * stride, (double)B);
            }
        }
    }
    if (push) xchg_push_commit(xd, seq, gridDim.x * gridDim.y);
}

// ------------------------------------------------------------------------------------------ projection on tcgen05
// z_out = (z - mu) P + mu,  P = V_sel V_sel^T  (pkl_file_model_updated.py:503-511: project on the selected components and
// back), both contractions on the tensor pipe in ONE kernel, 3xTF32 split precision as above:
//   GEMM 1  P[128 x 128]    = W W^T,      W[f][t] = V[f][idx[t]]   (K = selected components, zero-padded to a multiple of 8)
//   GEMM 2  D[128 rows x 128] = Zc P,     Zc = z - mu               (K = features, two halves of 64 through one 64 KiB buffer)
// A CTA owns 128 batch rows of one view.  W is gathered into the canonical K-major tile (hi | lo), P comes back from TMEM
// (lane = row n, 128 columns), is split again and written over the W tile as GEMM 2's B operand -- B[n][k] = P[k][n] = P[n][k]
// by symmetry, so the row a thread reads from TMEM is the row it stores.  Replaces dpca_projector_kernel + the 32 x 32 SIMT
// tile of dpca_apply_kernel on the tac_dpca_project path (d <= 128).
constexpr int kPjThreads = 128;
constexpr int kPjTileB = 16 * (int)kTcKGroup;                 // 64 KiB: one 128 x 128 operand (hi or lo)
constexpr int kPjTileA = 8 * (int)kTcKGroup;                  // 32 KiB: 128 rows x 64 features (hi or lo)
constexpr int kPjSmemBytes = 2 * kPjTileB + 2 * kPjTileA + 1024;

__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

struct PjPtrs { const float* z[8]; float* out[8]; };

// grid (ceil(B / 128), nv); block 128.
__global__ void __launch_bounds__(kPjThreads, 1)
dpca_project_tc_kernel(const PjPtrs vp, long long B, int d, const float* __restrict__ vmat, const float* __restrict__ mu,
                       const int* __restrict__ idx, const int* __restrict__ cnt) {
    extern __shared__ unsigned char pj_smem_raw[];
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ uint32_t s_tmem;
    __shared__ float s_mu[kTcFeat];
    __shared__ int s_idx[kTcFeat];

    const int v = blockIdx.y, tid = threadIdx.x, warp = tid >> 5;
    const long long r0 = (long long)blockIdx.x * 128;
    const float* z = vp.z[v];
    float* out = vp.out[v];
    const float* V = vmat + (size_t)v * d * d;
    const int k = min(max(cnt[v], 0), d);

    const uint32_t raw = tc_smem_u32(pj_smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    unsigned char* bh = pj_smem_raw + (base - raw);               // W, then P:  hi | lo
    unsigned char* bl = bh + kPjTileB;
    unsigned char* ah = bl + kPjTileB;                            // Zc half:    hi | lo
    unsigned char* al = ah + kPjTileA;

    s_mu[tid] = tid < d ? mu[(size_t)v * d + tid] : 0.0f;
    s_idx[tid] = tid < k ? idx[(size_t)v * d + tid] : 0;
    if (k == 0) {                                                 // nothing selected in this view: z_out = mu
        __syncthreads();
        const long long row = r0 + tid;
        if (row < B)
            for (int c = 0; c < d; ++c) out[row * d + c] = s_mu[c];
        return;
    }
    if (tid == 0) {
        tc_mbar_init(&s_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     :: "r"(tc_smem_u32(&s_tmem)), "r"(256u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = s_tmem;

    // ---- GEMM 1: W tile (gather of the selected eigenvectors), item = (component quad q, feature f)
    const int ng1 = (k + 7) >> 3;
    for (int it = tid; it < ng1 * 2 * kTcFeat; it += kPjThreads) {
        const int f = it % kTcFeat, q = it / kTcFeat;
        float a[4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
            a[i] = (f < d && 4 * q + i < k) ? __ldg(V + (size_t)f * d + s_idx[4 * q + i]) : 0.0f;
        float4 h, l;
        h.x = tc_to_tf32(a[0]); h.y = tc_to_tf32(a[1]); h.z = tc_to_tf32(a[2]); h.w = tc_to_tf32(a[3]);
        l.x = tc_to_tf32(a[0] - h.x); l.y = tc_to_tf32(a[1] - h.y); l.z = tc_to_tf32(a[2] - h.z); l.w = tc_to_tf32(a[3] - h.w);
        const size_t off = (size_t)(q >> 1) * kTcKGroup + (size_t)(f >> 3) * kTcSBO + (size_t)(q & 1) * kTcLBO + (size_t)(f & 7) * 16;
        *reinterpret_cast<float4*>(bh + off) = h;
        *reinterpret_cast<float4*>(bl + off) = l;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (tid == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t ha = tc_smem_u32(bh), la = tc_smem_u32(bl);
        for (int g = 0; g < ng1; ++g) {
            const uint64_t dh = tc_make_desc(ha + g * kTcKGroup), dl = tc_make_desc(la + g * kTcKGroup);
            tc_mma(tmem, dh, dh, g > 0 ? 1u : 0u);
            tc_mma(tmem, dh, dl, 1u);
            tc_mma(tmem, dl, dh, 1u);
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                     :: "r"(tc_smem_u32(&s_bar)) : "memory");
    }
    tc_mbar_wait(&s_bar, 0u);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    // ---- P: TMEM (lane n = row n) -> split -> B operand tile over the W tile (its MMAs have retired)
    {
        const int n = tid;
#pragma unroll 1
        for (int c0 = 0; c0 < kTcFeat; c0 += 32) {
            uint32_t r[32];
            tc_ld32(tmem + ((uint32_t)(32 * warp) << 16) + (uint32_t)c0, r);
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
                const int kk = c0 + j;
                float a[4] = {__uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3])};
                float4 h, l;
                h.x = tc_to_tf32(a[0]); h.y = tc_to_tf32(a[1]); h.z = tc_to_tf32(a[2]); h.w = tc_to_tf32(a[3]);
                l.x = tc_to_tf32(a[0] - h.x); l.y = tc_to_tf32(a[1] - h.y); l.z = tc_to_tf32(a[2] - h.z); l.w = tc_to_tf32(a[3] - h.w);
                const size_t off = (size_t)(kk >> 3) * kTcKGroup + (size_t)(n >> 3) * kTcSBO + (size_t)((kk & 7) >> 2) * kTcLBO + (size_t)(n & 7) * 16;
                *reinterpret_cast<float4*>(bh + off) = h;
                *reinterpret_cast<float4*>(bl + off) = l;
            }
        }
    }

    // ---- GEMM 2: D = Zc P in two feature halves; thread <-> batch row
    const long long row = r0 + tid;
    const float* zr = z + row * d;
    const bool vec = (d & 3) == 0 && ((reinterpret_cast<uintptr_t>(z) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0;
#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
        if (64 * half >= d) break;                                // uniform: no features in this half
        if (half == 1) {                                          // the MMAs that read the A buffer must have retired
            tc_mbar_wait(&s_bar, 1u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        }
#pragma unroll 4
        for (int kq = 0; kq < 16; ++kq) {
            const int f0 = 64 * half + 4 * kq;
            float a[4];
            if (vec && row < B && f0 + 3 < d) {
                const float4 q = __ldg(reinterpret_cast<const float4*>(zr + f0));
                a[0] = q.x - s_mu[f0]; a[1] = q.y - s_mu[f0 + 1]; a[2] = q.z - s_mu[f0 + 2]; a[3] = q.w - s_mu[f0 + 3];
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) a[i] = (row < B && f0 + i < d) ? __ldg(zr + f0 + i) - s_mu[f0 + i] : 0.0f;
            }
            float4 h, l;
            h.x = tc_to_tf32(a[0]); h.y = tc_to_tf32(a[1]); h.z = tc_to_tf32(a[2]); h.w = tc_to_tf32(a[3]);
            l.x = tc_to_tf32(a[0] - h.x); l.y = tc_to_tf32(a[1] - h.y); l.z = tc_to_tf32(a[2] - h.z); l.w = tc_to_tf32(a[3] - h.w);
            const size_t off = (size_t)(kq >> 1) * kTcKGroup + (size_t)(tid >> 3) * kTcSBO + (size_t)(kq & 1) * kTcLBO + (size_t)(tid & 7) * 16;
            *reinterpret_cast<float4*>(ah + off) = h;
            *reinterpret_cast<float4*>(al + off) = l;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t aha = tc_smem_u32(ah), ala = tc_smem_u32(al), bha = tc_smem_u32(bh), bla = tc_smem_u32(bl);
            for (int g = 0; g < 8; ++g) {
                const int kg = 8 * half + g;
                const uint64_t dah = tc_make_desc(aha + g * kTcKGroup), dal = tc_make_desc(ala + g * kTcKGroup);
                const uint64_t dbh = tc_make_desc(bha + kg * kTcKGroup), dbl = tc_make_desc(bla + kg * kTcKGroup);
                tc_mma(tmem + 128u, dah, dbh, (half > 0 || g > 0) ? 1u : 0u);
                tc_mma(tmem + 128u, dah, dbl, 1u);
                tc_mma(tmem + 128u, dal, dbh, 1u);
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                         :: "r"(tc_smem_u32(&s_bar)) : "memory");
        }
    }
    tc_mbar_wait(&s_bar, d > 64 ? 0u : 1u);                       // third (or, with one half, second) completion of the barrier
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    // ---- epilogue: + mu, one batch row per thread
#pragma unroll 1
    for (int c0 = 0; c0 < kTcFeat; c0 += 32) {
        if (c0 >= d) break;                                       // uniform
        uint32_t r[32];
        tc_ld32(tmem + ((uint32_t)(32 * warp) << 16) + 128u + (uint32_t)c0, r);
        if (row < B) {
            float* o = out + row * d + c0;
            if (vec && c0 + 32 <= d) {
#pragma unroll
                for (int j = 0; j < 32; j += 4)
                    *reinterpret_cast<float4*>(o + j) = make_float4(__uint_as_float(r[j]) + s_mu[c0 + j], __uint_as_float(r[j + 1]) + s_mu[c0 + j + 1],
                                                                     __uint_as_float(r[j + 2]) + s_mu[c0 + j + 2], __uint_as_float(r[j + 3]) + s_mu[c0 + j + 3]);
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (c0 + j < d) o[j] = __uint_as_float(r[j]) + s_mu[c0 + j];
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"(256u) : "memory");
}

inline size_t dpca_tc_workspace_bytes(int nv, long long B, int d) {
    if (d > kTcFeat) return 0;
    const long long nsplit = (B + kTcRowsPerCta - 1) / kTcRowsPerCta;
    // Gram partials [nv][nsplit][128][128] float | column sums [nv][nsplit][128] double | shifts [nv][nsplit][128] float
    return (size_t)nv * (size_t)nsplit * ((size_t)kTcFeat * kTcFeat * sizeof(float) + kTcFeat * (sizeof(double) + sizeof(float)));
}

}  // namespace tac

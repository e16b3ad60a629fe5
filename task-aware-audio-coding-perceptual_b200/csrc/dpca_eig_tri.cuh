// Symmetric eigen-solve of the DPCA covariance for d <= 128, one CTA (1024 threads) per view, float64.
//
//   C = S - s s^T / n  (shared memory, 128 KiB)
//   1. Householder tridiagonalisation  Q^T C Q = T      (reflectors parked in the caller's workspace)
//   2. split T where |b_i| <= 4 eps |T| into unreduced blocks
//   3. eigenvalues by multisection on the Sturm count of each block (8 threads per eigenvalue, product
//      form of the leading-minor recurrence, rescaled every 8 steps -> no division on the critical path)
//   4. eigenvectors of T by the twisted factorisation (Fernando / Parlett-Dhillon "getvec"): forward and
//      backward pivots of T - lambda I, twist at argmin |gamma_k|; one thread per eigenvector;
//      Gram-Schmidt only inside clusters closer than 1e-9 |T| (rare)
//   5. back-transformation V = Q Z, descending order, float32 outputs in data_pca's layout
//
// Replaces the thin SVD of data_pca (pkl_file_model_updated.py:37-42) on the all-reduced statistics:
// singular values are sqrt(eig(Zc^T Zc)); measured 6.7 ms -> ~0.15 ms against the one-sided Jacobi
// kernel it supersedes (which stays for 128 < d <= 256).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "xchg.cuh"

namespace tac {

constexpr int kTriThreads = 1024;
constexpr int kTriMaxD = 128;
constexpr int kTriParts = kTriThreads / kTriMaxD;          // 8 row-groups in the mat-vec / update mapping

struct TriShared {
    double a[kTriMaxD];            // diagonal of T
    double b[kTriMaxD];            // sub-diagonal of T (b[i] couples i and i+1)
    double b2[kTriMaxD];
    double vec_v[kTriMaxD];
    double vec_p[kTriMaxD];      // p, then w = p - K v
    double vec_x[kTriMaxD];      // next column x' (index = row of the current trailing matrix; [0] = 0)
    double part[kTriParts][kTriMaxD];
    double lam[kTriMaxD];
    double red[40];
    double scal[8];                // alpha, beta, K, fro, tn
    int blk_s[kTriMaxD], blk_e[kTriMaxD];
    double beta[kTriMaxD];         // Householder scalars (0: identity)
    int ord[kTriMaxD], inv[kTriMaxD];
    int flag;
};

__device__ __forceinline__ double tri_warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// 1/sqrt(a) and 1/a from the ~20-bit hardware seeds + two Newton steps (float64-accurate to a few ulp)
__device__ __forceinline__ double tri_rsqrt(double a) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a));
    const double h = 0.5 * a;
#pragma unroll
    for (int it = 0; it < 2; ++it) { const double e = fma(-h * y, y, 0.5); y = fma(y, e, y); }
    return y;
}
__device__ __forceinline__ double tri_rcp(double a) {
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a));
#pragma unroll
    for (int it = 0; it < 3; ++it) { const double e = fma(-a, y, 1.0); y = fma(y, e, y); }
    return y;
}

// number of eigenvalues of T[s..e] smaller than x
__device__ __forceinline__ int tri_sturm(const double* __restrict__ a, const double* __restrict__ b2, int s, int e, double x) {
    // p_i = (a_i - x) p_{i-1} - b_{i-1}^2 p_{i-2}; the count is the number of sign changes of p_0 = 1, p_1, ...
    // (an exact zero carries the sign bit it was rounded with; the total over a zero crossing is unchanged)
    double pm = 1.0, p = a[s] - x;
    int cnt = __double2hiint(p) < 0 ? 1 : 0;
    int i = s + 1;
    while (i <= e) {
        // operands of the next 8 steps first (independent of the recurrence), then the dependent chain
        double av[8], bv[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int ii = min(i + u, e);
            av[u] = a[ii] - x; bv[u] = b2[ii - 1];
        }
        const int nst = min(8, e - i + 1);
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            if (u < nst) {
                const double pn = fma(av[u], p, -bv[u] * pm);
                cnt += (unsigned)(__double2hiint(pn) ^ __double2hiint(p)) >> 31;
                pm = p; p = pn;
            }
        }
        i += nst;
        // rescale: keep max(|p|, |pm|) near 1 (exponent arithmetic only)
        const int hp = __double2hiint(p) & 0x7ff00000, hm = __double2hiint(pm) & 0x7ff00000;
        const int he = max(hp, hm);
        if (he != 0) {
            int se = 0x3ff00000 - he;                                   // biased exponent of 2^-ex, before the bias is re-added
            se = max(-(1000 << 20), min(1000 << 20, se));
            const double sc = __hiloint2double(se + 0x3ff00000, 0);
            p *= sc; pm *= sc;
        }
    }
    return cnt;
}

// offset of Householder vector k (length d - k - 1) in the packed reflector store
__host__ __device__ __forceinline__ int tri_refl_off(int k, int d) { return k * (d - 1) - (k * (k - 1)) / 2; }

// phase time stamps (clock64) of view 0, read back by tools through tac_debug_eig_profile()
__device__ long long g_tri_prof[8];
#define TRI_STAMP(i) do { if (blockIdx.x == 0 && threadIdx.x == 0) g_tri_prof[i] = clock64(); } while (0)

// Workspace of the split solve, per view, in doubles (fixed offsets, d <= 128).
constexpr int kTriWsRefl = 0;                    // packed reflectors, d (d - 1) / 2 <= 8128
constexpr int kTriWsA = 8192;                    // a, b, b2, beta: 4 x 128
constexpr int kTriWsLam = kTriWsA + 512;         // eigenvalues of the scaled T, slot order
constexpr int kTriWsScal = kTriWsLam + 128;      // scal[8]
constexpr int kTriWsBlk = kTriWsScal + 8;        // blk_s[128], blk_e[128] (ints)
constexpr int kTriWsVec = 9216;                  // eigenvectors of C, [slot][row] float64
constexpr int kTriWsDoubles = kTriWsVec + kTriMaxD * kTriMaxD;

template <bool SPLIT>
__global__ void __launch_bounds__(kTriThreads, 1)
dpca_eig_tri_kernel(double* stats, int d, float* __restrict__ s_out, float* __restrict__ v_out,
                    float* __restrict__ mu_out, int* __restrict__ rank_out, double* __restrict__ gws, const XchgDev xd) {
    extern __shared__ __align__(16) unsigned char tri_smem_raw[];
    double* A = reinterpret_cast<double*>(tri_smem_raw);            // [d][d]; later Z[row][col] with leading dimension d + 1
    const int ldz = d + 1;                                          // odd stride: a lane-per-row access pattern is conflict free
    double* refl = A + (size_t)d * (d + 1);                         // packed Householder vectors: v_k at tri_refl_off(k, d), length d-k-1
    __shared__ TriShared sh;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int view = blockIdx.x;
    const size_t stride = 1 + d + (size_t)d * d;
    double* st = stats + view * stride;
    if (xd.world > 1) {
        // The pop of the step's exchange happens HERE: this CTA is resident from the start of the step (high-priority
        // stream), waits for the peers' flags itself and adds the world's slots of ITS view in rank order into `stats`,
        // so no reduction kernel has to find an SM between the loss kernel's CTAs and nothing else waits for the peers.
        const unsigned seq = xchg_pop_seq(xd);
        xchg_wait(xd, seq);
        __syncthreads();
        for (size_t e = tid; e < stride; e += kTriThreads) st[e] = xchg_sum(xd, seq, (size_t)view * stride + e);
        __syncthreads();
        if (tid == 0) xchg_pop_commit(xd, seq, gridDim.x);
    }
    double* Lp = gws + (size_t)view * (SPLIT ? kTriWsDoubles : 2 * d * d);   // [i][slot] forward multipliers b_i / D+_i
    double* Um = Lp + (size_t)d * d;                                // [i][slot] backward multipliers b_{i-1} / D-_i
    const double n = st[0];
    const double inv_n = n > 0.0 ? 1.0 / n : 0.0;
    const double eps = 2.220446049250313e-16;

    TRI_STAMP(0);
    // ---- C (symmetric, full storage) and its Frobenius norm; the solve runs on C / |C|_F
    double fro_part = 0.0;
    for (int e = tid; e < d * d; e += kTriThreads) {
        const int i = e / d, j = e % d;
        const int lo = i < j ? i : j, hi = i < j ? j : i;           // read one triangle: exactly symmetric
        const double c = st[1 + d + (size_t)lo * d + hi] - st[1 + lo] * st[1 + hi] * inv_n;
        A[e] = c;
        fro_part = fma(c, c, fro_part);
    }
    for (int i = tid; i < d; i += kTriThreads) mu_out[(size_t)view * d + i] = (float)(st[1 + i] * inv_n);
    fro_part = tri_warp_sum(fro_part);
    if (lane == 0) sh.red[warp] = fro_part;
    __syncthreads();
    {
        double f = 0.0;
#pragma unroll
        for (int w = 0; w < kTriThreads / 32; ++w) f += sh.red[w];
        const double fro = sqrt(f);
        const double isc = fro > 0.0 ? 1.0 / fro : 0.0;
        for (int e = tid; e < d * d; e += kTriThreads) A[e] *= isc;
        if (tid == 0) sh.scal[3] = fro;
    }
    __syncthreads();

    TRI_STAMP(1);
    // ---- 1. Householder tridiagonalisation (column k: x = A[k+1.., k]).  TWO block barriers and ONE pass over the trailing
    // matrix per column: the rank-2 update of column k and the mat-vec q' = A22'' x' of column k + 1 are the same loop (the
    // next column x' is known as soon as v and w are: x'_i = c0_i - (v_i w_0 + w_i v_0), c0 = old first column).  The four
    // warps that own the vectors meet at a 128-thread named barrier between "p.v" and "K, w, x'"; the other 28 warps only
    // see the two block barriers.  (The three-barrier, two-pass form this replaces took 0.29 ms at d = 128.)
    const int ci = tid % kTriMaxD;                 // row / column index owned in the 128 x 8 mapping
    const int cp = tid / kTriMaxD;                 // part 0..7
    if (d >= 3) {                                  // q = A22 x and sigma = x.x of column 0
        const int m = d - 1;
        const double* x = A + 1;
        if (ci < m) {
            const double* col0 = A + (size_t)d + 1 + ci;
            double a0 = 0.0, a1 = 0.0;
            int j = cp;
            for (; j + kTriParts < m; j += 2 * kTriParts) {
                a0 = fma(col0[(size_t)j * d], x[j], a0);
                a1 = fma(col0[(size_t)(j + kTriParts) * d], x[j + kTriParts], a1);
            }
            if (j < m) a0 = fma(col0[(size_t)j * d], x[j], a0);
            sh.part[cp][ci] = a0 + a1;
        }
        if (warp < 4) {
            const double xv = tid < m ? x[tid] : 0.0;
            const double s2 = tri_warp_sum(xv * xv);
            if (lane == 0) sh.red[warp] = s2;
        }
    }
    __syncthreads();
    for (int k = 0; k + 2 < d; ++k) {
        const int m = d - k - 1;
        const double* x = A + (size_t)k * d + (k + 1);              // x_i, i < m (row k == column k by symmetry)
        // P2 (threads < 128): scalars (redundantly, fast reciprocal square root), v, p, p.v | K, w, next column, its sigma
        if (warp < 4) {
            const double sigma = (sh.red[0] + sh.red[1]) + (sh.red[2] + sh.red[3]);
            const double x0 = x[0];
            const double tail = sigma - x0 * x0;
            double alpha = 0.0, beta = 0.0;
            if (sigma > 1e-280 && tail > 1e-30 * sigma) {           // otherwise the column is already tridiagonal: H = I
                const double rs = tri_rsqrt(sigma);
                double nrm = sigma * rs;
                nrm = fma(fma(-nrm, nrm, sigma), 0.5 * rs, nrm);
                alpha = x0 >= 0.0 ? -nrm : nrm;
                beta = tri_rcp(fma(-alpha, x0, sigma));             // 2 / v.v
            }
            double vi = 0.0, pi = 0.0, c0 = 0.0;
            if (tid < m) {
                double qi = 0.0;
#pragma unroll
                for (int p = 0; p < kTriParts; ++p) qi += sh.part[p][tid];
                c0 = A[(size_t)(k + 1) * d + (k + 1) + tid];        // first column of A22 before the update
                vi = x[tid] - (tid == 0 ? alpha : 0.0);
                pi = beta * (qi - alpha * c0);
                refl[tri_refl_off(k, d) + tid] = beta != 0.0 ? vi : 0.0;
            }
            sh.vec_v[tid] = vi; sh.vec_p[tid] = pi;
            const double pv = tri_warp_sum(vi * pi);
            if (lane == 0) sh.red[8 + warp] = pv;
            if (tid == 0) {
                sh.a[k] = A[(size_t)k * d + k];
                sh.b[k] = beta != 0.0 ? alpha : x0;
                sh.beta[k] = beta;
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");
            // w = p - K v, K = beta/2 p.v; next column x'_i (row i >= 1 of A22) with the rounding sequence of the update below
            const double K = 0.5 * beta * ((sh.red[8] + sh.red[9]) + (sh.red[10] + sh.red[11]));
            const double v0 = sh.vec_v[0];
            const double w0 = sh.vec_p[0] - K * v0;
            const double wi = pi - K * vi;
            double xn = 0.0;
            if (tid >= 1 && tid < m) xn = __dsub_rn(c0, __dadd_rn(__dmul_rn(vi, w0), __dmul_rn(wi, v0)));
            asm volatile("bar.sync 1, 128;" ::: "memory");           // every p_0 has been read before p turns into w
            sh.vec_p[tid] = wi;
            sh.vec_x[tid] = xn;
            const double s2 = tri_warp_sum(xn * xn);
            if (lane == 0) sh.red[warp] = s2;                       // sigma of column k + 1
        }
        __syncthreads();
        // P3: A22 -= v w^T + w v^T, and in the same pass the partial q' = A22'' x' of the next column (rows / columns >= 1)
        if (ci < m) {
            const double v_i = sh.vec_v[ci];
            const double w_i = sh.vec_p[ci];
            double* col0 = A + (size_t)(k + 1) * d + (k + 1) + ci;
            double a0 = 0.0, a1 = 0.0;
            int j = cp;
            for (; j + kTriParts < m; j += 2 * kTriParts) {
                const int j1 = j + kTriParts;
                const double va = sh.vec_v[j], vb = sh.vec_v[j1];
                const double wa = sh.vec_p[j], wb = sh.vec_p[j1];
                const double ca = col0[(size_t)j * d], cb = col0[(size_t)j1 * d];
                // exactly symmetric in (i, j): both triangles stay bit-identical.  (An index-ordered two-FMA form, also exactly
                // symmetric, measured SLOWER on B200: 519 k -> 590 k cycles for the tridiagonalisation -- the selects cost more
                // than the two float64 operations they save.)
                const double na = __dsub_rn(ca, __dadd_rn(__dmul_rn(v_i, wa), __dmul_rn(w_i, va)));
                const double nb = __dsub_rn(cb, __dadd_rn(__dmul_rn(v_i, wb), __dmul_rn(w_i, vb)));
                col0[(size_t)j * d] = na;
                col0[(size_t)j1 * d] = nb;
                a0 = fma(na, sh.vec_x[j], a0);                       // vec_x[0] == 0: column 0 is x', not part of A22''
                a1 = fma(nb, sh.vec_x[j1], a1);
            }
            if (j < m) {
                const double va = sh.vec_v[j], wa = sh.vec_p[j];
                const double na = __dsub_rn(col0[(size_t)j * d], __dadd_rn(__dmul_rn(v_i, wa), __dmul_rn(w_i, va)));
                col0[(size_t)j * d] = na;
                a0 = fma(na, sh.vec_x[j], a0);
            }
            if (ci >= 1) sh.part[cp][ci - 1] = a0 + a1;
        }
        __syncthreads();
    }
    TRI_STAMP(2);
    if (tid == 0) {
        if (d >= 2) {
            sh.a[d - 2] = A[(size_t)(d - 2) * d + (d - 2)];
            sh.b[d - 2] = A[(size_t)(d - 2) * d + (d - 1)];
            sh.beta[d - 2] = 0.0;
        }
        sh.a[d - 1] = A[(size_t)(d - 1) * d + (d - 1)];
        sh.b[d - 1] = 0.0;
        sh.beta[d - 1] = 0.0;
    }
    __syncthreads();
    // ---- 2. scale T to unit max-norm, split into unreduced blocks
    if (warp == 0) {
        double tn = 0.0;
        for (int i = lane; i < d; i += 32) tn = fmax(tn, fmax(fabs(sh.a[i]), i + 1 < d ? fabs(sh.b[i]) : 0.0));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) tn = fmax(tn, __shfl_xor_sync(0xffffffffu, tn, o));
        const double it = tn > 0.0 ? 1.0 / tn : 0.0;
        for (int i = lane; i < d; i += 32) {
            const double av = sh.a[i] * it;
            double bv = sh.b[i] * it;
            if (i + 1 >= d || fabs(bv) <= 4.0 * eps) bv = 0.0;      // split point after i
            sh.a[i] = av; sh.b[i] = bv; sh.b2[i] = bv * bv;
        }
        __syncwarp();
        if (lane == 0) {
            sh.scal[5] = tn * sh.scal[3];                            // eigenvalues of C = scal[5] * eigenvalues of the scaled T
            sh.scal[4] = tn > 0.0 ? 1.0 : 0.0;
            int s = 0;
            for (int i = 0; i < d; ++i) {
                if (i == d - 1 || sh.b[i] == 0.0) {
                    for (int j = s; j <= i; ++j) { sh.blk_s[j] = s; sh.blk_e[j] = i; }
                    s = i + 1;
                }
            }
            sh.flag = 0;
        }
    }
    __syncthreads();
    if constexpr (SPLIT) {       // hand T, the block structure and the reflectors to dpca_tri_vectors_kernel
        double* w = Lp;
        const int nrefl = tri_refl_off(d - 1 > 0 ? d - 1 : 0, d);
        for (int i = tid; i < nrefl; i += kTriThreads) w[kTriWsRefl + i] = refl[i];
        if (tid < d) {
            w[kTriWsA + tid] = sh.a[tid];
            w[kTriWsA + kTriMaxD + tid] = sh.b[tid];
            w[kTriWsA + 2 * kTriMaxD + tid] = sh.b2[tid];
            w[kTriWsA + 3 * kTriMaxD + tid] = sh.beta[tid];
            int* blk = reinterpret_cast<int*>(w + kTriWsBlk);
            blk[tid] = sh.blk_s[tid]; blk[kTriMaxD + tid] = sh.blk_e[tid];
        }
        if (tid < 8) w[kTriWsScal + tid] = sh.scal[tid];
        TRI_STAMP(3);
        return;
    }
    const double tn = sh.scal[4];

    TRI_STAMP(3);
    // ---- 3. eigenvalues: multisection on the Sturm count.  8 threads per slot (slot j = (j - s)-th smallest eigenvalue
    // of block [s, e]) evaluate 8 interior points of the bracket per round, so the bracket shrinks 9x per round and all
    // 1024 threads keep the float64 pipe busy (a Sturm count is one dependent chain of d steps).
    {
        const int slot = min(tid >> 3, d - 1), sub = tid & 7;        // spare groups recompute slot d - 1
        const int s = sh.blk_s[slot], e = sh.blk_e[slot], kk = slot - s;
        double gl = 1e300, gu = -1e300;
        for (int i = s; i <= e; ++i) {
            const double r = (i > s ? fabs(sh.b[i - 1]) : 0.0) + (i < e ? fabs(sh.b[i]) : 0.0);
            gl = fmin(gl, sh.a[i] - r); gu = fmax(gu, sh.a[i] + r);
        }
        double lo = gl - 4.0 * eps, hi = gu + 4.0 * eps;
        const unsigned gshift = (unsigned)(lane & ~7);
        for (int it = 0; it < 19; ++it) {                            // 9^19 > 2^60
            const double step = (hi - lo) * (1.0 / 9.0);
            const double x = fma(step, (double)(sub + 1), lo);
            const int c = tri_sturm(sh.a, sh.b2, s, e, x);
            const unsigned below = (__ballot_sync(0xffffffffu, c <= kk) >> gshift) & 0xffu;   // points still left of the eigenvalue
            const int q = __popc(below);                              // counts are monotone in x: the first q points
            const double nlo = q > 0 ? fma(step, (double)q, lo) : lo;
            const double nhi = q < 8 ? fma(step, (double)(q + 1), lo) : hi;
            lo = nlo; hi = nhi;
        }
        if (sub == 0 && (tid >> 3) < d) sh.lam[slot] = 0.5 * (lo + hi);
    }
    // Z[row][slot] = 0 (the tridiagonalised matrix is no longer needed)
    double* Z = A;
    __syncthreads();
    for (int e = tid; e < d * ldz; e += kTriThreads) Z[e] = 0.0;
    __syncthreads();

    TRI_STAMP(4);
    // ---- 4. eigenvectors of T: twisted factorisation, one thread per slot
    if (tid < d) {
        const int slot = tid;
        const int s = sh.blk_s[slot], e = sh.blk_e[slot];
        const double l = sh.lam[slot];
        const double pivmin = 1e-290;
        // backward pivots D-_i (multipliers Um[i] = b_{i-1} / D-_i); the pivots themselves are parked in Lp
        double dm = sh.a[e] - l;
        for (int i = e; i > s; --i) {
            if (fabs(dm) < pivmin) dm = -pivmin;
            const double r = tri_rcp(dm);
            Lp[(size_t)i * d + slot] = dm;
            Um[(size_t)i * d + slot] = sh.b[i - 1] * r;
            dm = fma(-sh.b2[i - 1], r, sh.a[i - 1] - l);
        }
        // dm == D-_s.  forward sweep: D+_i, gamma_i = D+_i + D-_i - (a_i - l)
        double dp = sh.a[s] - l;
        double best = fabs(dp + dm - (sh.a[s] - l));
        int r = s;
        for (int i = s; i < e; ++i) {
            if (fabs(dp) < pivmin) dp = -pivmin;
            const double dmn = Lp[(size_t)(i + 1) * d + slot];       // D-_{i+1}
            const double rc = tri_rcp(dp);
            Lp[(size_t)i * d + slot] = sh.b[i] * rc;                 // forward multiplier (overwrites the parked pivot of row i)
            dp = fma(-sh.b2[i], rc, sh.a[i + 1] - l);
            const double g = fabs(dp + dmn - (sh.a[i + 1] - l));
            if (g < best) { best = g; r = i + 1; }
        }
        // z_r = 1; downwards with the forward multipliers, upwards with the backward ones
        double zz = 1.0, nrm = 1.0;
        Z[(size_t)r * ldz + slot] = 1.0;
        for (int i = r - 1; i >= s; --i) {
            zz = -Lp[(size_t)i * d + slot] * zz;
            Z[(size_t)i * ldz + slot] = zz;
            nrm = fma(zz, zz, nrm);
        }
        zz = 1.0;
        for (int i = r + 1; i <= e; ++i) {
            zz = -Um[(size_t)i * d + slot] * zz;
            Z[(size_t)i * ldz + slot] = zz;
            nrm = fma(zz, zz, nrm);
        }
        const double inv = tri_rsqrt(nrm);
        for (int i = s; i <= e; ++i) Z[(size_t)i * ldz + slot] *= inv;
        // cluster detection (ascending inside a block)
        if (slot > s && l - sh.lam[slot - 1] <= 1e-9 * tn) sh.flag = 1;
    }
    __syncthreads();
    if (sh.flag) {                                                   // rare: Gram-Schmidt inside clusters, warp 0, sequential
        if (warp == 0) {
            for (int k = 1; k < d; ++k) {
                const int s = sh.blk_s[k];
                bool changed = false;
                for (int j = k - 1; j >= s && sh.lam[k] - sh.lam[j] <= 1e-9 * tn * (double)(k - j); --j) {
                    double dot = 0.0;
                    for (int i = lane; i < d; i += 32) dot = fma(Z[(size_t)i * ldz + j], Z[(size_t)i * ldz + k], dot);
                    dot = tri_warp_sum(dot);
                    for (int i = lane; i < d; i += 32) Z[(size_t)i * ldz + k] -= dot * Z[(size_t)i * ldz + j];
                    changed = true;
                    __syncwarp();
                }
                if (changed) {
                    double nn = 0.0;
                    for (int i = lane; i < d; i += 32) nn = fma(Z[(size_t)i * ldz + k], Z[(size_t)i * ldz + k], nn);
                    nn = sqrt(tri_warp_sum(nn));
                    const double sc = nn > 1e-8 ? 1.0 / nn : 0.0;
                    for (int i = lane; i < d; i += 32) Z[(size_t)i * ldz + k] *= sc;
                    __syncwarp();
                }
            }
        }
        __syncthreads();
    }

    TRI_STAMP(5);
    // ---- 5. back-transformation  Z <- H_0 H_1 ... H_{d-3} Z  (reverse order).  Columns are independent: a warp keeps
    // four columns of Z in registers (lane <-> rows lane, lane + 32, ...), reads each packed reflector from shared
    // memory, reduces the four dot products with shuffles and updates in place -- no block barrier inside the loop.
    {
        double z[4][4];
        int colj[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            colj[j] = warp + 32 * j;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = lane + 32 * i;
                z[j][i] = (colj[j] < d && r < d) ? Z[(size_t)r * ldz + colj[j]] : 0.0;
            }
        }
        for (int k = d - 3; k >= 0; --k) {
            const double beta = sh.beta[k];
            if (beta == 0.0) continue;                               // uniform
            const double* vk = refl + tri_refl_off(k, d) - (k + 1);  // vk[r] = component for row r > k
            double v[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = lane + 32 * i;
                v[i] = (r > k && r < d) ? vk[r] : 0.0;
            }
            double t[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                double acc = 0.0;
#pragma unroll
                for (int i = 0; i < 4; ++i) acc = fma(v[i], z[j][i], acc);
                t[j] = acc;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
                for (int j = 0; j < 4; ++j) t[j] += __shfl_xor_sync(0xffffffffu, t[j], o);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const double tb = t[j] * beta;
#pragma unroll
                for (int i = 0; i < 4; ++i) z[j][i] = fma(-v[i], tb, z[j][i]);
            }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = lane + 32 * i;
                if (colj[j] < d && r < d) Z[(size_t)r * ldz + colj[j]] = z[j][i];
            }
    }
    __syncthreads();

    TRI_STAMP(6);
    // ---- outputs: descending order by counting, s = sqrt(lambda), null directions -> zero columns
    const double fro = sh.scal[3];
    const double null_thr = 1e-14 * fro;
    if (tid < d) {
        const double l = sh.lam[tid] * sh.scal[5];
        sh.lam[tid] = (l == l && fabs(l) < 1e300) ? l : 0.0;         // NaN-safe ranks (see dpca_tri_finish_kernel)
    }
    __syncthreads();
    if (tid < d) {
        const double lj = sh.lam[tid];
        int rk = 0;
        for (int i = 0; i < d; ++i) { const double li = sh.lam[i]; rk += (li > lj) || (li == lj && i < tid); }
        sh.ord[tid] = rk;
        sh.inv[rk] = tid;
    }
    __syncthreads();
    const int r_eff = (int)(n < (double)d ? n : (double)d);          // torch.svd(some=True): min(B, d) values
    float* so = s_out + (size_t)view * d;
    float* vo = v_out + (size_t)view * d * d;
    if (tid < d) {
        const int rk = sh.ord[tid];
        const double l = sh.lam[tid];
        so[rk] = (rk < r_eff && l > 0.0) ? (float)sqrt(l) : 0.0f;
    }
    for (int e = tid; e < d * d; e += kTriThreads) {
        const int row = e / d, rk = e % d;                           // coalesced stores: consecutive threads -> consecutive ranks
        // inverse permutation: column whose rank is rk
        const int col = sh.inv[rk];
        vo[e] = (float)(sh.lam[col] > null_thr ? Z[(size_t)row * ldz + col] : 0.0);
    }
    if (tid == 0) rank_out[view] = r_eff;
    TRI_STAMP(7);
}

// ------------------------------------------------------------------------------------------ split solve, stages 2 and 3
// Eigenpairs are independent once T is known, and every stage below is a dependent float64 chain, so one SM is the wrong
// place for 128 of them: dpca_tri_vectors_kernel gives each eigenpair its own WARP, eight warps per CTA, d / 8 CTAs per
// view.  Measured on B200 (d = 128): multisection + twisted factorisation + back-transformation 725 k cycles on one SM
// -> under 100 k cycles spread over 16 SMs.
constexpr int kVecWarps = 8;
constexpr int kVecThreads = kVecWarps * 32;

struct VecShared {
    double a[kTriMaxD], b[kTriMaxD], b2[kTriMaxD], beta[kTriMaxD];
    double scal[8];
    int blk_s[kTriMaxD], blk_e[kTriMaxD];
};

__global__ void __launch_bounds__(kVecThreads, 1)
dpca_tri_vectors_kernel(int d, double* __restrict__ gws) {
    extern __shared__ __align__(16) unsigned char vec_smem_raw[];
    double* refl = reinterpret_cast<double*>(vec_smem_raw);                       // packed reflectors (<= 8128)
    double* scratch = refl + 8192;                                                // per warp: Dp | Dm | Lp | Um, 128 each
    __shared__ VecShared sh;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int view = blockIdx.y;
    double* w = gws + (size_t)view * kTriWsDoubles;
    const int nrefl = tri_refl_off(d - 1 > 0 ? d - 1 : 0, d);
    for (int i = tid; i < nrefl; i += kVecThreads) refl[i] = w[kTriWsRefl + i];
    if (tid < d) {
        sh.a[tid] = w[kTriWsA + tid];
        sh.b[tid] = w[kTriWsA + kTriMaxD + tid];
        sh.b2[tid] = w[kTriWsA + 2 * kTriMaxD + tid];
        sh.beta[tid] = w[kTriWsA + 3 * kTriMaxD + tid];
        const int* blk = reinterpret_cast<const int*>(w + kTriWsBlk);
        sh.blk_s[tid] = blk[tid]; sh.blk_e[tid] = blk[kTriMaxD + tid];
    }
    if (tid < 8) sh.scal[tid] = w[kTriWsScal + tid];
    __syncthreads();
    const int slot = blockIdx.x * kVecWarps + warp;
    if (slot >= d) return;                                   // no block barrier below
    const double eps = 2.220446049250313e-16;
    const int s = sh.blk_s[slot], e = sh.blk_e[slot], kk = slot - s;

    // ---- eigenvalue: multisection, the 32 lanes evaluate 32 interior points of the bracket per round (33^12 > 2^60)
    double gl = 1e300, gu = -1e300;
    for (int i = s + lane; i <= e; i += 32) {
        const double r = (i > s ? fabs(sh.b[i - 1]) : 0.0) + (i < e ? fabs(sh.b[i]) : 0.0);
        gl = fmin(gl, sh.a[i] - r); gu = fmax(gu, sh.a[i] + r);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        gl = fmin(gl, __shfl_xor_sync(0xffffffffu, gl, o));
        gu = fmax(gu, __shfl_xor_sync(0xffffffffu, gu, o));
    }
    double lo = gl - 4.0 * eps, hi = gu + 4.0 * eps;
    for (int it = 0; it < 12; ++it) {
        const double step = (hi - lo) * (1.0 / 33.0);
        const double x = fma(step, (double)(lane + 1), lo);
        const int c = tri_sturm(sh.a, sh.b2, s, e, x);
        const int q = __popc(__ballot_sync(0xffffffffu, c <= kk));    // counts are monotone in x: the first q points
        const double nlo = q > 0 ? fma(step, (double)q, lo) : lo;
        const double nhi = q < 32 ? fma(step, (double)(q + 1), lo) : hi;
        lo = nlo; hi = nhi;
    }
    const double l = 0.5 * (lo + hi);

    // ---- eigenvector of T: twisted factorisation.  Lane 0 runs the backward sweep, lane 1 the forward sweep (they are
    // independent), all lanes pick the twist index, lanes 0 / 1 run the two substitutions.
    double* Dp = scratch + (size_t)warp * 4 * kTriMaxD;
    double* Dm = Dp + kTriMaxD;
    double* Lp = Dm + kTriMaxD;
    double* Um = Lp + kTriMaxD;
    const double pivmin = 1e-290;
    if (lane == 0) {
        double dm = sh.a[e] - l;
        for (int i = e; i > s; --i) {
            if (fabs(dm) < pivmin) dm = -pivmin;
            const double r = tri_rcp(dm);
            Dm[i] = dm;
            Um[i] = sh.b[i - 1] * r;
            dm = fma(-sh.b2[i - 1], r, sh.a[i - 1] - l);
        }
        Dm[s] = dm;
    } else if (lane == 1) {
        double dp = sh.a[s] - l;
        Dp[s] = dp;
        for (int i = s; i < e; ++i) {
            if (fabs(dp) < pivmin) dp = -pivmin;
            const double rc = tri_rcp(dp);
            Lp[i] = sh.b[i] * rc;
            dp = fma(-sh.b2[i], rc, sh.a[i + 1] - l);
            Dp[i + 1] = dp;
        }
    }
    __syncwarp();
    double best = 1e300;
    int r = e;
    for (int i = s + lane; i <= e; i += 32) {               // gamma_i = D+_i + D-_i - (a_i - l); first minimum wins
        const double g = fabs(Dp[i] + Dm[i] - (sh.a[i] - l));
        if (g < best) { best = g; r = i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int orr = __shfl_xor_sync(0xffffffffu, r, o);
        if (ob < best || (ob == best && orr < r)) { best = ob; r = orr; }
    }
    __syncwarp();
    double* zc = Dp;                                         // the column of Z, rows s..e (D+ is no longer needed)
    if (lane == 0) {
        double zz = 1.0;
        zc[r] = 1.0;
        for (int i = r - 1; i >= s; --i) { zz = -Lp[i] * zz; zc[i] = zz; }
    } else if (lane == 1) {
        double zz = 1.0;
        for (int i = r + 1; i <= e; ++i) { zz = -Um[i] * zz; zc[i] = zz; }
    }
    __syncwarp();
    double z[4];
    double nrm = 0.0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int row = lane + 32 * i;
        z[i] = (row >= s && row <= e) ? zc[row] : 0.0;
        nrm = fma(z[i], z[i], nrm);
    }
    nrm = tri_warp_sum(nrm);
    const double inv = tri_rsqrt(nrm);
#pragma unroll
    for (int i = 0; i < 4; ++i) z[i] *= inv;

    // ---- back-transformation  z <- H_0 H_1 ... H_{d-3} z
    for (int k = d - 3; k >= 0; --k) {
        const double beta = sh.beta[k];
        if (beta == 0.0) continue;                           // uniform
        const double* vk = refl + tri_refl_off(k, d) - (k + 1);
        double v[4];
        double acc = 0.0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int row = lane + 32 * i;
            v[i] = (row > k && row < d) ? vk[row] : 0.0;
            acc = fma(v[i], z[i], acc);
        }
        acc = tri_warp_sum(acc);
        const double tb = acc * beta;
#pragma unroll
        for (int i = 0; i < 4; ++i) z[i] = fma(-v[i], tb, z[i]);
    }
    double* vec = w + kTriWsVec + (size_t)slot * kTriMaxD;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int row = lane + 32 * i;
        if (row < d) vec[row] = z[i];
    }
    if (lane == 0) w[kTriWsLam + slot] = l;
}

// Stage 3 (one CTA per view): Gram-Schmidt inside clusters closer than 1e-9 |T| (rare; on the back-transformed vectors,
// which is the same thing because Q is orthogonal), descending order by counting, s = sqrt(lambda), float32 outputs.
__global__ void __launch_bounds__(kTriThreads, 1)
dpca_tri_finish_kernel(const double* __restrict__ stats, int d, float* __restrict__ s_out, float* __restrict__ v_out,
                       int* __restrict__ rank_out, double* __restrict__ gws) {
    __shared__ double lam[kTriMaxD];
    __shared__ int blk_s[kTriMaxD], ord[kTriMaxD], inv[kTriMaxD];
    __shared__ double scal[8];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int view = blockIdx.x;
    double* w = gws + (size_t)view * kTriWsDoubles;
    double* V = w + kTriWsVec;                               // [slot][row]
    const double n = stats[(size_t)view * (1 + d + (size_t)d * d)];
    if (tid < d) {
        // non-finite statistics (NaN / Inf latents) must not reach the rank-by-counting below: every comparison with a NaN
        // is false, two slots would share a rank and inv[] would keep an unwritten entry -> wild column index
        const double l = w[kTriWsLam + tid];
        lam[tid] = (l == l && fabs(l) < 1e300) ? l : 0.0;
        blk_s[tid] = reinterpret_cast<const int*>(w + kTriWsBlk)[tid];
    }
    if (tid < 8) scal[tid] = w[kTriWsScal + tid];
    __syncthreads();
    const double tn = scal[4];
    // Numerically null eigenvalues (rank-deficient batch, B < d: up to d - B + 1 of them in one cluster) are emitted as zero
    // columns below, so their vectors need no orthogonalisation -- without this a 16-row batch at d = 128 spent 2.7 ms in
    // the sequential Gram-Schmidt of a 113-fold "cluster" (cfg 5).  lam is still in units of the scaled T here.
    const double null_scaled = scal[5] > 0.0 ? 1e-14 * scal[3] / scal[5] : 0.0;
    const int clustered = __syncthreads_or(tid < d && tid > blk_s[tid] && lam[tid] > null_scaled && lam[tid] - lam[tid - 1] <= 1e-9 * tn);
    if (clustered) {
        if (warp == 0) {
            for (int k = 1; k < d; ++k) {
                const int s = blk_s[k];
                bool changed = false;
                if (!(lam[k] > null_scaled)) continue;                   // null direction: zero column (warp-uniform)
                for (int j = k - 1; j >= s && lam[j] > null_scaled && lam[k] - lam[j] <= 1e-9 * tn * (double)(k - j); --j) {
                    double dot = 0.0;
                    for (int i = lane; i < d; i += 32) dot = fma(V[(size_t)j * kTriMaxD + i], V[(size_t)k * kTriMaxD + i], dot);
                    dot = tri_warp_sum(dot);
                    for (int i = lane; i < d; i += 32) V[(size_t)k * kTriMaxD + i] -= dot * V[(size_t)j * kTriMaxD + i];
                    changed = true;
                    __syncwarp();
                }
                if (changed) {
                    double nn = 0.0;
                    for (int i = lane; i < d; i += 32) nn = fma(V[(size_t)k * kTriMaxD + i], V[(size_t)k * kTriMaxD + i], nn);
                    nn = sqrt(tri_warp_sum(nn));
                    const double sc = nn > 1e-8 ? 1.0 / nn : 0.0;
                    for (int i = lane; i < d; i += 32) V[(size_t)k * kTriMaxD + i] *= sc;
                    __syncwarp();
                }
            }
        }
        __syncthreads();
    }
    const double fro = scal[3];
    const double null_thr = 1e-14 * fro;
    if (tid < d) {
        const double l = lam[tid] * scal[5];
        lam[tid] = (l == l && fabs(l) < 1e300) ? l : 0.0;
    }
    __syncthreads();
    if (tid < d) {
        const double lj = lam[tid];
        int rk = 0;
        for (int i = 0; i < d; ++i) { const double li = lam[i]; rk += (li > lj) || (li == lj && i < tid); }
        ord[tid] = rk;
        inv[rk] = tid;
    }
    __syncthreads();
    const int r_eff = (int)(n < (double)d ? n : (double)d);          // torch.svd(some=True): min(B, d) values
    float* so = s_out + (size_t)view * d;
    float* vo = v_out + (size_t)view * d * d;
    if (tid < d) {
        const int rk = ord[tid];
        const double l = lam[tid];
        so[rk] = (rk < r_eff && l > 0.0) ? (float)sqrt(l) : 0.0f;
    }
    for (int e = tid; e < d * d; e += kTriThreads) {
        const int rk = e / d, row = e % d;                           // coalesced reads of one eigenvector per d threads
        const int col = inv[rk];
        vo[(size_t)row * d + rk] = (float)(lam[col] > null_thr ? V[(size_t)col * kTriMaxD + row] : 0.0);
    }
    if (tid == 0) rank_out[view] = r_eff;
}

inline size_t dpca_tri_vectors_smem_bytes() { return (size_t)(8192 + kVecWarps * 4 * kTriMaxD) * sizeof(double); }

#ifndef TAC_EIG_SPLIT
#define TAC_EIG_SPLIT 1        // 1: tridiagonalise (1 CTA / view) -> eigenpairs on d/8 CTAs per view -> order + outputs;  0: one CTA per view
#endif
inline size_t dpca_eig_tri_workspace_bytes(int nv, int d) {
    const size_t per = TAC_EIG_SPLIT ? (size_t)kTriWsDoubles : (size_t)2 * d * d;
    return (size_t)nv * per * sizeof(double);
}
inline size_t dpca_eig_tri_smem_bytes(int d) { return ((size_t)d * (d + 1) + (size_t)tri_refl_off(d - 1 > 0 ? d - 1 : 0, d) + d) * sizeof(double); }

}  // namespace tac

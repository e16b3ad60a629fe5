// K3 -- DPCA: sufficient statistics (Gram), symmetric eigen-decomposition, global top-k selection
// across views and projection.  Everything stays on the device: no host synchronisation between
// the statistics (all-reducible across ranks), the eigen-solve, the selection and the projection.
//
//   stats  [nv][1 + d + d*d] float64 : n, sum z, sum z z^T                (tac_dpca_stats)
//   eig    one CTA per view, parallel one-sided Jacobi in float64 on C = S - s s^T / n, columns in
//          shared memory (d <= 128) or in the caller's workspace (d <= 256)   (tac_dpca_eig)
//   select rank-by-counting over the concatenated singular values          (tac_dpca_select)
//   project z_out = (z - mu) (V_sel V_sel^T) + mu                           (tac_dpca_project)
//
// The Gram matrix has two implementations: TAC_GRAM_FP64 (SIMT, exact products, float64 sums) and
// TAC_GRAM_TF32X3 (tcgen05 tensor cores, see dpca_tc.cuh).
#include "tac_common.cuh"
#include "dpca_tc.cuh"
#include "dpca_eig_tri.cuh"

namespace tac {

constexpr int kMaxViews = 8;
constexpr int kMaxD = 256;

struct ViewPtrs { const float* z[kMaxViews]; float* out[kMaxViews]; };

// ------------------------------------------------------------------------------------------ stats (float64 SIMT)
// grid (tiles_i * tiles_j, nv); CTA computes a 32x32 tile of sum_b z z^T over all rows.
__global__ void __launch_bounds__(256) dpca_gram_fp64_kernel(const ViewPtrs vp, long long B, int d, double* __restrict__ stats) {
    __shared__ float si[32][33], sj[32][33];
    const int v = blockIdx.y;
    const int nt = (d + 31) / 32;
    const int ti = blockIdx.x / nt, tj = blockIdx.x % nt;
    if (tj < ti) return;                                   // symmetric: upper tiles only, mirrored on store
    const float* z = vp.z[v];
    double* st = stats + (size_t)v * (1 + d + (size_t)d * d);
    const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;       // 16 x 16 threads, 2 x 2 outputs each
    double acc[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
    double csum = 0.0;                                      // column sums (tiles with ti == tj only)
    for (long long r0 = 0; r0 < B; r0 += 32) {
        for (int e = threadIdx.x; e < 32 * 32; e += 256) {
            const int r = e / 32, c = e % 32;
            const long long row = r0 + r;
            const int ci = ti * 32 + c, cj = tj * 32 + c;
            si[r][c] = (row < B && ci < d) ? __ldg(z + row * d + ci) : 0.0f;
            sj[r][c] = (row < B && cj < d) ? __ldg(z + row * d + cj) : 0.0f;
        }
        __syncthreads();
#pragma unroll 8
        for (int r = 0; r < 32; ++r) {
            const double a0 = si[r][ty * 2], a1 = si[r][ty * 2 + 1];
            const double b0 = sj[r][tx * 2], b1 = sj[r][tx * 2 + 1];
            acc[0][0] = fma(a0, b0, acc[0][0]); acc[0][1] = fma(a0, b1, acc[0][1]);
            acc[1][0] = fma(a1, b0, acc[1][0]); acc[1][1] = fma(a1, b1, acc[1][1]);
        }
        if (ti == tj && threadIdx.x < 32) {
            for (int r = 0; r < 32; ++r) csum += (double)si[r][threadIdx.x];
        }
        __syncthreads();
    }
    double* S = st + 1 + d;
    for (int a = 0; a < 2; ++a)
        for (int b = 0; b < 2; ++b) {
            const int i = ti * 32 + ty * 2 + a, j = tj * 32 + tx * 2 + b;
            if (i < d && j < d) { S[(size_t)i * d + j] = acc[a][b]; S[(size_t)j * d + i] = acc[a][b]; }
        }
    if (ti == tj && threadIdx.x < 32) {
        const int i = ti * 32 + threadIdx.x;
        if (i < d) st[1 + i] = csum;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) st[0] = (double)B;
}

// ------------------------------------------------------------------------------------------ eigen-solve
// One-sided (Hestenes) Jacobi on the symmetric PSD matrix C held column-wise in A (ld = d):
// rotate column pairs until all columns are mutually orthogonal; then column j equals
// lambda_j * v_j, i.e. lambda_j = |a_j| and v_j = a_j / |a_j|.
constexpr int kEigThreads = 1024;
constexpr int kLanesPerPair = 16;

__global__ void __launch_bounds__(kEigThreads, 1) dpca_eig_kernel(const double* __restrict__ stats, int d, float* __restrict__ s_out,
                                                                   float* __restrict__ v_out, float* __restrict__ mu_out,
                                                                   int* __restrict__ rank_out, double* __restrict__ gws, int use_gws) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ int s_rot;
    __shared__ double s_lam[kMaxD];
    __shared__ int s_ord[kMaxD];
    const int v = blockIdx.x;
    const size_t stride = 1 + d + (size_t)d * d;
    const double* st = stats + v * stride;
    double* A = use_gws ? gws + (size_t)v * d * d : reinterpret_cast<double*>(smem_raw);
    const double n = st[0];
    const double inv_n = n > 0.0 ? 1.0 / n : 0.0;
    // C = S - s s^T / n   (column-major == row-major: symmetric)
    for (int e = threadIdx.x; e < d * d; e += blockDim.x) {
        const int i = e / d, j = e % d;
        A[e] = st[1 + d + e] - st[1 + i] * st[1 + j] * inv_n;
    }
    for (int i = threadIdx.x; i < d; i += blockDim.x) mu_out[(size_t)v * d + i] = (float)(st[1 + i] * inv_n);
    __syncthreads();

    const int dd = (d + 1) & ~1;                 // even player count (a dummy column sits out)
    const int npairs = dd / 2;
    const int sub = threadIdx.x % kLanesPerPair;
    const int slot = threadIdx.x / kLanesPerPair;
    const int nslots = blockDim.x / kLanesPerPair;
    const double tol = 1e-13;
    const unsigned hmask = 0xFFFFu << (16 * ((threadIdx.x % 32) / kLanesPerPair));
    // columns whose norm falls below 1e-14 * |C|_F are numerically null directions: never rotate them
    __shared__ double s_fro;
    if (threadIdx.x < 32) {
        double acc = 0.0;
        for (int e = threadIdx.x; e < d * d; e += 32) acc = fma(A[e], A[e], acc);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (threadIdx.x == 0) s_fro = acc;
    }
    __syncthreads();
    const double null2 = 1e-28 * s_fro;
    for (int sweep = 0; sweep < 30; ++sweep) {
        if (threadIdx.x == 0) s_rot = 0;
        __syncthreads();
        for (int step = 0; step < dd - 1; ++step) {
            for (int pi = slot; pi < npairs; pi += nslots) {
                int p, q;
                if (pi == 0) { p = dd - 1; q = step % (dd - 1); }
                else { p = (step + pi) % (dd - 1); q = (step - pi + (dd - 1)) % (dd - 1); }
                if (p < d && q < d) {                      // uniform across the 16 lanes of the pair
                    double* ap = A + (size_t)p * d;
                    double* aq = A + (size_t)q * d;
                    double al = 0.0, be = 0.0, ga = 0.0;
                    for (int r = sub; r < d; r += kLanesPerPair) {
                        const double x = ap[r], y = aq[r];
                        al = fma(x, x, al); be = fma(y, y, be); ga = fma(x, y, ga);
                    }
#pragma unroll
                    for (int o = kLanesPerPair / 2; o > 0; o >>= 1) {
                        al += __shfl_xor_sync(hmask, al, o, kLanesPerPair);
                        be += __shfl_xor_sync(hmask, be, o, kLanesPerPair);
                        ga += __shfl_xor_sync(hmask, ga, o, kLanesPerPair);
                    }
                    if (fabs(ga) > tol * sqrt(al * be) && ga != 0.0 && al > null2 && be > null2) {
                        const double zeta = (be - al) / (2.0 * ga);
                        const double t = (zeta >= 0.0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                        const double c = 1.0 / sqrt(1.0 + t * t), sn = c * t;
                        for (int r = sub; r < d; r += kLanesPerPair) {
                            const double x = ap[r], y = aq[r];
                            ap[r] = c * x - sn * y;
                            aq[r] = sn * x + c * y;
                        }
                        if (sub == 0) s_rot = 1;
                    }
                }
            }
            __syncthreads();
        }
        const int any = s_rot;
        __syncthreads();
        if (!any) break;
    }
    // eigenvalues = column norms; order descending by counting
    for (int j = threadIdx.x / 32; j < d; j += blockDim.x / 32) {
        const int lane = threadIdx.x % 32;
        double acc = 0.0;
        for (int r = lane; r < d; r += 32) { const double x = A[(size_t)j * d + r]; acc = fma(x, x, acc); }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) s_lam[j] = sqrt(acc);
    }
    __syncthreads();
    for (int j = threadIdx.x; j < d; j += blockDim.x) {
        const double lj = s_lam[j];
        int rk = 0;
        for (int i = 0; i < d; ++i) { const double li = s_lam[i]; rk += (li > lj) || (li == lj && i < j); }
        s_ord[j] = rk;
    }
    __syncthreads();
    const int r_eff = (int)(n < (double)d ? n : (double)d);       // torch.svd(some=True): min(B, d) values
    float* so = s_out + (size_t)v * d;
    float* vo = v_out + (size_t)v * d * d;
    for (int j = threadIdx.x; j < d; j += blockDim.x) {
        const int rk = s_ord[j];
        so[rk] = rk < r_eff ? (float)sqrt(s_lam[j]) : 0.0f;
    }
    for (int e = threadIdx.x; e < d * d; e += blockDim.x) {
        const int j = e / d, r = e % d;                             // column j of A, row r
        const double lam = s_lam[j];
        // numerically null columns carry no direction (normalising their rounding noise would give a
        // vector that is NOT orthogonal to the range): emit a zero column, which projects to nothing --
        // exactly what the reference's arbitrary null-space vector contributes ((z - mu) v = 0)
        const double val = lam * lam > null2 ? A[e] / lam : 0.0;
        vo[(size_t)r * d + s_ord[j]] = (float)val;                  // v[r][rank] : eigenvector in column `rank`
    }
    if (threadIdx.x == 0) rank_out[v] = r_eff;
}

// ------------------------------------------------------------------------------------------ selection
// Global top-dim_p of the concatenated singular values (first rank[v] of every view), split per view.
__global__ void __launch_bounds__(1024) dpca_select_kernel(const float* __restrict__ s, const int* __restrict__ rank, int nv, int d,
                                                          int dim_p, int* __restrict__ idx, int* __restrict__ cnt) {
    __shared__ float sv[kMaxViews * kMaxD];
    __shared__ int sview[kMaxViews * kMaxD], sloc[kMaxViews * kMaxD], sgr[kMaxViews * kMaxD];
    __shared__ int s_n;
    // concatenated list of the views' first rank[v] singular values, built by all threads (one thread doing it serially
    // cost ~14 us of dependent global loads on the latency-bound DPCA chain)
    __shared__ int s_off[kMaxViews + 1];
    if (threadIdx.x == 0) {
        int n = 0;
        for (int v = 0; v < nv; ++v) { s_off[v] = n; n += max(0, min(rank[v], d)); }
        s_off[nv] = n;
        s_n = n;
    }
    for (int v = threadIdx.x; v < nv; v += blockDim.x) cnt[v] = 0;
    __syncthreads();
    for (int e = threadIdx.x; e < nv * d; e += blockDim.x) {
        const int v = e / d, i = e - v * d;
        if (i < s_off[v + 1] - s_off[v]) {
            const int n = s_off[v] + i;
            sv[n] = s[(size_t)v * d + i]; sview[n] = v; sloc[n] = i;
        }
    }
    __syncthreads();
    const int n = s_n;
    for (int e = threadIdx.x; e < n; e += blockDim.x) {
        const float me = sv[e];
        int g = 0;
        for (int i = 0; i < n; ++i) { const float o = sv[i]; g += (o > me) || (o == me && i < e); }
        sgr[e] = g;                                     // global descending rank (stable)
    }
    __syncthreads();
    for (int e = threadIdx.x; e < n; e += blockDim.x) {
        if (sgr[e] < dim_p) {
            int pos = 0;                                 // order inside the view follows the global order
            for (int i = 0; i < n; ++i) pos += (sview[i] == sview[e]) && (sgr[i] < sgr[e]);
            idx[(size_t)sview[e] * d + pos] = sloc[e];
            atomicAdd(cnt + sview[e], 1);
        }
    }
}

// ------------------------------------------------------------------------------------------ projection
// P_v = V_sel V_sel^T  (d x d, float32).  grid (d, nv), block d threads (<= 256).
__global__ void dpca_projector_kernel(const float* __restrict__ vmat, const int* __restrict__ idx, const int* __restrict__ cnt,
                                      int d, float* __restrict__ P) {
    __shared__ float rowi[kMaxD];
    __shared__ int sidx[kMaxD];
    const int v = blockIdx.y, i = blockIdx.x, j = threadIdx.x;
    const float* V = vmat + (size_t)v * d * d;
    const int k = cnt[v];
    for (int t = threadIdx.x; t < k; t += blockDim.x) {
        const int c = idx[(size_t)v * d + t];
        sidx[t] = c;
        rowi[t] = V[(size_t)i * d + c];
    }
    __syncthreads();
    if (j < d) {
        float acc = 0.0f;
        for (int t = 0; t < k; ++t) acc = fmaf(rowi[t], __ldg(V + (size_t)j * d + sidx[t]), acc);
        P[((size_t)v * d + i) * d + j] = acc;
    }
}

// z_out = (z - mu) P + mu.  grid (ceil(B/32), ceil(d/32), nv), 256 threads, 32x32 output tile.
__global__ void __launch_bounds__(256) dpca_apply_kernel(const ViewPtrs vp, long long B, int d, const float* __restrict__ P,
                                                        const float* __restrict__ mu, int add_mean) {
    __shared__ float sz[32][33], sp[32][33];
    const int v = blockIdx.z;
    const float* z = vp.z[v];
    float* out = vp.out[v];
    const float* Pv = P + (size_t)v * d * d;
    const float* m = mu + (size_t)v * d;
    const long long r0 = (long long)blockIdx.x * 32;
    const int c0 = blockIdx.y * 32;
    const int tx = threadIdx.x % 32, ty = threadIdx.x / 32;          // 32 x 8, 4 rows per thread
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int k0 = 0; k0 < d; k0 += 32) {
        for (int e = threadIdx.x; e < 1024; e += 256) {
            const int r = e / 32, c = e % 32;
            const long long row = r0 + r;
            const int kk = k0 + c;
            sz[r][c] = (row < B && kk < d) ? __ldg(z + row * d + kk) - __ldg(m + kk) : 0.0f;
            const int pr = k0 + r, pc = c0 + c;
            sp[r][c] = (pr < d && pc < d) ? __ldg(Pv + (size_t)pr * d + pc) : 0.0f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 32; ++kk) {
            const float pv = sp[kk][tx];
#pragma unroll
            for (int a = 0; a < 4; ++a) acc[a] = fmaf(sz[ty * 4 + a][kk], pv, acc[a]);
        }
        __syncthreads();
    }
    const int col = c0 + tx;
    if (col < d) {
        const float mc = add_mean ? __ldg(m + col) : 0.0f;
        for (int a = 0; a < 4; ++a) {
            const long long row = r0 + ty * 4 + a;
            if (row < B) out[row * d + col] = acc[a] + mc;
        }
    }
}

// ------------------------------------------------------------------------------------------ variance top-k
__global__ void __launch_bounds__(256) var_topk_kernel(const double* __restrict__ stats, int d, int k, float* __restrict__ var,
                                                      int* __restrict__ top_idx) {
    __shared__ double sv[kMaxD * 4];
    const double n = stats[0];
    for (int j = threadIdx.x; j < d; j += blockDim.x) {
        const double s = stats[1 + j];
        const double q = stats[1 + d + (size_t)j * d + j];
        const double vv = n > 1.0 ? (q - s * s / n) / (n - 1.0) : 0.0;     // unbiased, torch.var default
        sv[j] = vv;
        var[j] = (float)vv;
    }
    __syncthreads();
    for (int j = threadIdx.x; j < d; j += blockDim.x) {
        const double me = sv[j];
        int rk = 0;
        for (int i = 0; i < d; ++i) rk += (sv[i] > me) || (sv[i] == me && i < j);
        if (rk < k) top_idx[rk] = j;
    }
}

// ------------------------------------------------------------------------------------------ nuclear-norm regulariser
// pkl_file_models.py:943-947 / pkl_file_model_updated.py:526-532:  z -= mean; z /= |z|_F; nuc = |z|_* / B.
// With the singular values s of the centred batch (tac_dpca_eig): |z|_F = sqrt(sum s^2), |z|_* = sum s.
// out[v] = sum s / (|z|_F * n);  aux[v] = {sum s, |z|_F, n}.  grid (nv), block 256.
__global__ void __launch_bounds__(256) nuclear_kernel(const float* __restrict__ s, const int* __restrict__ rank, const double* __restrict__ stats,
                                                     int d, float* __restrict__ out, double* __restrict__ aux) {
    __shared__ double sh[2][256];
    const int v = blockIdx.x;
    const int r = min(rank[v], d);
    double a = 0.0, q = 0.0;
    for (int i = threadIdx.x; i < r; i += 256) { const double x = (double)s[(size_t)v * d + i]; a += x; q = fma(x, x, q); }
    sh[0][threadIdx.x] = a; sh[1][threadIdx.x] = q;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) { sh[0][threadIdx.x] += sh[0][threadIdx.x + o]; sh[1][threadIdx.x] += sh[1][threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double n = stats[(size_t)v * (1 + d + (size_t)d * d)];
        const double fro = sqrt(sh[1][0]);
        out[v] = (fro > 0.0 && n > 0.0) ? (float)(sh[0][0] / (fro * n)) : 0.0f;
        if (aux) { aux[3 * v] = sh[0][0]; aux[3 * v + 1] = fro; aux[3 * v + 2] = n; }
    }
}

// Gm_v = (V diag(1 / s_i) V^T - (sum s / fro^2) I) / (fro * n):  d nuc / d z = (z - mean) Gm  (null directions carry the zero
// sub-gradient).  grid (d, nv), block d rounded up to a warp.
__global__ void nuclear_grad_matrix_kernel(const float* __restrict__ s, const float* __restrict__ vmat, const int* __restrict__ rank,
                                           const double* __restrict__ aux, int d, float* __restrict__ gm) {
    __shared__ float rowi[kMaxD];
    const int v = blockIdx.y, i = blockIdx.x, j = threadIdx.x;
    const float* V = vmat + (size_t)v * d * d;
    const int r = min(rank[v], d);
    const double sum_s = aux[3 * v], fro = aux[3 * v + 1], n = aux[3 * v + 2];
    const float smax = s[(size_t)v * d];
    for (int t = threadIdx.x; t < r; t += blockDim.x) {
        const float sv = s[(size_t)v * d + t];
        rowi[t] = sv > 1e-6f * smax ? V[(size_t)i * d + t] / sv : 0.0f;
    }
    __syncthreads();
    if (j < d) {
        float acc = 0.0f;
        for (int t = 0; t < r; ++t) acc = fmaf(rowi[t], __ldg(V + (size_t)j * d + t), acc);
        double g = (double)acc - (i == j ? sum_s / (fro * fro) : 0.0);
        gm[((size_t)v * d + i) * d + j] = (fro > 0.0 && n > 0.0) ? (float)(g / (fro * n)) : 0.0f;
    }
}

static size_t eig_smem_bytes(int d) { return (size_t)d * d * sizeof(double); }

static int dpca_tc_stats(const float* const* z, int nv, long long B, int d, double* stats, void* ws, size_t ws_bytes,
                         const XchgDev& xd, cudaStream_t st) {
    TAC_REQUIRE(d <= kTcFeat, "TAC_GRAM_TF32X3 supports d <= %d (got %d); use TAC_GRAM_FP64", kTcFeat, d);
    const size_t need = dpca_tc_workspace_bytes(nv, B, d);
    if (!ws || ws_bytes < need) return set_error(TAC_EWORKSPACE, "tac_dpca_stats needs %zu workspace bytes", need);
    TAC_ALIGNED16(ws, "workspace");
    const float* p[8];
    for (int i = 0; i < 8; ++i) p[i] = z[i < nv ? i : 0];
    const int nsplit = (int)((B + kTcRowsPerCta - 1) / kTcRowsPerCta);
    float* partial = reinterpret_cast<float*>(ws);
    double* lsum = reinterpret_cast<double*>(partial + (size_t)nv * nsplit * kTcFeat * kTcFeat);       // 8-byte aligned: 64 KiB multiples
    float* lmean = reinterpret_cast<float*>(lsum + (size_t)nv * nsplit * kTcFeat);
    TAC_CHECK_CUDA(cudaFuncSetAttribute(dpca_gram_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes));
    dpca_gram_tc_kernel<<<dim3(nsplit, nv), kTcThreads, kTcSmemBytes, st>>>(p[0], p[1], p[2], p[3], p[4], p[5], p[6], p[7], B, d,
                                                                           partial, lsum, lmean); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    dpca_gram_finish_kernel<<<dim3(d, nv), kTcFeat, 0, st>>>(partial, lsum, lmean, nsplit, B, d, stats, xd); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

}  // namespace tac

using namespace tac;

extern "C" size_t tac_dpca_workspace_bytes(int nv, int64_t B, int d) {
    if (nv <= 0 || d <= 0 || B <= 0) return 0;
    // eig (global-memory columns for d > 128) | projector P | tensor-core Gram scratch
    size_t eig = d <= kTriMaxD ? dpca_eig_tri_workspace_bytes(nv, d) : (size_t)nv * d * d * sizeof(double);
    size_t proj = (size_t)nv * d * d * sizeof(float);
    size_t tc = dpca_tc_workspace_bytes(nv, B, d);
    size_t m = eig > proj ? eig : proj;
    m = m > tc ? m : tc;
    return (m + 255) & ~(size_t)255;
}

static int check_stats_xchg(const tac_xchg* x, int nv, int d) {
    if (!x || x->world <= 1) return TAC_OK;
    TAC_REQUIRE(x->world <= kXchgMaxWorld && x->rank >= 0 && x->rank < x->world, "bad exchange world / rank");
    TAC_REQUIRE(x->payload == (int64_t)nv * (1 + d + (int64_t)d * d), "exchange payload %lld != nv * (1 + d + d*d) = %lld",
                (long long)x->payload, (long long)nv * (1 + d + (long long)d * d));
    for (int r = 0; r < x->world; ++r) TAC_REQUIRE(x->peer[r] != nullptr, "peer[%d] is null", r);
    return TAC_OK;
}

extern "C" int tac_dpca_stats(const float* const* z, int nv, int64_t B, int d, int algo, double* stats, void* workspace,
                              size_t workspace_bytes, tac_stream_t stream) {
    return tac_dpca_stats_push(z, nv, B, d, algo, stats, workspace, workspace_bytes, nullptr, stream);
}

extern "C" int tac_dpca_stats_push(const float* const* z, int nv, int64_t B, int d, int algo, double* stats, void* workspace,
                                   size_t workspace_bytes, const tac_xchg* x, tac_stream_t stream) {
    TAC_REQUIRE(z && stats, "null argument");
    {
        int rc = check_stats_xchg(x, nv, d);
        if (rc) return rc;
    }
    const XchgDev xd = xchg_dev(x);
    TAC_REQUIRE(nv >= 1 && nv <= kMaxViews, "nv must be in [1, %d]", kMaxViews);
    TAC_REQUIRE(B >= 1 && d >= 1 && d <= kMaxD, "need B >= 1 and 1 <= d <= %d", kMaxD);
    TAC_REQUIRE(algo == TAC_GRAM_TF32X3 || algo == TAC_GRAM_FP64, "unknown algo %d", algo);
    ViewPtrs vp;
    for (int v = 0; v < nv; ++v) {
        TAC_REQUIRE(z[v] != nullptr, "z[%d] is null", v);
        vp.z[v] = z[v]; vp.out[v] = nullptr;
    }
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (algo == TAC_GRAM_TF32X3) {
        return dpca_tc_stats(vp.z, nv, B, d, stats, workspace, workspace_bytes, xd, st);
    }
    const int nt = (d + 31) / 32;
    dpca_gram_fp64_kernel<<<dim3(nt * nt, nv), 256, 0, st>>>(vp, B, d, stats); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    if (xd.world > 1) return tac_xchg_push(x, stats, stream);       // float64 SIMT Gram: the generic push kernel follows
    return TAC_OK;
}

extern "C" int tac_dpca_eig(const double* stats, int nv, int d, float* s, float* v, float* mu, int* rank, void* workspace,
                            size_t workspace_bytes, tac_stream_t stream) {
    // without an exchange the statistics are only read
    return tac_dpca_eig_pop(const_cast<double*>(stats), nv, d, s, v, mu, rank, workspace, workspace_bytes, nullptr, stream);
}

extern "C" int tac_dpca_eig_pop(double* stats, int nv, int d, float* s, float* v, float* mu, int* rank, void* workspace,
                                size_t workspace_bytes, const tac_xchg* x, tac_stream_t stream) {
    TAC_REQUIRE(stats && s && v && mu && rank, "null argument");
    {
        int rc = check_stats_xchg(x, nv, d);
        if (rc) return rc;
    }
    const XchgDev xd = xchg_dev(x);
    TAC_REQUIRE(nv >= 1 && nv <= kMaxViews, "nv must be in [1, %d]", kMaxViews);
    TAC_REQUIRE(d >= 1 && d <= kMaxD, "need 1 <= d <= %d", kMaxD);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (d <= kTriMaxD) {                       // tridiagonalisation + multisection + twisted factorisation (dpca_eig_tri.cuh)
        const size_t need = dpca_eig_tri_workspace_bytes(nv, d);
        if (!workspace || workspace_bytes < need)
            return set_error(TAC_EWORKSPACE, "tac_dpca_eig with d=%d needs %zu workspace bytes", d, need);
        TAC_ALIGNED16(workspace, "workspace");
        const size_t tsmem = dpca_eig_tri_smem_bytes(d);
        TAC_SMEM_OPT_IN(dpca_eig_tri_kernel<false>, dpca_eig_tri_smem_bytes(kTriMaxD));
        TAC_SMEM_OPT_IN(dpca_eig_tri_kernel<true>, dpca_eig_tri_smem_bytes(kTriMaxD));
        TAC_SMEM_OPT_IN(dpca_tri_vectors_kernel, dpca_tri_vectors_smem_bytes());
        double* gws = reinterpret_cast<double*>(workspace);
#if TAC_EIG_SPLIT
        // 1. tridiagonalise (one CTA per view)  2. one warp per eigenpair on d/8 CTAs per view  3. order + outputs
        dpca_eig_tri_kernel<true><<<nv, kTriThreads, tsmem, st>>>(stats, d, s, v, mu, rank, gws, xd); TAC_LAUNCHED(1);
        TAC_CHECK_CUDA(cudaGetLastError());
        dpca_tri_vectors_kernel<<<dim3((d + kVecWarps - 1) / kVecWarps, nv), kVecThreads, dpca_tri_vectors_smem_bytes(), st>>>(d, gws); TAC_LAUNCHED(1);
        TAC_CHECK_CUDA(cudaGetLastError());
        dpca_tri_finish_kernel<<<nv, kTriThreads, 0, st>>>(stats, d, s, v, rank, gws); TAC_LAUNCHED(1);
#else
        dpca_eig_tri_kernel<false><<<nv, kTriThreads, tsmem, st>>>(stats, d, s, v, mu, rank, gws, xd); TAC_LAUNCHED(1);
#endif
        TAC_CHECK_CUDA(cudaGetLastError());
        return TAC_OK;
    }
    if (xd.world > 1) {                                             // Jacobi path (d > 128): generic pop kernel in front
        int rc = tac_xchg_pop(x, stats, stream);
        if (rc) return rc;
    }
    const size_t smem = eig_smem_bytes(d);
    const int use_gws = smem > 160 * 1024;
    if (use_gws) {
        const size_t need = (size_t)nv * d * d * sizeof(double);
        if (!workspace || workspace_bytes < need)
            return set_error(TAC_EWORKSPACE, "tac_dpca_eig with d=%d needs %zu workspace bytes", d, need);
    }
    TAC_CHECK_CUDA(cudaFuncSetAttribute(dpca_eig_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    dpca_eig_kernel<<<nv, kEigThreads, use_gws ? 0 : smem, st>>>(stats, d, s, v, mu, rank, reinterpret_cast<double*>(workspace), use_gws); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

extern "C" int tac_dpca_select(const float* s, const int* rank, int nv, int d, int dim_p, int* idx, int* cnt, tac_stream_t stream) {
    TAC_REQUIRE(s && rank && idx && cnt, "null argument");
    TAC_REQUIRE(nv >= 1 && nv <= kMaxViews && d >= 1 && d <= kMaxD, "bad nv / d");
    TAC_REQUIRE(dim_p >= 0, "dim_p must be >= 0");
    dpca_select_kernel<<<1, 1024, 0, reinterpret_cast<cudaStream_t>(stream)>>>(s, rank, nv, d, dim_p, idx, cnt); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

extern "C" int tac_dpca_project(const float* const* z, int nv, int64_t B, int d, const float* v, const float* mu, const int* idx,
                                const int* cnt, float* const* z_out, void* workspace, size_t workspace_bytes, tac_stream_t stream) {
    TAC_REQUIRE(z && v && mu && idx && cnt && z_out && workspace, "null argument");
    TAC_REQUIRE(nv >= 1 && nv <= kMaxViews && d >= 1 && d <= kMaxD && B >= 1, "bad nv / d / B");
    const size_t need = (size_t)nv * d * d * sizeof(float);
    if (workspace_bytes < need) return set_error(TAC_EWORKSPACE, "tac_dpca_project needs %zu workspace bytes", need);
    ViewPtrs vp;
    for (int i = 0; i < nv; ++i) {
        TAC_REQUIRE(z[i] && z_out[i], "z[%d] / z_out[%d] is null", i, i);
        vp.z[i] = z[i]; vp.out[i] = z_out[i];
    }
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
#ifndef TAC_PROJECT_TC
#define TAC_PROJECT_TC 1
#endif
    if (TAC_PROJECT_TC && d <= kTcFeat) {          // both contractions on tcgen05, one launch (dpca_tc.cuh)
        PjPtrs pp;
        for (int i = 0; i < 8; ++i) { pp.z[i] = vp.z[i < nv ? i : 0]; pp.out[i] = vp.out[i < nv ? i : 0]; }
        TAC_SMEM_OPT_IN(dpca_project_tc_kernel, kPjSmemBytes);
        dpca_project_tc_kernel<<<dim3((unsigned)((B + 127) / 128), nv), kPjThreads, kPjSmemBytes, st>>>(pp, B, d, v, mu, idx, cnt);
        TAC_LAUNCHED(1);
        TAC_CHECK_CUDA(cudaGetLastError());
        return TAC_OK;
    }
    float* P = reinterpret_cast<float*>(workspace);
    dpca_projector_kernel<<<dim3(d, nv), ((d + 31) / 32) * 32, 0, st>>>(v, idx, cnt, d, P); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    dpca_apply_kernel<<<dim3((unsigned)((B + 31) / 32), (d + 31) / 32, nv), 256, 0, st>>>(vp, B, d, P, mu, 1); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

extern "C" int tac_var_topk(const double* stats, int d, int k, float* var, int* top_idx, tac_stream_t stream) {
    TAC_REQUIRE(stats && var && top_idx, "null argument");
    TAC_REQUIRE(d >= 1 && d <= kMaxD * 4 && k >= 1 && k <= d, "need 1 <= k <= d <= %d", kMaxD * 4);
    var_topk_kernel<<<1, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(stats, d, k, var, top_idx); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

// Development aid, NOT part of the ABI and not in the default build (build.py -DTAC_DEBUG_EXPORTS): clock64 stamps of the
// last tac_dpca_eig (d <= 128), view 0.
#ifdef TAC_DEBUG_EXPORTS
extern "C" TAC_API int tac_debug_eig_profile(long long* out8) {
    TAC_CHECK_CUDA(cudaMemcpyFromSymbol(out8, tac::g_tri_prof, 8 * sizeof(long long)));
    return TAC_OK;
}
#endif

extern "C" int tac_nuclear_from_s(const float* s, const int* rank, const double* stats, int nv, int d, float* out, double* aux,
                                  tac_stream_t stream) {
    TAC_REQUIRE(s && rank && stats && out, "null argument");
    TAC_REQUIRE(nv >= 1 && nv <= kMaxViews && d >= 1 && d <= kMaxD, "bad nv / d");
    nuclear_kernel<<<nv, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(s, rank, stats, d, out, aux); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

extern "C" int tac_nuclear_grad_matrix(const float* s, const float* v, const int* rank, const double* aux, int nv, int d, float* gm,
                                       tac_stream_t stream) {
    TAC_REQUIRE(s && v && rank && aux && gm, "null argument");
    TAC_REQUIRE(nv >= 1 && nv <= kMaxViews && d >= 1 && d <= kMaxD, "bad nv / d");
    nuclear_grad_matrix_kernel<<<dim3(d, nv), ((d + 31) / 32) * 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(s, v, rank, aux, d, gm);
    TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

extern "C" int tac_dpca_apply(const float* const* z, int nv, int64_t B, int d, const float* P, const float* mu, int add_mean,
                              float* const* z_out, tac_stream_t stream) {
    TAC_REQUIRE(z && P && mu && z_out, "null argument");
    TAC_REQUIRE(nv >= 1 && nv <= kMaxViews && d >= 1 && d <= kMaxD && B >= 1, "bad nv / d / B");
    ViewPtrs vp;
    for (int i = 0; i < nv; ++i) {
        TAC_REQUIRE(z[i] && z_out[i], "z[%d] / z_out[%d] is null", i, i);
        vp.z[i] = z[i]; vp.out[i] = z_out[i];
    }
    dpca_apply_kernel<<<dim3((unsigned)((B + 31) / 32), (d + 31) / 32, nv), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        vp, B, d, P, mu, add_mean ? 1 : 0); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

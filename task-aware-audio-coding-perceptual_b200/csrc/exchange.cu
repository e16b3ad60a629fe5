// The step's cross-rank exchange as plain peer stores over NVLink (xchg.cuh): generic push / pop kernels, the tail of
// the loss chain (push sums -> pop -> loss scalar in ONE single-CTA kernel), and the CUDA-IPC allocator for callers
// without a symmetric-memory allocator.
#include <stdlib.h>
#include "tac_common.cuh"
#include "xchg.cuh"

namespace tac {

__global__ void __launch_bounds__(256) xchg_push_kernel(const XchgDev x, const double* __restrict__ src) {
    const unsigned seq = xchg_push_seq(x);
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < x.payload; i += (long long)gridDim.x * 256)
        xchg_store(x, seq, (size_t)i, src[i]);
    xchg_push_commit(x, seq, gridDim.x);
}

__global__ void __launch_bounds__(256) xchg_pop_kernel(const XchgDev x, double* __restrict__ dst) {
    const unsigned seq = xchg_pop_seq(x);
    xchg_wait(x, seq);
    __syncthreads();
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < x.payload; i += (long long)gridDim.x * 256)
        dst[i] = xchg_sum(x, seq, (size_t)i);
    __syncthreads();
    if (threadIdx.x == 0) xchg_pop_commit(x, seq, gridDim.x);
}

// One CTA of 64 threads.  sums[n] (this rank) -> all ranks; global sums back into `sums`; loss = sums . inv.
__global__ void __launch_bounds__(64) loss_finalize_kernel(const XchgDev x, double* sums, int n, const double* __restrict__ inv,
                                                          double* __restrict__ loss) {
    __shared__ double sh[64];
    const int j = threadIdx.x;
    double v = j < n ? sums[j] : 0.0;
    if (x.world > 1) {
        const unsigned ps = xchg_push_seq(x);
        if (j < n) xchg_store(x, ps, (size_t)j, v);
        xchg_push_commit(x, ps, 1);
        const unsigned qs = xchg_pop_seq(x);
        xchg_wait(x, qs);
        __syncthreads();
        if (j < n) { v = xchg_sum(x, qs, (size_t)j); sums[j] = v; }
        __syncthreads();
        if (j == 0) xchg_pop_commit(x, qs, 1);
    }
    sh[j] = j < n ? v * inv[j] : 0.0;
    __syncthreads();
    if (j == 0) {
        double acc = 0.0;
        for (int i = 0; i < n; ++i) acc += sh[i];                  // fixed order
        loss[0] = acc;
    }
}

static int check_xchg(const tac_xchg* x) {
    TAC_REQUIRE(x != nullptr, "null exchange descriptor");
    TAC_REQUIRE(x->world >= 1 && x->world <= kXchgMaxWorld, "world must be in [1, %d]", kXchgMaxWorld);
    TAC_REQUIRE(x->rank >= 0 && x->rank < x->world, "rank %d outside world of %d", x->rank, x->world);
    TAC_REQUIRE(x->payload >= 1, "payload must be positive");
    for (int r = 0; r < x->world; ++r) {
        TAC_REQUIRE(x->peer[r] != nullptr, "peer[%d] is null", r);
        TAC_ALIGNED16(x->peer[r], "peer buffer");
    }
    return TAC_OK;
}

// -1: not set by tac_xchg_set_timeout -> environment / default
static std::atomic<long long> g_xchg_timeout_ns{-1};

unsigned long long xchg_timeout_ns() {
    const long long v = g_xchg_timeout_ns.load(std::memory_order_relaxed);
    if (v >= 0) return (unsigned long long)v;
    static const unsigned long long from_env = [] {
        double sec = 600.0;
        if (const char* e = getenv("TAC_XCHG_TIMEOUT_S")) {
            char* end = nullptr;
            const double t = strtod(e, &end);
            if (end != e && t >= 0.0) sec = t;
        }
        return (unsigned long long)(sec * 1e9);
    }();
    return from_env;
}

}  // namespace tac

using namespace tac;

extern "C" int tac_xchg_set_timeout(double seconds) {
    TAC_REQUIRE(seconds == seconds && seconds < 1e9, "timeout must be a number of seconds below 1e9 (0: unbounded, < 0: default)");
    g_xchg_timeout_ns.store(seconds < 0.0 ? -1ll : (long long)(seconds * 1e9), std::memory_order_relaxed);
    return TAC_OK;
}

extern "C" int tac_xchg_status(const tac_xchg* x, unsigned* timed_out) {
    TAC_REQUIRE(timed_out != nullptr, "null argument");
    *timed_out = 0;
    if (!x || x->world <= 1) return TAC_OK;
    int rc = check_xchg(x);
    if (rc) return rc;
    TAC_CHECK_CUDA(cudaMemcpy(timed_out, reinterpret_cast<const unsigned*>(x->peer[x->rank]) + 4, sizeof(unsigned),
                              cudaMemcpyDeviceToHost));
    return TAC_OK;
}

extern "C" size_t tac_xchg_bytes(int world, int64_t payload) {
    if (world < 1 || world > kXchgMaxWorld || payload < 1) return 0;
    return kXchgSlotOff + (size_t)2 * world * (size_t)payload * sizeof(double);
}

extern "C" int tac_xchg_push(const tac_xchg* x, const double* src, tac_stream_t stream) {
    int rc = check_xchg(x);
    if (rc) return rc;
    TAC_REQUIRE(src != nullptr, "null argument");
    if (x->world == 1) return TAC_OK;
    long long g = (x->payload + 255) / 256;
    if (g > 64) g = 64;
    xchg_push_kernel<<<(unsigned)g, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(xchg_dev(x), src); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

extern "C" int tac_xchg_pop(const tac_xchg* x, double* dst, tac_stream_t stream) {
    int rc = check_xchg(x);
    if (rc) return rc;
    TAC_REQUIRE(dst != nullptr, "null argument");
    if (x->world == 1) return TAC_OK;
    long long g = (x->payload + 255) / 256;
    if (g > 64) g = 64;
    xchg_pop_kernel<<<(unsigned)g, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(xchg_dev(x), dst); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

extern "C" int tac_loss_finalize(double* sums, int n, const double* inv, const tac_xchg* x, double* loss, tac_stream_t stream) {
    TAC_REQUIRE(sums && inv && loss, "null argument");
    TAC_REQUIRE(n >= 1 && n <= 64, "n must be in [1, 64]");
    if (x && x->world > 1) {
        int rc = check_xchg(x);
        if (rc) return rc;
        TAC_REQUIRE(x->payload >= n, "exchange payload %lld < n = %d", (long long)x->payload, n);
    }
    loss_finalize_kernel<<<1, 64, 0, reinterpret_cast<cudaStream_t>(stream)>>>(xchg_dev(x), sums, n, inv, loss); TAC_LAUNCHED(1);
    TAC_CHECK_CUDA(cudaGetLastError());
    return TAC_OK;
}

extern "C" int tac_ipc_alloc(size_t bytes, void** ptr, unsigned char* handle64) {
    TAC_REQUIRE(ptr && handle64 && bytes > 0, "null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    void* p = nullptr;
    TAC_CHECK_CUDA(cudaMalloc(&p, bytes));
    cudaError_t e = cudaMemset(p, 0, bytes);
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {
        cudaFree(p);
        return set_error(TAC_ECUDA, "tac_ipc_alloc: %s", cudaGetErrorString(e));
    }
    memcpy(handle64, &h, 64);
    *ptr = p;
    return TAC_OK;
}

extern "C" int tac_ipc_open(const unsigned char* handle64, void** ptr) {
    TAC_REQUIRE(ptr && handle64, "null argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    TAC_CHECK_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return TAC_OK;
}

extern "C" int tac_ipc_close(void* ptr) {
    TAC_REQUIRE(ptr, "null argument");
    TAC_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
    return TAC_OK;
}

extern "C" int tac_ipc_free(void* ptr) {
    TAC_REQUIRE(ptr, "null argument");
    TAC_CHECK_CUDA(cudaFree(ptr));
    return TAC_OK;
}

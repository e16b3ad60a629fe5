// Shared host-side helpers of libtac_spectral: error channel, argument checks, table cache API.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <string.h>
#include <atomic>

#include "../../include/tac_spectral.h"
#include "fft_core.cuh"

namespace tac {

int set_error(int code, const char* fmt, ...);   // stores the message thread-locally, returns code

#define TAC_CHECK_CUDA(expr)                                                                  \
    do {                                                                                      \
        cudaError_t _e = (expr);                                                              \
        if (_e != cudaSuccess)                                                                \
            return tac::set_error(TAC_ECUDA, "%s failed: %s (%s:%d)", #expr,                  \
                                  cudaGetErrorString(_e), __FILE__, __LINE__);                \
    } while (0)

#define TAC_REQUIRE(cond, ...)                                                                \
    do {                                                                                      \
        if (!(cond)) return tac::set_error(TAC_EINVAL, __VA_ARGS__);                          \
    } while (0)

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is a PER-DEVICE property of a kernel: the opt-in is remembered per
// (call site, device), thread-safely, so that a second GPU used from the same process gets it too.
#define TAC_SMEM_OPT_IN(kern, bytes)                                                          \
    do {                                                                                      \
        static std::atomic<unsigned long long> _tac_done{0ull};                               \
        int _tac_dev = 0;                                                                     \
        TAC_CHECK_CUDA(cudaGetDevice(&_tac_dev));                                             \
        if (!((_tac_done.load(std::memory_order_acquire) >> (_tac_dev & 63)) & 1ull)) {       \
            TAC_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes))); \
            _tac_done.fetch_or(1ull << (_tac_dev & 63), std::memory_order_release);           \
        }                                                                                     \
    } while (0)

// Same once-per-(call site, device) guard for the shared-memory carve-out hint (percent of the SM's L1/shared array).
#define TAC_CARVEOUT_ONCE(kern, percent)                                                      \
    do {                                                                                      \
        static std::atomic<unsigned long long> _tac_done{0ull};                               \
        int _tac_dev = 0;                                                                     \
        TAC_CHECK_CUDA(cudaGetDevice(&_tac_dev));                                             \
        if (!((_tac_done.load(std::memory_order_acquire) >> (_tac_dev & 63)) & 1ull)) {       \
            TAC_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, (int)(percent))); \
            _tac_done.fetch_or(1ull << (_tac_dev & 63), std::memory_order_release);           \
        }                                                                                     \
    } while (0)

#define TAC_ALIGNED16(p, name)                                                                \
    do {                                                                                      \
        if ((reinterpret_cast<uintptr_t>(p) & 15u) != 0)                                      \
            return tac::set_error(TAC_EALIGN, "%s must be 16-byte aligned", name);            \
    } while (0)

inline bool supported_nfft(int n) {
    return n == 64 || n == 128 || n == 256 || n == 512 || n == 1024 || n == 2048;
}

// Looks up (building on first use) the device tables of one FFT configuration on the current device.
int get_tables(int n_fft, int win, int normalized, FftTables* out);
void free_blu_tables();      // chirp-z tables of stft_ex.cu (tac_shutdown)

inline int64_t num_frames(int64_t T, int n, int hop, int center) {
    int64_t Tp = T + (center ? 2 * (int64_t)(n / 2) : 0);
    if (Tp < n || hop <= 0) return 0;
    return 1 + (Tp - n) / hop;
}

constexpr int kNumSMs = 148;

// Kernel-launch counter behind tac_launch_count(): every kernel (not memset) this library enqueues.
void count_launches(int n);
#define TAC_LAUNCHED(n) tac::count_launches(n)

}  // namespace tac

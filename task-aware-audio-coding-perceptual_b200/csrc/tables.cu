// Error channel + per-device cache of the immutable FFT tables (window*scale, pass twiddles,
// post-processing twiddles), generated on the host in float64.
#include <math.h>
#include <atomic>
#include <map>
#include <mutex>
#include <tuple>
#include <vector>

#include "tac_common.cuh"

namespace tac {

static thread_local char g_err[512] = "";
static std::atomic<unsigned long long> g_launches{0};

void count_launches(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }

int set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

namespace {
struct Entry { void* base; FftTables t; };
std::mutex g_mu;
std::map<std::tuple<int, int, int, int>, Entry> g_cache;   // (device, n, win, normalized)

void plan_split(int n, int* A, int* BQ) {
    switch (n) {
        case 2048: *A = 32; *BQ = 32; break;
        case 1024: *A = 16; *BQ = 32; break;
        case 512: *A = 16; *BQ = 16; break;
        case 256: *A = 8; *BQ = 16; break;
        case 128: *A = 8; *BQ = 8; break;
        default: *A = 4; *BQ = 8; break;     // 64
    }
}
}  // namespace

int get_tables(int n, int win, int normalized, FftTables* out) {
    if (!supported_nfft(n)) return set_error(TAC_EINVAL, "unsupported n_fft %d (64..2048, power of two)", n);
    if (win <= 0 || win > n) return set_error(TAC_EINVAL, "win_length %d must be in [1, n_fft=%d]", win, n);
    int dev = 0;
    TAC_CHECK_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(g_mu);
    auto key = std::make_tuple(dev, n, win, normalized ? 1 : 0);
    auto it = g_cache.find(key);
    if (it != g_cache.end()) { *out = it->second.t; return TAC_OK; }

    const int M = n / 2;
    int A, BQ;
    plan_split(n, &A, &BQ);
    // periodic Hann of `win` samples, centred inside n_fft like torch.stft pads it
    std::vector<double> w(n, 0.0);
    const int left = (n - win) / 2;
    double ss = 0.0;
    for (int i = 0; i < win; ++i) {
        double v = 0.5 - 0.5 * cos(2.0 * M_PI * (double)i / (double)win);
        w[left + i] = v;
        ss += v * v;
    }
    const double scale = 0.5 * (normalized ? 1.0 / sqrt(ss) : 1.0);   // 0.5: real-FFT post-process
    std::vector<float2> host(3 * (size_t)M);
    float2* win2 = host.data();
    float2* tw = host.data() + M;
    float2* pt = host.data() + 2 * (size_t)M;
    for (int m = 0; m < M; ++m) win2[m] = make_float2((float)(w[2 * m] * scale), (float)(w[2 * m + 1] * scale));
    for (int k1 = 0; k1 < A; ++k1)
        for (int b = 0; b < BQ; ++b) {
            long e = ((long)b * k1) % M;
            double ang = 2.0 * M_PI * (double)e / (double)M;
            tw[k1 * BQ + b] = make_float2((float)cos(ang), (float)sin(ang));
        }
    for (int k = 0; k < M; ++k) {
        double ang = 2.0 * M_PI * (double)k / (double)n;
        pt[k] = make_float2((float)cos(ang), (float)sin(ang));
    }
    void* base = nullptr;
    TAC_CHECK_CUDA(cudaMalloc(&base, host.size() * sizeof(float2)));
    cudaError_t e = cudaMemcpy(base, host.data(), host.size() * sizeof(float2), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        cudaFree(base);
        return set_error(TAC_ECUDA, "table upload failed: %s", cudaGetErrorString(e));
    }
    Entry en;
    en.base = base;
    en.t.win2 = reinterpret_cast<const float2*>(base);
    en.t.tw = en.t.win2 + M;
    en.t.pt = en.t.win2 + 2 * (size_t)M;
    g_cache[key] = en;
    *out = en.t;
    return TAC_OK;
}

}  // namespace tac

extern "C" {

const char* tac_last_error(void) { return tac::g_err; }
int tac_version(void) { return 101; }
uint64_t tac_launch_count(void) { return tac::g_launches.load(std::memory_order_relaxed); }

int tac_prepare_fft(int n_fft, int win_length, int normalized) {
    tac::FftTables t;
    return tac::get_tables(n_fft, win_length, normalized, &t);
}

int tac_shutdown(void) {
    std::lock_guard<std::mutex> lk(tac::g_mu);
    int cur = 0;
    cudaGetDevice(&cur);
    for (auto& kv : tac::g_cache) {
        cudaSetDevice(std::get<0>(kv.first));
        cudaFree(kv.second.base);
    }
    tac::g_cache.clear();
    cudaSetDevice(cur);
    tac::free_blu_tables();
    return TAC_OK;
}

int64_t tac_stft_num_frames(int64_t T, int n_fft, int hop, int center) {
    return tac::num_frames(T, n_fft, hop, center);
}

}  // extern "C"

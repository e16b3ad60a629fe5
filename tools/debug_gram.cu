// Debug harness (not product): runs dpca_gram_tc_kernel on exactly representable inputs and compares
// with a host Gram.  nvcc -gencode arch=compute_100a,code=sm_100a -I. tools/debug_gram.cu -o /tmp/debug_gram
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cmath>
#include "../task-aware-audio-coding-perceptual_b200/csrc/dpca_tc.cuh"
using namespace tac;
int main(int argc, char** argv) {
    const int B = argc > 1 ? atoi(argv[1]) : 64, d = argc > 2 ? atoi(argv[2]) : 128;
    const int mode = argc > 3 ? atoi(argv[3]) : 0;
    std::vector<float> z((size_t)B * d);
    srand(1);
    for (int b = 0; b < B; ++b)
        for (int f = 0; f < d; ++f)
            z[(size_t)b * d + f] = mode == 0 ? (float)((rand() % 7) - 3) : (f == b % d ? 1.0f : 0.0f);
    std::vector<double> stats(1 + d + (size_t)d * d, 0.0);
    stats[0] = 1.0;                                   // mean = 0
    float* dz; double* dst; float* dp;
    const int nsplit = (B + kTcRowsPerCta - 1) / kTcRowsPerCta;
    cudaMalloc(&dz, z.size() * 4); cudaMalloc(&dst, stats.size() * 8); cudaMalloc(&dp, (size_t)nsplit * 128 * 128 * 4);
    cudaMemcpy(dz, z.data(), z.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dst, stats.data(), stats.size() * 8, cudaMemcpyHostToDevice);
    cudaMemset(dp, 0xff, (size_t)nsplit * 128 * 128 * 4);
    cudaFuncSetAttribute(dpca_gram_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes);
    dpca_gram_tc_kernel<<<dim3(nsplit, 1), kTcThreads, kTcSmemBytes>>>(dz, dz, dz, dz, dz, dz, dz, dz, B, d, dst, dp);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel: %s\n", cudaGetErrorString(e));
    std::vector<float> p((size_t)nsplit * 128 * 128);
    cudaMemcpy(p.data(), dp, p.size() * 4, cudaMemcpyDeviceToHost);
    double maxerr = 0; int bad = 0;
    for (int i = 0; i < d; ++i)
        for (int j = 0; j < d; ++j) {
            double ref = 0;
            for (int b = 0; b < B; ++b) ref += (double)z[(size_t)b * d + i] * z[(size_t)b * d + j];
            double got = 0;
            for (int s = 0; s < nsplit; ++s) got += p[((size_t)s * 128 + i) * 128 + j];
            double er = fabs(got - ref);
            if (er > 1e-3) { if (bad < 12) printf("  G[%d][%d] got %g ref %g\n", i, j, got, ref); ++bad; }
            if (er > maxerr || std::isnan(er)) maxerr = er;
        }
    printf("B=%d d=%d mode=%d: max err %g, bad %d / %d\n", B, d, mode, maxerr, bad, d * d);
    printf("row0: "); for (int j = 0; j < 12; ++j) printf("%g ", p[j]); printf("\n");
    return 0;
}

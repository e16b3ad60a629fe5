// Micro-benchmark (B200): issue rate and latency of packed FP32 (FFMA2 / FADD2) against scalar FFMA / FADD at
// 1, 2, 4 warps per scheduler.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o f32x2_bench f32x2_bench.cu
#include <cuda_runtime.h>
#include <cstdio>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void upk(u64 v, float& a, float& b) { asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 r; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ float fma1(float a, float b, float c) { float r; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
__device__ __forceinline__ float add1(float a, float b) { float r; asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }

constexpr int ITERS = 2048;
// MODE 0: scalar FFMA, ILP chains; 1: FFMA2; 2: scalar FADD; 3: FADD2; 4: scalar FFMA x2 (same flops as mode 1)
template <int MODE, int ILP>
__global__ void bench(float* out, long long* cyc, float seed) {
    float a[2 * ILP];
    u64 p[ILP];
#pragma unroll
    for (int i = 0; i < 2 * ILP; ++i) a[i] = seed + threadIdx.x * 1e-3f + i;
#pragma unroll
    for (int i = 0; i < ILP; ++i) p[i] = pk(a[2 * i], a[2 * i + 1]);
    const float m = 0.999f + seed * 1e-6f, c = 1e-3f;
    const u64 m2 = pk(m, m), c2 = pk(c, c);
    long long t0 = clock64();
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) {
            if (MODE == 0) a[i] = fma1(a[i], m, c);
            if (MODE == 1) p[i] = fma2(p[i], m2, c2);
            if (MODE == 2) a[i] = add1(a[i], c);
            if (MODE == 3) p[i] = add2(p[i], c2);
            if (MODE == 4) { a[2 * i] = fma1(a[2 * i], m, c); a[2 * i + 1] = fma1(a[2 * i + 1], m, c); }
        }
    }
    long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < ILP; ++i) { float x, y; upk(p[i], x, y); s += x + y + a[2 * i] + a[2 * i + 1]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int MODE, int ILP>
void run(const char* name, int warps_per_sm, float* out, long long* cyc) {
    const int threads = warps_per_sm * 32;
    bench<MODE, ILP><<<148, threads>>>(out, cyc, 1.0f);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    bench<MODE, ILP><<<148, threads>>>(out, cyc, 1.0f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    const double inst = (double)ITERS * ILP * (MODE == 4 ? 2 : 1);                 // per warp
    const double per_smsp = inst * (warps_per_sm / 4.0) / (double)h;                // warp-instr / clk / SMSP
    const double flop_per_inst = (MODE == 1 ? 128 : MODE == 3 ? 64 : MODE == 2 ? 32 : 64);
    printf("{\"mode\": \"%s\", \"ilp\": %d, \"warps_per_smsp\": %.1f, \"cycles\": %lld, \"inst_per_clk_smsp\": %.3f, "
           "\"flop_per_clk_sm\": %.1f, \"ms\": %.4f}\n", name, ILP, warps_per_sm / 4.0, h, per_smsp,
           per_smsp * 4 * flop_per_inst, ms);
}

int main() {
    float* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
    for (int w : {4, 8, 16}) {
        run<0, 8>("ffma", w, out, cyc);
        run<1, 8>("ffma2", w, out, cyc);
        run<4, 8>("ffma_x2", w, out, cyc);
        run<2, 8>("fadd", w, out, cyc);
        run<3, 8>("fadd2", w, out, cyc);
    }
    run<0, 1>("ffma_lat", 4, out, cyc);
    run<1, 1>("ffma2_lat", 4, out, cyc);
    run<2, 1>("fadd_lat", 4, out, cyc);
    run<3, 1>("fadd2_lat", 4, out, cyc);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}

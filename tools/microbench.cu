// Pipe-rate microbenchmarks for sm_100a (development aid): FFMA, FFMA2 (fma.rn.f32x2), MUFU.EX2, F2F, DFMA.
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 4096
#define CHAINS 16

__global__ void k_ffma(float* out, float a, float b) {
    float v[CHAINS];
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) v[i] = threadIdx.x + i;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) v[i] = fmaf(v[i], a, b);
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_ffma2(float* out, float a, float b) {
    unsigned long long v[CHAINS], A, B;
    asm("mov.b64 %0, {%1, %1};" : "=l"(A) : "f"(a));
    asm("mov.b64 %0, {%1, %1};" : "=l"(B) : "f"(b));
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) { float x = threadIdx.x + i; asm("mov.b64 %0, {%1, %1};" : "=l"(v[i]) : "f"(x)); }
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(v[i]) : "l"(A), "l"(B));
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) { float lo, hi; asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v[i])); s += lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_mufu(float* out, float a) {
    float v[CHAINS];
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) v[i] = (threadIdx.x + i) * 1e-3f;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(v[i]));
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_f2f(double* out, float a) {
    float v[CHAINS];
    double d[CHAINS];
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) v[i] = (threadIdx.x + i) * 1e-3f;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) { asm volatile("cvt.f64.f32 %0, %1;" : "=d"(d[i]) : "f"(v[i])); asm volatile("cvt.rn.f32.f64 %0, %1;" : "=f"(v[i]) : "d"(d[i])); }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) s += d[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_dfma(double* out, double a, double b) {
    double v[CHAINS];
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) v[i] = threadIdx.x + i;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) v[i] = fma(v[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// FFMA + MUFU mixed: does MUFU issue overlap with a saturated FMA pipe?
__global__ void k_mix(float* out, float a, float b) {
    float v[CHAINS], m[4];
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) v[i] = threadIdx.x + i;
#pragma unroll
    for (int i = 0; i < 4; ++i) m[i] = i * 1e-3f;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) { v[i] = fmaf(v[i], a, b); if ((i & 3) == 3) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(m[i >> 2])); }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + m[0] + m[1] + m[2] + m[3];
}

template <typename F>
float timeit(F f) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); return ms;
}

int main() {
    int dev; cudaGetDevice(&dev);
    cudaDeviceProp p; cudaGetDeviceProperties(&p, dev);
    int clk_khz; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, dev);
    printf("%s SMs=%d clock=%d kHz\n", p.name, p.multiProcessorCount, clk_khz);
    int blocks = p.multiProcessorCount * 4, threads = 256;
    void* buf; cudaMalloc(&buf, (size_t)blocks * threads * 8);
    double lanes = (double)blocks * threads * ITERS * CHAINS;
    float ms;
    ms = timeit([&] { k_ffma<<<blocks, threads>>>((float*)buf, 1.0001f, 0.5f); });
    printf("FFMA   : %.3f ms  %.2f T lane-fma/s  (%.1f per clk per SM @1.965GHz)\n", ms, lanes / ms * 1e-9, lanes / (ms * 1e-3) / 1.965e9 / p.multiProcessorCount);
    ms = timeit([&] { k_ffma2<<<blocks, threads>>>((float*)buf, 1.0001f, 0.5f); });
    printf("FFMA2  : %.3f ms  %.2f T lane-fma/s  (%.1f per clk per SM)\n", ms, 2 * lanes / ms * 1e-9, 2 * lanes / (ms * 1e-3) / 1.965e9 / p.multiProcessorCount);
    ms = timeit([&] { k_mufu<<<blocks, threads>>>((float*)buf, 1.0f); });
    printf("MUFU   : %.3f ms  %.2f T ex2/s  (%.1f per clk per SM)\n", ms, lanes / ms * 1e-9, lanes / (ms * 1e-3) / 1.965e9 / p.multiProcessorCount);
    ms = timeit([&] { k_f2f<<<blocks, threads>>>((double*)buf, 1.0f); });
    printf("F2F x2 : %.3f ms  %.2f T cvt/s  (%.1f per clk per SM)\n", ms, 2 * lanes / ms * 1e-9, 2 * lanes / (ms * 1e-3) / 1.965e9 / p.multiProcessorCount);
    ms = timeit([&] { k_dfma<<<blocks, threads>>>((double*)buf, 1.0001, 0.5); });
    printf("DFMA   : %.3f ms  %.2f T dfma/s  (%.1f per clk per SM)\n", ms, lanes / ms * 1e-9, lanes / (ms * 1e-3) / 1.965e9 / p.multiProcessorCount);
    ms = timeit([&] { k_mix<<<blocks, threads>>>((float*)buf, 1.0001f, 0.5f); });
    printf("FFMA+MUFU(4:1): %.3f ms  %.2f T lane-fma/s (%.1f fma per clk per SM)\n", ms, lanes / ms * 1e-9, lanes / (ms * 1e-3) / 1.965e9 / p.multiProcessorCount);
    return 0;
}

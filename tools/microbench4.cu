// FFMA2 dependent-issue latency and FFMA2/MUFU co-issue on sm_100a (development aid).
//   k_chain<C>: every warp runs C independent FFMA2 chains; launched with W warps per SM sub-partition (block = 128*W
//   threads, one block per SM).  Reports FFMA2 lane-FMAs per clock per SM: with C*W large it is the pipe rate, with
//   C*W small it exposes the dependent-issue latency (clocks per FFMA2 = 2*128/rate * ... see printout).
//   k_mix<C, M>: the same chains with M MUFU.EX2 (independent values) interleaved per C*8 FFMA2 -- the leaf-step mix.
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 2048

template <int C>
__global__ void k_chain(float* out, float a, float b, long long* clk) {
    unsigned long long v[C], A, B;
    asm("mov.b64 %0, {%1, %1};" : "=l"(A) : "f"(a));
    asm("mov.b64 %0, {%1, %1};" : "=l"(B) : "f"(b));
#pragma unroll
    for (int i = 0; i < C; ++i) { float x = threadIdx.x + i; asm("mov.b64 %0, {%1, %1};" : "=l"(v[i]) : "f"(x)); }
    long long t0 = clock64();
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int rep = 0; rep < 8; ++rep)
#pragma unroll
            for (int i = 0; i < C; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(v[i]) : "l"(A), "l"(B));
    }
    long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < C; ++i) { float lo, hi; asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v[i])); s += lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}

template <int C, int M>
__global__ void k_mix(float* out, float a, float b, long long* clk) {
    unsigned long long v[C], A, B;
    float e[8];
    asm("mov.b64 %0, {%1, %1};" : "=l"(A) : "f"(a));
    asm("mov.b64 %0, {%1, %1};" : "=l"(B) : "f"(b));
#pragma unroll
    for (int i = 0; i < C; ++i) { float x = threadIdx.x + i; asm("mov.b64 %0, {%1, %1};" : "=l"(v[i]) : "f"(x)); }
#pragma unroll
    for (int i = 0; i < 8; ++i) e[i] = (threadIdx.x + i) * 1e-3f;
    long long t0 = clock64();
    for (int it = 0; it < ITERS; ++it) {
        int m = 0;
#pragma unroll
        for (int rep = 0; rep < 8; ++rep) {
#pragma unroll
            for (int i = 0; i < C; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(v[i]) : "l"(A), "l"(B));
            // M MUFUs spread over the 8 repetitions
#pragma unroll
            for (int k = 0; k < (M + 7 - rep) / 8; ++k) { asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(e[m & 7])); ++m; }
        }
    }
    long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < C; ++i) { float lo, hi; asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v[i])); s += lo + hi; }
#pragma unroll
    for (int i = 0; i < 8; ++i) s += e[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}

template <typename K>
static void run(const char* name, K kernel, int chains, int mufu, int warps_per_sched, float* out, long long* clk) {
    int threads = 128 * warps_per_sched;
    kernel<<<148, threads>>>(out, 1.0000001f, 1e-9f, clk);
    cudaDeviceSynchronize();
    kernel<<<148, threads>>>(out, 1.0000001f, 1e-9f, clk);
    cudaDeviceSynchronize();
    long long c;
    cudaMemcpy(&c, clk, sizeof(c), cudaMemcpyDeviceToHost);
    double ffma2_per_warp = (double)ITERS * 8 * chains;
    double lane_fma_per_clk_sm = ffma2_per_warp * 64.0 * 4 * warps_per_sched / (double)c;
    printf("%-8s chains=%2d mufu/8rep=%2d warps/sched=%d : %.1f clk per FFMA2 per warp, %.1f lane-FMA/clk/SM (%.0f%% of 128)\n",
           name, chains, mufu, warps_per_sched, (double)c / ffma2_per_warp, lane_fma_per_clk_sm,
           100.0 * lane_fma_per_clk_sm / 128.0);
}

int main() {
    float* out;
    long long* clk;
    cudaMalloc(&out, 148 * 1024 * sizeof(float));
    cudaMalloc(&clk, sizeof(long long));
    for (int w : {1, 2, 4}) {
        run("chain", k_chain<1>, 1, 0, w, out, clk);
        run("chain", k_chain<2>, 2, 0, w, out, clk);
        run("chain", k_chain<4>, 4, 0, w, out, clk);
        run("chain", k_chain<8>, 8, 0, w, out, clk);
        run("chain", k_chain<16>, 16, 0, w, out, clk);
    }
    for (int w : {1, 2, 4}) {
        run("mix", k_mix<4, 8>, 4, 8, w, out, clk);      // 32 FFMA2 : 8 MUFU  (one skewed phase)
        run("mix", k_mix<8, 16>, 8, 16, w, out, clk);    // 64 FFMA2 : 16 MUFU (one leaf step)
        run("mix", k_mix<8, 8>, 8, 8, w, out, clk);      // 64 FFMA2 : 8 MUFU
        run("mix", k_mix<16, 16>, 16, 16, w, out, clk);  // 128 : 16 (flat layout ratio)
    }
    return 0;
}

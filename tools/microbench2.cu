// FFMA2 operand-form throughput on sm_100a (development aid): which FFMA2 forms run at full FMA-pipe rate?
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 2048

// form A: pair * scalar-register-broadcast + pair, 8 accumulators, 2 z pairs x 4 scalars per step (our loop shape)
__global__ void k_scalar_reg(float* out, const float* in) {
    float2 z[8][2], t[4][2];
    float w[8][4];
    for (int i = 0; i < 8; ++i) { for (int p = 0; p < 2; ++p) z[i][p] = make_float2(in[i * 2 + p], in[i * 2 + p + 1]); for (int q = 0; q < 4; ++q) w[i][q] = in[32 + i * 4 + q + threadIdx.x % 2]; }
    for (int q = 0; q < 4; ++q) for (int p = 0; p < 2; ++p) t[q][p] = make_float2(0.f, 0.f);
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int p = 0; p < 2; ++p)
#pragma unroll
                for (int q = 0; q < 4; ++q) t[q][p] = __ffma2_rn(z[i][p], make_float2(w[i][q], w[i][q]), t[q][p]);
    }
    float s = 0;
    for (int q = 0; q < 4; ++q) for (int p = 0; p < 2; ++p) s += t[q][p].x + t[q][p].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// form B: same but the scalar comes from constant memory via a uniform register
__constant__ float cw[64];
__global__ void k_scalar_ur(float* out, const float* in) {
    float2 z[8][2], t[4][2];
    for (int i = 0; i < 8; ++i) for (int p = 0; p < 2; ++p) z[i][p] = make_float2(in[i * 2 + p], in[i * 2 + p + 1]);
    for (int q = 0; q < 4; ++q) for (int p = 0; p < 2; ++p) t[q][p] = make_float2(0.f, 0.f);
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int p = 0; p < 2; ++p)
#pragma unroll
                for (int q = 0; q < 4; ++q) { float w = cw[(i * 4 + q + it) & 63]; t[q][p] = __ffma2_rn(z[i][p], make_float2(w, w), t[q][p]); }
    }
    float s = 0;
    for (int q = 0; q < 4; ++q) for (int p = 0; p < 2; ++p) s += t[q][p].x + t[q][p].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// form C: plain FFMA, scalar z * scalar w (register) + acc, 16 accumulators (4 rows x 4 SVs)
__global__ void k_ffma_reg(float* out, const float* in) {
    float z[8][4], t[4][4], w[8][4];
    for (int i = 0; i < 8; ++i) for (int r = 0; r < 4; ++r) { z[i][r] = in[i * 4 + r]; w[i][r] = in[32 + i * 4 + r + threadIdx.x % 2]; }
    for (int q = 0; q < 4; ++q) for (int r = 0; r < 4; ++r) t[q][r] = 0.f;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int q = 0; q < 4; ++q) t[q][r] = fmaf(z[i][r], w[i][q], t[q][r]);
    }
    float s = 0;
    for (int q = 0; q < 4; ++q) for (int r = 0; r < 4; ++r) s += t[q][r];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// form D: pair * pair + pair with distinct registers (duplicated weights)
__global__ void k_pair_reg(float* out, const float* in) {
    float2 z[8][2], t[4][2], w[8][4];
    for (int i = 0; i < 8; ++i) { for (int p = 0; p < 2; ++p) z[i][p] = make_float2(in[i * 2 + p], in[i * 2 + p + 1]); for (int q = 0; q < 4; ++q) w[i][q] = make_float2(in[32 + i * 4 + q], in[33 + i * 4 + q + threadIdx.x % 2]); }
    for (int q = 0; q < 4; ++q) for (int p = 0; p < 2; ++p) t[q][p] = make_float2(0.f, 0.f);
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int p = 0; p < 2; ++p)
#pragma unroll
                for (int q = 0; q < 4; ++q) t[q][p] = __ffma2_rn(z[i][p], w[i][q], t[q][p]);
    }
    float s = 0;
    for (int q = 0; q < 4; ++q) for (int p = 0; p < 2; ++p) s += t[q][p].x + t[q][p].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename K>
void run(const char* name, K k, int threads, int blocks_per_sm, double lane_fma_per_thread, float* out, float* in) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    int blocks = 148 * blocks_per_sm;
    k<<<blocks, threads>>>(out, in); cudaDeviceSynchronize();
    cudaEventRecord(a); k<<<blocks, threads>>>(out, in); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    double rate = lane_fma_per_thread * threads * blocks / (ms * 1e-3);
    printf("%-28s warps/SM=%2d  %.3f ms  %.1f lane-FMA/clk/SM (@1.92GHz)  err=%s\n", name, threads * blocks_per_sm / 32, ms, rate / 1.92e9 / 148, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    float *out, *in; cudaMalloc(&out, 148 * 8 * 1024 * 4); cudaMalloc(&in, 4096); cudaMemset(in, 0, 4096);
    double per = (double)ITERS * 64 * 2;   // 64 FFMA2 (= 128 lane-FMA) per iteration
    for (int w : {1, 2, 4}) {
        run("FFMA2 pair*scalarReg+pair", k_scalar_reg, 256, w, per, out, in);
        run("FFMA2 pair*scalarUR+pair", k_scalar_ur, 256, w, per, out, in);
        run("FFMA  4x4 regs", k_ffma_reg, 256, w, per, out, in);
        run("FFMA2 pair*pair+pair", k_pair_reg, 256, w, per, out, in);
    }
    return 0;
}

// FFMA2 operand patterns (development aid): which source-operand mix keeps fma.rn.f32x2 below 128 lane-FMA/clk/SM?
//   P0: acc[c] = W * Z + acc[c]           one 64-bit weight pair, one 64-bit z pair
//   P1: acc[c] = W * (z,z) + acc[c]       z as a broadcast 32-bit scalar (SASS R.F32), one scalar
//   P2: as P1, 32 different scalars (one per FFMA2, the kernel's standardised features)
//   P3: as P2, plus the weight pair changes every 4 FFMA2 (8 different pairs in registers)
//   P4: as P3 with 16 accumulators instead of 8
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 4096

template <int PAT, int C>
__global__ void __launch_bounds__(128, 4) k_pat(float* out, const float* in) {
    float2 acc[C], w[8];
    float z[32];
#pragma unroll
    for (int i = 0; i < C; ++i) acc[i] = make_float2(threadIdx.x + i, 1.f);
#pragma unroll
    for (int i = 0; i < 8; ++i) w[i] = make_float2(in[i], in[i + 8]);
#pragma unroll
    for (int i = 0; i < 32; ++i) z[i] = in[16 + i];
    const float2 Z = make_float2(in[50], in[51]);
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int k = 0; k < 32; ++k) {
            const int c = k % C;
            if (PAT == 0) acc[c] = __ffma2_rn(w[0], Z, acc[c]);
            else if (PAT == 1) acc[c] = __ffma2_rn(w[0], make_float2(z[0], z[0]), acc[c]);
            else if (PAT == 2) acc[c] = __ffma2_rn(w[0], make_float2(z[k], z[k]), acc[c]);
            else acc[c] = __ffma2_rn(w[(k / 4) % 8], make_float2(z[k], z[k]), acc[c]);
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < C; ++i) s += acc[i].x + acc[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename K>
static void run(const char* name, K kernel, float* out, const float* in) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    kernel<<<148 * 4, 128>>>(out, in);
    cudaEventRecord(e0);
    kernel<<<148 * 4, 128>>>(out, in);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    double lane_fma = 148.0 * 4 * 128 * (double)ITERS * 32 * 2;
    printf("%-40s %.3f ms  %.1f lane-FMA/clk/SM at 1.9 GHz\n", name, ms, lane_fma / (ms * 1e-3) / 1.9e9 / 148.0);
}

int main() {
    float *out, *in;
    cudaMalloc(&out, 148 * 4 * 128 * sizeof(float));
    cudaMalloc(&in, 64 * sizeof(float));
    float h[64];
    for (int i = 0; i < 64; ++i) h[i] = 1e-3f * (i + 1);
    cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice);
    run("P0 pair * pair", k_pat<0, 8>, out, in);
    run("P1 pair * scalar (one)", k_pat<1, 8>, out, in);
    run("P2 pair * scalar (32 regs)", k_pat<2, 8>, out, in);
    run("P3 8 pairs * 32 scalars, 8 acc", k_pat<3, 8>, out, in);
    run("P4 8 pairs * 32 scalars, 16 acc", k_pat<3, 16>, out, in);
    run("P5 8 pairs * 32 scalars, 4 acc", k_pat<3, 4>, out, in);
    return 0;
}
